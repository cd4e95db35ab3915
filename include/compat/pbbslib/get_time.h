// compat: the reference's timer (pbbslib/get_time.h:9-81) lives in wharfmh_b200.hpp
#include "../../wharfmh_b200.hpp"
