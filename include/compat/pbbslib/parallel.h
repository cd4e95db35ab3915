// compat: the slice of pbbslib/parallel.h the reference's callers use (parallel_for, num_workers) lives in wharfmh_b200.hpp
#include "../../wharfmh_b200.hpp"
