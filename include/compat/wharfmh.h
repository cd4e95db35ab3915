// compat/wharfmh.h - lets a translation unit written against the reference (`#include <wharfmh.h>`, graph/wharfmh.h)
// compile unchanged: put include/compat before the reference's own include directories and link libwharfmh_b200.so.
// The reference's headers open namespace std at global scope (libs/compressed_trees/common/IO.h:22) and its drivers
// rely on it (`string`, `cout`, `endl` unqualified), so this forwarding header does the same; wharfmh_b200.hpp
// itself does not.
#ifndef WHARFMH_B200_COMPAT_WHARFMH_H
#define WHARFMH_B200_COMPAT_WHARFMH_H

#include "../wharfmh_b200.hpp"

using namespace std;

static const string default_file_name = "";   // libs/compressed_trees/graph/api.h:14 (the drivers' -f default)

#endif
