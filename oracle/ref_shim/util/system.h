/* Stand-in for MVE's <util/system.h>: byte-order macro and byteswap used by libs/util/numpy_io.h.  TEST INFRASTRUCTURE ONLY. */
#ifndef ORACLE_REF_SHIM_UTIL_SYSTEM_H
#define ORACLE_REF_SHIM_UTIL_SYSTEM_H
#include <cerrno>
#include <sstream>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>
#include "util/exception.h"
#define HOST_BYTEORDER_LE 1
namespace util { template <int N> inline void byteswap(void *) {} }
#endif
