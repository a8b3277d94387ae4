/*
 * shim.h — force-included (-include) when compiling the UNMODIFIED reference
 * sources under /root/reference/source on Linux.  It supplies what the
 * reference expects from <windows.h> / math/base.h (whose platform check is an
 * #error on Linux, reference source/math/base.h:49-51).  TEST INFRASTRUCTURE:
 * nothing in the product includes this file.
 */
#ifndef FS2_REF_SHIM_H_
#define FS2_REF_SHIM_H_

#define __BASE_TYPES_H__ /* skips reference math/base.h and window/base_types.h */

#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>
#include <algorithm>
#include <memory>
#include <utility>

namespace base {
typedef int64_t int64;
typedef int32_t int32;
typedef int16_t int16;
typedef int8_t int8;
typedef uint64_t uint64;
typedef uint32_t uint32;
typedef uint16_t uint16;
typedef uint8_t uint8;
typedef float float32;
typedef double float64;
}  // namespace base

/* constants of reference math/base.h:80-102 */
#define BASE_PI (3.14159262f)
#define BASE_2PI (6.28318524f)
#define BASE_INFINITY (1.0e15f)
#define BASE_EPSILON (1.0e-5f)
#define BASE_LOG2 (0.3010299956639f)
#define BASE_MAX_INT64 (0x7FFFFFFFFFFFFFFF)
#define BASE_MAX_INT32 (0x7FFFFFFF)
#define BASE_MAX_INT16 (0x7FFF)
#define BASE_MAX_INT8 (0x7F)
#define BASE_MAX_UINT64 (0xFFFFFFFFFFFFFFFF)
#define BASE_MAX_UINT32 (0xFFFFFFFF)
#define BASE_MAX_UINT16 (0xFFFF)
#define BASE_MAX_UINT8 (0xFF)
#define BASE_MIN_INT64 (-BASE_MAX_INT64 - 1)
#define BASE_MIN_INT32 (-BASE_MAX_INT32 - 1)
#define BASE_MIN_INT16 (-BASE_MAX_INT16 - 1)
#define BASE_MIN_INT8 (-BASE_MAX_INT8 - 1)
#define BASE_PARAM_CHECK (1)

#define MAX_PATH 260
#define BASE_PLATFORM_MACOSX /* selects the gettimeofday branch, engine.cpp:22 */
#define sscanf_s sscanf      /* the extra size argument is ignored */

/* Win32 min/max macros promote through the ternary operator; mirror that
 * (used at reference scene.cpp:39 and material.cpp:333).  The result type is
 * decayed: for two lvalues of one type the conditional is an lvalue, and
 * returning it by reference would dangle. */
#include <type_traits>
namespace base {
template <class A, class B>
inline typename std::decay<decltype(true ? std::declval<A>() : std::declval<B>())>::type min(A a, B b) {
  return a < b ? a : b;
}
template <class A, class B>
inline typename std::decay<decltype(true ? std::declval<A>() : std::declval<B>())>::type max(A a, B b) {
  return a > b ? a : b;
}
}  // namespace base

#endif
