// check.h — error plumbing: CUDA runtime/driver checks that RETURN a status instead of exiting.
// Replaces the reference's cutilSafeCallNoSync / cutilCheckMsg (cutil_inline_runtime.h:267-293,343-360),
// which print and exit(-1).
#pragma once
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include "b2r.h"

namespace b2r {

// Thread-local detail string returned by b2r_last_error_string().
char* last_error_buf();
constexpr int kLastErrorLen = 512;

inline b2r_status_t fail(b2r_status_t s, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buf(), kLastErrorLen, fmt, ap);
  va_end(ap);
  return s;
}

}  // namespace b2r

#define B2R_CUDA_CHECK(expr)                                                                             \
  do {                                                                                                   \
    cudaError_t _e = (expr);                                                                             \
    if (_e != cudaSuccess) {                                                                             \
      return ::b2r::fail(B2R_ERR_CUDA, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
    }                                                                                                    \
  } while (0)

#define B2R_CU_CHECK(expr)                                                                      \
  do {                                                                                          \
    CUresult _r = (expr);                                                                       \
    if (_r != CUDA_SUCCESS) {                                                                   \
      return ::b2r::fail(B2R_ERR_CUDA, "%s:%d: %s -> CUresult %d", __FILE__, __LINE__, #expr, (int)_r); \
    }                                                                                           \
  } while (0)

#define B2R_CHECK(expr)                 \
  do {                                  \
    b2r_status_t _s = (expr);           \
    if (_s != B2R_OK) return _s;        \
  } while (0)

#define B2R_REQUIRE(cond, ...)                                   \
  do {                                                           \
    if (!(cond)) return ::b2r::fail(B2R_ERR_INVALID, __VA_ARGS__); \
  } while (0)
