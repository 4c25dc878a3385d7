// oracle.cpp — CPU oracles and host-side data generation (no CUDA dependency).
// Restates what the reference's sumreduceCPU / minreduceCPU / maxreduceCPU do (reduction.cpp:214-249:
// Kahan-compensated sum, linear min, linear max) with defined semantics for the cases the reference leaves
// open: integer sums wrap (two's complement), NaNs propagate through min/max, count==0 gives the identity.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>

#include "../util/check.h"
#include "b2r.h"
#include "rng.h"

namespace b2r {

char* last_error_buf() {
  static thread_local char buf[kLastErrorLen] = {0};
  return buf;
}

namespace {

template <typename T>
T min_nan(T a, T b) {  // NaN-propagating
  if (a != a) return a;
  if (b != b) return b;
  return b < a ? b : a;
}
template <typename T>
T max_nan(T a, T b) {
  if (a != a) return a;
  if (b != b) return b;
  return b > a ? b : a;
}

// Neumaier (Kahan-Babuska) summation in long double: error is O(eps_ld) independent of count.
template <typename T>
long double kahan_sum(const T* x, uint64_t n) {
  long double s = 0.0L, c = 0.0L;
  for (uint64_t i = 0; i < n; ++i) {
    long double v = (long double)x[i];
    long double t = s + v;
    if (std::isfinite((double)t)) {
      if (fabsl(s) >= fabsl(v)) c += (s - t) + v;
      else c += (v - t) + s;
    }
    s = t;
  }
  return s + c;
}

template <typename T>
void reduce_fp(const T* in, T* out, uint64_t n, b2r_op_t op) {
  if (op == B2R_SUM) {
    *out = (T)kahan_sum(in, n);
  } else if (op == B2R_MIN) {
    T m = std::numeric_limits<T>::infinity();
    for (uint64_t i = 0; i < n; ++i) m = min_nan(m, in[i]);
    *out = m;
  } else {
    T m = -std::numeric_limits<T>::infinity();
    for (uint64_t i = 0; i < n; ++i) m = max_nan(m, in[i]);
    *out = m;
  }
}

void reduce_i32(const int32_t* in, int32_t* out, uint64_t n, b2r_op_t op) {
  if (op == B2R_SUM) {
    uint32_t s = 0;
    for (uint64_t i = 0; i < n; ++i) s += (uint32_t)in[i];
    *out = (int32_t)s;
  } else if (op == B2R_MIN) {
    int32_t m = std::numeric_limits<int32_t>::max();
    for (uint64_t i = 0; i < n; ++i) m = in[i] < m ? in[i] : m;
    *out = m;
  } else {
    int32_t m = std::numeric_limits<int32_t>::min();
    for (uint64_t i = 0; i < n; ++i) m = in[i] > m ? in[i] : m;
    *out = m;
  }
}

void reduce_bf16(const uint16_t* in, uint16_t* out, uint64_t n, b2r_op_t op) {
  if (op == B2R_SUM) {
    // every bf16 is exactly a double; a double accumulator is exact for < 2^45 elements of equal exponent,
    // and Neumaier keeps the rest. One rounding at the end (RNE), like the device path (fp32 accumulate).
    long double s = 0.0L, c = 0.0L;
    for (uint64_t i = 0; i < n; ++i) {
      long double v = (long double)bf16_bits_to_f32(in[i]);
      long double t = s + v;
      if (std::isfinite((double)t)) {
        if (fabsl(s) >= fabsl(v)) c += (s - t) + v;
        else c += (v - t) + s;
      }
      s = t;
    }
    *out = f32_to_bf16_bits((float)(s + c));
  } else if (op == B2R_MIN) {
    float m = std::numeric_limits<float>::infinity();
    for (uint64_t i = 0; i < n; ++i) m = min_nan(m, bf16_bits_to_f32(in[i]));
    *out = f32_to_bf16_bits(m);
  } else {
    float m = -std::numeric_limits<float>::infinity();
    for (uint64_t i = 0; i < n; ++i) m = max_nan(m, bf16_bits_to_f32(in[i]));
    *out = f32_to_bf16_bits(m);
  }
}

template <typename T>
void combine_fp(T* acc, const T* in, uint64_t n, b2r_op_t op) {
  if (op == B2R_SUM) for (uint64_t i = 0; i < n; ++i) acc[i] = acc[i] + in[i];
  else if (op == B2R_MIN) for (uint64_t i = 0; i < n; ++i) acc[i] = min_nan(acc[i], in[i]);
  else for (uint64_t i = 0; i < n; ++i) acc[i] = max_nan(acc[i], in[i]);
}

}  // namespace
}  // namespace b2r

using namespace b2r;

extern "C" {

size_t b2r_dtype_size(b2r_dtype_t dtype) {
  switch (dtype) {
    case B2R_INT32: return 4;
    case B2R_FLOAT32: return 4;
    case B2R_FLOAT64: return 8;
    case B2R_BFLOAT16: return 2;
  }
  return 0;
}

const char* b2r_status_string(b2r_status_t s) {
  switch (s) {
    case B2R_OK: return "ok";
    case B2R_ERR_INVALID: return "invalid argument";
    case B2R_ERR_CUDA: return "CUDA error";
    case B2R_ERR_UNSUPPORTED: return "unsupported on this machine";
    case B2R_ERR_TIMEOUT: return "cross-GPU wait timed out";
    case B2R_ERR_NOMEM: return "out of symmetric heap / workspace";
    case B2R_ERR_SYSTEM: return "system call failed";
    case B2R_ERR_PENDING: return "recorded; waiting for the other loopback ranks";
  }
  return "unknown status";
}

const char* b2r_last_error_string(void) { return last_error_buf(); }

const char* b2r_dtype_name(b2r_dtype_t d) {
  switch (d) {
    case B2R_INT32: return "int";
    case B2R_FLOAT32: return "float";
    case B2R_FLOAT64: return "double";
    case B2R_BFLOAT16: return "bf16";
  }
  return "?";
}
const char* b2r_op_name(b2r_op_t o) {
  switch (o) {
    case B2R_SUM: return "SUM";
    case B2R_MIN: return "MIN";
    case B2R_MAX: return "MAX";
  }
  return "?";
}
const char* b2r_algo_name(b2r_algo_t a) {
  switch (a) {
    case B2R_ALGO_AUTO: return "auto";
    case B2R_ALGO_ONESHOT: return "oneshot";
    case B2R_ALGO_TWOSHOT: return "twoshot";
    case B2R_ALGO_TREE: return "tree";
    case B2R_ALGO_NVLS: return "nvls";
    case B2R_ALGO_PUSH: return "push";
    case B2R_ALGO_RING: return "ring";
    case B2R_ALGO_LL: return "ll";
    case B2R_ALGO_POSTED: return "posted";
  }
  return "?";
}
int b2r_version(void) { return B2R_VERSION_MAJOR * 100 + B2R_VERSION_MINOR; }

b2r_status_t b2r_reduce_host(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op) {
  B2R_REQUIRE(out != nullptr, "b2r_reduce_host: out is NULL");
  B2R_REQUIRE(in != nullptr || count == 0, "b2r_reduce_host: in is NULL with count=%llu", (unsigned long long)count);
  B2R_REQUIRE(op == B2R_SUM || op == B2R_MIN || op == B2R_MAX, "b2r_reduce_host: bad op %d", (int)op);
  switch (dtype) {
    case B2R_INT32: reduce_i32((const int32_t*)in, (int32_t*)out, count, op); return B2R_OK;
    case B2R_FLOAT32: reduce_fp((const float*)in, (float*)out, count, op); return B2R_OK;
    case B2R_FLOAT64: reduce_fp((const double*)in, (double*)out, count, op); return B2R_OK;
    case B2R_BFLOAT16: reduce_bf16((const uint16_t*)in, (uint16_t*)out, count, op); return B2R_OK;
  }
  return fail(B2R_ERR_INVALID, "b2r_reduce_host: bad dtype %d", (int)dtype);
}

b2r_status_t b2r_combine_host(void* acc, const void* in, uint64_t count, b2r_dtype_t dtype, b2r_op_t op) {
  B2R_REQUIRE((acc && in) || count == 0, "b2r_combine_host: NULL buffer");
  B2R_REQUIRE(op == B2R_SUM || op == B2R_MIN || op == B2R_MAX, "b2r_combine_host: bad op %d", (int)op);
  switch (dtype) {
    case B2R_INT32: {
      int32_t* a = (int32_t*)acc;
      const int32_t* b = (const int32_t*)in;
      if (op == B2R_SUM) for (uint64_t i = 0; i < count; ++i) a[i] = (int32_t)((uint32_t)a[i] + (uint32_t)b[i]);
      else if (op == B2R_MIN) for (uint64_t i = 0; i < count; ++i) a[i] = b[i] < a[i] ? b[i] : a[i];
      else for (uint64_t i = 0; i < count; ++i) a[i] = b[i] > a[i] ? b[i] : a[i];
      return B2R_OK;
    }
    case B2R_FLOAT32: combine_fp((float*)acc, (const float*)in, count, op); return B2R_OK;
    case B2R_FLOAT64: combine_fp((double*)acc, (const double*)in, count, op); return B2R_OK;
    case B2R_BFLOAT16: {
      // element-wise bf16: combine in fp32, round back each step (matches the device collectives, which keep
      // bf16 on the wire and combine all P contributions in fp32 before ONE rounding — see coll kernels; the
      // oracle for those uses b2r_combine_host on fp32 mirrors instead).
      uint16_t* a = (uint16_t*)acc;
      const uint16_t* b = (const uint16_t*)in;
      for (uint64_t i = 0; i < count; ++i) {
        float x = bf16_bits_to_f32(a[i]), y = bf16_bits_to_f32(b[i]);
        float r = op == B2R_SUM ? x + y : (op == B2R_MIN ? min_nan(x, y) : max_nan(x, y));
        a[i] = f32_to_bf16_bits(r);
      }
      return B2R_OK;
    }
  }
  return fail(B2R_ERR_INVALID, "b2r_combine_host: bad dtype %d", (int)dtype);
}

b2r_status_t b2r_fill_host(void* dst, uint64_t count, b2r_dtype_t dtype, uint64_t seed, uint64_t first, int mode) {
  B2R_REQUIRE(dst != nullptr || count == 0, "b2r_fill_host: dst is NULL");
  B2R_REQUIRE(mode >= 0 && mode <= 2, "b2r_fill_host: bad mode %d", mode);
  switch (dtype) {
    case B2R_INT32: for (uint64_t i = 0; i < count; ++i) ((int32_t*)dst)[i] = gen_i32(seed, first + i, mode); return B2R_OK;
    case B2R_FLOAT32: for (uint64_t i = 0; i < count; ++i) ((float*)dst)[i] = gen_f32(seed, first + i, mode); return B2R_OK;
    case B2R_FLOAT64: for (uint64_t i = 0; i < count; ++i) ((double*)dst)[i] = gen_f64(seed, first + i, mode); return B2R_OK;
    case B2R_BFLOAT16: for (uint64_t i = 0; i < count; ++i) ((uint16_t*)dst)[i] = gen_bf16(seed, first + i, mode); return B2R_OK;
  }
  return fail(B2R_ERR_INVALID, "b2r_fill_host: bad dtype %d", (int)dtype);
}

}  // extern "C"
