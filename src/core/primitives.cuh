// primitives.cuh — sm_100a device building blocks of the reduction family.
//
// What the reference does with scalar loads, a shared-memory tree and a `volatile` warp-synchronous tail
// (reduction_kernel.cu:91-126) is done here with: 128/256-bit streaming loads or cp.async.bulk (TMA) tiles,
// register accumulation, redux.sync / shfl.sync warp reduction, one shared-memory hop across warps,
// a DSMEM hop across the CTAs of a thread-block cluster, and a release/acquire ticket across clusters.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdint>

#include "b2r.h"

namespace b2r {

// ------------------------------------------------------------------------------------------------ loads
struct __align__(16) V16 { uint32_t w[4]; };
struct __align__(32) V32 { uint32_t w[8]; };

// Streaming read-only loads: no L1 allocation (each byte is touched once). `volatile` keeps the unrolled batch of
// loads together in program order (a plain asm gets sunk to its first use, which serialises load -> add).
__device__ __forceinline__ V16 ldg_stream(const V16* p) {
  V16 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];"
      : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3])
      : "l"(p));
  return v;
}
// 256-bit global load (PTX 8.8, sm_100+): SASS LDG.E.256.
__device__ __forceinline__ V32 ldg_stream(const V32* p) {
  V32 v;
  asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3]), "=r"(v.w[4]), "=r"(v.w[5]), "=r"(v.w[6]), "=r"(v.w[7])
      : "l"(p));
  return v;
}
__device__ __forceinline__ V16 lds128(uint32_t smem_addr) {
  V16 v;
  asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3])
               : "r"(smem_addr));
  return v;
}

__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ------------------------------------------------------------------------------------------------ mbarrier + bulk copy (TMA engine, 1-D)
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  // labels inside { } are scoped in PTX, so this can be inlined any number of times
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "B2R_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra B2R_DONE;\n\t"
      "bra B2R_WAIT;\n\t"
      "B2R_DONE:\n\t"
      "}" ::"r"(bar), "r"(parity)
      : "memory");
}
// global -> shared bulk async copy; completion is signalled as transaction bytes on `bar`. SASS: UBLKCP.
__device__ __forceinline__ void bulk_g2s(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_dst),
               "l"(gsrc), "r"(bytes), "r"(bar)
               : "memory");
}

// ------------------------------------------------------------------------------------------------ scalar op functors
template <int O>
struct OpI32 {
  __device__ __forceinline__ static int32_t identity() {
    return O == B2R_SUM ? 0 : (O == B2R_MIN ? 2147483647 : (-2147483647 - 1));
  }
  __device__ __forceinline__ static int32_t apply(int32_t a, int32_t b) {
    if (O == B2R_SUM) return (int32_t)((uint32_t)a + (uint32_t)b);  // two's-complement wrap, no UB
    if (O == B2R_MIN) return min(a, b);
    return max(a, b);
  }
  __device__ __forceinline__ static int32_t warp(int32_t v) {  // REDUX.{SUM,MIN,MAX}.S32
    int32_t r;
    if (O == B2R_SUM) asm volatile("redux.sync.add.s32 %0, %1, 0xffffffff;" : "=r"(r) : "r"(v));
    else if (O == B2R_MIN) asm volatile("redux.sync.min.s32 %0, %1, 0xffffffff;" : "=r"(r) : "r"(v));
    else asm volatile("redux.sync.max.s32 %0, %1, 0xffffffff;" : "=r"(r) : "r"(v));
    return r;
  }
};

template <int O>
struct OpF32 {
  __device__ __forceinline__ static float identity() {
    return O == B2R_SUM ? 0.0f : (O == B2R_MIN ? __int_as_float(0x7f800000) : __int_as_float(0xff800000));
  }
  __device__ __forceinline__ static float apply(float a, float b) {
    float r;
    if (O == B2R_SUM) return a + b;
    if (O == B2R_MIN) asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));  // FMNMX.NAN: NaN propagates
    else asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
  }
  __device__ __forceinline__ static float warp(float v) {
    if (O == B2R_SUM) {
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
      return v;
    }
    float r;  // sm_100a: CREDUX.{MIN,MAX}.F32.NAN — one instruction for the whole warp
    if (O == B2R_MIN) asm volatile("redux.sync.min.NaN.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
    else asm volatile("redux.sync.max.NaN.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
    return r;
  }
};

template <int O>
struct OpF64 {
  __device__ __forceinline__ static double identity() {
    return O == B2R_SUM ? 0.0 : (O == B2R_MIN ? __longlong_as_double(0x7ff0000000000000ll)
                                              : __longlong_as_double((long long)0xfff0000000000000ull));
  }
  __device__ __forceinline__ static double apply(double a, double b) {
    if (O == B2R_SUM) return a + b;
    if (O == B2R_MIN) return (b < a || b != b) ? b : a;  // NaN in either operand propagates
    return (b > a || b != b) ? b : a;
  }
  __device__ __forceinline__ static double warp(double v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = apply(v, __shfl_xor_sync(0xffffffffu, v, d));
    return v;
  }
};

// ------------------------------------------------------------------------------------------------ per-(dtype,op) reducers
// acc_t   : type of cross-thread partials (int32 / float / double)
// Thread  : per-thread register accumulator fed with 16-byte vectors or single elements
template <int D, int O>
struct Reducer;

template <int O>
struct Reducer<B2R_INT32, O> {
  using acc_t = int32_t;
  using F = OpI32<O>;
  static constexpr int kElemBytes = 4;
  static constexpr bool kIsInt = true;
  static constexpr bool kPackedSum = (O == B2R_SUM);  // grid stage in one 64-bit atomic (finish_grid_packed)
  struct Thread {
    int32_t a[4];
    __device__ __forceinline__ void init() {
#pragma unroll
      for (int j = 0; j < 4; ++j) a[j] = F::identity();
    }
    __device__ __forceinline__ void add16(const uint32_t* w) {
#pragma unroll
      for (int j = 0; j < 4; ++j) a[j] = F::apply(a[j], (int32_t)w[j]);
    }
    __device__ __forceinline__ void add_elem(const void* base, uint64_t i) {
      a[0] = F::apply(a[0], ((const int32_t*)base)[i]);
    }
    __device__ __forceinline__ acc_t collapse() const { return F::apply(F::apply(a[0], a[1]), F::apply(a[2], a[3])); }
  };
  __device__ __forceinline__ static void store(void* out, acc_t v) { *(int32_t*)out = v; }
};

template <int O>
struct Reducer<B2R_FLOAT32, O> {
  using acc_t = float;
  using F = OpF32<O>;
  static constexpr int kElemBytes = 4;
  static constexpr bool kIsInt = false;
  static constexpr bool kPackedSum = false;
  struct Thread {
    float a[4];
    __device__ __forceinline__ void init() {
#pragma unroll
      for (int j = 0; j < 4; ++j) a[j] = F::identity();
    }
    __device__ __forceinline__ void add16(const uint32_t* w) {
#pragma unroll
      for (int j = 0; j < 4; ++j) a[j] = F::apply(a[j], __uint_as_float(w[j]));
    }
    __device__ __forceinline__ void add_elem(const void* base, uint64_t i) {
      a[0] = F::apply(a[0], ((const float*)base)[i]);
    }
    __device__ __forceinline__ acc_t collapse() const { return F::apply(F::apply(a[0], a[1]), F::apply(a[2], a[3])); }
  };
  __device__ __forceinline__ static void store(void* out, acc_t v) { *(float*)out = v; }
};

template <int O>
struct Reducer<B2R_FLOAT64, O> {
  using acc_t = double;
  using F = OpF64<O>;
  static constexpr int kElemBytes = 8;
  static constexpr bool kIsInt = false;
  static constexpr bool kPackedSum = false;
  struct Thread {
    double a[2];
    __device__ __forceinline__ void init() { a[0] = a[1] = F::identity(); }
    __device__ __forceinline__ void add16(const uint32_t* w) {
      a[0] = F::apply(a[0], __hiloint2double((int)w[1], (int)w[0]));
      a[1] = F::apply(a[1], __hiloint2double((int)w[3], (int)w[2]));
    }
    __device__ __forceinline__ void add_elem(const void* base, uint64_t i) {
      a[0] = F::apply(a[0], ((const double*)base)[i]);
    }
    __device__ __forceinline__ acc_t collapse() const { return F::apply(a[0], a[1]); }
  };
  __device__ __forceinline__ static void store(void* out, acc_t v) { *(double*)out = v; }
};

// bf16 SUM: widen each half-word to fp32 (a shift / a mask — bf16 is the top half of an fp32), accumulate in
// fp32, round ONCE when the final scalar is stored.
template <>
struct Reducer<B2R_BFLOAT16, B2R_SUM> {
  using acc_t = float;
  using F = OpF32<B2R_SUM>;
  static constexpr int kElemBytes = 2;
  static constexpr bool kIsInt = false;
  static constexpr bool kPackedSum = false;
  struct Thread {
    float a[8];
    __device__ __forceinline__ void init() {
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] = 0.0f;
    }
    __device__ __forceinline__ void add16(const uint32_t* w) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        a[2 * j] += __uint_as_float(w[j] << 16);
        a[2 * j + 1] += __uint_as_float(w[j] & 0xffff0000u);
      }
    }
    __device__ __forceinline__ void add_elem(const void* base, uint64_t i) {
      a[0] += __uint_as_float(((uint32_t)((const uint16_t*)base)[i]) << 16);
    }
    __device__ __forceinline__ acc_t collapse() const {
      return ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
    }
  };
  __device__ __forceinline__ static void store(void* out, acc_t v) { *(__nv_bfloat16*)out = __float2bfloat16_rn(v); }
};

// bf16 MIN/MAX: stay packed (HMNMX2.BF16.NAN handles two elements per instruction); widen only at the end.
template <int O>
struct Reducer<B2R_BFLOAT16, O> {
  using acc_t = float;
  using F = OpF32<O>;
  static constexpr int kElemBytes = 2;
  static constexpr bool kIsInt = false;
  static constexpr bool kPackedSum = false;
  __device__ __forceinline__ static uint32_t packed_identity() { return O == B2R_MIN ? 0x7f807f80u : 0xff80ff80u; }
  __device__ __forceinline__ static uint32_t apply2(uint32_t a, uint32_t b) {
    uint32_t r;
    if (O == B2R_MIN) asm("min.NaN.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    else asm("max.NaN.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
  }
  struct Thread {
    uint32_t a[4];
    __device__ __forceinline__ void init() {
#pragma unroll
      for (int j = 0; j < 4; ++j) a[j] = packed_identity();
    }
    __device__ __forceinline__ void add16(const uint32_t* w) {
#pragma unroll
      for (int j = 0; j < 4; ++j) a[j] = apply2(a[j], w[j]);
    }
    __device__ __forceinline__ void add_elem(const void* base, uint64_t i) {
      uint32_t e = ((const uint16_t*)base)[i];
      a[0] = apply2(a[0], e | (e << 16));
    }
    __device__ __forceinline__ acc_t collapse() const {
      uint32_t m = apply2(apply2(a[0], a[1]), apply2(a[2], a[3]));
      return F::apply(__uint_as_float(m << 16), __uint_as_float(m & 0xffff0000u));
    }
  };
  __device__ __forceinline__ static void store(void* out, acc_t v) { *(__nv_bfloat16*)out = __float2bfloat16_rn(v); }
};

// ------------------------------------------------------------------------------------------------ block reduce
// Result is valid in every lane of warp 0. `scratch` holds 32 acc_t. Fixed combination order -> deterministic.
template <class R>
__device__ __forceinline__ typename R::acc_t block_reduce(typename R::acc_t v, typename R::acc_t* scratch) {
  using F = typename R::F;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
  v = F::warp(v);
  if (nwarps == 1) return v;
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  if (warp == 0) {
    v = lane < nwarps ? scratch[lane] : F::identity();
    v = F::warp(v);
  }
  __syncthreads();  // scratch may be reused by the caller
  return v;
}

__device__ __forceinline__ uint32_t atom_add_acq_rel_gpu(unsigned* p, unsigned v) {
  uint32_t old;
  asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}

template <typename T>
__device__ __forceinline__ T ld_cg(const T* p) {
  return __ldcg(p);
}

}  // namespace b2r
