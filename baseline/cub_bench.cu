// cub_bench.cu — BASELINE (not product): cub::DeviceReduce {Sum,Min,Max} x {int32, fp32, fp64, bf16} over a size
// sweep, timed exactly like bench_reduce (CUDA events, rotating inputs > L2). BASELINE.md §5, SURVEY §7.1.
//   cub_bench [--type=int|float|double|bf16] [--method=SUM|MIN|MAX] [--min=1024] [--max=268435456] [--iters=50] [--csv=path]
#include <cuda_bf16.h>

#include <cub/cub.cuh>

#include "../bench/bench_common.h"

using namespace b2rbench;

struct Bf16ToFloat {
  __host__ __device__ float operator()(const __nv_bfloat16& v) const { return __bfloat162float(v); }
};

template <typename InIt, typename OutT>
double time_cub(b2r_op_t op, InIt* ins, int copies, OutT* d_out, size_t n, int iters) {
  void* tmp = nullptr;
  size_t tmp_bytes = 0;
  auto call = [&](InIt in) {
    if (op == B2R_SUM) return cub::DeviceReduce::Sum(tmp, tmp_bytes, in, d_out, (int64_t)n);
    if (op == B2R_MIN) return cub::DeviceReduce::Min(tmp, tmp_bytes, in, d_out, (int64_t)n);
    return cub::DeviceReduce::Max(tmp, tmp_bytes, in, d_out, (int64_t)n);
  };
  BENCH_CUDA(call(ins[0]));
  BENCH_CUDA(cudaMalloc(&tmp, std::max<size_t>(tmp_bytes, 16)));
  for (int i = 0; i < 3; ++i) BENCH_CUDA(call(ins[i % copies]));
  BENCH_CUDA(cudaDeviceSynchronize());
  ClockLog::Window clocks_window;
  EventTimer et;
  et.start(nullptr);
  for (int i = 0; i < iters; ++i) call(ins[i % copies]);
  et.stop(nullptr);
  const double ms = et.ms() / iters;
  cudaFree(tmp);
  return ms;
}

int main(int argc, char** argv) {
  Flags f;
  if (!f.parse(argc, argv)) { fprintf(stderr, "%s\n", f.error().c_str()); return 1; }
  b2r_dtype_t dt = B2R_INT32;
  b2r_op_t op = B2R_SUM;
  if (!parse_dtype(f.str("type", "int"), &dt) || !parse_op(f.str("method", "SUM"), &op)) { fprintf(stderr, "bad --type/--method\n"); return 1; }
  const uint64_t nmin = (uint64_t)f.i64("min", 1024), nmax = (uint64_t)f.i64("max", 1ll << 28);
  const int iters = (int)f.i64("iters", 50);
  FILE* csv = f.has("csv") ? fopen(f.str("csv").c_str(), "a") : nullptr;
  BENCH_CUDA(cudaSetDevice((int)f.i64("device", 0)));
  ClockLog::get().start(1);
  printf("# cub::DeviceReduce %s %s\n# %12s %10s %10s\n", b2r_dtype_name(dt), b2r_op_name(op), "elements", "us", "GB/s");
  for (uint64_t n = nmin; n <= nmax; n *= 4) {
    const size_t bytes = n * b2r_dtype_size(dt);
    const int copies = rotating_copies(bytes);
    char* d_in = nullptr;
    void* d_out = nullptr;
    BENCH_CUDA(cudaMalloc(&d_in, bytes * copies));
    BENCH_CUDA(cudaMalloc(&d_out, 64));
    for (int c = 0; c < copies; ++c) BENCH_B2R(b2r_fill_device(d_in + c * bytes, n, dt, 99 + c, 0, 0, nullptr));
    double ms = 0;
    const int it = n > (1ull << 26) ? std::min(iters, 20) : iters;
    if (dt == B2R_INT32) {
      std::vector<const int*> ins(copies);
      for (int c = 0; c < copies; ++c) ins[c] = (const int*)(d_in + c * bytes);
      ms = time_cub(op, ins.data(), copies, (int*)d_out, n, it);
    } else if (dt == B2R_FLOAT32) {
      std::vector<const float*> ins(copies);
      for (int c = 0; c < copies; ++c) ins[c] = (const float*)(d_in + c * bytes);
      ms = time_cub(op, ins.data(), copies, (float*)d_out, n, it);
    } else if (dt == B2R_FLOAT64) {
      std::vector<const double*> ins(copies);
      for (int c = 0; c < copies; ++c) ins[c] = (const double*)(d_in + c * bytes);
      ms = time_cub(op, ins.data(), copies, (double*)d_out, n, it);
    } else {
      using It = cub::TransformInputIterator<float, Bf16ToFloat, const __nv_bfloat16*>;  // fp32 accumulate, like ours
      std::vector<It> ins;
      for (int c = 0; c < copies; ++c) ins.push_back(It((const __nv_bfloat16*)(d_in + c * bytes), Bf16ToFloat()));
      ms = time_cub(op, ins.data(), copies, (float*)d_out, n, it);
    }
    printf("  %12llu %10.2f %10.1f\n", (unsigned long long)n, ms * 1e3, bytes / ms * 1e-6);
    if (csv) fprintf(csv, "cub,%s,%s,%llu,%.3f,%.3f\n", b2r_dtype_name(dt), b2r_op_name(op), (unsigned long long)n, ms * 1e3, bytes / ms * 1e-6);
    cudaFree(d_in);
    cudaFree(d_out);
  }
  if (csv) fclose(csv);
  ClockLog::get().stop();
  return ClockLog::get().print(stdout) ? 0 : 1;
}
