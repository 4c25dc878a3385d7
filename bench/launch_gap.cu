// launch_gap.cu — what a kernel boundary costs on this machine (diagnostic for profiles/single_gpu.md, not a product path).
// A 64 MiB reduction is 9.1 us of HBM time but 16.4 us per launch in plain stream order; this tool measures the fixed part:
//   empty kernels back to back (1 CTA and 592 CTAs x 512 threads), in plain stream order and with programmatic dependent
//   launch (griddepcontrol.launch_dependents at entry, griddepcontrol.wait before the first instruction that matters).
#include <cuda_runtime.h>

#include <cstdio>

#define CK(x)                                                                                \
  do {                                                                                       \
    cudaError_t e_ = (x);                                                                    \
    if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 2; } \
  } while (0)

__global__ void __launch_bounds__(512) empty_kernel(int* sink, int pdl) {
  if (pdl) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
  }
  if (sink && threadIdx.x == 0 && blockIdx.x == 0xffffffffu) *sink = 1;
}

int main() {
  CK(cudaSetDevice(0));
  cudaStream_t s;
  CK(cudaStreamCreate(&s));
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  const int iters = 2000;
  for (int pdl = 0; pdl <= 1; ++pdl)
    for (int shape = 0; shape < 4; ++shape) {
      const int ctas = shape == 0 ? 1 : (shape == 1 ? 148 : (shape == 2 ? 296 : 592));
      const int threads = shape == 0 ? 32 : 512;
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(ctas);
      cfg.blockDim = dim3(threads);
      cfg.stream = s;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = pdl ? attr : nullptr;
      cfg.numAttrs = pdl ? 1 : 0;
      for (int i = 0; i < 20; ++i) CK(cudaLaunchKernelEx(&cfg, empty_kernel, (int*)nullptr, pdl));
      CK(cudaStreamSynchronize(s));
      CK(cudaEventRecord(a, s));
      for (int i = 0; i < iters; ++i) CK(cudaLaunchKernelEx(&cfg, empty_kernel, (int*)nullptr, pdl));
      CK(cudaEventRecord(b, s));
      CK(cudaEventSynchronize(b));
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, a, b));
      printf("empty kernel %4d CTAs x %3d threads, %s: %.2f us per launch\n", ctas, threads, pdl ? "programmatic launch" : "plain stream order   ", ms * 1e3 / iters);
    }
  return 0;
}
