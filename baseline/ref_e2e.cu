// ref_e2e.cu — end-to-end timing of the UNMODIFIED reference reduction, step by step.
//
// The reference's stock binary copies its input to the GPU once, outside its timer (reduction.cpp:721-726), so it prints no
// end-to-end figure. This driver links the reference's own, untouched reduction_kernel.cu (baseline/_ref/src, a byte-for-
// byte copy) through its public API (reduction.h:15-25) and repeats PER STEP exactly the sequence reduction.cpp runs once:
//     cudaMemcpy H2D of the whole input            (reduction.cpp:725)
//     pass 1: sumreduce<T>(n, threads, blocks, 6)  (reduction.cpp:323, geometry from getNumBlocksAndThreads :272-291)
//     pass 2: in-place sumreduce over the block partials until one value is left (reduction.cpp:344-357)
//     cudaMemcpy D2H of the scalar                 (reduction.cpp:377-381)
// None of this repository's kernels or library code is involved. The geometry helper below restates the ten lines of
// getNumBlocksAndThreads (it is `static`-free but lives in reduction.cpp next to main(), which cannot be linked twice).
//
//   ref_e2e --type=int|float|double --method=SUM|MIN|MAX --n=16777216 --steps=20 --warmup=3 [--device=0] [--pageable]
// Host memory: pinned by default (the same rule as this repository's own end-to-end arm); --pageable uses malloc like the
// stock sample does (reduction.cpp:696).
// Prints one line:  REF_E2E gbs=<GB/s> ms_per_step=<ms> h2d_bytes=<n> d2h_bytes=<n> result_ok=<0|1> host=<pinned|pageable>
#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "reduction.h"

#define CK(x)                                                                                  \
  do {                                                                                         \
    cudaError_t e_ = (x);                                                                      \
    if (e_ != cudaSuccess) {                                                                   \
      fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_));       \
      exit(2);                                                                                 \
    }                                                                                          \
  } while (0)

// The kernel file's wrappers call this helper, which the reference defines in reduction.cpp next to main() (reduction.cpp:74-78);
// a driver that does not link reduction.cpp has to provide it (one line, restated).
extern "C" bool isPow2(unsigned int x) { return (x & (x - 1)) == 0; }

static unsigned next_pow2(unsigned x) {
  --x;
  x |= x >> 1; x |= x >> 2; x |= x >> 4; x |= x >> 8; x |= x >> 16;
  return ++x;
}
// kernel 6 geometry: two elements per thread on load, grid capped at maxBlocks
static void geometry(int n, int max_blocks, int max_threads, int* blocks, int* threads) {
  *threads = (n < max_threads * 2) ? (int)next_pow2((unsigned)(n + 1) / 2) : max_threads;
  *blocks = (n + (*threads * 2 - 1)) / (*threads * 2);
  if (*blocks > max_blocks) *blocks = max_blocks;
}

enum Method { M_SUM, M_MIN, M_MAX };

template <class T>
static void launch(Method m, int n, int threads, int blocks, T* in, T* out) {
  if (m == M_SUM) sumreduce<T>(n, threads, blocks, 6, in, out);
  else if (m == M_MIN) minreduce<T>(n, threads, blocks, 6, in, out);
  else maxreduce<T>(n, threads, blocks, 6, in, out);
}

template <class T>
static int run(Method m, int n, int steps, int warmup, bool pageable, bool is_int) {
  const size_t bytes = (size_t)n * sizeof(T);
  const int kMaxThreads = 256, kMaxBlocks = 64;  // the sample's defaults (reduction.cpp:666-668)
  T* h_in[2];
  for (int c = 0; c < 2; ++c) {
    if (pageable) h_in[c] = (T*)malloc(bytes);
    else CK(cudaMallocHost((void**)&h_in[c], bytes));
    for (int i = 0; i < n; ++i) h_in[c][i] = is_int ? (T)(rand() & 0xFF) : (T)((rand() & 0xFF) / (double)RAND_MAX);  // reduction.cpp:698-705
  }
  int blocks = 0, threads = 0;
  geometry(n, kMaxBlocks, kMaxThreads, &blocks, &threads);
  T *d_in = nullptr, *d_out = nullptr;
  CK(cudaMalloc((void**)&d_in, bytes));
  CK(cudaMalloc((void**)&d_out, blocks * sizeof(T)));
  T result = 0;
  auto step = [&](int i) {
    CK(cudaMemcpy(d_in, h_in[i & 1], bytes, cudaMemcpyHostToDevice));
    launch<T>(m, n, threads, blocks, d_in, d_out);
    int s = blocks;
    while (s > 1) {
      int t2 = 0, b2 = 0;
      geometry(s, kMaxBlocks, kMaxThreads, &b2, &t2);
      launch<T>(m, s, t2, b2, d_out, d_out);
      s = (s + (t2 * 2 - 1)) / (t2 * 2);
    }
    CK(cudaMemcpy(&result, d_out, sizeof(T), cudaMemcpyDeviceToHost));
  };
  for (int i = 0; i < warmup; ++i) step(i);
  CK(cudaDeviceSynchronize());
  const auto t0 = std::chrono::steady_clock::now();
  for (int i = 0; i < steps; ++i) step(i);
  CK(cudaDeviceSynchronize());
  const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  // check the last step against a plain host loop
  const T* h = h_in[(steps - 1) & 1];
  double want = m == M_SUM ? 0.0 : (double)h[0];
  for (int i = 0; i < n; ++i) want = m == M_SUM ? want + (double)h[i] : (m == M_MIN ? (h[i] < want ? h[i] : want) : (h[i] > want ? h[i] : want));
  const double tol = is_int ? 0.0 : (sizeof(T) == 4 ? 1e-8 * n : 1e-9);
  const bool ok = (double)result >= want - tol && (double)result <= want + tol;
  printf("REF_E2E gbs=%.4f ms_per_step=%.5f h2d_bytes=%zu d2h_bytes=%zu result_ok=%d host=%s steps=%d\n", 1e-9 * (double)bytes * steps / sec,
         sec * 1e3 / steps, bytes, sizeof(T), ok ? 1 : 0, pageable ? "pageable" : "pinned", steps);
  return ok ? 0 : 1;
}

int main(int argc, char** argv) {
  std::string type = "int", method = "SUM";
  int n = 1 << 24, steps = 20, warmup = 3, device = 0;
  bool pageable = false;
  for (int i = 1; i < argc; ++i) {
    const char* a = argv[i];
    if (!strncmp(a, "--type=", 7)) type = a + 7;
    else if (!strncmp(a, "--method=", 9)) method = a + 9;
    else if (!strncmp(a, "--n=", 4)) n = atoi(a + 4);
    else if (!strncmp(a, "--steps=", 8)) steps = atoi(a + 8);
    else if (!strncmp(a, "--warmup=", 9)) warmup = atoi(a + 9);
    else if (!strncmp(a, "--device=", 9)) device = atoi(a + 9);
    else if (!strcmp(a, "--pageable")) pageable = true;
  }
  CK(cudaSetDevice(device));
  const Method m = method == "MIN" ? M_MIN : (method == "MAX" ? M_MAX : M_SUM);
  if (type == "float") return run<float>(m, n, steps, warmup, pageable, false);
  if (type == "double") return run<double>(m, n, steps, warmup, pageable, false);
  return run<int>(m, n, steps, warmup, pageable, true);
}
