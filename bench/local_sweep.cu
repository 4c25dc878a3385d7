// local_sweep.cu — geometry sweep of the single-GPU kernel family on this machine (tuning tool, not a product
// path). Measures every (variant, threads, ctas/SM, unroll, stages, cluster) combination with CUDA events over
// rotating inputs larger than L2 and prints GB/s; the winners become tuned_defaults() in src/core/reduce_local.cu.
// The reference's own sweep (`--shmoo`, reduction.cpp:576-657) is advertised but aborts; this is its working heir.
//
//   local_sweep [--type=int|float|double|bf16] [--method=SUM|MIN|MAX] [--n=16777216] [--iters=50] [--csv=path] [--quick]
#include <algorithm>
#include <string>

#include "bench_common.h"

using namespace b2rbench;

struct Row {
  int variant, threads, cps, unroll, stages, cluster;
  double gbs, us;
  int dyn_pct = 0, tile_rounds = 0;
};

int main(int argc, char** argv) {
  Flags f;
  if (!f.parse(argc, argv)) { fprintf(stderr, "%s\n", f.error().c_str()); return 1; }
  b2r_dtype_t dt = B2R_INT32;
  b2r_op_t op = B2R_SUM;
  if (!parse_dtype(f.str("type", "int"), &dt) || !parse_op(f.str("method", "SUM"), &op)) {
    fprintf(stderr, "bad --type/--method\n");
    return 1;
  }
  const uint64_t n = (uint64_t)f.i64("n", 1 << 24);
  const int iters = (int)f.i64("iters", 50);
  const bool quick = f.has("quick");
  const int pdl = f.has("pdl") ? (int)f.i64("pdl", B2R_PDL_INPUT_READY) : 0;
  const size_t bytes = n * b2r_dtype_size(dt);
  const int copies = rotating_copies(bytes, (size_t)1 << 30);

  BENCH_CUDA(cudaSetDevice((int)f.i64("device", 0)));
  cudaDeviceProp prop;
  BENCH_CUDA(cudaGetDeviceProperties(&prop, 0));
  printf("# device %s, %d SMs; pdl=%d; %s %s n=%llu (%.1f MiB) x %d rotating copies, %d iters\n", prop.name, prop.multiProcessorCount, pdl,
         b2r_dtype_name(dt), b2r_op_name(op), (unsigned long long)n, bytes / 1048576.0, copies, iters);

  char* d_in = nullptr;
  void* d_out = nullptr;
  BENCH_CUDA(cudaMalloc(&d_in, bytes * copies + 256));
  BENCH_CUDA(cudaMalloc(&d_out, 64));
  for (int c = 0; c < copies; ++c) BENCH_B2R(b2r_fill_device(d_in + c * bytes, n, dt, 1234 + c, 0, 0, nullptr));
  b2r_workspace_t ws;
  BENCH_B2R(b2r_workspace_create(&ws, 0));
  BENCH_CUDA(cudaDeviceSynchronize());

  std::vector<Row> rows;
  EventTimer et;
  int dyn_pct = -1, tile_rounds = 1;  // the classic grid below measures the static split
  auto run = [&](int variant, int threads, int cps, int unroll, int stages, int cluster) {
    b2r_launch_cfg_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.dyn_pct = dyn_pct;
    cfg.variant = variant;
    cfg.threads = threads;
    cfg.ctas = cps * prop.multiProcessorCount;
    cfg.unroll = unroll;
    cfg.stages = stages;
    cfg.cluster = cluster;
    cfg.pdl = pdl;
    for (int i = 0; i < 3; ++i) {
      if (b2r_reduce_local_ex(d_in + (i % copies) * bytes, d_out, n, dt, op, ws, nullptr, &cfg) != B2R_OK) return;
    }
    if (cudaDeviceSynchronize() != cudaSuccess) { fprintf(stderr, "launch failed: %s\n", cudaGetErrorString(cudaGetLastError())); return; }
    et.start(nullptr);
    for (int i = 0; i < iters; ++i) b2r_reduce_local_ex(d_in + (i % copies) * bytes, d_out, n, dt, op, ws, nullptr, &cfg);
    et.stop(nullptr);
    const double ms = et.ms() / iters;
    Row r{variant, threads, cps, unroll, stages, cluster, bytes / ms * 1e-6, ms * 1e3};
    r.dyn_pct = dyn_pct;
    r.tile_rounds = tile_rounds;
    rows.push_back(r);
  };

  if (f.has("dyn")) {
    // --dyn: the dynamically scheduled tail (share of the input x tile size) against the static split, plain stream order,
    // for the register-streaming geometries that matter at this size
    for (int pct : {-1, 10, 25, 40, 60, 90})
      for (int rounds : {1, 2, 4}) {
        if (pct < 0 && rounds > 1) continue;
        dyn_pct = pct;
        tile_rounds = rounds;
        setenv("B2R_DYN_TILE_ROUNDS", std::to_string(rounds).c_str(), 1);
        b2r_refresh_env();
        for (int threads : {256, 512})
          for (int cps : {2, 4, 8}) {
            if (threads * cps > 2048 || threads * cps < 1024) continue;
            for (int unroll : {4, 8}) run(B2R_VARIANT_LDG128, threads, cps, unroll, 0, 1);
          }
        run(B2R_VARIANT_LDG256, 512, 2, 4, 0, 1);
        run(B2R_VARIANT_LDG256, 256, 4, 4, 0, 1);
      }
    std::sort(rows.begin(), rows.end(), [](const Row& a, const Row& b) { return a.gbs > b.gbs; });
    printf("%-8s %7s %6s %6s %8s %6s %10s %9s\n", "variant", "threads", "cta/SM", "unroll", "dyn_pct", "rounds", "GB/s", "us");
    for (size_t i = 0; i < rows.size(); ++i) {
      const Row& r = rows[i];
      if (i < 30 || r.dyn_pct < 0)
        printf("%-8s %7d %6d %6d %8d %6d %10.1f %9.2f\n", variant_name(r.variant), r.threads, r.cps, r.unroll, r.dyn_pct, r.tile_rounds, r.gbs, r.us);
    }
    b2r_workspace_destroy(ws);
    cudaFree(d_in);
    cudaFree(d_out);
    return 0;
  }

  const int thr_ldg[] = {128, 256, 512};
  const int cps_ldg[] = {1, 2, 4, 8, 16};
  for (int variant : {B2R_VARIANT_LDG128, B2R_VARIANT_LDG256})
    for (int threads : thr_ldg)
      for (int cps : cps_ldg) {
        if (threads * cps > 2048) continue;
        if (quick && threads * cps < 512) continue;
        for (int unroll : {4, 8}) {
          if (variant == B2R_VARIANT_LDG256 && unroll == 8) continue;
          run(variant, threads, cps, unroll, 0, 1);
        }
      }
  const int thr_tma[] = {96, 160, 288, 416};
  for (int threads : thr_tma)
    for (int cps : {1, 2, 3, 4})
      for (int stages : {2, 4, 8}) {
        const int tile = (threads - 32) * 64;
        if ((size_t)cps * (stages * tile + 16 * stages + 1024) > 220 * 1024) continue;
        if (threads * cps > 2048) continue;
        run(B2R_VARIANT_TMA, threads, cps, 4, stages, 1);
      }
  // cluster DSMEM combine on top of the best few
  std::sort(rows.begin(), rows.end(), [](const Row& a, const Row& b) { return a.gbs > b.gbs; });
  std::vector<Row> best(rows.begin(), rows.begin() + std::min<size_t>(3, rows.size()));
  for (const Row& r : best)
    for (int cluster : {2, 4, 8}) run(r.variant, r.threads, r.cps, r.unroll, r.stages, cluster);
  std::sort(rows.begin(), rows.end(), [](const Row& a, const Row& b) { return a.gbs > b.gbs; });

  FILE* csv = f.has("csv") ? fopen(f.str("csv").c_str(), "a") : nullptr;
  printf("%-8s %7s %6s %6s %6s %7s %10s %9s\n", "variant", "threads", "cta/SM", "unroll", "stages", "cluster", "GB/s", "us");
  for (size_t i = 0; i < rows.size(); ++i) {
    const Row& r = rows[i];
    if (i < 25 || i + 5 >= rows.size())
      printf("%-8s %7d %6d %6d %6d %7d %10.1f %9.2f\n", variant_name(r.variant), r.threads, r.cps, r.unroll, r.stages, r.cluster, r.gbs, r.us);
    if (csv)
      fprintf(csv, "%s,%s,%llu,%s,%d,%d,%d,%d,%d,%.1f,%.2f\n", b2r_dtype_name(dt), b2r_op_name(op), (unsigned long long)n,
              variant_name(r.variant), r.threads, r.cps, r.unroll, r.stages, r.cluster, r.gbs, r.us);
  }
  if (csv) fclose(csv);
  b2r_workspace_destroy(ws);
  cudaFree(d_in);
  cudaFree(d_out);
  return 0;
}
