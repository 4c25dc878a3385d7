// bench_reduce.cu — single-GPU reduction benchmark with the reference's command line and output format.
//
// Counterpart of the reference's driver (cuda/C/src/reduction/reduction.cpp: main :84-204, runTest* :661-1034,
// benchmarkReduce* :297-568, shmoo :576-657). Same flags (doc block :31-41 and parsing :91-94,124-134,672-684):
//   --method=SUM|MIN|MAX  (required; case-insensitive here, the reference's strcmp is case-sensitive)
//   --type=int|float|double|bf16   --n=<elements>   --threads=<per CTA>   --kernel=<0 auto,1 ldg128,2 ldg256,3 tma>
//   --maxblocks=<CTAs>   --cpufinal   --cputhresh=<N>   --shmoo   --device=<N>   --qatest   --noprompt   --prompt
// and new ones:
//   --iters=<100>  --warmup=<3>  --norotate (time one L2-resident buffer like the reference does)  --cluster=<1|2|4|8>
//   --pdl (overlap consecutive launches)  --unroll=<4|8>  --stages=<N>  --seed=<N>  --fill=sdk|full|signed  --csv=<path>  --jsonl=<path>  --nolog
// It prints the reference's lines ("Reduction, Throughput = ... GB/s, Time = ... s, Size = ... Elements, NumDevsUsed = 1,
// Workgroup = ...", GPU/CPU result, PASSED/FAILED) so scripts written against the sample keep working, followed by the
// device-event measurement over rotating inputs (the honest HBM number) and the fraction of the measured roofline.
//
// Reference defects deliberately NOT reproduced: the min/max out-of-bounds reads, `+=` host fall-back
// for min/max under --cpufinal, --kernel=0..5 launching nothing, --shmoo aborting, 32-bit sizes.
#include <cmath>
#include <cstdint>

#include "bench_common.h"

using namespace b2rbench;

namespace {

struct Options {
  b2r_dtype_t dtype = B2R_INT32;
  b2r_op_t op = B2R_SUM;
  uint64_t n = 1ull << 24;
  int iters = 100, warmup = 3;
  bool rotate = true, cpufinal = false;
  int cputhresh = 1;
  int fill = 0;
  uint64_t seed = 2012;
  b2r_launch_cfg_t cfg;
  double hbm_peak_gbs = 6542.1;  // MEASURED_PEAKS.json on this pool; override with --peak
};

struct Result {
  double host_ms;   // reference-style: host stopwatch around sync / launch / sync, averaged
  double event_ms;  // CUDA events around `iters` back-to-back launches over rotating inputs
  unsigned char gpu[8], cpu[8];
  bool ok;
  b2r_launch_cfg_t geom;
};

double to_double(const unsigned char* v, b2r_dtype_t d) {
  switch (d) {
    case B2R_INT32: { int32_t x; memcpy(&x, v, 4); return x; }
    case B2R_FLOAT32: { float x; memcpy(&x, v, 4); return x; }
    case B2R_FLOAT64: { double x; memcpy(&x, v, 8); return x; }
    case B2R_BFLOAT16: { uint16_t b; memcpy(&b, v, 2); uint32_t u = (uint32_t)b << 16; float x; memcpy(&x, &u, 4); return x; }
  }
  return 0;
}

bool validate(const Options& o, const unsigned char* gpu, const unsigned char* cpu) {
  const double g = to_double(gpu, o.dtype), c = to_double(cpu, o.dtype);
  if (o.dtype == B2R_INT32 || o.op != B2R_SUM) return (std::isnan(g) && std::isnan(c)) || g == c;
  // thresholds of the reference (reduction.cpp:750-780): float 1e-8*n, double 1e-12 — absolute, for its small inputs;
  // scaled by the magnitude of the sum for other fills.
  const double scale = std::max(1.0, std::fabs(c));
  if (o.dtype == B2R_FLOAT32) return std::fabs(g - c) < 1e-8 * (double)std::max<uint64_t>(o.n, 1) * scale;
  if (o.dtype == B2R_FLOAT64) return std::fabs(g - c) < 1e-12 * scale * std::max(1.0, std::log2((double)std::max<uint64_t>(o.n, 2)));
  return std::fabs(g - c) <= std::fabs(c) * 0.0078125 + 1e-30;  // bf16: one rounding at the end (2^-7 relative)
}

Result run_one(const Options& o, b2r_workspace_t ws, Log* log) {
  Result r;
  memset(&r, 0, sizeof(r));
  const size_t esz = b2r_dtype_size(o.dtype);
  const size_t bytes = o.n * esz;
  const int copies = o.rotate ? rotating_copies(bytes) : 1;

  char* d_in = nullptr;
  char* d_out = nullptr;
  BENCH_CUDA(cudaMalloc(&d_in, std::max<size_t>(bytes, 16) * copies));
  BENCH_CUDA(cudaMalloc(&d_out, 8 * 4096));
  // input 0 is generated on the host and copied (the reference's path, reduction.cpp:698-726) so the CPU oracle sees
  // the same bits; the other rotating copies are generated on the device.
  std::vector<unsigned char> h_in(std::max<size_t>(bytes, 16));
  BENCH_B2R(b2r_fill_host(h_in.data(), o.n, o.dtype, o.seed, 0, o.fill));
  BENCH_CUDA(cudaMemcpy(d_in, h_in.data(), bytes, cudaMemcpyHostToDevice));
  for (int c = 1; c < copies; ++c) BENCH_B2R(b2r_fill_device(d_in + (size_t)c * bytes, o.n, o.dtype, o.seed + c, 0, o.fill, nullptr));
  BENCH_B2R(b2r_plan_local(o.n, o.dtype, o.op, 0, &o.cfg, &r.geom));

  auto reduce_once = [&](const char* in) {
    if (!o.cpufinal && o.cputhresh <= 1) {
      BENCH_B2R(b2r_reduce_local_ex(in, d_out, o.n, o.dtype, o.op, ws, nullptr, &o.cfg));
      return;
    }
    // --cpufinal / --cputhresh: the GPU produces `parts` partial scalars, the host finishes with the SAME operator
    // (the reference's host fall-back adds even for min/max, reduction.cpp:426-429).
    uint64_t parts = o.cpufinal ? (uint64_t)std::max(1, r.geom.ctas) : (uint64_t)o.cputhresh;
    parts = std::min<uint64_t>(std::min<uint64_t>(parts, 4096), std::max<uint64_t>(o.n, 1));
    const uint64_t per = (o.n + parts - 1) / parts;
    std::vector<unsigned char> h_part(parts * 8);
    uint64_t used = 0;
    for (uint64_t p = 0; p < parts; ++p) {
      const uint64_t first = p * per;
      if (first >= o.n) break;
      const uint64_t cnt = std::min(per, o.n - first);
      BENCH_B2R(b2r_reduce_local(in + first * esz, d_out + used * esz, cnt, o.dtype, o.op, ws, nullptr));
      ++used;
    }
    BENCH_CUDA(cudaMemcpy(h_part.data(), d_out, used * esz, cudaMemcpyDeviceToHost));
    unsigned char fin[8];
    BENCH_B2R(b2r_reduce_host(h_part.data(), fin, used, o.dtype, o.op));
    BENCH_CUDA(cudaMemcpy(d_out, fin, esz, cudaMemcpyHostToDevice));
  };

  for (int i = 0; i < o.warmup; ++i) reduce_once(d_in + (size_t)(i % copies) * bytes);
  BENCH_CUDA(cudaDeviceSynchronize());

  ClockLog::Window clocks_window;  // clock samples from here on count as "during a timed region"
  // (a) the reference's measurement: host stopwatch, device sync on both sides of every iteration
  HostTimer ht;
  for (int i = 0; i < o.iters; ++i) {
    BENCH_CUDA(cudaDeviceSynchronize());
    ht.start();
    reduce_once(d_in + (size_t)(i % copies) * bytes);
    BENCH_CUDA(cudaDeviceSynchronize());
    ht.stop();
  }
  r.host_ms = ht.average_ms();

  // (b) device time of `iters` back-to-back launches
  EventTimer et;
  et.start(nullptr);
  for (int i = 0; i < o.iters; ++i) reduce_once(d_in + (size_t)(i % copies) * bytes);
  et.stop(nullptr);
  r.event_ms = et.ms() / o.iters;

  // result of input 0 against the oracle
  reduce_once(d_in);
  BENCH_CUDA(cudaMemcpy(r.gpu, d_out, esz, cudaMemcpyDeviceToHost));
  BENCH_B2R(b2r_reduce_host(h_in.data(), r.cpu, o.n, o.dtype, o.op));
  r.ok = validate(o, r.gpu, r.cpu);
  (void)log;
  cudaFree(d_in);
  cudaFree(d_out);
  return r;
}

void print_value(Log& log, const char* who, const unsigned char* v, b2r_dtype_t d) {
  if (d == B2R_INT32) log.print("%s result = %d\n", who, (int)to_double(v, d));
  else log.print("%s result = %.12g\n", who, to_double(v, d));
}

}  // namespace

int main(int argc, char** argv) {
  Flags f;
  QA qa;
  if (!f.parse(argc, argv)) {
    fprintf(stderr, "%s\n", f.error().c_str());
    return 1;
  }
  qa.start(argc, argv, f);
  Log log;
  if (!f.has("nolog")) {
    log.open(f.str("logfile", "reduction.txt").c_str());
    log.open_master(f.str("masterlog", "SdkMasterLog.csv").c_str());
  }

  Options o;
  memset(&o.cfg, 0, sizeof(o.cfg));
  if (!f.has("method") && !f.has("op")) {
    fprintf(stderr, "MISSING --method FLAG.\nYou must provide --method={ SUM | MIN | MAX }.\n");
    return 1;
  }
  if (!parse_op(f.str("method", f.str("op")), &o.op)) {
    fprintf(stderr, "unknown --method=%s; expected SUM, MIN or MAX\n", f.str("method", f.str("op")).c_str());
    return 1;
  }
  if (!parse_dtype(f.str("type", "int"), &o.dtype)) {
    fprintf(stderr, "unknown --type=%s; expected int, float, double or bf16\n", f.str("type").c_str());
    return 1;
  }
  o.n = (uint64_t)f.i64("n", 1ll << 24);
  o.iters = (int)f.i64("iters", 100);
  o.warmup = (int)std::max<long long>(1, f.i64("warmup", 3));
  o.rotate = !f.has("norotate");
  o.cpufinal = f.has("cpufinal");
  o.cputhresh = (int)f.i64("cputhresh", 1);
  o.seed = (uint64_t)f.i64("seed", 2012);
  o.hbm_peak_gbs = f.f64("peak", o.hbm_peak_gbs);
  const std::string fill = f.str("fill", "sdk");
  o.fill = fill == "full" ? 1 : (fill == "signed" ? 2 : 0);
  o.cfg.variant = parse_variant(f.str("kernel", "0"));
  o.cfg.threads = (int)f.i64("threads", 0);
  o.cfg.ctas = (int)f.i64("maxblocks", f.i64("ctas", 0));
  o.cfg.cluster = (int)f.i64("cluster", 0);
  o.cfg.unroll = (int)f.i64("unroll", 0);
  o.cfg.stages = (int)f.i64("stages", 0);
  o.cfg.pdl = f.has("pdl") ? (int)f.i64("pdl", B2R_PDL_INPUT_READY) : 0;  // --pdl: the timed launches reduce buffers filled long ago (early streaming); --pdl=1: safe flavour

  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    log.print("No CUDA device available.\n");
    return qa.finish(QA_WAIVED);
  }
  const int dev = (int)f.i64("device", 0);
  BENCH_CUDA(cudaSetDevice(dev));
  cudaDeviceProp prop;
  BENCH_CUDA(cudaGetDeviceProperties(&prop, dev));
  log.print("Using Device %d: %s (%d SMs, sm_%d%d)\n\n", dev, prop.name, prop.multiProcessorCount, prop.major, prop.minor);
  if (prop.major < 10) {
    log.print("This library targets sm_100a (Blackwell); device is sm_%d%d.\n", prop.major, prop.minor);
    return qa.finish(QA_WAIVED);
  }
  b2r_workspace_t ws;
  BENCH_B2R(b2r_workspace_create(&ws, dev));
  ClockLog::get().start(ndev);

  FILE* csv = f.has("csv") ? fopen(f.str("csv").c_str(), "a") : nullptr;
  FILE* jsonl = f.has("jsonl") ? fopen(f.str("jsonl").c_str(), "a") : nullptr;
  auto emit = [&](const Options& oo, const Result& r) {
    const size_t bytes = oo.n * b2r_dtype_size(oo.dtype);
    const double gbs = bytes / r.event_ms * 1e-6, gibs = bytes / r.event_ms * 1e3 / 1073741824.0;
    if (csv)
      fprintf(csv, "%s,%s,%llu,%s,%d,%d,%d,%.3f,%.3f,%.3f,%.4f,%s\n", b2r_dtype_name(oo.dtype), b2r_op_name(oo.op),
              (unsigned long long)oo.n, variant_name(r.geom.variant), r.geom.threads, r.geom.ctas, r.geom.cluster, r.event_ms * 1e3,
              gbs, gibs, gbs / oo.hbm_peak_gbs, r.ok ? "PASSED" : "FAILED");
    if (jsonl)
      fprintf(jsonl,
              "{\"bench\":\"reduce_local\",\"dtype\":\"%s\",\"op\":\"%s\",\"n\":%llu,\"bytes\":%zu,\"variant\":\"%s\",\"threads\":%d,"
              "\"ctas\":%d,\"cluster\":%d,\"us_event\":%.3f,\"us_host\":%.3f,\"GBps\":%.3f,\"GiBps\":%.3f,\"frac_hbm_measured\":%.4f,"
              "\"rotating\":%s,\"ok\":%s}\n",
              b2r_dtype_name(oo.dtype), b2r_op_name(oo.op), (unsigned long long)oo.n, bytes, variant_name(r.geom.variant), r.geom.threads,
              r.geom.ctas, r.geom.cluster, r.event_ms * 1e3, r.host_ms * 1e3, gbs, gibs, gbs / oo.hbm_peak_gbs, oo.rotate ? "true" : "false",
              r.ok ? "true" : "false");
  };

  bool all_ok = true;
  if (f.has("shmoo")) {
    // Working replacement of the reference's disabled sweep (reduction.cpp:576-657): n = 2^10 .. 2^30 (bounded by
    // --n if given) for every kernel variant; one line per size, columns = variants, values = GB/s (device events).
    const uint64_t nmax = f.has("n") ? o.n : (1ull << 28);
    log.print("Reduction shmoo: %s %s, GB/s by kernel variant (device events, rotating inputs)\n", b2r_dtype_name(o.dtype), b2r_op_name(o.op));
    log.print("%12s %10s %10s %10s %10s\n", "Elements", "auto", "ldg128", "ldg256", "tma");
    for (uint64_t n = 1024; n <= nmax; n *= 4) {
      log.print("%12llu", (unsigned long long)n);
      for (int v = 0; v <= 3; ++v) {
        Options oo = o;
        oo.n = n;
        oo.cfg.variant = v;
        oo.iters = std::min(o.iters, n > (1ull << 26) ? 20 : 100);
        Result r = run_one(oo, ws, &log);
        all_ok = all_ok && r.ok;
        log.print(" %10.1f", n * b2r_dtype_size(o.dtype) / r.event_ms * 1e-6);
        emit(oo, r);
      }
      log.print("\n");
    }
  } else {
    log.print("METHOD: %s\n", b2r_op_name(o.op));
    log.print("%llu elements\n", (unsigned long long)o.n);
    Result r = run_one(o, ws, &log);
    log.print("%d threads (max)\n", r.geom.threads);
    log.print("%d blocks\n\n", r.geom.ctas);
    const size_t bytes = o.n * b2r_dtype_size(o.dtype);
    const double host_s = r.host_ms * 1e-3;
    log.print_master("Reduction, Throughput = %.4f GB/s, Time = %.5f s, Size = %llu Elements, NumDevsUsed = %d, Workgroup = %u\n",
                     1.0e-9 * (double)bytes / host_s, host_s, (unsigned long long)o.n, 1, (unsigned)r.geom.threads);
    log.print("Device-event time (back-to-back, %s): %.3f us -> %.2f GB/s = %.2f GiB/s (%.1f%% of measured HBM copy roofline %.1f GB/s)\n",
              o.rotate ? "rotating inputs > L2" : "single L2-resident input", r.event_ms * 1e3, bytes / r.event_ms * 1e-6,
              bytes / r.event_ms * 1e3 / 1073741824.0, 100.0 * bytes / r.event_ms * 1e-6 / o.hbm_peak_gbs, o.hbm_peak_gbs);
    log.print("Kernel: %s, %d CTAs x %d threads, cluster %d, unroll %d, pdl %d\n", variant_name(r.geom.variant), r.geom.ctas, r.geom.threads,
              r.geom.cluster, r.geom.unroll, r.geom.pdl);
    log.print("\n");
    print_value(log, "GPU", r.gpu, o.dtype);
    print_value(log, "CPU", r.cpu, o.dtype);
    log.print("\n");
    all_ok = r.ok;
    emit(o, r);
  }
  if (csv) fclose(csv);
  if (jsonl) fclose(jsonl);
  b2r_workspace_destroy(ws);
  ClockLog::get().stop();
  if (!ClockLog::get().print(stdout)) all_ok = false;  // a slowdown during the timed region invalidates the number
  cudaDeviceReset();
  return qa.finish(all_ok ? QA_PASSED : QA_FAILED);
}
