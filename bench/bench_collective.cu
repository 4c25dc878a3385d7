// bench_collective.cu — multi-GPU reduce / allreduce benchmark: the reference's MPI_Reduce loop on NVLink.
//
// Counterpart of mpi/reduce.c (main :9-108). Default mode reproduces it: a fixed TOTAL problem (512 Mi int32 / 256 Mi fp64 =
// 2 GiB each, mpi/constants.h:1-2) is sharded over the P ranks (count = total / P, reduce.c:43-44), one untimed warm-up,
// then RETRY_COUNT=5 x {MAX, MIN, SUM} x {INT, DOUBLE} element-wise reductions to root 0, printing the reference's rows
//     # DATATYPE OP NODES GB/sec
//     INT MAX 8    123.456          <- (total bytes / time) / 2^30, the reference's unit (reduce.c:79,93)
// so mpi/getAvgs.sh-style tooling and makePlots.gp work unchanged on the output. Differences by design (SURVEY §2.7
// items 8-10): a device barrier before every timed call, device-event timing, MAX over ranks instead of rank 0's clock;
// decimal GB/s and bus bandwidth go to the JSONL/CSV sink, not into the reference-format rows.
//
//   bench_collective [--gpus=N] [--total-mib=2048] [--retries=5] [--types=INT,DOUBLE[,BF16,FLOAT]] [--coll=reduce|allreduce]
//                    [--algo=auto|oneshot|twoshot|tree|nvls|push|ring|ll|posted] [--root=0] [--check] [--jsonl=path] [--raw=path]
//   bench_collective --tune=ring|push|nvls|twoshot [--coll=] [--bytes=] [--type=] [--iters=]   (in-process launch-shape grid)
//   bench_collective --sweep [--min-bytes=1024] [--max-bytes=4294967296] [--types=...] [--colls=reduce,allreduce]
//                    [--algos=auto,oneshot,twoshot,tree,nvls,push,ring,ll,posted] [--iters=20] [--jsonl=path] [--csv=path]
// Runs as ONE process driving all GPUs (b2r_comm_init_all); buffers come from the symmetric heap (zero copy).
#include <algorithm>
#include <cmath>
#include <sstream>
#include <thread>

#include "bench_common.h"

using namespace b2rbench;

namespace {

struct Rank {
  int dev;
  b2r_comm_t comm;
  cudaStream_t stream;
  cudaEvent_t e0, e1;
  void* send;
  void* recv;
};

std::vector<std::string> split(const std::string& s, char sep = ',') {
  std::vector<std::string> out;
  std::stringstream ss(s);
  std::string item;
  while (std::getline(ss, item, sep))
    if (!item.empty()) out.push_back(item);
  return out;
}

const char* ref_type_name(b2r_dtype_t d) {
  switch (d) {
    case B2R_INT32: return "INT";
    case B2R_FLOAT32: return "FLOAT";
    case B2R_FLOAT64: return "DOUBLE";
    case B2R_BFLOAT16: return "BF16";
  }
  return "?";
}
bool parse_ref_type(const std::string& s, b2r_dtype_t* d) { return parse_dtype(s, d); }

b2r_algo_t parse_algo(const std::string& s) {
  if (s == "oneshot") return B2R_ALGO_ONESHOT;
  if (s == "twoshot") return B2R_ALGO_TWOSHOT;
  if (s == "tree") return B2R_ALGO_TREE;
  if (s == "nvls") return B2R_ALGO_NVLS;
  if (s == "push") return B2R_ALGO_PUSH;
  if (s == "ring") return B2R_ALGO_RING;
  if (s == "ll") return B2R_ALGO_LL;
  if (s == "posted") return B2R_ALGO_POSTED;
  return B2R_ALGO_AUTO;
}

// One timed collective (or `iters` back-to-back): returns milliseconds per call, max over ranks; <0 on failure.
// One host thread per GPU issues that GPU's launches (like one MPI rank per GPU would), so the host launch rate of a
// single thread driving P devices does not masquerade as collective latency.
double timed(std::vector<Rank>& R, bool allreduce, uint64_t count, b2r_dtype_t dt, b2r_op_t op, int root, b2r_algo_t algo, int iters) {
  const int P = (int)R.size();
  ClockLog::Window clocks_window;  // clock samples taken from here to the end of the call count as "during a timed region"
  std::vector<int> failed(P, 0);
  std::vector<std::thread> th;
  for (int i = 0; i < P; ++i) {
    th.emplace_back([&, i] {
      Rank& r = R[i];
      cudaSetDevice(r.dev);
      if (b2r_barrier(r.comm, r.stream) != B2R_OK) { failed[i] = 1; return; }  // every rank enters the timed region together
      cudaEventRecord(r.e0, r.stream);
      for (int k = 0; k < iters; ++k) {
        b2r_status_t st = allreduce ? b2r_allreduce_ex(r.send, r.recv, count, dt, op, r.comm, r.stream, algo)
                                    : b2r_reduce_ex(r.send, r.recv, count, dt, op, root, r.comm, r.stream, algo);
        if (st != B2R_OK) {
          fprintf(stderr, "collective failed on rank %d: %s (%s)\n", i, b2r_status_string(st), b2r_last_error_string());
          failed[i] = 1;
          return;
        }
      }
      cudaEventRecord(r.e1, r.stream);
    });
  }
  for (auto& t : th) t.join();
  double worst = 0;
  for (int i = 0; i < P; ++i) {
    Rank& r = R[i];
    cudaSetDevice(r.dev);
    if (b2r_comm_check(r.comm, r.stream) != B2R_OK) {
      fprintf(stderr, "rank %d: %s\n", i, b2r_last_error_string());
      return -1;
    }
    if (failed[i]) return -1;
    float ms = 0;
    cudaEventElapsedTime(&ms, r.e0, r.e1);
    worst = std::max(worst, (double)ms / iters);
  }
  return worst;
}

// Fills the send buffers, runs once, and compares the result at root (and at every rank for allreduce) with the CPU
// element-wise oracle (exact for integers and min/max; NVLS / tree fp sums within a relative bound).
bool verify(std::vector<Rank>& R, bool allreduce, uint64_t count, b2r_dtype_t dt, b2r_op_t op, int root, b2r_algo_t algo) {
  const size_t esz = b2r_dtype_size(dt);
  const uint64_t n = std::min<uint64_t>(count, 1u << 22);  // oracle on the first 4 Mi elements
  const int P = (int)R.size();
  std::vector<std::vector<unsigned char>> h(P, std::vector<unsigned char>(n * esz));
  for (int r = 0; r < P; ++r) {
    cudaSetDevice(R[r].dev);
    BENCH_B2R(b2r_fill_device(R[r].send, count, dt, 4242 + r, 0, 2, R[r].stream));
    BENCH_B2R(b2r_fill_host(h[r].data(), n, dt, 4242 + r, 0, 2));
  }
  if (timed(R, allreduce, count, dt, op, root, algo, 1) < 0) return false;
  std::vector<const void*> srcs(P);
  for (int r = 0; r < P; ++r) srcs[r] = h[r].data();
  std::vector<unsigned char> want(n * esz), got(n * esz);
  BENCH_B2R(b2r_host_reduce_threads(srcs.data(), want.data(), n, dt, op, P, 0, B2R_ALGO_ONESHOT));
  bool ok = true;
  for (int r = 0; r < P; ++r) {
    if (!allreduce && r != root) continue;
    cudaSetDevice(R[r].dev);
    BENCH_CUDA(cudaMemcpy(got.data(), R[r].recv, n * esz, cudaMemcpyDeviceToHost));
    if (memcmp(got.data(), want.data(), n * esz) == 0) continue;
    // not bit-identical: acceptable only for fp sums whose order differs (tree, NVLS)
    double worst = 0;
    for (uint64_t i = 0; i < n; ++i) {
      double a = 0, b = 0;
      if (dt == B2R_INT32) { a = ((int32_t*)got.data())[i]; b = ((int32_t*)want.data())[i]; }
      else if (dt == B2R_FLOAT32) { a = ((float*)got.data())[i]; b = ((float*)want.data())[i]; }
      else if (dt == B2R_FLOAT64) { a = ((double*)got.data())[i]; b = ((double*)want.data())[i]; }
      else {
        uint32_t x = (uint32_t)((uint16_t*)got.data())[i] << 16, y = (uint32_t)((uint16_t*)want.data())[i] << 16;
        float fa, fb;
        memcpy(&fa, &x, 4);
        memcpy(&fb, &y, 4);
        a = fa;
        b = fb;
      }
      worst = std::max(worst, std::fabs(a - b) / std::max(1.0, std::fabs(b)));
    }
    const double tol = dt == B2R_INT32 ? 0 : (dt == B2R_FLOAT64 ? 1e-13 : (dt == B2R_FLOAT32 ? 1e-5 : 2e-2));
    if (worst > tol || op != B2R_SUM) {
      fprintf(stderr, "VERIFY FAILED rank %d %s %s count %llu: max rel err %.3g\n", r, ref_type_name(dt), b2r_op_name(op), (unsigned long long)count, worst);
      ok = false;
    }
  }
  return ok;
}

}  // namespace

int main(int argc, char** argv) {
  Flags f;
  if (!f.parse(argc, argv)) { fprintf(stderr, "%s\n", f.error().c_str()); return 1; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { fprintf(stderr, "no CUDA device\n"); return 2; }
  const int P = (int)f.i64("gpus", ndev);
  if (P < 1 || P > ndev) { fprintf(stderr, "--gpus=%d but %d devices are visible\n", P, ndev); return 1; }
  ClockLog::get().start(P);
  const bool nvlink = f.has("nvlink");  // --nvlink: measured NVLink payload bytes per call (NVML counters) next to every sweep row
  const bool tune = f.has("tune");
  const bool sweep = f.has("sweep") || tune;
  const uint64_t total_bytes = (uint64_t)f.i64("total-mib", 2048) << 20;
  const uint64_t max_bytes = tune ? (uint64_t)f.i64("bytes", 256ll << 20)
                             : ((sweep || f.has("plan")) ? (uint64_t)f.i64("max-bytes", 1ll << 32) : total_bytes / P);
  const int root = (int)f.i64("root", 0);

  b2r_comm_config_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.heap_bytes = 2 * (max_bytes + 4096) + (1 << 20);
  cfg.use_multicast = f.has("no-multicast") ? 0 : -1;
  cfg.symmetric_only = 1;
  cfg.timeout_ms = (int)f.i64("timeout-ms", 20000);
  cfg.max_ctas = (int)f.i64("ctas", 0);
  std::vector<b2r_comm_t> comms(P);
  std::vector<int> devs(P);
  for (int i = 0; i < P; ++i) devs[i] = i;
  BENCH_B2R(b2r_comm_init_all(comms.data(), P, devs.data(), &cfg));
  int has_mc = 0;
  b2r_comm_has_multicast(comms[0], &has_mc);
  std::vector<Rank> R(P);
  for (int i = 0; i < P; ++i) {
    R[i].dev = i;
    R[i].comm = comms[i];
    BENCH_CUDA(cudaSetDevice(i));
    BENCH_CUDA(cudaStreamCreateWithFlags(&R[i].stream, cudaStreamNonBlocking));
    BENCH_CUDA(cudaEventCreate(&R[i].e0));
    BENCH_CUDA(cudaEventCreate(&R[i].e1));
    BENCH_B2R(b2r_symm_alloc(comms[i], max_bytes + 4096, &R[i].send));
    BENCH_B2R(b2r_symm_alloc(comms[i], max_bytes + 4096, &R[i].recv));
  }
  fprintf(stderr, "# b200reduce bench_collective: %d GPUs, multicast(NVLS) %s, %s\n", P, has_mc ? "available" : "not available",
          sweep ? "size sweep" : "reference loop");

  FILE* jsonl = f.has("jsonl") ? fopen(f.str("jsonl").c_str(), "a") : nullptr;
  FILE* csv = f.has("csv") ? fopen(f.str("csv").c_str(), "a") : nullptr;
  FILE* raw = f.has("raw") ? fopen(f.str("raw").c_str(), "a") : nullptr;
  std::vector<b2r_dtype_t> types;
  for (auto& t : split(f.str("types", sweep ? "INT,DOUBLE,BF16" : "INT,DOUBLE"))) {
    b2r_dtype_t d;
    if (!parse_ref_type(t, &d)) { fprintf(stderr, "unknown type %s\n", t.c_str()); return 1; }
    types.push_back(d);
  }
  bool all_ok = true;

  if (f.has("plan")) {
    // --plan=<file>: an explicit list of measurements run in ONE process (communicator set-up is the expensive part at 8
    // GPUs). One measurement per line, '#' starts a comment:
    //     <label> <reduce|allreduce> <type> <op> <algo> <bytes> [B2R_KNOB=value ...]
    // The knobs are exported for that line only (the library re-reads them: b2r_refresh_env). Output: one row per line in
    // the sweep format, prefixed by the label. --check verifies every row of at most --check-max bytes against the oracle.
    FILE* pf = fopen(f.str("plan").c_str(), "r");
    if (!pf) { fprintf(stderr, "cannot open plan %s\n", f.str("plan").c_str()); return 1; }
    const int iters = (int)f.i64("iters", 20);
    const uint64_t check_max = (uint64_t)f.i64("check-max", 64ll << 20);
    printf("# %-22s %-9s %-6s %-4s %-8s %14s %10s %10s %10s  knobs\n", "label", "coll", "dtype", "op", "algo", "bytes", "us", "algbw GB/s", "busbw GB/s");
    char linebuf[1024];
    while (fgets(linebuf, sizeof(linebuf), pf)) {
      std::string line(linebuf);
      const size_t hash = line.find('#');
      if (hash != std::string::npos) line = line.substr(0, hash);
      std::stringstream ss(line);
      std::vector<std::string> tok;
      std::string t;
      while (ss >> t) tok.push_back(t);
      if (tok.size() < 6) continue;
      const bool allreduce = tok[1] == "allreduce";
      b2r_dtype_t dt;
      b2r_op_t op;
      if (!parse_dtype(tok[2], &dt) || !parse_op(tok[3], &op)) { fprintf(stderr, "bad plan line: %s", linebuf); continue; }
      const b2r_algo_t algo = parse_algo(tok[4]);
      const uint64_t bytes = strtoull(tok[5].c_str(), nullptr, 0);
      if (bytes == 0 || bytes > max_bytes) { fprintf(stderr, "plan line skipped (bytes %llu outside (0, %llu]): %s", (unsigned long long)bytes, (unsigned long long)max_bytes, linebuf); continue; }
      std::vector<std::string> names;
      std::string knobs;
      for (size_t k = 6; k < tok.size(); ++k) {
        const size_t eq = tok[k].find('=');
        if (eq == std::string::npos) continue;
        names.push_back(tok[k].substr(0, eq));
        setenv(names.back().c_str(), tok[k].substr(eq + 1).c_str(), 1);
        knobs += tok[k] + " ";
      }
      b2r_refresh_env();
      const uint64_t count = bytes / b2r_dtype_size(dt);
      bool row_ok = true;
      if (algo == B2R_ALGO_LL) {
        uint64_t ll_max = 0;
        b2r_comm_ll_max_bytes(R[0].comm, &ll_max);
        if (bytes > ll_max) row_ok = false;
      }
      if (algo == B2R_ALGO_POSTED) {
        uint64_t pmax = 0;
        b2r_comm_posted_max_bytes(R[0].comm, &pmax);
        if (allreduce || bytes > pmax) row_ok = false;
      }
      if (algo == B2R_ALGO_NVLS && !has_mc) row_ok = false;  // fp min / max: the keyed form
      double ms = -1;
      if (row_ok) {
        if (f.has("check") && bytes <= check_max && !verify(R, allreduce, count, dt, op, root, algo)) {
          all_ok = false;
          printf("  %-22s VERIFY FAILED\n", tok[0].c_str());
        }
        const int it = bytes >= (1ull << 30) ? std::max(3, iters / 4) : (bytes >= (64ull << 20) ? std::max(5, iters / 2) : (bytes >= (4ull << 20) ? iters * 2 : iters * 5));
        if (timed(R, allreduce, count, dt, op, root, algo, 3) >= 0) ms = timed(R, allreduce, count, dt, op, root, algo, it);
      }
      if (ms > 0) {
        const double algbw = bytes / ms * 1e-6;
        const double busbw = allreduce ? algbw * 2.0 * (P - 1) / P : algbw;
        const b2r_algo_t used = algo == B2R_ALGO_AUTO ? b2r_select_algo(allreduce, bytes, dt, op, P, has_mc) : algo;
        printf("  %-22s %-9s %-6s %-4s %-8s %14llu %10.2f %10.2f %10.2f  %s\n", tok[0].c_str(), allreduce ? "allreduce" : "reduce", ref_type_name(dt),
               b2r_op_name(op), b2r_algo_name(used), (unsigned long long)bytes, ms * 1e3, algbw, busbw, knobs.c_str());
        if (csv)
          fprintf(csv, "b2r,%s,%s,%s,%s,%d,%llu,%.3f,%.3f,%.3f,%s,%s\n", allreduce ? "allreduce" : "reduce", ref_type_name(dt), b2r_op_name(op),
                  b2r_algo_name(used), P, (unsigned long long)bytes, ms * 1e3, algbw, busbw, tok[0].c_str(), knobs.c_str());
      } else {
        printf("  %-22s %-9s %-6s %-4s %-8s %14llu  skipped or failed  %s\n", tok[0].c_str(), allreduce ? "allreduce" : "reduce", ref_type_name(dt),
               b2r_op_name(op), tok[4].c_str(), (unsigned long long)bytes, knobs.c_str());
        if (ms < 0 && row_ok) {  // a failed collective leaves the error word set: clear it so the next rows can run
          for (auto& r : R) { cudaSetDevice(r.dev); b2r_comm_check(r.comm, r.stream); }
        }
      }
      fflush(stdout);
      for (auto& nme : names) unsetenv(nme.c_str());
      b2r_refresh_env();
    }
    fclose(pf);
  } else if (tune) {
    // In-process launch-shape sweep: the library reads its B2R_* knobs from the environment at every call, so one process
    // (one communicator set-up, the expensive part at 8 GPUs) can walk a whole grid. --tune=ring|push|nvls|twoshot
    //   [--coll=reduce|allreduce] [--bytes=268435456] [--type=FLOAT] [--iters=10]
    const std::string which = f.str("tune", "ring");
    const b2r_algo_t algo = parse_algo(which);
    const bool allreduce = f.str("coll", which == "ring" ? "reduce" : "allreduce") == "allreduce";
    b2r_dtype_t dt = f.has("types") && !types.empty() ? types[0] : B2R_FLOAT32;
    if (f.has("type")) parse_dtype(f.str("type"), &dt);
    const uint64_t bytes = max_bytes, count = bytes / b2r_dtype_size(dt);
    const int iters = (int)f.i64("iters", 10);
    struct Knob { const char* name; std::vector<int> values; };
    std::vector<Knob> knobs;
    // ring: total time ~ (steps per CTA + P - 2) x step period, so a long chain wants MANY short steps spread over many CTAs
    // (8 GPUs, 256 MiB: 32 CTAs x 512 KiB is 16 steps + 6 of pipeline fill; 128 x 64 KiB is 32 + 6 at the same bytes per fence round);
    // 256 threads x >= 148 CTAs puts two CTAs on an SM, so one streams while the other sits in its release fence
    if (which == "ring") knobs = {{"B2R_RING_CTAS", {32, 64, 128, 256}}, {"B2R_RING_SUBBLOCKS", {1, 2, 4, 8}}, {"B2R_COLL_THREADS", {256, 512}}};
    else if (which == "push") knobs = {{"B2R_PUSH_CTAS", {32, 64, 128}}, {"B2R_COLL_THREADS", {256, 512}}};
    else if (which == "nvls") knobs = {{"B2R_COLL_CTAS", {8, 16, 32, 64, 128}}, {"B2R_COLL_THREADS", {128, 256, 512}}, {"B2R_COLL_UNROLL", {4, 8}}};
    else knobs = {{"B2R_COLL_CTAS", {32, 64, 128, 148}}, {"B2R_COLL_THREADS", {256, 512}}, {"B2R_COLL_UNROLL", {4, 8}}};
    printf("# tune %s %s %s %llu bytes on %d GPUs: knobs -> us, algbw GB/s\n", which.c_str(), allreduce ? "allreduce" : "reduce", ref_type_name(dt),
           (unsigned long long)bytes, P);
    std::vector<size_t> idx(knobs.size(), 0);
    for (;;) {
      std::string label;
      for (size_t k = 0; k < knobs.size(); ++k) {
        setenv(knobs[k].name, std::to_string(knobs[k].values[idx[k]]).c_str(), 1);
        label += std::string(knobs[k].name + 4) + "=" + std::to_string(knobs[k].values[idx[k]]) + " ";
      }
      b2r_refresh_env();  // the library caches B2R_* lookups; make it read the new knob values
      if (timed(R, allreduce, count, dt, B2R_SUM, root, algo, 2) < 0) return 3;
      const double ms = timed(R, allreduce, count, dt, B2R_SUM, root, algo, iters);
      if (ms < 0) return 3;
      printf("  %-60s %10.2f %10.2f\n", label.c_str(), ms * 1e3, bytes / ms * 1e-6);
      fflush(stdout);
      size_t k = 0;
      while (k < knobs.size() && ++idx[k] == knobs[k].values.size()) idx[k++] = 0;
      if (k == knobs.size()) break;
    }
    for (auto& kn : knobs) unsetenv(kn.name);
    b2r_refresh_env();
  } else if (!sweep) {
    const bool allreduce = f.str("coll", "reduce") == "allreduce";
    const b2r_algo_t algo = parse_algo(f.str("algo", "auto"));
    const int retries = (int)f.i64("retries", 5);
    const b2r_op_t ops[3] = {B2R_MAX, B2R_MIN, B2R_SUM};  // the reference's order (reduce.c:26-28)
    // data: per-rank seeded "full range" values, like genrand_int32 / genrand_res53 with seed {rank, ...} (reduce.c:38-57)
    for (b2r_dtype_t dt : types) {
      const uint64_t count = total_bytes / b2r_dtype_size(dt) / P;
      for (int r = 0; r < P; ++r) {
        cudaSetDevice(R[r].dev);
        BENCH_B2R(b2r_fill_device(R[r].send, count, dt, 0x123 + r, 0, 1, R[r].stream));
      }
      if (timed(R, allreduce, count, dt, B2R_SUM, root, algo, 1) < 0) return 3;  // warm-up (reduce.c:61-64)
    }
    printf("# DATATYPE OP NODES GB/sec\n");
    if (raw) fprintf(raw, "# DATATYPE OP NODES GB/sec\n");
    for (int x = 0; x < retries; ++x) {
      for (b2r_dtype_t dt : types) {
        const uint64_t count = total_bytes / b2r_dtype_size(dt) / P;
        for (b2r_op_t op : ops) {
          for (auto& r : R) {  // bzero(recv) like reduce.c:74
            cudaSetDevice(r.dev);
            cudaMemsetAsync(r.recv, 0, count * b2r_dtype_size(dt), r.stream);
          }
          const double ms = timed(R, allreduce, count, dt, op, root, algo, 1);
          if (ms < 0) return 3;
          const double gib = (double)total_bytes / (ms * 1e-3) / 1073741824.0;
          printf("%s %s %d %10.3lf\n", ref_type_name(dt), b2r_op_name(op), P, gib);
          if (raw) fprintf(raw, "%s %s %d %10.3lf\n", ref_type_name(dt), b2r_op_name(op), P, gib);
          if (jsonl) {
            const uint64_t bytes = count * b2r_dtype_size(dt);
            fprintf(jsonl,
                    "{\"bench\":\"reduce_c_loop\",\"impl\":\"b2r\",\"coll\":\"%s\",\"dtype\":\"%s\",\"op\":\"%s\",\"gpus\":%d,\"bytes_per_rank\":%llu,"
                    "\"total_bytes\":%llu,\"ms\":%.5f,\"total_GiBps\":%.3f,\"total_GBps\":%.3f,\"algbw_GBps\":%.3f,\"algo\":\"%s\"}\n",
                    allreduce ? "allreduce" : "reduce", ref_type_name(dt), b2r_op_name(op), P, (unsigned long long)bytes,
                    (unsigned long long)total_bytes, ms, gib, total_bytes / ms * 1e-6, bytes / ms * 1e-6,
                    b2r_algo_name(algo == B2R_ALGO_AUTO ? b2r_select_algo(allreduce, bytes, dt, op, P, has_mc) : algo));
          }
        }
      }
    }
    if (f.has("check"))
      for (b2r_dtype_t dt : types)
        for (b2r_op_t op : ops) all_ok = verify(R, allreduce, total_bytes / b2r_dtype_size(dt) / P, dt, op, root, algo) && all_ok;
  } else {
    const uint64_t min_bytes = (uint64_t)f.i64("min-bytes", 1024);
    const int iters = (int)f.i64("iters", 20);
    printf("# %-9s %-6s %-4s %-8s %14s %10s %10s %10s\n", "coll", "dtype", "op", "algo", "bytes", "us", "algbw GB/s", "busbw GB/s");
    for (auto& coll : split(f.str("colls", "reduce,allreduce"))) {
      const bool allreduce = coll == "allreduce";
      for (b2r_dtype_t dt : types) {
        for (auto& opname : split(f.str("ops", "SUM"))) {
          b2r_op_t op;
          if (!parse_op(opname, &op)) continue;
          for (auto& algoname : split(f.str("algos", "auto"))) {
            const b2r_algo_t algo = parse_algo(algoname);
            for (uint64_t bytes = min_bytes; bytes <= max_bytes; bytes *= (uint64_t)f.i64("factor", 4)) {
              const uint64_t count = bytes / b2r_dtype_size(dt);
              if (count == 0) continue;
              if (algo == B2R_ALGO_NVLS && !has_mc) continue;
              if (algo == B2R_ALGO_RING && allreduce && !(getenv("B2R_EXPERIMENTAL") && atoi(getenv("B2R_EXPERIMENTAL")))) continue;
              if (algo == B2R_ALGO_LL) {
                uint64_t ll_max = 0;
                b2r_comm_ll_max_bytes(R[0].comm, &ll_max);
                if (bytes > ll_max) continue;
              }
              if (algo == B2R_ALGO_POSTED) {
                uint64_t pmax = 0;
                b2r_comm_posted_max_bytes(R[0].comm, &pmax);
                if (allreduce || bytes > pmax) continue;
              }
              if ((algo == B2R_ALGO_ONESHOT || algo == B2R_ALGO_TREE) && bytes > (64ull << 20)) continue;  // pointless at this size
              if (f.has("check") && bytes <= (256ull << 20)) all_ok = verify(R, allreduce, count, dt, op, root, algo) && all_ok;
              const int it = bytes >= (1ull << 30) ? std::max(3, iters / 4) : (bytes >= (64ull << 20) ? std::max(5, iters / 2) : iters * 5);
              if (timed(R, allreduce, count, dt, op, root, algo, 3) < 0) return 3;
              std::vector<unsigned long long> tx0(P, 0), rx0(P, 0), tx1(P, 0), rx1(P, 0);
              bool have_nvl = nvlink;
              for (int r = 0; r < P && have_nvl; ++r) have_nvl = ClockLog::get().nvlink_bytes(r, &tx0[r], &rx0[r]);
              const double ms = timed(R, allreduce, count, dt, op, root, algo, it);
              if (ms < 0) return 3;
              if (have_nvl) std::this_thread::sleep_for(std::chrono::milliseconds(30));  // the counters are published with a delay
              for (int r = 0; r < P && have_nvl; ++r) have_nvl = ClockLog::get().nvlink_bytes(r, &tx1[r], &rx1[r]);
              const double algbw = bytes / ms * 1e-6;
              const double busbw = allreduce ? algbw * 2.0 * (P - 1) / P : algbw;  // nccl-tests convention
              const b2r_algo_t used = algo == B2R_ALGO_AUTO ? b2r_select_algo(allreduce, bytes, dt, op, P, has_mc) : algo;
              printf("  %-9s %-6s %-4s %-8s %14llu %10.2f %10.2f %10.2f", coll.c_str(), ref_type_name(dt), b2r_op_name(op), b2r_algo_name(used),
                     (unsigned long long)bytes, ms * 1e3, algbw, busbw);
              if (have_nvl) {
                // per call, busiest GPU in each direction, as a multiple of the message size (root ingress of an ideal reduce = 1.0)
                double txm = 0, rxm = 0;
                for (int r = 0; r < P; ++r) {
                  txm = std::max(txm, (double)(tx1[r] - tx0[r]) / it);
                  rxm = std::max(rxm, (double)(rx1[r] - rx0[r]) / it);
                }
                printf("   nvlink tx %.3fx rx %.3fx of the message (busiest GPU; root rx %.3fx)", txm / bytes, rxm / bytes,
                       (double)(rx1[root] - rx0[root]) / it / bytes);
              }
              printf("\n");
              fflush(stdout);
              if (csv)
                fprintf(csv, "b2r,%s,%s,%s,%s,%d,%llu,%.3f,%.3f,%.3f\n", coll.c_str(), ref_type_name(dt), b2r_op_name(op), b2r_algo_name(used), P,
                        (unsigned long long)bytes, ms * 1e3, algbw, busbw);
              if (jsonl)
                fprintf(jsonl,
                        "{\"bench\":\"collective_sweep\",\"impl\":\"b2r\",\"coll\":\"%s\",\"dtype\":\"%s\",\"op\":\"%s\",\"algo\":\"%s\",\"gpus\":%d,"
                        "\"bytes\":%llu,\"us\":%.3f,\"algbw_GBps\":%.3f,\"busbw_GBps\":%.3f}\n",
                        coll.c_str(), ref_type_name(dt), b2r_op_name(op), b2r_algo_name(used), P, (unsigned long long)bytes, ms * 1e3, algbw, busbw);
            }
          }
        }
      }
    }
  }
  if (jsonl) fclose(jsonl);
  if (csv) fclose(csv);
  if (raw) fclose(raw);
  ClockLog::get().stop();
  if (!ClockLog::get().print(stdout)) all_ok = false;  // a slowdown during a timed region invalidates the run
  for (auto& r : R) {
    cudaSetDevice(r.dev);
    cudaStreamSynchronize(r.stream);
  }
  for (auto& c : comms) b2r_comm_destroy(c);
  if (f.has("check")) printf("%s\n", all_ok ? "PASSED" : "FAILED");
  return all_ok ? 0 : 1;
}
