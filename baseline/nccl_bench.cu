// nccl_bench.cu — BASELINE (not product): ncclReduce / ncclAllReduce timed exactly like bin/bench_collective
// (one process, all GPUs, CUDA events per device, max over ranks, same sizes, same output columns).
//   nccl_bench [--gpus=N] --sweep [--min-bytes=1024] [--max-bytes=4294967296] [--types=INT,DOUBLE,BF16] [--colls=reduce,allreduce]
//              [--ops=SUM] [--iters=20] [--csv=path] [--jsonl=path]
//   nccl_bench [--gpus=N] [--total-mib=2048] [--retries=5] [--types=INT,DOUBLE]      (the reduce.c loop, reference row format)
#include <nccl.h>

#include <algorithm>
#include <sstream>
#include <thread>

#include "../bench/bench_common.h"

using namespace b2rbench;

#define NCCLCHECK(x)                                                                       \
  do {                                                                                     \
    ncclResult_t r_ = (x);                                                                 \
    if (r_ != ncclSuccess) {                                                               \
      fprintf(stderr, "%s:%d NCCL error %s\n", __FILE__, __LINE__, ncclGetErrorString(r_)); \
      exit(2);                                                                             \
    }                                                                                      \
  } while (0)

struct Rank {
  int dev;
  ncclComm_t comm;
  cudaStream_t stream;
  cudaEvent_t e0, e1;
  void *send, *recv;
};

static std::vector<std::string> split(const std::string& s) {
  std::vector<std::string> out;
  std::stringstream ss(s);
  std::string item;
  while (std::getline(ss, item, ','))
    if (!item.empty()) out.push_back(item);
  return out;
}
static ncclDataType_t nccl_type(b2r_dtype_t d) {
  switch (d) {
    case B2R_INT32: return ncclInt32;
    case B2R_FLOAT32: return ncclFloat32;
    case B2R_FLOAT64: return ncclFloat64;
    case B2R_BFLOAT16: return ncclBfloat16;
  }
  return ncclInt32;
}
static ncclRedOp_t nccl_op(b2r_op_t o) { return o == B2R_SUM ? ncclSum : (o == B2R_MIN ? ncclMin : ncclMax); }
static const char* ref_type_name(b2r_dtype_t d) { return d == B2R_INT32 ? "INT" : d == B2R_FLOAT32 ? "FLOAT" : d == B2R_FLOAT64 ? "DOUBLE" : "BF16"; }

// one host thread per GPU, like bench_collective (and like one MPI rank per GPU)
static double timed(std::vector<Rank>& R, bool allreduce, uint64_t count, b2r_dtype_t dt, b2r_op_t op, int root, int iters) {
  const int P = (int)R.size();
  ClockLog::Window clocks_window;
  std::vector<std::thread> th;
  for (int i = 0; i < P; ++i) {
    th.emplace_back([&, i] {
      Rank& r = R[i];
      cudaSetDevice(r.dev);
      NCCLCHECK(ncclAllReduce(r.recv, r.recv, 1, ncclInt32, ncclSum, r.comm, r.stream));  // entry barrier
      cudaEventRecord(r.e0, r.stream);
      for (int k = 0; k < iters; ++k) {
        if (allreduce) NCCLCHECK(ncclAllReduce(r.send, r.recv, count, nccl_type(dt), nccl_op(op), r.comm, r.stream));
        else NCCLCHECK(ncclReduce(r.send, r.recv, count, nccl_type(dt), nccl_op(op), root, r.comm, r.stream));
      }
      cudaEventRecord(r.e1, r.stream);
    });
  }
  for (auto& t : th) t.join();
  double worst = 0;
  for (auto& r : R) {
    cudaSetDevice(r.dev);
    BENCH_CUDA(cudaStreamSynchronize(r.stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, r.e0, r.e1);
    worst = std::max(worst, (double)ms / iters);
  }
  return worst;
}

int main(int argc, char** argv) {
  Flags f;
  if (!f.parse(argc, argv)) { fprintf(stderr, "%s\n", f.error().c_str()); return 1; }
  int ndev = 0;
  BENCH_CUDA(cudaGetDeviceCount(&ndev));
  const int P = (int)f.i64("gpus", ndev);
  const bool sweep = f.has("sweep");
  const uint64_t total_bytes = (uint64_t)f.i64("total-mib", 2048) << 20;
  const uint64_t max_bytes = sweep ? (uint64_t)f.i64("max-bytes", 1ll << 32) : total_bytes / P;
  const int root = (int)f.i64("root", 0);
  std::vector<int> devs(P);
  for (int i = 0; i < P; ++i) devs[i] = i;
  std::vector<ncclComm_t> comms(P);
  NCCLCHECK(ncclCommInitAll(comms.data(), P, devs.data()));
  ClockLog::get().start(P);
  std::vector<Rank> R(P);
  for (int i = 0; i < P; ++i) {
    R[i].dev = i;
    R[i].comm = comms[i];
    BENCH_CUDA(cudaSetDevice(i));
    BENCH_CUDA(cudaStreamCreateWithFlags(&R[i].stream, cudaStreamNonBlocking));
    BENCH_CUDA(cudaEventCreate(&R[i].e0));
    BENCH_CUDA(cudaEventCreate(&R[i].e1));
    BENCH_CUDA(cudaMalloc(&R[i].send, max_bytes + 4096));
    BENCH_CUDA(cudaMalloc(&R[i].recv, max_bytes + 4096));
    BENCH_B2R(b2r_fill_device(R[i].send, max_bytes / 4, B2R_INT32, 0x123 + i, 0, 0, R[i].stream));
  }
  int ver = 0;
  ncclGetVersion(&ver);
  fprintf(stderr, "# nccl_bench: NCCL %d, %d GPUs\n", ver, P);
  FILE* jsonl = f.has("jsonl") ? fopen(f.str("jsonl").c_str(), "a") : nullptr;
  FILE* csv = f.has("csv") ? fopen(f.str("csv").c_str(), "a") : nullptr;
  std::vector<b2r_dtype_t> types;
  for (auto& t : split(f.str("types", sweep ? "INT,DOUBLE,BF16" : "INT,DOUBLE"))) {
    b2r_dtype_t d;
    if (!parse_dtype(t, &d)) return 1;
    types.push_back(d);
  }
  if (!sweep) {
    const bool allreduce = f.str("coll", "reduce") == "allreduce";
    const int retries = (int)f.i64("retries", 5);
    const b2r_op_t ops[3] = {B2R_MAX, B2R_MIN, B2R_SUM};
    for (b2r_dtype_t dt : types) timed(R, allreduce, total_bytes / b2r_dtype_size(dt) / P, dt, B2R_SUM, root, 1);
    printf("# DATATYPE OP NODES GB/sec\n");
    for (int x = 0; x < retries; ++x)
      for (b2r_dtype_t dt : types)
        for (b2r_op_t op : ops) {
          const uint64_t count = total_bytes / b2r_dtype_size(dt) / P;
          const double ms = timed(R, allreduce, count, dt, op, root, 1);
          const double gib = (double)total_bytes / (ms * 1e-3) / 1073741824.0;
          printf("%s %s %d %10.3lf\n", ref_type_name(dt), b2r_op_name(op), P, gib);
          if (jsonl)
            fprintf(jsonl,
                    "{\"bench\":\"reduce_c_loop\",\"impl\":\"nccl\",\"coll\":\"%s\",\"dtype\":\"%s\",\"op\":\"%s\",\"gpus\":%d,\"bytes_per_rank\":%llu,"
                    "\"total_bytes\":%llu,\"ms\":%.5f,\"total_GiBps\":%.3f,\"algbw_GBps\":%.3f}\n",
                    allreduce ? "allreduce" : "reduce", ref_type_name(dt), b2r_op_name(op), P, (unsigned long long)(count * b2r_dtype_size(dt)),
                    (unsigned long long)total_bytes, ms, gib, count * b2r_dtype_size(dt) / ms * 1e-6);
        }
  } else {
    const uint64_t min_bytes = (uint64_t)f.i64("min-bytes", 1024);
    const int iters = (int)f.i64("iters", 20);
    printf("# %-9s %-6s %-4s %-8s %14s %10s %10s %10s\n", "coll", "dtype", "op", "algo", "bytes", "us", "algbw GB/s", "busbw GB/s");
    for (auto& coll : split(f.str("colls", "reduce,allreduce"))) {
      const bool allreduce = coll == "allreduce";
      for (b2r_dtype_t dt : types)
        for (auto& opname : split(f.str("ops", "SUM"))) {
          b2r_op_t op;
          if (!parse_op(opname, &op)) continue;
          for (uint64_t bytes = min_bytes; bytes <= max_bytes; bytes *= (uint64_t)f.i64("factor", 4)) {
            const uint64_t count = bytes / b2r_dtype_size(dt);
            if (!count) continue;
            const int it = bytes >= (1ull << 30) ? std::max(3, iters / 4) : (bytes >= (64ull << 20) ? std::max(5, iters / 2) : iters * 5);
            timed(R, allreduce, count, dt, op, root, 3);
            const double ms = timed(R, allreduce, count, dt, op, root, it);
            const double algbw = bytes / ms * 1e-6;
            const double busbw = allreduce ? algbw * 2.0 * (P - 1) / P : algbw;
            printf("  %-9s %-6s %-4s %-8s %14llu %10.2f %10.2f %10.2f\n", coll.c_str(), ref_type_name(dt), b2r_op_name(op), "nccl",
                   (unsigned long long)bytes, ms * 1e3, algbw, busbw);
            fflush(stdout);
            if (csv)
              fprintf(csv, "nccl,%s,%s,%s,nccl,%d,%llu,%.3f,%.3f,%.3f\n", coll.c_str(), ref_type_name(dt), b2r_op_name(op), P, (unsigned long long)bytes,
                      ms * 1e3, algbw, busbw);
            if (jsonl)
              fprintf(jsonl,
                      "{\"bench\":\"collective_sweep\",\"impl\":\"nccl\",\"coll\":\"%s\",\"dtype\":\"%s\",\"op\":\"%s\",\"algo\":\"nccl\",\"gpus\":%d,"
                      "\"bytes\":%llu,\"us\":%.3f,\"algbw_GBps\":%.3f,\"busbw_GBps\":%.3f}\n",
                      coll.c_str(), ref_type_name(dt), b2r_op_name(op), P, (unsigned long long)bytes, ms * 1e3, algbw, busbw);
          }
        }
    }
  }
  if (jsonl) fclose(jsonl);
  if (csv) fclose(csv);
  ClockLog::get().stop();
  const bool clocks_ok = ClockLog::get().print(stdout);
  for (auto& r : R) {
    cudaSetDevice(r.dev);
    cudaStreamSynchronize(r.stream);
    ncclCommDestroy(r.comm);
  }
  return clocks_ok ? 0 : 1;
}
