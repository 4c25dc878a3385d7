// reduce_gpu_mpi.cu — the reference's mpi/reduce.c loop with one MPI rank per GPU and the data never leaving the GPUs.
// Counterpart of the vendored simpleMPI wiring (simpleMPI.cpp:46-112: MPI_Scatter -> H2D -> kernel -> D2H -> host sum ->
// MPI_Reduce of one float), turned inside out: MPI (here the bundled mini-MPI, any MPI would do) is used ONLY to learn
// rank/size and to hand every rank the bootstrap id; buffers are symmetric device memory and MPI_Reduce's job is done by
// b2r_reduce over NVLink.
//   make examples && bin/b2r_mpirun -np 8 bin/reduce_gpu_mpi [total_mib=2048] [retries=5]
// Output: the reference's rows "# DATATYPE OP NODES GB/sec" / "INT MAX 8   1234.567" (GiB/s of the whole problem).
#include <cuda_runtime.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <vector>

#include "b2r.h"
#include "mpi.h"

#define CK(x)                                                                                          \
  do {                                                                                                 \
    b2r_status_t s_ = (x);                                                                             \
    if (s_ != B2R_OK) {                                                                                \
      fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, b2r_status_string(s_), b2r_last_error_string()); \
      MPI_Abort(MPI_COMM_WORLD, 2);                                                                    \
    }                                                                                                  \
  } while (0)

int main(int argc, char** argv) {
  if (MPI_Init(&argc, &argv) != MPI_SUCCESS) return 1;
  int rank = 0, size = 1;
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  MPI_Comm_size(MPI_COMM_WORLD, &size);
  const size_t total_bytes = (size_t)(argc > 1 ? atol(argv[1]) : 2048) << 20;
  const int retries = argc > 2 ? atoi(argv[2]) : 5;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < size) {
    if (rank == 0) fprintf(stderr, "need %d GPUs, found %d\n", size, ndev);
    MPI_Abort(MPI_COMM_WORLD, 3);
  }

  // the only thing MPI carries: a bootstrap id chosen by rank 0 (an int is enough to name the rendezvous socket)
  int id = rank == 0 ? (int)getpid() : 0;
  if (rank == 0) for (int r = 1; r < size; ++r) MPI_Send(&id, 1, MPI_INT, r, 0, MPI_COMM_WORLD);
  else MPI_Recv(&id, 1, MPI_INT, 0, 0, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
  char boot[64];
  snprintf(boot, sizeof(boot), "gpu-mpi-%d", id);

  const size_t per_rank = total_bytes / size;
  b2r_comm_config_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.heap_bytes = 2 * per_rank + (1 << 20);
  cfg.use_multicast = -1;
  cfg.symmetric_only = 1;
  b2r_comm_t comm;
  CK(b2r_comm_init_rank(&comm, size, rank, rank, boot, &cfg));
  void *send = nullptr, *recv = nullptr;
  CK(b2r_symm_alloc(comm, per_rank, &send));
  CK(b2r_symm_alloc(comm, per_rank, &recv));
  cudaStream_t stream;
  cudaStreamCreate(&stream);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);

  int failures = 0;
  const b2r_dtype_t types[2] = {B2R_INT32, B2R_FLOAT64};
  const char* tnames[2] = {"INT", "DOUBLE"};
  const b2r_op_t ops[3] = {B2R_MAX, B2R_MIN, B2R_SUM};
  if (rank == 0) printf("# DATATYPE OP NODES GB/sec\n");
  for (int x = -1; x < retries; ++x) {  // x == -1: warm-up pass (reduce.c:61-64)
    for (int t = 0; t < 2; ++t) {
      const uint64_t count = per_rank / b2r_dtype_size(types[t]);
      CK(b2r_fill_device(send, count, types[t], 0x123 + rank, 0, 1, stream));
      for (int o = 0; o < 3; ++o) {
        cudaMemsetAsync(recv, 0, per_rank, stream);
        CK(b2r_barrier(comm, stream));
        cudaEventRecord(e0, stream);
        CK(b2r_reduce(send, recv, count, types[t], ops[o], 0, comm, stream));
        cudaEventRecord(e1, stream);
        CK(b2r_comm_check(comm, stream));
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        double mine = ms, worst = 0;  // max over ranks, not rank 0's clock
        MPI_Reduce(&mine, &worst, 1, MPI_DOUBLE, MPI_MAX, 0, MPI_COMM_WORLD);
        if (rank == 0 && x >= 0) printf("%s %s %d %10.3lf\n", tnames[t], b2r_op_name(ops[o]), size, (double)total_bytes / (worst * 1e-3) / 1073741824.0);
        if (rank == 0 && x == retries - 1) {
          // self-check of the last pass (the reference's reduce.c has none): the first 1 Mi elements at the root against the
          // element-wise host oracle over every rank's regenerated input (integers and min/max exact; fp64 SUM to 1e-12)
          const uint64_t n = count < (1u << 20) ? count : (1u << 20);
          const size_t esz = b2r_dtype_size(types[t]);
          std::vector<unsigned char> want(n * esz), in(n * esz), got(n * esz);
          b2r_fill_host(want.data(), n, types[t], 0x123, 0, 1);
          for (int r = 1; r < size; ++r) {
            b2r_fill_host(in.data(), n, types[t], 0x123 + r, 0, 1);
            b2r_combine_host(want.data(), in.data(), n, types[t], ops[o]);
          }
          cudaMemcpy(got.data(), recv, n * esz, cudaMemcpyDeviceToHost);
          bool ok = memcmp(got.data(), want.data(), n * esz) == 0;
          if (!ok && types[t] == B2R_FLOAT64 && ops[o] == B2R_SUM) {  // the switch (NVLS) adds in its own order
            ok = true;
            for (uint64_t i = 0; i < n; ++i) {
              const double a = ((const double*)got.data())[i], b = ((const double*)want.data())[i];
              if (fabs(a - b) > 1e-12 * (fabs(b) > 1 ? fabs(b) : 1)) ok = false;
            }
          }
          if (!ok) ++failures;
        }
      }
    }
  }
  if (rank == 0) printf("%s\n", failures ? "FAILED" : "PASSED");
  CK(b2r_comm_destroy(comm));
  MPI_Finalize();
  return failures ? 1 : 0;
}
