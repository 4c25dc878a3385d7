/* tree_reduce.c — the assignment's "hand-rolled tree reduce vs MPI_Reduce" comparison (BASELINE.json north star /
 * config 1). The snapshot of the reference contains only the MPI_Reduce side (mpi/reduce.c; SURVEY finds no
 * MPI_Send/MPI_Recv anywhere), so the custom tree is written here from scratch: a binomial tree over MPI_Send/MPI_Recv,
 * ceil(log2 P) rounds, partner = rank +/- 2^k in a root-relative numbering, element-wise combine at the receiver.
 * Same loop and row format as reduce.c (RETRY x {MAX,MIN,SUM} x {INT,DOUBLE}); every result is checked against
 * MPI_Reduce bit for bit (integers; doubles combine in the same pairing order in both only for MIN/MAX, so SUM is
 * checked to 1e-9 relative).
 *   b2r_mpirun -np 4 bin/tree_reduce [total_elements_per_type] [retries]
 * Output: "# IMPL DATATYPE OP NODES GB/sec" then rows "TREE INT SUM 4 1.234" / "LIB INT SUM 4 2.345" (GiB/s of the
 * TOTAL problem, like reduce.c:79). */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "mpi.h"

static void combine(void* acc, const void* in, int n, MPI_Datatype t, MPI_Op op) {
  int i;
  if (t == MPI_INT) {
    int* a = (int*)acc;
    const int* b = (const int*)in;
    for (i = 0; i < n; i++) a[i] = op == MPI_SUM ? (int)((unsigned)a[i] + (unsigned)b[i]) : (op == MPI_MIN ? (b[i] < a[i] ? b[i] : a[i]) : (b[i] > a[i] ? b[i] : a[i]));
  } else {
    double* a = (double*)acc;
    const double* b = (const double*)in;
    for (i = 0; i < n; i++) a[i] = op == MPI_SUM ? a[i] + b[i] : (op == MPI_MIN ? (b[i] < a[i] ? b[i] : a[i]) : (b[i] > a[i] ? b[i] : a[i]));
  }
}

/* binomial tree reduce to `root`; recv is only written at root; tmp and acc are scratch of n elements */
static void tree_reduce(const void* send, void* recv, void* acc, void* tmp, int n, MPI_Datatype t, MPI_Op op, int root, int rank, int size) {
  const size_t esz = t == MPI_INT ? sizeof(int) : sizeof(double);
  const int pos = (rank - root + size) % size;
  int bit;
  memcpy(acc, send, (size_t)n * esz);
  for (bit = 1; bit < size; bit <<= 1) {
    if (pos & bit) {
      MPI_Send(acc, n, t, (pos - bit + root) % size, bit, MPI_COMM_WORLD);
      break;
    } else if (pos + bit < size) {
      MPI_Recv(tmp, n, t, (pos + bit + root) % size, bit, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
      combine(acc, tmp, n, t, op);
    }
  }
  if (rank == root) memcpy(recv, acc, (size_t)n * esz);
}

static unsigned long long splitmix(unsigned long long* s) {
  unsigned long long z = (*s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

int main(int argc, char** argv) {
  int rank, size, x, i, o, errors = 0;
  long total = argc > 1 ? atol(argv[1]) : (1L << 24);
  int retries = argc > 2 ? atoi(argv[2]) : 5;
  const MPI_Op ops[3] = {MPI_MAX, MPI_MIN, MPI_SUM};
  const char* names[3] = {"MAX", "MIN", "SUM"};
  if (MPI_Init(&argc, &argv) != MPI_SUCCESS) return 1;
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  MPI_Comm_size(MPI_COMM_WORLD, &size);
  {
    const int n = (int)(total / size);
    unsigned long long seed = 0x123 + (unsigned long long)rank;
    int* si = (int*)malloc((size_t)n * sizeof(int));
    double* sd = (double*)malloc((size_t)n * sizeof(double));
    void* r_tree = malloc((size_t)n * sizeof(double));
    void* r_lib = malloc((size_t)n * sizeof(double));
    void* acc = malloc((size_t)n * sizeof(double));
    void* tmp = malloc((size_t)n * sizeof(double));
    for (i = 0; i < n; i++) si[i] = (int)(unsigned)(splitmix(&seed) >> 32);
    for (i = 0; i < n; i++) sd[i] = (double)(splitmix(&seed) >> 11) / 9007199254740992.0;
    MPI_Reduce(si, r_lib, n, MPI_INT, MPI_SUM, 0, MPI_COMM_WORLD); /* warm-up, like reduce.c:61-64 */
    tree_reduce(si, r_tree, acc, tmp, n, MPI_INT, MPI_SUM, 0, rank, size);
    if (rank == 0) printf("# IMPL DATATYPE OP NODES GB/sec\n");
    for (x = 0; x < retries; x++) {
      int t;
      for (t = 0; t < 2; t++) {
        const MPI_Datatype dt = t == 0 ? MPI_INT : MPI_DOUBLE;
        const void* send = t == 0 ? (const void*)si : (const void*)sd;
        const double bytes = (double)n * size * (t == 0 ? sizeof(int) : sizeof(double));
        for (o = 0; o < 3; o++) {
          double t0, t1, t2;
          MPI_Barrier(MPI_COMM_WORLD);
          t0 = MPI_Wtime();
          tree_reduce(send, r_tree, acc, tmp, n, dt, ops[o], 0, rank, size);
          MPI_Barrier(MPI_COMM_WORLD);
          t1 = MPI_Wtime();
          MPI_Reduce(send, r_lib, n, dt, ops[o], 0, MPI_COMM_WORLD);
          MPI_Barrier(MPI_COMM_WORLD);
          t2 = MPI_Wtime();
          if (rank == 0) {
            printf("TREE %s %s %d %10.3lf\n", t == 0 ? "INT" : "DOUBLE", names[o], size, bytes / (t1 - t0) / 1073741824.0);
            printf("LIB %s %s %d %10.3lf\n", t == 0 ? "INT" : "DOUBLE", names[o], size, bytes / (t2 - t1) / 1073741824.0);
            if (t == 0 || ops[o] != MPI_SUM) {
              if (memcmp(r_tree, r_lib, (size_t)n * (t == 0 ? sizeof(int) : sizeof(double))) != 0) errors++;
            } else {
              for (i = 0; i < n; i++)
                if (fabs(((double*)r_tree)[i] - ((double*)r_lib)[i]) > 1e-9 * (1.0 + fabs(((double*)r_lib)[i]))) { errors++; break; }
            }
          }
        }
      }
    }
    if (rank == 0) printf("%s\n", errors ? "FAILED" : "PASSED");
    free(si); free(sd); free(r_tree); free(r_lib); free(acc); free(tmp);
  }
  MPI_Finalize();
  return errors ? 1 : 0;
}
