/*
 * b2r.h — public C API of b200reduce, a Blackwell-native (sm_100a) reduction library.
 *
 * Capability map against the reference (szabodabo/CUDA-MPI-Reductions):
 *   - b2r_reduce_local          <-> sumreduce/minreduce/maxreduce<T>   (cuda/C/src/reduction/reduction.h:15-25)
 *   - b2r_reduce / b2r_allreduce <-> MPI_Reduce(..., root 0, MPI_COMM_WORLD) (mpi/reduce.c:76,90)
 *   - b2r_reduce_scalar_sharded <-> simpleMultiGPU's per-GPU partial + host combine (simpleMultiGPU.cpp:212-293)
 *   - b2r_reduce_host / b2r_fill <-> sumreduceCPU/minreduceCPU/maxreduceCPU (reduction.cpp:214-249),
 *                                    MT19937 input generation (mpi/externalfunctions.h:45-179)
 *
 * All sizes are 64-bit element counts (the reference is limited to int).
 * All functions return a b2r_status_t; nothing calls exit() (reference: cutil_inline_runtime.h:267-283).
 * Streams are passed as void* (a cudaStream_t / CUstream); NULL is the legacy default stream.
 */
#ifndef B2R_H_
#define B2R_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2R_VERSION_MAJOR 0
#define B2R_VERSION_MINOR 1
#define B2R_MAX_RANKS 16

typedef enum {
  B2R_OK = 0,
  B2R_ERR_INVALID = 1,     /* bad argument (null pointer, misaligned element pointer, unknown enum) */
  B2R_ERR_CUDA = 2,        /* a CUDA runtime/driver call failed; see b2r_last_error_string() */
  B2R_ERR_UNSUPPORTED = 3, /* feature not available on this machine (no peer access, no multicast, ...) */
  B2R_ERR_TIMEOUT = 4,     /* a cross-GPU wait exceeded its deadline (peer died / fault injected) */
  B2R_ERR_NOMEM = 5,       /* symmetric heap / workspace exhausted */
  B2R_ERR_SYSTEM = 6,      /* socket / shm / fork failure in the bootstrap */
  B2R_ERR_PENDING = 7      /* loopback group: call recorded, launch happens when the last rank calls */
} b2r_status_t;

typedef enum { B2R_INT32 = 0, B2R_FLOAT32 = 1, B2R_FLOAT64 = 2, B2R_BFLOAT16 = 3 } b2r_dtype_t;
typedef enum { B2R_SUM = 0, B2R_MIN = 1, B2R_MAX = 2 } b2r_op_t;

/* Kernel variants of the single-GPU family (the reference's --kernel=N selected among SDK kernels
 * 0..6, reduction_kernel.cu:278-299; here N selects among *our* variants). */
typedef enum {
  B2R_VARIANT_AUTO = 0,
  B2R_VARIANT_LDG128 = 1, /* 128-bit ld.global.nc.L1::no_allocate, unrolled, register accumulate */
  B2R_VARIANT_LDG256 = 2, /* 256-bit ld.global (sm_100+) */
  B2R_VARIANT_TMA = 3     /* cp.async.bulk global->smem ring + mbarrier, producer warp / consumer warps */
} b2r_variant_t;

/* Collective algorithms. */
typedef enum {
  B2R_ALGO_AUTO = 0,
  B2R_ALGO_ONESHOT = 1, /* every rank (or root) loads all peers' vectors, combines in registers */
  B2R_ALGO_TWOSHOT = 2, /* reduce-scatter by P2P loads + gather by P2P stores */
  B2R_ALGO_TREE = 3,    /* hand-written binomial tree, log2(P) rounds of pairwise P2P combine */
  B2R_ALGO_NVLS = 4,    /* multimem.ld_reduce / multimem.st through the NVSwitch (needs multicast); fp32 / fp64 min and max
                           travel as order-preserving integer keys, encoded by one extra local pass */
  B2R_ALGO_PUSH = 5,    /* scatter by P2P stores into the owners' staging, reduce locally, push the result: no remote loads */
  B2R_ALGO_RING = 6,    /* reduce only: store-forwarding chain that ends at the root (one neighbour per GPU, no remote loads) */
  B2R_ALGO_LL = 7,      /* small messages (see b2r_comm_ll_max_bytes): flag-in-data packets pushed into the consumers' slots, no
                           barrier at either end; AUTO's choice up to 64 KiB (allreduce) / 256 KiB (reduce) */
  B2R_ALGO_POSTED = 8   /* reduce only, mid-size messages (see b2r_comm_posted_max_bytes): every rank stages its contribution in
                           its own heap and returns, the root reduces the staged copies through the switch; no barrier at either end,
                           fp min/max travel as order-preserving integer keys */
} b2r_algo_t;

typedef struct {
  int variant;     /* b2r_variant_t; 0 = auto */
  int threads;     /* threads per CTA; 0 = auto */
  int ctas;        /* total CTAs; 0 = auto (k x SM count, persistent) */
  int cluster;     /* thread-block-cluster size 1,2,4,8; 0 = auto */
  int stages;      /* TMA ring depth; 0 = auto */
  int tile_bytes;  /* TMA tile bytes (multiple of 16*32*consumer warps); 0 = auto */
  int unroll;      /* LDG loads in flight per thread: 4 or 8; 0 = auto */
  int pdl;         /* programmatic dependent launch (PDL) between consecutive kernels of the stream:
                        0  default: flavour 1, unless the environment says B2R_PDL=0 (then plain stream order);
                        1  B2R_PDL_SAFE: the launch may be scheduled while the previous kernel is still finishing, but it
                           waits for that kernel's completion (and the visibility of its writes) before its first load, so
                           it is correct whatever produced the input. Hides launch latency only;
                        2  B2R_PDL_INPUT_READY: the caller DECLARES that the input was complete before the previous kernel
                           in the stream started (e.g. back-to-back reductions of buffers filled long ago). The launch
                           streams its input while the previous kernel drains and waits only before it touches the
                           workspace and `out`. Never implied by the environment: reading an input that the previous
                           kernel is still writing would be a silent race;
                       <0  off, whatever the environment says. */
  int dyn_pct;     /* share (percent, 1..90) of the input that CTAs claim tile by tile at the end of the launch instead of by
                      the static split: evens out CTAs that started late. 0 = auto, <0 = static split only. Ignored by the
                      TMA variant, by clusters and by pdl = 2. Results do not depend on it for integers; for floating-point
                      types they are bit-reproducible for a given (count, alignment, geometry, dyn_pct). */
} b2r_launch_cfg_t;

#define B2R_PDL_OFF (-1)
#define B2R_PDL_SAFE 1
#define B2R_PDL_INPUT_READY 2

size_t b2r_dtype_size(b2r_dtype_t dtype);
const char* b2r_status_string(b2r_status_t s);
const char* b2r_last_error_string(void); /* thread-local detail of the last non-OK status */
const char* b2r_dtype_name(b2r_dtype_t d);
const char* b2r_op_name(b2r_op_t o);
const char* b2r_algo_name(b2r_algo_t a);
int b2r_version(void);

/* ------------------------------------------------------------------ host oracle + data generation */

/* CPU oracle. int32: two's-complement wrap-around sum (bit-exact target). fp32/fp64 sum: Kahan-Babuska
 * compensated in long double, rounded once. bf16: accumulated in double, rounded once to bf16 (RNE).
 * min/max propagate NaN (a NaN anywhere gives NaN), -0 < +0 is NOT distinguished. count==0 gives the
 * identity (0 / type max / type lowest; +inf / -inf for floating types). */
b2r_status_t b2r_reduce_host(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op);

/* Element-wise host combine: acc[i] = op(acc[i], in[i]); used as the oracle for reduce/allreduce. */
b2r_status_t b2r_combine_host(void* acc, const void* in, uint64_t count, b2r_dtype_t dtype, b2r_op_t op);

/* Counter-based generator (splitmix64 of (seed, index)); identical on host and device.
 * mode 0: "sdk"  — small values like the SDK sample (ints 0..255, fp (0..255)/2^31; reduction.cpp:698-705)
 * mode 1: "full" — full-range ints, fp in [0,1) with 53/24/8 bits (mpi/reduce.c:51-57)
 * mode 2: "signed" — fp in [-1,1), ints full range */
b2r_status_t b2r_fill_host(void* dst, uint64_t count, b2r_dtype_t dtype, uint64_t seed, uint64_t first_index, int mode);
b2r_status_t b2r_fill_device(void* dst, uint64_t count, b2r_dtype_t dtype, uint64_t seed, uint64_t first_index, int mode,
                             void* stream);

/* ------------------------------------------------------------------ single-GPU reduction */

typedef struct b2r_workspace* b2r_workspace_t;

/* A workspace holds the per-launch scratch (cluster partials, arrival ticket, optional trace stamps).
 * One workspace must not be used by two launches that can overlap on the GPU (use one per stream). */
b2r_status_t b2r_workspace_create(b2r_workspace_t* ws, int device);
b2r_status_t b2r_workspace_destroy(b2r_workspace_t ws);

/* out[0] = op over in[0..count). `in` and `out` are device pointers of `dtype` (element-aligned; any
 * element offset is fine, vector alignment is handled inside). One kernel launch, no host sync.
 * Deterministic: the same (count, pointer alignment, launch geometry) gives bit-identical results. */
b2r_status_t b2r_reduce_local(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                              b2r_workspace_t ws, void* stream);
b2r_status_t b2r_reduce_local_ex(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                                 b2r_workspace_t ws, void* stream, const b2r_launch_cfg_t* cfg);

/* The same reduction for callers that cannot allocate device memory. (No counterpart in the reference: its sample
 * cudaMallocs its block-partial buffer in the driver, reduction.cpp:721-722.) Who needs it: a process whose device
 * memory is entirely owned by another allocator, a latency-critical loop that may not run anything that synchronises the
 * device (cudaMalloc / cudaFree do), or a hook / callback whose contract forbids allocation calls. Instead of a
 * b2r_workspace the kernel's scratch (block partials + ticket) is a static 8448-byte __device__ array inside libb2r.so,
 * resident on every GPU that loads the library until the process exits.
 *  - Limit: at most B2R_NOALLOC_MAX_PARTIALS block partials fit; a geometry that needs more (cfg->ctas / cluster above
 *    it) returns B2R_ERR_INVALID. With cfg->ctas = 0 (or cfg = NULL) the library's own choice is capped at the limit, so
 *    every count is served.
 *  - This call makes no cudaMalloc / cudaFree / allocating library call. It enqueues a 4-byte memset of the ticket (the
 *    only scratch word that must start at zero; the array otherwise keeps the previous call's contents) and the kernel.
 *  - All calls on one GPU share the array: calls that overlap on the GPU race. Keep them on one stream.
 *  - The first call on each device does a one-time set-up with that device current (looks up the array's device
 *    address); issue it before any stream capture begins.
 *  - Launches on the CURRENT device; variants ldg128 / ldg256 only; cfg may be NULL. */
#define B2R_NOALLOC_MAX_PARTIALS 1024
b2r_status_t b2r_reduce_local_noalloc(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, void* stream,
                                      const b2r_launch_cfg_t* cfg);

/* Transform-reduce (vendored scalarProd analogue, scalarProd_kernel.cu:33-80): out[0] = sum_i a[i]*b[i].
 * fp32/fp64/bf16 (bf16 accumulates in fp32) and int32 (wrap). */
b2r_status_t b2r_dot_local(const void* a, const void* b, void* out, uint64_t count, b2r_dtype_t dtype,
                           b2r_workspace_t ws, void* stream);

/* The launch geometry the dispatcher would pick (for logging / tests). Re-reads the B2R_* geometry variables first. */
b2r_status_t b2r_plan_local(uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int device, const b2r_launch_cfg_t* hint,
                            b2r_launch_cfg_t* chosen);

/* The single-GPU launch path reads B2R_VARIANT / _THREADS / _CTAS_PER_SM / _CLUSTER / _UNROLL / _STAGES / _PDL once, at the
 * first launch (seven getenv calls cost 2-3 us per launch, a third of a 64 MiB reduction). A program that changes them at
 * run time calls this (or b2r_plan_local) afterwards. */
void b2r_refresh_env(void);

/* In-kernel phase stamps (vendored `clock` sample analogue, clock_kernel.cu:18-54). When enabled every CTA
 * records %globaltimer at start / after streaming loop / after block+cluster combine / at exit. */
b2r_status_t b2r_workspace_trace_enable(b2r_workspace_t ws, int enable);
/* Copies stamps of the last launch: stamps[4*cta + phase] in ns; returns number of CTAs in *n_ctas. */
b2r_status_t b2r_workspace_trace_read(b2r_workspace_t ws, uint64_t* stamps, int max_ctas, int* n_ctas);
/* The stamps of the launch `launches_back` traced launches ago (0 = the last one; a ring of 8 launches is kept), so the
 * period between consecutive launches and the overlap of one launch with the next can be read off, not only one launch. */
b2r_status_t b2r_workspace_trace_read_at(b2r_workspace_t ws, int launches_back, uint64_t* stamps, int max_ctas, int* n_ctas);

/* Pipelined host->device + reduce: `host_in` is (ideally pinned) host memory; it is streamed to the GPU in
 * chunks that are reduced while the next chunk is in flight; out_host receives the scalar. Synchronous. */
b2r_status_t b2r_reduce_from_host(const void* host_in, void* out_host, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                                  b2r_workspace_t ws, void* stream);

/* ------------------------------------------------------------------ guard-page buffers (out-of-bounds detector)
 * The reference's only memory checker is cutil's emulation-mode bank checker (bank_checker.h:25-60); where compute-sanitizer
 * cannot be used, the MMU can: b2r_guard_alloc returns `bytes` of device memory placed so that the byte right after the
 * buffer (at_end = 1) or right before it (at_end = 0) is reserved, UNMAPPED address space. A kernel that strays past that
 * edge faults (cudaErrorIllegalAddress) instead of quietly touching a neighbour. Needs the VMM driver API. */
b2r_status_t b2r_guard_alloc(uint64_t bytes, int device, int at_end, void** ptr);
b2r_status_t b2r_guard_free(void* ptr);

/* ------------------------------------------------------------------ communicator */

typedef struct b2r_comm* b2r_comm_t;

typedef struct {
  uint64_t heap_bytes;   /* symmetric heap per rank (0 = default 256 MiB + control region) */
  int use_multicast;     /* -1 auto (use if supported), 0 never, 1 require */
  int timeout_ms;        /* cross-GPU wait deadline, 0 = default (10000), <0 = never time out */
  int max_ctas;          /* CTAs per collective kernel (0 = auto) */
  int symmetric_only;    /* 1 = every buffer passed to a collective comes from b2r_symm_alloc on every rank: messages are
                            never staged nor chunked (one launch per collective whatever the size). All ranks must agree. */
  int reserved[3];
} b2r_comm_config_t;

/* Single process, several devices (vendored MonteCarloMultiGPU "one thread, one stream per GPU" mode,
 * MonteCarloMultiGPU.cpp:244-300). comms[i] drives devices[i]. Calls on different comms never block the host,
 * so one thread may issue the same collective on every comm in turn. */
b2r_status_t b2r_comm_init_all(b2r_comm_t* comms, int ndev, const int* devices, const b2r_comm_config_t* cfg);

/* One process per GPU (mpi/reduce.c style). `bootstrap_id` names an abstract Unix socket on which rank 0
 * listens; VMM handles travel as file descriptors (SCM_RIGHTS). Every rank passes the same id. */
b2r_status_t b2r_comm_init_rank(b2r_comm_t* comm, int nranks, int rank, int device, const char* bootstrap_id,
                                const b2r_comm_config_t* cfg);

/* Loopback group: `nranks` virtual ranks on ONE device. Collective calls are recorded per rank and fused into a
 * single cooperative launch (gridDim.y = rank) when the last rank of the group has called; earlier callers get
 * B2R_ERR_PENDING. Exists so the cross-rank protocol can be tested on a one-GPU box. */
b2r_status_t b2r_comm_init_loopback(b2r_comm_t* comms, int nranks, int device, const b2r_comm_config_t* cfg);

b2r_status_t b2r_comm_destroy(b2r_comm_t comm);
b2r_status_t b2r_comm_rank(b2r_comm_t comm, int* rank);
b2r_status_t b2r_comm_size(b2r_comm_t comm, int* nranks);
b2r_status_t b2r_comm_device(b2r_comm_t comm, int* device);
b2r_status_t b2r_comm_has_multicast(b2r_comm_t comm, int* yes);
/* The workspace the communicator's scalar-sharded reductions run on (owned by the communicator; for tracing). */
b2r_status_t b2r_comm_workspace(b2r_comm_t comm, b2r_workspace_t* ws);

/* Symmetric allocation: every rank must call with the same size in the same order. Buffers obtained here are
 * addressable by every peer (zero-copy collectives); other device pointers are staged through the heap. */
b2r_status_t b2r_symm_alloc(b2r_comm_t comm, uint64_t bytes, void** ptr);
b2r_status_t b2r_symm_free_all(b2r_comm_t comm); /* resets the bump allocator */
/* Address of `local_ptr` (inside the symmetric heap) as mapped for `peer` in this process. */
b2r_status_t b2r_symm_peer_ptr(b2r_comm_t comm, const void* local_ptr, int peer, void** peer_ptr);

/* Reads (and clears) the device-side error word: B2R_OK, or B2R_ERR_TIMEOUT if a wait hit its deadline.
 * Synchronises `stream`. */
b2r_status_t b2r_comm_check(b2r_comm_t comm, void* stream);

/* ------------------------------------------------------------------ collectives (MPI_Reduce semantics)
 * recvbuf[j] = op over ranks r of sendbuf_r[j], j < count. b2r_reduce: valid at `root` only (recvbuf may be
 * NULL elsewhere). sendbuf is not modified unless sendbuf == recvbuf (in place). All ranks call in the same
 * order with the same count/dtype/op/root/algo. fp sums are deterministic on the P2P paths (bit-reproducible run to run):
 * ascending rank order for one-shot / two-shot / push / LL, chain order (farthest rank first) for the ring, pairwise order
 * rotated by the root for the tree; the NVLS and posted paths' order is defined by the switch (integers and min / max are exact
 * on every path).
 * Stream capture: the element-wise collectives number their calls on the HOST (barrier epochs, LL / posted call numbers are kernel
 * arguments), so a captured graph that contains one must not be replayed — the replay would reuse call numbers its peers have
 * already seen. Only b2r_reduce_scalar_sharded keeps its call number on the device and replays correctly. */
b2r_status_t b2r_reduce(const void* sendbuf, void* recvbuf, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int root,
                        b2r_comm_t comm, void* stream);
b2r_status_t b2r_allreduce(const void* sendbuf, void* recvbuf, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                           b2r_comm_t comm, void* stream);
b2r_status_t b2r_reduce_ex(const void* sendbuf, void* recvbuf, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                           int root, b2r_comm_t comm, void* stream, b2r_algo_t algo);
b2r_status_t b2r_allreduce_ex(const void* sendbuf, void* recvbuf, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                              b2r_comm_t comm, void* stream, b2r_algo_t algo);

/* Sharded scalar reduce: every rank owns `count` local elements; out[0] at root = op over ALL ranks' elements.
 * One kernel per rank: the streaming reduction, then the rank's last CTA stores its partial into root's mailbox
 * over NVLink and root's last CTA combines the P partials in rank order. `all` != 0 delivers the scalar to
 * every rank (scalar allreduce). */
b2r_status_t b2r_reduce_scalar_sharded(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                                       int root, int all, b2r_comm_t comm, void* stream);

/* Same with an explicit launch configuration for the local streaming part (variant, threads, ctas, pdl, ...). */
b2r_status_t b2r_reduce_scalar_sharded_ex(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                                          int root, int all, b2r_comm_t comm, void* stream, const b2r_launch_cfg_t* cfg);

/* End-to-end flavour (vendored simpleMultiGPU's H2D + kernel + D2H + host loop, simpleMultiGPU.cpp:227-293, without the host
 * loop): this rank's shard is `count` elements of (pinned) HOST memory. It is copied to the GPU in chunks by the copy engine
 * while earlier chunks are reduced; one more fused launch folds the chunk scalars and combines them across the GPUs over
 * NVLink; the scalar is written to out_host (an element of `dtype`) on root, or on every rank with `all`. Synchronous. */
b2r_status_t b2r_reduce_scalar_sharded_from_host(const void* host_in, void* out_host, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                                                 int root, int all, b2r_comm_t comm, void* stream);

b2r_status_t b2r_barrier(b2r_comm_t comm, void* stream);

/* Largest message (bytes per rank) the LL path accepts on a communicator of `nranks` ranks; 0 if nranks is out of range. */
uint64_t b2r_ll_max_bytes(int nranks);
/* The same for an existing communicator: larger with the 64-CTA LL region behind the user heap (B2R_LL_BIG_MB at creation,
 * default 16 MiB; 0 = none). */
b2r_status_t b2r_comm_ll_max_bytes(b2r_comm_t comm, uint64_t* bytes);
/* Largest message (bytes per rank) B2R_ALGO_POSTED accepts on this communicator: a quarter of the work region (B2R_WORK_MB at
 * creation, default 320), at most 64 MiB. */
b2r_status_t b2r_comm_posted_max_bytes(b2r_comm_t comm, uint64_t* bytes);

/* Which algorithm AUTO resolves to (pure function; used by tests and the harness). */
b2r_algo_t b2r_select_algo(int is_allreduce, uint64_t bytes, b2r_dtype_t dtype, b2r_op_t op, int nranks,
                           int has_multicast);

/* ------------------------------------------------------------------ host (CPU) collectives
 * The same algorithms executed by CPU threads over ordinary memory: the "host loopback transport" used for
 * protocol tests without a GPU and as the engine of the bundled mini-MPI (MPI_Reduce / custom tree on CPU).
 * bufs[r] is rank r's send buffer; out receives the result (root's recvbuf). */
b2r_status_t b2r_host_reduce_threads(const void* const* sendbufs, void* recvbuf, uint64_t count, b2r_dtype_t dtype,
                                     b2r_op_t op, int nranks, int root, b2r_algo_t algo);

/* Multi-PROCESS flavour of the host transport: every process opens the same POSIX shared-memory segment (rank 0 creates it
 * and removes it on destroy) and the collectives run the same algorithms with a process-shared barrier. Used by the Python
 * HostCommunicator (CPU tensors under torch.distributed/gloo) and, through the same class, by the bundled mini-MPI.
 * root >= 0: reduce to root; root == -1: allreduce. `shm_name` must start with '/' and be the same on every rank. */
typedef struct b2r_host_comm* b2r_host_comm_t;
b2r_status_t b2r_host_comm_init(b2r_host_comm_t* comm, const char* shm_name, int rank, int nranks, uint64_t window_bytes);
b2r_status_t b2r_host_comm_destroy(b2r_host_comm_t comm);
b2r_status_t b2r_host_comm_reduce(b2r_host_comm_t comm, const void* sendbuf, void* recvbuf, uint64_t count, b2r_dtype_t dtype,
                                  b2r_op_t op, int root, b2r_algo_t algo);
b2r_status_t b2r_host_comm_barrier(b2r_host_comm_t comm);

#ifdef __cplusplus
}
#endif
#endif /* B2R_H_ */
