/*
 * b200validate.h — C ABI of libb200validate.so, the B200-native (sm_100a) replacement for the
 * CUDA side of the GPU Operator's cuda-vectoradd node validation.
 *
 * This is the drop-in boundary: what a cgo (`import "C"`), ctypes or JNI binding of the validator
 * would bind.  Plain C types only; no C++ exceptions, no exit(), no callbacks cross it.
 *
 * Provenance of the "replaces" notes.  The reference repository is a single README; the code on
 * this path lives in the chart it installs with no version pin (reference README.md:96,101-110):
 * NVIDIA/cuda-samples `vectorAdd.cu` and NVIDIA/gpu-operator `validator/main.go`.  Neither is on
 * disk here, so upstream units are named by SURVEY.md §8(a) row (a3..a10) rather than file:line;
 * the README lines that reach them are cited where one exists.
 *
 * Return convention (SURVEY.md §8b "Errors"): 0 = OK; > 0 = a cudaError_t value;
 * < 0 = one of the B200V_E_* library codes below.  b200v_strerror() names any of them.
 *
 * Ownership: the caller owns every host pointer and out-struct it passes; `_device` entry points
 * take device pointers the caller owns; everything else on the device (buffers, streams, events,
 * NCCL communicators) belongs to the library and is released before the call returns or by
 * b200v_shutdown().  Every entry point selects its device itself (cudaSetDevice) and is
 * re-entrant across different devices.
 */
#ifndef B200VALIDATE_H
#define B200VALIDATE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200V_VERSION_MAJOR 0
#define B200V_VERSION_MINOR 1

/* Library error codes (negative; positive values are cudaError_t). */
#define B200V_OK            0
#define B200V_E_BADARG     (-1)  /* null pointer, bad op, bad size */
#define B200V_E_MISMATCH   (-2)  /* device result differs from the check; index in the out-struct */
#define B200V_E_NCCL       (-3)  /* NCCL failed; ncclResult_t in the out-struct / last_nccl_error */
#define B200V_E_NODEVICE   (-4)  /* no CUDA device visible */
#define B200V_E_NONCCL     (-5)  /* libnccl.so.2 could not be loaded */
#define B200V_E_NOMEM      (-6)  /* host allocation failed */
#define B200V_E_TOOLARGE   (-7)  /* n exceeds the fixed capacity of a no-allocation entry point */
#define B200V_E_UNSUPPORTED (-8) /* the devices or the driver lack a required feature (e.g. multicast) */
#define B200V_E_DRIVER     (-9)  /* a CUDA driver API call failed; CUresult in b200v_last_driver_error */
#define B200V_E_TIMEOUT    (-10) /* a cross-GPU barrier did not complete in time */

/* Elementwise ops.  dst = f(x, y, s), fp32, round-to-nearest-even, denormals kept (no -ftz):
 *   COPY dst=x   SCALE dst=s*x   ADD dst=x+y   TRIAD dst=fmaf(s,y,x) (one FFMA)
 *   VECTORADD dst=x+y+0.0f (the sample's kernel)
 *   TRIAD_UNFUSED dst=x+(s*y) with two roundings. */
#define B200V_OP_COPY          0
#define B200V_OP_SCALE         1
#define B200V_OP_ADD           2
#define B200V_OP_TRIAD         3
#define B200V_OP_VECTORADD     4
#define B200V_OP_TRIAD_UNFUSED 5
#define B200V_OP_COUNT         6

/* b200v_stream fills y from seed + this, so y is never a shifted copy of x (n <= 2^32). */
#define B200V_SEED_Y_OFFSET 0x100000000ull

typedef struct b200v_info {
    char     name[256];
    int      cc_major, cc_minor;
    int      sm_count;
    int      mem_clock_khz, mem_bus_bits;   /* nominal HBM GB/s = 2*clock*bits/8 */
    int      sm_clock_khz;
    uint64_t l2_bytes;
    uint64_t hbm_bytes;
} b200v_info;

typedef struct b200v_timing {
    float    h2d_ms;      /* first H2D enqueue -> last H2D done */
    float    kernel_ms;   /* sum of kernel durations */
    float    d2h_ms;      /* first D2H enqueue -> last D2H done */
    float    total_ms;    /* whole call on the device timeline */
    int      grid, block; /* launch shape of the (first) kernel */
    int      chunks;      /* pipeline chunks used for the transfers */
    int      kernel_launches;
} b200v_timing;

typedef struct b200v_stream_result {
    float    best_ms, median_ms, mean_ms;
    double   best_gbs, median_gbs;     /* algorithmic bytes / time */
    uint64_t bytes_per_launch;         /* algorithmic bytes one launch moves */
    uint64_t digest;                   /* b200v digest of the op's output array */
    uint64_t mismatches;               /* device re-check of the output against a scalar kernel */
    uint64_t first_mismatch;           /* element index, valid when mismatches > 0 */
    int      l2_resident;              /* 1 when the op's working set fits the device L2 */
    int      grid, block;
    int      kernel_launches;          /* timed launches of the op kernel: iters x the batch size
                                          (small arrays are timed as batches of back-to-back
                                          launches lasting >= ~200 us; times are per launch) */
} b200v_stream_result;

typedef struct b200v_ar_row {
    uint64_t bytes;          /* message bytes per rank */
    float    best_ms, median_ms;   /* max over ranks of the per-rank device time */
    double   algbw_gbs;      /* bytes / best time */
    double   busbw_gbs;      /* algbw * 2(n-1)/n */
    uint64_t mismatches;     /* elements differing from the closed form, over the check before the
                                timed launches and the check after them */
    int      nccl_result;    /* ncclResult_t of the failing call, 0 otherwise */
    int      algo;           /* hand-written probe: the kernel that ran, B200V_AR_ALGO_*; 0 for NCCL */
} b200v_ar_row;

#define B200V_AR_ALGO_MULTIMEM 1   /* two-shot, reduction inside the NVSwitch */
#define B200V_AR_ALGO_P2P      2   /* peer loads + peer stores */
#define B200V_AR_ALGO_PUSH     3   /* two phases of peer stores */
#define B200V_AR_ALGO_ONESHOT  4   /* every GPU reduces the whole (small) vector itself */
#define B200V_AR_ALGO_HYBRID   5   /* one grid: part of the vector through the switch, the rest by peer stores */

/* ---- library / device queries ----------------------------------------------------------------- */

/* No upstream counterpart (new with the C ABI). */
void        b200v_version(int *major, int *minor);
const char *b200v_strerror(int code);   /* static string; never NULL */

/* Replaces cudaGetDeviceCount / the nvidia-smi listing the runbook checks by hand
 * (reference README.md:152, 158-168). */
int b200v_device_count(int *n);
int b200v_device_info(int dev, b200v_info *out);

/* Pinned host memory for callers that want the fast transfer path (the sample uses malloc;
 * SURVEY.md §8 a5).  No upstream counterpart. */
int b200v_host_alloc(void **ptr, size_t bytes);
int b200v_host_free(void *ptr);
/* The same with a hint.  B200V_HOST_WRITE_ONLY: the CPU only ever writes this buffer (the input
 * vectors), so it may be write-combined memory when the library's host-memory policy
 * (B200V_HOST_POLICY=pinned|wc|thp) says that transfers faster; reading such a buffer from the
 * CPU is very slow.  Free with b200v_host_free.  No upstream counterpart. */
#define B200V_HOST_WRITE_ONLY 1
int b200v_host_alloc_ex(void **ptr, size_t bytes, int flags);

/* ---- the validation workload -------------------------------------------------------------------- */

/* Replaces the body of the sample's main() between host init and the host verification loop
 * (SURVEY.md §8 a3 + a5): cudaMalloc x3, H2D x2, vectorAdd<<<>>>, D2H, cudaFree x3 — deployed by
 * reference README.md:101-110.  hA, hB, hC are HOST pointers (pinned or pageable), n elements
 * each.  hC[i] = hA[i] + hB[i] + 0.0f.  Transfers are chunked and overlapped with the kernel on
 * three streams.  n == 0 is OK and a no-op.  timing may be NULL. */
int b200v_vectoradd(int dev, const float *hA, const float *hB, float *hC, size_t n,
                    b200v_timing *timing);

/* Has no counterpart in the reference or in the sample it deploys.  The same computation as
 * b200v_vectoradd (hC[i] = hA[i] + hB[i] + 0.0f on host buffers) for callers that cannot
 * allocate device memory: a validator re-run as a health probe on a GPU that is already serving a
 * tenant, where HBM may be fully committed (cudaMalloc would fail and a healthy GPU would be
 * reported broken) and where cudaMalloc/cudaFree would stall the tenant by synchronising the
 * device.  It calls no allocator, creates no stream or event, and stages through a fixed array
 * inside the library: 3 x B200V_NOALLOC_MAX_ELEMS floats (12 MiB) that every GPU which loads the
 * library holds until the process exits.
 *   - n > B200V_NOALLOC_MAX_ELEMS returns B200V_E_TOOLARGE and computes nothing.
 *   - `stream` is the caller's cudaStream_t (NULL = default stream); the call enqueues two H2D
 *     copies, the kernel and one D2H copy on it and returns after synchronising it.
 *   - The first call on each GPU also performs that GPU's one-time set-up (looking up the
 *     array's device address); make it before starting any stream capture.  CUDA loads modules
 *     lazily, so that first call is also when the 12 MiB are reserved on the GPU (and when the
 *     context is created, if the process had none): on a GPU that is already full it can fail
 *     with cudaErrorMemoryAllocation.  Make it when the probe process starts, before the tenant
 *     fills HBM (or start the process with CUDA_MODULE_LOADING=EAGER); later calls reserve
 *     nothing.
 *   - On a CUDA error after the first copy was enqueued the call waits for the stream before
 *     returning, so the caller's host buffers are never still in use when it returns.
 *   - All calls on one GPU share the array, which still holds the previous call's data: calls
 *     that overlap in time on the same GPU race with each other, so serialise them.  No clearing
 *     is needed between calls (every element read was written earlier in the same call). */
#define B200V_NOALLOC_MAX_ELEMS (1u << 20)
int b200v_vectoradd_noalloc(int dev, const float *hA, const float *hB, float *hC, size_t n,
                            void *stream);

/* Same contract for any B200V_OP_* on host buffers: hDst = op(hX, hY, s).  hY is ignored (may be
 * NULL) for COPY and SCALE.  New (STREAM sweep). */
int b200v_stream_host(int dev, int op, float *hDst, const float *hX, const float *hY, float s,
                      size_t n, b200v_timing *timing);

/* The kernel alone on DEVICE pointers the caller owns, enqueued on `stream` (a cudaStream_t cast
 * to void*; NULL = the default stream) and not synchronised.  Replaces the
 * `vectorAdd<<<blocksPerGrid, threadsPerBlock>>>(d_A, d_B, d_C, numElements)` launch
 * (SURVEY.md §8 a3).  Any alignment and any n are accepted. */
int b200v_vectoradd_device(int dev, const float *dA, const float *dB, float *dC, size_t n,
                           void *stream);
int b200v_stream_device(int dev, int op, float *dDst, const float *dX, const float *dY, float s,
                        size_t n, void *stream);

/* Which of the compiled kernels b200v_stream_device would launch for this op, these device
 * pointers (their alignment and overlap decide) and this n: its name is written to buf, e.g.
 * "b200v::ew_tile_kernel<VECTORADD,VEC=8,UNROLL=1,BLOCK=1024> (t8u1b1024)".  Launches nothing.
 * For reports that must name the kernel that actually ran.  No upstream counterpart. */
int b200v_kernel_name(int op, const float *dDst, const float *dX, const float *dY, size_t n,
                      char *buf, size_t buflen);

/* The sample's own launch shape — one element per thread, ceil(n/256) blocks of 256 threads,
 * bounds-checked — kept as the in-library cross-check for the tuned kernels.
 * n must fit the 2^31-1 block limit times 256. */
int b200v_vectoradd_device_sampleshape(int dev, const float *dA, const float *dB, float *dC,
                                       size_t n, void *stream);

/* Counter-based input generator on the device: X[i] = (splitmix64(seed+i) >> 40) * 2^-24
 * (SURVEY.md §8d cfg1).  Replaces the sample's rand() host init for arrays too
 * large to fill on one host core. */
int b200v_fill_device(int dev, float *dX, size_t n, uint64_t seed, void *stream);

/* Digest of n 32-bit words on the device: sum_i mix64(w[i] + GOLDEN*(i+1)) mod 2^64 — order
 * independent, position sensitive.  Allocates an 8-byte device accumulator for the call.
 * Synchronises `stream` before returning.  New (SURVEY.md §8c "memcmp + digest"). */
int b200v_digest_device(int dev, const void *dWords, size_t n, uint64_t *digest, void *stream);

/* Device-side bitwise check of dDst against op(dX, dY, s) recomputed by the scalar sample-shape
 * kernel: counts differing elements and reports the lowest differing index.  Allocates 16 bytes
 * of device scratch for the call and synchronises `stream`.  Returns B200V_E_MISMATCH when any
 * element differs.  Device analogue of the sample's verification loop. */
int b200v_check_device(int dev, int op, const float *dDst, const float *dX, const float *dY,
                       float s, size_t n, uint64_t *mismatches, uint64_t *first_mismatch,
                       void *stream);

/* One row of the STREAM sweep, self-contained: allocates three n-element arrays on `dev`, fills
 * them from `seed` (x: seed, y: seed + B200V_SEED_Y_OFFSET), runs `warmup` untimed and `iters` timed launches of `op`
 * with CUDA events, checks and digests the output, frees everything.  New (SURVEY.md §8 a9,
 * §8d cfg2). */
int b200v_stream(int dev, int op, size_t n, float s, uint64_t seed, int warmup, int iters,
                 b200v_stream_result *result);

/* ---- NVLink health probe (the only NCCL use) --------------------------- */

/* Single process, one communicator per listed device (ncclCommInitAll).  In-place fp32 sum
 * all-reduce, message sizes min_bytes, 2*min_bytes, ... <= max_bytes; rank r contributes the
 * integer-valued (float)((i*(2r+1) + r) & 15), so the sum is exact in any order and is checked
 * bit for bit against the closed form on the device, once before the timed launches and once
 * more (re-filled) after them.  Small messages are timed as batches of back-to-back launches.
 * rows must hold max_rows entries; *nrows receives the count written. */
int b200v_allreduce_probe(int ndev, const int *devs, size_t min_bytes, size_t max_bytes,
                          int warmup, int iters, b200v_ar_row *rows, int max_rows, int *nrows);

/* One-process-per-GPU form.  Rank 0 calls b200v_nccl_unique_id, the launcher distributes the 128
 * bytes, every rank calls b200v_allreduce_probe_rank (collective).  Times are this rank's; the
 * caller takes the max over ranks. */
#define B200V_NCCL_UNIQUE_ID_BYTES 128
int b200v_nccl_unique_id(unsigned char id[B200V_NCCL_UNIQUE_ID_BYTES]);
int b200v_allreduce_probe_rank(int dev, int rank, int nranks,
                               const unsigned char id[B200V_NCCL_UNIQUE_ID_BYTES],
                               size_t min_bytes, size_t max_bytes, int warmup, int iters,
                               b200v_ar_row *rows, int max_rows, int *nrows);
int b200v_nccl_version(int *version);
int b200v_last_nccl_error(void);

/* A second, hand-written NVLink probe (nothing upstream): the same sweep, input
 * pattern, bit-exact check and row format as b200v_allreduce_probe, but the all-reduce is one
 * sm_100a kernel per GPU over a CUDA multicast object and peer mappings.  Which kernel runs
 * depends on the message size and the GPU count (row.algo says which; b200v_nvls_algo_for
 * answers without a GPU): one-shot multimem.ld_reduce for small messages (up to 2 MiB / 1 MiB /
 * 512 KiB at 2 / 4 / 8 GPUs), two phases of peer stores for larger messages on 2 GPUs,
 * multimem.ld_reduce + multimem.st (reduction inside the NVSwitch) from 4 GPUs;
 * B200V_NVLS_ALGO=multimem|p2p|push|oneshot|hybrid forces one.  Every size is checked twice, before the
 * timed launches and again after them.  Single process, one host thread per GPU, 2 <= ndev
 * distinct devices (a multicast object cannot be made for one), sizes are multiples of 16 bytes.  B200V_E_UNSUPPORTED when the GPUs or the driver have no multicast support,
 * B200V_E_TIMEOUT if a peer GPU never reaches a barrier (bounded spin, default 2 s). */
int b200v_nvls_allreduce_probe(int ndev, const int *devs, size_t min_bytes, size_t max_bytes,
                               int warmup, int iters, b200v_ar_row *rows, int max_rows, int *nrows);
int b200v_last_driver_error(void);
/* Which kernel (B200V_AR_ALGO_*) the hand-written all-reduce uses for a message of `bytes` bytes
 * over `nranks` GPUs, given the environment (B200V_NVLS_ALGO, B200V_NVLS_ONESHOT_MAX).  Host logic
 * only: touches no device, so it also runs where there is none. */
int b200v_nvls_algo_for(int nranks, size_t bytes);

/* One-process-per-GPU form of the same probe.  Every rank calls it with the same `rendezvous`
 * string — the name of an abstract-namespace Unix socket on which rank 0 hands the multicast
 * object's file descriptor to the other ranks of this host (and, for the 2-rank peer-to-peer
 * kernel, relays every rank's buffer descriptor to every other rank).  Collective; the
 * times are this rank's and the caller takes the max over ranks, as with
 * b200v_allreduce_probe_rank.  A rank that never shows up yields B200V_E_TIMEOUT after 60 s. */
int b200v_nvls_allreduce_probe_rank(int dev, int rank, int nranks, const char *rendezvous,
                                    size_t min_bytes, size_t max_bytes, int warmup, int iters,
                                    b200v_ar_row *rows, int max_rows, int *nrows);

/* ---- all-reduce of a CALLER-OWNED device buffer ------------------------
 * The probes above fill, reduce and check buffers of their own.  These entry points reduce a
 * buffer the caller filled and will read back, so a checker outside the library (tests/,
 * bench.py: the CPU oracle's closed form) can compare the reduced bytes.  One process (or host
 * thread) per GPU; create and destroy are collective; every rank must issue the same sequence of
 * all-reduce calls with the same counts.  In-place fp32 sum over `count` floats at dBuf (device
 * memory on the communicator's GPU), enqueued on `stream` (cudaStream_t, NULL = default stream)
 * and not synchronised.  No upstream counterpart. */

/* NCCL: ncclCommInitRank with the id from b200v_nccl_unique_id, then ncclAllReduce on dBuf. */
typedef struct b200v_ar_comm b200v_ar_comm;
int b200v_ar_comm_create(int dev, int rank, int nranks,
                         const unsigned char id[B200V_NCCL_UNIQUE_ID_BYTES], b200v_ar_comm **comm);
int b200v_ar_comm_allreduce(b200v_ar_comm *comm, float *dBuf, size_t count, void *stream);
int b200v_ar_comm_destroy(b200v_ar_comm *comm);

/* The hand-written kernels.  They work on the library's own NVLink-mapped allocation (sized for
 * max_bytes at create; 2 <= nranks), so each call copies dBuf in, reduces, and copies the result back, all on
 * `stream`; any count with 4 * count <= max_bytes is accepted (the tail is zero-padded to a whole
 * 16-byte vector), else B200V_E_TOOLARGE.  The two-phase kernels add in rank order, so the
 * result equals the sequential sum r = 0..n-1 bit for bit on every rank; the switch-side
 * reduction's order is the hardware's.  b200v_nvls_comm_check synchronises the device and returns
 * B200V_E_TIMEOUT if any kernel gave up waiting for a peer. */
typedef struct b200v_nvls_comm b200v_nvls_comm;
int b200v_nvls_comm_create(int dev, int rank, int nranks, const char *rendezvous, size_t max_bytes,
                           b200v_nvls_comm **comm);
int b200v_nvls_comm_allreduce(b200v_nvls_comm *comm, float *dBuf, size_t count, void *stream);
int b200v_nvls_comm_check(b200v_nvls_comm *comm);
int b200v_nvls_comm_destroy(b200v_nvls_comm *comm);

/* Releases cached per-device state (streams, events, staging buffers). */
int b200v_shutdown(void);

#ifdef __cplusplus
}
#endif
#endif /* B200VALIDATE_H */
