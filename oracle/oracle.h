/*
 * oracle.h — CPU restatement of the cuda-vectoradd validation path.  TEST INFRASTRUCTURE ONLY.
 *
 * Nothing under oracle/ is part of the product: only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library, and there only as the
 * checker (or as the timed CPU baseline), never as the thing shipped.
 *
 * PARITY UNPINNED BY THE REFERENCE: /root/reference is one README (no code, no tests, no golden
 * vectors).  The algorithm restated here is the NVIDIA/cuda-samples
 * `vectorAdd` sample that the GPU Operator chart installed at reference README.md:101-110
 * deploys as its validator workload (chart version unpinned, README.md:96,101).  That sample is
 * NOT on this machine, so it is restated from its published behaviour; the only pin is the
 * self-derived regression vector in SURVEY.md §8(c) / BASELINE.md §5 (tests/golden/).
 */
#ifndef B200V_ORACLE_H
#define B200V_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* STREAM op codes — same numbering as include/b200validate.h (B200V_OP_*). */
enum { ORACLE_OP_COPY = 0, ORACLE_OP_SCALE = 1, ORACLE_OP_ADD = 2, ORACLE_OP_TRIAD = 3,
       ORACLE_OP_VECTORADD = 4, ORACLE_OP_TRIAD_UNFUSED = 5 };

/* Sample main() host init: glibc rand() stream at its default seed (1),
 * drawn interleaved A[i], B[i].  srand(1) is called first so repeated calls in one process
 * reproduce what a fresh process sees. */
void oracle_fill_rand(float *A, float *B, size_t n);

/* Counter-based generator for the large configs:
 * X[i] = (splitmix64(seed + i) >> 40) * 2^-24, in [0,1). */
void oracle_fill_hash(float *X, size_t n, uint64_t seed);
uint64_t oracle_splitmix64(uint64_t x);

/* Device kernel semantics: C[i] = A[i] + B[i] + 0.0f. */
void oracle_vadd(const float *A, const float *B, float *C, size_t n);

/* Sample main() host verification loop:
 * fabs(A[i] + B[i] - C[i]) > 1e-5 fails.  Returns the first failing index, or -1 if none. */
long long oracle_verify_tolerance(const float *A, const float *B, const float *C, size_t n);

/* STREAM ops.  dst = f(x, y, s):
 *   COPY  dst = x          SCALE dst = s*x        ADD dst = x + y
 *   TRIAD dst = fmaf(s, y, x)   (one rounding; the device emits one FFMA)
 *   TRIAD_UNFUSED dst = x + (s*y) with both roundings
 *   VECTORADD dst = x + y + 0.0f */
void oracle_stream(int op, float *dst, const float *x, const float *y, float s, size_t n);

/* FNV-1a-64 folded one little-endian 32-bit word at a time (SURVEY.md §8c pin). */
uint64_t oracle_fnv1a64_words(const uint32_t *w, size_t n);

/* Order-independent, position-sensitive digest used for device-side checks at sizes where an
 * FNV chain would serialise: sum over i of mix64(w[i] + GOLDEN*(i+1)) mod 2^64. */
uint64_t oracle_digest(const uint32_t *w, size_t n);
/* Same, for a piece of a larger array whose first word has index first_index; pieces add up. */
uint64_t oracle_digest_from(const uint32_t *w, size_t n, uint64_t first_index);

/* All-reduce probe closed form: rank r holds (float)((i*(2r+1) + r) & 15);
 * the sum over nranks is an exactly representable integer.  (SURVEY's (i+r)&7 sums to the
 * constant 28 over 8 ranks, which would not catch a misplaced element.) */
float oracle_allreduce_expected(size_t i, int nranks);

/* Multi-threaded (pthread, static contiguous partitioning) forms for the timed CPU baseline. */
void oracle_vadd_mt(const float *A, const float *B, float *C, size_t n, int threads);
long long oracle_verify_tolerance_mt(const float *A, const float *B, const float *C, size_t n,
                                     int threads);
void oracle_fill_hash_mt(float *X, size_t n, uint64_t seed, int threads);
void oracle_stream_mt(int op, float *dst, const float *x, const float *y, float s, size_t n,
                      int threads);

#ifdef __cplusplus
}
#endif
#endif
