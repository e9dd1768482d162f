/*
 * oracle.c — CPU restatement of the cuda-vectoradd validation path.  TEST INFRASTRUCTURE ONLY.
 * See oracle.h for the scope rule and the "parity unpinned" statement.
 *
 * Build: gcc -O2 -ffp-contract=off -fPIC -shared -pthread (oracle/Makefile).  -ffp-contract=off
 * keeps gcc from fusing a*b+c on its own, so the only fused multiply-add is the explicit fmaf.
 */
#include "oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ---- generators -------------------------------------------------------------------------- */

/* `h_A[i] = rand()/(float)RAND_MAX; h_B[i] = rand()/(float)RAND_MAX;` with no
 * srand() in the sample, i.e. glibc's default seed 1. */
void oracle_fill_rand(float *A, float *B, size_t n) {
    srand(1);
    for (size_t i = 0; i < n; ++i) {
        A[i] = rand() / (float)RAND_MAX;
        B[i] = rand() / (float)RAND_MAX;
    }
}

uint64_t oracle_splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

/* top 24 bits of splitmix64(seed + i), scaled by 2^-24 -> [0,1) exactly. */
void oracle_fill_hash(float *X, size_t n, uint64_t seed) {
    for (size_t i = 0; i < n; ++i)
        X[i] = (float)(uint32_t)(oracle_splitmix64(seed + (uint64_t)i) >> 40) * 0x1p-24f;
}

/* ---- the sample's arithmetic ---------------------------------------------------------------- */

/* `C[i] = A[i] + B[i] + 0.0f`.  The trailing + 0.0f is kept: it turns a -0.0f
 * sum into +0.0f, exactly as the device code does. */
void oracle_vadd(const float *A, const float *B, float *C, size_t n) {
    for (size_t i = 0; i < n; ++i) C[i] = A[i] + B[i] + 0.0f;
}

/* `if (fabs(h_A[i] + h_B[i] - h_C[i]) > 1e-5)` -> "Result verification failed
 * at element %d!".  float arithmetic, fabs and the comparison in double, as in the sample. */
long long oracle_verify_tolerance(const float *A, const float *B, const float *C, size_t n) {
    for (size_t i = 0; i < n; ++i)
        if (fabs(A[i] + B[i] - C[i]) > 1e-5) return (long long)i;
    return -1;
}

/* SURVEY.md §8 a9 / §7 hard part 3: triad is defined as fmaf(s, y, x) on both sides. */
void oracle_stream(int op, float *dst, const float *x, const float *y, float s, size_t n) {
    switch (op) {
    case ORACLE_OP_COPY:
        for (size_t i = 0; i < n; ++i) dst[i] = x[i];
        break;
    case ORACLE_OP_SCALE:
        for (size_t i = 0; i < n; ++i) dst[i] = s * x[i];
        break;
    case ORACLE_OP_ADD:
        for (size_t i = 0; i < n; ++i) dst[i] = x[i] + y[i];
        break;
    case ORACLE_OP_TRIAD:
        for (size_t i = 0; i < n; ++i) dst[i] = fmaf(s, y[i], x[i]);
        break;
    case ORACLE_OP_TRIAD_UNFUSED:
        for (size_t i = 0; i < n; ++i) dst[i] = x[i] + s * y[i]; /* -ffp-contract=off */
        break;
    case ORACLE_OP_VECTORADD:
        oracle_vadd(x, y, dst, n);
        break;
    default:
        break;
    }
}

/* ---- digests ------------------------------------------------------------------------------ */

uint64_t oracle_fnv1a64_words(const uint32_t *w, size_t n) {
    uint64_t h = 0xcbf29ce484222325ull;
    for (size_t i = 0; i < n; ++i) h = (h ^ (uint64_t)w[i]) * 0x100000001b3ull;
    return h;
}

static inline uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

uint64_t oracle_digest(const uint32_t *w, size_t n) { return oracle_digest_from(w, n, 0); }

/* The digest is a plain sum over elements, so an array can be digested in pieces: the digest of
 * the whole is the sum (mod 2^64) of oracle_digest_from(piece, len, index of its first element). */
uint64_t oracle_digest_from(const uint32_t *w, size_t n, uint64_t first_index) {
    uint64_t acc = 0;
    for (size_t i = 0; i < n; ++i)
        acc += mix64((uint64_t)w[i] + 0x9E3779B97F4A7C15ull * (first_index + (uint64_t)i + 1));
    return acc;
}

float oracle_allreduce_expected(size_t i, int nranks) {
    float s = 0.0f;
    for (int r = 0; r < nranks; ++r) s += (float)((i * (size_t)(2 * r + 1) + (size_t)r) & 15);
    return s;
}

/* ---- threaded forms for the CPU baseline --------------------------------------------------- */

typedef struct {
    int kind; /* 0 vadd/stream, 1 verify, 2 fill */
    int op;
    float *dst;
    const float *x, *y;
    float s;
    uint64_t seed;
    size_t lo, hi;
    long long bad;
} job_t;

static void *job_main(void *p) {
    job_t *j = (job_t *)p;
    size_t n = j->hi - j->lo;
    if (j->kind == 0) {
        oracle_stream(j->op, j->dst + j->lo, j->x + j->lo, j->y ? j->y + j->lo : NULL, j->s, n);
    } else if (j->kind == 1) {
        long long b = oracle_verify_tolerance(j->x + j->lo, j->y + j->lo, j->dst + j->lo, n);
        j->bad = b < 0 ? -1 : b + (long long)j->lo;
    } else {
        for (size_t i = j->lo; i < j->hi; ++i)
            j->dst[i] = (float)(uint32_t)(oracle_splitmix64(j->seed + (uint64_t)i) >> 40) * 0x1p-24f;
    }
    return NULL;
}

static long long run_jobs(job_t proto, size_t n, int threads) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    if ((size_t)threads > n) threads = n ? (int)n : 1;
    pthread_t tid[256];
    job_t jobs[256];
    size_t per = n / (size_t)threads, rem = n % (size_t)threads, at = 0;
    for (int t = 0; t < threads; ++t) {
        jobs[t] = proto;
        jobs[t].lo = at;
        at += per + ((size_t)t < rem ? 1 : 0);
        jobs[t].hi = at;
        jobs[t].bad = -1;
        if (t > 0) pthread_create(&tid[t], NULL, job_main, &jobs[t]);
    }
    job_main(&jobs[0]);
    long long bad = jobs[0].bad;
    for (int t = 1; t < threads; ++t) {
        pthread_join(tid[t], NULL);
        if (bad < 0 && jobs[t].bad >= 0) bad = jobs[t].bad;
    }
    return bad;
}

void oracle_stream_mt(int op, float *dst, const float *x, const float *y, float s, size_t n,
                      int threads) {
    job_t j;
    memset(&j, 0, sizeof j);
    j.kind = 0; j.op = op; j.dst = dst; j.x = x; j.y = y; j.s = s;
    run_jobs(j, n, threads);
}

void oracle_vadd_mt(const float *A, const float *B, float *C, size_t n, int threads) {
    oracle_stream_mt(ORACLE_OP_VECTORADD, C, A, B, 0.0f, n, threads);
}

long long oracle_verify_tolerance_mt(const float *A, const float *B, const float *C, size_t n,
                                     int threads) {
    job_t j;
    memset(&j, 0, sizeof j);
    j.kind = 1; j.dst = (float *)C; j.x = A; j.y = B;
    return run_jobs(j, n, threads);
}

void oracle_fill_hash_mt(float *X, size_t n, uint64_t seed, int threads) {
    job_t j;
    memset(&j, 0, sizeof j);
    j.kind = 2; j.dst = X; j.seed = seed;
    run_jobs(j, n, threads);
}
