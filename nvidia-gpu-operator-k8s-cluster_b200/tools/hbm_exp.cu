// hbm_exp.cu — round-2 HBM experiments behind DESIGN.md §4 (not part of the library):
//
//   b200v-hbm-exp phase [log2_n]   does the relative placement of the three arrays matter?  A, B and C
//                                  are carved out of one slab at base offsets that differ by odd
//                                  multiples of 2 MiB + 4 KiB, 64 KiB, ... — if DRAM channel/bank
//                                  aliasing between the three streams cost bandwidth, some placement
//                                  would beat the cudaMalloc default.
//   b200v-hbm-exp burst [log2_n]   1-read:1-write ops (copy, scale) with shapes that separate the read
//                                  burst from the write burst: 32/64/128 KiB loaded per CTA before the
//                                  first store, and stores carrying an L2 evict-first policy.
// One CSV row per measurement: name,op,offsets/shape,median_ms,median_gbs,best_gbs.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../csrc/ew_kernels.cuh"

using namespace b200v;

#define CK(x)                                                                         \
    do {                                                                              \
        cudaError_t e_ = (x);                                                         \
        if (e_ != cudaSuccess) {                                                      \
            std::fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            std::exit(2);                                                             \
        }                                                                             \
    } while (0)

using ew_fn = void (*)(float *, const float *, const float *, float, size_t);

// Stores with an L2 evict-first policy (loads plain): nothing written is read again, so the lines
// need not displace anything on their way to DRAM.
template <int OP, int UNROLL, int BLOCK>
__global__ void __launch_bounds__(BLOCK)
ew_tile_stef_kernel(float *__restrict__ dst, const float *__restrict__ x, const float *__restrict__ y, float s, size_t n) {
    constexpr bool needs_y = ew_traits<OP>::needs_y;
    const uint64_t pol = make_evict_first_policy();
    const size_t e0 = ((size_t)blockIdx.x * BLOCK * UNROLL + threadIdx.x) * 4;
    if (((size_t)blockIdx.x + 1) * BLOCK * UNROLL * 4 > n) return;      // full tiles only (n is a power of two here)
    fvec<4> a[UNROLL], b[UNROLL];
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) a[j] = ew_load<4, LD_PLAIN>(x + e0 + (size_t)j * BLOCK * 4, 0);
    if constexpr (needs_y) {
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) b[j] = ew_load<4, LD_PLAIN>(y + e0 + (size_t)j * BLOCK * 4, 0);
    }
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
        float r[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) r[k] = ew_apply<OP>(a[j].v[k], needs_y ? b[j].v[k] : 0.0f, s);
        asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;"
                     :: "l"(dst + e0 + (size_t)j * BLOCK * 4), "f"(r[0]), "f"(r[1]), "f"(r[2]), "f"(r[3]), "l"(pol) : "memory");
    }
}

static float time_kernel(ew_fn fn, unsigned grid, int block, float *d, const float *x, const float *y, size_t n,
                         int iters, float *best) {
    static cudaEvent_t ev[64];
    static bool init = false;
    if (!init) { for (auto &e : ev) CK(cudaEventCreate(&e)); init = true; }
    for (int i = 0; i < 3; ++i) fn<<<grid, block>>>(d, x, y, 3.0f, n);
    for (int i = 0; i < iters; ++i) {
        CK(cudaEventRecord(ev[2 * i]));
        fn<<<grid, block>>>(d, x, y, 3.0f, n);
        CK(cudaEventRecord(ev[2 * i + 1]));
    }
    CK(cudaDeviceSynchronize());
    std::vector<float> ms((size_t)iters);
    for (int i = 0; i < iters; ++i) CK(cudaEventElapsedTime(&ms[(size_t)i], ev[2 * i], ev[2 * i + 1]));
    std::sort(ms.begin(), ms.end());
    *best = ms[0];
    return ms[(size_t)iters / 2];
}

int main(int argc, char **argv) {
    const std::string mode = argc > 1 ? argv[1] : "phase";
    const int log2n = argc > 2 ? std::atoi(argv[2]) : 28;
    const size_t n = (size_t)1 << log2n, bytes = n * 4;
    const int iters = 20;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    int memclk = 0;
    CK(cudaDeviceGetAttribute(&memclk, cudaDevAttrMemoryClockRate, 0));
    std::fprintf(stderr, "# %s, bus %d bit, mem clock %d kHz -> nominal %.1f GB/s\n", prop.name, prop.memoryBusWidth, memclk,
                 2.0 * memclk * 1e3 * prop.memoryBusWidth / 8 / 1e9);
    const size_t slack = (size_t)64 << 20;
    char *slab;
    CK(cudaMalloc(&slab, 3 * (bytes + slack)));
    float *x0 = (float *)slab, *y0 = (float *)(slab + bytes + slack), *d0 = (float *)(slab + 2 * (bytes + slack));
    fill_hash_kernel<<<prop.multiProcessorCount * 8, 256>>>((float *)slab, 3 * (bytes + slack) / 4, 0xB200);
    CK(cudaDeviceSynchronize());
    std::printf("name,op,shape,median_ms,median_gbs,best_gbs\n");

    if (mode == "phase") {
        const size_t K = 1024, M = 1024 * 1024;
        const size_t offs[] = {0, 4 * K, 64 * K, 2 * M + 4 * K, 6 * M + 4 * K, 10 * M + 4 * K, 1 * M, 32 * M + 4 * K, 256, 1 * K};
        const unsigned grid = (unsigned)(n / 8192);
        for (int op : {B200V_OP_VECTORADD, B200V_OP_COPY}) {
            ew_fn fn = op == B200V_OP_COPY ? (ew_fn)ew_tile_kernel<B200V_OP_COPY, 8, 1, 1024, LD_PLAIN, ST_PLAIN, 2>
                                           : (ew_fn)ew_tile_kernel<B200V_OP_VECTORADD, 8, 1, 1024, LD_PLAIN, ST_PLAIN, 2>;
            for (size_t oy : offs)
                for (size_t od : offs) {
                    if (op == B200V_OP_COPY && oy != 0) continue;
                    float best = 0;
                    const float med = time_kernel(fn, grid, 1024, (float *)((char *)d0 + od), x0, (float *)((char *)y0 + oy), n, iters, &best);
                    const double b = (double)n * ew_bytes_per_elem(op);
                    std::printf("phase,%s,y+%zu_d+%zu,%.5f,%.1f,%.1f\n", op == B200V_OP_COPY ? "copy" : "vectoradd", oy, od, med,
                                b / med / 1e6, b / best / 1e6);
                    std::fflush(stdout);
                }
        }
        return 0;
    }
    if (mode == "gap") {
        // What separates two back-to-back launches of the default kernel: K launches on a stream, the
        // same K as nodes of one CUDA graph, timed as a whole; per-launch time against ncu's 445.7 us.
        const int K = 200;
        ew_fn fn = (ew_fn)ew_tile_kernel<B200V_OP_VECTORADD, 8, 1, 1024, LD_PLAIN, ST_PLAIN, 2>;
        const unsigned grid = (unsigned)(n / 8192);
        cudaStream_t st;
        CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        for (int i = 0; i < 10; ++i) fn<<<grid, 1024, 0, st>>>(d0, x0, y0, 3.0f, n);
        for (int rep = 0; rep < 3; ++rep) {
            CK(cudaEventRecord(e0, st));
            for (int i = 0; i < K; ++i) fn<<<grid, 1024, 0, st>>>(d0, x0, y0, 3.0f, n);
            CK(cudaEventRecord(e1, st));
            CK(cudaStreamSynchronize(st));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            std::printf("gap,vectoradd,stream_x%d,%.5f,%.1f,0\n", K, ms / K, 12.0 * n / (ms / K) / 1e6);
        }
        cudaGraph_t g; cudaGraphExec_t ge;
        CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        for (int i = 0; i < K; ++i) fn<<<grid, 1024, 0, st>>>(d0, x0, y0, 3.0f, n);
        CK(cudaStreamEndCapture(st, &g));
        CK(cudaGraphInstantiate(&ge, g, 0));
        CK(cudaGraphLaunch(ge, st)); CK(cudaStreamSynchronize(st));
        for (int rep = 0; rep < 3; ++rep) {
            CK(cudaEventRecord(e0, st));
            CK(cudaGraphLaunch(ge, st));
            CK(cudaEventRecord(e1, st));
            CK(cudaStreamSynchronize(st));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            std::printf("gap,vectoradd,graph_x%d,%.5f,%.1f,0\n", K, ms / K, 12.0 * n / (ms / K) / 1e6);
        }
        return 0;
    }
    // burst: shapes for the 1R:1W ops
    struct Shape { const char *name; ew_fn copy, scale; int block; size_t tile; };
#define TILE_SHAPE(V, U, B, M) {"tile_v" #V "_u" #U "_b" #B, ew_tile_kernel<B200V_OP_COPY, V, U, B, LD_PLAIN, ST_PLAIN, M>, \
                                ew_tile_kernel<B200V_OP_SCALE, V, U, B, LD_PLAIN, ST_PLAIN, M>, B, (size_t)V * U * B}
#define STEF_SHAPE(U, B) {"stef_v4_u" #U "_b" #B, ew_tile_stef_kernel<B200V_OP_COPY, U, B>, ew_tile_stef_kernel<B200V_OP_SCALE, U, B>, B, (size_t)4 * U * B}
    const Shape shapes[] = {
        TILE_SHAPE(8, 1, 1024, 2), TILE_SHAPE(8, 2, 1024, 1), TILE_SHAPE(8, 4, 1024, 1), TILE_SHAPE(8, 1, 256, 8),
        TILE_SHAPE(8, 4, 256, 1), TILE_SHAPE(8, 8, 256, 1), TILE_SHAPE(8, 8, 512, 1), TILE_SHAPE(4, 8, 1024, 1),
        STEF_SHAPE(2, 1024), STEF_SHAPE(4, 1024), STEF_SHAPE(8, 512), STEF_SHAPE(2, 256),
    };
    for (const Shape &sh : shapes)
        for (int op : {B200V_OP_COPY, B200V_OP_SCALE}) {
            float best = 0;
            const float med = time_kernel(op == B200V_OP_COPY ? sh.copy : sh.scale, (unsigned)(n / sh.tile), sh.block, d0, x0, y0, n, iters, &best);
            const double b = (double)n * 8;
            std::printf("burst,%s,%s,%.5f,%.1f,%.1f\n", op == B200V_OP_COPY ? "copy" : "scale", sh.name, med, b / med / 1e6, b / best / 1e6);
            std::fflush(stdout);
        }
    return 0;
}
