// tune.cu — sweeps the launch shapes and cache policies of the elementwise kernels on the GPU it
// runs on and prints one CSV row per (kernel variant, grid) so the product defaults in
// csrc/ew_launch.cu can be chosen from measurements.  Development tool, not shipped in the image.
//
//   b200v-tune [log2_n=28] [iters=10] [quick|tiles]
//
// Every row is checked against the scalar sample-shape kernel before it is timed.
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

struct Cfg {
    std::string name;
    int op;
    ew_fn fn;
    int block;
    size_t tile_elems;   // elements per tile (vec: BLOCK*UNROLL*VEC; bulk: TILE_F)
    size_t smem;
    bool bulk;
    bool oneshot_only = false;   // ew_tile_kernel / ew_clc_kernel: grid is always one CTA per tile
};

static std::vector<Cfg> g_cfgs;

static const char *ld_name(int p) { return p == LD_PLAIN ? "ldplain" : p == LD_NC_NA ? "ldncna" : "ldncnaef"; }
static const char *st_name(int p) { return p == ST_PLAIN ? "stplain" : p == ST_CS ? "stcs" : "stna"; }
static const char *op_name(int op) {
    static const char *n[] = {"copy", "scale", "add", "triad", "vectoradd", "triad_unfused"};
    return n[op];
}

template <int OP, int VEC, int UNROLL, int BLOCK, int LD, int ST>
static void reg_vec() {
    char buf[128];
    std::snprintf(buf, sizeof buf, "vec%d_u%d_b%d_%s_%s", VEC * 32, UNROLL, BLOCK, ld_name(LD), st_name(ST));
    g_cfgs.push_back({buf, OP, ew_vec_kernel<OP, VEC, UNROLL, BLOCK, LD, ST>, BLOCK,
                      (size_t)BLOCK * UNROLL * VEC, 0, false});
}

template <int OP, int TILE_F, int STAGES, int BLOCK>
static void reg_bulk() {
    char buf[128];
    std::snprintf(buf, sizeof buf, "bulk_t%d_s%d_b%d", TILE_F, STAGES, BLOCK);
    g_cfgs.push_back({buf, OP, ew_bulk_kernel<OP, TILE_F, STAGES, BLOCK>, BLOCK, (size_t)TILE_F,
                      ew_bulk_smem_bytes<OP, TILE_F, STAGES>(), true});
}

template <int OP, int VEC, int UNROLL, int BLOCK, int LD, int ST, int MINB>
static void reg_tile() {
    char buf[128];
    std::snprintf(buf, sizeof buf, "tile%d_u%d_b%d_m%d_%s_%s", VEC * 32, UNROLL, BLOCK, MINB, ld_name(LD), st_name(ST));
    g_cfgs.push_back({buf, OP, ew_tile_kernel<OP, VEC, UNROLL, BLOCK, LD, ST, MINB>, BLOCK,
                      (size_t)BLOCK * UNROLL * VEC, 0, false, true});
    std::snprintf(buf, sizeof buf, "clc%d_u%d_b%d_m%d_%s_%s", VEC * 32, UNROLL, BLOCK, MINB, ld_name(LD), st_name(ST));
    g_cfgs.push_back({buf, OP, ew_clc_kernel<OP, VEC, UNROLL, BLOCK, LD, ST, MINB>, BLOCK,
                      (size_t)BLOCK * UNROLL * VEC, 0, false, true});
}

template <int OP, int LD, int ST>
static void reg_tile_shapes() {
    reg_tile<OP, 8, 1, 256, LD, ST, 1>(); reg_tile<OP, 8, 1, 256, LD, ST, 8>();
    reg_tile<OP, 8, 2, 256, LD, ST, 1>(); reg_tile<OP, 8, 2, 256, LD, ST, 4>();
    reg_tile<OP, 8, 4, 256, LD, ST, 1>(); reg_tile<OP, 8, 1, 512, LD, ST, 1>();
    reg_tile<OP, 8, 1, 512, LD, ST, 4>(); reg_tile<OP, 8, 2, 512, LD, ST, 1>();
    reg_tile<OP, 8, 1, 1024, LD, ST, 1>(); reg_tile<OP, 8, 1, 1024, LD, ST, 2>();
    reg_tile<OP, 8, 2, 1024, LD, ST, 1>(); reg_tile<OP, 8, 1, 128, LD, ST, 1>();
    reg_tile<OP, 8, 1, 128, LD, ST, 16>(); reg_tile<OP, 8, 2, 128, LD, ST, 1>();
    reg_tile<OP, 4, 2, 256, LD, ST, 1>(); reg_tile<OP, 4, 4, 256, LD, ST, 1>();
    reg_tile<OP, 4, 2, 512, LD, ST, 1>(); reg_tile<OP, 4, 1, 256, LD, ST, 8>();
}

template <int OP, int LD, int ST>
static void reg_vec_shapes() {
    reg_vec<OP, 4, 1, 256, LD, ST>(); reg_vec<OP, 4, 2, 256, LD, ST>();
    reg_vec<OP, 4, 4, 256, LD, ST>(); reg_vec<OP, 4, 8, 256, LD, ST>();
    reg_vec<OP, 4, 2, 512, LD, ST>(); reg_vec<OP, 4, 4, 512, LD, ST>();
    reg_vec<OP, 4, 4, 128, LD, ST>(); reg_vec<OP, 4, 8, 128, LD, ST>();
    reg_vec<OP, 4, 2, 1024, LD, ST>(); reg_vec<OP, 4, 4, 1024, LD, ST>();
    reg_vec<OP, 8, 1, 256, LD, ST>(); reg_vec<OP, 8, 2, 256, LD, ST>();
    reg_vec<OP, 8, 4, 256, LD, ST>(); reg_vec<OP, 8, 1, 512, LD, ST>();
    reg_vec<OP, 8, 2, 512, LD, ST>(); reg_vec<OP, 8, 2, 128, LD, ST>();
    reg_vec<OP, 8, 4, 128, LD, ST>(); reg_vec<OP, 8, 1, 1024, LD, ST>();
    reg_vec<OP, 8, 2, 1024, LD, ST>();
}

template <int OP>
static void reg_bulk_shapes() {
    reg_bulk<OP, 2048, 3, 128>(); reg_bulk<OP, 2048, 4, 128>(); reg_bulk<OP, 2048, 6, 128>();
    reg_bulk<OP, 4096, 3, 256>(); reg_bulk<OP, 4096, 4, 256>(); reg_bulk<OP, 4096, 6, 256>();
    reg_bulk<OP, 8192, 3, 256>(); reg_bulk<OP, 8192, 4, 256>(); reg_bulk<OP, 4096, 4, 128>();
    reg_bulk<OP, 8192, 4, 512>(); reg_bulk<OP, 16384, 3, 512>();
}

static void register_all(bool quick, bool tiles_only) {
    reg_tile_shapes<B200V_OP_VECTORADD, LD_PLAIN, ST_PLAIN>();
    reg_tile_shapes<B200V_OP_VECTORADD, LD_NC_NA, ST_PLAIN>();
    reg_tile_shapes<B200V_OP_COPY, LD_PLAIN, ST_PLAIN>();
    reg_tile_shapes<B200V_OP_TRIAD, LD_PLAIN, ST_PLAIN>();
    reg_tile_shapes<B200V_OP_SCALE, LD_PLAIN, ST_PLAIN>();
    reg_tile_shapes<B200V_OP_ADD, LD_PLAIN, ST_PLAIN>();
    if (tiles_only) {   // keep two reference rows from the first sweep
        reg_vec<B200V_OP_VECTORADD, 8, 1, 256, LD_PLAIN, ST_PLAIN>();
        reg_vec<B200V_OP_VECTORADD, 8, 2, 1024, LD_PLAIN, ST_PLAIN>();
        return;
    }
    reg_vec_shapes<B200V_OP_VECTORADD, LD_NC_NA_EF, ST_CS>();
    reg_bulk_shapes<B200V_OP_VECTORADD>();
    if (!quick) {
        reg_vec_shapes<B200V_OP_VECTORADD, LD_PLAIN, ST_PLAIN>();
        reg_vec_shapes<B200V_OP_VECTORADD, LD_NC_NA, ST_PLAIN>();
        reg_vec_shapes<B200V_OP_VECTORADD, LD_NC_NA, ST_CS>();
        reg_vec_shapes<B200V_OP_VECTORADD, LD_NC_NA_EF, ST_NA>();
        reg_vec_shapes<B200V_OP_VECTORADD, LD_PLAIN, ST_CS>();
    }
    reg_vec_shapes<B200V_OP_COPY, LD_NC_NA_EF, ST_CS>();
    reg_vec_shapes<B200V_OP_TRIAD, LD_NC_NA_EF, ST_CS>();
    reg_vec_shapes<B200V_OP_SCALE, LD_NC_NA_EF, ST_CS>();
    reg_vec_shapes<B200V_OP_ADD, LD_NC_NA_EF, ST_CS>();
    if (!quick) {
        reg_vec_shapes<B200V_OP_COPY, LD_PLAIN, ST_PLAIN>();
        reg_vec_shapes<B200V_OP_TRIAD, LD_PLAIN, ST_PLAIN>();
    }
    reg_bulk_shapes<B200V_OP_COPY>();
    reg_bulk_shapes<B200V_OP_TRIAD>();
}

struct Row {
    std::string name;
    int op, grid, block, ctas_per_sm;
    float best_ms, median_ms;
    double best_gbs, median_gbs;
    unsigned long long bad;
};

int main(int argc, char **argv) {
    const int log2n = argc > 1 ? std::atoi(argv[1]) : 28;
    const int iters = argc > 2 ? std::atoi(argv[2]) : 10;
    const bool quick = argc > 3 && std::strcmp(argv[3], "quick") == 0;
    const bool tiles_only = argc > 3 && std::strcmp(argv[3], "tiles") == 0;
    const size_t n = (size_t)1 << log2n;

    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    std::fprintf(stderr, "# %s, %d SMs, L2 %d MiB, n=2^%d, iters=%d\n", prop.name, sms,
                 prop.l2CacheSize >> 20, log2n, iters);

    float *x, *y, *d, *ref;
    unsigned long long *d_out;
    CK(cudaMalloc(&x, n * 4)); CK(cudaMalloc(&y, n * 4)); CK(cudaMalloc(&d, n * 4)); CK(cudaMalloc(&ref, n * 4));
    CK(cudaMalloc(&d_out, 16));
    fill_hash_kernel<<<sms * 8, 256>>>(x, n, 0xB200);
    fill_hash_kernel<<<sms * 8, 256>>>(y, n, 0xB201);
    CK(cudaDeviceSynchronize());

    register_all(quick, tiles_only);
    std::vector<cudaEvent_t> ev((size_t)iters * 2);
    for (auto &e : ev) CK(cudaEventCreate(&e));

    // baseline: the driver's device-to-device copy
    {
        std::vector<float> ms;
        for (int i = 0; i < 3 + iters; ++i) {
            CK(cudaEventRecord(ev[0]));
            CK(cudaMemcpyAsync(d, x, n * 4, cudaMemcpyDeviceToDevice));
            CK(cudaEventRecord(ev[1]));
            CK(cudaEventSynchronize(ev[1]));
            float t; CK(cudaEventElapsedTime(&t, ev[0], ev[1]));
            if (i >= 3) ms.push_back(t);
        }
        std::sort(ms.begin(), ms.end());
        std::printf("name,op,grid,block,ctas_per_sm,best_ms,median_ms,best_gbs,median_gbs,mismatches\n");
        std::printf("cudaMemcpyD2D,copy,0,0,0,%.4f,%.4f,%.1f,%.1f,0\n", ms[0], ms[ms.size() / 2],
                    n * 8.0 / ms[0] / 1e6, n * 8.0 / ms[ms.size() / 2] / 1e6);
    }

    std::vector<Row> rows;
    int last_ref_op = -1;
    const int grid_mults[] = {0, 1, 2, 3, 4, 6, 8, 12, 16, 32};
    for (const Cfg &c : g_cfgs) {
        if (c.smem > (size_t)prop.sharedMemPerBlockOptin) continue;
        if (c.smem) CK(cudaFuncSetAttribute((const void *)c.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c.smem));
        int occ = 0;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void *)c.fn, c.block, c.smem));
        if (occ < 1) continue;
        if (c.op != last_ref_op) {   // reference output from the sample-shape scalar kernel
            const unsigned blocks = (unsigned)((n + 255) / 256);
            switch (c.op) {
            case B200V_OP_COPY: ew_sampleshape_kernel<B200V_OP_COPY><<<blocks, 256>>>(ref, x, y, 3.0f, n); break;
            case B200V_OP_SCALE: ew_sampleshape_kernel<B200V_OP_SCALE><<<blocks, 256>>>(ref, x, y, 3.0f, n); break;
            case B200V_OP_ADD: ew_sampleshape_kernel<B200V_OP_ADD><<<blocks, 256>>>(ref, x, y, 3.0f, n); break;
            case B200V_OP_TRIAD: ew_sampleshape_kernel<B200V_OP_TRIAD><<<blocks, 256>>>(ref, x, y, 3.0f, n); break;
            default: ew_sampleshape_kernel<B200V_OP_VECTORADD><<<blocks, 256>>>(ref, x, y, 3.0f, n); break;
            }
            CK(cudaDeviceSynchronize());
            last_ref_op = c.op;
        }
        const size_t tiles = (n + c.tile_elems - 1) / c.tile_elems;
        for (int m : grid_mults) {
            if (m > occ) continue;                  // more CTAs than fit is just a second wave
            if (m == 0 && c.bulk) continue;         // the bulk ring needs several tiles per CTA
            if (m != 0 && c.oneshot_only) continue;
            const size_t g = m == 0 ? tiles : std::min(tiles, (size_t)sms * m);
            CK(cudaMemsetAsync(d, 0xFF, n * 4));
            c.fn<<<(unsigned)g, c.block, c.smem>>>(d, x, y, 3.0f, n);
            CK(cudaGetLastError());
            // correctness: bitwise equal to the reference (COPY check of d against ref)
            const unsigned long long init[2] = {0ull, ~0ull};
            CK(cudaMemcpy(d_out, init, 16, cudaMemcpyHostToDevice));
            check_kernel<B200V_OP_COPY><<<sms * 8, 256>>>(d, ref, nullptr, 0.f, n, d_out);
            unsigned long long res[2];
            CK(cudaMemcpy(res, d_out, 16, cudaMemcpyDeviceToHost));
            for (int i = 0; i < 2; ++i) c.fn<<<(unsigned)g, c.block, c.smem>>>(d, x, y, 3.0f, n);
            for (int i = 0; i < iters; ++i) {
                CK(cudaEventRecord(ev[2 * i]));
                c.fn<<<(unsigned)g, c.block, c.smem>>>(d, x, y, 3.0f, n);
                CK(cudaEventRecord(ev[2 * i + 1]));
            }
            CK(cudaDeviceSynchronize());
            std::vector<float> ms((size_t)iters);
            for (int i = 0; i < iters; ++i) CK(cudaEventElapsedTime(&ms[i], ev[2 * i], ev[2 * i + 1]));
            std::sort(ms.begin(), ms.end());
            const double bytes = (double)n * ew_bytes_per_elem(c.op);
            Row r{c.name, c.op, (int)g, c.block, m, ms[0], ms[(size_t)iters / 2],
                  bytes / ms[0] / 1e6, bytes / ms[(size_t)iters / 2] / 1e6, res[0]};
            rows.push_back(r);
            std::printf("%s,%s,%d,%d,%d,%.4f,%.4f,%.1f,%.1f,%llu\n", r.name.c_str(), op_name(r.op), r.grid,
                        r.block, r.ctas_per_sm, r.best_ms, r.median_ms, r.best_gbs, r.median_gbs, r.bad);
            std::fflush(stdout);
        }
    }

    for (int op = 0; op < B200V_OP_COUNT; ++op) {
        std::vector<Row> sel;
        for (const Row &r : rows) if (r.op == op) sel.push_back(r);
        if (sel.empty()) continue;
        std::sort(sel.begin(), sel.end(), [](const Row &a, const Row &b) { return a.median_gbs > b.median_gbs; });
        std::fprintf(stderr, "# top for %s (median GB/s)\n", op_name(op));
        for (size_t i = 0; i < std::min<size_t>(12, sel.size()); ++i)
            std::fprintf(stderr, "#   %-40s grid=%-8d ctas/sm=%-2d median %.1f best %.1f bad=%llu\n",
                         sel[i].name.c_str(), sel[i].grid, sel[i].ctas_per_sm, sel[i].median_gbs,
                         sel[i].best_gbs, sel[i].bad);
    }
    return 0;
}
