// ew_launch.cu — picks and launches the elementwise kernel for one (op, pointers, n).
//
// Dispatch rules (results are identical on every path; only the speed differs):
//   * all three pointers 32-byte aligned, dst not overlapping an input  -> tuned vector kernel
//   * 16-byte aligned only                                               -> 128-bit vector kernel
//   * anything else (misaligned base, in-place)                          -> grid-stride scalar
// The default shape is the winner of tools/tune.cu on a B200 (profiles/r01_tune.csv, summarised in
// profiles/r01_tune_top.txt); the
// environment variable B200V_EW_VARIANT selects another compiled variant for A/B runs.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "ew_kernels.cuh"

namespace b200v {

namespace {

using ew_fn = void (*)(float *, const float *, const float *, float, size_t);

struct Variant {
    const char *name;
    int vec, unroll, block;
    bool bulk;
    int tile_f;          // bulk only
    size_t smem;         // bulk only (for the two-input ops; one-input ops use half + barriers)
    int ctas_per_sm;     // grid = min(tiles, sm_count * ctas_per_sm); 0 = one block per tile
    ew_fn fn[B200V_OP_COUNT];
    size_t smem_op[B200V_OP_COUNT];
};

#define EW_VEC_ROW(VEC, UNROLL, BLOCK, LD, ST)                                        \
    {ew_vec_kernel<B200V_OP_COPY, VEC, UNROLL, BLOCK, LD, ST>,                         \
     ew_vec_kernel<B200V_OP_SCALE, VEC, UNROLL, BLOCK, LD, ST>,                        \
     ew_vec_kernel<B200V_OP_ADD, VEC, UNROLL, BLOCK, LD, ST>,                          \
     ew_vec_kernel<B200V_OP_TRIAD, VEC, UNROLL, BLOCK, LD, ST>,                        \
     ew_vec_kernel<B200V_OP_VECTORADD, VEC, UNROLL, BLOCK, LD, ST>,                    \
     ew_vec_kernel<B200V_OP_TRIAD_UNFUSED, VEC, UNROLL, BLOCK, LD, ST>}

#define EW_BULK_ROW(TILE_F, STAGES, BLOCK)                                            \
    {ew_bulk_kernel<B200V_OP_COPY, TILE_F, STAGES, BLOCK>,                             \
     ew_bulk_kernel<B200V_OP_SCALE, TILE_F, STAGES, BLOCK>,                            \
     ew_bulk_kernel<B200V_OP_ADD, TILE_F, STAGES, BLOCK>,                              \
     ew_bulk_kernel<B200V_OP_TRIAD, TILE_F, STAGES, BLOCK>,                            \
     ew_bulk_kernel<B200V_OP_VECTORADD, TILE_F, STAGES, BLOCK>,                        \
     ew_bulk_kernel<B200V_OP_TRIAD_UNFUSED, TILE_F, STAGES, BLOCK>}

#define EW_BULK_SMEM(TILE_F, STAGES)                                                  \
    {ew_bulk_smem_bytes<B200V_OP_COPY, TILE_F, STAGES>(),                              \
     ew_bulk_smem_bytes<B200V_OP_SCALE, TILE_F, STAGES>(),                             \
     ew_bulk_smem_bytes<B200V_OP_ADD, TILE_F, STAGES>(),                               \
     ew_bulk_smem_bytes<B200V_OP_TRIAD, TILE_F, STAGES>(),                             \
     ew_bulk_smem_bytes<B200V_OP_VECTORADD, TILE_F, STAGES>(),                         \
     ew_bulk_smem_bytes<B200V_OP_TRIAD_UNFUSED, TILE_F, STAGES>()}

#define EW_TILE_ROW(K, VEC, UNROLL, BLOCK, MINB)                                               \
    {K<B200V_OP_COPY, VEC, UNROLL, BLOCK, LD_PLAIN, ST_PLAIN, MINB>,                             \
     K<B200V_OP_SCALE, VEC, UNROLL, BLOCK, LD_PLAIN, ST_PLAIN, MINB>,                            \
     K<B200V_OP_ADD, VEC, UNROLL, BLOCK, LD_PLAIN, ST_PLAIN, MINB>,                              \
     K<B200V_OP_TRIAD, VEC, UNROLL, BLOCK, LD_PLAIN, ST_PLAIN, MINB>,                            \
     K<B200V_OP_VECTORADD, VEC, UNROLL, BLOCK, LD_PLAIN, ST_PLAIN, MINB>,                        \
     K<B200V_OP_TRIAD_UNFUSED, VEC, UNROLL, BLOCK, LD_PLAIN, ST_PLAIN, MINB>}

// Variant 0 is the default for 32-byte aligned data, variant 1 for data that is only 16-byte
// aligned.  Measured on a B200 at n = 2^28 (profiles/README.md): every good shape plateaus at the
// same ~7.05 TB/s (DRAM-bound); what loses is (a) a static grid-stride partition over persistent
// CTAs (3-9 %: nothing rebalances the per-SM speed spread), (b) many small tiles in flight (8 KiB
// tiles at full occupancy: -1.3 %, DRAM page locality), (c) streaming cache hints (-1 %).  The
// default is therefore the lean one-tile-per-CTA kernel with 1024-thread CTAs (32 KiB contiguous
// per array per CTA, 2 CTAs per SM).  ctas_per_sm == 0 means "one CTA per tile".
// The other shapes stay compiled for A/B runs (B200V_EW_VARIANT): the grid-stride kernel in its
// one-shot and persistent forms, the cluster-launch-control work-stealing kernel and the TMA ring.
Variant g_variants[] = {
    {"t8u1b1024", 8, 1, 1024, false, 0, 0, 0, EW_TILE_ROW(ew_tile_kernel, 8, 1, 1024, 2), {}},
    {"t4u2b1024", 4, 2, 1024, false, 0, 0, 0, EW_TILE_ROW(ew_tile_kernel, 4, 2, 1024, 2), {}},
    {"v8u1b256-oneshot", 8, 1, 256, false, 0, 0, 0, EW_VEC_ROW(8, 1, 256, LD_PLAIN, ST_PLAIN), {}},
    {"clc8u1b1024", 8, 1, 1024, false, 0, 0, 0, EW_TILE_ROW(ew_clc_kernel, 8, 1, 1024, 2), {}},
    {"v8u2b256-persist8", 8, 2, 256, false, 0, 0, 8, EW_VEC_ROW(8, 2, 256, LD_NC_NA_EF, ST_CS), {}},
    {"v4u4b256-persist8", 4, 4, 256, false, 0, 0, 8, EW_VEC_ROW(4, 4, 256, LD_NC_NA_EF, ST_CS), {}},
    {"bulk2048s3b128", 4, 0, 128, true, 2048, 0, 2, EW_BULK_ROW(2048, 3, 128), EW_BULK_SMEM(2048, 3)},
};
constexpr int kNumVariants = (int)(sizeof(g_variants) / sizeof(g_variants[0]));

int env_variant() {
    static int v = [] {
        const char *e = std::getenv("B200V_EW_VARIANT");
        if (!e || !*e) return -1;
        for (int i = 0; i < kNumVariants; ++i)
            if (std::strcmp(e, g_variants[i].name) == 0) return i;
        char *end = nullptr;
        long k = std::strtol(e, &end, 10);
        return (end && *end == 0 && k >= 0 && k < kNumVariants) ? (int)k : -1;
    }();
    return v;
}

bool overlaps(const float *a, const float *b, size_t n) {
    if (!a || !b) return false;
    const uintptr_t pa = (uintptr_t)a, pb = (uintptr_t)b, len = (uintptr_t)n * 4;
    return pa < pb + len && pb < pa + len;
}

int grid_for(size_t tiles, int sm_count, int ctas_per_sm) {
    size_t g = tiles;
    if (ctas_per_sm > 0) {
        const size_t cap = (size_t)sm_count * (size_t)ctas_per_sm;
        if (g > cap) g = cap;
    }
    if (g < 1) g = 1;
    if (g > 0x7fffffffull) g = 0x7fffffffull;
    return (int)g;
}

}  // namespace

// Which compiled shape launch_ew runs for these pointers: -1 = the scalar grid-stride kernel,
// otherwise an index into g_variants.  One place, so the launch and its description cannot drift.
static int pick_variant(const float *dst, const float *x, const float *y, size_t n) {
    const uintptr_t bits = (uintptr_t)dst | (uintptr_t)x | (uintptr_t)y;
    const bool alias = overlaps(dst, x, n) || overlaps(dst, y, n);
    if (alias || (bits & 15)) return -1;
    int vi = env_variant();
    if (vi < 0) vi = (bits & 31) ? 1 : 0;
    if ((bits & 31) && g_variants[vi].vec == 8) vi = 1;
    return vi;
}

int describe_ew(int op, const float *dst, const float *x, const float *y, size_t n, char *buf, size_t len) {
    if (op < 0 || op >= B200V_OP_COUNT || !buf || len == 0) return B200V_E_BADARG;
    static const char *ops[B200V_OP_COUNT] = {"COPY", "SCALE", "ADD", "TRIAD", "VECTORADD", "TRIAD_UNFUSED"};
    if (!ew_needs_y(op)) y = nullptr;
    const int vi = pick_variant(dst, x, y, n);
    if (vi < 0) {
        std::snprintf(buf, len, "b200v::ew_scalar_kernel<%s>", ops[op]);
        return B200V_OK;
    }
    const Variant &v = g_variants[vi];
    const char *fam = v.bulk ? "ew_bulk_kernel" : std::strncmp(v.name, "clc", 3) == 0 ? "ew_clc_kernel"
                    : (v.name[0] == 't' ? "ew_tile_kernel" : "ew_vec_kernel");
    if (v.bulk) std::snprintf(buf, len, "b200v::%s<%s,tile=%d,block=%d> (%s)", fam, ops[op], v.tile_f, v.block, v.name);
    else std::snprintf(buf, len, "b200v::%s<%s,VEC=%d,UNROLL=%d,BLOCK=%d> (%s)", fam, ops[op], v.vec, v.unroll, v.block, v.name);
    return B200V_OK;
}

int launch_ew(const DeviceCtx &ctx, int op, float *dst, const float *x, const float *y, float s,
              size_t n, cudaStream_t stream, int *grid_out, int *block_out) {
    if (op < 0 || op >= B200V_OP_COUNT) return B200V_E_BADARG;
    if (n == 0) {
        if (grid_out) *grid_out = 0;
        if (block_out) *block_out = 0;
        return B200V_OK;
    }
    if (!dst || !x || (ew_needs_y(op) && !y)) return B200V_E_BADARG;
    if (!ew_needs_y(op)) y = nullptr;

    int vi = pick_variant(dst, x, y, n);

    int grid, block;
    if (vi < 0) {
        // scalar grid-stride path: 8 elements per thread per pass keeps the grid persistent-sized
        block = 256;
        grid = grid_for((n + 255) / 256, ctx.sm_count, 8);
        static const ew_fn scalar[B200V_OP_COUNT] = {
            ew_scalar_kernel<B200V_OP_COPY>, ew_scalar_kernel<B200V_OP_SCALE>,
            ew_scalar_kernel<B200V_OP_ADD>, ew_scalar_kernel<B200V_OP_TRIAD>,
            ew_scalar_kernel<B200V_OP_VECTORADD>, ew_scalar_kernel<B200V_OP_TRIAD_UNFUSED>};
        scalar[op]<<<grid, block, 0, stream>>>(dst, x, y, s, n);
    } else {
        Variant &v = g_variants[vi];
        block = v.block;
        if (v.bulk) {
            const size_t tiles = ((n & ~(size_t)3) + v.tile_f - 1) / v.tile_f;
            grid = grid_for(tiles, ctx.sm_count, v.ctas_per_sm);
            const size_t smem = v.smem_op[op];
            B200V_CUDA(cudaFuncSetAttribute((const void *)v.fn[op],
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            v.fn[op]<<<grid, block, smem, stream>>>(dst, x, y, s, n);
        } else {
            const size_t tile = (size_t)v.block * v.unroll * v.vec;   // elements per tile
            grid = grid_for((n + tile - 1) / tile, ctx.sm_count, v.ctas_per_sm);
            v.fn[op]<<<grid, block, 0, stream>>>(dst, x, y, s, n);
        }
    }
    if (grid_out) *grid_out = grid;
    if (block_out) *block_out = block;
    B200V_CUDA(cudaGetLastError());
    return B200V_OK;
}

int launch_sampleshape(int op, float *dst, const float *x, const float *y, float s, size_t n,
                       cudaStream_t stream) {
    if (op < 0 || op >= B200V_OP_COUNT) return B200V_E_BADARG;
    if (n == 0) return B200V_OK;
    if (!dst || !x || (ew_needs_y(op) && !y)) return B200V_E_BADARG;
    const size_t blocks = (n + 255) / 256;
    if (blocks > 0x7fffffffull) return B200V_E_BADARG;
    static const ew_fn fn[B200V_OP_COUNT] = {
        ew_sampleshape_kernel<B200V_OP_COPY>, ew_sampleshape_kernel<B200V_OP_SCALE>,
        ew_sampleshape_kernel<B200V_OP_ADD>, ew_sampleshape_kernel<B200V_OP_TRIAD>,
        ew_sampleshape_kernel<B200V_OP_VECTORADD>, ew_sampleshape_kernel<B200V_OP_TRIAD_UNFUSED>};
    fn[op]<<<(unsigned)blocks, 256, 0, stream>>>(dst, x, y, s, n);
    B200V_CUDA(cudaGetLastError());
    return B200V_OK;
}

}  // namespace b200v
