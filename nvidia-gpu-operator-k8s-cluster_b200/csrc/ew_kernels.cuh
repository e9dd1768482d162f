// ew_kernels.cuh — the elementwise fp32 kernels of the validation workload, for sm_100a.
//
// vectorAdd and the STREAM quartet (§8 a9) are the same shape of work: read one
// or two fp32 arrays once, write one.  There is no reuse and no contraction, so nothing here uses
// shared-memory tiling for reuse or tensor cores; the whole job is keeping HBM3e busy:
//   * 128-bit (LDG.E.128) or 256-bit (LDG.E.256, new on sm_100) fully coalesced accesses,
//   * UNROLL independent loads per array issued before the first use (memory-level parallelism),
//   * streaming cache policy (no L1 allocation, L2 evict-first) because nothing is read twice,
//   * a grid that is a whole multiple of the SM count, walking tiles with a grid stride.
// A second path (ew_bulk_kernel) moves the same tiles with the TMA engine's 1-D bulk copies
// (cp.async.bulk + mbarrier ring) so that one elected thread keeps the queue deep.
//
// Included by capi.cu (product) and tools/tune.cu (sweep); both compile for sm_100a only.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/b200validate.h"

namespace b200v {

// ---- arithmetic ---------------------------------------------------------------------------------
// Explicit _rn intrinsics: nvcc never contracts them, so TRIAD is exactly one FFMA and
// TRIAD_UNFUSED exactly FMUL+FADD, whatever -fmad says.  VECTORADD keeps the sample's "+ 0.0f"
// (it maps a -0.0f sum to +0.0f; the compiler may not fold it).
template <int OP>
__device__ __forceinline__ float ew_apply(float x, float y, float s) {
    if constexpr (OP == B200V_OP_COPY) return x;
    else if constexpr (OP == B200V_OP_SCALE) return __fmul_rn(s, x);
    else if constexpr (OP == B200V_OP_ADD) return __fadd_rn(x, y);
    else if constexpr (OP == B200V_OP_TRIAD) return __fmaf_rn(s, y, x);
    else if constexpr (OP == B200V_OP_VECTORADD) return __fadd_rn(__fadd_rn(x, y), 0.0f);
    else return __fadd_rn(x, __fmul_rn(s, y));
}

template <int OP>
struct ew_traits {
    static constexpr bool needs_y = (OP >= B200V_OP_ADD);
    // Algorithmic bytes per element, STREAM convention: no write-allocate.
    static constexpr int bytes_per_elem = needs_y ? 12 : 8;
};

__host__ __device__ inline int ew_bytes_per_elem(int op) { return op >= B200V_OP_ADD ? 12 : 8; }
__host__ __device__ inline bool ew_needs_y(int op) { return op >= B200V_OP_ADD; }

// ---- vector load / store with cache policy -----------------------------------------------------
enum { LD_PLAIN = 0, LD_NC_NA = 1, LD_NC_NA_EF = 2 };   // load policies
enum { ST_PLAIN = 0, ST_CS = 1, ST_NA = 2 };            // store policies

template <int VEC>
struct alignas(VEC * 4) fvec { float v[VEC]; };

// L2 evict-first is a load modifier only on the 256-bit form; 128-bit loads carry it as a cache
// policy operand made once per thread with createpolicy.
__device__ __forceinline__ uint64_t make_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

template <int VEC, int LDPOL>
__device__ __forceinline__ fvec<VEC> ew_load(const float *p, uint64_t pol) {
    fvec<VEC> r;
    (void)pol;
    if constexpr (VEC == 4) {
        if constexpr (LDPOL == LD_PLAIN)
            asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]) : "l"(p));
        else if constexpr (LDPOL == LD_NC_NA)
            asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]) : "l"(p));
        else
            asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
                         : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]) : "l"(p), "l"(pol));
    } else {
        static_assert(VEC == 8, "VEC is 4 or 8");
        if constexpr (LDPOL == LD_PLAIN)
            asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]),
                           "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7]) : "l"(p));
        else if constexpr (LDPOL == LD_NC_NA)
            asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]),
                           "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7]) : "l"(p));
        else
            asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.f32 "
                         "{%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]),
                           "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7]) : "l"(p));
    }
    return r;
}

template <int VEC, int STPOL>
__device__ __forceinline__ void ew_store(float *p, const fvec<VEC> &r) {
    if constexpr (VEC == 4) {
        if constexpr (STPOL == ST_PLAIN)
            asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};"
                         :: "l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]) : "memory");
        else if constexpr (STPOL == ST_CS)
            asm volatile("st.global.cs.v4.f32 [%0], {%1,%2,%3,%4};"
                         :: "l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]) : "memory");
        else
            asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
                         :: "l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]) : "memory");
    } else {
        if constexpr (STPOL == ST_PLAIN)
            asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                         :: "l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]),
                            "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7]) : "memory");
        else if constexpr (STPOL == ST_CS)
            asm volatile("st.global.cs.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                         :: "l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]),
                            "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7]) : "memory");
        else
            asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                         :: "l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]),
                            "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7]) : "memory");
    }
}

// ---- the vector kernel ---------------------------------------------------------------------------
// Work is cut into tiles of BLOCK*UNROLL vectors.  Block b walks tiles b, b+grid, b+2*grid, ...;
// inside a tile thread t owns vectors t, t+BLOCK, ... so every warp-wide access is one contiguous
// 32*VEC*4-byte run.  All loads of a tile are issued before the first arithmetic instruction.
// Requires dst, x, y aligned to VEC*4 bytes; the n % VEC tail elements are done by scalar lanes.
// With grid == number of tiles the loop body runs once per block (the one-shot shape).
template <int OP, int VEC, int UNROLL, int BLOCK, int LDPOL, int STPOL>
__global__ void __launch_bounds__(BLOCK)
ew_vec_kernel(float *__restrict__ dst, const float *__restrict__ x, const float *__restrict__ y,
              float s, size_t n) {
    constexpr bool needs_y = ew_traits<OP>::needs_y;
    constexpr size_t TILE = (size_t)BLOCK * UNROLL;          // vectors per tile
    const size_t n_vec = n / VEC;
    const size_t full_tiles = n_vec / TILE;
    const uint64_t pol = (LDPOL == LD_NC_NA_EF && VEC == 4) ? make_evict_first_policy() : 0;

    for (size_t t = blockIdx.x; t < full_tiles; t += gridDim.x) {
        const size_t v0 = t * TILE + threadIdx.x;
        fvec<VEC> a[UNROLL], b[UNROLL];
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) a[j] = ew_load<VEC, LDPOL>(x + (v0 + (size_t)j * BLOCK) * VEC, pol);
        if constexpr (needs_y) {
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) b[j] = ew_load<VEC, LDPOL>(y + (v0 + (size_t)j * BLOCK) * VEC, pol);
        }
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) {
            fvec<VEC> r;
#pragma unroll
            for (int k = 0; k < VEC; ++k) r.v[k] = ew_apply<OP>(a[j].v[k], needs_y ? b[j].v[k] : 0.0f, s);
            ew_store<VEC, STPOL>(dst + (v0 + (size_t)j * BLOCK) * VEC, r);
        }
    }

    // Ragged end: the vectors past the last full tile, then the n % VEC scalar elements.  One
    // block takes them (the one whose turn it would have been), guarded per vector.
    if (blockIdx.x == (unsigned)(full_tiles % gridDim.x)) {
        for (size_t v = full_tiles * TILE + threadIdx.x; v < n_vec; v += BLOCK) {
            fvec<VEC> a = ew_load<VEC, LDPOL>(x + v * VEC, pol), b, r;
            if constexpr (needs_y) b = ew_load<VEC, LDPOL>(y + v * VEC, pol);
#pragma unroll
            for (int k = 0; k < VEC; ++k) r.v[k] = ew_apply<OP>(a.v[k], needs_y ? b.v[k] : 0.0f, s);
            ew_store<VEC, STPOL>(dst + v * VEC, r);
        }
        const size_t i = n_vec * VEC + threadIdx.x;
        if (threadIdx.x < VEC && i < n) dst[i] = ew_apply<OP>(x[i], needs_y ? y[i] : 0.0f, s);
    }
}

// ---- the one-tile-per-CTA kernel ---------------------------------------------------------------------
// The measured winner is "one tile per CTA" (the hardware block scheduler balances the SMs), and
// for that shape the grid-stride loop above is dead weight: its trip-count division and the
// ragged-end ownership test cost ~80 scalar instructions before the first load issues, in each of
// the 131072 CTAs of a 2^28-element launch.  This kernel is that shape with nothing else in it:
// CTA b owns elements [b*TILE*VEC, (b+1)*TILE*VEC); every CTA but the last takes the unguarded
// path.  Launch with grid = ceil(n / (BLOCK*UNROLL*VEC)) (>= 1).  MINB is the occupancy hint.
template <int OP, int VEC, int UNROLL, int BLOCK, int LDPOL, int STPOL, int MINB = 1>
__global__ void __launch_bounds__(BLOCK, MINB)
ew_tile_kernel(float *__restrict__ dst, const float *__restrict__ x, const float *__restrict__ y,
               float s, size_t n) {
    constexpr bool needs_y = ew_traits<OP>::needs_y;
    constexpr size_t TILE = (size_t)BLOCK * UNROLL;          // vectors per tile
    const uint64_t pol = (LDPOL == LD_NC_NA_EF && VEC == 4) ? make_evict_first_policy() : 0;
    const size_t v0 = (size_t)blockIdx.x * TILE + threadIdx.x;
    const size_t e0 = v0 * VEC;

    if (((size_t)blockIdx.x + 1) * TILE * VEC <= n) {
        fvec<VEC> a[UNROLL], b[UNROLL];
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) a[j] = ew_load<VEC, LDPOL>(x + e0 + (size_t)j * BLOCK * VEC, pol);
        if constexpr (needs_y) {
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) b[j] = ew_load<VEC, LDPOL>(y + e0 + (size_t)j * BLOCK * VEC, pol);
        }
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) {
            fvec<VEC> r;
#pragma unroll
            for (int k = 0; k < VEC; ++k) r.v[k] = ew_apply<OP>(a[j].v[k], needs_y ? b[j].v[k] : 0.0f, s);
            ew_store<VEC, STPOL>(dst + e0 + (size_t)j * BLOCK * VEC, r);
        }
        return;
    }
    // last CTA: guarded vectors, then the n % VEC scalar elements
    const size_t n_vec = n / VEC;
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
        const size_t v = v0 + (size_t)j * BLOCK;
        if (v < n_vec) {
            fvec<VEC> a = ew_load<VEC, LDPOL>(x + v * VEC, pol), b, r;
            if constexpr (needs_y) b = ew_load<VEC, LDPOL>(y + v * VEC, pol);
#pragma unroll
            for (int k = 0; k < VEC; ++k) r.v[k] = ew_apply<OP>(a.v[k], needs_y ? b.v[k] : 0.0f, s);
            ew_store<VEC, STPOL>(dst + v * VEC, r);
        }
    }
    const size_t i = n_vec * VEC + threadIdx.x;
    if (threadIdx.x < VEC && i < n) dst[i] = ew_apply<OP>(x[i], needs_y ? y[i] : 0.0f, s);
}

// ---- persistent kernel with hardware work stealing (cluster launch control) ---------------------------
// Blackwell's answer to "persistent CTAs, but dynamically balanced": the grid is still one CTA per
// tile, but a running CTA that finishes its tile asks the hardware scheduler to cancel a CTA that
// has not started yet (clusterlaunchcontrol.try_cancel) and processes that CTA's tile itself.  The
// CTA-launch cost is paid ~once per SM slot instead of once per tile, and unlike a static
// grid-stride partition the work still flows to whichever SM is free.  Launch exactly like
// ew_tile_kernel.  Thread 0 issues the request one tile ahead; the 16-byte response lands in
// shared memory and completes an mbarrier all threads wait on.
template <int OP, int VEC, int UNROLL, int BLOCK, int LDPOL, int STPOL, int MINB = 1>
__global__ void __launch_bounds__(BLOCK, MINB)
ew_clc_kernel(float *__restrict__ dst, const float *__restrict__ x, const float *__restrict__ y,
              float s, size_t n) {
    constexpr bool needs_y = ew_traits<OP>::needs_y;
    constexpr size_t TILE = (size_t)BLOCK * UNROLL;
    __shared__ __align__(16) uint4 clc_resp[2];
    __shared__ __align__(8) uint64_t clc_bar[2];
    const uint64_t pol = (LDPOL == LD_NC_NA_EF && VEC == 4) ? make_evict_first_policy() : 0;

    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"((uint32_t)__cvta_generic_to_shared(&clc_bar[0])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"((uint32_t)__cvta_generic_to_shared(&clc_bar[1])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    auto request = [&](int slot) {      // thread 0 only
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&clc_bar[slot]);
        const uint32_t resp = (uint32_t)__cvta_generic_to_shared(&clc_resp[slot]);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], 16;" :: "r"(bar) : "memory");
        asm volatile("clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes.b128 [%0], [%1];"
                     :: "r"(resp), "r"(bar) : "memory");
    };

    unsigned tile = blockIdx.x;
    int slot = 0;
    uint32_t phase[2] = {0, 0};
    if (threadIdx.x == 0) request(0);
    for (;;) {
        // ---- process `tile` ----
        const size_t v0 = (size_t)tile * TILE + threadIdx.x;
        const size_t e0 = v0 * VEC;
        if (((size_t)tile + 1) * TILE * VEC <= n) {
            fvec<VEC> a[UNROLL], b[UNROLL];
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) a[j] = ew_load<VEC, LDPOL>(x + e0 + (size_t)j * BLOCK * VEC, pol);
            if constexpr (needs_y) {
#pragma unroll
                for (int j = 0; j < UNROLL; ++j) b[j] = ew_load<VEC, LDPOL>(y + e0 + (size_t)j * BLOCK * VEC, pol);
            }
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) {
                fvec<VEC> r;
#pragma unroll
                for (int k = 0; k < VEC; ++k) r.v[k] = ew_apply<OP>(a[j].v[k], needs_y ? b[j].v[k] : 0.0f, s);
                ew_store<VEC, STPOL>(dst + e0 + (size_t)j * BLOCK * VEC, r);
            }
        } else {
            const size_t n_vec = n / VEC;
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) {
                const size_t v = v0 + (size_t)j * BLOCK;
                if (v < n_vec) {
                    fvec<VEC> a = ew_load<VEC, LDPOL>(x + v * VEC, pol), b, r;
                    if constexpr (needs_y) b = ew_load<VEC, LDPOL>(y + v * VEC, pol);
#pragma unroll
                    for (int k = 0; k < VEC; ++k) r.v[k] = ew_apply<OP>(a.v[k], needs_y ? b.v[k] : 0.0f, s);
                    ew_store<VEC, STPOL>(dst + v * VEC, r);
                }
            }
            const size_t i = n_vec * VEC + threadIdx.x;
            if (threadIdx.x < VEC && i < n) dst[i] = ew_apply<OP>(x[i], needs_y ? y[i] : 0.0f, s);
        }
        // ---- next tile: the response requested while this one was in flight ----
        {
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&clc_bar[slot]);
            asm volatile(
                "{\n.reg .pred p;\nCLCW_%=:\n"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                "@p bra CLCD_%=;\nbra CLCW_%=;\nCLCD_%=:\n}\n" :: "r"(bar), "r"(phase[slot]) : "memory");
            phase[slot] ^= 1;
        }
        const uint32_t resp = (uint32_t)__cvta_generic_to_shared(&clc_resp[slot]);
        uint32_t ok = 0, next = 0;
        asm volatile(
            "{\n.reg .pred p;\n.reg .b128 r;\n"
            "ld.shared.b128 r, [%2];\n"
            "clusterlaunchcontrol.query_cancel.is_canceled.pred.b128 p, r;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "@p clusterlaunchcontrol.query_cancel.get_first_ctaid::x.b32.b128 %1, r;\n"
            "}\n" : "=r"(ok), "=r"(next) : "r"(resp) : "memory");
        if (!ok) break;
        tile = next;
        // every thread has read this slot's response before it is reused two rounds from now
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        slot ^= 1;
        if (threadIdx.x == 0) request(slot);
    }
}

// ---- scalar kernels --------------------------------------------------------------------------------
// (1) The sample's own shape: one element per thread, bounds-checked.  Used as
//     the in-library cross-check of the vector kernels and for the sampleshape entry point.
template <int OP>
__global__ void __launch_bounds__(256)
ew_sampleshape_kernel(float *dst, const float *x, const float *y, float s, size_t n) {
    const size_t i = (size_t)blockDim.x * blockIdx.x + threadIdx.x;
    if (i < n) dst[i] = ew_apply<OP>(x[i], ew_traits<OP>::needs_y ? y[i] : 0.0f, s);
}

// (2) Grid-stride scalar path for pointers that are not 16-byte aligned (any of the three), and
//     for dst aliasing x or y (no .nc loads here).  Still fully coalesced, 4 bytes per lane.
template <int OP>
__global__ void __launch_bounds__(256)
ew_scalar_kernel(float *dst, const float *x, const float *y, float s, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockDim.x * blockIdx.x + threadIdx.x; i < n; i += stride)
        dst[i] = ew_apply<OP>(x[i], ew_traits<OP>::needs_y ? y[i] : 0.0f, s);
}

// ---- TMA bulk-copy path ------------------------------------------------------------------------------
// Each CTA owns a ring of STAGES shared-memory slots of TILE_F floats per input array.  Thread 0
// is the producer: it arms the slot's mbarrier with the expected byte count and issues one
// cp.async.bulk (global -> shared) per input array.  All threads consume: wait on the slot's
// mbarrier, compute in place (the result overwrites the x slot), make the writes visible to the
// async proxy, barrier, and thread 0 issues one cp.async.bulk (shared -> global) for the tile.
// A slot is refilled only after the bulk store that read it has drained
// (cp.async.bulk.wait_group.read).  Requires 16-byte aligned pointers; tiles shorter than TILE_F
// (the last one) use a byte count rounded down to 16 and leave the < 4-element tail to scalar lanes.
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                 :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gmem_src, uint32_t bytes,
                                         uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        :: "r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *gmem_dst, const void *smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 :: "l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <int OP, int TILE_F, int STAGES, int BLOCK>
__global__ void __launch_bounds__(BLOCK)
ew_bulk_kernel(float *__restrict__ dst, const float *__restrict__ x, const float *__restrict__ y,
               float s, size_t n) {
    constexpr bool needs_y = ew_traits<OP>::needs_y;
    constexpr int NARR = needs_y ? 2 : 1;
    static_assert(TILE_F % (4 * BLOCK) == 0, "tile is a whole number of float4 per thread");
    constexpr int PER_THREAD = TILE_F / (4 * BLOCK);

    extern __shared__ __align__(128) unsigned char smem_raw[];
    float *slots = reinterpret_cast<float *>(smem_raw);                    // [STAGES][NARR][TILE_F]
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + (size_t)STAGES * NARR * TILE_F * 4);

    const size_t n4 = n & ~(size_t)3;                       // elements the bulk path may touch
    const size_t tiles = (n4 + TILE_F - 1) / TILE_F;
    // Tiles of this block: blockIdx.x, blockIdx.x + grid, ...
    const size_t my_tiles = tiles > blockIdx.x ? (tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

    if (threadIdx.x == 0) {
        for (int i = 0; i < STAGES; ++i) mbar_init(&full[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    auto tile_elems = [&](size_t k) -> uint32_t {          // elements in this block's k-th tile
        const size_t t = blockIdx.x + k * gridDim.x;
        const size_t lo = t * TILE_F;
        return (uint32_t)((n4 - lo) < (size_t)TILE_F ? (n4 - lo) : (size_t)TILE_F);
    };
    auto issue = [&](size_t k) {                           // producer: fill slot k % STAGES
        const int slot = (int)(k % STAGES);
        const size_t lo = (blockIdx.x + k * gridDim.x) * (size_t)TILE_F;
        const uint32_t bytes = tile_elems(k) * 4u;
        float *sx = slots + (size_t)slot * NARR * TILE_F;
        mbar_expect_tx(&full[slot], bytes * NARR);
        bulk_g2s(sx, x + lo, bytes, &full[slot]);
        if constexpr (needs_y) bulk_g2s(sx + TILE_F, y + lo, bytes, &full[slot]);
    };

    if (threadIdx.x == 0) {
        const size_t pre = my_tiles < (size_t)(STAGES - 1) ? my_tiles : (size_t)(STAGES - 1);
        for (size_t k = 0; k < pre; ++k) issue(k);
    }

    for (size_t k = 0; k < my_tiles; ++k) {
        const int slot = (int)(k % STAGES);
        const uint32_t parity = (uint32_t)((k / STAGES) & 1);
        float *sx = slots + (size_t)slot * NARR * TILE_F;
        const uint32_t elems = tile_elems(k);
        mbar_wait(&full[slot], parity);
#pragma unroll
        for (int j = 0; j < PER_THREAD; ++j) {
            const int e = (j * BLOCK + threadIdx.x) * 4;
            if ((uint32_t)e < elems) {
                float4 a = *reinterpret_cast<float4 *>(sx + e), b = make_float4(0, 0, 0, 0), r;
                if constexpr (needs_y) b = *reinterpret_cast<float4 *>(sx + TILE_F + e);
                r.x = ew_apply<OP>(a.x, b.x, s); r.y = ew_apply<OP>(a.y, b.y, s);
                r.z = ew_apply<OP>(a.z, b.z, s); r.w = ew_apply<OP>(a.w, b.w, s);
                *reinterpret_cast<float4 *>(sx + e) = r;
            }
        }
        fence_async_smem();
        __syncthreads();
        if (threadIdx.x == 0) {
            const size_t lo = (blockIdx.x + k * gridDim.x) * (size_t)TILE_F;
            bulk_s2g(dst + lo, sx, elems * 4u);
            bulk_commit();
            // Slot (k-1) % STAGES was stored in the previous iteration; once every group but the
            // one just committed has finished reading shared memory it can take tile k+STAGES-1.
            if (k + STAGES - 1 < my_tiles) {
                bulk_wait_read<1>();
                issue(k + STAGES - 1);
            }
        }
    }
    if (threadIdx.x == 0) bulk_wait_all();

    // n % 4 tail elements.
    if (blockIdx.x == 0 && threadIdx.x < 4) {
        const size_t i = n4 + threadIdx.x;
        if (i < n) dst[i] = ew_apply<OP>(x[i], needs_y ? y[i] : 0.0f, s);
    }
}

template <int OP, int TILE_F, int STAGES>
constexpr size_t ew_bulk_smem_bytes() {
    return (size_t)STAGES * (ew_traits<OP>::needs_y ? 2 : 1) * TILE_F * 4 + (size_t)STAGES * 8;
}

// ---- generator, digest, check ---------------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ float hash_to_unit(uint64_t h) {
    return (float)(uint32_t)(h >> 40) * 0x1p-24f;          // exact: 24-bit integer times 2^-24
}

// X[i] = hash_to_unit(splitmix64(seed + i)).  Write-only, 4 bytes per element.
static __global__ void __launch_bounds__(256)
fill_hash_kernel(float *X, size_t n, uint64_t seed) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockDim.x * blockIdx.x + threadIdx.x; i < n; i += stride)
        X[i] = hash_to_unit(splitmix64(seed + (uint64_t)i));
}

// acc += sum_i mix64(w[i] + GOLDEN*(i+1)).  One 64-bit atomic per block after a warp-shuffle and
// shared-memory reduction.  `acc` is scratch the caller zeroes before the launch.
static __global__ void __launch_bounds__(256)
digest_kernel(const uint32_t *__restrict__ w, size_t n, unsigned long long *acc) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    unsigned long long local = 0;
    for (size_t i = (size_t)blockDim.x * blockIdx.x + threadIdx.x; i < n; i += stride)
        local += mix64((uint64_t)__ldg(w + i) + 0x9E3779B97F4A7C15ull * ((uint64_t)i + 1));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_down_sync(0xffffffffu, local, o);
    __shared__ unsigned long long warp_sum[8];
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = local;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += warp_sum[k];
        atomicAdd(acc, t);
    }
}

// Bitwise compare dst against op(x, y, s) recomputed one element per lane.  out[0] += number of
// differing elements, out[1] = min differing index (caller presets out = {0, ~0ull}).
template <int OP>
__global__ void __launch_bounds__(256)
check_kernel(const float *__restrict__ dst, const float *__restrict__ x, const float *__restrict__ y,
             float s, size_t n, unsigned long long *out) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    unsigned long long bad = 0, first = ~0ull;
    for (size_t i = (size_t)blockDim.x * blockIdx.x + threadIdx.x; i < n; i += stride) {
        const float want = ew_apply<OP>(x[i], ew_traits<OP>::needs_y ? y[i] : 0.0f, s);
        if (__float_as_uint(want) != __float_as_uint(dst[i])) {
            ++bad;
            if ((unsigned long long)i < first) first = i;
        }
    }
    if (__any_sync(0xffffffffu, bad != 0)) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            bad += __shfl_down_sync(0xffffffffu, bad, o);
            const unsigned long long f2 = __shfl_down_sync(0xffffffffu, first, o);
            first = f2 < first ? f2 : first;
        }
        if ((threadIdx.x & 31) == 0 && bad) {
            atomicAdd(out, bad);
            atomicMin(out + 1, first);
        }
    }
}

}  // namespace b200v
