// nvls_allreduce.cu — hand-written all-reduce kernels over NVLink/NVSwitch, as a second NVLink
// health probe beside NCCL, and the all-reduce of a caller-owned device buffer
// that the parity tests and bench.py compare with the closed form computed on the CPU.
//
// Every GPU holds one allocation of identical layout — barrier flags | data | scratch — bound to
// one multicast object (cuMulticastCreate / cuMulticastBindMem), so one address names "this offset
// on all GPUs", and mapped into every peer, so each GPU can also load from and store to any other
// GPU's copy directly.  Four kernels share that memory (one launch per GPU per all-reduce):
//
//   multimem (two-shot, switch-side reduction; default from 4 GPUs above the one-shot sizes)
//       GPU r owns slice r.  multimem.ld_reduce.add.v4.f32 makes the NVSwitch read that offset from
//       all N GPUs and return the sum; multimem.st.v4.f32 makes it write the sum back to all N.
//       Per GPU about (1 + 1/N) S bytes each way instead of a ring's 2(N-1)/N S.
//   one-shot (small messages)
//       every GPU multimem.ld_reduce's the WHOLE vector into its own output: N times the traffic,
//       but one cross-GPU barrier instead of two and no second pass — latency, not bandwidth.
//       Inputs are double-buffered (two slots used alternately), which is what makes the closing
//       barrier unnecessary: a slot is rewritten only two calls later, and a rank can only be in
//       call k+1 once every rank has finished reading in call k.
//   p2p (pull + push, plain peer loads and stores)
//       GPU r loads slice r from every GPU, adds in rank order, stores the sum to every GPU.
//   push (two-phase, stores only; default at 2 GPUs above the one-shot sizes)
//       phase 1: every GPU stores its copy of slice p into GPU p's scratch (no loads cross NVLink,
//       so nothing waits a round trip and no read requests compete with data); barrier;
//       phase 2: GPU r adds its own slice r and the N-1 copies in its scratch, in rank order, and
//       stores the sum to every GPU.  2(N-1)/N S per direction, all of it posted writes.
//
// A fifth kernel, hybrid, runs multimem and push on disjoint parts of one vector in one grid; it
// measured no faster than either and stays selectable for A/B only.
//
// Cross-GPU barriers use multimem.red.release.add on a per-CTA flag word that exists on every GPU
// (the release is the barrier's only fence), then spin on the local copy (bounded:
// B200V_NVLS_TIMEOUT_MS, default 2000 — a peer that never arrives yields B200V_E_TIMEOUT, never a
// hung GPU).  The count a CTA waits for lives in a small
// per-GPU array (`epochs`) that only that CTA index touches, so kernels with different grids can
// follow each other freely.  B200V_NVLS_ALGO=multimem|p2p|push|oneshot|hybrid forces one kernel.
//
// Everything from the CUDA driver API is resolved through cudaGetDriverEntryPoint, so the library
// keeps having no link-time dependency on libcuda.
#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda.h>   // types and enums; every driver function is fetched at run time
#include <new>
#include <poll.h>
#include <sys/socket.h>
#include <sys/uio.h>
#include <sys/un.h>
#include <unistd.h>
#include <thread>
#include <vector>

#include "ar_common.cuh"
#include "common.cuh"
#include "rendezvous.h"

namespace b200v {
namespace {

constexpr size_t kFlagBytes = 64 * 1024;      // barrier words live in front of the data
constexpr int kMaxCtas = 4096;
constexpr int kThreads = 512;
// One-shot messages are at most this long; the data region always holds two such input slots.
constexpr size_t kOneShotCap = (size_t)4 << 20;
constexpr int kMaxPeers = 8;

struct Drv {
    CUresult (*DeviceGet)(CUdevice *, int) = nullptr;
    CUresult (*DeviceGetAttribute)(int *, CUdevice_attribute, CUdevice) = nullptr;
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle *, const CUmulticastObjectProp *) = nullptr;
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t,
                                 size_t, unsigned long long) = nullptr;
    CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
    CUresult (*MulticastGetGranularity)(size_t *, const CUmulticastObjectProp *, CUmulticastGranularity_flags) = nullptr;
    CUresult (*MemCreate)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long) = nullptr;
    CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*MemAddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t) = nullptr;
    CUresult (*MemGetAllocationGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags) = nullptr;
    bool ok = false;
};

template <typename F>
bool drv_sym(const char *name, F &out) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult st;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &st) != cudaSuccess || st != cudaDriverEntryPointSuccess || !p) {
        (void)cudaGetLastError();
        return false;
    }
    out = reinterpret_cast<F>(p);
    return true;
}

const Drv &drv() {
    static Drv d = [] {
        Drv x;
        x.ok = drv_sym("cuDeviceGet", x.DeviceGet) && drv_sym("cuDeviceGetAttribute", x.DeviceGetAttribute) &&
               drv_sym("cuMulticastCreate", x.MulticastCreate) && drv_sym("cuMulticastAddDevice", x.MulticastAddDevice) &&
               drv_sym("cuMulticastBindMem", x.MulticastBindMem) && drv_sym("cuMulticastUnbind", x.MulticastUnbind) &&
               drv_sym("cuMulticastGetGranularity", x.MulticastGetGranularity) && drv_sym("cuMemCreate", x.MemCreate) &&
               drv_sym("cuMemRelease", x.MemRelease) && drv_sym("cuMemAddressReserve", x.MemAddressReserve) &&
               drv_sym("cuMemAddressFree", x.MemAddressFree) && drv_sym("cuMemMap", x.MemMap) &&
               drv_sym("cuMemUnmap", x.MemUnmap) && drv_sym("cuMemSetAccess", x.MemSetAccess) &&
               drv_sym("cuMemGetAllocationGranularity", x.MemGetAllocationGranularity);
        return x;
    }();
    return d;
}

// ---- device side -------------------------------------------------------------------------------------

__device__ __forceinline__ unsigned long long now_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// What every kernel needs to find the shared allocation and to synchronise with its peers.
struct ArArgs {
    unsigned char *mc_base;          // this GPU's mapping of the multicast object (flags at offset 0)
    unsigned char *uc_base;          // this GPU's own memory, same layout
    unsigned *epochs;                // per-CTA barrier counts already consumed (private to this GPU)
    int *err;                        // set to 1 / 2 when the first / second barrier timed out
    unsigned long long timeout_ns;
    size_t data_off, scratch_off;    // byte offsets of the data and scratch regions
    int rank, nranks;
};

struct PeerPtrs {
    float *p[kMaxPeers];
};

// All CTAs with this blockIdx on all GPUs meet here.  mc_flag names the flag word on every GPU,
// uc_flag is this GPU's own copy.  Returns false on timeout.
__device__ __forceinline__ bool gpu_barrier(unsigned *mc_flag, unsigned *uc_flag, unsigned target,
                                            unsigned long long timeout_ns) {
    __shared__ int ok_s;
    __syncthreads();
    if (threadIdx.x == 0) {
        // The CTA's stores were ordered before this point by the bar.sync above; the release on the
        // arrival (a system-scope fence followed by the reduction) publishes them, cumulatively,
        // before any peer can observe the count.  No separate __threadfence_system() is needed.
        asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" :: "l"(mc_flag), "r"(1u) : "memory");
        const unsigned long long t0 = now_ns();
        int ok = 1;
        unsigned polls = 0;
        for (;;) {
            unsigned v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(uc_flag) : "memory");
            if ((int)(v - target) >= 0) break;
            if ((++polls & 1023u) == 0 && now_ns() - t0 > timeout_ns) { ok = 0; break; }
        }
        ok_s = ok;
    }
    __syncthreads();
    return ok_s != 0;
}

// Opening barrier of a kernel that has not written anything yet (its inputs were produced by earlier
// work on the stream, complete at the kernel boundary): the arrival is sent first, with nothing to
// wait for on this side, and the CTA's epoch word is fetched while it is in flight.  Returns the
// epoch the CTA started from through *epoch_out; false on timeout.
__device__ __forceinline__ bool gpu_barrier_open(const unsigned *epochs, unsigned *mc_flag, unsigned *uc_flag,
                                                 unsigned nranks, unsigned long long timeout_ns, unsigned *epoch_out) {
    __shared__ int ok_o;
    __shared__ unsigned epoch_o;
    if (threadIdx.x == 0) {
        asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" :: "l"(mc_flag), "r"(1u) : "memory");
        const unsigned epoch = epochs[blockIdx.x];
        const unsigned target = epoch + nranks;
        const unsigned long long t0 = now_ns();
        int ok = 1;
        unsigned polls = 0;
        for (;;) {
            unsigned v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(uc_flag) : "memory");
            if ((int)(v - target) >= 0) break;
            if ((++polls & 1023u) == 0 && now_ns() - t0 > timeout_ns) { ok = 0; break; }
        }
        ok_o = ok;
        epoch_o = epoch;
    }
    __syncthreads();
    *epoch_out = epoch_o;
    return ok_o != 0;
}

// The barrier count this CTA index has consumed so far (same value on every GPU, because every
// GPU runs the same sequence of launches).
__device__ __forceinline__ unsigned load_epoch(const ArArgs &a) {
    __shared__ unsigned epoch_s;
    if (threadIdx.x == 0) epoch_s = a.epochs[blockIdx.x];
    __syncthreads();
    return epoch_s;
}
__device__ __forceinline__ void store_epoch(const ArArgs &a, unsigned e) {
    if (threadIdx.x == 0) a.epochs[blockIdx.x] = e;
}

__device__ __forceinline__ float4 mm_ld_reduce(const float *mc) {
    float4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc) : "memory");
    return r;
}
__device__ __forceinline__ void mm_st(float *mc, const float4 &r) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(mc), "f"(r.x), "f"(r.y), "f"(r.z), "f"(r.w) : "memory");
}

// STRONG = .relaxed.sys accesses; otherwise ordinary weak ones.  Weak is safe for data that was
// complete before this kernel's opening barrier (L1 is invalidated at launch and nothing in the
// kernel has read those lines before); data a peer writes DURING the kernel is read .relaxed.sys.
template <bool STRONG>
__device__ __forceinline__ float4 ld_f4(const float *p) {
    float4 r;
    if constexpr (STRONG)
        asm volatile("ld.global.relaxed.sys.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
    else
        asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
    return r;
}
template <bool STRONG>
__device__ __forceinline__ void st_f4(float *p, const float4 &r) {
    if constexpr (STRONG)
        asm volatile("st.global.relaxed.sys.v4.f32 [%0], {%1,%2,%3,%4};"
                     :: "l"(p), "f"(r.x), "f"(r.y), "f"(r.z), "f"(r.w) : "memory");
    else
        asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};"
                     :: "l"(p), "f"(r.x), "f"(r.y), "f"(r.z), "f"(r.w) : "memory");
}

__device__ __forceinline__ void add4(float4 &s, const float4 &v) {
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
}

// ---- multimem, two-shot: in-place sum of `len` float4 vectors starting at vector `vbase` of the data
// region.  `cta` of `nctas` is this CTA's share of the work; the barrier flag and the epoch word
// are always the ones of blockIdx.x, so several bodies can share one grid (hybrid kernel below).
template <int UNROLL>
__device__ __forceinline__ void multimem_body(const ArArgs &a, size_t vbase, size_t len, unsigned cta, unsigned nctas) {
    unsigned *mc_flag = reinterpret_cast<unsigned *>(a.mc_base) + blockIdx.x;
    unsigned *uc_flag = reinterpret_cast<unsigned *>(a.uc_base) + blockIdx.x;
    float *mc = reinterpret_cast<float *>(a.mc_base + a.data_off) + 4 * vbase;
    const unsigned nr = (unsigned)a.nranks;

    // (1) every GPU has entered this kernel, so every GPU's input (produced earlier on its stream)
    //     is complete and visible
    unsigned epoch;
    if (!gpu_barrier_open(a.epochs, mc_flag, uc_flag, nr, a.timeout_ns, &epoch)) {
        if (threadIdx.x == 0) atomicExch(a.err, 1);
        return;
    }
    const size_t per = (len + (size_t)nr - 1) / (size_t)nr;
    const size_t lo = per * (size_t)a.rank < len ? per * (size_t)a.rank : len;
    const size_t hi = lo + per < len ? lo + per : len;
    // CTA b walks tiles b, b+nctas, ... of its GPU's slice; a tile is UNROLL x kThreads float4s, and a
    // thread issues all UNROLL switch-side reductions before the first broadcast store.
    constexpr size_t TILE = (size_t)UNROLL * kThreads;
    for (size_t t0 = lo + (size_t)cta * TILE; t0 < hi; t0 += (size_t)nctas * TILE) {
        float4 acc[UNROLL];
        if (t0 + TILE <= hi) {
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) acc[j] = mm_ld_reduce(mc + 4 * (t0 + (size_t)j * kThreads + threadIdx.x));
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) mm_st(mc + 4 * (t0 + (size_t)j * kThreads + threadIdx.x), acc[j]);
        } else {
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) {
                const size_t v = t0 + (size_t)j * kThreads + threadIdx.x;
                if (v < hi) mm_st(mc + 4 * v, mm_ld_reduce(mc + 4 * v));
            }
        }
    }

    // (2) nobody leaves before every GPU's stores have landed everywhere
    if (!gpu_barrier(mc_flag, uc_flag, epoch + 2u * nr, a.timeout_ns)) {
        if (threadIdx.x == 0) atomicExch(a.err, 2);
    }
    store_epoch(a, epoch + 2u * nr);
}

template <int UNROLL>
__global__ void __launch_bounds__(kThreads)
nvls_allreduce_kernel(const __grid_constant__ ArArgs a, size_t n4) {
    multimem_body<UNROLL>(a, 0, n4, blockIdx.x, gridDim.x);
}

// ---- one-shot: out[v] = sum over GPUs of in[v], every GPU computing the whole vector ------------------
// `in_off` is the byte offset of the input slot inside the allocation (the same on every GPU);
// `out` is local memory.  One barrier: see the file comment for why none is needed at the end.
__global__ void __launch_bounds__(kThreads)
oneshot_allreduce_kernel(ArArgs a, size_t in_off, float *__restrict__ out, size_t n4) {
    unsigned *mc_flag = reinterpret_cast<unsigned *>(a.mc_base) + blockIdx.x;
    unsigned *uc_flag = reinterpret_cast<unsigned *>(a.uc_base) + blockIdx.x;
    const float *mc = reinterpret_cast<const float *>(a.mc_base + in_off);
    // gpu_barrier_open: this kernel has written nothing before its only barrier.
    unsigned epoch;
    const bool ok = gpu_barrier_open(a.epochs, mc_flag, uc_flag, (unsigned)a.nranks, a.timeout_ns, &epoch);
    store_epoch(a, epoch + (unsigned)a.nranks);
    if (!ok) {
        if (threadIdx.x == 0) atomicExch(a.err, 1);
        return;
    }
    for (size_t v = (size_t)blockIdx.x * kThreads + threadIdx.x; v < n4; v += (size_t)gridDim.x * kThreads)
        st_f4<false>(out + 4 * v, mm_ld_reduce(mc + 4 * v));
}

// ---- p2p: plain peer loads and stores, pull the slice from everyone, push the sum to everyone ---------
// GPU r reads slice r from every GPU (N-1 of them across NVLink), adds in rank order, and stores the
// sum into slice r of every GPU: 2(N-1)/N x S per direction, one hop, no per-step synchronisation.
template <int NR, int UNROLL, bool STRONG>
__global__ void __launch_bounds__(kThreads)
p2p_allreduce_kernel(const __grid_constant__ ArArgs a, const __grid_constant__ PeerPtrs data, size_t n4) {
    unsigned *mc_flag = reinterpret_cast<unsigned *>(a.mc_base) + blockIdx.x;
    unsigned *uc_flag = reinterpret_cast<unsigned *>(a.uc_base) + blockIdx.x;
    unsigned epoch;
    if (!gpu_barrier_open(a.epochs, mc_flag, uc_flag, (unsigned)NR, a.timeout_ns, &epoch)) {
        if (threadIdx.x == 0) atomicExch(a.err, 1);
        return;
    }
    const size_t per = (n4 + (size_t)NR - 1) / (size_t)NR;
    const size_t lo = per * (size_t)a.rank < n4 ? per * (size_t)a.rank : n4;
    const size_t hi = lo + per < n4 ? lo + per : n4;
    constexpr size_t TILE = (size_t)UNROLL * kThreads;
    for (size_t t0 = lo + (size_t)blockIdx.x * TILE; t0 < hi; t0 += (size_t)gridDim.x * TILE) {
        float4 in[UNROLL][NR];
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) {
            const size_t v = t0 + (size_t)j * kThreads + threadIdx.x;
            if (v < hi) {
#pragma unroll
                for (int p = 0; p < NR; ++p) in[j][p] = ld_f4<STRONG>(data.p[p] + 4 * v);
            }
        }
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) {
            const size_t v = t0 + (size_t)j * kThreads + threadIdx.x;
            if (v < hi) {
                float4 s = in[j][0];
#pragma unroll
                for (int p = 1; p < NR; ++p) add4(s, in[j][p]);
#pragma unroll
                for (int p = 0; p < NR; ++p) st_f4<STRONG>(data.p[p] + 4 * v, s);
            }
        }
    }
    if (!gpu_barrier(mc_flag, uc_flag, epoch + 2u * (unsigned)NR, a.timeout_ns)) {
        if (threadIdx.x == 0) atomicExch(a.err, 2);
    }
    store_epoch(a, epoch + 2u * (unsigned)NR);
}

// ---- push: two phases of posted stores, nothing is loaded across NVLink -------------------------------
// Slot of sender s in receiver p's scratch (p never sends to itself): N-1 slots of `per` vectors.
__device__ __forceinline__ int push_slot(int s, int p) { return s < p ? s : s - 1; }

template <int NR, int UNROLL>
__device__ __forceinline__ void push_body(const ArArgs &a, const PeerPtrs &data, const PeerPtrs &scratch, size_t vbase,
                                          size_t len, unsigned cta, unsigned nctas) {
    unsigned *mc_flag = reinterpret_cast<unsigned *>(a.mc_base) + blockIdx.x;
    unsigned *uc_flag = reinterpret_cast<unsigned *>(a.uc_base) + blockIdx.x;
    const unsigned epoch = load_epoch(a);
    const int rank = a.rank;
    const size_t per = (len + (size_t)NR - 1) / (size_t)NR;
    const float *mine = data.p[rank] + 4 * vbase;
    constexpr size_t TILE = (size_t)UNROLL * kThreads;

    // Phase 1 needs no opening barrier: it reads only this GPU's own input (complete, by stream
    // order) and writes only peers' scratch, which every GPU finished reading before anyone could
    // leave the previous launch.  CTA b handles the same offsets of every slice in both phases,
    // so the per-CTA barrier below is all the ordering phase 2 needs.
    for (size_t t0 = (size_t)cta * TILE; t0 < per; t0 += (size_t)nctas * TILE) {
        float4 v[UNROLL][NR - 1];
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) {
            const size_t o = t0 + (size_t)j * kThreads + threadIdx.x;
#pragma unroll
            for (int q = 1; q < NR; ++q) {
                const int p = (rank + q) % NR;
                const size_t idx = (size_t)p * per + o;
                if (o < per && idx < len) v[j][q - 1] = ld_f4<false>(mine + 4 * idx);
            }
        }
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) {
            const size_t o = t0 + (size_t)j * kThreads + threadIdx.x;
#pragma unroll
            for (int q = 1; q < NR; ++q) {
                const int p = (rank + q) % NR;
                const size_t idx = (size_t)p * per + o;
                if (o < per && idx < len)
                    st_f4<true>(scratch.p[p] + 4 * ((size_t)push_slot(rank, p) * per + o), v[j][q - 1]);
            }
        }
    }
    if (!gpu_barrier(mc_flag, uc_flag, epoch + (unsigned)NR, a.timeout_ns)) {
        if (threadIdx.x == 0) atomicExch(a.err, 1);
        return;
    }
    // Phase 2: own slice + the N-1 copies that arrived, added in rank order, stored to every GPU.
    const float *inbox = scratch.p[rank];
    for (size_t t0 = (size_t)cta * TILE; t0 < per; t0 += (size_t)nctas * TILE) {
        float4 in[UNROLL][NR];
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) {
            const size_t o = t0 + (size_t)j * kThreads + threadIdx.x;
            const size_t idx = (size_t)rank * per + o;
            if (o < per && idx < len) {
#pragma unroll
                for (int s = 0; s < NR; ++s)
                    in[j][s] = s == rank ? ld_f4<false>(mine + 4 * idx)
                                         : ld_f4<true>(inbox + 4 * ((size_t)push_slot(s, rank) * per + o));
            }
        }
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) {
            const size_t o = t0 + (size_t)j * kThreads + threadIdx.x;
            const size_t idx = (size_t)rank * per + o;
            if (o < per && idx < len) {
                float4 s = in[j][0];
#pragma unroll
                for (int r = 1; r < NR; ++r) add4(s, in[j][r]);
#pragma unroll
                for (int p = 0; p < NR; ++p) st_f4<true>(data.p[p] + 4 * (vbase + idx), s);
            }
        }
    }
    if (!gpu_barrier(mc_flag, uc_flag, epoch + 2u * (unsigned)NR, a.timeout_ns)) {
        if (threadIdx.x == 0) atomicExch(a.err, 2);
    }
    store_epoch(a, epoch + 2u * (unsigned)NR);
}

template <int NR, int UNROLL>
__global__ void __launch_bounds__(kThreads)
push_allreduce_kernel(const __grid_constant__ ArArgs a, const __grid_constant__ PeerPtrs data,
                      const __grid_constant__ PeerPtrs scratch, size_t n4) {
    push_body<NR, UNROLL>(a, data, scratch, 0, n4, blockIdx.x, gridDim.x);
}

// ---- hybrid: both of the above in one grid, on disjoint parts of the vector ----------------------------
// The switch-side reduction tops out near 540-570 GB/s per direction per GPU and SM-issued stores
// near 700, and neither fills the link alone.  CTAs [0, mm_ctas) reduce vectors [0, m4) through the
// switch while the remaining CTAs push-reduce vectors [m4, n4): two independent all-reduces that
// share one launch and overlap on the wire.  Every CTA keeps its own barrier flag (blockIdx.x),
// so the two halves never wait on each other.
template <int NR, int UM, int UP>
__global__ void __launch_bounds__(kThreads)
hybrid_allreduce_kernel(const __grid_constant__ ArArgs a, const __grid_constant__ PeerPtrs data,
                        const __grid_constant__ PeerPtrs scratch, size_t n4, size_t m4, unsigned mm_ctas) {
    if (blockIdx.x < mm_ctas) multimem_body<UM>(a, 0, m4, blockIdx.x, mm_ctas);
    else push_body<NR, UP>(a, data, scratch, m4, n4 - m4, blockIdx.x - mm_ctas, gridDim.x - mm_ctas);
}

// ---- host side ------------------------------------------------------------------------------------------

struct NvlsRank {
    int dev = -1;
    int rank = -1;          // rank in the collective
    CUdevice cudev = 0;
    CUmemGenericAllocationHandle mem = 0;
    CUdeviceptr uc = 0, mc = 0;
    bool bound = false, uc_mapped = false, mc_mapped = false;
    cudaStream_t st = nullptr;
    unsigned long long *d_bad = nullptr;
    int *d_err = nullptr;
    unsigned *d_epochs = nullptr;
    int sm_count = 0;
    std::vector<cudaEvent_t> ev;
    // Address of every rank's allocation as seen from this rank's process (index = rank); own
    // entry = uc.  One process: the other ranks' own mappings.  One process per GPU: their
    // allocations imported and mapped here.  Empty when unavailable (more ranks than kMaxPeers).
    std::vector<CUdeviceptr> peer_base;
};

struct NvlsTeam {
    std::vector<NvlsRank> ranks;         // the ranks THIS process drives
    int nranks = 0;
    CUmemGenericAllocationHandle mc_handle = 0;
    size_t size = 0;                     // whole allocation: flags | data | scratch
    size_t data_cap = 0;                 // bytes in the data region (and in the scratch region)
    bool have_mc = false;
    bool peers_mapped = false;
    std::vector<CUmemGenericAllocationHandle> peer_mem;   // imported peer allocations (per-rank form)
    std::vector<CUdeviceptr> peer_va;
    unsigned long oneshot_calls = 0;     // parity of the one-shot input slot (same on every rank)
};

#define B200V_CU(expr)                                   \
    do {                                                 \
        CUresult r__ = (expr);                           \
        if (r__ != CUDA_SUCCESS) {                       \
            g_last_cu = (int)r__;                        \
            std::fprintf(stderr, "b200v: CUDA driver call failed with CUresult %d at %s:%d: %s\n", (int)r__, \
                         "nvls_allreduce.cu", __LINE__, #expr);                                             \
            return r__ == CUDA_ERROR_NOT_SUPPORTED ? B200V_E_UNSUPPORTED : B200V_E_DRIVER; \
        }                                                \
    } while (0)

int g_last_cu = 0;

size_t data_cap_for(size_t data_bytes) {
    const size_t want = std::max(data_bytes, 2 * kOneShotCap);
    return (want + 255) / 256 * 256;
}

void destroy_team(NvlsTeam &t) {
    const Drv &d = drv();
    for (NvlsRank &r : t.ranks) {
        if (r.dev < 0) continue;
        cudaSetDevice(r.dev);
        if (r.st) cudaStreamSynchronize(r.st);
        cudaDeviceSynchronize();     // callers' streams may still hold work that touches the mappings
    }
    for (CUdeviceptr va : t.peer_va) {
        d.MemUnmap(va, t.size);
        d.MemAddressFree(va, t.size);
    }
    for (CUmemGenericAllocationHandle h : t.peer_mem) d.MemRelease(h);
    t.peer_va.clear();
    t.peer_mem.clear();
    for (NvlsRank &r : t.ranks) {
        if (r.dev < 0) continue;
        cudaSetDevice(r.dev);
        if (r.mc_mapped) { d.MemUnmap(r.mc, t.size); }
        if (r.mc) d.MemAddressFree(r.mc, t.size);
        if (r.bound) d.MulticastUnbind(t.mc_handle, r.cudev, 0, t.size);
        if (r.uc_mapped) d.MemUnmap(r.uc, t.size);
        if (r.uc) d.MemAddressFree(r.uc, t.size);
        if (r.mem) d.MemRelease(r.mem);
        for (cudaEvent_t e : r.ev) cudaEventDestroy(e);
        if (r.d_bad) cudaFree(r.d_bad);
        if (r.d_err) cudaFree(r.d_err);
        if (r.d_epochs) cudaFree(r.d_epochs);
        if (r.st) cudaStreamDestroy(r.st);
    }
    if (t.have_mc) d.MemRelease(t.mc_handle);
    t.ranks.clear();
}

// Per-rank objects that need no peer: stream, counters, events.
int init_rank_local(NvlsRank &r, int iters) {
    const Drv &d = drv();
    B200V_CUDA(cudaSetDevice(r.dev));
    B200V_CUDA(cudaFree(nullptr));   // make sure the primary context exists
    B200V_CU(d.DeviceGet(&r.cudev, r.dev));
    int mc_ok = 0;
    B200V_CU(d.DeviceGetAttribute(&mc_ok, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, r.cudev));
    if (!mc_ok) return B200V_E_UNSUPPORTED;
    B200V_CUDA(cudaDeviceGetAttribute(&r.sm_count, cudaDevAttrMultiProcessorCount, r.dev));
    B200V_CUDA(cudaStreamCreateWithFlags(&r.st, cudaStreamNonBlocking));
    B200V_CUDA(cudaMalloc(&r.d_bad, 8));
    B200V_CUDA(cudaMalloc(&r.d_err, 4));
    B200V_CUDA(cudaMemset(r.d_err, 0, 4));
    B200V_CUDA(cudaMalloc(&r.d_epochs, kMaxCtas * sizeof(unsigned)));
    B200V_CUDA(cudaMemset(r.d_epochs, 0, kMaxCtas * sizeof(unsigned)));
    r.ev.resize((size_t)iters * 2);
    for (auto &e : r.ev) B200V_CUDA(cudaEventCreate(&e));
    return B200V_OK;
}

int team_geometry(NvlsTeam &t, int nranks, size_t data_bytes, CUmulticastObjectProp &mprop,
                  CUmemAllocationProp &aprop, CUdevice first, bool shareable, size_t &gran) {
    const Drv &d = drv();
    t.nranks = nranks;
    t.data_cap = data_cap_for(data_bytes);
    std::memset(&mprop, 0, sizeof mprop);
    mprop.numDevices = (unsigned)nranks;
    mprop.size = kFlagBytes + 2 * t.data_cap;
    if (shareable) mprop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    gran = 0;
    B200V_CU(d.MulticastGetGranularity(&gran, &mprop, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    std::memset(&aprop, 0, sizeof aprop);
    aprop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    aprop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    aprop.location.id = first;
    if (shareable) aprop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;   // same kind as the multicast object
    size_t agran = 0;
    B200V_CU(d.MemGetAllocationGranularity(&agran, &aprop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    gran = std::max(gran, agran);
    t.size = (mprop.size + gran - 1) / gran * gran;     // every rank computes the same size
    mprop.size = t.size;
    return B200V_OK;
}

// Allocate this rank's memory, bind it to the multicast object, map both views, zero the flags.
int bind_and_map(NvlsTeam &t, NvlsRank &r, CUmemAllocationProp aprop, size_t gran,
                 const std::vector<CUdevice> &readers) {
    const Drv &d = drv();
    B200V_CUDA(cudaSetDevice(r.dev));
    aprop.location.id = r.cudev;
    B200V_CU(d.MemCreate(&r.mem, t.size, &aprop, 0));
    B200V_CU(d.MulticastBindMem(t.mc_handle, 0, r.mem, 0, t.size, 0));
    r.bound = true;
    CUmemAccessDesc acc;
    std::memset(&acc, 0, sizeof acc);
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = r.cudev;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    B200V_CU(d.MemAddressReserve(&r.uc, t.size, gran, 0, 0));
    B200V_CU(d.MemMap(r.uc, t.size, 0, r.mem, 0));
    r.uc_mapped = true;
    // every GPU listed in `readers` may load from and store to this mapping (one process: all of
    // the team, for the peer kernels; one process per GPU: just the owner — peers map their own view)
    std::vector<CUmemAccessDesc> all(readers.size(), acc);
    for (size_t j = 0; j < readers.size(); ++j) all[j].location.id = readers[j];
    B200V_CU(d.MemSetAccess(r.uc, t.size, all.data(), all.size()));
    B200V_CU(d.MemAddressReserve(&r.mc, t.size, gran, 0, 0));
    B200V_CU(d.MemMap(r.mc, t.size, 0, t.mc_handle, 0));
    r.mc_mapped = true;
    B200V_CU(d.MemSetAccess(r.mc, t.size, &acc, 1));
    B200V_CUDA(cudaMemsetAsync((void *)r.uc, 0, kFlagBytes, r.st));
    B200V_CUDA(cudaStreamSynchronize(r.st));
    return B200V_OK;
}

// One process drives every GPU of the team.
int build_team(NvlsTeam &t, int ndev, const int *devs, size_t data_bytes, int iters) {
    const Drv &d = drv();
    if (!d.ok) return B200V_E_UNSUPPORTED;
    t.ranks.resize((size_t)ndev);
    for (int i = 0; i < ndev; ++i) {
        NvlsRank &r = t.ranks[(size_t)i];
        r.dev = devs[i];
        r.rank = i;
        int rc = init_rank_local(r, iters);
        if (rc != B200V_OK) return rc;
    }
    CUmulticastObjectProp mprop;
    CUmemAllocationProp aprop;
    size_t gran = 0;
    int rc = team_geometry(t, ndev, data_bytes, mprop, aprop, t.ranks[0].cudev, false, gran);
    if (rc != B200V_OK) return rc;
    B200V_CU(d.MulticastCreate(&t.mc_handle, &mprop));
    t.have_mc = true;
    for (NvlsRank &r : t.ranks) B200V_CU(d.MulticastAddDevice(t.mc_handle, r.cudev));
    std::vector<CUdevice> everyone;
    for (NvlsRank &r : t.ranks) everyone.push_back(r.cudev);
    for (NvlsRank &r : t.ranks) {
        rc = bind_and_map(t, r, aprop, gran, everyone);
        if (rc != B200V_OK) return rc;
    }
    if (ndev <= kMaxPeers) {
        for (NvlsRank &r : t.ranks)
            for (NvlsRank &o : t.ranks) r.peer_base.push_back(o.uc);
        t.peers_mapped = true;
    }
    return B200V_OK;
}

// ---- one process per GPU (see rendezvous.h for how the processes meet) --------------------------------
// This process's one rank of an nranks-wide team whose other ranks live in other processes.
int build_team_rank(NvlsTeam &t, Rendezvous &rv, int dev, int rank, int nranks, size_t data_bytes, int iters) {
    const Drv &d = drv();
    CUresult (*Export)(void *, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
    CUresult (*Import)(CUmemGenericAllocationHandle *, void *, CUmemAllocationHandleType) = nullptr;
    if (!d.ok || !drv_sym("cuMemExportToShareableHandle", Export) || !drv_sym("cuMemImportFromShareableHandle", Import))
        return B200V_E_UNSUPPORTED;
    t.ranks.resize(1);
    NvlsRank &r = t.ranks[0];
    r.dev = dev;
    r.rank = rank;
    int rc = init_rank_local(r, iters);
    if (rc != B200V_OK) return rc;
    CUmulticastObjectProp mprop;
    CUmemAllocationProp aprop;
    size_t gran = 0;
    rc = team_geometry(t, nranks, data_bytes, mprop, aprop, r.cudev, true, gran);
    if (rc != B200V_OK) return rc;

    int fd = -1;
    if (rank == 0) {
        B200V_CU(d.MulticastCreate(&t.mc_handle, &mprop));
        t.have_mc = true;
        B200V_CU(Export(&fd, t.mc_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    }
    const bool shared = rv.share_fd(fd);
    if (rank == 0) {
        if (fd >= 0) ::close(fd);
    } else if (shared) {
        CUresult ir = Import(&t.mc_handle, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
        ::close(fd);
        B200V_CU(ir);
        t.have_mc = true;
    }
    if (!shared) return B200V_E_TIMEOUT;
    B200V_CU(d.MulticastAddDevice(t.mc_handle, r.cudev));
    if (!rv.barrier()) return B200V_E_TIMEOUT;          // every device is in the team before anyone binds

    rc = bind_and_map(t, r, aprop, gran, {r.cudev});
    if (rc != B200V_OK) return rc;

    // For the peer kernels every rank also maps every other rank's allocation: each exports its
    // allocation as a descriptor, rank 0 relays them all around.
    if (nranks <= kMaxPeers) {
        CUmemAccessDesc acc;
        std::memset(&acc, 0, sizeof acc);
        acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc.location.id = r.cudev;
        acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        int my_fd = -1;
        B200V_CU(Export(&my_fd, r.mem, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
        std::vector<int> fds;
        const bool ok = rv.exchange_fds(my_fd, fds);
        rc = ok ? B200V_OK : B200V_E_TIMEOUT;
        std::vector<CUdeviceptr> base((size_t)nranks, 0);
        base[(size_t)rank] = r.uc;
        for (int k = 0; ok && k < nranks && rc == B200V_OK; ++k) {
            if (k == rank) continue;
            CUmemGenericAllocationHandle h = 0;
            CUdeviceptr va = 0;
            if (Import(&h, (void *)(uintptr_t)fds[(size_t)k], CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) != CUDA_SUCCESS) {
                rc = B200V_E_DRIVER;
                break;
            }
            t.peer_mem.push_back(h);
            if (d.MemAddressReserve(&va, t.size, gran, 0, 0) != CUDA_SUCCESS) { rc = B200V_E_DRIVER; break; }
            if (d.MemMap(va, t.size, 0, h, 0) != CUDA_SUCCESS) {
                d.MemAddressFree(va, t.size);
                rc = B200V_E_DRIVER;
                break;
            }
            t.peer_va.push_back(va);
            if (d.MemSetAccess(va, t.size, &acc, 1) != CUDA_SUCCESS) { rc = B200V_E_DRIVER; break; }
            base[(size_t)k] = va;
        }
        for (int k = 0; k < (int)fds.size(); ++k)
            if (fds[(size_t)k] >= 0) ::close(fds[(size_t)k]);      // includes my_fd (fds[rank])
        if (fds.empty() && my_fd >= 0) ::close(my_fd);
        if (rc != B200V_OK) return rc;
        r.peer_base = base;
        t.peers_mapped = true;
    }
    if (!rv.barrier()) return B200V_E_TIMEOUT;          // every flag word is zero before the first launch
    return B200V_OK;
}

int grid_fill(const NvlsRank &r, size_t n) {
    size_t g = (n + 255) / 256;
    const size_t cap = (size_t)r.sm_count * 8;
    if (g > cap) g = cap;
    return g < 1 ? 1 : (int)g;
}

// ---- algorithm choice and launch ------------------------------------------------------------------------

enum Algo { ALGO_MULTIMEM = 0, ALGO_P2P = 1, ALGO_PUSH = 2, ALGO_ONESHOT = 3, ALGO_HYBRID = 4 };

const char *algo_name(int a) {
    switch (a) {
    case ALGO_MULTIMEM: return "multimem";
    case ALGO_P2P: return "p2p";
    case ALGO_PUSH: return "push";
    case ALGO_HYBRID: return "hybrid";
    default: return "oneshot";
    }
}

struct Tuning {
    int forced = -1;                 // B200V_NVLS_ALGO
    int ctas = 0;                    // B200V_NVLS_CTAS (0 = per-algorithm default)
    int unroll = 0;                  // B200V_NVLS_UNROLL (0 = per-algorithm default)
    size_t oneshot_max = 0;          // B200V_NVLS_ONESHOT_MAX bytes (messages up to this go one-shot)
    int hybrid_pct = 0;              // B200V_NVLS_HYBRID_PCT: share of the vector reduced in the switch (0 = default)
    int hybrid_mm_ctas = 0;          // B200V_NVLS_HYBRID_MM_CTAS: CTAs on the switch-side part (0 = default)
    unsigned long long timeout_ns = 2000ull * 1000000ull;
};

Tuning read_tuning() {
    Tuning t;
    if (const char *e = std::getenv("B200V_NVLS_ALGO")) {
        if (!std::strcmp(e, "multimem")) t.forced = ALGO_MULTIMEM;
        else if (!std::strcmp(e, "p2p")) t.forced = ALGO_P2P;
        else if (!std::strcmp(e, "push")) t.forced = ALGO_PUSH;
        else if (!std::strcmp(e, "oneshot")) t.forced = ALGO_ONESHOT;
        else if (!std::strcmp(e, "hybrid")) t.forced = ALGO_HYBRID;
    }
    if (const char *e = std::getenv("B200V_NVLS_HYBRID_PCT")) {
        const int v = std::atoi(e);
        if (v >= 1 && v <= 99) t.hybrid_pct = v;
    }
    if (const char *e = std::getenv("B200V_NVLS_HYBRID_MM_CTAS")) {
        const int v = std::atoi(e);
        if (v >= 1 && v <= 128) t.hybrid_mm_ctas = v;
    }
    if (const char *e = std::getenv("B200V_NVLS_CTAS")) {
        const int v = std::atoi(e);
        if (v >= 1 && v <= kMaxCtas) t.ctas = v;
    }
    if (const char *e = std::getenv("B200V_NVLS_UNROLL")) t.unroll = std::atoi(e);
    t.oneshot_max = 0;               // 0 = by rank count, see choose_algo
    if (const char *e = std::getenv("B200V_NVLS_ONESHOT_MAX")) t.oneshot_max = (size_t)std::atoll(e);
    if (t.oneshot_max > kOneShotCap) t.oneshot_max = kOneShotCap;
    if (const char *e = std::getenv("B200V_NVLS_TIMEOUT_MS")) {
        const long v = std::atol(e);
        if (v > 0) t.timeout_ns = (unsigned long long)v * 1000000ull;
    }
    return t;
}

bool peer_algo_compiled(int nranks) { return nranks == 2 || nranks == 4 || nranks == 8; }

// Which kernel reduces a message of `bytes` bytes over `nranks` GPUs.  Every rank evaluates this
// with the same arguments and the same environment, so all ranks launch the same kernel.
int choose_algo(const Tuning &tn, const NvlsTeam &t, size_t bytes) {
    const int n = t.nranks;
    const bool peers = t.peers_mapped && peer_algo_compiled(n);
    if (tn.forced == ALGO_ONESHOT) return bytes <= kOneShotCap ? ALGO_ONESHOT : ALGO_MULTIMEM;
    if (tn.forced == ALGO_MULTIMEM) return ALGO_MULTIMEM;
    if (tn.forced == ALGO_P2P) return peers ? ALGO_P2P : ALGO_MULTIMEM;
    if (tn.forced == ALGO_PUSH) return peers ? ALGO_PUSH : ALGO_MULTIMEM;
    if (tn.forced == ALGO_HYBRID) return (peers && n >= 4) ? ALGO_HYBRID : ALGO_MULTIMEM;
    // One-shot moves N times the bytes but saves a barrier and a pass: it wins while the message is
    // latency-bound.  Measured crossovers with the lean barriers (profiles/README.md, us per call):
    //   2 GPUs  1 MiB 13.9 vs 15.9 (push), 2 MiB 18.5 vs 19.2, 4 MiB 28.3 vs 20.6   -> 2 MiB
    //   4 GPUs  1 MiB 16.4 vs 17.5 (multimem), 2 MiB 23.3 vs 19.5                   -> 1 MiB
    //   8 GPUs  512 KiB 16.0 vs 16.9 (multimem), 1 MiB 22.4 vs 18.0                 -> 512 KiB
    const size_t oneshot_max = tn.oneshot_max ? tn.oneshot_max
                             : (n == 2 ? (size_t)2 << 20 : n <= 4 ? (size_t)1 << 20 : (size_t)512 << 10);
    if (bytes <= oneshot_max) return ALGO_ONESHOT;
    // Large messages: with 2 GPUs the switch has nothing to reduce, so the peer-store kernel wins
    // (694 vs 408 GB/s at 1 GiB).  From 4 GPUs the switch-side reduction moves fewer bytes per GPU:
    // at 4 it ties push at 1 GiB (680-688 both) and is 2-6 % ahead from 2 to 16 MiB; at 8 it is far
    // ahead (842 GB/s).
    if (n == 2 && peers) return ALGO_PUSH;
    return ALGO_MULTIMEM;
}

// Enqueue one all-reduce of n4 float4 vectors for rank r on `st`.  Two-shot kernels work in place
// on the data region; the one-shot kernel reads input slot (oneshot_calls & 1) of the data region
// and writes `oneshot_out` (local memory; the scratch region when null).
int launch_allreduce(const Tuning &tn, NvlsTeam &t, NvlsRank &r, int algo, size_t n4, unsigned long oneshot_call,
                     float *oneshot_out, cudaStream_t st) {
    ArArgs a;
    a.mc_base = (unsigned char *)r.mc;
    a.uc_base = (unsigned char *)r.uc;
    a.epochs = r.d_epochs;
    a.err = r.d_err;
    a.timeout_ns = tn.timeout_ns;
    a.data_off = kFlagBytes;
    a.scratch_off = kFlagBytes + t.data_cap;
    a.rank = r.rank;
    a.nranks = t.nranks;
    PeerPtrs data, scratch;
    std::memset(&data, 0, sizeof data);
    std::memset(&scratch, 0, sizeof scratch);
    for (size_t i = 0; i < r.peer_base.size() && i < (size_t)kMaxPeers; ++i) {
        data.p[i] = reinterpret_cast<float *>((unsigned char *)r.peer_base[i] + a.data_off);
        scratch.p[i] = reinterpret_cast<float *>((unsigned char *)r.peer_base[i] + a.scratch_off);
    }
    B200V_CUDA(cudaSetDevice(r.dev));
    const int n = t.nranks;
    int ctas;
    // Peer kernels: every SM for large messages (measured at 2 GPUs, 1 GiB: 148 CTAs 682 GB/s, 64 CTAs
    // 623, 16 CTAs 482), fewer for small ones, where each extra CTA is one more barrier word to
    // reduce across the switch (1 MiB: 8 CTAs 20.3 us, 32 CTAs 21.0 us).
    const int peer_ctas = (int)std::min<size_t>((size_t)r.sm_count, std::max<size_t>(8, (n4 * 16 + (128u << 10) - 1) / (128u << 10)));
    switch (algo) {
    case ALGO_ONESHOT: {
        // one vector per thread while that fits a wave of CTAs; few CTAs keep the barrier cheap
        ctas = (int)std::min<size_t>((n4 + kThreads - 1) / kThreads, (size_t)std::min(r.sm_count, 64));
        if (tn.ctas) ctas = std::min(tn.ctas, r.sm_count);
        if (ctas < 1) ctas = 1;
        const size_t in_off = a.data_off + (oneshot_call & 1ul) * kOneShotCap;
        float *out = oneshot_out ? oneshot_out : reinterpret_cast<float *>(a.uc_base + a.scratch_off);
        oneshot_allreduce_kernel<<<ctas, kThreads, 0, st>>>(a, in_off, out, n4);
        break;
    }
    case ALGO_P2P: {
        ctas = tn.ctas ? tn.ctas : peer_ctas;
        auto kern = n == 2 ? p2p_allreduce_kernel<2, 4, true>
                  : n == 4 ? p2p_allreduce_kernel<4, 2, true> : p2p_allreduce_kernel<8, 1, true>;
        kern<<<ctas, kThreads, 0, st>>>(a, data, n4);
        break;
    }
    case ALGO_PUSH: {
        ctas = tn.ctas ? tn.ctas : peer_ctas;
        const int u = tn.unroll;
        if (n == 2) {
            auto kern = u == 1 ? push_allreduce_kernel<2, 1> : u == 2 ? push_allreduce_kernel<2, 2>
                      : u == 8 ? push_allreduce_kernel<2, 8> : push_allreduce_kernel<2, 4>;
            kern<<<ctas, kThreads, 0, st>>>(a, data, scratch, n4);
        } else if (n == 4) {
            auto kern = u == 1 ? push_allreduce_kernel<4, 1> : u == 4 ? push_allreduce_kernel<4, 4>
                      : push_allreduce_kernel<4, 2>;
            kern<<<ctas, kThreads, 0, st>>>(a, data, scratch, n4);
        } else {
            auto kern = u == 2 ? push_allreduce_kernel<8, 2> : push_allreduce_kernel<8, 1>;
            kern<<<ctas, kThreads, 0, st>>>(a, data, scratch, n4);
        }
        break;
    }
    case ALGO_HYBRID: {
        const int pct = tn.hybrid_pct ? tn.hybrid_pct : 83;
        const unsigned mm = (unsigned)std::min(tn.hybrid_mm_ctas ? tn.hybrid_mm_ctas : 16, r.sm_count / 2);
        ctas = tn.ctas ? std::max(tn.ctas, (int)mm + 1) : r.sm_count;
        // the switch-side part is a whole number of (ranks x 512-vector) units so both parts stay aligned
        size_t m4 = n4 * (size_t)pct / 100;
        m4 -= m4 % ((size_t)n * kThreads);
        if (m4 == 0 || m4 >= n4) {            // too small to split: all of it through the switch
            nvls_allreduce_kernel<4><<<(int)mm, kThreads, 0, st>>>(a, n4);
            break;
        }
        if (n == 4) hybrid_allreduce_kernel<4, 4, 2><<<ctas, kThreads, 0, st>>>(a, data, scratch, n4, m4, mm);
        else hybrid_allreduce_kernel<8, 4, 1><<<ctas, kThreads, 0, st>>>(a, data, scratch, n4, m4, mm);
        break;
    }
    default: {
        // Few CTAs are enough to keep the NVLink ports busy, and more only add barrier traffic and
        // switch-side contention: measured on 8 x B200 (busBW GB/s at 16 MiB / 64 MiB / 1 GiB)
        //   16 CTAs 458 / 697 / 839     48 CTAs 407 / 666 / 838     148 CTAs 419 / 590 / 829
        ctas = tn.ctas ? tn.ctas : std::min(r.sm_count, 16);
        const int u = tn.unroll;
        auto kern = u == 1 ? nvls_allreduce_kernel<1> : u == 2 ? nvls_allreduce_kernel<2>
                  : u == 8 ? nvls_allreduce_kernel<8> : nvls_allreduce_kernel<4>;
        kern<<<ctas, kThreads, 0, st>>>(a, n4);
        break;
    }
    }
    B200V_CUDA(cudaGetLastError());
    return B200V_OK;
}

// ---- the sweep --------------------------------------------------------------------------------------------

struct RowSample {
    std::vector<float> ms;            // per timed sample, already divided by the launches in it
    unsigned long long bad = 0;
    int algo = 0;
};

// The whole sweep for ONE rank: every rank of the collective runs the same sequence of launches, so
// nothing but the kernels' own barriers keeps them in step.  Runs on its own host thread when this
// process drives several GPUs, so that no GPU waits for the host to get round to it.
int sweep_rank(NvlsTeam &team, NvlsRank &r, const Tuning &tn, size_t min_bytes, size_t max_bytes, int warmup,
               int iters, int max_rows, std::vector<RowSample> &rows) {
    B200V_CUDA(cudaSetDevice(r.dev));
    const int nranks = team.nranks;
    unsigned long oneshot_call = 0;
    unsigned char *uc = (unsigned char *)r.uc;
    float *data = reinterpret_cast<float *>(uc + kFlagBytes);
    float *scratch = reinterpret_cast<float *>(uc + kFlagBytes + team.data_cap);
    auto check_err = [&]() -> int {
        B200V_CUDA(cudaStreamSynchronize(r.st));
        int e = 0;
        B200V_CUDA(cudaMemcpy(&e, r.d_err, 4, cudaMemcpyDeviceToHost));
        return e ? B200V_E_TIMEOUT : B200V_OK;
    };
    int nr = 0;
    for (size_t bytes = min_bytes; bytes <= max_bytes && nr < max_rows; bytes *= 2) {
        rows.emplace_back();
        RowSample &row = rows.back();
        const size_t count_f = bytes / 4, n4 = bytes / 16;
        const int algo = choose_algo(tn, team, bytes);
        row.algo = algo;
        const bool oneshot = algo == ALGO_ONESHOT;
        auto allreduce = [&]() -> int {
            int rc = launch_allreduce(tn, team, r, algo, n4, oneshot_call, nullptr, r.st);
            if (oneshot) ++oneshot_call;
            return rc;
        };
        // fill (both input slots for the one-shot kernel), one all-reduce, bitwise check
        auto fill_reduce_check = [&]() -> int {
            ar_fill_kernel<<<grid_fill(r, count_f), 256, 0, r.st>>>(data, count_f, r.rank);
            if (oneshot)
                ar_fill_kernel<<<grid_fill(r, count_f), 256, 0, r.st>>>(data + kOneShotCap / 4, count_f, r.rank);
            B200V_CUDA(cudaGetLastError());
            B200V_CUDA(cudaMemsetAsync(r.d_bad, 0, 8, r.st));
            int rc = allreduce();
            if (rc != B200V_OK) return rc;
            ar_check_kernel<<<grid_fill(r, count_f), 256, 0, r.st>>>(oneshot ? scratch : data, count_f, nranks, r.d_bad);
            B200V_CUDA(cudaGetLastError());
            rc = check_err();
            if (rc != B200V_OK) return rc;
            unsigned long long bad = 0;
            B200V_CUDA(cudaMemcpy(&bad, r.d_bad, 8, cudaMemcpyDeviceToHost));
            row.bad += bad;
            return B200V_OK;
        };
        int rc = fill_reduce_check();
        if (rc != B200V_OK) return rc;

        for (int i = 0; i < warmup; ++i) {
            rc = allreduce();
            if (rc != B200V_OK) return rc;
        }
        // A small all-reduce lasts a few microseconds, comparable to the resolution of an event
        // pair: each timed sample is `reps` back-to-back launches lasting >= ~100 us in total.
        const double est_us = (double)bytes / 300e3 + 8.0;
        const int reps = (int)std::min(16.0, std::max(1.0, 100.0 / est_us));
        for (int it = 0; it < iters; ++it) {
            B200V_CUDA(cudaEventRecord(r.ev[2 * it], r.st));
            for (int k = 0; k < reps; ++k) {
                rc = allreduce();
                if (rc != B200V_OK) return rc;
            }
            B200V_CUDA(cudaEventRecord(r.ev[2 * it + 1], r.st));
        }
        rc = check_err();
        if (rc != B200V_OK) return rc;
        row.ms.resize((size_t)iters);
        for (int it = 0; it < iters; ++it) {
            float t = 0.f;
            B200V_CUDA(cudaEventElapsedTime(&t, r.ev[2 * it], r.ev[2 * it + 1]));
            row.ms[(size_t)it] = t / (float)reps;
        }
        // and once more after the back-to-back launches: a barrier race that only shows under
        // queued launches would corrupt this result
        rc = fill_reduce_check();
        if (rc != B200V_OK) return rc;
        ++nr;
        if (bytes > max_bytes / 2) break;
    }
    return B200V_OK;
}

int sweep_team(NvlsTeam &team, size_t min_bytes, size_t max_bytes, int warmup, int iters, b200v_ar_row *rows,
               int max_rows, int *nrows) {
    const Tuning tn = read_tuning();
    const size_t nl = team.ranks.size();
    std::vector<std::vector<RowSample>> per_rank(nl);
    std::vector<int> rcs(nl, B200V_OK);
    if (nl == 1) {
        rcs[0] = sweep_rank(team, team.ranks[0], tn, min_bytes, max_bytes, warmup, iters, max_rows, per_rank[0]);
    } else {
        std::vector<std::thread> th;
        for (size_t i = 0; i < nl; ++i)
            th.emplace_back([&, i] {
                rcs[i] = sweep_rank(team, team.ranks[i], tn, min_bytes, max_bytes, warmup, iters, max_rows, per_rank[i]);
            });
        for (auto &x : th) x.join();
    }
    for (int rc : rcs)
        if (rc != B200V_OK) return rc;
    const size_t nr = per_rank[0].size();
    size_t bytes = min_bytes;
    for (size_t k = 0; k < nr; ++k, bytes *= 2) {
        b200v_ar_row *row = &rows[k];
        std::memset(row, 0, sizeof(*row));
        row->bytes = bytes;
        std::vector<float> ms((size_t)iters, 0.f);
        for (size_t i = 0; i < nl; ++i) {
            row->mismatches += per_rank[i][k].bad;
            for (int it = 0; it < iters; ++it) ms[(size_t)it] = std::max(ms[(size_t)it], per_rank[i][k].ms[(size_t)it]);
        }
        std::sort(ms.begin(), ms.end());
        row->best_ms = ms[0];
        row->median_ms = ms[(size_t)iters / 2];
        row->algbw_gbs = (double)bytes / (row->best_ms * 1e-3) / 1e9;
        row->busbw_gbs = row->algbw_gbs * 2.0 * (team.nranks - 1) / team.nranks;
        row->algo = per_rank[0][k].algo + 1;         // B200V_AR_ALGO_*
    }
    *nrows = (int)nr;
    return B200V_OK;
}

bool bad_probe_args(size_t min_bytes, size_t max_bytes, int warmup, int iters, const void *rows,
                    int max_rows, const int *nrows) {
    return !rows || !nrows || max_rows < 1 || iters < 1 || warmup < 0 || min_bytes < 16 || (min_bytes & 15) ||
           max_bytes < min_bytes;
}

}  // namespace
}  // namespace b200v

using namespace b200v;

extern "C" int b200v_last_driver_error(void) { return g_last_cu; }

// Host logic only (no device is touched): the kernel the probes and the comm pick for a message.
extern "C" int b200v_nvls_algo_for(int nranks, size_t bytes) {
    if (nranks < 2 || bytes == 0) return B200V_E_BADARG;
    NvlsTeam t;
    t.nranks = nranks;
    t.peers_mapped = nranks <= kMaxPeers;
    return choose_algo(read_tuning(), t, (bytes + 15) / 16 * 16) + 1;
}

extern "C" int b200v_nvls_allreduce_probe(int ndev, const int *devs, size_t min_bytes, size_t max_bytes,
                                          int warmup, int iters, b200v_ar_row *rows, int max_rows,
                                          int *nrows) {
    // a multicast object needs at least two devices (cuMulticastCreate refuses one)
    if (ndev < 2 || !devs || bad_probe_args(min_bytes, max_bytes, warmup, iters, rows, max_rows, nrows))
        return B200V_E_BADARG;
    *nrows = 0;
    int count = 0;
    int rc = b200v_device_count(&count);
    if (rc != B200V_OK) return rc;
    for (int i = 0; i < ndev; ++i) {
        if (devs[i] < 0 || devs[i] >= count) return B200V_E_BADARG;
        for (int j = 0; j < i; ++j)
            if (devs[j] == devs[i]) return B200V_E_BADARG;     // one rank per GPU: ranks wait on each other
    }

    NvlsTeam team;
    struct Cleanup {
        NvlsTeam &t;
        ~Cleanup() { destroy_team(t); }
    } cleanup{team};
    rc = build_team(team, ndev, devs, max_bytes, iters);
    if (rc != B200V_OK) return rc;
    return sweep_team(team, min_bytes, max_bytes, warmup, iters, rows, max_rows, nrows);
}

extern "C" int b200v_nvls_allreduce_probe_rank(int dev, int rank, int nranks, const char *rendezvous,
                                               size_t min_bytes, size_t max_bytes, int warmup, int iters,
                                               b200v_ar_row *rows, int max_rows, int *nrows) {
    if (nranks < 2 || rank < 0 || rank >= nranks || !rendezvous || !*rendezvous ||
        bad_probe_args(min_bytes, max_bytes, warmup, iters, rows, max_rows, nrows))
        return B200V_E_BADARG;
    *nrows = 0;
    int count = 0;
    int rc = b200v_device_count(&count);
    if (rc != B200V_OK) return rc;
    if (dev < 0 || dev >= count) return B200V_E_BADARG;

    Rendezvous rv;
    NvlsTeam team;
    struct Cleanup {
        NvlsTeam &t;
        Rendezvous &rv;
        ~Cleanup() {
            rv.barrier();        // nobody unbinds while a peer's kernel may still be running
            destroy_team(t);
            rv.close_all();
        }
    } cleanup{team, rv};
    rc = rv.open(rendezvous, rank, nranks);
    if (rc != B200V_OK) return rc;
    rc = build_team_rank(team, rv, dev, rank, nranks, max_bytes, iters);
    if (rc != B200V_OK) return rc;
    return sweep_team(team, min_bytes, max_bytes, warmup, iters, rows, max_rows, nrows);
}

// ---- all-reduce of a caller-owned device buffer -------------------------------------------------------------

struct b200v_nvls_comm {
    NvlsTeam team;
    Rendezvous rv;
    Tuning tn;
    size_t max_bytes = 0;
};

extern "C" int b200v_nvls_comm_create(int dev, int rank, int nranks, const char *rendezvous, size_t max_bytes,
                                      b200v_nvls_comm **out) {
    if (!out) return B200V_E_BADARG;
    *out = nullptr;
    if (nranks < 2 || rank < 0 || rank >= nranks || !rendezvous || !*rendezvous || max_bytes < 16)
        return B200V_E_BADARG;
    int count = 0;
    int rc = b200v_device_count(&count);
    if (rc != B200V_OK) return rc;
    if (dev < 0 || dev >= count) return B200V_E_BADARG;
    b200v_nvls_comm *c = new (std::nothrow) b200v_nvls_comm();
    if (!c) return B200V_E_NOMEM;
    c->tn = read_tuning();
    c->max_bytes = (max_bytes + 15) / 16 * 16;
    rc = c->rv.open(rendezvous, rank, nranks);
    if (rc == B200V_OK) rc = build_team_rank(c->team, c->rv, dev, rank, nranks, c->max_bytes, 0);
    if (rc != B200V_OK) {
        c->rv.barrier();
        destroy_team(c->team);
        c->rv.close_all();
        delete c;
        return rc;
    }
    *out = c;
    return B200V_OK;
}

extern "C" int b200v_nvls_comm_allreduce(b200v_nvls_comm *c, float *dBuf, size_t count, void *stream) {
    if (!c || c->team.ranks.empty()) return B200V_E_BADARG;
    if (count == 0) return B200V_OK;
    if (!dBuf) return B200V_E_BADARG;
    const size_t n4 = (count + 3) / 4, padded = n4 * 16;
    if (padded > c->max_bytes) return B200V_E_TOOLARGE;
    NvlsRank &r = c->team.ranks[0];
    cudaStream_t st = (cudaStream_t)stream;
    B200V_CUDA(cudaSetDevice(r.dev));
    const int algo = choose_algo(c->tn, c->team, padded);
    unsigned char *uc = (unsigned char *)r.uc;
    unsigned char *in = uc + kFlagBytes + (algo == ALGO_ONESHOT ? (c->team.oneshot_calls & 1ul) * kOneShotCap : 0);
    unsigned char *res = algo == ALGO_ONESHOT ? uc + kFlagBytes + c->team.data_cap : in;
    // The kernels work on the library's own NVLink-mapped allocation: stage the caller's buffer in
    // (zero-padding to whole float4 vectors), reduce, stage the result out.  All on `stream`.
    B200V_CUDA(cudaMemcpyAsync(in, dBuf, count * 4, cudaMemcpyDeviceToDevice, st));
    if (padded > count * 4) B200V_CUDA(cudaMemsetAsync(in + count * 4, 0, padded - count * 4, st));
    int rc = launch_allreduce(c->tn, c->team, r, algo, n4, c->team.oneshot_calls, nullptr, st);
    if (rc != B200V_OK) return rc;
    if (algo == ALGO_ONESHOT) ++c->team.oneshot_calls;
    B200V_CUDA(cudaMemcpyAsync(dBuf, res, count * 4, cudaMemcpyDeviceToDevice, st));
    return B200V_OK;
}

extern "C" int b200v_nvls_comm_check(b200v_nvls_comm *c) {
    if (!c || c->team.ranks.empty()) return B200V_E_BADARG;
    NvlsRank &r = c->team.ranks[0];
    B200V_CUDA(cudaSetDevice(r.dev));
    B200V_CUDA(cudaDeviceSynchronize());
    int e = 0;
    B200V_CUDA(cudaMemcpy(&e, r.d_err, 4, cudaMemcpyDeviceToHost));
    return e ? B200V_E_TIMEOUT : B200V_OK;
}

extern "C" int b200v_nvls_comm_destroy(b200v_nvls_comm *c) {
    if (!c) return B200V_OK;
    c->rv.barrier();             // nobody unbinds while a peer's kernel may still be running
    destroy_team(c->team);
    c->rv.close_all();
    delete c;
    return B200V_OK;
}
