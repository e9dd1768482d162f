// nvls_allreduce.cu — a hand-written all-reduce over NVLink/NVSwitch multicast, as a second NVLink
// health probe beside NCCL.  Single process, one kernel per GPU.
//
// How it moves data.  Every GPU's buffer is bound to one multicast object (cuMulticastCreate /
// cuMulticastBindMem), so one address names "this offset on all GPUs".  GPU r owns slice r of the
// vector and, for each 16 bytes of it, issues
//     multimem.ld_reduce.add.v4.f32   -> the NVSwitch reads that offset from all N GPUs, adds, and
//                                        returns the sum (the reduction happens in the switch)
//     multimem.st.v4.f32              -> the switch writes the sum back to that offset on all N
// so each GPU sends and receives about one vector's worth of bytes per all-reduce instead of the
// 2(N-1)/N x N of a ring, and no GPU ever touches a peer's memory with ordinary loads or stores.
// Two cross-GPU barriers bracket the work (inputs ready / results landed); they use
// multimem.red.add on a per-CTA flag word that exists on every GPU, then spin on the local copy.
//
// With 2 GPUs the switch has nothing to reduce and a second kernel is used instead
// (p2p_allreduce_kernel below): plain peer loads and stores over NVLink, same slicing, same
// barriers.  B200V_NVLS_ALGO=multimem|p2p forces either.
//
// Everything from the CUDA driver API is resolved through cudaGetDriverEntryPoint, so the library
// keeps having no link-time dependency on libcuda.  The spin loops are bounded (B200V_NVLS_TIMEOUT_MS,
// default 2000): a peer that never arrives yields B200V_E_TIMEOUT, never a hung GPU.
#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda.h>   // types and enums; every driver function is fetched at run time
#include <poll.h>
#include <sys/socket.h>
#include <sys/uio.h>
#include <sys/un.h>
#include <unistd.h>
#include <vector>

#include "ar_common.cuh"
#include "common.cuh"
#include "rendezvous.h"

namespace b200v {
namespace {

constexpr size_t kFlagBytes = 64 * 1024;      // barrier words live in front of the data
constexpr int kMaxCtas = 4096;
constexpr int kThreads = 512;

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

// All CTAs with this blockIdx on all GPUs meet here.  mc_flag names the flag word on every GPU,
// uc_flag is this GPU's own copy.  Returns false on timeout.
__device__ __forceinline__ bool gpu_barrier(unsigned *mc_flag, unsigned *uc_flag, unsigned target,
                                            unsigned long long timeout_ns) {
    __shared__ int ok_s;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
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

// In-place sum all-reduce of n4 float4 vectors that live at the same offset on every GPU.
// mc_base / uc_base: this GPU's mapping of the multicast object / of its own memory; flags first.
template <int UNROLL>
__global__ void __launch_bounds__(kThreads)
nvls_allreduce_kernel(unsigned char *mc_base, unsigned char *uc_base, size_t n4, int rank, int nranks,
                      unsigned epoch, unsigned long long timeout_ns, int *err) {
    unsigned *mc_flag = reinterpret_cast<unsigned *>(mc_base) + blockIdx.x;
    unsigned *uc_flag = reinterpret_cast<unsigned *>(uc_base) + blockIdx.x;
    float *mc = reinterpret_cast<float *>(mc_base + kFlagBytes);

    // (1) every GPU has entered this kernel, so every GPU's input (produced earlier on its stream)
    //     is complete and visible
    if (!gpu_barrier(mc_flag, uc_flag, epoch + (unsigned)nranks, timeout_ns)) {
        if (threadIdx.x == 0) atomicExch(err, 1);
        return;
    }
    const size_t per = (n4 + (size_t)nranks - 1) / (size_t)nranks;
    const size_t lo = per * (size_t)rank < n4 ? per * (size_t)rank : n4;
    const size_t hi = lo + per < n4 ? lo + per : n4;
    // CTA b walks tiles b, b+grid, ... of its GPU's slice; a tile is UNROLL x kThreads float4s, and a
    // thread issues all UNROLL switch-side reductions before the first broadcast store.
    constexpr size_t TILE = (size_t)UNROLL * kThreads;
    for (size_t t0 = lo + (size_t)blockIdx.x * TILE; t0 < hi; t0 += (size_t)gridDim.x * TILE) {
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
    if (!gpu_barrier(mc_flag, uc_flag, epoch + 2u * (unsigned)nranks, timeout_ns)) {
        if (threadIdx.x == 0) atomicExch(err, 2);
    }
}

// The same all-reduce with plain peer loads and stores over NVLink instead of switch-side
// reduction: GPU r reads slice r from every GPU (N-1 of them across NVLink), adds in rank order,
// and stores the sum into slice r of every GPU.  Each GPU then moves 2(N-1)/N x S per direction —
// what a ring moves, but in one hop and with no per-step synchronisation — which beats the
// multicast form at 2 GPUs (683 vs 394 GB/s), where the switch has nothing to reduce and the
// multicast form still sends the local copy on a round trip through it.  From 4 GPUs the
// multicast form is ahead (678 vs 661 GB/s at 4).
constexpr int kMaxPeers = 8;
struct PeerPtrs {
    float *data[kMaxPeers];
};

// STRONG = .relaxed.sys accesses; otherwise ordinary weak ones.  Weak is safe here: inputs were
// complete before the opening barrier (acquire.sys + bar.sync), nothing in this kernel has read
// these lines before (L1 is invalidated at launch), and the closing barrier's release.sys follows
// a __threadfence_system().
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

template <int NR, int UNROLL, bool STRONG>
__global__ void __launch_bounds__(kThreads)
p2p_allreduce_kernel(PeerPtrs peers, unsigned char *mc_base, unsigned char *uc_base, size_t n4, int rank,
                     unsigned epoch, unsigned long long timeout_ns, int *err) {
    unsigned *mc_flag = reinterpret_cast<unsigned *>(mc_base) + blockIdx.x;
    unsigned *uc_flag = reinterpret_cast<unsigned *>(uc_base) + blockIdx.x;
    if (!gpu_barrier(mc_flag, uc_flag, epoch + (unsigned)NR, timeout_ns)) {
        if (threadIdx.x == 0) atomicExch(err, 1);
        return;
    }
    const size_t per = (n4 + (size_t)NR - 1) / (size_t)NR;
    const size_t lo = per * (size_t)rank < n4 ? per * (size_t)rank : n4;
    const size_t hi = lo + per < n4 ? lo + per : n4;
    constexpr size_t TILE = (size_t)UNROLL * kThreads;
    for (size_t t0 = lo + (size_t)blockIdx.x * TILE; t0 < hi; t0 += (size_t)gridDim.x * TILE) {
        float4 in[UNROLL][NR];
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) {
            const size_t v = t0 + (size_t)j * kThreads + threadIdx.x;
            if (v < hi) {
#pragma unroll
                for (int p = 0; p < NR; ++p) in[j][p] = ld_f4<STRONG>(peers.data[p] + 4 * v);
            }
        }
#pragma unroll
        for (int j = 0; j < UNROLL; ++j) {
            const size_t v = t0 + (size_t)j * kThreads + threadIdx.x;
            if (v < hi) {
                float4 s = in[j][0];
#pragma unroll
                for (int p = 1; p < NR; ++p) {
                    s.x += in[j][p].x; s.y += in[j][p].y; s.z += in[j][p].z; s.w += in[j][p].w;
                }
#pragma unroll
                for (int p = 0; p < NR; ++p) st_f4<STRONG>(peers.data[p] + 4 * v, s);
            }
        }
    }
    if (!gpu_barrier(mc_flag, uc_flag, epoch + 2u * (unsigned)NR, timeout_ns)) {
        if (threadIdx.x == 0) atomicExch(err, 2);
    }
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
    int sm_count = 0;
    std::vector<cudaEvent_t> ev;
};

struct NvlsTeam {
    std::vector<NvlsRank> ranks;
    CUmemGenericAllocationHandle mc_handle = 0;
    size_t size = 0;
    bool have_mc = false;
    // Address of every rank's buffer as seen from this process (index = rank), for the
    // peer-to-peer kernel.  One process: the ranks' own mappings.  One process per GPU: the other
    // ranks' allocations imported (peer_mem) and mapped here (peer_va).  Empty when unavailable.
    std::vector<CUdeviceptr> peer_uc;
    std::vector<CUmemGenericAllocationHandle> peer_mem;
    std::vector<CUdeviceptr> peer_va;
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

void destroy_team(NvlsTeam &t) {
    const Drv &d = drv();
    for (NvlsRank &r : t.ranks) {
        if (r.dev < 0) continue;
        cudaSetDevice(r.dev);
        if (r.st) cudaStreamSynchronize(r.st);
    }
    for (CUdeviceptr va : t.peer_va) {
        d.MemUnmap(va, t.size);
        d.MemAddressFree(va, t.size);
    }
    for (CUmemGenericAllocationHandle h : t.peer_mem) d.MemRelease(h);
    t.peer_va.clear();
    t.peer_mem.clear();
    t.peer_uc.clear();
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
        if (r.st) cudaStreamDestroy(r.st);
    }
    if (t.have_mc) d.MemRelease(t.mc_handle);
    t.ranks.clear();
}

int build_team(NvlsTeam &t, int ndev, const int *devs, size_t data_bytes, int iters) {
    const Drv &d = drv();
    if (!d.ok) return B200V_E_UNSUPPORTED;
    t.ranks.resize((size_t)ndev);
    for (int i = 0; i < ndev; ++i) {
        NvlsRank &r = t.ranks[(size_t)i];
        r.dev = devs[i];
        r.rank = i;
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
        r.ev.resize((size_t)iters * 2);
        for (auto &e : r.ev) B200V_CUDA(cudaEventCreate(&e));
    }

    CUmulticastObjectProp mprop;
    std::memset(&mprop, 0, sizeof mprop);
    mprop.numDevices = (unsigned)ndev;
    mprop.size = kFlagBytes + data_bytes;
    size_t gran = 0;
    B200V_CU(d.MulticastGetGranularity(&gran, &mprop, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    CUmemAllocationProp aprop;
    std::memset(&aprop, 0, sizeof aprop);
    aprop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    aprop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    aprop.location.id = t.ranks[0].cudev;
    size_t agran = 0;
    B200V_CU(d.MemGetAllocationGranularity(&agran, &aprop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    gran = std::max(gran, agran);
    t.size = (mprop.size + gran - 1) / gran * gran;
    mprop.size = t.size;
    B200V_CU(d.MulticastCreate(&t.mc_handle, &mprop));
    t.have_mc = true;
    for (NvlsRank &r : t.ranks) B200V_CU(d.MulticastAddDevice(t.mc_handle, r.cudev));

    for (NvlsRank &r : t.ranks) {
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
        // every GPU of the team may load from and store to this GPU's buffer (the peer-to-peer
        // kernel does; the multicast kernel only touches its own)
        std::vector<CUmemAccessDesc> all((size_t)ndev, acc);
        for (int j = 0; j < ndev; ++j) all[(size_t)j].location.id = t.ranks[(size_t)j].cudev;
        B200V_CU(d.MemSetAccess(r.uc, t.size, all.data(), (size_t)ndev));
        B200V_CU(d.MemAddressReserve(&r.mc, t.size, gran, 0, 0));
        B200V_CU(d.MemMap(r.mc, t.size, 0, t.mc_handle, 0));
        r.mc_mapped = true;
        B200V_CU(d.MemSetAccess(r.mc, t.size, &acc, 1));
        B200V_CUDA(cudaMemsetAsync((void *)r.uc, 0, kFlagBytes, r.st));
        B200V_CUDA(cudaStreamSynchronize(r.st));
    }
    for (NvlsRank &r : t.ranks) t.peer_uc.push_back(r.uc);
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
    B200V_CUDA(cudaSetDevice(r.dev));
    B200V_CUDA(cudaFree(nullptr));
    B200V_CU(d.DeviceGet(&r.cudev, r.dev));
    int mc_ok = 0;
    B200V_CU(d.DeviceGetAttribute(&mc_ok, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, r.cudev));
    if (!mc_ok) return B200V_E_UNSUPPORTED;
    B200V_CUDA(cudaDeviceGetAttribute(&r.sm_count, cudaDevAttrMultiProcessorCount, r.dev));
    B200V_CUDA(cudaStreamCreateWithFlags(&r.st, cudaStreamNonBlocking));
    B200V_CUDA(cudaMalloc(&r.d_bad, 8));
    B200V_CUDA(cudaMalloc(&r.d_err, 4));
    B200V_CUDA(cudaMemset(r.d_err, 0, 4));
    r.ev.resize((size_t)iters * 2);
    for (auto &e : r.ev) B200V_CUDA(cudaEventCreate(&e));

    CUmulticastObjectProp mprop;
    std::memset(&mprop, 0, sizeof mprop);
    mprop.numDevices = (unsigned)nranks;
    mprop.size = kFlagBytes + data_bytes;
    mprop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t gran = 0;
    B200V_CU(d.MulticastGetGranularity(&gran, &mprop, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    CUmemAllocationProp aprop;
    std::memset(&aprop, 0, sizeof aprop);
    aprop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    aprop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    aprop.location.id = r.cudev;
    aprop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;   // same kind as the multicast object
    size_t agran = 0;
    B200V_CU(d.MemGetAllocationGranularity(&agran, &aprop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    gran = std::max(gran, agran);
    t.size = (mprop.size + gran - 1) / gran * gran;     // every rank computes the same size
    mprop.size = t.size;

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
    B200V_CU(d.MemSetAccess(r.uc, t.size, &acc, 1));
    B200V_CU(d.MemAddressReserve(&r.mc, t.size, gran, 0, 0));
    B200V_CU(d.MemMap(r.mc, t.size, 0, t.mc_handle, 0));
    r.mc_mapped = true;
    B200V_CU(d.MemSetAccess(r.mc, t.size, &acc, 1));
    B200V_CUDA(cudaMemsetAsync((void *)r.uc, 0, kFlagBytes, r.st));
    B200V_CUDA(cudaStreamSynchronize(r.st));

    // For the peer-to-peer kernel (2 ranks, or forced) every rank also maps every other rank's
    // buffer: each exports its allocation as a descriptor, rank 0 relays them all around.  All ranks
    // take this branch or none (same nranks, same environment).
    const char *algo = std::getenv("B200V_NVLS_ALGO");
    const bool want_peers = algo ? !std::strcmp(algo, "p2p") && (nranks == 2 || nranks == 4 || nranks == 8)
                                 : nranks == 2;
    if (want_peers) {
        int my_fd = -1;
        B200V_CU(Export(&my_fd, r.mem, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
        std::vector<int> fds;
        const bool ok = rv.exchange_fds(my_fd, fds);
        int rc = ok ? B200V_OK : B200V_E_TIMEOUT;
        std::vector<CUdeviceptr> uc((size_t)nranks, 0);
        uc[(size_t)rank] = r.uc;
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
            uc[(size_t)k] = va;
        }
        for (int k = 0; k < (int)fds.size(); ++k)
            if (fds[(size_t)k] >= 0) ::close(fds[(size_t)k]);      // includes my_fd (fds[rank])
        if (fds.empty() && my_fd >= 0) ::close(my_fd);
        if (rc != B200V_OK) return rc;
        t.peer_uc = uc;
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

}  // namespace
}  // namespace b200v

using namespace b200v;

extern "C" int b200v_last_driver_error(void) { return g_last_cu; }

// The sweep over message sizes, for the ranks this process drives (all of them in the
// single-process probe, one in the per-rank probe).  Every rank of the collective runs the same
// sequence of launches, so the barrier epochs stay in step without any host-side exchange.
static int sweep_team(NvlsTeam &team, int nranks, size_t min_bytes, size_t max_bytes, int warmup, int iters,
                      b200v_ar_row *rows, int max_rows, int *nrows) {
    const int ndev = nranks;
    // every rank's buffer is addressable from here (one process, or imported peer mappings)
    const bool all_local = (int)team.peer_uc.size() == nranks && nranks <= kMaxPeers;
    int rc = B200V_OK;

    // Few CTAs are enough to keep the NVLink ports busy, and more only add barrier traffic and
    // switch-side contention: measured on 8 x B200 (busBW GB/s at 16 MiB / 64 MiB / 1 GiB)
    //   16 CTAs 458 / 697 / 839     48 CTAs 407 / 666 / 838     148 CTAs 419 / 590 / 829
    int ctas = 16;
    for (const NvlsRank &r : team.ranks) ctas = std::min(ctas, r.sm_count);
    if (const char *e = std::getenv("B200V_NVLS_CTAS")) {
        const int v = std::atoi(e);
        if (v >= 1 && v <= kMaxCtas) ctas = v;
    }
    if ((size_t)ctas * 4 > kFlagBytes) ctas = (int)(kFlagBytes / 4);
    int unroll = 4;
    if (const char *e = std::getenv("B200V_NVLS_UNROLL")) unroll = std::atoi(e);
    unsigned long long timeout_ns = 2000ull * 1000000ull;
    if (const char *e = std::getenv("B200V_NVLS_TIMEOUT_MS")) {
        const long v = std::atol(e);
        if (v > 0) timeout_ns = (unsigned long long)v * 1000000ull;
    }

    unsigned epoch = 0;   // every launch consumes 2*ndev counts of each flag word
    // Algorithm: switch-side reduction (multimem) from 4 GPUs up; with 2 GPUs there is nothing for
    // the switch to reduce and the plain peer-to-peer form wins.  B200V_NVLS_ALGO=multimem|p2p
    // overrides (p2p is compiled for 2, 4 and 8 GPUs).
    // Measured busBW at 1 GiB (GB/s): 2 GPUs p2p 683 / multimem 394; 4 GPUs p2p 661 / multimem 678.
    bool use_p2p = ndev == 2 && all_local;
    if (const char *e = std::getenv("B200V_NVLS_ALGO")) {
        if (!std::strcmp(e, "p2p")) use_p2p = all_local && (ndev == 2 || ndev == 4 || ndev == 8);
        else if (!std::strcmp(e, "multimem")) use_p2p = false;
    }
    PeerPtrs peers;
    std::memset(&peers, 0, sizeof peers);
    for (int i = 0; all_local && i < ndev && i < kMaxPeers; ++i)
        peers.data[i] = reinterpret_cast<float *>((unsigned char *)team.peer_uc[(size_t)i] + kFlagBytes);
    if (use_p2p && !std::getenv("B200V_NVLS_CTAS")) ctas = std::min(team.ranks[0].sm_count, 64);

    auto allreduce_all = [&](size_t n4) -> int {
        for (NvlsRank &r : team.ranks) {
            const int i = r.rank;
            B200V_CUDA(cudaSetDevice(r.dev));
            if (use_p2p) {
                // Measured at 2 GPUs: weak vs .relaxed.sys accesses and twice the loads in flight all
                // land on the same 680-686 GB/s — the link is the limit — so one shape per rank count.
                auto kern = ndev == 2 ? p2p_allreduce_kernel<2, 4, true>
                          : ndev == 4 ? p2p_allreduce_kernel<4, 2, true> : p2p_allreduce_kernel<8, 1, true>;
                kern<<<ctas, kThreads, 0, r.st>>>(peers, (unsigned char *)r.mc, (unsigned char *)r.uc, n4, i, epoch,
                                                  timeout_ns, r.d_err);
            } else {
                auto kern = unroll == 1 ? nvls_allreduce_kernel<1> : unroll == 2 ? nvls_allreduce_kernel<2>
                          : unroll == 8 ? nvls_allreduce_kernel<8> : nvls_allreduce_kernel<4>;
                kern<<<ctas, kThreads, 0, r.st>>>((unsigned char *)r.mc, (unsigned char *)r.uc, n4, i, ndev, epoch,
                                                  timeout_ns, r.d_err);
            }
            B200V_CUDA(cudaGetLastError());
        }
        epoch += 2u * (unsigned)ndev;
        return B200V_OK;
    };
    auto sync_and_check_err = [&]() -> int {
        int worst = 0;
        for (NvlsRank &r : team.ranks) {
            B200V_CUDA(cudaSetDevice(r.dev));
            B200V_CUDA(cudaStreamSynchronize(r.st));
            int e = 0;
            B200V_CUDA(cudaMemcpy(&e, r.d_err, 4, cudaMemcpyDeviceToHost));
            worst = std::max(worst, e);
        }
        return worst ? B200V_E_TIMEOUT : B200V_OK;
    };

    int nr = 0;
    for (size_t bytes = min_bytes; bytes <= max_bytes && nr < max_rows; bytes *= 2) {
        b200v_ar_row *row = &rows[nr];
        std::memset(row, 0, sizeof(*row));
        row->bytes = bytes;
        const size_t count_f = bytes / 4, n4 = bytes / 16;

        for (NvlsRank &r : team.ranks) {
            const int i = r.rank;
            B200V_CUDA(cudaSetDevice(r.dev));
            float *data = reinterpret_cast<float *>((unsigned char *)r.uc + kFlagBytes);
            ar_fill_kernel<<<grid_fill(r, count_f), 256, 0, r.st>>>(data, count_f, i);
            B200V_CUDA(cudaGetLastError());
            B200V_CUDA(cudaMemsetAsync(r.d_bad, 0, 8, r.st));
        }
        rc = allreduce_all(n4);
        if (rc != B200V_OK) return rc;
        for (NvlsRank &r : team.ranks) {
            const int i = r.rank;
            B200V_CUDA(cudaSetDevice(r.dev));
            const float *data = reinterpret_cast<const float *>((unsigned char *)r.uc + kFlagBytes);
            ar_check_kernel<<<grid_fill(r, count_f), 256, 0, r.st>>>(data, count_f, ndev, r.d_bad);
            B200V_CUDA(cudaGetLastError());
        }
        rc = sync_and_check_err();
        if (rc != B200V_OK) return rc;
        for (NvlsRank &r : team.ranks) {
            B200V_CUDA(cudaSetDevice(r.dev));
            unsigned long long bad = 0;
            B200V_CUDA(cudaMemcpy(&bad, r.d_bad, 8, cudaMemcpyDeviceToHost));
            row->mismatches += bad;
        }

        for (int i = 0; i < warmup; ++i) {
            rc = allreduce_all(n4);
            if (rc != B200V_OK) return rc;
        }
        for (int it = 0; it < iters; ++it) {
            for (NvlsRank &r : team.ranks) {
                B200V_CUDA(cudaSetDevice(r.dev));
                B200V_CUDA(cudaEventRecord(r.ev[2 * it], r.st));
            }
            rc = allreduce_all(n4);
            if (rc != B200V_OK) return rc;
            for (NvlsRank &r : team.ranks) {
                B200V_CUDA(cudaSetDevice(r.dev));
                B200V_CUDA(cudaEventRecord(r.ev[2 * it + 1], r.st));
            }
        }
        rc = sync_and_check_err();
        if (rc != B200V_OK) return rc;
        std::vector<float> ms((size_t)iters, 0.f);
        for (int it = 0; it < iters; ++it)
            for (NvlsRank &r : team.ranks) {
                float t = 0.f;
                B200V_CUDA(cudaEventElapsedTime(&t, r.ev[2 * it], r.ev[2 * it + 1]));
                ms[(size_t)it] = std::max(ms[(size_t)it], t);
            }
        std::sort(ms.begin(), ms.end());
        row->best_ms = ms[0];
        row->median_ms = ms[(size_t)iters / 2];
        row->algbw_gbs = (double)bytes / (row->best_ms * 1e-3) / 1e9;
        row->busbw_gbs = row->algbw_gbs * 2.0 * (ndev - 1) / ndev;
        ++nr;
        if (bytes > max_bytes / 2) break;
    }
    *nrows = nr;
    return B200V_OK;
}

static bool bad_probe_args(size_t min_bytes, size_t max_bytes, int warmup, int iters, const void *rows,
                           int max_rows, const int *nrows) {
    return !rows || !nrows || max_rows < 1 || iters < 1 || warmup < 0 || min_bytes < 16 || (min_bytes & 15) ||
           max_bytes < min_bytes;
}

extern "C" int b200v_nvls_allreduce_probe(int ndev, const int *devs, size_t min_bytes, size_t max_bytes,
                                          int warmup, int iters, b200v_ar_row *rows, int max_rows,
                                          int *nrows) {
    if (ndev < 2 || !devs || bad_probe_args(min_bytes, max_bytes, warmup, iters, rows, max_rows, nrows))
        return B200V_E_BADARG;
    *nrows = 0;
    int count = 0;
    int rc = b200v_device_count(&count);
    if (rc != B200V_OK) return rc;
    for (int i = 0; i < ndev; ++i)
        if (devs[i] < 0 || devs[i] >= count) return B200V_E_BADARG;

    NvlsTeam team;
    struct Cleanup {
        NvlsTeam &t;
        ~Cleanup() { destroy_team(t); }
    } cleanup{team};
    rc = build_team(team, ndev, devs, max_bytes, iters);
    if (rc != B200V_OK) return rc;
    return sweep_team(team, ndev, min_bytes, max_bytes, warmup, iters, rows, max_rows, nrows);
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
    return sweep_team(team, nranks, min_bytes, max_bytes, warmup, iters, rows, max_rows, nrows);
}
