// capi.cu — the extern "C" surface of libb200validate.so (include/b200validate.h), except the
// NCCL probe which lives in nccl_probe.cu.  No CPU fallback exists anywhere in this library: with
// no usable device every compute entry point returns B200V_E_NODEVICE or the CUDA error.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <sys/mman.h>
#include <vector>

#include "common.cuh"
#include "copy_pool.h"
#include "ew_kernels.cuh"

namespace b200v {

namespace {
struct Slot {
    DeviceCtx ctx;
    std::mutex mu;
};
std::mutex g_map_mu;
std::map<int, std::unique_ptr<Slot>> g_ctx;

int init_ctx(DeviceCtx &c, int dev) {
    c.dev = dev;
    B200V_CUDA(cudaDeviceGetAttribute(&c.sm_count, cudaDevAttrMultiProcessorCount, dev));
    int l2 = 0;
    B200V_CUDA(cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, dev));
    c.l2_bytes = (size_t)l2;
    B200V_CUDA(cudaStreamCreateWithFlags(&c.s_in, cudaStreamNonBlocking));
    B200V_CUDA(cudaStreamCreateWithFlags(&c.s_k, cudaStreamNonBlocking));
    B200V_CUDA(cudaStreamCreateWithFlags(&c.s_out, cudaStreamNonBlocking));
    for (int i = 0; i < kSlots; ++i) {
        B200V_CUDA(cudaEventCreateWithFlags(&c.in_done[i], cudaEventDisableTiming));
        B200V_CUDA(cudaEventCreateWithFlags(&c.k_done[i], cudaEventDisableTiming));
        B200V_CUDA(cudaEventCreateWithFlags(&c.out_done[i], cudaEventDisableTiming));
    }
    cudaEvent_t *timed[] = {&c.t_begin, &c.t_end, &c.h2d_first, &c.h2d_last, &c.d2h_first, &c.d2h_last};
    for (cudaEvent_t *e : timed) B200V_CUDA(cudaEventCreate(e));
    c.ready = true;
    return B200V_OK;
}

void free_slots(DeviceCtx &c) {
    for (int i = 0; i < kSlots; ++i) {
        if (c.slot_x[i]) cudaFree(c.slot_x[i]);
        if (c.slot_y[i]) cudaFree(c.slot_y[i]);
        if (c.slot_d[i]) cudaFree(c.slot_d[i]);
        c.slot_x[i] = c.slot_y[i] = c.slot_d[i] = nullptr;
    }
    c.slot_elems = 0;
}

int ensure_slots(DeviceCtx &c, size_t elems) {
    if (c.slot_elems >= elems) return B200V_OK;
    free_slots(c);
    for (int i = 0; i < kSlots; ++i) {
        B200V_CUDA(cudaMalloc(&c.slot_x[i], elems * 4));
        B200V_CUDA(cudaMalloc(&c.slot_y[i], elems * 4));
        B200V_CUDA(cudaMalloc(&c.slot_d[i], elems * 4));
    }
    c.slot_elems = elems;
    return B200V_OK;
}

void free_stage(DeviceCtx &c) {
    for (int i = 0; i < kSlots; ++i) {
        if (c.stage_x[i]) cudaFreeHost(c.stage_x[i]);
        if (c.stage_y[i]) cudaFreeHost(c.stage_y[i]);
        if (c.stage_d[i]) cudaFreeHost(c.stage_d[i]);
        c.stage_x[i] = c.stage_y[i] = c.stage_d[i] = nullptr;
    }
    c.stage_elems = 0;
}

int ensure_stage(DeviceCtx &c, size_t elems) {
    if (c.stage_elems >= elems) return B200V_OK;
    free_stage(c);
    for (int i = 0; i < kSlots; ++i) {
        B200V_CUDA(cudaHostAlloc((void **)&c.stage_x[i], elems * 4, cudaHostAllocPortable));
        B200V_CUDA(cudaHostAlloc((void **)&c.stage_y[i], elems * 4, cudaHostAllocPortable));
        B200V_CUDA(cudaHostAlloc((void **)&c.stage_d[i], elems * 4, cudaHostAllocPortable));
    }
    c.stage_elems = elems;
    return B200V_OK;
}

// True when the DMA engines can read/write this host pointer directly (cudaHostAlloc /
// cudaHostRegister memory); false for ordinary pageable memory.
bool is_pinned(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        (void)cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

void destroy_ctx(DeviceCtx &c) {
    if (!c.ready) return;
    cudaSetDevice(c.dev);
    free_slots(c);
    free_stage(c);
    for (int i = 0; i < kSlots; ++i) {
        cudaEventDestroy(c.in_done[i]);
        cudaEventDestroy(c.k_done[i]);
        cudaEventDestroy(c.out_done[i]);
    }
    cudaEvent_t timed[] = {c.t_begin, c.t_end, c.h2d_first, c.h2d_last, c.d2h_first, c.d2h_last};
    for (cudaEvent_t e : timed) cudaEventDestroy(e);
    for (cudaEvent_t e : c.k_ev) cudaEventDestroy(e);
    c.k_ev.clear();
    cudaStreamDestroy(c.s_in);
    cudaStreamDestroy(c.s_k);
    cudaStreamDestroy(c.s_out);
    c.ready = false;
}
}  // namespace

int acquire_ctx(int dev, DeviceCtx **out) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        (void)cudaGetLastError();
        return B200V_E_NODEVICE;
    }
    if (dev < 0 || dev >= count) return B200V_E_BADARG;
    B200V_CUDA(cudaSetDevice(dev));
    Slot *slot;
    {
        std::lock_guard<std::mutex> g(g_map_mu);
        auto &p = g_ctx[dev];
        if (!p) p.reset(new Slot());
        slot = p.get();
    }
    slot->mu.lock();
    if (!slot->ctx.ready) {
        int rc = init_ctx(slot->ctx, dev);
        if (rc != B200V_OK) {
            slot->mu.unlock();
            return rc;
        }
    }
    slot->ctx.owner = slot;
    *out = &slot->ctx;
    return B200V_OK;
}

void release_ctx(DeviceCtx *ctx) { static_cast<Slot *>(ctx->owner)->mu.unlock(); }

int destroy_all_ctx() {
    // The Slot objects themselves are never freed: a thread that has just taken its Slot* in
    // acquire_ctx (map lock already dropped) must still find a live mutex to lock.  Shutdown
    // releases what the slots hold — streams, events, device and pinned buffers — under each
    // slot's own lock; a later call on that device simply builds the context again.
    std::lock_guard<std::mutex> g(g_map_mu);
    for (auto &kv : g_ctx) {
        std::lock_guard<std::mutex> g2(kv.second->mu);
        destroy_ctx(kv.second->ctx);
    }
    return B200V_OK;
}

namespace {

// RAII holder so every early return releases the per-device lock.
struct CtxGuard {
    DeviceCtx *c = nullptr;
    ~CtxGuard() { if (c) release_ctx(c); }
};

int grid_1d(const DeviceCtx &c, size_t n, int per_sm) {
    size_t g = (n + 255) / 256;
    const size_t cap = (size_t)c.sm_count * per_sm;
    if (g > cap) g = cap;
    return g < 1 ? 1 : (int)g;
}

// Chunked, three-stream host pipeline shared by b200v_vectoradd and b200v_stream_host.
int host_pipeline(int dev, int op, float *hDst, const float *hX, const float *hY, float s, size_t n,
                  b200v_timing *timing) {
    if (timing) std::memset(timing, 0, sizeof(*timing));
    if (op < 0 || op >= B200V_OP_COUNT) return B200V_E_BADARG;
    const bool needs_y = ew_needs_y(op);
    if (n > 0 && (!hDst || !hX || (needs_y && !hY))) return B200V_E_BADARG;
    CtxGuard g;
    int rc = acquire_ctx(dev, &g.c);
    if (rc != B200V_OK) return rc;
    if (n == 0) return B200V_OK;
    DeviceCtx &c = *g.c;

    // B200V_CHUNK_LOG2 overrides the chunk size (elements, log2) for transfer tuning runs.
    static const size_t chunk_cap = [] {
        const char *e = std::getenv("B200V_CHUNK_LOG2");
        const int lg = e && *e ? std::atoi(e) : 0;
        return (lg >= 10 && lg <= 30) ? (size_t)1 << lg : kChunkElems;
    }();
    // Pageable buffers (malloc, Go slices, numpy) are staged through pinned memory by a pool of
    // host threads, one chunk ahead of the DMA; pinned buffers are handed to the DMA directly.
    const bool pg_x = !is_pinned(hX), pg_y = needs_y && !is_pinned(hY), pg_d = !is_pinned(hDst);
    const bool staged = (pg_x || pg_y || pg_d) && !std::getenv("B200V_NO_STAGING");
    const size_t chunk = std::min(n, staged && !std::getenv("B200V_CHUNK_LOG2") ? kStagedChunkElems : chunk_cap);
    // Chunk schedule: full-size chunks, then a geometric taper (each chunk at most half of what is
    // left, down to kMinChunkElems).  The H2D stream is the bottleneck and runs back to back from
    // t = 0, so the call ends one kernel + one D2H after the LAST chunk lands: keeping that last
    // chunk small removes ~1 ms (one 64 MiB D2H) from a 2^28-element call.
    std::vector<std::pair<size_t, size_t>> sched;   // (first element, length)
    for (size_t at = 0; at < n;) {
        const size_t left = n - at;
        size_t len = std::min(chunk, std::max(kMinChunkElems, left / 2));
        if (len > left) len = left;
        sched.emplace_back(at, len);
        at += len;
    }
    const size_t nchunks = sched.size();
    rc = ensure_slots(c, chunk);
    if (rc != B200V_OK) return rc;
    if (staged) {
        rc = ensure_stage(c, chunk);
        if (rc != B200V_OK) return rc;
    }
    CopyPool &pool = CopyPool::instance();
    while (c.k_ev.size() < 2 * nchunks) {
        cudaEvent_t e;
        B200V_CUDA(cudaEventCreate(&e));
        c.k_ev.push_back(e);
    }

    // Whatever path leaves this function, nothing enqueued here may still be touching the caller's
    // host buffers afterwards: drain the three streams on every exit (a no-op after the normal
    // synchronisation below).
    struct Drain {
        DeviceCtx &c;
        ~Drain() {
            cudaStreamSynchronize(c.s_in);
            cudaStreamSynchronize(c.s_k);
            cudaStreamSynchronize(c.s_out);
            (void)cudaGetLastError();
        }
    } drain{c};

    int grid = 0, block = 0;
    B200V_CUDA(cudaEventRecord(c.t_begin, c.s_in));
    B200V_CUDA(cudaStreamWaitEvent(c.s_k, c.t_begin, 0));
    B200V_CUDA(cudaStreamWaitEvent(c.s_out, c.t_begin, 0));
    for (size_t k = 0; k < nchunks; ++k) {
        const int slot = (int)(k % kSlots);
        const size_t lo = sched[k].first, len = sched[k].second;
        const float *src_x = hX + lo, *src_y = needs_y ? hY + lo : nullptr;
        float *dst_h = hDst + lo;
        if (staged) {
            if (k >= (size_t)kSlots) {
                // this slot's staging buffers still belong to chunk k - kSlots: its result must be
                // handed to the caller, and its H2D must be off the wire, before they are reused
                if (pg_d) {
                    B200V_CUDA(cudaEventSynchronize(c.out_done[slot]));
                    pool.copy(hDst + sched[k - kSlots].first, c.stage_d[slot], sched[k - kSlots].second * 4);
                }
                if (pg_x || pg_y) B200V_CUDA(cudaEventSynchronize(c.in_done[slot]));
            }
            if (pg_x) { pool.copy(c.stage_x[slot], hX + lo, len * 4); src_x = c.stage_x[slot]; }
            if (pg_y) { pool.copy(c.stage_y[slot], hY + lo, len * 4); src_y = c.stage_y[slot]; }
            if (pg_d) dst_h = c.stage_d[slot];
        }
        // inputs of this slot are free once the kernel that read them (chunk k - kSlots) is done
        if (k >= (size_t)kSlots) B200V_CUDA(cudaStreamWaitEvent(c.s_in, c.k_done[slot], 0));
        if (k == 0) B200V_CUDA(cudaEventRecord(c.h2d_first, c.s_in));
        B200V_CUDA(cudaMemcpyAsync(c.slot_x[slot], src_x, len * 4, cudaMemcpyHostToDevice, c.s_in));
        if (needs_y)
            B200V_CUDA(cudaMemcpyAsync(c.slot_y[slot], src_y, len * 4, cudaMemcpyHostToDevice, c.s_in));
        B200V_CUDA(cudaEventRecord(c.in_done[slot], c.s_in));
        if (k + 1 == nchunks) B200V_CUDA(cudaEventRecord(c.h2d_last, c.s_in));

        B200V_CUDA(cudaStreamWaitEvent(c.s_k, c.in_done[slot], 0));
        // the output slot is free once its previous contents (chunk k - kSlots) reached the host
        if (k >= (size_t)kSlots) B200V_CUDA(cudaStreamWaitEvent(c.s_k, c.out_done[slot], 0));
        B200V_CUDA(cudaEventRecord(c.k_ev[2 * k], c.s_k));
        int gk = 0, bk = 0;
        rc = launch_ew(c, op, c.slot_d[slot], c.slot_x[slot], c.slot_y[slot], s, len, c.s_k, &gk, &bk);
        if (rc != B200V_OK) return rc;
        if (k == 0) { grid = gk; block = bk; }
        B200V_CUDA(cudaEventRecord(c.k_ev[2 * k + 1], c.s_k));
        B200V_CUDA(cudaEventRecord(c.k_done[slot], c.s_k));

        B200V_CUDA(cudaStreamWaitEvent(c.s_out, c.k_done[slot], 0));
        if (k == 0) B200V_CUDA(cudaEventRecord(c.d2h_first, c.s_out));
        B200V_CUDA(cudaMemcpyAsync(dst_h, c.slot_d[slot], len * 4, cudaMemcpyDeviceToHost, c.s_out));
        B200V_CUDA(cudaEventRecord(c.out_done[slot], c.s_out));
    }
    B200V_CUDA(cudaEventRecord(c.d2h_last, c.s_out));
    B200V_CUDA(cudaEventRecord(c.t_end, c.s_out));
    if (staged && pg_d) {   // hand over the results still sitting in the staging ring, oldest first
        for (size_t k = nchunks > (size_t)kSlots ? nchunks - kSlots : 0; k < nchunks; ++k) {
            const int slot = (int)(k % kSlots);
            B200V_CUDA(cudaEventSynchronize(c.out_done[slot]));
            pool.copy(hDst + sched[k].first, c.stage_d[slot], sched[k].second * 4);
        }
    }
    B200V_CUDA(cudaStreamSynchronize(c.s_out));
    B200V_CUDA(cudaStreamSynchronize(c.s_k));
    B200V_CUDA(cudaStreamSynchronize(c.s_in));

    if (timing) {
        B200V_CUDA(cudaEventElapsedTime(&timing->total_ms, c.t_begin, c.t_end));
        B200V_CUDA(cudaEventElapsedTime(&timing->h2d_ms, c.h2d_first, c.h2d_last));
        B200V_CUDA(cudaEventElapsedTime(&timing->d2h_ms, c.d2h_first, c.d2h_last));
        float sum = 0.f;
        for (size_t k = 0; k < nchunks; ++k) {
            float ms = 0.f;
            B200V_CUDA(cudaEventElapsedTime(&ms, c.k_ev[2 * k], c.k_ev[2 * k + 1]));
            sum += ms;
        }
        timing->kernel_ms = sum;
        timing->grid = grid;
        timing->block = block;
        timing->chunks = (int)nchunks;
        timing->kernel_launches = (int)nchunks;
    }
    return B200V_OK;
}

template <int OP>
void launch_check(const DeviceCtx &c, const float *dst, const float *x, const float *y, float s,
                  size_t n, unsigned long long *out, cudaStream_t st) {
    check_kernel<OP><<<grid_1d(c, n, 8), 256, 0, st>>>(dst, x, y, s, n, out);
}

int run_check(const DeviceCtx &c, int op, const float *dst, const float *x, const float *y, float s,
              size_t n, unsigned long long *d_out, cudaStream_t st) {
    switch (op) {
    case B200V_OP_COPY: launch_check<B200V_OP_COPY>(c, dst, x, y, s, n, d_out, st); break;
    case B200V_OP_SCALE: launch_check<B200V_OP_SCALE>(c, dst, x, y, s, n, d_out, st); break;
    case B200V_OP_ADD: launch_check<B200V_OP_ADD>(c, dst, x, y, s, n, d_out, st); break;
    case B200V_OP_TRIAD: launch_check<B200V_OP_TRIAD>(c, dst, x, y, s, n, d_out, st); break;
    case B200V_OP_VECTORADD: launch_check<B200V_OP_VECTORADD>(c, dst, x, y, s, n, d_out, st); break;
    case B200V_OP_TRIAD_UNFUSED: launch_check<B200V_OP_TRIAD_UNFUSED>(c, dst, x, y, s, n, d_out, st); break;
    default: return B200V_E_BADARG;
    }
    B200V_CUDA(cudaGetLastError());
    return B200V_OK;
}

// Device scratch freed on every exit path.
struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
};

int digest_impl(const DeviceCtx &c, const void *dWords, size_t n, uint64_t *digest, cudaStream_t st) {
    DevBuf acc;
    B200V_CUDA(cudaMalloc(&acc.p, 8));
    B200V_CUDA(cudaMemsetAsync(acc.p, 0, 8, st));
    if (n > 0) {
        digest_kernel<<<grid_1d(c, n, 8), 256, 0, st>>>((const uint32_t *)dWords, n,
                                                        (unsigned long long *)acc.p);
        B200V_CUDA(cudaGetLastError());
    }
    B200V_CUDA(cudaMemcpyAsync(digest, acc.p, 8, cudaMemcpyDeviceToHost, st));
    B200V_CUDA(cudaStreamSynchronize(st));
    return B200V_OK;
}

int check_impl(const DeviceCtx &c, int op, const float *dDst, const float *dX, const float *dY,
               float s, size_t n, uint64_t *mismatches, uint64_t *first, cudaStream_t st) {
    DevBuf out;
    B200V_CUDA(cudaMalloc(&out.p, 16));
    const unsigned long long init[2] = {0ull, ~0ull};
    B200V_CUDA(cudaMemcpyAsync(out.p, init, 16, cudaMemcpyHostToDevice, st));
    if (n > 0) {
        int rc = run_check(c, op, dDst, dX, dY, s, n, (unsigned long long *)out.p, st);
        if (rc != B200V_OK) return rc;
    }
    unsigned long long res[2] = {0, 0};
    B200V_CUDA(cudaMemcpyAsync(res, out.p, 16, cudaMemcpyDeviceToHost, st));
    B200V_CUDA(cudaStreamSynchronize(st));
    if (mismatches) *mismatches = res[0];
    if (first) *first = res[1];
    return res[0] ? B200V_E_MISMATCH : B200V_OK;
}

}  // namespace
}  // namespace b200v

using namespace b200v;

extern "C" {

void b200v_version(int *major, int *minor) {
    if (major) *major = B200V_VERSION_MAJOR;
    if (minor) *minor = B200V_VERSION_MINOR;
}

const char *b200v_strerror(int code) {
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    switch (code) {
    case B200V_OK: return "ok";
    case B200V_E_BADARG: return "bad argument";
    case B200V_E_MISMATCH: return "result verification failed";
    case B200V_E_NCCL: return "NCCL error";
    case B200V_E_NODEVICE: return "no CUDA device";
    case B200V_E_NONCCL: return "libnccl.so.2 not loadable";
    case B200V_E_NOMEM: return "host allocation failed";
    case B200V_E_TOOLARGE: return "size exceeds the fixed capacity of the no-allocation entry point";
    case B200V_E_UNSUPPORTED: return "not supported by these devices or this driver";
    case B200V_E_DRIVER: return "CUDA driver API call failed";
    case B200V_E_TIMEOUT: return "cross-GPU barrier timed out";
    default: return "unknown b200v error";
    }
}

int b200v_device_count(int *n) {
    if (!n) return B200V_E_BADARG;
    *n = 0;
    int c = 0;
    cudaError_t e = cudaGetDeviceCount(&c);
    if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) {
        (void)cudaGetLastError();
        return B200V_E_NODEVICE;
    }
    B200V_CUDA(e);
    *n = c;
    return c > 0 ? B200V_OK : B200V_E_NODEVICE;
}

int b200v_device_info(int dev, b200v_info *out) {
    if (!out) return B200V_E_BADARG;
    std::memset(out, 0, sizeof(*out));
    int count = 0;
    int rc = b200v_device_count(&count);
    if (rc != B200V_OK) return rc;
    if (dev < 0 || dev >= count) return B200V_E_BADARG;
    cudaDeviceProp p;
    B200V_CUDA(cudaGetDeviceProperties(&p, dev));
    std::snprintf(out->name, sizeof(out->name), "%s", p.name);
    out->cc_major = p.major;
    out->cc_minor = p.minor;
    out->sm_count = p.multiProcessorCount;
    out->l2_bytes = (uint64_t)p.l2CacheSize;
    out->hbm_bytes = (uint64_t)p.totalGlobalMem;
    out->mem_bus_bits = p.memoryBusWidth;
    B200V_CUDA(cudaDeviceGetAttribute(&out->mem_clock_khz, cudaDevAttrMemoryClockRate, dev));
    B200V_CUDA(cudaDeviceGetAttribute(&out->sm_clock_khz, cudaDevAttrClockRate, dev));
    return B200V_OK;
}

// Host buffers the DMA engines can use directly.  Three ways to get one, chosen per call by
// B200V_HOST_POLICY (default below; profiles/README.md has the measurements behind it):
//   pinned  cudaHostAlloc, portable
//   wc      the same, write-combined, for buffers the CPU only writes (B200V_HOST_WRITE_ONLY):
//           device reads of them need no CPU-cache snoop; the CPU must never read them
//   thp     anonymous memory advised into transparent huge pages, touched, then cudaHostRegister'ed:
//           512 x fewer pages for the IOMMU and the DMA page tables to walk
// The registry remembers how each pointer was made so b200v_host_free can undo it.
namespace {
struct HostBlock { int kind; void *raw; size_t raw_bytes; };
std::mutex g_host_mu;
std::map<void *, HostBlock> g_host_blocks;
enum { HOST_PINNED = 0, HOST_WC = 1, HOST_THP = 2 };

int host_policy() {
    const char *e = std::getenv("B200V_HOST_POLICY");
    if (e && !std::strcmp(e, "wc")) return HOST_WC;
    if (e && !std::strcmp(e, "thp")) return HOST_THP;
    return HOST_PINNED;
}
}  // namespace

int b200v_host_alloc_ex(void **ptr, size_t bytes, int flags) {
    if (!ptr) return B200V_E_BADARG;
    *ptr = nullptr;
    if (bytes == 0) return B200V_OK;
    int kind = host_policy();
    if (kind == HOST_WC && !(flags & B200V_HOST_WRITE_ONLY)) kind = HOST_PINNED;   // the CPU reads results
    HostBlock blk{kind, nullptr, 0};
    void *p = nullptr;
    if (kind == HOST_THP) {
        const size_t huge = (size_t)2 << 20;
        blk.raw_bytes = (bytes + huge - 1) / huge * huge + huge;
        blk.raw = ::mmap(nullptr, blk.raw_bytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        if (blk.raw == MAP_FAILED) return B200V_E_NOMEM;
        p = (void *)(((uintptr_t)blk.raw + huge - 1) / huge * huge);
        ::madvise(p, blk.raw_bytes - huge, MADV_HUGEPAGE);
        for (size_t off = 0; off < bytes; off += 4096) ((volatile char *)p)[off] = 0;   // fault in before pinning
        cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
        if (e != cudaSuccess) {
            (void)cudaGetLastError();
            ::munmap(blk.raw, blk.raw_bytes);
            return (int)e;
        }
    } else {
        B200V_CUDA(cudaHostAlloc(&p, bytes, cudaHostAllocPortable | (kind == HOST_WC ? cudaHostAllocWriteCombined : 0)));
    }
    {
        std::lock_guard<std::mutex> g(g_host_mu);
        g_host_blocks[p] = blk;
    }
    *ptr = p;
    return B200V_OK;
}

int b200v_host_alloc(void **ptr, size_t bytes) { return b200v_host_alloc_ex(ptr, bytes, 0); }

int b200v_host_free(void *ptr) {
    if (!ptr) return B200V_OK;
    HostBlock blk{HOST_PINNED, nullptr, 0};
    {
        std::lock_guard<std::mutex> g(g_host_mu);
        auto it = g_host_blocks.find(ptr);
        if (it != g_host_blocks.end()) {
            blk = it->second;
            g_host_blocks.erase(it);
        }
    }
    if (blk.kind == HOST_THP) {
        B200V_CUDA(cudaHostUnregister(ptr));
        ::munmap(blk.raw, blk.raw_bytes);
        return B200V_OK;
    }
    B200V_CUDA(cudaFreeHost(ptr));
    return B200V_OK;
}

int b200v_vectoradd(int dev, const float *hA, const float *hB, float *hC, size_t n,
                    b200v_timing *timing) {
    return host_pipeline(dev, B200V_OP_VECTORADD, hC, hA, hB, 0.0f, n, timing);
}

int b200v_stream_host(int dev, int op, float *hDst, const float *hX, const float *hY, float s,
                      size_t n, b200v_timing *timing) {
    return host_pipeline(dev, op, hDst, hX, hY, s, n, timing);
}

int b200v_stream_device(int dev, int op, float *dDst, const float *dX, const float *dY, float s,
                        size_t n, void *stream) {
    CtxGuard g;
    int rc = acquire_ctx(dev, &g.c);
    if (rc != B200V_OK) return rc;
    return launch_ew(*g.c, op, dDst, dX, dY, s, n, (cudaStream_t)stream, nullptr, nullptr);
}

int b200v_kernel_name(int op, const float *dDst, const float *dX, const float *dY, size_t n, char *buf,
                      size_t buflen) {
    return describe_ew(op, dDst, dX, dY, n, buf, buflen);
}

int b200v_vectoradd_device(int dev, const float *dA, const float *dB, float *dC, size_t n,
                           void *stream) {
    return b200v_stream_device(dev, B200V_OP_VECTORADD, dC, dA, dB, 0.0f, n, stream);
}

int b200v_vectoradd_device_sampleshape(int dev, const float *dA, const float *dB, float *dC,
                                       size_t n, void *stream) {
    CtxGuard g;
    int rc = acquire_ctx(dev, &g.c);
    if (rc != B200V_OK) return rc;
    return launch_sampleshape(B200V_OP_VECTORADD, dC, dA, dB, 0.0f, n, (cudaStream_t)stream);
}

int b200v_fill_device(int dev, float *dX, size_t n, uint64_t seed, void *stream) {
    CtxGuard g;
    int rc = acquire_ctx(dev, &g.c);
    if (rc != B200V_OK) return rc;
    if (n == 0) return B200V_OK;
    if (!dX) return B200V_E_BADARG;
    fill_hash_kernel<<<grid_1d(*g.c, n, 8), 256, 0, (cudaStream_t)stream>>>(dX, n, seed);
    B200V_CUDA(cudaGetLastError());
    return B200V_OK;
}

int b200v_digest_device(int dev, const void *dWords, size_t n, uint64_t *digest, void *stream) {
    if (!digest || (n > 0 && !dWords)) return B200V_E_BADARG;
    CtxGuard g;
    int rc = acquire_ctx(dev, &g.c);
    if (rc != B200V_OK) return rc;
    return digest_impl(*g.c, dWords, n, digest, (cudaStream_t)stream);
}

int b200v_check_device(int dev, int op, const float *dDst, const float *dX, const float *dY,
                       float s, size_t n, uint64_t *mismatches, uint64_t *first_mismatch,
                       void *stream) {
    if (op < 0 || op >= B200V_OP_COUNT) return B200V_E_BADARG;
    if (n > 0 && (!dDst || !dX || (ew_needs_y(op) && !dY))) return B200V_E_BADARG;
    CtxGuard g;
    int rc = acquire_ctx(dev, &g.c);
    if (rc != B200V_OK) return rc;
    return check_impl(*g.c, op, dDst, dX, dY, s, n, mismatches, first_mismatch, (cudaStream_t)stream);
}

int b200v_stream(int dev, int op, size_t n, float s, uint64_t seed, int warmup, int iters,
                 b200v_stream_result *r) {
    if (!r || op < 0 || op >= B200V_OP_COUNT || iters < 1 || warmup < 0 || n == 0) return B200V_E_BADARG;
    std::memset(r, 0, sizeof(*r));
    CtxGuard g;
    int rc = acquire_ctx(dev, &g.c);
    if (rc != B200V_OK) return rc;
    DeviceCtx &c = *g.c;
    cudaStream_t st = c.s_k;
    const bool needs_y = ew_needs_y(op);

    DevBuf x, y, d;
    B200V_CUDA(cudaMalloc(&x.p, n * 4));
    if (needs_y) B200V_CUDA(cudaMalloc(&y.p, n * 4));
    B200V_CUDA(cudaMalloc(&d.p, n * 4));
    const int gfill = grid_1d(c, n, 8);
    fill_hash_kernel<<<gfill, 256, 0, st>>>((float *)x.p, n, seed);
    if (needs_y) fill_hash_kernel<<<gfill, 256, 0, st>>>((float *)y.p, n, seed + B200V_SEED_Y_OFFSET);
    B200V_CUDA(cudaGetLastError());
    B200V_CUDA(cudaMemsetAsync(d.p, 0xFF, n * 4, st));

    for (int i = 0; i < warmup; ++i) {
        rc = launch_ew(c, op, (float *)d.p, (const float *)x.p, (const float *)y.p, s, n, st, nullptr, nullptr);
        if (rc != B200V_OK) return rc;
    }
    std::vector<cudaEvent_t> ev((size_t)iters * 2);
    for (auto &e : ev) B200V_CUDA(cudaEventCreate(&e));
    struct EvFree {
        std::vector<cudaEvent_t> &v;
        ~EvFree() { for (auto e : v) cudaEventDestroy(e); }
    } evfree{ev};
    // A launch of a small array lasts a few microseconds, comparable to the gap between two
    // launches and to the resolution of an event pair.  Each timed sample is therefore `reps`
    // back-to-back launches sized to last >= ~200 us (one launch once the array is large), and the
    // sample's time is divided by reps.
    const double est_us = (double)n * ew_bytes_per_elem(op) / 5.0e6 + 2.0;   // at ~5 TB/s, + launch
    const int reps = (int)std::min(256.0, std::max(1.0, 200.0 / est_us));
    for (int i = 0; i < iters; ++i) {
        B200V_CUDA(cudaEventRecord(ev[2 * i], st));
        for (int k = 0; k < reps; ++k) {
            rc = launch_ew(c, op, (float *)d.p, (const float *)x.p, (const float *)y.p, s, n, st, &r->grid, &r->block);
            if (rc != B200V_OK) return rc;
        }
        B200V_CUDA(cudaEventRecord(ev[2 * i + 1], st));
    }
    B200V_CUDA(cudaStreamSynchronize(st));
    std::vector<float> ms((size_t)iters);
    double sum = 0;
    for (int i = 0; i < iters; ++i) {
        B200V_CUDA(cudaEventElapsedTime(&ms[i], ev[2 * i], ev[2 * i + 1]));
        ms[i] /= (float)reps;
        sum += ms[i];
    }
    std::sort(ms.begin(), ms.end());
    r->best_ms = ms[0];
    r->median_ms = ms[(size_t)iters / 2];
    r->mean_ms = (float)(sum / iters);
    r->bytes_per_launch = (uint64_t)n * (uint64_t)ew_bytes_per_elem(op);
    r->best_gbs = (double)r->bytes_per_launch / (r->best_ms * 1e-3) / 1e9;
    r->median_gbs = (double)r->bytes_per_launch / (r->median_ms * 1e-3) / 1e9;
    r->kernel_launches = iters * reps;
    r->l2_resident = ((size_t)(needs_y ? 3 : 2) * n * 4 <= c.l2_bytes) ? 1 : 0;

    rc = check_impl(c, op, (const float *)d.p, (const float *)x.p, (const float *)y.p, s, n,
                    &r->mismatches, &r->first_mismatch, st);
    if (rc != B200V_OK && rc != B200V_E_MISMATCH) return rc;
    int rc2 = digest_impl(c, d.p, n, &r->digest, st);
    if (rc2 != B200V_OK) return rc2;
    return rc;
}

int b200v_shutdown(void) { return destroy_all_ctx(); }

}  // extern "C"
