// nccl_probe.cu — the NVLink health probe: an NCCL fp32 sum all-reduce sweep (SURVEY.md §8 a10,
// §8d cfg4).  This is the only place the library touches NCCL, and it does so through dlopen so
// that libb200validate.so loads (and validates a GPU) on images that ship no NCCL at all.
//
// Inputs are integer-valued floats — rank r holds (float)((i*(2r+1) + r) & 15) — so the reduced value is
// exact whatever order the ring, tree or NVSwitch reduction adds in, and the result is checked
// bit for bit against the closed form on the device.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <mutex>
#include <nccl.h>
#include <new>   // types and enums only; every function is resolved with dlsym
#include <vector>

#include "ar_common.cuh"
#include "common.cuh"

namespace b200v {
namespace {

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                              cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    // optional (NCCL >= 2.19): buffers from NCCL's own allocator, registered with the communicator,
    // let NVLS / copy-engine paths work on the user buffer directly instead of staging through FIFOs
    ncclResult_t (*MemAlloc)(void **, size_t) = nullptr;
    ncclResult_t (*MemFree)(void *) = nullptr;
    ncclResult_t (*CommRegister)(ncclComm_t, void *, size_t, void **) = nullptr;
    ncclResult_t (*CommDeregister)(ncclComm_t, void *) = nullptr;
};

NcclApi g_nccl;
std::once_flag g_nccl_once;
int g_last_nccl = 0;

template <typename F>
bool sym(void *h, const char *name, F &out) {
    out = reinterpret_cast<F>(dlsym(h, name));
    return out != nullptr;
}

const NcclApi *nccl() {
    std::call_once(g_nccl_once, [] {
        // Order matters when the host process carries an NCCL of its own (a Python process with
        // torch does): two builds share the soname libnccl.so.2, and whichever is mapped first is
        // the one every later DT_NEEDED resolves to.  So: the library the caller names
        // (B200V_NCCL_LIB), else one the process has already loaded, else the system's — and
        // never RTLD_GLOBAL, every symbol is fetched from the handle.
        void *h = nullptr;
        if (const char *path = std::getenv("B200V_NCCL_LIB")) h = dlopen(path, RTLD_NOW | RTLD_LOCAL);
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL | RTLD_NOLOAD);
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
        if (!h) return;
        NcclApi a;
        a.handle = h;
        bool ok = sym(h, "ncclGetVersion", a.GetVersion) && sym(h, "ncclGetUniqueId", a.GetUniqueId) &&
                  sym(h, "ncclCommInitRank", a.CommInitRank) && sym(h, "ncclCommInitAll", a.CommInitAll) &&
                  sym(h, "ncclCommDestroy", a.CommDestroy) && sym(h, "ncclAllReduce", a.AllReduce) &&
                  sym(h, "ncclGroupStart", a.GroupStart) && sym(h, "ncclGroupEnd", a.GroupEnd);
        sym(h, "ncclMemAlloc", a.MemAlloc);
        sym(h, "ncclMemFree", a.MemFree);
        sym(h, "ncclCommRegister", a.CommRegister);
        sym(h, "ncclCommDeregister", a.CommDeregister);
        if (ok) g_nccl = a;
    });
    return g_nccl.handle ? &g_nccl : nullptr;
}

#define B200V_NCCL(expr, rowptr)                                  \
    do {                                                          \
        ncclResult_t r__ = (expr);                                \
        if (r__ != ncclSuccess) {                                 \
            g_last_nccl = (int)r__;                               \
            if (rowptr) (rowptr)->nccl_result = (int)r__;         \
            return B200V_E_NCCL;                                  \
        }                                                         \
    } while (0)

struct Rank {
    int dev = -1, rank = -1;
    ncclComm_t comm = nullptr;
    cudaStream_t st = nullptr;
    float *buf = nullptr;
    size_t buf_bytes = 0;
    bool nccl_mem = false;        // buf came from ncclMemAlloc
    void *reg_handle = nullptr;   // ncclCommRegister handle
    unsigned long long *d_bad = nullptr;
    int sm_count = 0;
    std::vector<cudaEvent_t> ev;
};

void destroy_ranks(const NcclApi *api, std::vector<Rank> &rs) {
    for (Rank &r : rs) {
        if (r.dev < 0) continue;
        cudaSetDevice(r.dev);
        if (r.st) cudaStreamSynchronize(r.st);
        if (r.comm && r.reg_handle && api->CommDeregister) api->CommDeregister(r.comm, r.reg_handle);
        if (r.comm) api->CommDestroy(r.comm);
        for (cudaEvent_t e : r.ev) cudaEventDestroy(e);
        if (r.buf && r.nccl_mem) api->MemFree(r.buf);
        else if (r.buf) cudaFree(r.buf);
        if (r.d_bad) cudaFree(r.d_bad);
        if (r.st) cudaStreamDestroy(r.st);
    }
    rs.clear();
}

bool plain_buffers() {
    const char *e = std::getenv("B200V_NCCL_PLAIN_BUFFERS");
    return e && *e && *e != '0';
}

int setup_rank(const NcclApi *api, Rank &r, size_t max_bytes, int iters) {
    B200V_CUDA(cudaSetDevice(r.dev));
    B200V_CUDA(cudaDeviceGetAttribute(&r.sm_count, cudaDevAttrMultiProcessorCount, r.dev));
    B200V_CUDA(cudaStreamCreateWithFlags(&r.st, cudaStreamNonBlocking));
    r.buf_bytes = max_bytes;
    if (api->MemAlloc && api->MemFree && !plain_buffers()) {
        void *p = nullptr;
        if (api->MemAlloc(&p, max_bytes) == ncclSuccess && p) {
            r.buf = (float *)p;
            r.nccl_mem = true;
        }
    }
    if (!r.buf) B200V_CUDA(cudaMalloc(&r.buf, max_bytes));
    B200V_CUDA(cudaMalloc(&r.d_bad, 8));
    r.ev.resize((size_t)iters * 2);
    for (auto &e : r.ev) B200V_CUDA(cudaEventCreate(&e));
    return B200V_OK;
}

// After the communicator exists: register the buffer with it (best effort; the probe is still
// valid unregistered, only slower).
void register_buffer(const NcclApi *api, Rank &r) {
    if (!r.nccl_mem || !api->CommRegister || !api->CommDeregister) return;
    cudaSetDevice(r.dev);
    void *h = nullptr;
    if (api->CommRegister(r.comm, r.buf, r.buf_bytes, &h) == ncclSuccess) r.reg_handle = h;
}

int grid_for(const Rank &r, size_t n) {
    size_t g = (n + 255) / 256;
    const size_t cap = (size_t)r.sm_count * 8;
    if (g > cap) g = cap;
    return g < 1 ? 1 : (int)g;
}

// One all-reduce across the local ranks, grouped when this process drives more than one device.
int allreduce_once(const NcclApi *api, std::vector<Rank> &rs, size_t count, b200v_ar_row *row) {
    if (rs.size() > 1) B200V_NCCL(api->GroupStart(), row);
    for (Rank &r : rs) {
        B200V_CUDA(cudaSetDevice(r.dev));
        B200V_NCCL(api->AllReduce(r.buf, r.buf, count, ncclFloat, ncclSum, r.comm, r.st), row);
    }
    if (rs.size() > 1) B200V_NCCL(api->GroupEnd(), row);
    return B200V_OK;
}

int sweep(const NcclApi *api, std::vector<Rank> &rs, int nranks, size_t min_bytes, size_t max_bytes,
          int warmup, int iters, b200v_ar_row *rows, int max_rows, int *nrows) {
    int nr = 0;
    for (size_t bytes = min_bytes; bytes <= max_bytes && nr < max_rows; bytes *= 2) {
        b200v_ar_row *row = &rows[nr];
        std::memset(row, 0, sizeof(*row));
        row->bytes = bytes;
        const size_t count = bytes / 4;

        // correctness pass: fill, one all-reduce, bitwise check against the closed form
        auto fill_reduce_check = [&]() -> int {
            for (Rank &r : rs) {
                B200V_CUDA(cudaSetDevice(r.dev));
                ar_fill_kernel<<<grid_for(r, count), 256, 0, r.st>>>(r.buf, count, r.rank);
                B200V_CUDA(cudaGetLastError());
                B200V_CUDA(cudaMemsetAsync(r.d_bad, 0, 8, r.st));
            }
            int rc2 = allreduce_once(api, rs, count, row);
            if (rc2 != B200V_OK) return rc2;
            for (Rank &r : rs) {
                B200V_CUDA(cudaSetDevice(r.dev));
                ar_check_kernel<<<grid_for(r, count), 256, 0, r.st>>>(r.buf, count, nranks, r.d_bad);
                B200V_CUDA(cudaGetLastError());
                unsigned long long bad = 0;
                B200V_CUDA(cudaMemcpyAsync(&bad, r.d_bad, 8, cudaMemcpyDeviceToHost, r.st));
                B200V_CUDA(cudaStreamSynchronize(r.st));
                row->mismatches += bad;
            }
            return B200V_OK;
        };
        int rc = fill_reduce_check();
        if (rc != B200V_OK) return rc;

        // timing pass (values keep accumulating in place; they are not read again).  A small
        // all-reduce lasts a few microseconds, comparable to the resolution of an event pair, so a
        // timed sample is `reps` back-to-back calls lasting >= ~100 us in total (the same rule as
        // the hand-written probe, so the two tables compare like with like).
        for (int i = 0; i < warmup; ++i) {
            rc = allreduce_once(api, rs, count, row);
            if (rc != B200V_OK) return rc;
        }
        const double est_us = (double)bytes / 300e3 + 8.0;
        const int reps = (int)std::min(16.0, std::max(1.0, 100.0 / est_us));
        for (int i = 0; i < iters; ++i) {
            for (Rank &r : rs) {
                B200V_CUDA(cudaSetDevice(r.dev));
                B200V_CUDA(cudaEventRecord(r.ev[2 * i], r.st));
            }
            for (int k = 0; k < reps; ++k) {
                rc = allreduce_once(api, rs, count, row);
                if (rc != B200V_OK) return rc;
            }
            for (Rank &r : rs) {
                B200V_CUDA(cudaSetDevice(r.dev));
                B200V_CUDA(cudaEventRecord(r.ev[2 * i + 1], r.st));
            }
        }
        for (Rank &r : rs) {
            B200V_CUDA(cudaSetDevice(r.dev));
            B200V_CUDA(cudaStreamSynchronize(r.st));
        }
        std::vector<float> ms((size_t)iters, 0.f);
        for (int i = 0; i < iters; ++i)
            for (Rank &r : rs) {
                float t = 0.f;
                B200V_CUDA(cudaEventElapsedTime(&t, r.ev[2 * i], r.ev[2 * i + 1]));
                ms[i] = std::max(ms[i], t / (float)reps);   // max over the ranks this process drives
            }
        // and once more after the queued launches, re-filled, so the last state is checked too
        rc = fill_reduce_check();
        if (rc != B200V_OK) return rc;
        std::sort(ms.begin(), ms.end());
        row->best_ms = ms[0];
        row->median_ms = ms[(size_t)iters / 2];
        row->algbw_gbs = (double)bytes / (row->best_ms * 1e-3) / 1e9;
        row->busbw_gbs = row->algbw_gbs * 2.0 * (nranks - 1) / nranks;
        ++nr;
        if (bytes > max_bytes / 2) break;   // next doubling would overflow or pass max
    }
    *nrows = nr;
    return B200V_OK;
}

bool bad_sweep_args(size_t min_bytes, size_t max_bytes, int warmup, int iters, const void *rows,
                    int max_rows, const int *nrows) {
    return !rows || !nrows || max_rows < 1 || iters < 1 || warmup < 0 || min_bytes < 4 ||
           (min_bytes & 3) || max_bytes < min_bytes;
}

}  // namespace
}  // namespace b200v

using namespace b200v;

extern "C" {

int b200v_nccl_version(int *version) {
    if (!version) return B200V_E_BADARG;
    const NcclApi *api = nccl();
    if (!api) return B200V_E_NONCCL;
    B200V_NCCL(api->GetVersion(version), (b200v_ar_row *)nullptr);
    return B200V_OK;
}

int b200v_last_nccl_error(void) { return g_last_nccl; }

int b200v_nccl_unique_id(unsigned char id[B200V_NCCL_UNIQUE_ID_BYTES]) {
    static_assert(sizeof(ncclUniqueId) == B200V_NCCL_UNIQUE_ID_BYTES, "ncclUniqueId is 128 bytes");
    if (!id) return B200V_E_BADARG;
    const NcclApi *api = nccl();
    if (!api) return B200V_E_NONCCL;
    ncclUniqueId u;
    B200V_NCCL(api->GetUniqueId(&u), (b200v_ar_row *)nullptr);
    std::memcpy(id, &u, sizeof(u));
    return B200V_OK;
}

int b200v_allreduce_probe(int ndev, const int *devs, size_t min_bytes, size_t max_bytes, int warmup,
                          int iters, b200v_ar_row *rows, int max_rows, int *nrows) {
    if (ndev < 1 || !devs || bad_sweep_args(min_bytes, max_bytes, warmup, iters, rows, max_rows, nrows))
        return B200V_E_BADARG;
    *nrows = 0;
    int count = 0;
    int rc = b200v_device_count(&count);
    if (rc != B200V_OK) return rc;
    for (int i = 0; i < ndev; ++i)
        if (devs[i] < 0 || devs[i] >= count) return B200V_E_BADARG;
    const NcclApi *api = nccl();
    if (!api) return B200V_E_NONCCL;

    std::vector<Rank> rs((size_t)ndev);
    struct Cleanup {
        const NcclApi *api;
        std::vector<Rank> &rs;
        ~Cleanup() { destroy_ranks(api, rs); }
    } cleanup{api, rs};
    for (int i = 0; i < ndev; ++i) {
        rs[i].dev = devs[i];
        rs[i].rank = i;
        rc = setup_rank(api, rs[i], max_bytes, iters);
        if (rc != B200V_OK) return rc;
    }
    std::vector<ncclComm_t> comms((size_t)ndev);
    B200V_NCCL(api->CommInitAll(comms.data(), ndev, devs), (b200v_ar_row *)nullptr);
    for (int i = 0; i < ndev; ++i) {
        rs[i].comm = comms[i];
        register_buffer(api, rs[i]);
    }
    return sweep(api, rs, ndev, min_bytes, max_bytes, warmup, iters, rows, max_rows, nrows);
}

int b200v_allreduce_probe_rank(int dev, int rank, int nranks,
                               const unsigned char id[B200V_NCCL_UNIQUE_ID_BYTES], size_t min_bytes,
                               size_t max_bytes, int warmup, int iters, b200v_ar_row *rows,
                               int max_rows, int *nrows) {
    if (!id || nranks < 1 || rank < 0 || rank >= nranks ||
        bad_sweep_args(min_bytes, max_bytes, warmup, iters, rows, max_rows, nrows))
        return B200V_E_BADARG;
    *nrows = 0;
    int count = 0;
    int rc = b200v_device_count(&count);
    if (rc != B200V_OK) return rc;
    if (dev < 0 || dev >= count) return B200V_E_BADARG;
    const NcclApi *api = nccl();
    if (!api) return B200V_E_NONCCL;

    std::vector<Rank> rs(1);
    struct Cleanup {
        const NcclApi *api;
        std::vector<Rank> &rs;
        ~Cleanup() { destroy_ranks(api, rs); }
    } cleanup{api, rs};
    rs[0].dev = dev;
    rs[0].rank = rank;
    rc = setup_rank(api, rs[0], max_bytes, iters);
    if (rc != B200V_OK) return rc;
    ncclUniqueId u;
    std::memcpy(&u, id, sizeof(u));
    B200V_NCCL(api->CommInitRank(&rs[0].comm, nranks, u, rank), (b200v_ar_row *)nullptr);
    register_buffer(api, rs[0]);
    return sweep(api, rs, nranks, min_bytes, max_bytes, warmup, iters, rows, max_rows, nrows);
}

// ---- all-reduce of a caller-owned device buffer -----------------------------------------------------

}  // extern "C"

struct b200v_ar_comm {
    int dev = -1;
    ncclComm_t comm = nullptr;
};

extern "C" {

int b200v_ar_comm_create(int dev, int rank, int nranks, const unsigned char id[B200V_NCCL_UNIQUE_ID_BYTES],
                         b200v_ar_comm **out) {
    if (!out) return B200V_E_BADARG;
    *out = nullptr;
    if (!id || nranks < 1 || rank < 0 || rank >= nranks) return B200V_E_BADARG;
    int count = 0;
    int rc = b200v_device_count(&count);
    if (rc != B200V_OK) return rc;
    if (dev < 0 || dev >= count) return B200V_E_BADARG;
    const NcclApi *api = nccl();
    if (!api) return B200V_E_NONCCL;
    B200V_CUDA(cudaSetDevice(dev));
    b200v_ar_comm *c = new (std::nothrow) b200v_ar_comm();
    if (!c) return B200V_E_NOMEM;
    c->dev = dev;
    ncclUniqueId u;
    std::memcpy(&u, id, sizeof(u));
    ncclResult_t r = api->CommInitRank(&c->comm, nranks, u, rank);
    if (r != ncclSuccess) {
        g_last_nccl = (int)r;
        delete c;
        return B200V_E_NCCL;
    }
    *out = c;
    return B200V_OK;
}

int b200v_ar_comm_allreduce(b200v_ar_comm *c, float *dBuf, size_t count, void *stream) {
    if (!c || !c->comm) return B200V_E_BADARG;
    if (count == 0) return B200V_OK;
    if (!dBuf) return B200V_E_BADARG;
    const NcclApi *api = nccl();
    if (!api) return B200V_E_NONCCL;
    B200V_CUDA(cudaSetDevice(c->dev));
    B200V_NCCL(api->AllReduce(dBuf, dBuf, count, ncclFloat, ncclSum, c->comm, (cudaStream_t)stream),
               (b200v_ar_row *)nullptr);
    return B200V_OK;
}

int b200v_ar_comm_destroy(b200v_ar_comm *c) {
    if (!c) return B200V_OK;
    const NcclApi *api = nccl();
    if (api && c->comm) {
        cudaSetDevice(c->dev);
        cudaDeviceSynchronize();
        api->CommDestroy(c->comm);
    }
    delete c;
    return B200V_OK;
}

}  // extern "C"
