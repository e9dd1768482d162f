// linkbw.cu — link measurements that the design documents quote as numbers, not sentences:
//
//   b200v-linkbw p2p  [ngpu]   cudaMemcpyPeerAsync between GPU pairs over NVLink, one direction and
//                              both directions at once — the "link rate" the all-reduce kernels are
//                              judged against;
//   b200v-linkbw pcie [ngpu]   host<->device copies: every GPU alone, every pair, groups of 4, all
//                              at once; H2D only, D2H only, both; with plain pinned, write-combined
//                              and transparent-huge-page-backed registered host buffers; and what the
//                              host cores themselves read/copy from the same DRAM — this names what
//                              limits the host-buffer (e2e) path when several GPUs run at once.
//
// One process, one host thread per GPU, one cudaMemcpyAsync per chunk (never the batched copy
// calls).  Prints one JSON object per line.  Not part of the library; built by `make tune`.
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#include <string>
#include <sys/mman.h>
#include <thread>
#include <vector>

#define CK(expr)                                                                                  \
    do {                                                                                          \
        cudaError_t e__ = (expr);                                                                 \
        if (e__ != cudaSuccess) {                                                                 \
            std::fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #expr, cudaGetErrorString(e__)); \
            std::exit(2);                                                                         \
        }                                                                                         \
    } while (0)

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---- p2p ---------------------------------------------------------------------------------------------
static void p2p_pair(int a, int b, size_t bytes, int iters) {
    void *pa = nullptr, *pb = nullptr, *qa = nullptr, *qb = nullptr;
    cudaStream_t sa, sb;
    cudaEvent_t a0, a1, b0, b1;
    CK(cudaSetDevice(a));
    cudaDeviceEnablePeerAccess(b, 0);
    (void)cudaGetLastError();
    CK(cudaMalloc(&pa, bytes)); CK(cudaMalloc(&qa, bytes));
    CK(cudaStreamCreateWithFlags(&sa, cudaStreamNonBlocking));
    CK(cudaEventCreate(&a0)); CK(cudaEventCreate(&a1));
    CK(cudaSetDevice(b));
    cudaDeviceEnablePeerAccess(a, 0);
    (void)cudaGetLastError();
    CK(cudaMalloc(&pb, bytes)); CK(cudaMalloc(&qb, bytes));
    CK(cudaStreamCreateWithFlags(&sb, cudaStreamNonBlocking));
    CK(cudaEventCreate(&b0)); CK(cudaEventCreate(&b1));
    auto run = [&](bool ab, bool ba, float &ms_ab, float &ms_ba) {
        for (int rep = 0; rep < 2; ++rep) {                  // first pass is the warm-up
            if (ab) { CK(cudaSetDevice(a)); CK(cudaEventRecord(a0, sa)); }
            if (ba) { CK(cudaSetDevice(b)); CK(cudaEventRecord(b0, sb)); }
            for (int i = 0; i < iters; ++i) {
                if (ab) CK(cudaMemcpyPeerAsync(pb, b, pa, a, bytes, sa));     // a pushes to b
                if (ba) CK(cudaMemcpyPeerAsync(qa, a, qb, b, bytes, sb));     // b pushes to a
            }
            if (ab) { CK(cudaSetDevice(a)); CK(cudaEventRecord(a1, sa)); }
            if (ba) { CK(cudaSetDevice(b)); CK(cudaEventRecord(b1, sb)); }
            if (ab) CK(cudaStreamSynchronize(sa));
            if (ba) CK(cudaStreamSynchronize(sb));
        }
        ms_ab = ms_ba = 0.f;
        if (ab) CK(cudaEventElapsedTime(&ms_ab, a0, a1));
        if (ba) CK(cudaEventElapsedTime(&ms_ba, b0, b1));
    };
    float uni = 0, dummy = 0, bi_ab = 0, bi_ba = 0;
    run(true, false, uni, dummy);
    run(true, true, bi_ab, bi_ba);
    const double gb = (double)bytes * iters / 1e9;
    std::printf("{\"mode\":\"p2p\",\"src\":%d,\"dst\":%d,\"bytes\":%zu,\"iters\":%d,\"uni_gbs\":%.1f,"
                "\"bi_ab_gbs\":%.1f,\"bi_ba_gbs\":%.1f}\n",
                a, b, bytes, iters, gb / (uni * 1e-3), gb / (bi_ab * 1e-3), gb / (bi_ba * 1e-3));
    std::fflush(stdout);
    CK(cudaSetDevice(a)); cudaFree(pa); cudaFree(qa); cudaStreamDestroy(sa);
    CK(cudaSetDevice(b)); cudaFree(pb); cudaFree(qb); cudaStreamDestroy(sb);
}

// ---- pcie ----------------------------------------------------------------------------------------------
enum HostKind { PINNED = 0, WRITE_COMBINED = 1, THP_REGISTERED = 2 };
static const char *kind_name(int k) { return k == PINNED ? "pinned" : k == WRITE_COMBINED ? "write_combined" : "thp_registered"; }

struct HostBuf {
    void *p = nullptr;
    size_t bytes = 0;
    int kind = PINNED;
    bool ok = false;
};

static HostBuf host_alloc(size_t bytes, int kind) {
    HostBuf h;
    h.bytes = bytes;
    h.kind = kind;
    if (kind == THP_REGISTERED) {
        const size_t huge = (size_t)2 << 20;
        void *raw = mmap(nullptr, bytes + huge, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        if (raw == MAP_FAILED) return h;
        void *al = (void *)(((uintptr_t)raw + huge - 1) / huge * huge);
        madvise(al, bytes, MADV_HUGEPAGE);
        std::memset(al, 1, bytes);                       // fault it in (as huge pages where the kernel agrees)
        if (cudaHostRegister(al, bytes, cudaHostRegisterPortable) != cudaSuccess) {
            (void)cudaGetLastError();
            return h;
        }
        h.p = al;
    } else {
        unsigned flags = cudaHostAllocPortable | (kind == WRITE_COMBINED ? cudaHostAllocWriteCombined : 0);
        if (cudaHostAlloc(&h.p, bytes, flags) != cudaSuccess) {
            (void)cudaGetLastError();
            return h;
        }
        if (kind != WRITE_COMBINED) std::memset(h.p, 1, bytes);
        else for (size_t i = 0; i < bytes; i += 4096) ((volatile char *)h.p)[i] = 1;
    }
    h.ok = true;
    return h;
}

struct Gpu {
    int dev;
    void *d_in = nullptr, *d_out = nullptr;
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaEvent_t e0, e1, f0, f1;
    HostBuf h_in[3], h_out[3];          // per host-memory kind
};

// All GPUs in `set` copy at once.  dir: 1 = H2D, 2 = D2H, 3 = both.  Returns aggregate GB/s over the
// wall time from the common start to the last GPU's completion; per[i] = that GPU's own rate(s).
static double run_set(std::vector<Gpu> &g, const std::vector<int> &set, int dir, int kind, size_t bytes,
                      int iters, std::vector<double> &per) {
    std::atomic<int> ready{0}, go{0};
    std::vector<std::thread> th;
    per.assign(set.size(), 0.0);
    std::vector<double> done(set.size(), 0.0);
    double t_start = 0;
    for (size_t k = 0; k < set.size(); ++k)
        th.emplace_back([&, k] {
            Gpu &x = g[(size_t)set[k]];
            CK(cudaSetDevice(x.dev));
            ready.fetch_add(1);
            while (!go.load()) {}
            if (dir & 1) CK(cudaEventRecord(x.e0, x.s_in));
            if (dir & 2) CK(cudaEventRecord(x.f0, x.s_out));
            for (int i = 0; i < iters; ++i) {
                if (dir & 1) CK(cudaMemcpyAsync(x.d_in, x.h_in[kind].p, bytes, cudaMemcpyHostToDevice, x.s_in));
                if (dir & 2) CK(cudaMemcpyAsync(x.h_out[kind].p, x.d_out, bytes, cudaMemcpyDeviceToHost, x.s_out));
            }
            if (dir & 1) CK(cudaEventRecord(x.e1, x.s_in));
            if (dir & 2) CK(cudaEventRecord(x.f1, x.s_out));
            if (dir & 1) CK(cudaStreamSynchronize(x.s_in));
            if (dir & 2) CK(cudaStreamSynchronize(x.s_out));
            done[k] = now_s();
            float a = 0, b = 0;
            if (dir & 1) CK(cudaEventElapsedTime(&a, x.e0, x.e1));
            if (dir & 2) CK(cudaEventElapsedTime(&b, x.f0, x.f1));
            const double gb = (double)bytes * iters / 1e9;
            per[k] = ((dir & 1) ? gb / (a * 1e-3) : 0.0) + ((dir & 2) ? gb / (b * 1e-3) : 0.0);
        });
    while (ready.load() < (int)set.size()) {}
    t_start = now_s();
    go.store(1);
    for (auto &t : th) t.join();
    double t_end = 0;
    for (double d : done) t_end = d > t_end ? d : t_end;
    const double total = (double)bytes * iters * set.size() * ((dir == 3) ? 2 : 1) / 1e9;
    return total / (t_end - t_start);
}

static void emit(const char *what, const std::vector<int> &set, int dir, int kind, double agg,
                 const std::vector<double> &per) {
    std::string gl, pl;
    for (size_t i = 0; i < set.size(); ++i) {
        gl += (i ? "," : "") + std::to_string(set[i]);
        char b[32];
        std::snprintf(b, sizeof b, "%s%.1f", i ? "," : "", per[i]);
        pl += b;
    }
    std::printf("{\"mode\":\"pcie\",\"what\":\"%s\",\"gpus\":[%s],\"dir\":\"%s\",\"host\":\"%s\",\"aggregate_gbs\":%.1f,"
                "\"per_gpu_gbs\":[%s]}\n",
                what, gl.c_str(), dir == 1 ? "h2d" : dir == 2 ? "d2h" : "both", kind_name(kind), agg, pl.c_str());
    std::fflush(stdout);
}

// What the host cores get from the same DRAM: a read-only pass (the shape of the sample's
// verification loop) and a copy, on `threads` threads over `bytes` each.
static void host_ceiling(int threads, size_t bytes) {
    std::vector<float *> src((size_t)threads), dst((size_t)threads);
    for (int t = 0; t < threads; ++t) {
        src[(size_t)t] = (float *)std::malloc(bytes);
        dst[(size_t)t] = (float *)std::malloc(bytes);
        std::memset(src[(size_t)t], 1, bytes);
        std::memset(dst[(size_t)t], 2, bytes);
    }
    for (int mode = 0; mode < 2; ++mode) {
        std::atomic<int> ready{0}, go{0};
        std::vector<std::thread> th;
        std::vector<double> sink((size_t)threads, 0.0);
        const int reps = 6;
        for (int t = 0; t < threads; ++t)
            th.emplace_back([&, t] {
                ready.fetch_add(1);
                while (!go.load()) {}
                for (int r = 0; r < reps; ++r) {
                    if (mode == 0) {
                        const float *p = src[(size_t)t];
                        float s0 = 0, s1 = 0, s2 = 0, s3 = 0;
                        for (size_t i = 0; i + 3 < bytes / 4; i += 4) { s0 += p[i]; s1 += p[i + 1]; s2 += p[i + 2]; s3 += p[i + 3]; }
                        sink[(size_t)t] += s0 + s1 + s2 + s3;
                    } else {
                        std::memcpy(dst[(size_t)t], src[(size_t)t], bytes);
                    }
                }
            });
        while (ready.load() < threads) {}
        const double t0 = now_s();
        go.store(1);
        for (auto &x : th) x.join();
        const double dt = now_s() - t0;
        const double gb = (double)bytes * reps * threads / 1e9;
        std::printf("{\"mode\":\"host\",\"what\":\"%s\",\"threads\":%d,\"bytes_per_thread\":%zu,\"gbs\":%.1f,"
                    "\"dram_traffic_gbs\":%.1f,\"sink\":%.0f}\n",
                    mode == 0 ? "read" : "memcpy", threads, bytes, gb / dt, (mode == 0 ? 1 : 2) * gb / dt, sink[0]);
        std::fflush(stdout);
    }
    for (int t = 0; t < threads; ++t) { std::free(src[(size_t)t]); std::free(dst[(size_t)t]); }
}

int main(int argc, char **argv) {
    const std::string mode = argc > 1 ? argv[1] : "p2p";
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (argc > 2) ndev = std::min(ndev, std::atoi(argv[2]));
    if (ndev < 1) { std::fprintf(stderr, "no device\n"); return 1; }
    if (mode == "p2p") {
        if (ndev < 2) { std::fprintf(stderr, "p2p needs two GPUs\n"); return 1; }
        p2p_pair(0, 1, (size_t)1 << 30, 10);
        p2p_pair(0, 1, (size_t)64 << 20, 40);
        if (ndev > 2) p2p_pair(0, ndev - 1, (size_t)1 << 30, 10);
        if (ndev > 4) p2p_pair(2, 5, (size_t)1 << 30, 10);
        return 0;
    }
    if (mode != "pcie") { std::fprintf(stderr, "usage: %s p2p|pcie [ngpu]\n", argv[0]); return 1; }

    const size_t bytes = (size_t)256 << 20;
    const int iters = 8;                                   // 2 GiB per direction per GPU per measurement
    std::vector<Gpu> g((size_t)ndev);
    for (int i = 0; i < ndev; ++i) {
        Gpu &x = g[(size_t)i];
        x.dev = i;
        CK(cudaSetDevice(i));
        CK(cudaMalloc(&x.d_in, bytes)); CK(cudaMalloc(&x.d_out, bytes));
        CK(cudaStreamCreateWithFlags(&x.s_in, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&x.s_out, cudaStreamNonBlocking));
        CK(cudaEventCreate(&x.e0)); CK(cudaEventCreate(&x.e1)); CK(cudaEventCreate(&x.f0)); CK(cudaEventCreate(&x.f1));
        for (int k = 0; k < 3; ++k) {
            x.h_in[k] = host_alloc(bytes, k);
            x.h_out[k] = host_alloc(bytes, k == WRITE_COMBINED ? PINNED : k);   // results are read by the CPU: never WC
        }
    }
    std::vector<double> per;
    std::vector<int> all;
    for (int i = 0; i < ndev; ++i) all.push_back(i);
    run_set(g, all, 3, PINNED, bytes, 2, per);             // warm-up: contexts, DMA engines, page tables

    for (int i = 0; i < ndev; ++i)
        for (int dir = 1; dir <= 3; ++dir) emit("alone", {i}, dir, PINNED, run_set(g, {i}, dir, PINNED, bytes, iters, per), per);
    for (int i = 0; i < ndev; ++i)
        for (int j = i + 1; j < ndev; ++j)
            for (int dir : {1, 3}) emit("pair", {i, j}, dir, PINNED, run_set(g, {i, j}, dir, PINNED, bytes, iters, per), per);
    if (ndev >= 4) {
        std::vector<std::vector<int>> quads = {{0, 1, 2, 3}};
        if (ndev >= 8) { quads.push_back({4, 5, 6, 7}); quads.push_back({0, 1, 4, 5}); quads.push_back({0, 2, 4, 6}); }
        for (auto &q : quads)
            for (int dir = 1; dir <= 3; ++dir) emit("quad", q, dir, PINNED, run_set(g, q, dir, PINNED, bytes, iters, per), per);
    }
    for (int kind = 0; kind < 3; ++kind) {
        bool ok = true;
        for (Gpu &x : g) ok = ok && x.h_in[kind].ok && x.h_out[kind].ok;
        if (!ok) {
            std::printf("{\"mode\":\"pcie\",\"what\":\"all\",\"host\":\"%s\",\"unavailable\":true}\n", kind_name(kind));
            continue;
        }
        for (int dir = 1; dir <= 3; ++dir) {
            emit("all", all, dir, kind, run_set(g, all, dir, kind, bytes, iters, per), per);
            if (kind != PINNED) emit("alone", {0}, dir, kind, run_set(g, {0}, dir, kind, bytes, iters, per), per);
        }
    }
    const int hw = (int)std::thread::hardware_concurrency();
    for (int t : {1, 8, 16, hw}) if (t >= 1 && t <= hw) host_ceiling(t, (size_t)128 << 20);
    return 0;
}
