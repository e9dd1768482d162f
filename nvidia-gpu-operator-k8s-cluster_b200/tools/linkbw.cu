// linkbw.cu — link measurements that the design documents quote as numbers, not sentences:
//
//   b200v-linkbw p2p  [ngpu]   cudaMemcpyPeerAsync between GPU pairs over NVLink, one direction and
//                              both directions at once — the "link rate" the all-reduce kernels are
//                              judged against;
//   b200v-linkbw p2pk          the same pair driven by SM kernels instead of the copy engines: 128- and
//                              256-bit stores/loads on the peer mapping, and bulk (TMA) copies of
//                              shared-memory tiles — which mechanism gets how close to the link rate;
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

// ---- p2pk: SM-driven peer copies (what an all-reduce kernel's stores and loads look like on the link) ----
// push: load local, store to the peer.  pull: load from the peer, store local.  128-bit and 256-bit
// accesses, and the bulk-copy (TMA) engine moving shared-memory tiles to the peer.
template <int VEC, bool PULL>
__global__ void __launch_bounds__(512) pk_copy_kernel(float *__restrict__ dst, const float *__restrict__ src, size_t n) {
    constexpr int U = 4;
    const size_t nv = n / VEC, stride = (size_t)gridDim.x * 512;
    for (size_t v0 = (size_t)blockIdx.x * 512 + threadIdx.x; v0 < nv; v0 += stride * U) {
        float r[U][VEC];
#pragma unroll
        for (int j = 0; j < U; ++j) {
            const size_t v = v0 + (size_t)j * stride;
            if (v < nv) {
                if constexpr (VEC == 4)
                    asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r[j][0]), "=f"(r[j][1]), "=f"(r[j][2]), "=f"(r[j][3]) : "l"(src + v * 4));
                else
                    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                 : "=f"(r[j][0]), "=f"(r[j][1]), "=f"(r[j][2]), "=f"(r[j][3]), "=f"(r[j][4]), "=f"(r[j][5]), "=f"(r[j][6]), "=f"(r[j][7])
                                 : "l"(src + v * 8));
            }
        }
#pragma unroll
        for (int j = 0; j < U; ++j) {
            const size_t v = v0 + (size_t)j * stride;
            if (v < nv) {
                if constexpr (VEC == 4)
                    asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(dst + v * 4), "f"(r[j][0]), "f"(r[j][1]), "f"(r[j][2]), "f"(r[j][3]) : "memory");
                else
                    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                                 :: "l"(dst + v * 8), "f"(r[j][0]), "f"(r[j][1]), "f"(r[j][2]), "f"(r[j][3]), "f"(r[j][4]), "f"(r[j][5]), "f"(r[j][6]), "f"(r[j][7]) : "memory");
            }
        }
    }
}

// One thread per CTA drives a ring of STAGES shared-memory tiles: bulk load (local global -> shared),
// wait, bulk store (shared -> peer global).  TILE bytes per transfer.
template <int TILE, int STAGES>
__global__ void __launch_bounds__(32) pk_bulk_kernel(char *__restrict__ dst, const char *__restrict__ src, size_t bytes) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) unsigned long long bar[STAGES];
    if (threadIdx.x != 0) return;
    auto sa = [](const void *p) { return (unsigned)__cvta_generic_to_shared(p); };
    for (int i = 0; i < STAGES; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(sa(&bar[i])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const size_t tiles = bytes / TILE;
    const size_t mine = tiles > blockIdx.x ? (tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    auto load = [&](size_t k) {
        const int s = (int)(k % STAGES);
        const size_t off = (blockIdx.x + k * gridDim.x) * (size_t)TILE;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(sa(&bar[s])), "r"((unsigned)TILE) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(sa(smem + (size_t)s * TILE)), "l"(src + off), "r"((unsigned)TILE), "r"(sa(&bar[s])) : "memory");
    };
    for (size_t k = 0; k < mine && k < (size_t)(STAGES - 1); ++k) load(k);
    for (size_t k = 0; k < mine; ++k) {
        const int s = (int)(k % STAGES);
        const unsigned parity = (unsigned)((k / STAGES) & 1);
        asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n"
                     :: "r"(sa(&bar[s])), "r"(parity) : "memory");
        const size_t off = (blockIdx.x + k * gridDim.x) * (size_t)TILE;
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(dst + off), "r"(sa(smem + (size_t)s * TILE)), "r"((unsigned)TILE) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        if (k + STAGES - 1 < mine) {
            asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");     // the slot stored last round is free
            load(k + STAGES - 1);
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

static void p2pk(int a, int b, size_t bytes, int iters) {
    float *la, *ra_on_b, *lb, *rb_on_a;      // a's source, a's target on b; b's source, b's target on a
    cudaStream_t sa, sb;
    cudaEvent_t a0, a1, b0, b1;
    CK(cudaSetDevice(a)); cudaDeviceEnablePeerAccess(b, 0); (void)cudaGetLastError();
    CK(cudaMalloc(&la, bytes)); CK(cudaMalloc(&rb_on_a, bytes)); CK(cudaMemset(la, 1, bytes));
    CK(cudaStreamCreateWithFlags(&sa, cudaStreamNonBlocking)); CK(cudaEventCreate(&a0)); CK(cudaEventCreate(&a1));
    CK(cudaSetDevice(b)); cudaDeviceEnablePeerAccess(a, 0); (void)cudaGetLastError();
    CK(cudaMalloc(&lb, bytes)); CK(cudaMalloc(&ra_on_b, bytes)); CK(cudaMemset(lb, 2, bytes));
    CK(cudaStreamCreateWithFlags(&sb, cudaStreamNonBlocking)); CK(cudaEventCreate(&b0)); CK(cudaEventCreate(&b1));
    CK(cudaDeviceSynchronize());
    const size_t n = bytes / 4;
    struct Var { const char *name; int kind; int ctas; };
    const Var vars[] = {{"push_v4", 0, 64}, {"push_v4", 0, 148}, {"push_v8", 1, 64}, {"push_v8", 1, 148}, {"pull_v4", 2, 148},
                        {"pull_v8", 3, 148}, {"bulk_push_16k_s4", 4, 64}, {"bulk_push_16k_s4", 4, 148}, {"bulk_push_32k_s4", 5, 148},
                        {"bulk_push_8k_s8", 6, 148}, {"bulk_pull_16k_s4", 7, 148}};
    for (int dev : {a, b}) {
        CK(cudaSetDevice(dev));
        CK(cudaFuncSetAttribute(pk_bulk_kernel<32768, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 32768));
        CK(cudaFuncSetAttribute(pk_bulk_kernel<16384, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 16384));
        CK(cudaFuncSetAttribute(pk_bulk_kernel<8192, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 8192));
    }
    for (const Var &v : vars) {
        auto launch = [&](int dev, cudaStream_t st, float *src_local, float *dst_remote) {
            CK(cudaSetDevice(dev));
            switch (v.kind) {
            case 0: pk_copy_kernel<4, false><<<v.ctas, 512, 0, st>>>(dst_remote, src_local, n); break;
            case 1: pk_copy_kernel<8, false><<<v.ctas, 512, 0, st>>>(dst_remote, src_local, n); break;
            case 2: pk_copy_kernel<4, true><<<v.ctas, 512, 0, st>>>(src_local, dst_remote, n); break;      // read the peer, write local
            case 3: pk_copy_kernel<8, true><<<v.ctas, 512, 0, st>>>(src_local, dst_remote, n); break;
            case 4: pk_bulk_kernel<16384, 4><<<v.ctas, 32, 4 * 16384, st>>>((char *)dst_remote, (const char *)src_local, bytes); break;
            case 5: pk_bulk_kernel<32768, 4><<<v.ctas, 32, 4 * 32768, st>>>((char *)dst_remote, (const char *)src_local, bytes); break;
            case 6: pk_bulk_kernel<8192, 8><<<v.ctas, 32, 8 * 8192, st>>>((char *)dst_remote, (const char *)src_local, bytes); break;
            default: pk_bulk_kernel<16384, 4><<<v.ctas, 32, 4 * 16384, st>>>((char *)src_local, (const char *)dst_remote, bytes); break;
            }
            CK(cudaGetLastError());
        };
        float res[2][2] = {{0, 0}, {0, 0}};
        for (int bi = 0; bi < 2; ++bi) {
            for (int rep = 0; rep < 2; ++rep) {
                CK(cudaSetDevice(a)); CK(cudaEventRecord(a0, sa));
                if (bi) { CK(cudaSetDevice(b)); CK(cudaEventRecord(b0, sb)); }
                for (int i = 0; i < iters; ++i) {
                    launch(a, sa, la, ra_on_b);
                    if (bi) launch(b, sb, lb, rb_on_a);
                }
                CK(cudaSetDevice(a)); CK(cudaEventRecord(a1, sa));
                if (bi) { CK(cudaSetDevice(b)); CK(cudaEventRecord(b1, sb)); }
                CK(cudaStreamSynchronize(sa));
                if (bi) CK(cudaStreamSynchronize(sb));
            }
            CK(cudaEventElapsedTime(&res[bi][0], a0, a1));
            if (bi) CK(cudaEventElapsedTime(&res[bi][1], b0, b1));
        }
        const double gb = (double)bytes * iters / 1e9;
        std::printf("{\"mode\":\"p2pk\",\"kernel\":\"%s\",\"ctas\":%d,\"bytes\":%zu,\"uni_gbs\":%.1f,\"bi_ab_gbs\":%.1f,\"bi_ba_gbs\":%.1f}\n",
                    v.name, v.ctas, bytes, gb / (res[0][0] * 1e-3), gb / (res[1][0] * 1e-3), gb / (res[1][1] * 1e-3));
        std::fflush(stdout);
    }
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
    if (mode == "p2pk") {
        if (ndev < 2) { std::fprintf(stderr, "p2pk needs two GPUs\n"); return 1; }
        p2pk(0, 1, (size_t)512 << 20, 10);
        return 0;
    }
    if (mode != "pcie") { std::fprintf(stderr, "usage: %s p2p|p2pk|pcie [ngpu]\n", argv[0]); return 1; }

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
