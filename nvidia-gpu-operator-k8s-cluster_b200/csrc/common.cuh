// common.cuh — per-device state and launch helpers shared by the C-ABI translation units.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>
#include <vector>

#include "../../include/b200validate.h"

namespace b200v {

// Returns from the enclosing function with the cudaError_t as a positive ABI code.
#define B200V_CUDA(expr)                                   \
    do {                                                   \
        cudaError_t e__ = (expr);                          \
        if (e__ != cudaSuccess) {                          \
            (void)cudaGetLastError();                      \
            return (int)e__;                               \
        }                                                  \
    } while (0)

// Host-path pipeline geometry: the transfers are cut into chunks that ride a ring of device
// slots, so device memory use is bounded whatever n is (3 arrays x SLOTS x CHUNK_ELEMS x 4 B).
// 64 MiB chunks measured best on PCIe Gen5 x16 (tools/e2e_tune.py: 2^22 -> 44.6 ms, 2^24 -> 42.8 ms,
// 2^26 -> 45.3 ms for a 2^28-element call); the ring then holds 3 x 3 x 64 MiB = 576 MiB.
constexpr size_t kChunkElems = (size_t)1 << 24;
constexpr size_t kMinChunkElems = (size_t)1 << 20;   // tail of the taper (4 MiB)
// Pageable callers are staged through pinned memory by host threads; smaller chunks keep the
// pinned footprint (3 x 3 x 16 MiB) and the first-chunk latency down.
constexpr size_t kStagedChunkElems = (size_t)1 << 22;
constexpr int kSlots = 3;

struct DeviceCtx {
    int dev = -1;
    bool ready = false;
    int sm_count = 0;
    size_t l2_bytes = 0;
    cudaStream_t s_in = nullptr, s_k = nullptr, s_out = nullptr;
    // ring slots for the host path (allocated on first use, sized to min(n, kChunkElems))
    float *slot_x[kSlots] = {}, *slot_y[kSlots] = {}, *slot_d[kSlots] = {};
    size_t slot_elems = 0;
    // pinned host staging for callers whose buffers are pageable (allocated on first such call)
    float *stage_x[kSlots] = {}, *stage_y[kSlots] = {}, *stage_d[kSlots] = {};
    size_t stage_elems = 0;
    cudaEvent_t in_done[kSlots] = {}, k_done[kSlots] = {}, out_done[kSlots] = {};
    cudaEvent_t t_begin = nullptr, t_end = nullptr, h2d_first = nullptr, h2d_last = nullptr,
                d2h_first = nullptr, d2h_last = nullptr;
    std::vector<cudaEvent_t> k_ev;   // per-chunk kernel start/stop pairs
    void *owner = nullptr;           // the lock-holding slot this context lives in
};

// Selects `dev`, creates its context on first use, and returns it locked for the caller's use.
// The per-device lock serialises host-path calls on one device; different devices run in parallel.
int acquire_ctx(int dev, DeviceCtx **out);
void release_ctx(DeviceCtx *ctx);
int destroy_all_ctx();

// Enqueue op on `stream`; picks the vector, bulk or scalar kernel from alignment and aliasing.
// grid/block (optional) receive the launch shape.
int launch_ew(const DeviceCtx &ctx, int op, float *dst, const float *x, const float *y, float s,
              size_t n, cudaStream_t stream, int *grid, int *block);
// Name of the kernel launch_ew would run for these pointers and this n (same decision code).
int describe_ew(int op, const float *dst, const float *x, const float *y, size_t n, char *buf, size_t len);
int launch_sampleshape(int op, float *dst, const float *x, const float *y, float s, size_t n,
                       cudaStream_t stream);

}  // namespace b200v
