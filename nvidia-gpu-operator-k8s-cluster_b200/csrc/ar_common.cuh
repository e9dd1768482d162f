// ar_common.cuh — the input pattern and the bitwise check shared by the two all-reduce probes
// (NCCL in nccl_probe.cu, the hand-written NVLink-multicast kernel in nvls_allreduce.cu).
#pragma once

#include <cstddef>
#include <cuda_runtime.h>

namespace b200v {

// Rank r contributes the integer-valued (float)((i*(2r+1) + r) & 15).  (i+r)&7 would do for 2 and
// 4 ranks but sums to the constant 28 over 8 ranks; the odd multiplier keeps the sum a function of
// i at every rank count.  Sums stay far below 2^24, so they are exact in any reduction order and
// the reduced buffer can be compared bit for bit with the closed form.
__host__ __device__ inline float ar_value(size_t i, int r) {
    return (float)((i * (size_t)(2 * r + 1) + (size_t)r) & 15);
}

static __global__ void __launch_bounds__(256) ar_fill_kernel(float *buf, size_t n, int rank) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockDim.x * blockIdx.x + threadIdx.x; i < n; i += stride)
        buf[i] = ar_value(i, rank);
}

// Counts elements whose bits differ from sum_{r<nranks} ar_value(i, r).  `bad` is an 8-byte
// device counter the caller zeroes first.
static __global__ void __launch_bounds__(256)
ar_check_kernel(const float *buf, size_t n, int nranks, unsigned long long *bad) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    unsigned long long local = 0;
    for (size_t i = (size_t)blockDim.x * blockIdx.x + threadIdx.x; i < n; i += stride) {
        float want = 0.0f;
        for (int r = 0; r < nranks; ++r) want += ar_value(i, r);
        local += __float_as_uint(want) != __float_as_uint(buf[i]);
    }
    if (local) atomicAdd(bad, local);
}

}  // namespace b200v
