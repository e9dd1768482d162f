"""Experiment: end-to-end vectorAdd with the kernel reading and writing PINNED HOST memory directly
over PCIe (zero-copy), against the chunked DMA pipeline of b200v_vectoradd.  n = 2^28."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import __graft_entry__ as e  # noqa: E402

capi = e.load_package().capi
n = 1 << int(os.environ.get("LOG2N", "28"))
hA = torch.rand(n, dtype=torch.float32).pin_memory()
hB = torch.rand(n, dtype=torch.float32).pin_memory()
hC = torch.zeros(n, dtype=torch.float32).pin_memory()
dA = torch.empty(n, dtype=torch.float32, device="cuda")
dB = torch.empty(n, dtype=torch.float32, device="cuda")
dC = torch.empty(n, dtype=torch.float32, device="cuda")
st = torch.cuda.current_stream().cuda_stream


def timed(fn, reps=4):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    return min(ts)


def check():
    k = 1 << 16
    return bool(torch.equal((hA[:k] + hB[:k]) + 0.0, hC[:k]) and torch.equal((hA[-k:] + hB[-k:]) + 0.0, hC[-k:]))


rows = []
t = timed(lambda: capi.vectoradd(hA, hB, hC, n))
rows.append(("dma pipeline (b200v_vectoradd)", t, check()))
hC.zero_()
t = timed(lambda: capi.vectoradd_device(hA.data_ptr(), hB.data_ptr(), hC.data_ptr(), n, stream=st))
rows.append(("zero-copy in+out", t, check()))
hC.zero_()


def in_dma_out_zero():
    dA.copy_(hA, non_blocking=True)
    dB.copy_(hB, non_blocking=True)
    capi.vectoradd_device(dA.data_ptr(), dB.data_ptr(), hC.data_ptr(), n, stream=st)


t = timed(in_dma_out_zero)
rows.append(("in: one DMA each, out: zero-copy (serial)", t, check()))
hC.zero_()


def in_zero_out_dma():
    capi.vectoradd_device(hA.data_ptr(), hB.data_ptr(), dC.data_ptr(), n, stream=st)
    hC.copy_(dC, non_blocking=True)


t = timed(in_zero_out_dma)
rows.append(("in: zero-copy, out: one DMA (serial)", t, check()))
t = timed(lambda: capi.vectoradd_device(hA.data_ptr(), hB.data_ptr(), dC.data_ptr(), n, stream=st))
rows.append(("kernel reading host only (2 GiB over PCIe)", t, True))
t = timed(lambda: capi.vectoradd_device(dA.data_ptr(), dB.data_ptr(), hC.data_ptr(), n, stream=st))
rows.append(("kernel writing host only (1 GiB over PCIe)", t, True))
for name, t, ok in rows:
    print("%-46s %7.2f ms  %6.1f GB/s algorithmic  ok=%s  variant=%s" % (
        name, t * 1e3, 12 * n / t / 1e9, ok, os.environ.get("B200V_EW_VARIANT", "default")))
