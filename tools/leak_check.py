"""Device-memory delta of 25 calls of each allocating entry point (after 3 warm calls)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import __graft_entry__ as e  # noqa: E402

capi = e.load_package().capi
n = (1 << 20) + 3
x = torch.rand(n, device="cuda")
a = np.ones(n, np.float32)
c = np.zeros(n, np.float32)
cases = {
    "stream": lambda: capi.stream(capi.OP_TRIAD, n, warmup=1, iters=2),
    "digest": lambda: capi.digest_device(x.data_ptr(), n),
    "check": lambda: capi.check_device(capi.OP_COPY, x.data_ptr(), x.data_ptr(), None, 0.0, n),
    "nccl_probe": lambda: capi.allreduce_probe([0], min_bytes=1 << 10, max_bytes=1 << 12, warmup=1, iters=2),
    "vectoradd": lambda: capi.vectoradd(a, a, c),
    "noalloc": lambda: capi.vectoradd_noalloc(a[:1 << 20], a[:1 << 20], c[:1 << 20]),
}
if capi.device_count() >= 2:
    cases["nvls_probe"] = lambda: capi.nvls_allreduce_probe([0, 1], min_bytes=1 << 10, max_bytes=1 << 12, warmup=1, iters=2)
for name, fn in cases.items():
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    f0, _ = torch.cuda.mem_get_info(0)
    for _ in range(25):
        fn()
    torch.cuda.synchronize()
    f1, _ = torch.cuda.mem_get_info(0)
    print("%-12s delta over 25 calls: %8.2f MiB" % (name, (f0 - f1) / 2**20))
