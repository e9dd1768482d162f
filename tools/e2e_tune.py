"""Transfer-pipeline tuning: e2e GB/s of b200v_vectoradd at n = 2^28 for several chunk sizes.
Each setting runs in its own process (the library reads B200V_CHUNK_LOG2 once)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r"""
import sys, time
sys.path.insert(0, %r)
import torch, __graft_entry__ as e
capi = e.load_package().capi
n = 1 << 28
hA = torch.rand(n, dtype=torch.float32).pin_memory(); hB = torch.rand(n, dtype=torch.float32).pin_memory()
hC = torch.empty(n, dtype=torch.float32, pin_memory=True)
for _ in range(2): capi.vectoradd(hA, hB, hC, n)
ts = []
for _ in range(5):
    t0 = time.perf_counter(); t = capi.vectoradd(hA, hB, hC, n); ts.append(time.perf_counter() - t0)
best = min(ts)
print("chunk_log2=%%s chunks=%%d best %%.2f ms  e2e %%.1f GB/s  h2d %%.2f ms d2h %%.2f ms kernels %%.2f ms" %% (
    sys.argv[1], t.chunks, best * 1e3, 12 * n / best / 1e9, t.h2d_ms, t.d2h_ms, t.kernel_ms))
""" % ROOT

for lg in (sys.argv[1:] or ["18", "20", "22", "24", "26", "28"]):
    env = dict(os.environ, B200V_CHUNK_LOG2=lg)
    subprocess.run([sys.executable, "-c", CHILD, lg], env=env, check=False)
