"""e2e GB/s of b200v_vectoradd at n = 2^28 with pageable (malloc'd) vs pinned host buffers."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import __graft_entry__ as e  # noqa: E402

capi = e.load_package().capi
n = 1 << int(os.environ.get("LOG2N", "28"))
rng = np.random.default_rng(0)
A = rng.random(n, dtype=np.float32)
B = rng.random(n, dtype=np.float32)
C = np.zeros(n, np.float32)
for kind in ("pageable", "pinned"):
    if kind == "pinned":
        pa, pb, pc = capi.PinnedArray(n), capi.PinnedArray(n), capi.PinnedArray(n)
        pa.array[:] = A
        pb.array[:] = B
        a, b, c = pa.array, pb.array, pc.array
    else:
        a, b, c = A, B, C
    capi.vectoradd(a, b, c)
    ts = []
    for _ in range(4):
        t0 = time.perf_counter()
        t = capi.vectoradd(a, b, c)
        ts.append(time.perf_counter() - t0)
    ok = bool((c[:1000] == (A[:1000] + B[:1000])).all() and (c[-1000:] == (A[-1000:] + B[-1000:])).all())
    print("%-9s best %.1f ms  e2e %.1f GB/s  (h2d %.1f ms, d2h %.1f ms, chunks %d) ok=%s" % (
        kind, min(ts) * 1e3, 12 * n / min(ts) / 1e9, t.h2d_ms, t.d2h_ms, t.chunks, ok))
