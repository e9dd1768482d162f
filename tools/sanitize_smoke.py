"""Small end-to-end exercise of every kernel family for compute-sanitizer (memcheck / racecheck):
vector, tile, CLC, TMA-bulk and scalar elementwise kernels at ragged sizes, generator, digest,
check, the no-allocation path.  Sizes are tiny so the tool's replay stays cheap."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r"""
import sys
sys.path.insert(0, %r)
import numpy as np, __graft_entry__ as e
capi = e.load_package().capi
rng = np.random.default_rng(1)
for n in (1, 7, 2049, 8193, 70001):
    x = rng.random(n, dtype=np.float32); y = rng.random(n, dtype=np.float32)
    for op in range(6):
        d = np.zeros(n, np.float32)
        capi.stream_host(op, d, x, y, 3.0)
    d = np.zeros(n, np.float32)
    capi.vectoradd_noalloc(x, y, d)
    assert (d == x + y).all()
r = capi.stream(capi.OP_TRIAD, 70001, warmup=1, iters=2)
assert r.mismatches == 0
print("SANITIZE_CHILD_OK", sys.argv[1])
""" % ROOT

rc = 0
for variant in ("t8u1b1024", "t4u2b1024", "v8u1b256-oneshot", "clc8u1b1024", "v8u2b256-persist8", "bulk2048s3b128"):
    env = dict(os.environ, B200V_EW_VARIANT=variant)
    p = subprocess.run([sys.executable, "-c", CHILD, variant], env=env)
    rc |= p.returncode
sys.exit(rc)
