"""Runs the hand-written NVLink-multicast all-reduce probe next to the NCCL probe on all visible
GPUs (or the first N given as argv[1]) and prints one line per message size."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

capi = entry.load_package().capi
n = int(sys.argv[1]) if len(sys.argv) > 1 else capi.device_count()
lo = 1 << int(os.environ.get("B200V_MIN_LOG2", "10"))
hi = 1 << int(os.environ.get("B200V_MAX_LOG2", "30"))
devs = list(range(n))
nvls = capi.nvls_allreduce_probe(devs, min_bytes=lo, max_bytes=hi, warmup=5, iters=20)
sys.stdout.flush()
if os.environ.get("B200V_SKIP_NCCL"):
    nccl = [capi.ArRow() for _ in nvls]
else:
    saved = os.dup(1)
    os.dup2(2, 1)                      # NCCL prints its banner on stdout
    try:
        nccl = capi.allreduce_probe(devs, min_bytes=lo, max_bytes=hi, warmup=5, iters=20)
    finally:
        os.dup2(saved, 1)
        os.close(saved)
for a, b in zip(nvls, nccl):
    print(json.dumps({"ranks": n, "bytes": int(a.bytes), "nvls_ms": round(a.best_ms, 5),
                      "nvls_busbw_gbs": round(a.busbw_gbs, 2), "nvls_mismatches": int(a.mismatches),
                      "nccl_ms": round(b.best_ms, 5), "nccl_busbw_gbs": round(b.busbw_gbs, 2),
                      "nccl_mismatches": int(b.mismatches), "ctas": os.environ.get("B200V_NVLS_CTAS", "sm_count")}))
