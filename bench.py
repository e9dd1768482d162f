#!/usr/bin/env python
"""bench.py — the validation workload's headline number on N B200s of one node.

    python bench.py --gpus N --steps K --warmup W            (own arm, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K ...   (CPU arm: the sample's host loop)

Workload (BASELINE.json configs[1]): vectorAdd, 2^28 fp32 elements per array (1 GiB each, 3 GiB
per GPU — far beyond the 126 MB L2, so no flush is needed between steps).  A step is one pass of
the hot path over one batch: one launch of the sm_100a vectorAdd kernel through the C ABI.
With N > 1 every rank validates its own GPU on its own arrays (weak scaling, no data-path
collective — SURVEY.md §8e); the value is the bytes all ranks moved over the slowest rank's time.

  value     device-resident throughput, algorithmic bytes (12 B/element) / CUDA-event time
  e2e       same metric through b200v_vectoradd with pinned HOST buffers: H2D, kernel, D2H inside
  roofline  the vectorAdd kernel against the measured HBM copy peak (MEASURED_PEAKS.json)
  cpu_baseline  the sample's host verification loop, one thread, timed on this box (N=1, rank 0)
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "vectorAdd/triad HBM GB/s per GPU & box"
UNIT = "GB/s"
LOG2_N = 28
BYTES_PER_ELEM = 12
SEED = 0xB200
FALLBACK_HBM_GBS = 6650.0          # /opt/skills/guides/B200_PROFILING.md fallback
NOMINAL_HBM_GBS = 8000.0           # north_star's "~8 TB/s"

THROTTLE_BITS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap",
                 0x8: "hw_slowdown", 0x10: "sync_boost", 0x20: "sw_thermal_slowdown",
                 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
                 0x100: "display_clock_setting"}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--log2n", type=int, default=LOG2_N, help="elements per array = 2^log2n")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-allreduce", action="store_true")
    return ap.parse_args()


def workload_config(n, world):
    return {"workload": "cuda-vectoradd validation, vectorAdd C=A+B+0.0f, 2^%d fp32 elements per array "
                        "(%d MiB each) per GPU" % (n.bit_length() - 1, n * 4 >> 20),
            "elements_per_array": n, "arrays": 3, "bytes_per_element": BYTES_PER_ELEM,
            "parallelism": "fanout%d (independent per-GPU shards, no collective)" % world,
            "l2": "working set %d MiB per GPU > L2, no flush between steps" % (3 * n * 4 >> 20)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    except Exception:
        return FALLBACK_HBM_GBS, "B200_PROFILING.md fallback (of fallback)"


def ncu_traffic(n):
    """DRAM bytes per launch of the vectorAdd kernel from the committed ncu --set full capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        return t["vectoradd"][str(n)]["dram_bytes_per_launch"]
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the GPU is loaded."""

    def __init__(self, uuid, period=0.002):
        super().__init__(daemon=True)
        self.period = period
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self.ok = False
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if isinstance(uuid, str) else uuid)
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # NVML missing: report that instead of inventing clocks
            self.err = repr(e)

    def run(self):
        if not self.ok:
            return
        while not self._stop_evt.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in THROTTLE_BITS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [],
                    "note": "NVML unavailable: %s" % getattr(self, "err", "no samples")}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons - {"gpu_idle"}), "samples": len(self.samples)}


def cpu_verify_loop(n, threads, passes, oracle):
    """Times the sample's host verification loop over `passes` passes of n
    elements.  Returns (best GB/s, median GB/s) at 12 B/element."""
    import numpy as np
    fill_threads = os.cpu_count() or 1
    A = oracle.fill_hash(n, SEED, threads=fill_threads)
    B = oracle.fill_hash(n, SEED + oracle.SEED_Y_OFFSET, threads=fill_threads)
    C = oracle.vadd(A, B, threads=fill_threads)
    times = []
    for _ in range(passes):
        t0 = time.perf_counter()
        bad = oracle.verify_tolerance(A, B, C, threads=threads)
        times.append(time.perf_counter() - t0)
        if bad != -1:
            raise RuntimeError("CPU verification loop failed at %d" % bad)
    del A, B, C
    gbs = sorted(n * BYTES_PER_ELEM / t / 1e9 for t in times)
    return gbs[-1], gbs[len(gbs) // 2], times


def run_reference(args):
    """The reference arm: the sample's own CPU code for this path — its host verification loop —
    on all host threads.  Rank 0 only; the other ranks exit without work."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import __graft_entry__ as entry
    oracle = entry.load_oracle()
    oracle.build()
    n = 1 << args.log2n
    threads = os.cpu_count() or 1
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    # bounded: at most ~60 s of CPU work whatever K says
    per_pass_est = n * BYTES_PER_ELEM / 2e9 / max(1, threads // 2)
    cap = max(3, int(60 / max(per_pass_est, 1e-3)))
    timed = min(steps, cap)
    _best, _med, times = cpu_verify_loop(n, threads, min(warmup, 3) + timed, oracle)
    times = times[min(warmup, 3):]
    total = sum(times)
    value = n * BYTES_PER_ELEM * len(times) / total / 1e9
    sample = "%d of the %d requested steps, each the full 2^%d-element verification loop" % (
        len(times), steps, args.log2n)
    line = {"impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT,
            "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
            "ms_per_step": round(1e3 * total / len(times), 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(n, args.gpus),
            "cpu_baseline": {"value": round(value, 3), "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": sample},
            "e2e": {"value": round(value, 3), "unit": UNIT, "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference repo is a README with no code; this arm times the CPU restatement "
                    "(oracle/) of the cuda-vectoradd sample's host verification loop"}
    print(json.dumps(line))
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)

    # stdout carries exactly one JSON line.  Libraries underneath (NCCL prints a version banner when
    # a communicator is created) write to fd 1 directly, so fd 1 points at stderr for the whole run
    # and the result goes to the saved descriptor at the end.
    sys.stdout.flush()
    result_fd = os.dup(1)
    os.dup2(2, 1)

    import numpy as np  # noqa: F401
    import torch
    import torch.distributed as dist
    import __graft_entry__ as entry

    pkg = entry.load_package()
    capi = pkg.capi
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    if not torch.cuda.is_available() or capi.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local)
    dev = local
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()

    def max_over_ranks(x):
        return pkg.fanout.max_over_ranks(x, device="cuda")

    n = 1 << args.log2n
    steps = max(1, args.steps)
    warmup = max(3, args.warmup)
    seed = pkg.fanout.rank_seed(SEED, rank)

    # ---- device-resident arrays (torch owns the memory and the stream; the kernels are ours) ----
    dA = torch.empty(n, dtype=torch.float32, device="cuda")
    dB = torch.empty(n, dtype=torch.float32, device="cuda")
    dC = torch.empty(n, dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    capi.fill_device(dA.data_ptr(), n, seed, dev=dev, stream=st)
    capi.fill_device(dB.data_ptr(), n, seed + capi.SEED_Y_OFFSET, dev=dev, stream=st)
    torch.cuda.synchronize()

    def timed_loop(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1)) * 1e-3      # seconds, max over ranks

    def step_vadd():
        capi.vectoradd_device(dA.data_ptr(), dB.data_ptr(), dC.data_ptr(), n, dev=dev, stream=st)

    def step_triad():
        capi.stream_device(capi.OP_TRIAD, dC.data_ptr(), dA.data_ptr(), dB.data_ptr(), 3.0, n, dev=dev, stream=st)

    for _ in range(warmup):
        step_vadd()
    uuid = "GPU-" + str(torch.cuda.get_device_properties(local).uuid)
    sampler = ClockSampler(uuid)
    sampler.start()
    t_vadd = timed_loop(step_vadd, steps)
    # keep the same launches going briefly so short timed regions still get clock samples under load
    t_cont = time.perf_counter()
    while len(sampler.samples) < 20 and time.perf_counter() - t_cont < 0.5:
        for _ in range(20):
            step_vadd()
        torch.cuda.synchronize()
    sampler.stop()
    clocks = sampler.summary()
    clocks["window"] = "timed region + <=0.5 s of the same launches"

    # the device result is checked before it is reported (not inside the timed region)
    bad, first = capi.check_device(capi.OP_VECTORADD, dC.data_ptr(), dA.data_ptr(), dB.data_ptr(), 0.0, n, dev=dev)
    if bad:
        raise SystemExit("rank %d: vectorAdd verification failed at element %d" % (rank, first))

    for _ in range(warmup):
        step_triad()
    t_triad = timed_loop(step_triad, steps)

    bytes_step = n * BYTES_PER_ELEM
    value = world * steps * bytes_step / t_vadd / 1e9
    triad = world * steps * bytes_step / t_triad / 1e9
    per_launch_s = t_vadd / steps
    achieved = bytes_step / per_launch_s / 1e9
    peak, peak_src = measured_peak()
    del dC

    # ---- end to end through the host-buffer entry point ----------------------------------------
    hA = torch.empty(n, dtype=torch.float32, pin_memory=True)
    hB = torch.empty(n, dtype=torch.float32, pin_memory=True)
    hC = torch.empty(n, dtype=torch.float32, pin_memory=True)
    hA.copy_(dA)
    hB.copy_(dB)
    torch.cuda.synchronize()
    del dA, dB
    e2e_steps = min(steps, 10)
    e2e_launches = 0
    for _ in range(2):
        capi.vectoradd(hA, hB, hC, n, dev=dev)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        tm = capi.vectoradd(hA, hB, hC, n, dev=dev)
        e2e_launches += tm.kernel_launches
    t_e2e = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e_value = world * e2e_steps * bytes_step / t_e2e / 1e9
    # spot-check the host result: first and last 64 Ki elements (IEEE add is what the kernel does)
    k = 1 << 16
    for sl in (slice(0, k), slice(n - k, n)):
        if not torch.equal((hA[sl] + hB[sl]) + 0.0, hC[sl]):
            raise SystemExit("rank %d: e2e result mismatch" % rank)

    # the same call with ordinary pageable buffers (what malloc, a Go slice or numpy hands over):
    # the library stages them through pinned memory with a host copy pool
    e2e_pageable = None
    if world == 1:
        import numpy as np
        pA, pB = hA.numpy().copy(), hB.numpy().copy()
        pC = np.empty(n, np.float32)
        capi.vectoradd(pA, pB, pC, n, dev=dev)
        t0 = time.perf_counter()
        for _ in range(3):
            capi.vectoradd(pA, pB, pC, n, dev=dev)
        e2e_pageable = {"value": round(3 * bytes_step / (time.perf_counter() - t0) / 1e9, 2), "unit": UNIT,
                        "steps": 3, "note": "pageable host buffers, staged by the library"}
        if pC[:k].tobytes() != hC[:k].numpy().tobytes():
            raise SystemExit("pageable e2e result mismatch")
        del pA, pB, pC

    # ---- NVLink probe (only place NCCL is used by the product) -----------------------------------
    allreduce = None
    if world > 1 and not args.no_allreduce:
        uid = torch.zeros(capi.NCCL_UNIQUE_ID_BYTES, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(capi.nccl_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, src=0)
        # NCCL announces its version on stdout when a communicator is created; stdout must carry
        # the one JSON line only, so the probe runs with fd 1 pointed at stderr.
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            rows = capi.allreduce_probe_rank(dev, rank, world, bytes(uid.cpu().numpy().tobytes()),
                                             min_bytes=1 << 20, max_bytes=1 << 30, warmup=5, iters=20)
        finally:
            os.dup2(saved_fd, 1)
            os.close(saved_fd)
        ms = torch.tensor([r.best_ms for r in rows], dtype=torch.float64, device="cuda")
        badrows = torch.tensor([float(r.mismatches) for r in rows], dtype=torch.float64, device="cuda")
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(badrows, op=dist.ReduceOp.SUM)
        table = []
        for r, t_ms, b in zip(rows, ms.tolist(), badrows.tolist()):
            alg = r.bytes / (t_ms * 1e-3) / 1e9
            table.append({"bytes": int(r.bytes), "ms": round(t_ms, 4), "algbw_gbs": round(alg, 1),
                          "busbw_gbs": round(alg * 2 * (world - 1) / world, 1), "mismatches": int(b)})
        top = table[-1]
        allreduce = {"op": "ncclAllReduce fp32 sum in place", "nccl": capi.nccl_version(),
                     "busbw_gbs_at_1GiB": top["busbw_gbs"], "frac_of_900": round(top["busbw_gbs"] / 900.0, 4),
                     "bit_exact": all(t["mismatches"] == 0 for t in table), "rows": table}

    # ---- hand-written NVLink-multicast all-reduce, one process per GPU like everything else here --
    # Every rank drives its own GPU; rank 0 hands the multicast object to the others over a Unix
    # socket (csrc/rendezvous.h).  Times are per rank on the device; the job's time is the max.
    nvls = None
    if world > 1 and not args.no_allreduce:
        torch.cuda.synchronize()
        name = "b200v-nvls-%s-%d" % (os.environ.get("MASTER_PORT", "0"), os.getppid())
        try:
            vrows = capi.nvls_allreduce_probe_rank(dev, rank, world, name, min_bytes=1 << 20, max_bytes=1 << 30,
                                                   warmup=5, iters=20)
            fail = 0.0
        except capi.B200VError as e:
            vrows, fail, why = [], 1.0, str(e)
        failed = torch.tensor([fail], dtype=torch.float64, device="cuda")
        dist.all_reduce(failed, op=dist.ReduceOp.MAX)
        if failed.item() > 0:
            nvls = {"unavailable": why if fail else "another rank failed"}
        else:
            vms = torch.tensor([r.best_ms for r in vrows], dtype=torch.float64, device="cuda")
            vbad = torch.tensor([float(r.mismatches) for r in vrows], dtype=torch.float64, device="cuda")
            dist.all_reduce(vms, op=dist.ReduceOp.MAX)
            dist.all_reduce(vbad, op=dist.ReduceOp.SUM)
            vt = []
            for r, t_ms, b in zip(vrows, vms.tolist(), vbad.tolist()):
                alg = r.bytes / (t_ms * 1e-3) / 1e9
                vt.append({"bytes": int(r.bytes), "ms": round(t_ms, 4), "algbw_gbs": round(alg, 1),
                           "busbw_gbs": round(alg * 2 * (world - 1) / world, 1), "mismatches": int(b)})
            nvls = {"op": ("peer load/store kernel" if world == 2 else "multimem.ld_reduce + multimem.st kernel")
                    + ", fp32 sum in place, one process per GPU",
                    "busbw_gbs_at_1GiB": vt[-1]["busbw_gbs"], "frac_of_900": round(vt[-1]["busbw_gbs"] / 900.0, 4),
                    "bit_exact": all(t["mismatches"] == 0 for t in vt), "gpu_launches": world * len(vt) * 26,
                    "rows": vt}

    # ---- CPU baseline: the sample's host verification loop, one thread, rank 0 at N=1 -------------
    cpu = None
    if world == 1 and rank == 0 and not args.no_cpu_baseline:
        oracle = entry.load_oracle()
        oracle.build()
        del hA, hB, hC
        passes = 40           # ~10 s of single-thread CPU work at ~0.25 s per pass
        best, med, times = cpu_verify_loop(n, 1, passes, oracle)
        cpu = {"value": round(med, 3), "best": round(best, 3), "unit": UNIT, "cores": 1, "kind": "port",
               "host_cores_available": os.cpu_count(),
               "sample": "%d passes of the sample's host verification loop over the full 2^%d-element "
                         "arrays (%.1f s of CPU work), single thread as in the sample" % (
                             passes, args.log2n, sum(times))}

    if rank == 0:
        line = {"metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world,
                "steps": steps, "warmup": warmup, "ms_per_step": round(1e3 * t_vadd / steps, 5),
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": workload_config(n, world),
                "per_gpu_gbs": round(value / world, 1), "triad_gbs": round(triad, 1),
                "frac_of_nominal_8000": round(value / world / NOMINAL_HBM_GBS, 4),
                "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                             "frac": round(achieved / peak, 4), "traffic": ncu_traffic(n),
                             "peak_source": peak_src, "kernel": "b200v::ew_vec_kernel<VECTORADD>",
                             "algorithmic_bytes_per_launch": bytes_step},
                "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "h2d_bytes_per_step": 8 * n,
                        "d2h_bytes_per_step": 4 * n, "steps": e2e_steps,
                        "api": "b200v_vectoradd (pinned host buffers, chunked 3-stream pipeline)"},
                "gpu_launches": steps * world, "e2e_gpu_launches": e2e_launches * world,
                "clocks": clocks}
        if e2e_pageable:
            line["e2e_pageable"] = e2e_pageable
        if cpu:
            line["cpu_baseline"] = cpu
        if allreduce:
            line["allreduce"] = allreduce
        if nvls:
            line["nvls_allreduce"] = nvls
        sys.stdout.flush()
        os.write(result_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
