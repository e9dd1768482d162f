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
  roofline  the vectorAdd kernel against the measured HBM copy peak (MEASURED_PEAKS.json), with the
            same launch held for >= 2 s (`sustained`) and the kernel named by the library's dispatch
  stream_sweep  (N=1) copy/scale/add/triad at 2^18 ... 2^32 elements, digests checked against the oracle
  cpu_baseline  the sample's host verification loop on every host thread (N=1, rank 0); the
            sample's own single-threaded form is reported beside it
  allreduce / nvls_allreduce  (N>1) the NVLink probes, 1 KiB - 1 GiB, and `allreduce_parity`: each
            rank's reduced buffer read back and compared with the oracle's closed form
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
    ap.add_argument("--no-stream-sweep", action="store_true")
    ap.add_argument("--e2e-only", action="store_true", help="diagnostic: time only the host-buffer path and print "
                    "a short line (not the bench contract)")
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


SWEEP_LOG2 = (18, 22, 24, 26, 28, 30, 32)
SWEEP_LIVE_ORACLE_MAX_LOG2 = 28      # larger sizes are compared with tests/golden/sweep_digests.json


def oracle_sweep_digests(oracle, log2_n, ops, scalar):
    """Digest of every op's output at 2^log2_n elements from the CPU oracle, in 2^24-element pieces
    on all host threads (the digest is a sum over elements, so pieces add up)."""
    import concurrent.futures
    n, piece = 1 << log2_n, 1 << 24

    def one(lo):
        m = min(piece, n - lo)
        x = oracle.fill_hash(m, SEED + lo)
        y = oracle.fill_hash(m, SEED + oracle.SEED_Y_OFFSET + lo)
        return [oracle.digest_from(oracle.stream(op, x, y, scalar), lo) for op in ops]
    with concurrent.futures.ThreadPoolExecutor(max_workers=os.cpu_count() or 1) as pool:
        parts = list(pool.map(one, range(0, n, piece)))
    return [sum(p[k] for p in parts) % 2**64 for k in range(len(ops))]


def run_stream_sweep(capi, oracle, dev, free_bytes):
    """BASELINE.json configs[2] in the driver-run line: median GB/s of the four STREAM ops through
    b200v_stream (allocates, fills, times, re-checks and digests on the device), each digest
    compared with the CPU oracle (live up to 2^28 elements, the committed piecewise golden above)."""
    ops = [(capi.OP_COPY, "copy"), (capi.OP_SCALE, "scale"), (capi.OP_ADD, "add"), (capi.OP_TRIAD, "triad")]
    try:
        with open(os.path.join(ROOT, "tests", "golden", "sweep_digests.json")) as f:
            golden = json.load(f)["log2_n"]
    except Exception:
        golden = {}
    info = capi.device_info(dev)
    nominal = 2.0 * info.mem_clock_khz * 1e3 * info.mem_bus_bits / 8.0 / 1e9
    rows, launches, bad = [], 0, 0
    for lg in SWEEP_LOG2:
        n = 1 << lg
        if 3 * 4 * n + (2 << 30) > free_bytes:
            rows.append({"log2_n": lg, "skipped": "needs %d GiB of device memory" % (3 * 4 * n >> 30)})
            continue
        live = oracle_sweep_digests(oracle, lg, [o for o, _ in ops], 3.0) if lg <= SWEEP_LIVE_ORACLE_MAX_LOG2 else None
        for k, (op, name) in enumerate(ops):
            r = capi.stream(op, n, s=3.0, seed=SEED, warmup=3, iters=10, dev=dev)
            launches += r.kernel_launches
            want = live[k] if live is not None else (int(golden[str(lg)][name], 16) if str(lg) in golden else None)
            ok = (want is not None and r.digest == want and r.mismatches == 0)
            bad += 0 if ok else 1
            row = {"op": name, "log2_n": lg, "median_gbs": round(r.median_gbs, 1), "best_gbs": round(r.best_gbs, 1),
                   "l2_resident": bool(r.l2_resident), "digest": "%016x" % r.digest, "digest_ok": ok,
                   "checked_vs": "oracle (live)" if live is not None else ("oracle (committed golden)" if want is not None else None)}
            if not r.l2_resident:
                row["frac_of_8000"] = round(r.median_gbs / NOMINAL_HBM_GBS, 4)
                row["frac_of_device_nominal"] = round(r.median_gbs / nominal, 4) if nominal > 0 else None
            rows.append(row)
    return {"api": "b200v_stream", "scalar": 3.0, "seed": SEED, "device_nominal_gbs": round(nominal, 1),
            "mem_bus_bits": info.mem_bus_bits, "mem_clock_khz": info.mem_clock_khz,
            "digest_failures": bad, "gpu_launches": launches, "rows": rows}


def ar_probe_launches(rows, warmup, iters):
    """Kernel launches of one rank in an all-reduce probe sweep, by the rule the library uses: per
    size two fill/reduce/check passes plus warm-up plus iters batches of `reps` launches."""
    total = 0
    for r in rows:
        reps = int(min(16.0, max(1.0, 100.0 / (r.bytes / 300e3 + 8.0))))
        total += 2 * 3 + warmup + iters * reps
    return total


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the GPU is loaded."""

    def __init__(self, uuid, period=0.002):
        super().__init__(daemon=True)
        self.period = period
        self.samples, self.reasons = [], set()
        self.power_w = 0.0
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
                self.power_w = max(self.power_w, self.nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
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
                "reasons": sorted(self.reasons - {"gpu_idle"}), "samples": len(self.samples),
                "power_w_max": round(self.power_w, 1)}


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

    def sum_over_ranks(x):
        return pkg.fanout.sum_over_ranks(x, device="cuda")

    n = 1 << args.log2n
    steps = max(1, args.steps)
    warmup = max(3, args.warmup)
    seed = pkg.fanout.rank_seed(SEED, rank)
    oracle = entry.load_oracle()          # the checker: digests, closed forms, the CPU baseline
    if local == 0:
        oracle.build()
    barrier()

    # ---- device-resident arrays (torch owns the memory and the stream; the kernels are ours) ----
    dA = torch.empty(n, dtype=torch.float32, device="cuda")
    dB = torch.empty(n, dtype=torch.float32, device="cuda")
    dC = torch.empty(n, dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    capi.fill_device(dA.data_ptr(), n, seed, dev=dev, stream=st)
    capi.fill_device(dB.data_ptr(), n, seed + capi.SEED_Y_OFFSET, dev=dev, stream=st)
    torch.cuda.synchronize()
    kernel = capi.kernel_name(capi.OP_VECTORADD, dC.data_ptr(), dA.data_ptr(), dB.data_ptr(), n)

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

    if args.e2e_only:
        hb = [capi.PinnedArray(n, write_only=True), capi.PinnedArray(n, write_only=True), capi.PinnedArray(n)]
        for h, d in ((hb[0], dA), (hb[1], dB)):
            tmp = d.cpu().numpy()
            h.array[:] = tmp                                   # CPU writes only: fine for write-combined
            del tmp
        for _ in range(2):
            capi.vectoradd(hb[0].array, hb[1].array, hb[2].array, n, dev=dev)
        barrier()
        t0 = time.perf_counter()
        reps = 6
        for _ in range(reps):
            capi.vectoradd(hb[0].array, hb[1].array, hb[2].array, n, dev=dev)
        t_local = time.perf_counter() - t0
        agg = pkg.fanout.job_throughput_gbs(reps * n * BYTES_PER_ELEM, t_local, device="cuda")
        barrier()
        ok = bool((hb[2].array[:4096] == (dA[:4096] + dB[:4096] + 0.0).cpu().numpy()).all())
        if rank == 0:
            os.write(result_fd, (json.dumps({"e2e_only": True, "n_gpus": world, "host_policy": os.environ.get("B200V_HOST_POLICY", "pinned"),
                                             "e2e_gbs": round(agg, 2), "per_gpu_gbs": round(agg / world, 2),
                                             "rank0_s_per_call": round(t_local / reps, 4), "ok": ok}) + "\n").encode())
        if world > 1:
            dist.destroy_process_group()
        return 0

    for _ in range(warmup):
        step_vadd()
    uuid = "GPU-" + str(torch.cuda.get_device_properties(local).uuid)
    sampler = ClockSampler(uuid)
    sampler.start()
    t_vadd = timed_loop(step_vadd, steps)
    bytes_step = n * BYTES_PER_ELEM

    # The same launch held for >= 2 s with the clock sampler still on: what the kernel sustains once
    # power and temperature have had time to react (the K-step region above lasts milliseconds).
    sustain_target_s = 2.25      # margin so the region is >= 2 s even if launches speed up a little
    k_sustain = max(steps, int(sustain_target_s / max(t_vadd / steps, 1e-6)) + 1)
    t_sustain = timed_loop(step_vadd, k_sustain)
    sampler.stop()
    clocks = sampler.summary()
    clocks["window"] = "timed region + the %.1f s sustained run of the same launch" % t_sustain
    sustained = {"seconds": round(t_sustain, 3), "launches": k_sustain,
                 "gbs_per_gpu": round(k_sustain * bytes_step / t_sustain / 1e9, 1),
                 "ms_per_launch": round(1e3 * t_sustain / k_sustain, 5)}

    # the device result is checked before it is reported (not inside the timed region)
    bad, first = capi.check_device(capi.OP_VECTORADD, dC.data_ptr(), dA.data_ptr(), dB.data_ptr(), 0.0, n, dev=dev)
    if bad:
        raise SystemExit("rank %d: vectorAdd verification failed at element %d" % (rank, first))

    for _ in range(warmup):
        step_triad()
    t_triad = timed_loop(step_triad, steps)

    value = pkg.fanout.aggregate_gbs([steps * bytes_step] * world, t_vadd)
    triad = pkg.fanout.aggregate_gbs([steps * bytes_step] * world, t_triad)
    per_launch_s = t_vadd / steps
    achieved = bytes_step / per_launch_s / 1e9
    peak, peak_src = measured_peak()
    del dC

    # ---- end to end through the host-buffer entry point ----------------------------------------
    # b200v_host_alloc_ex: the library's host-memory policy; the inputs are only ever written by the CPU
    hbuf = [capi.PinnedArray(n, write_only=True), capi.PinnedArray(n, write_only=True), capi.PinnedArray(n)]
    hA, hB, hC = (torch.from_numpy(h.array) for h in hbuf)
    hA.copy_(dA)
    hB.copy_(dB)
    torch.cuda.synchronize()
    # what the checks below need of the inputs, taken from the device copies: the CPU never READS
    # hA / hB (they may be write-combined memory, which is very slow to read back)
    k = 1 << 16
    expect = [((dA[sl] + dB[sl]) + 0.0).cpu() for sl in (slice(0, k), slice(n - k, n))]
    pA = dA.cpu().numpy() if world == 1 else None
    pB = dB.cpu().numpy() if world == 1 else None
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
    t_local = time.perf_counter() - t0
    e2e_value = pkg.fanout.job_throughput_gbs(e2e_steps * bytes_step, t_local, device="cuda")
    barrier()
    # spot-check the host result: first and last 64 Ki elements (IEEE add is what the kernel does);
    # the full-size byte-for-byte compare of this call is tests/test_gpu_parity.py::test_cfg1_host_path_full_size
    for sl, want in zip((slice(0, k), slice(n - k, n)), expect):
        if not torch.equal(want, hC[sl]):
            raise SystemExit("rank %d: e2e result mismatch" % rank)

    # the same call with ordinary pageable buffers (what malloc, a Go slice or numpy hands over):
    # the library stages them through pinned memory with a host copy pool
    e2e_pageable = None
    if world == 1:
        import numpy as np
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
    del hA, hB, hC
    for h in hbuf:
        h.close()

    # ---- STREAM sweep (configs[2]) in the driver-run line, N=1 only ---------------------------------
    stream_sweep = None
    if world == 1 and not args.no_stream_sweep:
        torch.cuda.empty_cache()
        free_bytes, _total = torch.cuda.mem_get_info(dev)
        stream_sweep = run_stream_sweep(capi, oracle, dev, free_bytes)

    # ---- NVLink probes (only place NCCL is used by the product) ------------------------------------
    # Parity first: each rank fills a buffer IT owns, the library reduces it in place, the rank reads
    # it back and compares it byte for byte with the oracle's closed form.  Run before the timed
    # sweep and again after it (three all-reduces back to back, nothing in between).
    import numpy as np
    PARITY_FLOATS = (256, 100003, (1 << 22) + 1, 1 << 26)      # 1 KiB, two ragged sizes, 256 MiB

    def parity_pass(comm, repeats):
        mism = 0
        for m in PARITY_FLOATS:
            buf = torch.from_numpy(oracle.allreduce_input(m, rank)).cuda()
            for _ in range(repeats):
                comm.allreduce(buf.data_ptr(), m, stream=st)
            torch.cuda.synchronize()
            if hasattr(comm, "check"):
                comm.check()
            want = oracle.allreduce_expected(m, world) * np.float32(world ** (repeats - 1))
            got = buf.cpu().numpy()
            mism += int(np.count_nonzero(got.view(np.uint32) != want.view(np.uint32)))
            del buf
        return int(sum_over_ranks(mism))

    def table_of(rows):
        ms = torch.tensor([r.best_ms for r in rows], dtype=torch.float64, device="cuda")
        badrows = torch.tensor([float(r.mismatches) for r in rows], dtype=torch.float64, device="cuda")
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(badrows, op=dist.ReduceOp.SUM)
        table = []
        for r, t_ms, b in zip(rows, ms.tolist(), badrows.tolist()):
            alg = r.bytes / (t_ms * 1e-3) / 1e9
            row = {"bytes": int(r.bytes), "us": round(t_ms * 1e3, 2), "algbw_gbs": round(alg, 2),
                   "busbw_gbs": round(alg * 2 * (world - 1) / world, 2), "mismatches": int(b)}
            if r.algo:
                row["kernel"] = capi.AR_ALGO_NAMES.get(r.algo, str(r.algo))
            table.append(row)
        return table

    allreduce = None
    if world > 1 and not args.no_allreduce:
        uid = torch.zeros(2 * capi.NCCL_UNIQUE_ID_BYTES, dtype=torch.uint8, device="cuda")
        if rank == 0:
            ids = capi.nccl_unique_id() + capi.nccl_unique_id()
            uid.copy_(torch.frombuffer(bytearray(ids), dtype=torch.uint8))
        dist.broadcast(uid, src=0)
        ids = bytes(uid.cpu().numpy().tobytes())
        comm = capi.ArComm(dev, rank, world, ids[:capi.NCCL_UNIQUE_ID_BYTES])
        parity_before = parity_pass(comm, 1)
        rows = capi.allreduce_probe_rank(dev, rank, world, ids[capi.NCCL_UNIQUE_ID_BYTES:],
                                         min_bytes=1 << 10, max_bytes=1 << 30, warmup=5, iters=20)
        parity_after = parity_pass(comm, 3)
        comm.close()
        table = table_of(rows)
        top = table[-1]
        allreduce = {"op": "ncclAllReduce fp32 sum in place", "nccl": capi.nccl_version(),
                     "busbw_gbs_at_1GiB": top["busbw_gbs"], "frac_of_900": round(top["busbw_gbs"] / 900.0, 4),
                     "bit_exact": all(t["mismatches"] == 0 for t in table),
                     "allreduce_parity": {"checked_vs": "oracle", "api": "b200v_ar_comm_allreduce on caller-owned buffers",
                                          "sizes_bytes": [4 * m for m in PARITY_FLOATS],
                                          "mismatches": parity_before + parity_after,
                                          "mismatches_before_timed": parity_before,
                                          "mismatches_after_timed_3_back_to_back": parity_after},
                     "rows": table}

    # ---- hand-written NVLink all-reduce, one process per GPU like everything else here -------------
    # Every rank drives its own GPU; rank 0 hands the multicast object to the others over a Unix
    # socket (csrc/rendezvous.h).  Times are per rank on the device; the job's time is the max.
    nvls = None
    if world > 1 and not args.no_allreduce:
        torch.cuda.synchronize()
        name = "b200v-nvls-%s-%d" % (os.environ.get("MASTER_PORT", "0"), os.getppid())
        why, vrows, pb, pa = "", [], 0, 0
        try:
            vcomm = capi.NvlsComm(dev, rank, world, name + "-comm", 4 * max(PARITY_FLOATS) + 64)
            pb = parity_pass(vcomm, 1)
            vrows = capi.nvls_allreduce_probe_rank(dev, rank, world, name, min_bytes=1 << 10, max_bytes=1 << 30,
                                                   warmup=5, iters=20)
            pa = parity_pass(vcomm, 3)
            vcomm.close()
            fail = 0.0
        except capi.B200VError as e:
            fail, why = 1.0, str(e)
        if sum_over_ranks(fail) > 0:
            nvls = {"unavailable": why if fail else "another rank failed"}
        else:
            vt = table_of(vrows)
            nvls = {"op": "hand-written sm_100a kernels over NVLink (see rows[].kernel), fp32 sum in place, one process per GPU",
                    "busbw_gbs_at_1GiB": vt[-1]["busbw_gbs"], "frac_of_900": round(vt[-1]["busbw_gbs"] / 900.0, 4),
                    "bit_exact": all(t["mismatches"] == 0 for t in vt),
                    "allreduce_parity": {"checked_vs": "oracle", "api": "b200v_nvls_comm_allreduce on caller-owned buffers",
                                         "sizes_bytes": [4 * m for m in PARITY_FLOATS],
                                         "mismatches": pb + pa, "mismatches_before_timed": pb,
                                         "mismatches_after_timed_3_back_to_back": pa},
                    "gpu_launches": world * ar_probe_launches(vrows, 5, 20), "rows": vt}

    # ---- CPU baseline: the sample's host verification loop on this box's host cores, N=1 ----------
    cpu = None
    if world == 1 and rank == 0 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        best, med, times = cpu_verify_loop(n, threads, 60, oracle)        # ~1-2 s on all threads
        best1, med1, times1 = cpu_verify_loop(n, 1, 30, oracle)           # ~8 s on one
        cpu = {"value": round(med, 3), "best": round(best, 3), "unit": UNIT, "cores": threads, "kind": "port",
               "sample": "60 passes of the sample's host verification loop over the full 2^%d-element arrays on "
                         "%d threads (%.1f s) — the same loop `--impl reference` times" % (args.log2n, threads, sum(times)),
               "single_thread": {"value": round(med1, 3), "best": round(best1, 3), "cores": 1,
                                 "note": "the sample's own form (one thread); reported for context, not the baseline",
                                 "sample": "30 passes, %.1f s" % sum(times1)}}

    if rank == 0:
        line = {"metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world,
                "steps": steps, "warmup": warmup, "ms_per_step": round(1e3 * t_vadd / steps, 5),
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": workload_config(n, world),
                "per_gpu_gbs": round(value / world, 1), "triad_gbs": round(triad, 1),
                "frac_of_nominal_8000": round(value / world / NOMINAL_HBM_GBS, 4),
                "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                             "frac": round(achieved / peak, 4), "traffic": ncu_traffic(n),
                             "traffic_source": "profiles/traffic.json (ncu --set full capture of this kernel, not "
                                               "re-measured in this run)",
                             "peak_source": peak_src, "kernel": kernel,
                             "algorithmic_bytes_per_launch": bytes_step, "sustained": sustained,
                             "sustained_frac": round(sustained["gbs_per_gpu"] / peak, 4)},
                "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "h2d_bytes_per_step": 8 * n,
                        "d2h_bytes_per_step": 4 * n, "steps": e2e_steps,
                        "api": "b200v_vectoradd (b200v_host_alloc buffers, chunked 3-stream pipeline)",
                        "result_check": "first and last 64 Ki elements here; the full-size compare is a GPU test"},
                "gpu_launches": steps * world, "e2e_gpu_launches": e2e_launches * world,
                "clocks": clocks}
        if e2e_pageable:
            line["e2e_pageable"] = e2e_pageable
        if stream_sweep:
            line["stream_sweep"] = stream_sweep
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
