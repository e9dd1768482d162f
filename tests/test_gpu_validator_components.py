"""GPU-box runs of the real `bin/nvidia-validator` (linked to libb200validate.so, no fake ABI) for
the components that round 1 had only exercised on the CPU with scripted stand-ins:

  a8  COMPONENT=driver / toolkit with the box's own nvidia-smi   (reference README.md:152-168)
  f1  COMPONENT=plugin with the allocatable count = visible GPUs (reference README.md:122)
  f2  B200V_MODE=stream -> *.prom -> COMPONENT=metrics            (reference README.md:107)
  cfg3 the product's own per-GPU fan-out: child rows, whole-node line, status files, .prom
"""
import json
import os
import shutil
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "nvidia-gpu-operator-k8s-cluster_b200", "bin", "nvidia-validator")


def run(tmp_path, *args, **env):
    e = dict(os.environ, OUTPUT_DIR=str(tmp_path))
    e.update({k: str(v) for k, v in env.items()})
    p = subprocess.run([EXE, *args], env=e, capture_output=True, text=True, timeout=900)
    return p.returncode, p.stdout, p.stderr


def ready(tmp_path):
    return sorted(f for f in os.listdir(tmp_path) if f.endswith("-ready"))


def test_driver_and_toolkit_components_with_the_real_nvidia_smi(capi, gpu, tmp_path):
    if not shutil.which("nvidia-smi"):
        pytest.skip("nvidia-smi is not on PATH on this box")
    for comp in ("driver", "toolkit"):
        (tmp_path / (comp + "-ready")).write_text("stale")
        # DRIVER_ROOT points nowhere, so the host-installed-driver branch (plain nvidia-smi) is taken
        rc, out, err = run(tmp_path, COMPONENT=comp, DRIVER_ROOT=str(tmp_path / "no-driver-root"))
        assert rc == 0, err
        assert "%s validation is successful" % comp in out
        assert "NVIDIA-SMI" in out and "Driver Version" in out       # the table the runbook reads
        assert (tmp_path / (comp + "-ready")).read_text() == ""       # stale file replaced
    assert ready(tmp_path) == ["driver-ready", "toolkit-ready"]
    # a driver that does not answer leaves no status file behind
    rc, _out, err = run(tmp_path, COMPONENT="driver", B200V_NVIDIA_SMI="false", DRIVER_ROOT=str(tmp_path / "x"))
    assert rc == 1 and "driver validation failed" in err
    assert ready(tmp_path) == ["toolkit-ready"]


def test_plugin_component_through_the_real_library(capi, gpu, tmp_path):
    rc, out, err = run(tmp_path, COMPONENT="plugin", B200V_ALLOCATABLE_GPUS=gpu)
    assert rc == 0, err
    assert "node advertises nvidia.com/gpu: %d" % gpu in out
    assert "Test PASSED" in out and "plugin validation is successful" in out
    assert ready(tmp_path) == ["plugin-ready"]
    # the kubectl-describe form, from a file
    alloc = tmp_path / "alloc.txt"
    alloc.write_text("Allocatable:\n  cpu: 192\n  nvidia.com/gpu:     %d\n" % gpu)
    rc, out, err = run(tmp_path, COMPONENT="plugin", B200V_ALLOCATABLE_FILE=alloc, B200V_INJECT_FAULT=12)
    assert rc == 1 and "Result verification failed at element 12!" in err
    assert ready(tmp_path) == []
    rc, _out, err = run(tmp_path, COMPONENT="plugin", B200V_ALLOCATABLE_GPUS=0, B200V_PLUGIN_RETRIES=2,
                        B200V_SLEEP_SECONDS=0)
    assert rc == 1 and "GPU resources are not discovered by the node" in err


def test_stream_sweep_prom_file_is_what_the_metrics_mode_exports(capi, gpu, tmp_path):
    rc, out, err = run(tmp_path, B200V_MODE="stream", B200V_MIN_LOG2=22, B200V_MAX_LOG2=26, B200V_WRITE_PROM=1)
    assert rc == 0, err
    rows = [json.loads(l) for l in out.splitlines()]
    assert len(rows) == 5 * 4 and all(r["mismatches"] == 0 for r in rows)
    hbm = {(r["op"], r["log2_n"]): r["median_gbs"] for r in rows if not r["l2_resident"]}
    assert hbm, "no HBM-bound row in the sweep"
    run(tmp_path, COMPONENT="plugin", B200V_ALLOCATABLE_GPUS=gpu, WITH_WORKLOAD="false")
    rc, metrics, err = run(tmp_path, COMPONENT="metrics", B200V_METRICS_ONESHOT=1)
    assert rc == 0, err
    exported = {}
    for line in metrics.splitlines():
        if line.startswith("b200v_hbm_bandwidth_gbs{"):
            labels, value = line.rsplit(" ", 1)
            op = labels.split('op="')[1].split('"')[0]
            lg = int(labels.split('log2_n="')[1].split('"')[0])
            exported[(op, lg)] = float(value)
    assert exported.keys() == hbm.keys()
    for key, v in hbm.items():
        assert abs(exported[key] - v) <= 0.051, key                  # same number, one decimal in the file
    assert "gpu_operator_node_device_plugin_devices_total %d" % gpu in metrics.splitlines()
    assert "gpu_operator_node_plugin_ready 1" in metrics.splitlines()
    # plausibility on a B200: HBM-bound add/triad rows are in the TB/s range
    assert max(v for (op, _lg), v in hbm.items() if op in ("add", "triad")) > 3000


def test_fanout_rows_and_aggregate_from_the_product(capi, gpu, tmp_path):
    """cfg3 through `nvidia-validator` itself: one child per GPU behind the start barrier (forced
    even on a one-GPU lease), 2^24 elements each, device-resident HBM figure per GPU, whole-node
    line, status files and the .prom file."""
    n = 1 << 24
    rc, out, err = run(tmp_path, COMPONENT="cuda", B200V_ELEMS=n, B200V_FANOUT_FORCE=1, B200V_WRITE_PROM=1)
    assert rc == 0, err
    rows = [json.loads(l) for l in out.splitlines() if l.startswith("{")]
    kids = sorted((r for r in rows if r["mode"] == "fanout-child"), key=lambda r: r["gpu"])
    assert [r["gpu"] for r in kids] == list(range(gpu))
    for r in kids:
        assert r["passed"] == 1 and r["elems"] == n and r["hbm_log2_n"] == 24 and r["hbm_mismatches"] == 0
        assert r["kernel_launches"] >= 1 and r["kernel_ms"] > 0 and r["total_ms"] >= r["kernel_ms"]
        assert r["hbm_median_gbs"] > 1000 and r["e2e_gbs"] > 1
    # every GPU ran the same seeded arrays? no: each has its own seed, so digests differ per GPU
    assert len({r["hbm_digest"] for r in kids}) == gpu
    (agg,) = [r for r in rows if r["mode"] == "fanout"]
    assert agg["gpus"] == gpu and agg["reported"] == gpu and len(agg["per_gpu_hbm_gbs"]) == gpu
    slowest = max(r["hbm_median_ms"] for r in kids)
    assert abs(agg["aggregate_hbm_gbs"] - gpu * 12 * n / (slowest * 1e-3) / 1e9) < 0.01 * agg["aggregate_hbm_gbs"] + 1
    assert out.count("Test PASSED") == gpu
    assert ready(tmp_path) == sorted(["cuda-gpu%d-ready" % i for i in range(gpu)] + ["cuda-ready"])
    prom = (tmp_path / "fanout-bandwidth.prom").read_text()
    assert prom.count("b200v_fanout_hbm_bandwidth_gbs{") == gpu and "b200v_fanout_hbm_aggregate_gbs " in prom
    rc, metrics, _err = run(tmp_path, COMPONENT="metrics", B200V_METRICS_ONESHOT=1)
    assert "b200v_fanout_hbm_aggregate_gbs %.1f" % agg["aggregate_hbm_gbs"] in metrics
