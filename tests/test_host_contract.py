"""CPU tests of the validator contract: exit codes, <OUTPUT_DIR>/*-ready status
files, the sample's output lines, per-GPU fan-out, WITH_WAIT retries, flag/env parsing.

The product host logic (host/validator_core.cpp) is linked against tests/host/fake_main.cpp, a
scripted stand-in for the C ABI, so the kubelet-facing behaviour is checked without a GPU.  The
shell harness here plays kubelet: it runs the entrypoint with the DaemonSet's environment and
looks at files and exit codes."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "nvidia-gpu-operator-k8s-cluster_b200")
BUILD = os.path.join(ROOT, "tests", "host", "build")
FAKE = os.path.join(BUILD, "fake-validator")


@pytest.fixture(scope="session")
def fake():
    os.makedirs(BUILD, exist_ok=True)
    srcs = [os.path.join(ROOT, "tests", "host", "fake_main.cpp"), os.path.join(PKG, "host", "validator_core.cpp")]
    newest = max(os.path.getmtime(p) for p in srcs + [os.path.join(PKG, "host", "validator_core.h")])
    if not os.path.exists(FAKE) or os.path.getmtime(FAKE) < newest:
        subprocess.check_call(["g++", "-O1", "-std=c++17", "-Wall", "-I", os.path.join(ROOT, "include"),
                               "-I", os.path.join(PKG, "host"), "-o", FAKE] + srcs)
    return FAKE


def run(fake, tmp_path, args=(), **env):
    e = {"PATH": os.environ["PATH"], "OUTPUT_DIR": str(tmp_path), "B200V_SLEEP_SECONDS": "0"}
    e.update({k: str(v) for k, v in env.items()})
    p = subprocess.run([fake] + list(args), env=e, capture_output=True, text=True, timeout=60)
    return p.returncode, p.stdout, p.stderr


def ready(tmp_path):
    return sorted(f for f in os.listdir(tmp_path) if f.endswith("-ready"))


def test_cuda_single_gpu_pass(fake, tmp_path):
    rc, out, err = run(fake, tmp_path, COMPONENT="cuda")
    assert rc == 0, err
    assert ready(tmp_path) == ["cuda-gpu0-ready", "cuda-ready"]
    assert os.path.getsize(tmp_path / "cuda-ready") == 0
    lines = out.splitlines()
    assert lines[0] == "[Vector addition of 50000 elements]"
    assert lines[1] == "Copy input data from the host memory to the CUDA device"
    assert lines[2] == "CUDA kernel launch with 25 blocks of 256 threads"
    assert lines[3] == "Copy output data from the CUDA device to the host memory"
    assert lines[4] == "Test PASSED" and lines[5] == "Done"


def test_cuda_failure_removes_stale_status_and_exits_1(fake, tmp_path):
    (tmp_path / "cuda-ready").write_text("")
    (tmp_path / "cuda-gpu0-ready").write_text("")
    (tmp_path / "driver-ready").write_text("")          # another component's file is not ours to touch
    rc, out, err = run(fake, tmp_path, COMPONENT="cuda", FAKE_MODE="wrong")
    assert rc == 1
    assert "Result verification failed at element 17!" in err
    assert "Test PASSED" not in out
    assert ready(tmp_path) == ["driver-ready"]


def test_with_workload_false_only_writes_the_status_file(fake, tmp_path):
    (tmp_path / "cuda-gpu3-ready").write_text("")      # stale per-GPU file from an earlier run
    rc, out, _err = run(fake, tmp_path, COMPONENT="cuda", WITH_WORKLOAD="false", FAKE_NDEV=0)
    assert rc == 0 and "disabled" in out and "Vector addition" not in out
    assert ready(tmp_path) == ["cuda-ready"]


def test_stream_sweep_fans_out_per_gpu(fake, tmp_path):
    rc, out, err = run(fake, tmp_path, B200V_MODE="stream", FAKE_NDEV=4, B200V_GPUS=4, B200V_MIN_LOG2=27,
                       B200V_MAX_LOG2=27, B200V_WRITE_PROM=1)
    assert rc == 0, err
    rows = [json.loads(line) for line in out.splitlines()]
    assert sorted(r["gpu"] for r in rows) == [g for g in range(4) for _ in range(4)]      # 4 ops per GPU
    assert sorted(f for f in os.listdir(tmp_path) if f.endswith(".prom")) == [
        "hbm-bandwidth-gpu%d.prom" % g for g in range(4)]
    rc, metrics, _err = run(fake, tmp_path, COMPONENT="metrics", B200V_METRICS_ONESHOT=1)
    assert metrics.count("# TYPE b200v_hbm_bandwidth_gbs gauge") == 1                       # one family header
    assert metrics.count('b200v_hbm_bandwidth_gbs{gpu="') == 16
    rc, _out, err = run(fake, tmp_path, B200V_MODE="stream", FAKE_NDEV=2, B200V_GPUS=4, B200V_MAX_LOG2=18)
    assert rc == 1 and "only 2 device(s) visible" in err
    rc, _out, err = run(fake, tmp_path, B200V_MODE="stream", FAKE_NDEV=2, B200V_GPUS=2, B200V_MAX_LOG2=18,
                        FAKE_MODE="wrong")
    assert rc == 1 and "stream sweep failed on gpu 0" in err and "stream sweep failed on gpu 1" in err


def test_probe_that_never_completes_is_cut_off(fake, tmp_path):
    import time
    t0 = time.time()
    rc, _out, err = run(fake, tmp_path, B200V_MODE="nccl", FAKE_NDEV=8, FAKE_MODE="hang", B200V_PROBE_TIMEOUT_S=1)
    assert rc == 1 and "probe timed out" in err and time.time() - t0 < 20


def test_cuda_error_from_library_is_reported(fake, tmp_path):
    rc, out, err = run(fake, tmp_path, COMPONENT="cuda", FAKE_MODE="cudaerr")
    assert rc == 1 and "illegal memory access" in err
    assert ready(tmp_path) == []


def test_no_device_fails_without_status_file(fake, tmp_path):
    rc, _out, err = run(fake, tmp_path, COMPONENT="cuda", FAKE_NDEV=0)
    assert rc == 1 and "no CUDA device" in err
    assert ready(tmp_path) == []


def test_fault_injection_hits_the_verification_loop(fake, tmp_path):
    rc, _out, err = run(fake, tmp_path, COMPONENT="cuda", B200V_INJECT_FAULT=4242)
    assert rc == 1 and "Result verification failed at element 4242!" in err
    assert ready(tmp_path) == []


def test_fanout_all_pass(fake, tmp_path):
    rc, out, err = run(fake, tmp_path, COMPONENT="cuda", FAKE_NDEV=4)
    assert rc == 0, err
    assert ready(tmp_path) == ["cuda-gpu0-ready", "cuda-gpu1-ready", "cuda-gpu2-ready", "cuda-gpu3-ready",
                               "cuda-ready"]
    assert out.count("Test PASSED") == 4
    assert "successful on 4 device(s)" in out


def test_fanout_one_gpu_fails(fake, tmp_path):
    rc, out, err = run(fake, tmp_path, COMPONENT="cuda", FAKE_NDEV=4, FAKE_MODE="wrong", FAKE_FAIL_DEVICE=2)
    assert rc == 1
    assert ready(tmp_path) == ["cuda-gpu0-ready", "cuda-gpu1-ready", "cuda-gpu3-ready"]
    assert "cuda validation failed on gpu 2 (exit code 1)" in err
    assert out.count("Test PASSED") == 3


def test_fanout_subset_and_oversubscription(fake, tmp_path):
    rc, _out, _err = run(fake, tmp_path, COMPONENT="cuda", FAKE_NDEV=8, B200V_GPUS=2)
    assert rc == 0 and ready(tmp_path) == ["cuda-gpu0-ready", "cuda-gpu1-ready", "cuda-ready"]
    rc, _out, err = run(fake, tmp_path, COMPONENT="cuda", FAKE_NDEV=2, B200V_GPUS=4)
    assert rc == 1 and "only 2 device(s) visible" in err
    assert ready(tmp_path) == []


def test_with_wait_retries_until_limit(fake, tmp_path):
    rc, _out, err = run(fake, tmp_path, COMPONENT="cuda", WITH_WAIT="true", FAKE_MODE="cudaerr",
                        B200V_MAX_RETRIES=3)
    assert rc == 1
    assert err.count("cuda is not ready, retrying after 0 seconds") == 2
    rc, _out, err = run(fake, tmp_path, COMPONENT="cuda", WITH_WAIT="false", FAKE_MODE="cudaerr")
    assert rc == 1 and "retrying" not in err


def test_flags_override_environment(fake, tmp_path):
    other = tmp_path / "elsewhere"
    other.mkdir()
    rc, out, _err = run(fake, tmp_path, args=["--component", "cuda", "--output-dir=%s" % other, "--elems", "4096"],
                        COMPONENT="bogus")
    assert rc == 0 and "[Vector addition of 4096 elements]" in out
    assert ready(other) == ["cuda-gpu0-ready", "cuda-ready"] and ready(tmp_path) == []
    rc, _out, err = run(fake, tmp_path, args=["--no-such-flag"], COMPONENT="cuda")
    assert rc == 1 and "flag provided but not defined: --no-such-flag" in err
    rc, _out, err = run(fake, tmp_path, args=["-c"], COMPONENT="cuda")
    assert rc == 1 and "flag needs an argument" in err


def test_component_validation(fake, tmp_path):
    rc, _out, err = run(fake, tmp_path)
    assert rc == 1 and "must not be empty string" in err
    for comp in ("mofed", "vfio-pci", "nvidia-fs", "nope"):
        rc, _out, err = run(fake, tmp_path, COMPONENT=comp)
        assert rc == 1 and "invalid component specified for validation: %s" % comp in err
    assert ready(tmp_path) == []


def test_driver_and_toolkit_use_nvidia_smi(fake, tmp_path):
    rc, out, _err = run(fake, tmp_path, COMPONENT="toolkit", B200V_NVIDIA_SMI="true")
    assert rc == 0 and ready(tmp_path) == ["toolkit-ready"] and "toolkit validation is successful" in out
    rc, out, _err = run(fake, tmp_path, COMPONENT="driver", B200V_NVIDIA_SMI="true")
    assert rc == 0 and ready(tmp_path) == ["driver-ready", "toolkit-ready"]
    rc, _out, err = run(fake, tmp_path, COMPONENT="driver", B200V_NVIDIA_SMI="false")
    assert rc == 1 and "driver validation failed" in err
    assert ready(tmp_path) == ["toolkit-ready"]          # stale driver-ready removed, toolkit's kept


def test_unwritable_output_dir_fails(fake, tmp_path):
    rc, _out, err = run(fake, tmp_path, COMPONENT="cuda", OUTPUT_DIR=str(tmp_path / "missing" / "dir"))
    assert rc == 1 and "unable to create status file" in err


def test_vectoradd_entrypoint(fake, tmp_path):
    rc, out, _err = run(fake, tmp_path, FAKE_ENTRY="vectoradd")
    assert rc == 0 and out.splitlines()[-2:] == ["Test PASSED", "Done"]
    assert ready(tmp_path) == []                          # the sample binary writes no status file
    rc, out, err = run(fake, tmp_path, FAKE_ENTRY="vectoradd", FAKE_MODE="wrong")
    assert rc == 1 and "Result verification failed at element 17!" in err
    rc, out, _err = run(fake, tmp_path, FAKE_ENTRY="vectoradd", B200V_ELEMS=0)
    assert rc == 0 and "[Vector addition of 0 elements]" in out


def test_stream_and_nccl_modes_emit_json_rows(fake, tmp_path):
    rc, out, _err = run(fake, tmp_path, B200V_MODE="stream", B200V_MIN_LOG2=18, B200V_MAX_LOG2=19)
    assert rc == 0
    rows = [json.loads(line) for line in out.splitlines()]
    assert [(r["op"], r["log2_n"]) for r in rows] == [(op, lg) for lg in (18, 19)
                                                      for op in ("copy", "scale", "add", "triad")]
    assert rows[0]["bytes_per_array"] == 1 << 20 and rows[0]["nominal_gbs"] == 8184
    rc, _out, _err = run(fake, tmp_path, B200V_MODE="stream", B200V_MAX_LOG2=18, FAKE_MODE="wrong")
    assert rc == 1
    rc, out, _err = run(fake, tmp_path, B200V_MODE="nccl", FAKE_NDEV=8, B200V_MIN_LOG2=10, B200V_MAX_LOG2=12)
    assert rc == 0
    rows = [json.loads(line) for line in out.splitlines()]
    assert [r["bytes"] for r in rows] == [1024, 2048, 4096] and rows[0]["ranks"] == 8
    assert abs(rows[0]["busbw_gbs"] - 400 * 2 * 7 / 8) < 1e-6
    rc, _out, err = run(fake, tmp_path, B200V_MODE="nccl", FAKE_NDEV=8, B200V_MIN_BUSBW_GBS=720,
                        B200V_MAX_LOG2=12)
    assert rc == 1 and "below the required 720.0 GB/s" in err
    rc, _out, _err = run(fake, tmp_path, B200V_MODE="nccl", FAKE_NDEV=2, FAKE_MODE="wrong", B200V_MAX_LOG2=12)
    assert rc == 1


# ---- plugin component and node-status-exporter metrics -----------------

def test_plugin_waits_for_allocatable_then_validates(fake, tmp_path):
    alloc = tmp_path / "allocatable"
    rc, out, err = run(fake, tmp_path, COMPONENT="plugin", B200V_ALLOCATABLE_GPUS=8)
    assert rc == 0, err
    assert "node advertises nvidia.com/gpu: 8" in out and "Test PASSED" in out
    assert ready(tmp_path) == ["plugin-ready"]
    # the field as `kubectl describe node` prints it (what README.md:122 greps for)
    alloc.write_text("Allocatable:\n  cpu: 192\n  nvidia.com/gpu:     8\n  pods: 110\n")
    rc, out, _err = run(fake, tmp_path, COMPONENT="plugin", B200V_ALLOCATABLE_FILE=alloc, WITH_WORKLOAD="false")
    assert rc == 0 and "nvidia.com/gpu: 8" in out and "Test PASSED" not in out
    alloc.write_text('{"status": {"allocatable": {"nvidia.com/gpu": "4"}}}')
    rc, out, _err = run(fake, tmp_path, COMPONENT="plugin", B200V_ALLOCATABLE_FILE=alloc)
    assert rc == 0 and "nvidia.com/gpu: 4" in out


def test_plugin_fails_when_no_gpu_is_advertised(fake, tmp_path):
    (tmp_path / "plugin-ready").write_text("")
    rc, _out, err = run(fake, tmp_path, COMPONENT="plugin", B200V_ALLOCATABLE_GPUS=0, B200V_PLUGIN_RETRIES=3)
    assert rc == 1
    assert err.count("GPU resources are not yet discovered by the node, retry") == 2
    assert "GPU resources are not discovered by the node" in err
    assert ready(tmp_path) == []
    rc, _out, err = run(fake, tmp_path, COMPONENT="plugin")
    assert rc == 1 and "no Kubernetes client in this build" in err
    rc, _out, err = run(fake, tmp_path, COMPONENT="plugin", B200V_ALLOCATABLE_GPUS=1, FAKE_MODE="wrong")
    assert rc == 1 and "Result verification failed" in err and ready(tmp_path) == []


def test_metrics_oneshot_reflects_status_files_and_bandwidth(fake, tmp_path):
    run(fake, tmp_path, COMPONENT="cuda", FAKE_NDEV=2)
    run(fake, tmp_path, COMPONENT="toolkit", B200V_NVIDIA_SMI="true")
    run(fake, tmp_path, B200V_MODE="stream", B200V_MIN_LOG2=27, B200V_MAX_LOG2=28, B200V_WRITE_PROM=1)
    run(fake, tmp_path, B200V_MODE="nccl", FAKE_NDEV=8, B200V_MAX_LOG2=11, B200V_WRITE_PROM=1)
    rc, out, _err = run(fake, tmp_path, COMPONENT="metrics", B200V_METRICS_ONESHOT=1)
    assert rc == 0
    lines = out.splitlines()
    assert "gpu_operator_node_cuda_ready 1" in lines and "gpu_operator_node_toolkit_ready 1" in lines
    assert "gpu_operator_node_driver_ready 0" in lines and "gpu_operator_node_plugin_ready 0" in lines
    import time
    ts = {l.split()[0]: int(l.split()[1]) for l in lines
          if "_validation_last_success_ts_seconds " in l and not l.startswith("#")}
    assert ts["gpu_operator_node_driver_validation_last_success_ts_seconds"] == 0
    assert abs(ts["gpu_operator_node_cuda_validation_last_success_ts_seconds"] - time.time()) < 300
    assert 'b200v_node_cuda_gpu_ready{gpu="0"} 1' in lines and 'b200v_node_cuda_gpu_ready{gpu="1"} 1' in lines
    # upstream's meaning of this gauge: GPUs the device plugin advertises — nothing yet, because
    # plugin validation has not run; the per-GPU cuda files do not feed it
    assert "gpu_operator_node_device_plugin_devices_total 0" in lines
    run(fake, tmp_path, COMPONENT="plugin", B200V_ALLOCATABLE_GPUS=8, WITH_WORKLOAD="false")
    _rc, out3, _err = run(fake, tmp_path, COMPONENT="metrics", B200V_METRICS_ONESHOT=1)
    assert "gpu_operator_node_device_plugin_devices_total 8" in out3.splitlines()
    assert "gpu_operator_node_plugin_ready 1" in out3.splitlines()
    assert any(l.startswith('b200v_hbm_bandwidth_gbs{gpu="0",op="triad",log2_n="28"} ') for l in lines)
    assert 'b200v_nvlink_allreduce_busbw_gbs{impl="nccl",ranks="8",bytes="2048"} 700.00' in lines
    # the hand-written NVLink-multicast probe is the same sweep under B200V_MODE=nvls
    rc, out2, _err = run(fake, tmp_path, B200V_MODE="nvls", FAKE_NDEV=8, B200V_MIN_LOG2=10, B200V_MAX_LOG2=11)
    assert rc == 0 and [json.loads(l)["mode"] for l in out2.splitlines()] == ["nvls", "nvls"]
    # every sample line is `name{labels} value` or `name value`, every family has HELP and TYPE
    for l in lines:
        assert l.startswith("#") or len(l.rsplit(" ", 1)) == 2
    assert out.count("# TYPE") >= 7


def test_metrics_http_endpoint(fake, tmp_path):
    import socket
    import urllib.request
    (tmp_path / "driver-ready").write_text("")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = {"PATH": os.environ["PATH"], "OUTPUT_DIR": str(tmp_path), "COMPONENT": "metrics",
           "B200V_METRICS_MAX_REQUESTS": "2"}
    p = subprocess.Popen([fake, "--metrics-port", str(port)], env=env, stdout=subprocess.PIPE, text=True)
    try:
        assert "listening on :%d/metrics" % port in p.stdout.readline()
        body = urllib.request.urlopen("http://127.0.0.1:%d/metrics" % port, timeout=10).read().decode()
        assert "gpu_operator_node_driver_ready 1" in body and "gpu_operator_node_cuda_ready 0" in body
        with pytest.raises(Exception):
            urllib.request.urlopen("http://127.0.0.1:%d/other" % port, timeout=10)
        assert p.wait(timeout=10) == 0
    finally:
        if p.poll() is None:
            p.kill()


def test_fanout_children_report_rows_and_the_parent_aggregates(fake, tmp_path):
    """cfg3 through the product's own fan-out: every child hands the parent one JSON row, the
    parent prints them, the whole-node line and (with B200V_WRITE_PROM) a .prom file."""
    rc, out, _err = run(fake, tmp_path, COMPONENT="cuda", FAKE_NDEV=4, B200V_ELEMS=1 << 22, B200V_WRITE_PROM=1)
    assert rc == 0
    rows = [json.loads(l) for l in out.splitlines() if l.startswith("{")]
    kids = [r for r in rows if r["mode"] == "fanout-child"]
    assert sorted(r["gpu"] for r in kids) == [0, 1, 2, 3]
    for r in kids:
        assert r["passed"] == 1 and r["elems"] == 1 << 22 and r["hbm_log2_n"] == 22
        assert r["hbm_mismatches"] == 0 and r["kernel_launches"] == 1
    (agg,) = [r for r in rows if r["mode"] == "fanout"]
    assert agg["gpus"] == 4 and agg["reported"] == 4 and len(agg["per_gpu_hbm_gbs"]) == 4
    # the fake times every launch at 1 ms: 4 GPUs x 12 B x 2^22 elements over 1 ms
    assert abs(agg["aggregate_hbm_gbs"] - 4 * 12 * (1 << 22) / 1e-3 / 1e9) < 0.2
    assert abs(agg["sum_hbm_gbs"] - sum(agg["per_gpu_hbm_gbs"])) < 0.5
    prom = (tmp_path / "fanout-bandwidth.prom").read_text()
    assert 'b200v_fanout_hbm_bandwidth_gbs{gpu="3"}' in prom and "b200v_fanout_hbm_aggregate_gbs " in prom
    assert ready(tmp_path) == ["cuda-gpu0-ready", "cuda-gpu1-ready", "cuda-gpu2-ready", "cuda-gpu3-ready",
                               "cuda-ready"]
    # the sample's size stays lean: no device-resident measurement unless asked for
    rc, out, _err = run(fake, tmp_path, COMPONENT="cuda", FAKE_NDEV=2)
    kids = [json.loads(l) for l in out.splitlines() if l.startswith('{"mode":"fanout-child"')]
    assert rc == 0 and len(kids) == 2 and all("hbm_log2_n" not in r for r in kids)
    rc, out, _err = run(fake, tmp_path, COMPONENT="cuda", FAKE_NDEV=2, B200V_FANOUT_HBM_LOG2=20)
    kids = [json.loads(l) for l in out.splitlines() if l.startswith('{"mode":"fanout-child"')]
    assert rc == 0 and all(r["hbm_log2_n"] == 20 for r in kids)


def test_fanout_barrier_does_not_wedge_on_a_failing_child(fake, tmp_path):
    """A child whose GPU fails still arrives at (or abandons) the barrier; the others run, the
    parent reports the failure and leaves no cuda-ready."""
    import time
    t0 = time.time()
    rc, out, err = run(fake, tmp_path, COMPONENT="cuda", FAKE_NDEV=3, FAKE_MODE="cudaerr", FAKE_FAIL_DEVICE=1,
                       B200V_FANOUT_BARRIER_TIMEOUT_S=20)
    assert rc == 1 and time.time() - t0 < 15
    assert "cuda validation failed on gpu 1" in err
    assert ready(tmp_path) == ["cuda-gpu0-ready", "cuda-gpu2-ready"]
    kids = {json.loads(l)["gpu"]: json.loads(l) for l in out.splitlines() if l.startswith('{"mode":"fanout-child"')}
    assert kids[0]["passed"] == 1 and kids[1]["passed"] == 0 and kids[2]["passed"] == 1
