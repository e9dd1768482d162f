"""The validator's host logic and the host copy pool under AddressSanitizer + UBSan (gcc), on the
CPU: representative scenarios of the contract tests must run clean (SURVEY.md §5 "sanitizers")."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "nvidia-gpu-operator-k8s-cluster_b200")
SAN = ["-fsanitize=address,undefined", "-fno-omit-frame-pointer", "-g", "-O1"]


def _clean(stderr):
    return "AddressSanitizer" not in stderr and "runtime error" not in stderr and "LeakSanitizer" not in stderr


@pytest.fixture(scope="module")
def san_fake(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("san") / "fake-validator-san")
    subprocess.check_call(["g++", "-std=c++17", "-Wall"] + SAN + [
        "-I", os.path.join(ROOT, "include"), "-I", os.path.join(PKG, "host"), "-o", exe,
        os.path.join(ROOT, "tests", "host", "fake_main.cpp"), os.path.join(PKG, "host", "validator_core.cpp")])
    return exe


def _run(exe, tmp_path, args=(), **env):
    e = {"PATH": os.environ["PATH"], "OUTPUT_DIR": str(tmp_path), "B200V_SLEEP_SECONDS": "0",
         "ASAN_OPTIONS": "detect_leaks=1:abort_on_error=0", "UBSAN_OPTIONS": "print_stacktrace=1"}
    e.update({k: str(v) for k, v in env.items()})
    p = subprocess.run([exe] + list(args), env=e, capture_output=True, text=True, timeout=120)
    return p.returncode, p.stdout, p.stderr


def test_validator_scenarios_run_clean_under_asan_ubsan(san_fake, tmp_path):
    scenarios = [
        (0, dict(COMPONENT="cuda")),
        (0, dict(COMPONENT="cuda", FAKE_NDEV=4)),
        (1, dict(COMPONENT="cuda", FAKE_NDEV=4, FAKE_MODE="wrong", FAKE_FAIL_DEVICE=1)),
        (1, dict(COMPONENT="cuda", FAKE_MODE="cudaerr", WITH_WAIT="true", B200V_MAX_RETRIES=2)),
        (1, dict(COMPONENT="cuda", B200V_INJECT_FAULT=49999)),
        (0, dict(COMPONENT="cuda", B200V_ELEMS=0)),
        # the fan-out pipes, the start barrier, child rows, the whole-node line and the .prom file
        (0, dict(COMPONENT="cuda", FAKE_NDEV=3, B200V_ELEMS=1 << 22, B200V_WRITE_PROM=1)),
        (0, dict(COMPONENT="cuda", FAKE_NDEV=1, B200V_FANOUT_FORCE=1, B200V_FANOUT_HBM_LOG2=20)),
        (1, dict(COMPONENT="cuda", FAKE_NDEV=3, FAKE_MODE="cudaerr", FAKE_FAIL_DEVICE=2, B200V_ELEMS=1 << 22)),
        (0, dict(B200V_MODE="stream", B200V_GPUS=2, FAKE_NDEV=2, B200V_MIN_LOG2=18, B200V_MAX_LOG2=18)),
        (0, dict(COMPONENT="plugin", B200V_ALLOCATABLE_GPUS=8)),
        (1, dict(COMPONENT="plugin", B200V_ALLOCATABLE_GPUS=0, B200V_PLUGIN_RETRIES=2)),
        (0, dict(COMPONENT="toolkit", B200V_NVIDIA_SMI="true")),
        (0, dict(COMPONENT="metrics", B200V_METRICS_ONESHOT=1)),
        (0, dict(B200V_MODE="stream", B200V_MIN_LOG2=18, B200V_MAX_LOG2=19, B200V_WRITE_PROM=1)),
        (0, dict(B200V_MODE="nccl", FAKE_NDEV=8, B200V_MAX_LOG2=12, B200V_WRITE_PROM=1)),
        (0, dict(B200V_MODE="nvls", FAKE_NDEV=8, B200V_MAX_LOG2=12)),
        (1, dict(COMPONENT="nope")),
        (1, dict()),
    ]
    for want_rc, env in scenarios:
        rc, _out, err = _run(san_fake, tmp_path, **env)
        assert rc == want_rc, (env, err)
        assert _clean(err), (env, err)
    rc, _out, err = _run(san_fake, tmp_path, args=["--component", "cuda", "--elems", "4097", "--with-wait=false"])
    assert rc == 0 and _clean(err), err
    rc, _out, err = _run(san_fake, tmp_path, args=["--elems"])
    assert rc == 1 and _clean(err), err


def test_copy_pool_runs_clean_under_tsan(tmp_path):
    exe = str(tmp_path / "copy_pool_test_tsan")
    subprocess.check_call(["g++", "-std=c++17", "-pthread", "-O1", "-g", "-fsanitize=thread",
                           "-I", os.path.join(PKG, "csrc"), "-o", exe,
                           os.path.join(ROOT, "tests", "host", "copy_pool_test.cpp")])
    p = subprocess.run([exe], env=dict(os.environ, B200V_COPY_THREADS="4"), capture_output=True, text=True,
                       timeout=600)
    assert p.returncode == 0 and "COPY_POOL_OK" in p.stdout, p.stdout + p.stderr
    assert "ThreadSanitizer" not in p.stderr, p.stderr


def test_copy_pool_runs_clean_under_asan_ubsan(tmp_path):
    exe = str(tmp_path / "copy_pool_test_san")
    subprocess.check_call(["g++", "-std=c++17", "-pthread", "-Wall"] + SAN + [
        "-I", os.path.join(PKG, "csrc"), "-o", exe, os.path.join(ROOT, "tests", "host", "copy_pool_test.cpp")])
    # the pool and its detached workers live until exit by design: leak checking is off for it
    p = subprocess.run([exe], env=dict(os.environ, ASAN_OPTIONS="detect_leaks=0", B200V_COPY_THREADS="4"),
                       capture_output=True, text=True, timeout=300)
    assert p.returncode == 0 and "COPY_POOL_OK" in p.stdout, p.stdout + p.stderr
    assert _clean(p.stderr), p.stderr
