"""CPU test of the host-side rendezvous the one-process-per-GPU NVLink probe uses
(csrc/rendezvous.h): real forked processes, an abstract-namespace socket, a file descriptor
handed from rank 0 to every other rank, reusable barriers, and a bounded failure when a rank never
shows up.  Built from the product header under ASan/UBSan; no GPU involved."""
import os
import subprocess
import time

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def rv_test(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("rv") / "rendezvous_test")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-g", "-Wall", "-fsanitize=address,undefined",
                           "-I", os.path.join(ROOT, "nvidia-gpu-operator-k8s-cluster_b200", "csrc"),
                           "-o", exe, os.path.join(ROOT, "tests", "host", "rendezvous_test.cpp")])
    return exe


@pytest.mark.parametrize("nranks", [2, 3, 8])
def test_all_ranks_meet_and_share_the_descriptor(rv_test, nranks):
    p = subprocess.run([rv_test, str(nranks)], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0 and "RENDEZVOUS_OK failures=0 bytes=%d" % nranks in p.stdout, p.stdout + p.stderr
    assert "Sanitizer" not in p.stderr and "runtime error" not in p.stderr


def test_a_missing_rank_fails_everyone_within_the_timeout(rv_test):
    t0 = time.time()
    p = subprocess.run([rv_test, "3", "1"], capture_output=True, text=True, timeout=120,
                       env=dict(os.environ, B200V_RV_TIMEOUT_MS="400"))
    assert p.returncode == 1 and "RENDEZVOUS_BAD failures=2" in p.stdout
    assert time.time() - t0 < 30
