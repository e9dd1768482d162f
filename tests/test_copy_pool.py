"""CPU test of the host copy pool that stages pageable buffers (csrc/copy_pool.h): built with g++
from the product header, no GPU involved."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_copy_pool_copies_every_size_class_exactly(tmp_path):
    exe = str(tmp_path / "copy_pool_test")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-pthread", "-Wall",
                           "-I", os.path.join(ROOT, "nvidia-gpu-operator-k8s-cluster_b200", "csrc"),
                           "-o", exe, os.path.join(ROOT, "tests", "host", "copy_pool_test.cpp")])
    for env_threads in (None, "1", "3", "16"):
        env = dict(os.environ)
        if env_threads:
            env["B200V_COPY_THREADS"] = env_threads
        out = subprocess.run([exe], env=env, capture_output=True, text=True, timeout=120)
        assert out.returncode == 0 and "COPY_POOL_OK" in out.stdout, out.stdout + out.stderr
        if env_threads:
            assert "threads=%s" % env_threads in out.stdout
