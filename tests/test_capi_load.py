"""CPU tests of the boundary: the shared library loads, exports exactly what include/b200validate.h
declares, and refuses to compute without a device (there is no CPU fallback)."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest


def header_functions(path):
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200v_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(capi):
    declared = header_functions(capi.HEADER_PATH)
    assert len(declared) >= 20
    lib = capi.lib()
    for name in declared:
        assert hasattr(lib, name), "libb200validate.so lacks %s" % name
    # and the binding mirrors the header one to one
    assert sorted(capi.SIGNATURES) == declared


def test_exports_are_plain_c_symbols(capi):
    out = subprocess.check_output(["nm", "-D", "--defined-only", capi.LIB_PATH], text=True)
    exported = sorted(line.split()[-1] for line in out.splitlines() if " T b200v_" in line)
    assert exported == header_functions(capi.HEADER_PATH)
    # nothing from torch or the oracle is linked into the product library
    needed = subprocess.check_output(["readelf", "-d", capi.LIB_PATH], text=True)
    assert "torch" not in needed and "oracle" not in needed and "libnccl" not in needed


def test_struct_layouts_match_header(capi):
    """Sizes computed by gcc from the header equal the ctypes mirrors."""
    src = r'''
#include <stdio.h>
#include "b200validate.h"
int main(void) {
  printf("%zu %zu %zu %zu\n", sizeof(b200v_info), sizeof(b200v_timing),
         sizeof(b200v_stream_result), sizeof(b200v_ar_row));
  return 0; }
'''
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "s.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "s")
        subprocess.check_call(["gcc", "-I", os.path.dirname(capi.HEADER_PATH), "-o", exe, c])
        sizes = [int(v) for v in subprocess.check_output([exe], text=True).split()]
    assert sizes == [ctypes.sizeof(capi.Info), ctypes.sizeof(capi.Timing),
                     ctypes.sizeof(capi.StreamResult), ctypes.sizeof(capi.ArRow)]


def test_version_and_strerror(capi):
    assert capi.version() == (0, 1)
    assert capi.strerror(0) == "ok"
    assert capi.strerror(capi.E_MISMATCH) == "result verification failed"
    assert capi.strerror(capi.E_NODEVICE) == "no CUDA device"
    assert capi.strerror(2)          # cudaErrorMemoryAllocation has a runtime string
    assert capi.strerror(-99) == "unknown b200v error"


def test_no_device_means_error_not_fallback(capi):
    """Without a GPU every compute entry point reports an error; nothing is computed on the CPU."""
    if capi.device_count() > 0:
        pytest.skip("a CUDA device is visible; the no-device path cannot be exercised here")
    A = np.ones(16, np.float32)
    B = np.ones(16, np.float32)
    C = np.full(16, -7.0, np.float32)
    with pytest.raises(capi.B200VError) as ei:
        capi.vectoradd(A, B, C)
    assert ei.value.code == capi.E_NODEVICE
    assert (C == -7.0).all()                     # output untouched
    with pytest.raises(capi.B200VError):
        capi.stream(capi.OP_TRIAD, 1024)
    with pytest.raises(capi.B200VError):
        capi.device_info(0)
    with pytest.raises(capi.B200VError):
        capi.allreduce_probe([0])


def test_missing_library_fails_loudly(pkg, monkeypatch, tmp_path):
    monkeypatch.setattr(pkg.capi, "_lib", None)
    monkeypatch.setattr(pkg.capi, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(OSError, match="no CPU fallback"):
        pkg.capi.lib()


def test_product_sources_do_not_reference_oracle():
    """The product tree never includes, links or imports anything under oracle/."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    prod = os.path.join(root, "nvidia-gpu-operator-k8s-cluster_b200")
    offenders = []
    for base, _dirs, files in os.walk(prod):
        if os.sep + "build" in base or os.sep + "lib" in base or os.sep + "bin" in base:
            continue
        for f in files:
            if f.endswith((".cu", ".cuh", ".cpp", ".h", ".py", ".go", "Makefile")):
                text = open(os.path.join(base, f), errors="replace").read()
                if re.search(r"oracle[/.]|liboracle|import oracle|from oracle", text):
                    offenders.append(os.path.join(base, f))
    assert offenders == []


def test_allreduce_kernel_choice_is_host_logic(capi, monkeypatch):
    """The size/rank-count table of the hand-written all-reduce, and its
    environment overrides, evaluated without any GPU."""
    for var in ("B200V_NVLS_ALGO", "B200V_NVLS_ONESHOT_MAX"):
        monkeypatch.delenv(var, raising=False)
    KiB, MiB = 1 << 10, 1 << 20
    assert capi.nvls_algo_for(2, 1 * KiB) == "oneshot" and capi.nvls_algo_for(2, 2 * MiB) == "oneshot"
    assert capi.nvls_algo_for(2, 2 * MiB + 16) == "push" and capi.nvls_algo_for(2, 1 << 30) == "push"
    assert capi.nvls_algo_for(4, 1 * MiB) == "oneshot" and capi.nvls_algo_for(4, 1 * MiB + 16) == "multimem"
    assert capi.nvls_algo_for(4, 1 << 30) == "multimem"
    assert capi.nvls_algo_for(8, 512 * KiB) == "oneshot" and capi.nvls_algo_for(8, 512 * KiB + 16) == "multimem"
    assert capi.nvls_algo_for(8, 1 << 30) == "multimem"
    assert capi.nvls_algo_for(3, 1 << 30) == "multimem"          # peer kernels are built for 2, 4 and 8
    assert capi.nvls_algo_for(16, 1 << 30) == "multimem"         # more ranks than peer mappings
    monkeypatch.setenv("B200V_NVLS_ALGO", "p2p")
    assert capi.nvls_algo_for(4, 1 * KiB) == "p2p" and capi.nvls_algo_for(3, 1 * KiB) == "multimem"
    monkeypatch.setenv("B200V_NVLS_ALGO", "oneshot")
    assert capi.nvls_algo_for(8, 4 * MiB) == "oneshot" and capi.nvls_algo_for(8, 4 * MiB + 16) == "multimem"
    monkeypatch.delenv("B200V_NVLS_ALGO")
    monkeypatch.setenv("B200V_NVLS_ONESHOT_MAX", str(64 * KiB))
    assert capi.nvls_algo_for(2, 64 * KiB) == "oneshot" and capi.nvls_algo_for(2, 128 * KiB) == "push"
    with pytest.raises(capi.B200VError):
        capi.nvls_algo_for(1, 1024)
