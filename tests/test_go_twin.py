"""The cgo driver (go/validator.go) cannot be compiled in this environment (no Go toolchain), so
this lint keeps it in step with the C++ twin the tests do build: same environment names, status
files, output lines and C-ABI symbols, declared against the real header, and structurally sane."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "nvidia-gpu-operator-k8s-cluster_b200")
GO = open(os.path.join(PKG, "go", "validator.go")).read()
CPP = open(os.path.join(PKG, "host", "validator_core.cpp")).read()
HEADER = open(os.path.join(ROOT, "include", "b200validate.h")).read()


def test_same_environment_names():
    names = set(re.findall(r'"((?:B200V|WITH|OUTPUT|COMPONENT|NODE|OPERATOR|DRIVER|METRICS)[A-Z0-9_]*)"', CPP))
    names -= {"B200V_INJECT_DEVICE", "B200V_MIN_BUSBW_GBS", "B200V_WRITE_PROM", "B200V_METRICS_ONESHOT",
              "B200V_METRICS_MAX_REQUESTS", "NVIDIA_VISIBLE_DEVICES"}        # test/diagnostic-only knobs
    missing = sorted(n for n in names if '"%s"' % n not in GO)
    assert missing == [], "go/validator.go does not read %s" % missing


def test_same_status_files_and_output_lines():
    for token in ("cuda-ready", "driver-ready", "toolkit-ready", "plugin-ready", "cuda-gpu%d-ready",
                  "[Vector addition of %d elements]", "Copy input data from the host memory to the CUDA device",
                  "CUDA kernel launch with %d blocks of %d threads",
                  "Copy output data from the CUDA device to the host memory", "Test PASSED", "Done",
                  "Result verification failed at element %d!", "is not ready, retrying after %d seconds",
                  "invalid -c <component-name> flag: must not be empty string",
                  "invalid component specified for validation: %s",
                  "GPU resources are not yet discovered by the node, retry: %d",
                  "gpu_operator_node_", "b200v_node_cuda_gpu_ready", "/run/nvidia/validations"):
        assert token in GO, token
    for comp in ('"cuda"', '"driver"', '"toolkit"', '"plugin"', '"metrics"', '"stream"', '"nccl"', '"nvls"'):
        assert comp in GO, comp


def test_every_c_symbol_it_calls_is_declared_in_the_header():
    called = set(re.findall(r"\bC\.(b200v_[a-z0-9_]+)\s*\(", GO))
    assert {"b200v_vectoradd", "b200v_device_count", "b200v_stream", "b200v_allreduce_probe",
            "b200v_nvls_allreduce_probe", "b200v_strerror"} <= called
    for sym in called:
        assert re.search(r"\b%s\s*\(" % sym, HEADER), "%s is not in b200validate.h" % sym
    for typ in set(re.findall(r"\bC\.(b200v_[a-z_]+)\b(?!\s*\()", GO)):
        assert ("} %s;" % typ) in HEADER, typ
    for const in set(re.findall(r"\bC\.(B200V_[A-Z_]+)\b", GO)):
        assert re.search(r"#define\s+%s\b" % const, HEADER), const


def test_structurally_sane():
    go = GO.split('import "C"', 1)[1]                        # after the cgo preamble
    imports = re.search(r"import \((.*?)\)", go, re.S).group(1).split()
    body = re.sub(r'"(?:\\.|[^"\\])*"', '""', go)            # drop string literals
    body = re.sub(r"//[^\n]*", "", body)
    for a, b in ("{}", "()", "[]"):
        assert body.count(a) == body.count(b), "unbalanced %s%s" % (a, b)
    for imp in imports:
        pkg_name = imp.strip('"').split("/")[-1]
        assert re.search(r"\b%s\." % re.escape(pkg_name), body), "unused import %s" % imp
    assert "package main" in GO and "func main()" in GO and "NOT COMPILED IN THIS ENVIRONMENT" in GO
