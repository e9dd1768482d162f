"""Lint of the wire format around the boundary: the manifests cannot be applied
here (no kubectl, no cluster), so they are parsed and checked against the contract the product
implements — executable names, COMPONENT values, status-file mount, resource and label names."""
import os

import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MANIFESTS = os.path.join(ROOT, "nvidia-gpu-operator-k8s-cluster_b200", "manifests")


def load(name):
    with open(os.path.join(MANIFESTS, name)) as f:
        return [d for d in yaml.safe_load_all(f) if d]


def test_workload_pods_run_vectoradd():
    for name, init, gpu_limit in (("cuda-workload-validation.yaml", "cuda-validation", None),
                                  ("plugin-workload-validation.yaml", "plugin-validation", 1)):
        (pod,) = load(name)
        assert pod["kind"] == "Pod" and pod["metadata"]["namespace"] == "gpu-operator-resources"
        assert pod["spec"]["restartPolicy"] == "OnFailure"
        assert pod["spec"]["nodeSelector"] == {"nvidia.com/gpu.present": "true"}
        (c,) = pod["spec"]["initContainers"]
        assert c["name"] == init and c["command"] == ["sh", "-c"] and c["args"] == ["vectorAdd"]
        limits = c.get("resources", {}).get("limits", {})
        assert limits.get("nvidia.com/gpu") == gpu_limit


def test_validator_daemonset_matches_the_product_contract():
    ds, exporter = load("operator-validator-daemonset.yaml")
    spec = ds["spec"]["template"]["spec"]
    inits = spec["initContainers"]
    assert [c["name"] for c in inits] == ["driver-validation", "toolkit-validation", "cuda-validation",
                                          "plugin-validation"]
    supported = {"driver", "toolkit", "cuda", "plugin"}
    for c in inits:
        assert c["command"] == ["sh", "-c"] and c["args"] == ["nvidia-validator"]
        env = {e["name"]: e.get("value") for e in c["env"]}
        assert env["COMPONENT"] in supported and env["WITH_WAIT"] in ("true", "false")
        mounts = {m["mountPath"] for m in c["volumeMounts"]}
        assert "/run/nvidia/validations" in mounts
    # The DaemonSet pod never exits, and Kubernetes reserves max(init requests, sum of app
    # requests) for a pod's whole life: its effective nvidia.com/gpu request must be 0, or the
    # validator would keep the node's GPUs away from every tenant (and deadlock a fresh node whose
    # device plugin waits for this pod's toolkit-ready).
    def gpus(c):
        r = c.get("resources", {})
        return max(int(r.get("limits", {}).get("nvidia.com/gpu", 0)),
                   int(r.get("requests", {}).get("nvidia.com/gpu", 0)))
    effective = max(max(gpus(c) for c in inits), sum(gpus(c) for c in spec["containers"]))
    assert effective == 0
    assert spec["runtimeClassName"] == "nvidia"
    for c in inits[2:]:
        env = {e["name"]: e.get("value") for e in c["env"]}
        assert env["NVIDIA_VISIBLE_DEVICES"] == "all"
    main = spec["containers"][0]
    assert main["lifecycle"]["preStop"]["exec"]["command"][-1] == "rm -f /run/nvidia/validations/*-ready"
    vols = {v["name"]: v for v in spec["volumes"]}
    assert vols["run-nvidia-validations"]["hostPath"]["path"] == "/run/nvidia/validations"
    # the exporter runs the same binary in metrics mode on upstream's port
    ex = exporter["spec"]["template"]["spec"]["containers"][0]
    assert ex["command"] == ["nvidia-validator"] and ex["args"][:2] == ["--component", "metrics"]
    assert ex["ports"][0]["containerPort"] == 8000


def test_status_files_written_by_the_product_match_the_prestop_glob():
    import fnmatch
    import re
    # every status-file name the product can create, read off its source
    src = open(os.path.join(ROOT, "nvidia-gpu-operator-k8s-cluster_b200", "host", "validator_core.cpp")).read()
    names = set(re.findall(r'"([a-z]+-ready)"', src)) | {"cuda-gpu%d-ready" % i for i in range(8)}
    assert {"cuda-ready", "driver-ready", "toolkit-ready", "plugin-ready"} <= names
    assert '"cuda-gpu" + std::to_string(i) + "-ready"' in src
    for name in names:
        assert fnmatch.fnmatch(name, "*-ready")


def test_dockerfile_installs_the_three_artefacts():
    text = open(os.path.join(MANIFESTS, "Dockerfile")).read()
    for dst in ("/usr/lib/libb200validate.so", "/usr/bin/vectorAdd", "/usr/bin/nvidia-validator"):
        assert dst in text
    assert "make -C nvidia-gpu-operator-k8s-cluster_b200 lib host" in text
