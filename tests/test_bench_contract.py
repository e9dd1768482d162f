"""CPU tests of bench.py's output contract: the reference arm runs here (it is CPU code), and the
recorded B200 lines under profiles/ are checked for the keys and invariants the driver reads."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
             "scaling", "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches"}


def test_reference_arm_prints_one_json_line():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1",
                        "--steps", "3", "--warmup", "1", "--log2n", "20"], capture_output=True, text=True,
                       timeout=300, env=dict(os.environ, RANK="0"))
    assert p.returncode == 0, p.stderr
    lines = p.stdout.strip().splitlines()
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert BASE_KEYS <= set(d) and d["impl"] == "reference" and d["unit"] == "GB/s"
    assert d["value"] > 0 and d["steps"] == 3 and d["higher_is_better"] is True
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in d["config"] and "model" not in d["config"]


def test_reference_arm_other_ranks_do_no_work():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, timeout=60, env=dict(os.environ, RANK="1"))
    assert p.returncode == 0 and p.stdout.strip() == ""


@pytest.mark.parametrize("name", ["r01_bench_n1_final.json", "r01_scale_n8.json", "r01_bench_n8_with_nvls.json"])
def test_recorded_b200_lines_keep_the_contract(name):
    with open(os.path.join(ROOT, "profiles", name)) as f:
        text = f.read().strip()
    assert len(text.splitlines()) == 1
    d = json.loads(text)
    assert BASE_KEYS <= set(d)
    n = d["config"]["elements_per_array"]
    assert d["metric"].startswith("vectorAdd") and d["unit"] == "GB/s" and d["dtype"] == "f32"
    assert d["scaling"] == "weak" and d["vs_baseline"] is None and d["data"] == "synthetic"
    # value = all ranks' algorithmic bytes over the slowest rank's time
    expect = d["n_gpus"] * 12 * n / (d["ms_per_step"] * 1e-3) / 1e9
    assert abs(d["value"] - expect) / expect < 1e-3
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-3
    assert r["algorithmic_bytes_per_launch"] == 12 * n
    assert r["traffic"] is None or 0.9 < r["traffic"] / (12 * n) < 1.1      # no wasted re-reads
    assert d["e2e"]["h2d_bytes_per_step"] == 8 * n and d["e2e"]["d2h_bytes_per_step"] == 4 * n
    assert d["e2e"]["value"] < d["value"]
    assert d["gpu_launches"] == d["steps"] * d["n_gpus"]
    c = d["clocks"]
    assert not {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} & set(c["reasons"])
    if d["n_gpus"] == 1:
        assert d["cpu_baseline"]["cores"] == 1 and d["cpu_baseline"]["kind"] == "port"
    else:
        assert d["allreduce"]["bit_exact"] is True


@pytest.mark.parametrize("name", ["r02_bench_n1.json", "r02_bench_n2.json", "r02_bench_n4.json", "r02_bench_n8.json"])
def test_round2_lines_carry_what_the_verdict_asked_for(name):
    with open(os.path.join(ROOT, "profiles", name)) as f:
        d = json.loads(f.read().strip())
    assert BASE_KEYS <= set(d)
    n = d["config"]["elements_per_array"]
    expect = d["n_gpus"] * 12 * n / (d["ms_per_step"] * 1e-3) / 1e9
    assert abs(d["value"] - expect) / expect < 1e-3
    r = d["roofline"]
    # the kernel is named by the library's dispatch, the traffic figure says where it comes from,
    # and the same launch was held for at least two seconds
    assert r["kernel"].startswith("b200v::ew_tile_kernel<VECTORADD,VEC=8,UNROLL=1,BLOCK=1024>")
    assert "profiles/traffic.json" in r["traffic_source"]
    assert r["sustained"]["seconds"] >= 1.99 and r["sustained"]["launches"] > 1000
    assert abs(r["sustained"]["gbs_per_gpu"] - r["achieved"]) / r["achieved"] < 0.02
    assert not {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} & set(d["clocks"]["reasons"])
    assert d["gpu_launches"] == d["steps"] * d["n_gpus"]
    if d["n_gpus"] == 1:
        sweep = d["stream_sweep"]
        assert sweep["digest_failures"] == 0 and sweep["mem_bus_bits"] == 7680
        rows = [x for x in sweep["rows"] if "op" in x]
        assert {x["log2_n"] for x in rows} == {18, 22, 24, 26, 28, 30, 32}
        assert {x["op"] for x in rows} == {"copy", "scale", "add", "triad"}
        assert all(x["digest_ok"] and x["checked_vs"].startswith("oracle") for x in rows)
        assert all(("frac_of_8000" in x) == (not x["l2_resident"]) for x in rows)
        c = d["cpu_baseline"]
        assert c["kind"] == "port" and c["cores"] > 1 and c["single_thread"]["cores"] == 1
        assert c["value"] > c["single_thread"]["value"]
    else:
        for key in ("allreduce", "nvls_allreduce"):
            a = d[key]
            assert a["bit_exact"] is True
            p = a["allreduce_parity"]
            assert p["checked_vs"] == "oracle" and p["mismatches"] == 0 and len(p["sizes_bytes"]) >= 3
            assert any(b % 16 for b in p["sizes_bytes"])                  # a ragged size is among them
            assert a["rows"][0]["bytes"] == 1024 and a["rows"][-1]["bytes"] == 1 << 30
        kernels = {x["kernel"] for x in d["nvls_allreduce"]["rows"]}
        assert "oneshot" in kernels and len(kernels) >= 2
