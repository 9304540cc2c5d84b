"""The measurement contract of bench.py, as far as it can be checked without a GPU:
  * the reference arm (`--impl reference`: the reference's own prover from oracle/_ref on the host cores) runs here and
    prints ONE JSON line with the contract's keys, the same `config` our arm prints, and `"impl": "reference"`;
  * the committed line of our arm (profiles/r02_bench_n1.json, written by `python bench.py` on a B200) carries every key
    the contract names, with consistent arithmetic (value = batch x steps / time, roofline.frac = achieved / peak).
"""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
             "data", "config", "e2e", "cpu_baseline"}


def _line(text):
    lines = [l for l in text.strip().splitlines() if l.startswith("{")]
    assert len(lines) == 1, text[-2000:]
    return json.loads(lines[0])


def test_reference_arm_runs_here_and_prints_the_contract_line():
    from oracle import zkref
    if not (zkref.available(mt=True) or zkref.available()):
        pytest.skip("oracle/_ref not built")
    env = dict(os.environ)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--no-cpu-throughput"], capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    j = _line(out.stdout)
    assert BASE_KEYS <= set(j) and j["impl"] == "reference"
    assert j["metric"] == "groth16_proofs_per_s" and j["unit"] == "proofs/s" and j["higher_is_better"] is True
    assert j["value"] > 0 and j["e2e"]["value"] == j["value"] and j["e2e"]["h2d_bytes_per_step"] == 0
    assert j["cpu_baseline"]["kind"] == "reference" and j["cpu_baseline"]["cores"] >= 1 and j["cpu_baseline"]["sample"]
    sys.path.insert(0, ROOT)
    import bench
    assert j["config"] == bench.common_config(1)  # both arms print the same config object


def test_committed_line_of_our_arm_has_every_contract_key():
    path = os.path.join(ROOT, "profiles", "r02_bench_n1.json")
    if not os.path.exists(path):
        pytest.skip("profiles/r02_bench_n1.json missing")
    j = json.load(open(path))
    assert BASE_KEYS | {"gpu_launches", "clocks", "roofline"} <= set(j)
    assert "workload" in j["config"] and "model" not in j["config"]
    assert j["n_gpus"] == 1 and j["warmup"] >= 3 and j["gpu_launches"] > 0
    batch = j["config"]["batch_per_gpu"]
    assert abs(j["value"] - 1e3 * batch / j["ms_per_step"]) / j["value"] < 1e-6
    e = j["e2e"]
    assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(e) and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    assert e["value"] != j["value"]
    r = j["roofline"]
    assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(r)
    assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9 and 0 < r["frac"] < 1.05
    assert abs(r["achieved"] - r["units_per_launch"] * 1360 / (r["ms_per_launch"] * 1e-3) / 1e9) / r["achieved"] < 1e-6
    c = j["cpu_baseline"]
    assert {"value", "unit", "cores", "kind", "sample"} <= set(c) and c["kind"] in ("reference", "port")
    assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(j["clocks"]) and not set(j["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    for k in ("msm_g1", "witness_map", "s22", "e2e_zklaim"):
        assert k in j and "error" not in j[k], k
