"""CPU: the bench contract.  The reference arm runs without a GPU (it is the oracle port on the host cores) and prints
the driver's JSON line; the record of the round's final GPU run kept under profiles/ carries every key and block the
contract names, on the same workload string as the reference arm."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "e2e", "cpu_baseline"}


def last_json_line(text):
    lines = [ln for ln in text.splitlines() if ln.startswith("{")]
    assert lines, text[-2000:]
    return json.loads(lines[-1])


def test_reference_arm_prints_the_contract_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--n-tx", "20000"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-3000:]
    d = last_json_line(r.stdout)
    assert BASE_KEYS <= set(d), sorted(BASE_KEYS - set(d))
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert abs(d["e2e"]["value"] - d["value"]) <= 1e-6 * d["value"]
    kept = json.load(open(os.path.join(ROOT, "profiles", "r2z_bench.json")))
    assert d["metric"] == kept["metric"] and d["unit"] == kept["unit"]
    assert d["config"]["workload"] == kept["config"]["workload"]          # same_config for the driver


def test_kept_gpu_record_has_every_block():
    d = json.load(open(os.path.join(ROOT, "profiles", "r2z_bench.json")))
    assert BASE_KEYS | {"clocks", "gpu_launches", "roofline"} <= set(d)
    assert d["gpu_launches"] > 0 and d["vs_baseline"] is None and d["data"] == "synthetic"
    assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(d["roofline"])
    assert abs(d["roofline"]["frac"] - d["roofline"]["achieved"] / d["roofline"]["peak"]) < 1e-9
    assert {"value", "unit", "cores", "kind", "sample"} <= set(d["cpu_baseline"])
    assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(d["e2e"])
    assert d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0
    assert d["e2e"]["value"] < d["value"]                                  # copies inside the timed region cost something
    assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(d["clocks"])
    for block in ("path_a", "c3", "live_loop", "ltv"):
        assert block in d, block
    for block in ("path_a", "c3", "ltv"):
        assert {"bound", "achieved", "peak", "frac"} <= set(d[block]["roofline"]), block
        assert "e2e" in d[block], block
    assert {"max", "p99", "p50"} <= set(d["ltv"]["error_vs_oracle"])
    n8 = json.load(open(os.path.join(ROOT, "profiles", "r2z_bench_n8.json")))
    assert n8["n_gpus"] == 8 and "c5" in n8 and n8["scaling"] == "weak"
