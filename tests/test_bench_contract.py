"""CPU: the bench line contract.  The committed B200 lines under profiles/ carry every key the driver reads, and bench.py's
command line still parses the way the driver calls it (no GPU needed for either)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
             "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"}


def _line(name):
    with open(os.path.join(ROOT, "profiles", name)) as fh:
        return json.loads(fh.read().strip().splitlines()[-1])


@pytest.mark.parametrize("name,n", [("r01_bench_cfg2_n1_run3.json", 1), ("r01_bench_cfg2_n2.json", 2), ("r01_bench_cfg2_n8.json", 8)])
def test_committed_bench_lines_carry_the_contract(name, n):
    d = _line(name)
    assert BASE_KEYS <= set(d), sorted(BASE_KEYS - set(d))
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert d["metric"].startswith("IBF bin-filter reads/s") and d["unit"] == "reads/s" and "reads/s" in json.dumps(base)
    assert d["n_gpus"] == n and d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["warmup"] >= 3 and d["data"] == "synthetic" and "workload" in d["config"] and "model" not in d["config"]
    assert d["gpu_launches"] > 0 and d["value"] > 0
    assert abs(d["value"] - n * d["config"]["reads_per_gpu"] / (d["ms_per_step"] / 1e3)) / d["value"] < 1e-6
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9 and 0 < r["frac"] < 1
    e = d["e2e"]
    assert e["unit"] == "reads/s" and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and e["value"] != d["value"]
    assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    if n == 1:
        c = d["cpu_baseline"]
        assert c["kind"] in ("reference", "port") and c["cores"] >= 1 and c["value"] > 0 and c["sample"]
        assert c.get("rows_equal_gpu") is True


def test_reference_arm_line():
    d = _line("r01_bench_cfg2_reference.json")
    assert d["impl"] == "reference" and d["unit"] == "reads/s" and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0


def test_bench_command_line_parses():
    for extra in ([], ["--impl", "reference"], ["--workload", "cfg3"], ["--workload", "cfg4"], ["--workload", "cfg5"]):
        r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--gpus", "1", "--steps", "2", "--warmup", "3", "--help"] + extra,
                           capture_output=True, text=True, timeout=120)
        assert r.returncode == 0 and "--dropin-batch" in r.stdout
