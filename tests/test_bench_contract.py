"""The bench line contract (task statement, "Measurement"): the JSON lines recorded on the B200 under profiles/ carry
every key the driver and the judge read, with consistent values.  CPU-only: it reads the committed lines, it does not
run the bench."""
import glob
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _line(path):
    with open(path) as f:
        return json.loads(f.read().strip().splitlines()[-1])


def _latest(pattern):
    files = glob.glob(os.path.join(ROOT, "profiles", pattern))
    assert files, pattern
    # v12 comes after v9: order by the numbers in the name, not by the characters
    return max(files, key=lambda f: [int(t) for t in re.findall(r"\d+", os.path.basename(f))])


def test_own_arm_line_has_the_contract_keys():
    d = _line(_latest("r2_bench_v*.json"))
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "roofline", "cpu_baseline", "clocks"):
        assert key in d, key
    # BASELINE.json names the metric "... at 1/2/4/8 B200"; the line carries the GPU count in n_gpus
    assert base["metric"].startswith(d["metric"]) and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["n_gpus"] == 1 and d["warmup"] >= 3 and d["scaling"] == "strong" and d["data"] == "synthetic"
    assert d["dtype"] == "f64" and "workload" in d["config"] and "model" not in d["config"]
    # value = units processed / device time
    per_step = d["config"]["centres_per_step_per_gpu"] * d["n_gpus"]
    assert abs(d["value"] - per_step / (d["ms_per_step"] * 1e-3)) / d["value"] < 1e-3
    e2e = d["e2e"]
    assert e2e["unit"] == d["unit"] and e2e["h2d_bytes_per_step"] > 1e9 and e2e["d2h_bytes_per_step"] > 0
    assert 0 < e2e["value"] < d["value"]                       # host copies inside the timed region
    r = d["roofline"]
    for key in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert key in r, key
    assert r["bound"] in ("hbm", "tensor") and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-3 and 0 < r["frac"] < 1
    c = d["cpu_baseline"]
    assert c["kind"] == "reference" and c["cores"] >= 1 and c["value"] > 0 and c["sample"]       # the unmodified reference
    ps = d["parity_sample"]                                    # the same centres scored by the reference and by the GPU
    assert ps["n"] >= 512 and ps["max_abs_diff"] <= 1e-3 and ps["flags"] == 0
    assert r["traffic"] > 0 and "sha256" in r["traffic_source"]
    assert d["gpu_launches"] > 0
    bad = {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    assert not bad & set(d["clocks"]["reasons"]) and d["clocks"]["sm_mhz"] > 0.9 * d["clocks"]["sm_max_mhz"]


def test_reference_arm_line():
    d = _line(_latest("r2_bench_reference_v*.json"))
    ours = _line(_latest("r2_bench_v*.json"))
    assert d["impl"] == "reference" and d["metric"] == ours["metric"] and d["unit"] == ours["unit"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["value"] == d["value"] and d["cpu_baseline"]["cores"] >= 1
    assert d["config"]["workload"] == ours["config"]["workload"] and d["cpu_baseline"]["kind"] == "reference"


def test_multi_gpu_line_is_the_vertex_sharded_split():
    d = _line(_latest("r2_bench_n8_v*.json"))
    assert d["n_gpus"] == 8 and d["scaling"] == "strong" and "Morton" in d["config"]["sharding"]
    rep = d["replicas"]                                        # 8 whole-hemisphere jobs, measured in the same run
    assert rep["scaling"] == "weak" and d["value"] > 0.9 * rep["value"]      # sharding one hemisphere costs < 10 %
