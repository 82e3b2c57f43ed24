"""The bench lines committed under profiles/ keep the driver's contract (keys, units, the tier's extra objects), and the product
sources never reach into oracle/ (the oracle is test infrastructure: tests, smoke() and bench.py's CPU legs only)."""
import glob
import json
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE = ["metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config"]


def line(name):
    with open(os.path.join(ROOT, "profiles", name)) as f:
        return json.loads(f.read().strip().splitlines()[-1])


# (r02_bench_n1_box8.json and r02_bench_n4.json are --no-extras runs -- device-resident points of the scaling box, no e2e leg)
@pytest.mark.parametrize("name,full", [("r02_final_bench_n1.json", True), ("r02_bench_n2.json", True), ("r02_bench_n8.json", True),
                                       ("r02_bench_n1_box8.json", False), ("r02_bench_n4.json", False)])
def test_own_arm_lines(name, full):
    d = line(name)
    for k in BASE + ["gpu_launches", "clocks", "roofline"] + (["e2e"] if full else []):
        assert k in d, (name, k)
    assert d["unit"] == "Gvoxels/s" and d["higher_is_better"] is True and d["dtype"] == "u8" and d["vs_baseline"] is None
    assert d["scaling"] in ("strong", "weak") and d["data"].startswith("synthetic")
    assert "workload" in d["config"] and "model" not in d["config"]
    assert d["warmup"] >= 3 and d["gpu_launches"] > 0
    assert abs(d["value"] - d["config"]["n"] ** 3 / (d["ms_per_step"] * 1e-3) / 1e9) < 1e-6 * d["value"]      # whole-job voxels / max-over-ranks time
    r = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in r, (name, k)
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    c = d["clocks"]
    assert c["sm_mhz"] > 0.9 * c["sm_max_mhz"] and not [x for x in c["reasons"] if "slowdown" in x]
    if not full:
        return
    e = d["e2e"]
    for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
        assert k in e, (name, k)
    assert e["h2d_bytes_per_step"] > 0 and e["value"] < d["value"]                                          # not the device-timed number repeated
    if d["n_gpus"] == 1 or "d2h_note" in e:
        assert e["d2h_bytes_per_step"] > 0
    else:       # the multi-GPU lines under profiles/ were measured before rg_multi_stats counted the gathered bytes: they describe them in words
        assert isinstance(e["d2h_bytes_per_step"], str) and "region voxels" in e["d2h_bytes_per_step"]
    if d["n_gpus"] == 1 and "cpu_baseline" in d:
        b = d["cpu_baseline"]
        for k in ("value", "unit", "cores", "kind", "sample"):
            assert k in b
        assert b["kind"] in ("port", "reference") and "bit-exact vs oracle" in d["config"]["check"]


def test_reference_arm_line():
    d = line("r02_final_bench_reference_arm.json")
    for k in BASE + ["impl", "cpu_baseline", "e2e"]:
        assert k in d, k
    own = line("r02_final_bench_n1.json")
    assert d["impl"] == "reference" and d["metric"] == own["metric"] and d["unit"] == own["unit"] and d["higher_is_better"] is True
    assert d["config"]["workload"] == own["config"]["workload"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["value"] == d["value"] and d["cpu_baseline"]["kind"] in ("port", "reference")


def test_product_sources_do_not_use_the_oracle():
    pkg = os.path.join(ROOT, "parallel-c-programs_b200")
    srcs = glob.glob(os.path.join(pkg, "rgb200", "*.py")) + glob.glob(os.path.join(pkg, "csrc", "*"))
    assert srcs
    for p in srcs:
        if not os.path.isfile(p) or p.endswith(".o"):
            continue
        with open(p, errors="replace") as f:
            for i, l in enumerate(f, 1):
                if "oracle" not in l:
                    continue
                # comments may name the oracle (which test compares against it); code may not include, import, open or load it
                assert not re.search(r"#\s*include|import |dlopen|CDLL|open\(|_ref", l), (p, i, l.strip())
