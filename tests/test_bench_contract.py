"""The bench.py JSON-line contract, checked on the committed lines of the last GPU run
(profiles/*_bench_*.json) and on bench.py's argument surface.  No GPU needed."""
import glob
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _lines(pattern):
    out = []
    for p in sorted(glob.glob(os.path.join(ROOT, "profiles", pattern))):
        txt = open(p).read().strip().splitlines()
        if txt:
            out.append((os.path.basename(p), json.loads(txt[-1])))
    return out


def test_own_arm_lines_carry_the_contract_keys():
    lines = [l for l in _lines("r01_*_bench_full.json") + _lines("r01_*_bench_*n8.json")]
    assert lines
    for name, d in lines[-3:]:
        for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                  "scaling", "vs_baseline", "dtype", "data", "config", "roofline", "e2e", "gpu_launches", "clocks"):
            assert k in d, (name, k)
        assert d["metric"] == "mfcc_frames_per_sec" and d["unit"] == "frames/s" and d["higher_is_better"] is True
        assert d["scaling"] == "weak" and d["vs_baseline"] is None and d["data"] == "synthetic"
        assert "workload" in d["config"] and "model" not in d["config"]
        r = d["roofline"]
        assert r["bound"] in ("hbm", "tensor") and r["unit"] == "GB/s"
        assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
        e = d["e2e"]
        assert e["unit"] == d["unit"] and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
        assert e["value"] < d["value"]                          # host copies are inside the e2e region
        assert d["gpu_launches"] == d["steps"]                  # one fused launch per step
        assert d["warmup"] >= 3
        assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}
        if d["n_gpus"] == 1:
            c = d["cpu_baseline"]
            assert c["kind"] in ("reference", "port") and c["cores"] >= 1 and c["value"] > 0 and c["sample"]


def test_round2_lines_carry_copy_ceiling_c4_and_baseline_flags():
    """Round 2: e2e next to its copy-only ceiling at every N, BASELINE configs[3] (C4, strong scaling) at
    every N, the CPU baseline's build flags and the oracle spot check of the timed launch."""
    lines = _lines("r02_*_bench_full.json") + _lines("r02_[c-z]_bench_n[248].json")
    assert len(lines) >= 4
    for name, d in lines:
        cc = d["e2e"]["copy_ceiling"]
        assert cc["unit"] == "frames/s" and cc["value"] > 0
        assert abs(cc["e2e_over_ceiling"] - d["e2e"]["value"] / cc["value"]) < 1e-9
        assert cc["e2e_over_ceiling"] >= 0.95, name                 # e2e >= 0.95 x the copy-only ceiling
        c4 = d["c4"]
        assert c4["scaling"] == "strong" and c4["unit"] == "frames/s" and c4["value"] > 0
        assert abs(c4["value"] - 16667 * 5645 / (c4["ms_per_pass"] * 1e-3)) < 1e-3 * c4["value"]
        if d["n_gpus"] == 1:
            c = d["cpu_baseline"]
            assert "-O3" in c["build_flags"] and "-march=native" in c["build_flags"]
            assert c["checker_build"]["value"] > 0
            if "spot_check" in c:
                assert c["spot_check"]["bit_exact"] is True
    ref = _lines("r02_*_bench_reference.json")
    assert ref and "-march=native" in ref[-1][1]["cpu_baseline"]["build_flags"]


def test_reference_arm_line():
    lines = _lines("r01_*_bench_reference.json")
    assert lines
    name, d = lines[-1]
    assert d["impl"] == "reference" and d["metric"] == "mfcc_frames_per_sec" and d["unit"] == "frames/s"
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["value"] == d["value"] and d["cpu_baseline"]["kind"] == "port"


def test_bench_cli_surface():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--help"], capture_output=True, text=True)
    assert out.returncode == 0
    for flag in ("--gpus", "--steps", "--warmup", "--impl", "--e2e-sweep", "--no-c4"):
        assert flag in out.stdout
