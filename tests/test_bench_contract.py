"""bench.py prints ONE JSON line with the contract's keys (reference arm on CPU here; B200 arm under -m gpu)."""
import json
import os
import subprocess
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COMMON = ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
          "data", "config", "e2e", "gpu_launches")


def _run(args, timeout):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, out.stdout[-2000:]
    return json.loads(lines[0])


def test_reference_arm_line():
    d = _run(["--impl", "reference", "--steps", "1", "--warmup", "1", "--workload", "large_pyramid"], 600)
    for k in COMMON + ("impl", "cpu_baseline"):
        assert k in d, k
    assert d["impl"] == "reference" and d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0 and d["e2e"]["value"] == d["value"]
    assert d["config"]["workload"] == "large_pyramid" and d["vs_baseline"] is None


@pytest.mark.gpu
def test_b200_arm_line():
    d = _run(["--steps", "8", "--warmup", "3", "--cpu-steps", "3", "--worlds", "256"], 900)
    for k in COMMON + ("roofline", "cpu_baseline", "clocks"):
        assert k in d, k
    r = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in r, k
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert d["config"]["workload"] == "many_pyramids" and d["dtype"] == "f32" and d["higher_is_better"] is True
    assert d["e2e"]["d2h_bytes_per_step"] > 0 and d["gpu_launches"] >= 3 * d["steps"]
    assert set(("sm_mhz", "sm_max_mhz", "reasons")) <= set(d["clocks"])
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] == 1
    b = d["batched"]                               # BASELINE.json configs[4] beside the headline
    assert b["worlds"] == 256 and b["worlds_per_gpu"] == 256 and b["scaling"] == "strong" and b["world_steps_per_s"] > 0
    assert b["stats"]["awake_bodies"] == 256 * 55
    assert d["roofline"]["traffic"] is None or d["roofline"]["traffic"] > 0
