"""bench.py host logic that needs no GPU: the CPU reference arm's line and the clock sampler's failure mode."""
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_reference_arm_prints_exactly_one_json_line():
    env = dict(os.environ, OMP_NUM_THREADS=str(min(4, os.cpu_count() or 1)))
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1", "--no-pairs"],
                       capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [x for x in r.stdout.splitlines() if x.strip()]
    assert len(lines) == 1, r.stdout
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["unit"] == "TFLOP/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["steps"] == 1 and line["warmup"] >= 3
    assert line["config"]["workload"] == "attn_plugin_op_B1_H4_N1024_M1024_d64_f16in_f16out"
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_do_no_work():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--no-pairs"],
                       capture_output=True, text=True, timeout=120, env=env, cwd=ROOT)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_clock_sampler_reports_missing_nvml_instead_of_raising():
    sys.path.insert(0, str(ROOT))
    import bench

    s = bench.ClockSampler("0000:ff:1f.0")      # no such device (and no driver in the CPU container)
    s.start()
    s.mark()
    out = s.stop()
    assert out["sm_mhz"] is None and out["samples"] == 0 and out["reasons"] and "nvml" in out["reasons"][0]
