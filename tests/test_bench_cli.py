"""bench.py's reference arm (the CPU restatement timed on the host cores) runs without a GPU and prints one
JSON line with the contract's keys."""
import json
import os
import subprocess
import sys

from conftest import ROOT


def test_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--mesh", "nsbench2", "--dim", "2",
                          "--refines", "1", "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["metric"] == "time_steps_per_s" and line["unit"] == "steps/s"
    assert line["higher_is_better"] is True and line["value"] > 0 and line["steps"] == 2 and line["warmup"] == 1
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert line["config"]["workload"].startswith("nsbench2 r1") and line["dtype"] == "f64"
    # the arm maps the host-only set-up library and the oracle, not the CUDA library
    assert "libstep35a_host.so" in line["native_libraries"] and "libstep35a_b200.so" not in line["native_libraries"]
    assert set(line["config"]) == {"workload", "dof", "n_vel_per_component", "n_pres", "cells", "preconditioner", "switches",
                                   "parallelism", "l2"}


def test_reference_arm_is_silent_on_other_ranks():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--mesh", "nsbench2",
                          "--dim", "2", "--refines", "0", "--steps", "1", "--warmup", "0"], capture_output=True, text=True,
                         timeout=120, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""
