"""CPU: bench.py's reference arm (`--impl reference`) and the `cpu_baseline` leg on the tiny configuration - the code path that
times the oracle port on the host cores must keep working when the oracle changes (it builds the model on the meta device)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_tiny_prints_contract_line():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--config", "tiny", "--steps", "2",
                          "--warmup", "1"], capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "train samples/s" and line["unit"] == "samples/s"
    assert line["value"] and line["value"] > 0 and line["higher_is_better"] is True
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1 and line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in line["config"]


def test_cpu_baseline_leg_tiny():
    sys.path.insert(0, ROOT)
    import argparse
    import bench
    args = argparse.Namespace(config="tiny", steps=1, warmup=0)
    r = bench.cpu_reference_step(args, budget_s=60.0, max_steps=1)
    assert r["value"] and r["value"] > 0 and r["kind"] == "port" and r["steps_timed"] == 1
