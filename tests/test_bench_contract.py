"""bench.py's reference arm (`--impl reference`): the JSON line the driver parses, on the smallest
workload so that it runs on the CPU box in seconds.  The GPU arm's line is checked on the box
(`profiles/bench_6M_*.json`); here only the contract of the arm that needs no device."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CMD = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "synthetic-200k-720p",
       "--steps", "2", "--warmup", "1"]


def _env(**kw):
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE")}
    env.update(kw)
    return env


def test_reference_arm_line_has_the_contract_keys():
    out = subprocess.run(CMD, cwd=ROOT, env=_env(), capture_output=True, text=True, timeout=300, check=True).stdout
    lines = [l for l in out.splitlines() if l.strip()]
    assert len(lines) == 1, out
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "frames_per_second" and d["unit"] == "frames/s"
    assert d["higher_is_better"] is True and d["vs_baseline"] is None and d["data"] == "synthetic"
    assert d["steps"] == 2 and d["warmup"] == 1 and d["n_gpus"] == 1 and d["value"] > 0
    assert d["config"]["workload"] == "synthetic-200k-720p" and d["config"]["gaussians"] == 200000
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0      # the CPU arm launches nothing: the b200 arm's count is the claim


def test_reference_arm_other_ranks_exit_without_work():
    r = subprocess.run(CMD, cwd=ROOT, env=_env(RANK="1", LOCAL_RANK="1", WORLD_SIZE="2"), capture_output=True,
                       text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""
