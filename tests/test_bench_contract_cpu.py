"""bench.py's contract, as far as it can be checked without a GPU: the reference arm (the reference's own CPU path
on the host cores) prints ONE JSON line with the agreed keys, only rank 0 runs it, and our arm refuses to run
without a CUDA device instead of falling back to anything."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BENCH = os.path.join(ROOT, "bench.py")


def run(args, **env):
    e = dict(os.environ, **{k: str(v) for k, v in env.items()})
    return subprocess.run([sys.executable, BENCH] + args, capture_output=True, text=True, timeout=600, env=e, cwd=ROOT)


def test_reference_arm_prints_one_contract_line():
    r = run(["--impl", "reference", "--gpus", "1", "--steps", "1", "--warmup", "0"])
    assert r.returncode == 0, r.stderr[-500:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1
    assert d["unit"] == "G interactions/s" and d["value"] > 0 and d["vs_baseline"] is None and d["data"] == "synthetic"
    assert d["config"]["workload"].startswith("N=1048576 bodies") and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and "CPU_compute" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0


def test_reference_arm_runs_on_rank_zero_only():
    r = run(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"], RANK=1, LOCAL_RANK=1, WORLD_SIZE=2)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_our_arm_needs_a_gpu_and_says_so():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: the arm runs (covered by the round-end bench)")
    r = run(["--steps", "1", "--warmup", "0"])
    assert r.returncode != 0 and "no CPU fallback" in r.stderr and not [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
