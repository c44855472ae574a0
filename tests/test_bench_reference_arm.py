"""The reference arm of bench.py (`--impl reference`) runs on host cores only: check, at a size that takes seconds, that it
prints the one JSON line the driver reads, with the keys the contract names, and that the CUDA arm refuses to run without a GPU
instead of falling back to anything."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, timeout=600):
    env = dict(os.environ, OMP_NUM_THREADS="1")
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, cwd=ROOT, env=env, timeout=timeout,
                          stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)


def test_reference_arm_prints_the_contract_line():
    p = _run(["--impl", "reference", "--n", "384", "--steps", "2", "--warmup", "1", "--ref-budget", "1"])
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    r = json.loads(lines[0])
    assert r["impl"] == "reference" and r["metric"] == "gibbs_sweeps_per_sec" and r["unit"] == "sweeps/s"
    assert r["higher_is_better"] is True and r["n_gpus"] == 1 and r["steps"] == 2 and r["warmup"] == 1
    assert r["value"] > 0 and r["ms_per_step"] > 0
    assert r["e2e"] == {"value": r["value"], "unit": r["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = r["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == r["value"] and cb["sample"]
    assert r["config"]["workload"] and "model" not in r["config"]
    for k, w in r.get("workloads", {}).items():
        assert w["value"] > 0, k


def test_reference_arm_of_the_ais_workload():
    p = _run(["--impl", "reference", "--workload", "ais", "--ais-n", "256", "--ais-particles", "2", "--steps", "1",
              "--warmup", "1", "--ref-budget", "1"])
    assert p.returncode == 0, p.stderr[-2000:]
    r = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][0])
    assert r["impl"] == "reference" and r["unit"] == "particle-sweeps/s" and r["value"] > 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_cuda_arm_fails_loudly_without_a_gpu():
    p = _run(["--n", "256", "--steps", "1", "--warmup", "3", "--no-cpu-baseline"])
    assert p.returncode != 0
    assert not [l for l in p.stdout.splitlines() if l.startswith("{")]
