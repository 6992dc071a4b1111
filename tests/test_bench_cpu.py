"""bench.py's contract where no GPU is needed: the reference arm (`--impl reference`) prints ONE JSON line with the
keys the driver reads, and the CUDA arm fails loudly (no CPU fallback, nothing printed as a result) without a GPU."""
import json
import os
import subprocess
import sys

import pytest

import common

ROOT = common.ROOT


def _run(args, timeout=600):
    env = dict(os.environ)
    env.pop("RANK", None)
    env.pop("WORLD_SIZE", None)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, cwd=ROOT, env=env, capture_output=True,
                          text=True, timeout=timeout)


def test_reference_arm_line():
    import refharness
    if not refharness.available():
        pytest.skip("oracle/_ref not built")
    r = _run(["--impl", "reference", "--workload", "c1", "--steps", "2", "--warmup", "1"])
    assert r.returncode == 0, r.stderr[-800:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "Msamples/s" and d["unit"] == "Msamples/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] >= 1
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["vs_baseline"] is None
    assert d["config"]["x_res"] == 256 and d["config"]["y_res"] == 256
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    e = d["e2e"]
    assert e["value"] == d["value"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0


def test_reference_arm_other_ranks_exit_quietly():
    """under torchrun only rank 0 runs the reference arm; the other ranks print nothing and exit 0"""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="29599")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--workload", "c1",
                        "--steps", "1", "--warmup", "0"], cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-800:]
    assert r.stdout.strip() == ""


def test_cuda_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = _run(["--workload", "c1", "--steps", "1", "--warmup", "0"], timeout=300)
    assert r.returncode != 0
    assert r.stdout.strip() == ""  # no result line from a machine that cannot run the CUDA path
