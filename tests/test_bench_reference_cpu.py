"""The reference arm of bench.py (`--impl reference`: the CPU restatement of the hot path on the host cores) prints ONE JSON
line with the keys the driver's contract names; under torchrun only rank 0 prints. Runs on a small sample, no GPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(env=None):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--cpu-rows", "300000", "--steps", "2",
                        "--warmup", "1"], capture_output=True, text=True, timeout=600, env=e)
    assert r.returncode == 0, r.stderr[-2000:]
    return [line for line in r.stdout.splitlines() if line.strip()]


def test_reference_arm_prints_the_contract_line():
    (line,) = run()
    d = json.loads(line)
    assert d["impl"] == "reference" and d["unit"] == "rows/s" and d["higher_is_better"] is True
    assert d["steps"] == 2 and d["warmup"] == 1 and d["n_gpus"] == 1 and d["gpu_launches"] == 0
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["data"] == "synthetic" and d["vs_baseline"] is None
    assert "workload" in d["config"] and d["config"]["rows"] == 300000
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["passes"] == 2 and cb["result_rows"] == 4
    assert d["e2e"] == {"value": d["value"], "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_stay_silent():
    assert run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}) == []


def test_our_arm_fails_loudly_without_a_gpu():
    """No CPU fallback: on a box without a CUDA device the product arm exits non-zero and prints no benchmark line."""
    import pytest
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "0"], capture_output=True,
                       text=True, timeout=600)
    assert r.returncode != 0
    assert not any(line.lstrip().startswith("{") for line in r.stdout.splitlines())
