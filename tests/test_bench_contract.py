"""bench.py's two arms, as far as a box without a GPU can see them: the reference arm (the oracle port on the
host cores) prints one JSON line with the contract's keys on the CUDA arm's own `config`; the CUDA arm refuses
to run without a GPU instead of measuring anything else."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench(*args):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, capture_output=True,
                          text=True, timeout=600)


def test_reference_arm_line():
    r = _bench("--impl", "reference", "--steps", "1", "--warmup", "1")
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["gpu_launches"] == 0
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "scaling", "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert key in line, key
    assert (line["n_gpus"], line["steps"], line["warmup"], line["higher_is_better"]) == (1, 1, 1, True)
    assert line["unit"] == "tokens/s" and line["vs_baseline"] is None and line["data"] == "synthetic"
    assert line["value"] > 0 and abs(line["ms_per_step"] * line["value"] / 1e3 - 8) < 1e-6   # 8 sequences per step
    cb, e2e = line["cpu_baseline"], line["e2e"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == line["value"] and "of 256 sequences" in cb["sample"]
    assert (e2e["value"], e2e["unit"], e2e["h2d_bytes_per_step"], e2e["d2h_bytes_per_step"]) == (line["value"], "tokens/s", 0, 0)
    # the same `config` object as the CUDA arm's line of this workload
    sys.path.insert(0, ROOT)
    import bench
    import workloads
    import argparse
    args = argparse.Namespace(workload="c3", shard_of=1)
    assert line["config"] == json.loads(json.dumps(bench.workload_config(args, workloads.build_tables("c3"), 1)))


def test_reference_arm_under_torchrun_prints_one_line():
    """Launched like the CUDA arm at N = 2: rank 0 alone measures and prints, the other rank exits 0 without work."""
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "bench.py"),
                        "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1"],
                       cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    assert lines[0]["impl"] == "reference" and lines[0]["n_gpus"] == 2
    assert lines[0]["config"]["parallelism"] == "kv-head-shard x2"


@pytest.mark.skipif(torch.cuda.is_available(), reason="needs a box without a GPU")
def test_cuda_arm_refuses_without_a_gpu():
    r = _bench("--steps", "1", "--warmup", "1")
    assert r.returncode != 0
    assert not any(l.startswith("{") for l in r.stdout.splitlines())   # no line, not even a partial one
