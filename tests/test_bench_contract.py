"""CPU: the parts of bench.py's contract that need no GPU — the reference arm
prints exactly one JSON line with the agreed keys, and the product arm refuses
to run (loudly, non-zero) on a machine without CUDA instead of falling back."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, timeout=600):
    env = dict(os.environ)
    env.pop("RANK", None)
    env.pop("WORLD_SIZE", None)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args],
                          capture_output=True, text=True, cwd=ROOT, env=env, timeout=timeout)


def test_reference_arm_prints_one_contract_line():
    out = _run("--impl", "reference", "--steps", "1", "--warmup", "1",
               "--workload", "gpt2-small-124M-adamw-fp32", "--cpu-sample-gb", "0.2")
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, out.stdout[-2000:]
    d = json.loads(lines[0])
    assert d["impl"] == "reference"
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step",
                "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config",
                "cpu_baseline", "e2e"):
        assert key in d, key
    assert d["metric"] == "flash_ckpt_save_plus_restore_GBps" and d["unit"] == "GB/s"
    assert d["value"] > 0 and d["higher_is_better"] is True and d["dtype"] == "u8"
    assert d["config"]["workload"] == "gpt2-small-124M-adamw-fp32"
    cb = d["cpu_baseline"]
    # the reference's own Python (shipped to baseline/_ref by build()), the
    # oracle port only where that is missing
    assert cb["kind"] in ("reference", "port")
    assert cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["config"]["same_config"] is True and d["config"]["ranks"] == 1
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"],
                        "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_product_arm_fails_loudly_without_cuda():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present")
    out = _run("--steps", "1", "--warmup", "1", timeout=300)
    assert out.returncode != 0
    assert not [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert "CUDA" in out.stderr or "cuda" in out.stderr


def test_reference_arm_under_torchrun_runs_on_every_rank():
    """N > 1: the driver launches the reference arm with torchrun like the
    product arm; every rank runs the reference's path on its own shard at the
    same time (gloo for the barriers) and rank 0 prints the one line."""
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    out = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
         "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "bench.py"),
         "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1",
         "--workload", "gpt2-small-124M-adamw-fp32"],
        capture_output=True, text=True, cwd=ROOT, env=env, timeout=900)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, out.stdout[-2000:]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["config"]["ranks"] == 2
    assert d["config"]["same_config"] is True and d["value"] > 0
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
