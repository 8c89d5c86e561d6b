"""bench.py's reference arm and JSON contract, on CPU (no GPU needed): the --impl reference line must carry the
metric / unit / config keys of the b200 arm, and the GPU arm must refuse to run without a device instead of
falling back to anything."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, env=None):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                          timeout=600, env=env)


def test_reference_arm_line():
    out = _run("--impl", "reference", "--steps", "1", "--warmup", "0", "--workload", "4096x2048x128")
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "Gpairs/s" and line["higher_is_better"] is True
    assert line["metric"].startswith("G L1 descriptor-pair comparisons/sec")
    assert line["value"] > 0 and line["gpu_launches"] == 0
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in line["config"]


def test_reference_arm_other_ranks_are_silent():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = _run("--impl", "reference", "--steps", "1", "--warmup", "0", "--gpus", "2", env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_gpu_arm_has_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present")
    out = _run("--steps", "1", "--warmup", "0", "--workload", "256x256x128")
    assert out.returncode != 0
    assert "no CUDA device" in (out.stderr + out.stdout)
