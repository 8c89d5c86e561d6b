"""Multi-GPU parity (needs >= 2 visible GPUs; skipped otherwise): the train set
sharded over GPUs gives bit-identical results to one GPU, (a) one process per GPU
over NCCL (torchrun), (b) inside the C ABI with SPECTAVI_GPUS=N (one process)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "mgpu_worker.py")


def _ngpu():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_one_process_per_gpu(world):
    if _ngpu() < world:
        pytest.skip("needs %d GPUs" % world)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
                          "--master-addr", "127.0.0.1", "--master-port", "29611", WORKER, "ranks"],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "MGPU_RANKS_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]


@pytest.mark.parametrize("gpus", [2, 8])
def test_single_process_c_abi(gpus):
    if _ngpu() < gpus:
        pytest.skip("needs %d GPUs" % gpus)
    env = dict(os.environ, SPECTAVI_GPUS=str(gpus))
    out = subprocess.run([sys.executable, WORKER, "abi"], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0 and "MGPU_ABI_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]
