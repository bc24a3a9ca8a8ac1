"""Multi-GPU correctness over NCCL (-m gpu; skipped on a box with one GPU): tools/nccl_parity.py under torchrun."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_nccl_sharded_step_equals_sequential():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu)")
    world = 2 if n < 4 else 4
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", "29531", os.path.join(ROOT, "tools", "nccl_parity.py")],
                       capture_output=True, text=True, timeout=900, cwd=ROOT)
    print(r.stdout[-3000:], r.stderr[-2000:])
    assert r.returncode == 0 and "nccl parity ok" in r.stdout
