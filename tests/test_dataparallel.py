"""Data-parallel mode: slice arithmetic on CPU; the NCCL equivalence check needs two GPUs (torchrun)."""
import os
import subprocess
import sys

import pytest

from featurenet_b200.dataparallel import split_batch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_split_batch_covers_the_global_batch():
    assert split_batch(128, 8) == ([16] * 8, [16 * r for r in range(8)])
    sizes, starts = split_batch(80, 3)
    assert sizes == [27, 27, 26] and starts == [0, 27, 54]
    sizes, starts = split_batch(3, 8)                       # ragged last batch smaller than the world
    assert sizes == [1, 1, 1, 0, 0, 0, 0, 0] and sum(sizes) == 3
    assert split_batch(0, 2) == ([0, 0], [0, 0])


@pytest.mark.gpu
def test_two_gpu_data_parallel_equals_single_gpu():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "dp_check.py")],
                         capture_output=True, text=True, timeout=600)
    assert "DP_CHECK_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
