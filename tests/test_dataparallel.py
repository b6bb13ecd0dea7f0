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


def _gloo_worker(rank, world, port, queue):
    import torch.distributed as dist
    from featurenet_b200.dataparallel import DataParallel, reduce_metrics
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        dp = DataParallel(dist)
        local, start = dp.slice_of(81)
        lo, hi = dp.shard_of(1101)
        # every rank contributes what it scored on its shard; everybody ends with the same totals
        totals = reduce_metrics(dist, 0.5 * (hi - lo), hi - lo)
        import numpy as np
        order = dp.broadcast_order(np.arange(10)[::-1] if rank == 0 else np.zeros(10), "cpu")
        queue.put((rank, local, start, lo, hi, totals, order.tolist()))
    finally:
        dist.destroy_process_group()


def test_slices_shards_and_metric_reduction_under_gloo_world_2():
    """The host logic of the data-parallel mode with two CPU ranks: slices of a global batch, shards of an evaluation
    set, the one all-reduce of (loss sum, correct count), rank 0's permutation reaching everybody."""
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as sock:
        sock.bind(("127.0.0.1", 0))
        port = sock.getsockname()[1]
    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, queue)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(queue.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, l0, s0, lo0, hi0, t0, o0), (_, l1, s1, lo1, hi1, t1, o1) = got
    assert (l0, s0, l1, s1) == (41, 0, 40, 41)
    assert (lo0, hi0, lo1, hi1) == (0, 551, 551, 1101)
    assert t0 == t1 == (0.5 * 1101, 1101)
    assert o0 == o1 == list(range(9, -1, -1))


@pytest.mark.gpu
def test_two_gpu_data_parallel_equals_single_gpu():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "dp_check.py")],
                         capture_output=True, text=True, timeout=600)
    assert "DP_CHECK_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
