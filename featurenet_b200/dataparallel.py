"""Data-parallel training of ONE large candidate over the GPUs of a box (NCCL over NVLink).

The reference's answer to "one model, several GPUs" is ``keras.utils.multi_gpu_model``
(/root/reference/model/keras_model.py:137-146): the batch is cut into one slice per GPU, every
tower runs on its slice with its own BatchNorm statistics, and the variables are shared.  The
equivalent here, one process per GPU (torchrun), lives INSIDE the ordinary engine -- ``fit``,
``evaluate``, callbacks and the best-checkpoint restore of ``helpers.train_model`` are the same code:

* every rank holds a replica (rank 0's initial weights are broadcast) and the whole dataset resident
  in its HBM; rank 0's epoch permutation is broadcast, so all ranks walk the same global batches;
* a global batch of G samples is cut into contiguous slices (the remainder goes to the first ranks);
  each rank runs forward, loss and backward on its slice with the loss scaled by 1/G, so the SUM
  of the per-rank gradients is the gradient of the global-batch mean;
* ONE ``all_reduce(SUM)`` over the flat ``grads`` arena (4 bytes x parameters: 21 MB for the
  ``keras`` template, <= 80 MB at the 20 M-parameter cap), issued between the captured backward pass and the
  captured optimizer launch -- on NVSwitch every GPU reaches every peer at full bandwidth, so a
  single large message is the bandwidth-optimal shape;
* the ordinary fused Adam launch on every rank (replicas stay bit-identical because the
  all-reduce result is), which also advances the epoch cursor by the GLOBAL batch;
* validation and ``evaluate`` are sharded by sample: each rank scores a contiguous share of the set and one
  all-reduce of (loss sum, correct count) gives every rank the same numbers, so callbacks
  (EarlyStopping, ReduceLROnPlateau, ModelCheckpoint) take the same decisions everywhere;
* BatchNorm moving statistics are per-rank during an epoch (as the towers' batch statistics are in the
  reference) and averaged over the ranks before every validation pass, so the evaluated model is one model.

Enabled by ``FNET_DP=1`` under torchrun (or ``Network.enable_data_parallel``).  Population training does
NOT use this (candidates are sharded whole, featurenet_b200.population); it is for the
single-large-candidate configuration only (BASELINE configs[4]).
"""
import os

import numpy as np


def split_batch(global_batch, world):
    """Rows per rank for one global batch: contiguous slices, remainder to the first ranks."""
    q, rem = divmod(int(global_batch), int(world))
    sizes = [q + (1 if r < rem else 0) for r in range(world)]
    starts = [sum(sizes[:r]) for r in range(world)]
    return sizes, starts


def reduce_metrics(dist, loss_sum, correct, device=None):
    """(sum of loss sums, sum of correct counts) over the ranks: one all-reduce of two doubles (counts are exact
    below 2^53)."""
    import torch
    packed = torch.tensor([float(loss_sum), float(correct)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(packed, op=dist.ReduceOp.SUM)
    return float(packed[0].item()), int(round(float(packed[1].item())))


def wanted():
    """True when this process was asked to train its candidates data-parallel (FNET_DP=1 under torchrun)."""
    if os.environ.get("FNET_DP", "0") != "1":
        return False
    try:
        import torch.distributed as dist
    except ImportError:
        return False
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


class DataParallel(object):
    """The distributed half of an Engine: who we are, and the three collectives of the mode."""

    def __init__(self, dist, rank=None, world=None):
        self.dist = dist
        self.rank = dist.get_rank() if rank is None else int(rank)
        self.world = dist.get_world_size() if world is None else int(world)
        self.allreduce_bytes = 0
        self.allreduce_calls = 0

    def slice_of(self, global_batch):
        sizes, starts = split_batch(global_batch, self.world)
        return sizes[self.rank], starts[self.rank]

    def shard_of(self, n):
        """This rank's contiguous share [lo, hi) of an n-row evaluation set."""
        sizes, starts = split_batch(n, self.world)
        return starts[self.rank], starts[self.rank] + sizes[self.rank]

    def broadcast_order(self, order, device):
        import torch
        t = torch.from_numpy(np.ascontiguousarray(order, dtype=np.int32)).to(device)
        self.dist.broadcast(t, src=0)
        return t


def multi_gpu_model(network, gpus=None):
    """keras.utils.multi_gpu_model for the engine: marks the network so that its engine trains data-parallel over the
    process group.  Raises when there is no group (the reference's branch catches that and carries on single-GPU)."""
    if not wanted():
        raise RuntimeError("multi gpu not available: set FNET_DP=1 and launch one process per GPU with torchrun")
    network.data_parallel = True
    return network
