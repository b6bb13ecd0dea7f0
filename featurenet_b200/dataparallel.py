"""Data-parallel training of ONE large candidate over the GPUs of a box (NCCL over NVLink).

The reference's answer to "one model, several GPUs" is ``keras.utils.multi_gpu_model``
(/root/reference/model/keras_model.py:137-146): the batch is cut into one slice per GPU, every
tower runs on its slice with its own BatchNorm statistics, and the variables are shared.  The
equivalent here, one process per GPU (torchrun):

* every rank holds a replica (identical initial weights, identical permutation) and the whole
  dataset resident in its HBM;
* a global batch of G samples is cut into contiguous slices (the remainder goes to the first ranks);
  each rank runs forward, loss and backward on its slice with the loss scaled by 1/G, so the SUM
  of the per-rank gradients is the gradient of the global-batch mean;
* ONE ``all_reduce(SUM)`` over the flat ``grads`` arena (4 bytes x parameters: 21 MB for the
  ``keras`` template, <= 80 MB at the 20 M-parameter cap) -- on NVSwitch every GPU reaches every
  peer at full bandwidth, so a single large message is the bandwidth-optimal shape;
* the ordinary fused Adam launch on every rank (replicas stay bit-identical because the
  all-reduce result is).

Population training does NOT use this (candidates are sharded whole, featurenet_b200.population);
it is for the single-large-candidate configuration only.
"""
import ctypes as C

import numpy as np
import torch


def split_batch(global_batch, world):
    """Rows per rank for one global batch: contiguous slices, remainder to the first ranks."""
    q, rem = divmod(int(global_batch), int(world))
    sizes = [q + (1 if r < rem else 0) for r in range(world)]
    starts = [sum(sizes[:r]) for r in range(world)]
    return sizes, starts


class DataParallelTrainer(object):
    def __init__(self, net, dist, rank, world):
        if net.optimizer is None:
            raise RuntimeError("compile() the model first")
        self.net, self.dist, self.rank, self.world = net, dist, rank, world
        self.eng = net._need_engine()
        self.graphs = {}
        # replicas must start identical: rank 0's weights win
        with torch.cuda.stream(self.eng.stream):
            dist.broadcast(self.eng.params, src=0)
            dist.broadcast(self.eng.stats, src=0)
        self.eng.stream.synchronize()

    def _programs(self, local, global_batch):
        """(forward+loss+backward, optimizer+bookkeeping) launch lists for a slice of `local` rows."""
        eng, L, sp = self.eng, self.eng.lib, self.eng.sp
        st = C.c_void_p(eng.state.data_ptr())
        opt = self.net.optimizer
        b1, b2 = (opt.beta_1, opt.beta_2) if opt.kind == "adam" else (0.9, 0.999)
        xdev, ydev, perm = eng.data["train"]
        row = int(np.prod(eng.shape[0]))
        first = [("step_begin", L.fnet_step_begin, (st, b1, b2, sp))]
        if local > 0:
            first.append(("gather_rows", L.fnet_gather_rows, (local, row, C.c_void_p(perm.data_ptr()), st,
                                                              C.c_void_p(xdev.data_ptr()), C.c_void_p(eng.act[0].data_ptr()), sp)))
            first.append(("gather_rows", L.fnet_gather_rows, (local, eng.n_out, C.c_void_p(perm.data_ptr()), st,
                                                              C.c_void_p(ydev.data_ptr()), C.c_void_p(eng.yb.data_ptr()), sp)))
            first += eng._forward_program(local, True)
            first.append(eng._loss_op(local, True, scale_batch=global_batch))
            first += eng._backward_program(local)
        else:
            first.append(("fill", L.fnet_fill, (eng.n_params, 0.0, C.c_void_p(eng.grads.data_ptr()), sp)))
        second = [eng._optimizer_op(), ("step_end", L.fnet_step_end, (st, global_batch, sp))]
        return first, second

    def prepare(self, x, y, global_batch):
        eng = self.eng
        sizes, _ = split_batch(global_batch, self.world)
        eng._ensure_capacity(max(max(sizes), 1), max(max(sizes), 1))
        eng._set_data("train", x, y)

    def set_epoch(self, order):
        """Same permutation on every rank; this rank's cursor starts at its slice of the first batch."""
        eng = self.eng
        with torch.cuda.stream(eng.stream):
            eng.data["train"][2].copy_(torch.from_numpy(np.asarray(order, dtype=np.int32)))
            eng.metrics.zero_()
        eng._metrics_seen = (0.0, 0)
        self._cursor_base = 0

    def step(self, global_batch):
        """One optimisation step on a global batch of `global_batch` rows starting at the epoch cursor."""
        eng = self.eng
        sizes, starts = split_batch(global_batch, self.world)
        local = sizes[self.rank]
        eng._sync_lr()
        key = (local, global_batch)
        if key not in self.graphs:
            need = max(eng._workspace_need(i, max(local, 1)) for i in range(len(eng.nodes)))
            if need > eng.ws_bytes:
                eng.stream.synchronize()
                eng._alloc_workspace(need)
                for fb, op in self.graphs.values():
                    eng.lib.fnet_graph_destroy(fb)
                    eng.lib.fnet_graph_destroy(op)
                self.graphs = {}
            first, second = self._programs(local, global_batch)
            self.graphs[key] = (eng._capture(first), eng._capture(second))
        fb, op = self.graphs[key]
        with torch.cuda.stream(eng.stream):
            # cursor of this rank = start of the global batch + offset of its slice
            eng.state[1:2].fill_(self._cursor_base + starts[self.rank])
        eng.lib.fnet_graph_launch(fb, eng.sp)
        with torch.cuda.stream(eng.stream):
            self.dist.all_reduce(eng.grads, op=self.dist.ReduceOp.SUM)     # NCCL, one message: the whole arena
        eng.lib.fnet_graph_launch(op, eng.sp)
        self._cursor_base += global_batch

    def epoch_metrics(self, n):
        """(mean loss, accuracy) of the training forward passes of this epoch, summed over ranks."""
        eng = self.eng
        with torch.cuda.stream(eng.stream):
            packed = torch.stack([eng.loss_sum[0], eng.correct[0].to(torch.float64)])
            self.dist.all_reduce(packed, op=self.dist.ReduceOp.SUM)
        eng.stream.synchronize()
        return float(packed[0].item()) / n, float(packed[1].item()) / n

    def fit_epoch(self, x, y, global_batch, order):
        n = x.shape[0]
        self.prepare(x, y, global_batch)
        self.set_epoch(order)
        for lo in range(0, n, global_batch):
            self.step(min(global_batch, n - lo))
        return self.epoch_metrics(n)
