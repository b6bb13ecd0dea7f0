"""CUDA executor of one candidate network: memory plan, launch programs, CUDA-graph steps, fit loop.

Replaces what Keras/TensorFlow do underneath /root/reference/helpers.py:105-111 (``model.fit``)
and tensorflow_generator.py:270 (``model.evaluate``): one ``session.run`` per step with a host
batch fed through feed_dict becomes one CUDA-graph launch per step on a device-resident
dataset.

Memory layout (all fp32, NHWC, allocated once per candidate through torch, sized for HBM):
  * ``params`` / ``grads`` / ``adam_m`` / ``adam_v``: four flat arenas with identical layout,
    every weight tensor at a 64-float (256 B) aligned offset, so the optimizer is one launch
    and (for the data-parallel mode) the gradient exchange is one all-reduce;
  * ``stats``: the non-trainable BatchNorm moving statistics;
  * one activation and one gradient buffer per layer output, sized for the largest batch;
  * one workspace shared by all split reductions (kernels of a step are serial on one stream);
  * ``state`` (int32[8]): iteration count, epoch cursor, lr, lr_t -- read by the kernels so a
    captured step is identical from launch to launch.

PyTorch is used for allocation, streams and host<->device copies only; every arithmetic
kernel is launched through the C ABI (featurenet_b200.lib).  No CPU execution path exists.
"""
import contextlib
import ctypes as C
import math
import sys
import threading
import time

import numpy as np
import torch

from . import lib as fl
from .layers import glorot_limit

ALIGN = 64            # floats; weight tensors start on 256-byte boundaries
GEMM_KINDS = ("conv2d", "sepconv2d", "dense")


def _prod(shape):
    n = 1
    for s in shape:
        n *= int(s)
    return n


class HistoryResult(object):
    """What keras ``fit`` returns: ``.history`` maps metric name -> list per epoch."""

    def __init__(self):
        self.history = {}
        self.epoch = []


class DeviceDataCache(object):
    """Datasets stay resident in HBM across candidates (the reference re-feeds every batch from host RAM)."""
    _store = {}

    _lock = threading.Lock()

    @classmethod
    def get(cls, array, device, dtype=torch.float32):
        key = (id(array), array.shape, str(array.dtype), str(device))
        with cls._lock:                   # candidates resident side by side ask for the same dataset at the same time
            hit = cls._store.get(key)
            if hit is not None and hit[0] is array:
                return hit[1]
            host = np.ascontiguousarray(array)
            with torch.cuda.device(device):
                t = torch.from_numpy(host).to(device=device, dtype=dtype, non_blocking=False).contiguous()
                torch.cuda.current_stream().synchronize()      # read by every engine's own stream from here on
            if len(cls._store) > 16:
                cls._store.clear()
            cls._store[key] = (array, t)
            return t

    @classmethod
    def clear(cls):
        cls._store.clear()


class SideLabel(str):
    """Label of a launch that runs on the engine's side stream (see Engine._backward_program)."""


class Engine(object):
    # inference batches are as large as this many bytes of activations allow (at most 1000 rows): results do not depend
    # on the batch, launches per evaluated image do; a bound on the whole footprint keeps co-resident candidates small
    EVAL_ACT_BYTES = 1 << 30

    def __init__(self, net, device=None, math_mode=None, seed=None):
        if not torch.cuda.is_available():
            raise fl.FnetError("featurenet_b200 needs a CUDA device: there is no CPU execution path")
        import os
        self.lib = fl.get()
        self.net = net
        self.device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        # "auto" (the default) = fp32 accuracy with the faster of the FFMA and the three-MMA tensor-core kernel per
        # product (FNET_MATH_AUTO); "fp32" forces FFMA everywhere, "tf32x3" the tensor core wherever it can run
        mode = (math_mode or os.environ.get("FNET_PRECISION", "auto")).lower()
        if mode == "bf16":
            raise NotImplementedError("FNET_PRECISION=bf16 is not built (it needs bf16 activation storage in every kernel); "
                                      "the reduced-precision tensor-core mode is 'tf32'")
        self.math = fl.MATH[mode]
        # Streams of this engine's OWN (fnet_stream_create), not torch.cuda.Stream(): torch hands out 32 pooled streams
        # per device round-robin, so with several candidates resident per GPU a long-lived engine ends up sharing a
        # stream with a later one -- and whatever that one launches while this one is capturing a graph on the shared
        # stream is recorded into the wrong graph instead of being executed.
        self._streams = []
        self.stream = self._new_stream()
        self.sp = C.c_void_p(self.stream.cuda_stream)
        # filter gradients run on a second stream, beside the data gradient of the same layer (both only read
        # dz); inside a captured step this becomes a fork/join of the CUDA graph, see _backward_program
        self.side = self._new_stream()
        self.sp2 = C.c_void_p(self.side.cuda_stream)
        self.overlap_wgrad = os.environ.get("FNET_NO_OVERLAP", "0") != "1"
        # folding act'(y) into the producing data gradient pays on the FFMA kernels; the tensor-core epilogues are
        # the critical path of their kernels (one tile's epilogue per main loop) and do not hide the extra loads
        # (in "auto" mode the planner asks the library, per consumer layer, which kind of kernel its data gradient gets)
        self.fuse_act_bwd = os.environ.get("FNET_NO_ACT_FUSION", "0") != "1" and self.math in (fl.MATH["fp32"], fl.MATH["auto"])
        self.nodes = list(net.layers)
        self.index = {id(n): i for i, n in enumerate(self.nodes)}
        self.src = [[self.index[id(t.node)] for t in n.inputs] for n in self.nodes]
        self.shape = [tuple(n.output.shape.as_list()[1:]) for n in self.nodes]
        self.n_out = _prod(self.shape[-1])
        head = self.nodes[-1]
        if head.kind != "dense" or head.cfg.get("activation") != "softmax":
            raise NotImplementedError("the loss kernel expects a Dense(softmax) output layer (keras_model.py:124)")
        self.seed = int(seed if seed is not None else int(os.environ.get("FNET_SEED", "0")) or
                        np.random.SeedSequence().entropy % (1 << 62))
        self.dropout_seed = (self.seed * 0x9E3779B97F4A7C15 + 0x1234567) & 0xFFFFFFFFFFFFFFFF
        self.capacity = 0             # rows the activation buffers hold (training AND inference batches)
        self.grad_capacity = 0        # rows the gradient buffers hold (training / attack batches only)
        self.generation = 0           # bumped whenever buffers that captured graphs point at are replaced
        self.graphs = {}
        self.programs = {}
        # host seconds per phase of this candidate's life (bench.py prints them inside candidate_cycle)
        self.timing = {"alloc_s": 0.0, "capture_s": 0.0, "train_s": 0.0, "eval_s": 0.0, "train_steps": 0, "eval_steps": 0}
        self.use_graphs = os.environ.get("FNET_NO_GRAPH", "0") != "1"
        self.poison = os.environ.get("FNET_POISON", "0") == "1"
        # FNET_FUSED=1: steps made of small-channel layers only run as ONE persistent cooperative kernel (fused.py,
        # csrc/fused_step.cu).  Correct and deterministic (tests/test_gpu_fused.py), but measured SLOWER than the
        # separate launches on the bench candidate (64 vs 55 us per step on B200: 11 phases of ~4 us, each two dependent
        # trips to L2 plus a grid barrier, against launches whose latencies overlap through programmatic dependent
        # launch), so it stays opt-in; profiles/r2_fused_phases.txt has the per-phase timers.
        self.fuse_steps = os.environ.get("FNET_FUSED", "0") == "1"
        # the classifier head and its loss as ONE launch (fnet_head_fwd_loss) instead of dense_fwd + softmax_cce
        self.fuse_head = os.environ.get("FNET_NO_HEAD_FUSION", "0") != "1"
        self._fused_keep, self.fused_summary = {}, {}
        self.dp = None                # featurenet_b200.dataparallel.DataParallel once enable_data_parallel() ran
        # A data-parallel step is two captured graphs around an eagerly issued NCCL all-reduce.  FNET_DP_SINGLE_GRAPH=1
        # captures the collective into one graph with the rest of the step: it replays correctly in a steady loop, but
        # mixed with the eager collectives of the mode (metric reduction, permutation broadcast) through
        # torch.distributed it deadlocked in tests/dp_check.py on 2 B200s -- kept as an experiment, off by default.
        self.dp_single_graph = os.environ.get("FNET_DP_SINGLE_GRAPH", "0") == "1"
        self._plan_params()
        self._init_weights()
        with self._on_stream():
            self.state = torch.zeros(8, dtype=torch.int32, device=self.device)
            self.state_eval = torch.zeros(8, dtype=torch.int32, device=self.device)
            # one 16-byte block: [loss sum (double bits), correct count] -> a single D2H copy reads both
            self.metrics = torch.zeros(2, dtype=torch.int64, device=self.device)
            self.loss_sum = self.metrics[0:1].view(torch.float64)
            self.correct = self.metrics[1:2]
            self.ticket = torch.zeros(4, dtype=torch.int32, device=self.device)      # fnet_head_fwd_loss's "who finishes last"
            self._metrics_host = torch.zeros(2, dtype=torch.int64).pin_memory()
            self._metrics_np = self._metrics_host.numpy()      # same pinned memory, cheap to read from Python
            self._metrics_seen = (0.0, 0)
        self._lr_on_device = None
        self.timeline = None          # when a list: (label, start_event, end_event) per launch (profiling mode)

    # Streams an engine is done with are parked here and leased again, never destroyed: torch remembers the streams a
    # pinned host block was used on and records an event on each when the block is freed (possibly long after this
    # engine), which must not meet a dead handle.  A stream has exactly one owner while it is leased.
    _idle_streams = {}
    _idle_lock = threading.Lock()

    @contextlib.contextmanager
    def _on_stream(self):
        """Everything an engine allocates or initialises with torch is created ON ITS OWN STREAM: torch's fills
        (zeros, arange, NaN poison) are kernels on the current stream, and the thread's default stream is shared by
        every candidate resident on the GPU -- a fill queued there can land after this engine's first copy into the
        same buffer.  It also ties the memory to this stream in torch's allocator, so a block an earlier engine freed
        is reused in stream order."""
        with torch.cuda.device(self.device), torch.cuda.stream(self.stream):
            yield

    def _new_stream(self):
        key = self.device.index if self.device.index is not None else torch.cuda.current_device()
        with Engine._idle_lock:
            idle = Engine._idle_streams.setdefault(key, [])
            handle = idle.pop() if idle else None
        if handle is None:
            handle = C.c_void_p()
            with torch.cuda.device(self.device):
                self.lib.fnet_stream_create(C.byref(handle))
        self._streams.append(handle)
        return torch.cuda.ExternalStream(handle.value, device=self.device)

    # ------------------------------------------------------------------ parameters
    def _plan_params(self):
        self.w_off, self.s_off = {}, {}          # (node, k) -> (offset, size) in params / stats arena
        p = s = 0
        for i, node in enumerate(self.nodes):
            for k, spec in enumerate(node.weights):
                if spec.trainable:
                    self.w_off[(i, k)] = (p, spec.size)
                    p += -(-spec.size // ALIGN) * ALIGN
                else:
                    self.s_off[(i, k)] = (s, spec.size)
                    s += -(-spec.size // ALIGN) * ALIGN
        self.n_params, self.n_stats = p, s
        with self._on_stream():
            mk = lambda n: torch.zeros(max(n, ALIGN), dtype=torch.float32, device=self.device)
            self.params, self.grads, self.adam_m, self.adam_v = mk(p), mk(p), mk(p), mk(p)
            self.stats = mk(s)

    def _wview(self, arena, i, k):
        off, size = (self.w_off if (i, k) in self.w_off else self.s_off)[(i, k)]
        return arena[off: off + size]

    def _wp(self, i, k, arena=None):
        """Device pointer of weight k of node i in the given arena (params by default, stats for BN moving)."""
        if (i, k) in self.w_off:
            base = self.params if arena is None else arena
            return C.c_void_p(base.data_ptr() + 4 * self.w_off[(i, k)][0])
        return C.c_void_p(self.stats.data_ptr() + 4 * self.s_off[(i, k)][0])

    def _init_weights(self):
        rng = np.random.default_rng(self.seed)
        weights = []
        for node in self.nodes:
            for spec in node.weights:
                if spec.init == "glorot_uniform":
                    lim = glorot_limit(spec.fans)
                    weights.append(rng.uniform(-lim, lim, size=spec.shape).astype(np.float32))
                elif spec.init == "ones":
                    weights.append(np.ones(spec.shape, np.float32))
                else:
                    weights.append(np.zeros(spec.shape, np.float32))
        self.set_weights(weights)

    def get_weights(self):
        """Keras order: layer by layer, each layer's weights in its own order."""
        self.stream.synchronize()
        p, s = self.params.cpu().numpy(), self.stats.cpu().numpy()
        out = []
        for i, node in enumerate(self.nodes):
            for k, spec in enumerate(node.weights):
                arena, table = (p, self.w_off) if spec.trainable else (s, self.s_off)
                off, size = table[(i, k)]
                out.append(arena[off: off + size].reshape(spec.shape).copy())
        return out

    def set_weights(self, weights):
        p = np.zeros(max(self.n_params, ALIGN), np.float32)
        s = np.zeros(max(self.n_stats, ALIGN), np.float32)
        it = iter(weights)
        for i, node in enumerate(self.nodes):
            for k, spec in enumerate(node.weights):
                w = np.asarray(next(it), dtype=np.float32)
                if w.shape != spec.shape:
                    raise ValueError("weight %s of layer %s has shape %r, expected %r" % (spec.name, node.name, w.shape, spec.shape))
                arena, table = (p, self.w_off) if spec.trainable else (s, self.s_off)
                off, size = table[(i, k)]
                arena[off: off + size] = w.reshape(-1)
        with torch.cuda.stream(self.stream):
            self.params.copy_(torch.from_numpy(p))
            self.stats.copy_(torch.from_numpy(s))
        self.stream.synchronize()

    def reset_optimizer(self):
        with torch.cuda.stream(self.stream):
            self.adam_m.zero_()
            self.adam_v.zero_()
            self.state.zero_()
        self._lr_on_device = None

    # ------------------------------------------------------------------ data-parallel mode (one large candidate)
    def enable_data_parallel(self, dist, rank=None, world=None):
        """Train this candidate over all ranks of the process group (see featurenet_b200/dataparallel.py)."""
        from .dataparallel import DataParallel
        self.dp = DataParallel(dist, rank, world)
        self.stream.synchronize()
        self._drop_graphs()
        with self._on_stream():
            dist.broadcast(self.params, src=0)            # replicas start identical: rank 0's weights win
            dist.broadcast(self.stats, src=0)
        self.stream.synchronize()
        # every rank draws its own dropout masks (the towers of multi_gpu_model do)
        self.dropout_seed = (self.dropout_seed + 0x9E3779B97F4A7C15 * self.dp.rank) & 0xFFFFFFFFFFFFFFFF
        return self

    def _allreduce_grads(self):
        """SUM of the flat gradient arena over the ranks, in stream order (captured into the step's graph)."""
        with self._on_stream():
            self.dp.dist.all_reduce(self.grads, op=self.dp.dist.ReduceOp.SUM)
        self.dp.allreduce_calls += 1
        self.dp.allreduce_bytes = 4 * self.n_params

    def _average_stats(self):
        """BatchNorm moving statistics: mean over the ranks (each rank saw its own slices)."""
        if self.n_stats:
            with self._on_stream():
                self.dp.dist.all_reduce(self.stats, op=self.dp.dist.ReduceOp.SUM)
                self.stats.mul_(1.0 / self.dp.world)

    def snapshot(self, into=None):
        """Device-to-device copy of the weights (ModelCheckpoint / SGDR restarts)."""
        with self._on_stream():
            if into is None:
                into = (torch.empty_like(self.params), torch.empty_like(self.stats))
            into[0].copy_(self.params)
            into[1].copy_(self.stats)
        return into

    def restore(self, snap):
        with torch.cuda.stream(self.stream):
            self.params.copy_(snap[0])
            self.stats.copy_(snap[1])

    def save(self, path):
        names = []
        arrays = {}
        ws = iter(self.get_weights())
        for node in self.nodes:
            for spec in node.weights:
                key = "%s/%s" % (node.name, spec.name)
                names.append(key)
                arrays[key] = next(ws)
        with open(path, "wb") as handle:
            np.savez(handle, __order__=np.array(names), **arrays)

    # ------------------------------------------------------------------ activations
    def _ensure_capacity(self, batch, grad_batch=0):
        """Activation buffers for `batch` rows, gradient buffers (and what only a backward pass needs: pooling taps,
        the separable convolution's inner gradient) for `grad_batch` rows.  Inference batches are larger than
        training batches, and several candidates share one GPU: gradient memory follows the training batch only."""
        if batch <= self.capacity and grad_batch <= self.grad_capacity:
            return
        t_alloc = time.perf_counter()
        self.stream.synchronize()             # launches that still use the old buffers
        self._drop_graphs()
        batch = self.capacity = max(batch, self.capacity)
        gbatch = self.grad_capacity = max(grad_batch, self.grad_capacity)
        dev = self.device
        with self._on_stream():
            f32 = lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)
            self.act, self.gact, self.aux = {}, {}, {}
            ws_bytes = 256
            for i, node in enumerate(self.nodes):
                if node.kind == "flatten":
                    continue                       # a view of its input (resolved in _buf)
                self.act[i] = f32(batch, *self.shape[i])
                if node.kind != "input" and gbatch:
                    self.gact[i] = f32(gbatch, *self.shape[i])
                if node.kind == "maxpool":
                    self.aux[i] = torch.empty((max(gbatch, 1),) + self.shape[i], dtype=torch.uint8, device=dev)
                elif node.kind == "sepconv2d":
                    mid = self.shape[i][:2] + (self.shape[self.src[i][0]][2],)
                    self.aux[i] = (f32(batch, *mid), f32(max(gbatch, 1), *mid))
                elif node.kind == "bn":
                    a = self.shape[i][node.cfg["axis"] - 1]
                    self.aux[i] = (f32(a), f32(a))
                ws_bytes = max(ws_bytes, self._workspace_need(i, batch))
            self._alloc_workspace(ws_bytes)
            self.yb = f32(batch, self.n_out)
            if self.poison:
                # debugging aid (FNET_POISON=1): every buffer starts as NaN, so a launch that reads what no launch of
                # the step wrote shows up as a NaN loss instead of depending on what the allocator handed out
                for store in (self.act, self.gact):
                    for t in store.values():
                        t.fill_(float("nan"))
                for t in self.aux.values():
                    for u in (t if isinstance(t, tuple) else (t,)):
                        u.fill_(float("nan")) if u.dtype == torch.float32 else u.fill_(255)
                self.ws.view(torch.float32).fill_(float("nan"))
        self.timing["alloc_s"] += time.perf_counter() - t_alloc

    def _alloc_workspace(self, need):
        """Two halves of `need` bytes: the main stream's scratch and the side stream's (they run concurrently)."""
        need = -(-int(need) // 256) * 256
        with self._on_stream():
            self.ws = torch.empty(2 * need, dtype=torch.uint8, device=self.device)
        self.ws_bytes = need
        self.generation += 1

    def _workspace_need(self, i, batch):
        node, need = self.nodes[i], 0
        if node.kind == "conv2d":
            need = self.lib.fnet_conv2d_workspace_bytes(C.byref(self._conv_desc(i, batch)))
        elif node.kind == "dwconv2d":
            need = self.lib.fnet_dwconv2d_workspace_bytes(C.byref(self._conv_desc(i, batch)))
        elif node.kind == "sepconv2d":
            dw, pw = self._sep_descs(i, batch)
            need = max(self.lib.fnet_dwconv2d_workspace_bytes(C.byref(dw)), self.lib.fnet_conv2d_workspace_bytes(C.byref(pw)))
        elif node.kind == "dense":
            m, k, n = self._dense_mkn(i, batch)
            need = max(self.lib.fnet_dense_workspace_bytes(m, k, n), self.lib.fnet_head_fwd_loss_workspace_bytes(m))
        elif node.kind == "bn":
            need = self.lib.fnet_bn_workspace_bytes(self.shape[i][node.cfg["axis"] - 1])
        return int(need)

    def _resolve(self, i, training):
        """Index of the buffer that holds node i's output (flatten, and dropout at inference, are views)."""
        node = self.nodes[i]
        while node.kind == "flatten" or (node.kind == "dropout" and not training):
            i = self.src[i][0]
            node = self.nodes[i]
        return i

    def _a(self, i, training):
        return C.c_void_p(self.act[self._resolve(i, training)].data_ptr())

    def _g(self, i):
        return C.c_void_p(self.gact[self._resolve(i, True)].data_ptr())

    # ------------------------------------------------------------------ descriptors
    def _conv_desc(self, i, batch, act=True):
        node, cfg = self.nodes[i], self.nodes[i].cfg
        h, w, c = self.shape[self.src[i][0]]
        ho, wo, k = self.shape[i]
        dil = cfg.get("dilation", (1, 1))
        return fl.Conv2dDesc(batch, h, w, c, ho, wo, k, cfg["kernel"][0], cfg["kernel"][1], cfg["strides"][0],
                             cfg["strides"][1], cfg["pad_top"], cfg["pad_left"], dil[0], dil[1],
                             fl.ACT[cfg["activation"]] if act else 0, self.math)

    def _sep_descs(self, i, batch):
        cfg = self.nodes[i].cfg
        h, w, c = self.shape[self.src[i][0]]
        ho, wo, k = self.shape[i]
        dw = fl.Conv2dDesc(batch, h, w, c, ho, wo, c, cfg["kernel"][0], cfg["kernel"][1], cfg["strides"][0],
                           cfg["strides"][1], cfg["pad_top"], cfg["pad_left"], 1, 1, 0, self.math)
        pw = fl.Conv2dDesc(batch, ho, wo, c, ho, wo, k, 1, 1, 1, 1, 0, 0, 1, 1, fl.ACT[cfg["activation"]], self.math)
        return dw, pw

    def _dense_mkn(self, i, batch):
        in_shape = self.shape[self.src[i][0]]
        return batch * _prod(in_shape[:-1]), in_shape[-1], self.shape[i][-1]

    def _pool_desc(self, i, batch):
        cfg = self.nodes[i].cfg
        h, w, c = self.shape[self.src[i][0]]
        ho, wo, _ = self.shape[i]
        return fl.Pool2dDesc(batch, h, w, c, ho, wo, cfg["pool"][0], cfg["pool"][1], cfg["strides"][0],
                             cfg["strides"][1], cfg["pad_top"], cfg["pad_left"],
                             fl.POOL_MAX if self.nodes[i].kind == "maxpool" else fl.POOL_AVG)

    def _bn_view(self, i, batch):
        shp, axis = self.shape[i], self.nodes[i].cfg["axis"] - 1
        return batch * _prod(shp[:axis]), shp[axis], _prod(shp[axis + 1:])

    # ------------------------------------------------------------------ programs
    def _forward_program(self, batch, training, head_logits=True, save_for_backward=None):
        """head_logits: the "out" layer stops at its logits (the loss kernel does the softmax and writes the
        probabilities in place); False makes the layer apply its own softmax (predict())."""
        L, sp, ws, wsb = self.lib, self.sp, C.c_void_p(self.ws.data_ptr()), self.ws_bytes
        ops = []
        add = lambda label, fn, *args: ops.append((label, fn, args))
        last = len(self.nodes) - 1
        for i, node in enumerate(self.nodes):
            kind, cfg = node.kind, node.cfg
            if kind in ("input", "flatten") or (kind == "dropout" and not training):
                continue
            x = self._a(self.src[i][0], training)
            y = self._a(i, training)
            n_out = batch * _prod(self.shape[i])
            if kind == "conv2d":
                d = self._conv_desc(i, batch)
                add("conv2d_fwd", L.fnet_conv2d_fwd, d, x, self._wp(i, 0), self._wp(i, 1), y, ws, wsb, sp)
            elif kind == "dwconv2d":
                d = self._conv_desc(i, batch)
                add("dwconv2d_fwd", L.fnet_dwconv2d_fwd, d, x, self._wp(i, 0), self._wp(i, 1), y, sp)
            elif kind == "sepconv2d":
                dw, pw = self._sep_descs(i, batch)
                mid = C.c_void_p(self.aux[i][0].data_ptr())
                add("dwconv2d_fwd", L.fnet_dwconv2d_fwd, dw, x, self._wp(i, 0), None, mid, sp)
                add("conv2d_fwd", L.fnet_conv2d_fwd, pw, mid, self._wp(i, 1), self._wp(i, 2), y, ws, wsb, sp)
            elif kind == "dense":
                m, k, n = self._dense_mkn(i, batch)
                act = 0 if (i == last and head_logits) else fl.ACT[cfg["activation"]]
                add("dense_fwd", L.fnet_dense_fwd, m, k, n, act, self.math, x, self._wp(i, 0),
                    self._wp(i, 1), y, ws, wsb, sp)
            elif kind in ("maxpool", "avgpool"):
                keep = training if save_for_backward is None else save_for_backward      # argmax taps for the backward
                arg = C.c_void_p(self.aux[i].data_ptr()) if kind == "maxpool" and keep else None
                add("pool2d_fwd", L.fnet_pool2d_fwd, self._pool_desc(i, batch), x, y, arg, sp)
            elif kind == "gap":
                h, w, c = self.shape[self.src[i][0]]
                add("gap_fwd", L.fnet_gap_fwd, batch, h * w, c, x, y, sp)
            elif kind == "dropout":
                add("dropout_fwd", L.fnet_dropout_fwd, n_out, cfg["rate"], self.dropout_seed, i,
                    C.c_void_p(self.state.data_ptr()), x, y, sp)
            elif kind == "bn":
                outer, a, inner = self._bn_view(i, batch)
                if training:
                    mean, rstd = self.aux[i]
                    add("bn_fwd_train", L.fnet_bn_fwd_train, outer, a, inner, cfg["epsilon"], cfg["momentum"], x,
                        self._wp(i, 0), self._wp(i, 1), y, C.c_void_p(mean.data_ptr()), C.c_void_p(rstd.data_ptr()),
                        self._wp(i, 2), self._wp(i, 3), ws, wsb, sp)
                else:
                    add("bn_fwd_infer", L.fnet_bn_fwd_infer, outer, a, inner, cfg["epsilon"], x, self._wp(i, 0),
                        self._wp(i, 1), self._wp(i, 2), self._wp(i, 3), y, sp)
            elif kind == "act":
                cols = self.shape[i][-1]
                add("act_fwd", L.fnet_act_fwd, fl.ACT[cfg["activation"]], n_out // cols, cols, x, y, sp)
            elif kind in ("add", "mul"):
                fn = L.fnet_add_fwd if kind == "add" else L.fnet_mul_fwd
                add(kind + "_fwd", fn, n_out, x, self._a(self.src[i][1], training), y, sp)
            elif kind == "concat":
                axis = cfg["axis"] - 1
                sa, sb = self.shape[self.src[i][0]], self.shape[self.src[i][1]]
                outer, inner = batch * _prod(sa[:axis]), _prod(sa[axis + 1:])
                ra, rb = sa[axis], sb[axis]
                add("concat_fwd", L.fnet_concat_rows_fwd, outer, ra, rb, inner, x, self._a(self.src[i][1], training), y, sp)
            elif kind == "pad":
                h, w, c = self.shape[self.src[i][0]]
                add("pad2d_fwd", L.fnet_pad2d_fwd, batch, h, w, c, cfg["pad"][0], cfg["pad"][1], x, y, sp)
            else:
                raise NotImplementedError(kind)
        return ops

    def _dgrad_on_tensor_core(self, i, batch):
        """True when layer i's data gradient runs on a tensor-core kernel: its epilogue is that kernel's critical path,
        so the activation derivative of the producer stays a separate pass there."""
        if self.nodes[i].kind == "dense":
            m, k, n = self._dense_mkn(i, batch)
            return self.lib.fnet_dense_bwd_data_on_tensor_core(m, k, n, self.math) == 1
        return self.lib.fnet_conv2d_bwd_data_on_tensor_core(C.byref(self._conv_desc(i, batch))) == 1

    def _needs_grad(self, input_grad=False):
        """True for nodes that have a trainable weight at or above them (a gradient must reach them);
        with input_grad, for every node: the gradient is wanted at the network input itself."""
        need = [bool(input_grad)] * len(self.nodes)
        for i, node in enumerate(self.nodes):
            has = any(spec.trainable for spec in node.weights)
            need[i] = need[i] or has or any(need[j] for j in self.src[i])
        return need

    def _backward_program(self, batch, training=True, input_grad=False):
        """training=False differentiates the INFERENCE graph (dropout is the identity, BatchNorm an affine map on
        its moving statistics); input_grad carries the gradient down to the network input and skips the filter
        gradients -- together they are what the adversarial attacks of the robustness fitness need."""
        L, sp, ws, wsb = self.lib, self.sp, C.c_void_p(self.ws.data_ptr()), self.ws_bytes
        ops = []
        add = lambda label, fn, *args: ops.append((label, fn, args))
        need = self._needs_grad(input_grad)
        written = set()
        ws2 = C.c_void_p(self.ws.data_ptr() + self.ws_bytes)
        forked = [False]

        def add_wgrad(overlap, label, fn, *args):
            """A filter-gradient launch (args end with ws, ws_bytes, stream).  With a data gradient to overlap
            with, it goes to the side stream (own scratch half) behind a fork; _run_ops/_capture turn the fork
            and the final join into event dependencies.  Nothing later in the step writes what it reads (the
            layer's input activation and dz) or reads what it writes (the grads arena) before the join."""
            if input_grad:
                return                                  # weights are constants for an attack
            if overlap and self.overlap_wgrad:
                ops.append(("fork", None, ()))
                ops.append((SideLabel(label), fn, args[:-3] + (ws2, wsb, self.sp2)))
                forked[0] = True
            else:
                ops.append((label, fn, args))

        def acc_flag(j):
            """0 for the first gradient piece that reaches buffer j, 1 for the following ones."""
            r = self._resolve(j, True)
            flag = 1 if r in written else 0
            written.add(r)
            return flag

        gw = lambda i, k: self._wp(i, k, self.grads)
        last = len(self.nodes) - 1
        # Activation derivative folded into the data gradient that PRODUCES the tensor: when an activated GEMM layer
        # has exactly one consumer and that consumer is a Conv2D / Dense, the consumer's data-gradient epilogue
        # multiplies by act'(y) and the separate act_bwd pass over the tensor disappears.
        fused_into = {}
        if training and self.fuse_act_bwd:
            consumers = {}
            for j in range(1, len(self.nodes)):
                if self.nodes[j].kind == "flatten":
                    continue
                for s in self.src[j]:
                    consumers.setdefault(self._resolve(s, True), []).append(j)
            for r, users in consumers.items():
                node_r = self.nodes[r]
                if not (len(users) == 1 and r != last and node_r.kind in GEMM_KINDS + ("dwconv2d",) and need[r]
                        and fl.ACT[node_r.cfg["activation"]] in (1, 2, 3)):
                    continue
                user = self.nodes[users[0]]
                if user.kind in ("conv2d", "dense") and len(self.src[users[0]]) == 1 and not self._dgrad_on_tensor_core(users[0], batch):
                    fused_into[r] = users[0]
                elif user.kind == "bn" and training:
                    fused_into[r] = users[0]          # BatchNorm's backward reads this layer's output anyway (fnet_bn_bwd_act)
                elif user.kind in ("add", "concat"):
                    # the merge's backward is a copy of (a slice of) dy into this layer's gradient buffer: the derivative
                    # rides along (act_bwd reading the merge's dy / fnet_concat_rows_bwd_act) instead of a pass of its own
                    fused_into[r] = users[0]

        def dx_epilogue(i, s0):
            """(act, y) for node i's data gradient when it carries the activation derivative of its producer."""
            r = self._resolve(s0, True)
            if fused_into.get(r) == i:
                return fl.ACT[self.nodes[r].cfg["activation"]], self._a(r, True)
            return None

        for i in range(last, 0, -1):
            node, kind, cfg = self.nodes[i], self.nodes[i].kind, self.nodes[i].cfg
            if kind == "flatten" or not need[i]:
                continue
            s0 = self.src[i][0]
            x, y, dy = self._a(s0, training), self._a(i, training), self._g(i)
            n_out = batch * _prod(self.shape[i])
            # a merge owes a gradient to whichever of its inputs has trainable weights upstream -- not only to the first
            # one (concat([network input, branch]) is how every FeatureNet cell stacks its branch onto its input)
            want_dx = any(need[j] for j in self.src[i]) if kind in ("add", "mul", "concat") else need[s0]
            if kind in GEMM_KINDS or kind == "dwconv2d":
                act = fl.ACT[cfg["activation"]]
                if act and i != last and i not in fused_into:   # the head's dz comes straight from the loss kernel
                    cols = self.shape[i][-1]
                    add("act_bwd", L.fnet_act_bwd, act, n_out // cols, cols, y, dy, dy, 0, sp)
            if kind == "conv2d":
                d = self._conv_desc(i, batch)
                add_wgrad(want_dx, "conv2d_bwd_filter", L.fnet_conv2d_bwd_filter, d, x, dy, gw(i, 0), gw(i, 1), ws, wsb, sp)
                if want_dx:
                    ep = dx_epilogue(i, s0)
                    if ep:
                        add("conv2d_bwd_data", L.fnet_conv2d_bwd_data_act, d, dy, self._wp(i, 0), self._g(s0), acc_flag(s0),
                            ep[0], ep[1], ws, wsb, sp)
                    else:
                        add("conv2d_bwd_data", L.fnet_conv2d_bwd_data, d, dy, self._wp(i, 0), self._g(s0), acc_flag(s0), ws, wsb, sp)
            elif kind == "dwconv2d":
                d = self._conv_desc(i, batch)
                add_wgrad(want_dx, "dwconv2d_bwd_filter", L.fnet_dwconv2d_bwd_filter, d, x, dy, gw(i, 0), gw(i, 1), ws, wsb, sp)
                if want_dx:
                    add("dwconv2d_bwd_data", L.fnet_dwconv2d_bwd_data, d, dy, self._wp(i, 0), self._g(s0), acc_flag(s0), sp)
            elif kind == "sepconv2d":
                dw, pw = self._sep_descs(i, batch)
                mid, gmid = (C.c_void_p(t.data_ptr()) for t in self.aux[i])
                add_wgrad(True, "conv2d_bwd_filter", L.fnet_conv2d_bwd_filter, pw, mid, dy, gw(i, 1), gw(i, 2), ws, wsb, sp)
                add("conv2d_bwd_data", L.fnet_conv2d_bwd_data, pw, dy, self._wp(i, 1), gmid, 0, ws, wsb, sp)
                add_wgrad(want_dx, "dwconv2d_bwd_filter", L.fnet_dwconv2d_bwd_filter, dw, x, gmid, gw(i, 0), None, ws, wsb, sp)
                if want_dx:
                    add("dwconv2d_bwd_data", L.fnet_dwconv2d_bwd_data, dw, gmid, self._wp(i, 0), self._g(s0), acc_flag(s0), sp)
            elif kind == "dense":
                m, k, n = self._dense_mkn(i, batch)
                add_wgrad(want_dx, "dense_bwd_filter", L.fnet_dense_bwd_filter, m, k, n, self.math, x, dy, gw(i, 0), gw(i, 1), ws, wsb, sp)
                if want_dx:
                    ep = dx_epilogue(i, s0)
                    if ep:
                        add("dense_bwd_data", L.fnet_dense_bwd_data_act, m, k, n, self.math, dy, self._wp(i, 0), self._g(s0),
                            acc_flag(s0), ep[0], ep[1], ws, wsb, sp)
                    else:
                        add("dense_bwd_data", L.fnet_dense_bwd_data, m, k, n, self.math, dy, self._wp(i, 0), self._g(s0),
                            acc_flag(s0), ws, wsb, sp)
            elif kind == "bn" and not training:
                if want_dx:
                    outer, a, inner = self._bn_view(i, batch)
                    add("bn_bwd_infer", L.fnet_bn_bwd_infer, outer, a, inner, cfg["epsilon"], dy, self._wp(i, 0), self._wp(i, 3),
                        self._g(s0), acc_flag(s0), sp)
            elif kind == "bn":
                # dgamma/dbeta are always owed; when nothing upstream trains, dx lands on dy in place
                outer, a, inner = self._bn_view(i, batch)
                mean, rstd = self.aux[i]
                dx, flag = (self._g(s0), acc_flag(s0)) if want_dx else (dy, 0)
                r0 = self._resolve(s0, True)
                if want_dx and fused_into.get(r0) == i:
                    add("bn_bwd", L.fnet_bn_bwd_act, outer, a, inner, x, dy, self._wp(i, 0), C.c_void_p(mean.data_ptr()),
                        C.c_void_p(rstd.data_ptr()), dx, gw(i, 0), gw(i, 1), flag, fl.ACT[self.nodes[r0].cfg["activation"]], ws, wsb, sp)
                else:
                    add("bn_bwd", L.fnet_bn_bwd, outer, a, inner, x, dy, self._wp(i, 0), C.c_void_p(mean.data_ptr()),
                        C.c_void_p(rstd.data_ptr()), dx, gw(i, 0), gw(i, 1), flag, ws, wsb, sp)
            elif not want_dx:
                continue
            elif kind in ("maxpool", "avgpool"):
                arg = C.c_void_p(self.aux[i].data_ptr()) if kind == "maxpool" else None
                add("pool2d_bwd", L.fnet_pool2d_bwd, self._pool_desc(i, batch), dy, arg, self._g(s0), acc_flag(s0), sp)
            elif kind == "gap":
                h, w, c = self.shape[s0]
                add("gap_bwd", L.fnet_gap_bwd, batch, h * w, c, dy, self._g(s0), acc_flag(s0), sp)
            elif kind == "dropout" and not training:
                add("add_bwd", L.fnet_accumulate, n_out, dy, self._g(s0), acc_flag(s0), sp)     # identity at inference
            elif kind == "dropout":
                add("dropout_bwd", L.fnet_dropout_bwd, n_out, cfg["rate"], self.dropout_seed, i,
                    C.c_void_p(self.state.data_ptr()), dy, self._g(s0), acc_flag(s0), sp)
            elif kind == "act":
                cols = self.shape[i][-1]
                add("act_bwd", L.fnet_act_bwd, fl.ACT[cfg["activation"]], n_out // cols, cols, y, dy, self._g(s0),
                    acc_flag(s0), sp)
            elif kind == "add":
                for j in self.src[i]:
                    if need[j]:
                        rj = self._resolve(j, True)
                        if fused_into.get(rj) == i:
                            cols = self.shape[rj][-1]
                            add("act_bwd", L.fnet_act_bwd, fl.ACT[self.nodes[rj].cfg["activation"]], n_out // cols, cols,
                                self._a(rj, True), dy, self._g(j), acc_flag(j), sp)
                        else:
                            add("add_bwd", L.fnet_accumulate, n_out, dy, self._g(j), acc_flag(j), sp)
            elif kind == "mul":
                a, b = self.src[i]
                if need[a]:
                    add("mul_bwd", L.fnet_mul_bwd, n_out, dy, self._a(b, training), self._g(a), acc_flag(a), sp)
                if need[b]:
                    add("mul_bwd", L.fnet_mul_bwd, n_out, dy, self._a(a, training), self._g(b), acc_flag(b), sp)
            elif kind == "concat":
                axis = cfg["axis"] - 1
                a, b = self.src[i]
                sa, sb = self.shape[a], self.shape[b]
                outer, inner = batch * _prod(sa[:axis]), _prod(sa[axis + 1:])
                ra, rb = sa[axis], sb[axis]
                if self._resolve(a, True) == self._resolve(b, True):
                    # the same tensor on both sides: two ordered row copies into one gradient buffer
                    if need[a]:
                        add("copy_rows", L.fnet_copy_rows, outer, ra + rb, 0, ra, 0, ra, inner, dy, self._g(a), acc_flag(a), sp)
                        add("copy_rows", L.fnet_copy_rows, outer, ra + rb, ra, rb, 0, rb, inner, dy, self._g(b), acc_flag(b), sp)
                elif need[a] or need[b]:
                    ga, fa = (self._g(a), acc_flag(a)) if need[a] else (None, 0)
                    gb, fb = (self._g(b), acc_flag(b)) if need[b] else (None, 0)
                    folded = []
                    for operand, wanted in ((a, need[a]), (b, need[b])):
                        r_op = self._resolve(operand, True)
                        if wanted and fused_into.get(r_op) == i:
                            folded.append((fl.ACT[self.nodes[r_op].cfg["activation"]], self._a(r_op, True)))
                        else:
                            folded.append((0, None))
                    if folded[0][0] or folded[1][0]:
                        add("concat_bwd", L.fnet_concat_rows_bwd_act, outer, ra, rb, inner, dy, ga, gb, fa, fb,
                            folded[0][0], folded[0][1], folded[1][0], folded[1][1], sp)
                    else:
                        add("concat_bwd", L.fnet_concat_rows_bwd, outer, ra, rb, inner, dy, ga, gb, fa, fb, sp)
            elif kind == "pad":
                h, w, c = self.shape[s0]
                add("pad2d_bwd", L.fnet_pad2d_bwd, batch, h, w, c, cfg["pad"][0], cfg["pad"][1], dy, self._g(s0), acc_flag(s0), sp)
            else:
                raise NotImplementedError(kind)
        if forked[0]:
            ops.append(("join", None, ()))
        return ops

    def _loss_op(self, batch, with_grad, scale_batch=None):
        """scale_batch: the batch the mean is taken over (the GLOBAL batch in data-parallel training)."""
        L = self.lib
        last = len(self.nodes) - 1
        dz = self._g(last) if with_grad else None
        return ("softmax_cce", L.fnet_softmax_cce,
                (batch, self.n_out, self._a(last, with_grad), C.c_void_p(self.yb.data_ptr()),
                 1.0 / (scale_batch or batch), 1, self._a(last, with_grad), dz,       # from logits, softmax in place
                 C.c_void_p(self.loss_sum.data_ptr()), C.c_void_p(self.correct.data_ptr()), self.sp))

    def _optimizer_op(self, advance=-1):
        """advance >= 0: the optimizer launch also closes the step (iteration += 1, cursor += advance)."""
        opt, L = self.net.optimizer, self.lib
        st = C.c_void_p(self.state.data_ptr())
        args4 = [C.c_void_p(t.data_ptr()) for t in (self.params, self.grads, self.adam_m, self.adam_v)]
        if opt.kind == "adam":
            return ("adam_step", L.fnet_adam_step, (self.n_params, args4[0], args4[1], args4[2], args4[3], st,
                                                    opt.beta_1, opt.beta_2, opt.epsilon, advance, self.sp))
        return ("sgd_step", L.fnet_sgd_step, (self.n_params, args4[0], args4[1], args4[2], st, opt.momentum,
                                              int(opt.nesterov), advance, self.sp))

    def _build_program(self, kind, batch):
        """kind: 'train' / 'eval' read the resident dataset through the permutation; '*_fed' expect xb/yb filled."""
        return self._fuse_head(self._build_program_plain(kind, batch))

    def _fuse_head(self, ops):
        """dense_fwd(head logits) directly followed by softmax_cce(from logits, in place) -> one fnet_head_fwd_loss."""
        if not self.fuse_head:
            return ops
        val = lambda p: p.value if isinstance(p, C.c_void_p) else p
        out, i = [], 0
        while i < len(ops):
            a = ops[i]
            b = ops[i + 1] if i + 1 < len(ops) else None
            if b is not None and str(a[0]) == "dense_fwd" and str(b[0]) == "softmax_cce":
                m, k, n, act = a[2][:4]
                x, w, bias, y = a[2][5:9]
                batch, classes, logits, targets, scale, from_logits, probs, dz, loss_sum, correct = b[2][:10]
                if (act == 0 and from_logits and m == batch and n == classes and val(y) == val(logits) == val(probs)
                        and k < 32768 and self.lib.fnet_fused_head_supported(m, k, n)):
                    out.append(("head_fwd_loss", self.lib.fnet_head_fwd_loss,
                                (m, k, n, x, w, bias, probs, targets, scale, dz, loss_sum, correct,
                                 C.c_void_p(self.ticket.data_ptr()), C.c_void_p(self.ws.data_ptr()), self.ws_bytes, self.sp)))
                    i += 2
                    continue
            out.append(a)
            i += 1
        return out

    def _build_program_plain(self, kind, batch):
        L, sp = self.lib, self.sp
        if kind == "grad_input":        # act[0] / yb are filled by the caller; inference graph, gradient down to the input
            return (self._forward_program(batch, False, save_for_backward=True) + [self._loss_op(batch, True)] +
                    self._backward_program(batch, training=False, input_grad=True))
        if kind == "logits":            # act[0] is filled by the caller; the head stops before its softmax
            return self._forward_program(batch, False)
        if kind == "grad_logits":       # as above, but the caller also fills the head's dz: no loss in between
            return (self._forward_program(batch, False, save_for_backward=True) +
                    self._backward_program(batch, training=False, input_grad=True))
        if kind.startswith("train_dp:"):
            return self._build_dp_program(int(kind.split(":")[1]), batch)
        training = kind.startswith("train")
        state = self.state if training else self.state_eval
        st = C.c_void_p(state.data_ptr())
        ops = []
        b1 = b2 = 0.0
        if training:
            opt = self.net.optimizer
            b1, b2 = (opt.beta_1, opt.beta_2) if opt.kind == "adam" else (0.9, 0.999)
        fused_gather = kind in ("train", "eval")         # inputs + targets (+ lr_t) in one launch
        ring_u8 = kind.startswith("train_ringu8_")
        ring_slot = int(kind[len("train_ringu8_" if ring_u8 else "train_ring"):]) if kind.startswith("train_ring") else None
        if ring_u8:
            # the batch was uploaded as uint8 pixels: scale on the device while moving it into place
            xs, ys = self.ring["x_u8"][ring_slot], self.ring["y"][ring_slot]
            ops.append(("feed_batch_u8", L.fnet_feed_batch_u8,
                        (batch, _prod(self.shape[0]), self.n_out, st, C.c_void_p(xs.data_ptr()), float(self.ring["scale"]),
                         C.c_void_p(ys.data_ptr()), C.c_void_p(self.act[0].data_ptr()), C.c_void_p(self.yb.data_ptr()),
                         1, b1, b2, sp)))
        elif ring_slot is not None:
            # pipelined host feed: the batch was uploaded into slot `ring_slot` of the ring while the previous step
            # ran; the same launch as the resident path moves it into place (identity order, cursor stays 0)
            xs, ys = self.ring["x"][ring_slot], self.ring["y"][ring_slot]
            ops.append(("gather_batch", L.fnet_gather_batch,
                        (batch, _prod(self.shape[0]), self.n_out, C.c_void_p(self.ring["order"].data_ptr()), st,
                         C.c_void_p(xs.data_ptr()), C.c_void_p(ys.data_ptr()), C.c_void_p(self.act[0].data_ptr()),
                         C.c_void_p(self.yb.data_ptr()), 1, b1, b2, sp)))
        elif training and not fused_gather:
            ops.append(("step_begin", L.fnet_step_begin, (st, b1, b2, sp)))
        if ring_slot is not None:
            pass
        elif fused_gather:
            xdev, ydev, perm = self.data[kind]
            ops.append(("gather_batch", L.fnet_gather_batch,
                        (batch, _prod(self.shape[0]), self.n_out, C.c_void_p(perm.data_ptr()), st,
                         C.c_void_p(xdev.data_ptr()), C.c_void_p(ydev.data_ptr()), C.c_void_p(self.act[0].data_ptr()),
                         C.c_void_p(self.yb.data_ptr()), 1 if training else 0, b1, b2, sp)))
        elif not kind.endswith("_fed"):
            xdev, ydev, perm = self.data[kind.split("_")[0]]
            row = _prod(self.shape[0])
            if kind == "train_aug":
                h, w, c = self.shape[0]
                tx, ty, flip = self.aug_dev
                ops.append(("gather_shift_flip", L.fnet_gather_shift_flip,
                            (batch, h, w, c, C.c_void_p(perm.data_ptr()), st, C.c_void_p(tx.data_ptr()),
                             C.c_void_p(ty.data_ptr()), C.c_void_p(flip.data_ptr()), C.c_void_p(xdev.data_ptr()),
                             C.c_void_p(self.act[0].data_ptr()), sp)))
            else:
                ops.append(("gather_rows", L.fnet_gather_rows, (batch, row, C.c_void_p(perm.data_ptr()), st,
                                                                C.c_void_p(xdev.data_ptr()), C.c_void_p(self.act[0].data_ptr()), sp)))
            ops.append(("gather_rows", L.fnet_gather_rows, (batch, self.n_out, C.c_void_p(perm.data_ptr()), st,
                                                            C.c_void_p(ydev.data_ptr()), C.c_void_p(self.yb.data_ptr()), sp)))
        ops += self._forward_program(batch, training)
        ops.append(self._loss_op(batch, training))
        if training:
            ops += self._backward_program(batch)
            ops.append(self._optimizer_op(advance=0 if ring_slot is not None else batch))   # the update also closes the step
        else:
            ops.append(("step_end", L.fnet_step_end, (st, batch, sp)))
        return ops

    def _build_dp_program(self, global_batch, local):
        """One data-parallel step: this rank's slice (`local` rows, possibly none) of a global batch, the gradient
        all-reduce, and the optimizer, which advances the epoch cursor by the GLOBAL batch."""
        L, sp = self.lib, self.sp
        st = C.c_void_p(self.state.data_ptr())
        opt = self.net.optimizer
        b1, b2 = (opt.beta_1, opt.beta_2) if opt.kind == "adam" else (0.9, 0.999)
        _, start = self.dp.slice_of(global_batch)
        xdev, ydev, perm = self.data["train"]
        ops = []
        if local > 0:
            # the cursor points at the global batch; this rank's slice starts `start` rows further into the permutation
            ops.append(("gather_batch", L.fnet_gather_batch,
                        (local, _prod(self.shape[0]), self.n_out, C.c_void_p(perm.data_ptr() + 4 * start), st,
                         C.c_void_p(xdev.data_ptr()), C.c_void_p(ydev.data_ptr()), C.c_void_p(self.act[0].data_ptr()),
                         C.c_void_p(self.yb.data_ptr()), 1, b1, b2, sp)))
            ops += self._forward_program(local, True)
            ops.append(self._loss_op(local, True, scale_batch=global_batch))
            ops += self._backward_program(local)
        else:
            ops.append(("step_begin", L.fnet_step_begin, (st, b1, b2, sp)))
            ops.append(("fill", L.fnet_fill, (self.n_params, 0.0, C.c_void_p(self.grads.data_ptr()), sp)))
        ops.append(("allreduce", None, ()))
        ops.append(self._optimizer_op(advance=global_batch))
        return ops

    def _fork(self):
        """Everything enqueued so far on the main stream happens before what the side stream gets next."""
        event = torch.cuda.Event()
        event.record(self.stream)
        self.side.wait_event(event)

    def _join(self):
        """What the side stream has been given happens before what the main stream gets next."""
        event = torch.cuda.Event()
        event.record(self.side)
        self.stream.wait_event(event)

    def _run_ops(self, ops):
        """Enqueue a launch list (eagerly, or into the capture that _capture opened)."""
        if self.timeline is None:
            for label, fn, args in ops:
                if fn is not None:
                    fn(*args)
                elif label == "fork":
                    self._fork()
                elif label == "allreduce":
                    self._allreduce_grads()
                else:
                    self._join()
            return
        # profiling: one event pair per launch, on the stream it runs on; side launches are serialised
        # (fork, launch, join) so that every kernel is timed alone
        for label, fn, args in ops:
            if label == "allreduce":
                self._allreduce_grads()
                continue
            if fn is None:
                continue
            on_side = isinstance(label, SideLabel)
            stream = self.side if on_side else self.stream
            if on_side:
                self._fork()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            fn(*args)
            e1.record(stream)
            if on_side:
                self._join()
            self.timeline.append((label, e0, e1))

    @staticmethod
    def op_work(label, args):
        """Algorithmic (flops, bytes) of one launch, the figures of SURVEY.md section 8(d) used for rooflines."""
        f = b = 0
        if label == "fused_step":
            from . import fused
            return fused.WORK.get(args[0].value, (0, 0))
        if label.startswith("conv2d") or label.startswith("dwconv2d"):
            d = args[0]
            m_out, m_in = d.batch * d.out_h * d.out_w, d.batch * d.in_h * d.in_w
            taps = d.kh * d.kw
            if label.startswith("conv2d"):
                f = 2 * m_out * d.out_c * taps * d.in_c
                b = 4 * (m_in * d.in_c + m_out * d.out_c + taps * d.in_c * d.out_c)
            else:
                f = 2 * m_out * d.in_c * taps
                b = 4 * (m_in * d.in_c + m_out * d.in_c + taps * d.in_c)
        elif label == "head_fwd_loss":
            m, k, n = args[0], args[1], args[2]
            f, b = 2 * m * k * n, 4 * (m * k + k * n + m * n) + 12 * m * n
        elif label.startswith("dense"):
            m, k, n = args[0], args[1], args[2]
            f, b = 2 * m * k * n, 4 * (m * k + k * n + m * n)
        elif label.startswith("pool2d"):
            d = args[0]
            b = 4 * d.channels * d.batch * (d.in_h * d.in_w + d.out_h * d.out_w) + d.batch * d.out_h * d.out_w * d.channels
        elif label.startswith("gap"):
            b = 4 * args[0] * args[2] * (args[1] + 1)
        elif label.startswith("bn_"):
            elems = args[0] * args[1] * args[2]
            b = 4 * elems * {"bn_fwd_train": 3, "bn_fwd_infer": 2, "bn_bwd": 5, "bn_bwd_infer": 2}[label]
        elif label in ("act_fwd", "act_bwd"):
            b = 4 * args[1] * args[2] * (2 if label == "act_fwd" else 3)
        elif label.startswith("dropout"):
            b = 4 * 2 * args[0]
        elif label in ("add_fwd", "mul_fwd", "mul_bwd"):
            b = 4 * 3 * args[0]
        elif label == "add_bwd":
            b = 4 * 2 * args[0]
        elif label == "copy_rows":
            b = 4 * 2 * args[0] * args[5] * args[6]
        elif label.startswith("concat"):
            b = 4 * 2 * args[0] * (args[1] + args[2]) * args[3]
        elif label.startswith("pad2d"):
            b = 4 * 2 * args[0] * args[1] * args[2] * args[3]
        elif label == "softmax_cce":
            b = 4 * 3 * args[0] * args[1]
        elif label == "adam_step":
            b = 28 * args[0]
        elif label == "sgd_step":
            b = 16 * args[0]
        elif label == "gather_rows":
            b = 4 * 2 * args[0] * args[1]
        elif label == "gather_batch":
            b = 4 * 2 * args[0] * (args[1] + args[2])
        elif label == "feed_batch_u8":
            b = args[0] * (5 * args[1] + 8 * args[2])
        elif label == "gather_shift_flip":
            b = 4 * 2 * args[0] * args[1] * args[2] * args[3]
        return f, b

    def profile_steps(self, kind, batch, steps, warmup=3):
        """Run `steps` eager steps with a CUDA-event pair around every launch (on the launching stream).

        Returns {label: dict(launches, ms_total, flops, bytes)}; used by bench.py for the roofline object.
        """
        for _ in range(warmup):
            self._step(kind, batch)
        self.stream.synchronize()
        self.timeline = []
        try:
            # park the stream for ~25 ms so the host enqueues ahead and launches run back to back:
            # otherwise every event pair would also time the host's launch latency
            with torch.cuda.stream(self.stream):
                torch.cuda._sleep(int(5e7))
            for _ in range(steps):
                self._step(kind, batch)
            self.stream.synchronize()
            ops = [op for op in self.programs[(kind, batch)] if op[1] is not None]
            table = {}
            for n, (label, e0, e1) in enumerate(self.timeline):
                row = table.setdefault(label, dict(launches=0, ms_total=0.0, flops=0, bytes=0))
                fl_, by_ = self.op_work(label, ops[n % len(ops)][2])
                row["launches"] += 1
                row["ms_total"] += e0.elapsed_time(e1)
                row["flops"] += fl_
                row["bytes"] += by_
        finally:
            self.timeline = None
        return table

    def _drop_graphs(self):
        for g in self.graphs.values():
            self.lib.fnet_graph_destroy(g)
        self.graphs, self.programs = {}, {}
        self._fused_keep, self.fused_summary = {}, {}
        self.generation += 1

    def _step(self, kind, batch):
        """Enqueue one step (captured CUDA graph after the first, eager run)."""
        ops = self._program(kind, batch)
        if not self.use_graphs or self.timeline is not None:
            self._run_ops(ops)
            return
        if self.dp is not None and kind.startswith("train_dp:") and not self.dp_single_graph:
            # the collective stays outside the graphs: [slice forward/backward] -> all-reduce -> [optimizer]
            first, second = self.graphs.get((kind + "/a", batch)), self.graphs.get((kind + "/b", batch))
            if first is None:
                at = [op[0] for op in ops].index("allreduce")
                first = self.graphs[(kind + "/a", batch)] = self._capture(ops[:at])
                second = self.graphs[(kind + "/b", batch)] = self._capture(ops[at + 1:])
            self.lib.fnet_graph_launch(first, self.sp)
            self._allreduce_grads()
            self.lib.fnet_graph_launch(second, self.sp)
            return
        self.lib.fnet_graph_launch(self._graph(kind, batch), self.sp)

    def _program(self, kind, batch):
        """The launch list of one step of `kind` at this batch size (built once)."""
        key = (kind, batch)
        ops = self.programs.get(key)
        if ops is None:
            # split plans depend on the batch (fewer rows -> more splits), so the shared workspace may have to grow
            need = max(self._workspace_need(i, max(batch, 1)) for i in range(len(self.nodes)))
            if need > self.ws_bytes:
                self.stream.synchronize()
                self._drop_graphs()                      # programs hold the old workspace pointer
                self._alloc_workspace(need)
            ops = self._build_program(kind, batch)
            if self.fuse_steps and batch > 0 and (kind.startswith("train") or kind == "eval") and ":" not in kind:
                # a step made of small-channel layers only becomes ONE launch of the fused step kernel (fused.py)
                from . import fused
                made = fused.try_fuse(self, ops)
                if made is not None:
                    launch, keep, summary = made
                    ops = [launch]
                    self._fused_keep[key] = keep
                    self.fused_summary[key] = summary
            self.programs[key] = ops
            self.launches_per_step = getattr(self, "launches_per_step", {})
            self.launches_per_step[key] = sum(1 for op in ops if op[1] is not None)
        return ops

    def _graph(self, kind, batch):
        """The captured step (captured on first use; nothing runs during the capture)."""
        ops = self._program(kind, batch)
        graph = self.graphs.get((kind, batch))
        if graph is None:
            graph = self.graphs[(kind, batch)] = self._capture(ops)
        return graph

    # While a thread records a graph (thread-local capture mode) it must not make "unsafe" CUDA calls: a stream or device
    # synchronisation, an allocation.  An Engine that the garbage collector finalises in the middle of a capture does
    # exactly that (release() waits for its streams), and the capture dies with "operation failed due to a previous error
    # during capture".  So finalisers that arrive on a capturing thread are parked and run when the capture is over.
    _capturing = threading.local()
    _parked = []
    _parked_lock = threading.Lock()

    @classmethod
    def _run_parked(cls):
        with cls._parked_lock:
            todo, cls._parked[:] = list(cls._parked), []
        for state in todo:
            ghost = object.__new__(cls)
            ghost.__dict__.update(state)
            try:
                ghost.release()
            except Exception:          # noqa: BLE001
                pass

    def _capture(self, ops):
        """Record a launch list into a CUDA graph (nothing executes until the graph is launched)."""
        Engine._capturing.depth = getattr(Engine._capturing, "depth", 0) + 1
        try:
            return self._capture_now(ops)
        finally:
            Engine._capturing.depth -= 1
            if Engine._capturing.depth == 0 and Engine._parked:
                Engine._run_parked()

    def _capture_now(self, ops):
        t0 = time.perf_counter()
        self.lib.fnet_graph_begin(self.sp)
        handle = C.c_void_p()
        try:
            self._run_ops(ops)
        except BaseException:
            # close the capture without letting its own error replace the one that matters; a fork without its join
            # leaves the side stream inside the (now invalid) capture, so both streams are replaced
            try:
                self.lib.fnet_graph_end(self.sp, C.byref(handle))
                if handle.value:
                    self.lib.fnet_graph_destroy(handle)
            except fl.FnetError:
                pass
            self.stream, self.side = self._new_stream(), self._new_stream()
            self.sp, self.sp2 = C.c_void_p(self.stream.cuda_stream), C.c_void_p(self.side.cuda_stream)
            self._drop_graphs()               # programs hold the old stream handles
            raise
        self.lib.fnet_graph_end(self.sp, C.byref(handle))
        self.timing["capture_s"] += time.perf_counter() - t0
        return handle

    # ------------------------------------------------------------------ host-side loop pieces
    def _sync_lr(self):
        """The optimizer's learning rate lives in state[2]; a change is one stream-ordered launch that carries the
        value as a kernel argument (SGDRScheduler changes it every batch: no host wait, no pinned scalar to keep alive)."""
        lr = float(self.net.optimizer.lr)
        if lr != self._lr_on_device:
            self.lib.fnet_state_set_lr(C.c_void_p(self.state.data_ptr()), lr, self.sp)
            self._lr_on_device = lr

    def _set_data(self, split, x, y):
        self.data = getattr(self, "data", {})
        xd = DeviceDataCache.get(x, self.device)
        yd = DeviceDataCache.get(y, self.device)
        old = self.data.get(split)
        if old is not None and old[0] is xd and old[1] is yd:
            return old
        n = xd.shape[0]
        with self._on_stream():
            perm = torch.arange(n, dtype=torch.int32, device=self.device)
        for key in [k for k in self.programs if k[0] == split]:
            if key in self.graphs:
                self.lib.fnet_graph_destroy(self.graphs.pop(key))
            self.programs.pop(key)
        self.data[split] = (xd, yd, perm)
        return self.data[split]

    def _fetch_metrics(self):
        """(loss sum, correct count) accumulated on the device so far: one 16-byte copy into pinned memory."""
        with torch.cuda.stream(self.stream):
            self._metrics_host.copy_(self.metrics, non_blocking=True)
        self.stream.synchronize()
        return float(self._metrics_host[0:1].view(torch.float64)[0]), int(self._metrics_host[1])

    def _read_metrics(self, n):
        loss_sum, correct = self._fetch_metrics()
        if self.dp is not None:       # every rank scored its slices / its shard: one all-reduce, same numbers everywhere
            from .dataparallel import reduce_metrics
            loss_sum, correct = reduce_metrics(self.dp.dist, loss_sum, correct, device=self.device)
        return loss_sum / n, correct / n

    def _zero_metrics(self, state):
        with torch.cuda.stream(self.stream):
            self.metrics.zero_()
            state[1:2].zero_()
        self._metrics_seen = (0.0, 0)

    def _check_xy(self, x, y):
        x, y = np.asarray(x), np.asarray(y)
        if tuple(x.shape[1:]) != self.shape[0]:
            raise ValueError("expected inputs of shape %r, got %r" % (self.shape[0], tuple(x.shape[1:])))
        if y.ndim != 2 or y.shape[1] != self.n_out or y.shape[0] != x.shape[0]:
            raise ValueError("expected one-hot targets of shape (%d, %d), got %r" % (x.shape[0], self.n_out, y.shape))
        return x, y

    def eval_batch_size(self, n):
        """Inference batch: results do not depend on it, so use large batches bounded by activation size."""
        per_image = 4 * sum(_prod(s) for i, s in enumerate(self.shape) if self.nodes[i].kind != "flatten")
        cap = max(32, self.EVAL_ACT_BYTES // max(per_image, 1))
        return int(min(n, cap, 1000))

    def evaluate(self, x, y, batch_size=32):
        """keras Model.evaluate -> [loss, accuracy] in inference phase (tensorflow_generator.py:270)."""
        t_eval = time.perf_counter()
        x, y = self._check_xy(x, y)
        n = x.shape[0]
        eb = self.eval_batch_size(n)
        self._ensure_capacity(eb)
        self._set_data("eval", x, y)
        # (the identity order the validation set was given in _set_data is never overwritten: nothing to upload)
        self._zero_metrics(self.state_eval)
        rows = n
        if self.dp is not None:
            # sharded by sample: this rank scores rows [lo, hi) of the set; _read_metrics sums over the ranks
            lo, hi = self.dp.shard_of(n)
            rows = hi - lo
            with self._on_stream():
                self.state_eval[1:2].fill_(lo)
        full = rows // eb
        if full and self.use_graphs and self.timeline is None:
            self.lib.fnet_graph_launch_n(self._graph("eval", eb), full, self.sp)
        else:
            for _ in range(full):
                self._step("eval", eb)
        if rows - full * eb:
            self._step("eval", rows - full * eb)
        loss, acc = self._read_metrics(n)
        self.timing["eval_s"] += time.perf_counter() - t_eval
        self.timing["eval_steps"] += -(-n // eb)
        return [loss, acc]

    def predict(self, x, batch_size=32):
        x = np.asarray(x, dtype=np.float32)
        n = x.shape[0]
        eb = self.eval_batch_size(n)
        self._ensure_capacity(eb)
        out = np.empty((n, self.n_out), np.float32)
        last = len(self.nodes) - 1
        for lo in range(0, n, eb):
            b = min(eb, n - lo)
            with torch.cuda.stream(self.stream):
                self.act[0][:b].copy_(torch.from_numpy(np.ascontiguousarray(x[lo: lo + b])))
                self._run_ops(self._forward_program(b, False, head_logits=False))
                self.stream.synchronize()
                out[lo: lo + b] = self.act[self._resolve(last, False)][:b].cpu().numpy()
        return out

    def train_on_batch(self, x, y):
        """keras Model.train_on_batch with HOST buffers: H2D copy, one optimisation step, D2H of (loss, acc)."""
        if self.net.optimizer is None:
            raise RuntimeError("compile() the model first")
        b = x.shape[0]
        self._ensure_capacity(b, b)
        self._sync_lr()
        xt = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
        yt = y if isinstance(y, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(y, dtype=np.float32))
        graph = self.graphs.get(("train_fed", b)) if (self.use_graphs and self.timeline is None) else None
        if graph is not None and xt.dtype == torch.float32 and yt.dtype == torch.float32 and xt.is_contiguous() \
                and yt.is_contiguous() and not xt.is_cuda and not yt.is_cuda:
            # steady state: both copies, the captured step, the 16-byte result read and the wait in ONE C call
            self.lib.fnet_graph_step_fed(graph, self.act[0].data_ptr(), xt.data_ptr(), xt.numel() * 4, self.yb.data_ptr(),
                                         yt.data_ptr(), yt.numel() * 4, self.metrics.data_ptr(),
                                         self._metrics_host.data_ptr(), 16, self.sp)
            loss_sum, correct = float(self._metrics_np.view(np.float64)[0]), int(self._metrics_np[1])
        else:
            with torch.cuda.stream(self.stream):
                self.act[0][:b].copy_(xt, non_blocking=True)
                self.yb[:b].copy_(yt, non_blocking=True)
            self._step("train_fed", b)
            loss_sum, correct = self._fetch_metrics()
        # the device accumulators keep running (no per-step zeroing launches); this step = the difference
        seen_loss, seen_correct = self._metrics_seen
        if loss_sum < seen_loss or correct < seen_correct:      # someone reset the accumulators (fit / evaluate)
            seen_loss, seen_correct = 0.0, 0
        self._metrics_seen = (loss_sum, correct)
        return [(loss_sum - seen_loss) / b, (correct - seen_correct) / b]

    def input_gradient(self, x, y):
        """dL/dx of the mean categorical cross-entropy of a batch, in INFERENCE mode (dropout off, BatchNorm on
        its moving statistics): what ART's ``KerasClassifier.loss_gradient`` hands to the attacks of the
        robustness fitness (tensorflow_generator.py:151-173).  x, y: host arrays or device tensors; returns a
        device tensor of x's shape that stays valid until the next call."""
        b = int(x.shape[0])
        self._prepare_input_gradient(b)
        xt = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
        yt = y if isinstance(y, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(y, dtype=np.float32))
        with torch.cuda.stream(self.stream):
            self.act[0][:b].copy_(xt, non_blocking=True)
            self.yb[:b].copy_(yt, non_blocking=True)
        self._step("grad_input", b)
        self.stream.synchronize()          # host sources may be temporaries
        return self.gact[0][:b]

    def logits(self, x):
        """The head's pre-softmax outputs in inference mode (ART's ``predict(x, logits=True)``), as a host array."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        n = x.shape[0]
        eb = self.eval_batch_size(n)
        self._ensure_capacity(eb)
        out = np.empty((n, self.n_out), np.float32)
        last = self._resolve(len(self.nodes) - 1, False)
        for lo in range(0, n, eb):
            b = min(eb, n - lo)
            with torch.cuda.stream(self.stream):
                self.act[0][:b].copy_(torch.from_numpy(x[lo: lo + b]), non_blocking=True)
            self._step("logits", b)
            self.stream.synchronize()
            out[lo: lo + b] = self.act[last][:b].cpu().numpy()
        return out

    def _prepare_input_gradient(self, b):
        """Buffers for a gradient that reaches the network input (training never needs dL/dx)."""
        self._ensure_capacity(b, b)
        if 0 not in self.gact or self.gact[0].shape[0] < self.grad_capacity:
            with self._on_stream():
                self.gact[0] = torch.empty((self.grad_capacity,) + tuple(self.shape[0]), dtype=torch.float32, device=self.device)
            for key in [k for k in self.programs if k[0] in ("grad_input", "grad_logits")]:
                if key in self.graphs:
                    self.lib.fnet_graph_destroy(self.graphs.pop(key))
                self.programs.pop(key)

    def logit_gradient(self, x, head_dz):
        """Vector-Jacobian product of the head's LOGITS Z in inference mode: row b of the result is
        sum_j head_dz[b, j] * dZ_j/dx[b].  A one-hot head_dz gives one row of ART's
        ``KerasClassifier.class_gradient(x, logits=True)``; CLEVER (model/metrics.py:309-314) wants
        grads[:, pred] - grads[:, target], i.e. head_dz = e_pred - e_target, in a single backward pass.
        Returns (gradient, logits): device tensors valid until the next call."""
        b = int(x.shape[0])
        self._prepare_input_gradient(b)
        last = self._resolve(len(self.nodes) - 1, True)
        xt = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
        zt = head_dz if isinstance(head_dz, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(head_dz, dtype=np.float32))
        with torch.cuda.stream(self.stream):
            self.act[0][:b].copy_(xt, non_blocking=True)
            self.gact[last][:b].copy_(zt, non_blocking=True)
        self._step("grad_logits", b)
        self.stream.synchronize()
        return self.gact[0][:b], self.act[self._resolve(len(self.nodes) - 1, False)][:b]

    def fit(self, x, y, batch_size=32, epochs=1, validation_data=None, shuffle=True, callbacks=(), verbose=0,
            permutations=None, initial_epoch=0, augment=None, **_):
        """keras Model.fit semantics: see module docstring.  ``permutations[e]`` overrides the shuffle (tests)."""
        from .callbacks import History, Callback
        if self.net.optimizer is None:
            raise RuntimeError("compile() the model first")
        x, y = self._check_xy(x, y)
        n = x.shape[0]
        batch_size = int(min(batch_size, n))
        val = None
        if validation_data is not None:
            val = self._check_xy(validation_data[0], validation_data[1])
        cap = max(batch_size, self.eval_batch_size(val[0].shape[0]) if val else 0)
        self._ensure_capacity(cap, batch_size)
        self._set_data("train", x, y)
        perm_dev = self.data["train"][2]
        train_kind = "train"
        if augment is not None:
            if len(self.shape[0]) != 3:
                raise ValueError("shift/flip augmentation needs image inputs [H, W, C]")
            train_kind = "train_aug"
            old = getattr(self, "aug_dev", None)
            if old is None or old[0].numel() != n:
                for key in [k for k in self.programs if k[0] == "train_aug"]:
                    if key in self.graphs:
                        self.lib.fnet_graph_destroy(self.graphs.pop(key))
                    self.programs.pop(key)
                with self._on_stream():
                    self.aug_dev = (torch.zeros(n, dtype=torch.float32, device=self.device),
                                    torch.zeros(n, dtype=torch.float32, device=self.device),
                                    torch.zeros(n, dtype=torch.uint8, device=self.device))
        history = History()
        cbs = [history] + list(callbacks)
        per_batch = [cb for cb in cbs if getattr(type(cb), "on_batch_end", None) not in (None, Callback.on_batch_end)]
        for cb in cbs:
            cb.set_model(self.net)
            cb.set_params({"epochs": epochs, "batch_size": batch_size, "samples": n, "steps": -(-n // batch_size)})
        self.net.stop_training = False
        for cb in cbs:
            cb.on_train_begin({})
        rng = np.random.default_rng((self.seed + 0x5EED) & 0xFFFFFFFF)
        steps = -(-n // batch_size)
        t_start = time.time()
        for epoch in range(initial_epoch, epochs):
            for cb in cbs:
                cb.on_epoch_begin(epoch, {})
            if permutations is not None:
                order = np.asarray(permutations[epoch], dtype=np.int32)
            elif shuffle:
                order = rng.permutation(n).astype(np.int32)
            else:
                order = np.arange(n, dtype=np.int32)
            with torch.cuda.stream(self.stream):
                if self.dp is not None:
                    perm_dev.copy_(self.dp.broadcast_order(order, self.device))      # rank 0's shuffle, everywhere
                else:
                    perm_dev.copy_(torch.from_numpy(order))
                if augment is not None:                 # this epoch's random shifts / flips, indexed by position
                    for dev_arr, host_arr in zip(self.aug_dev, augment.draw(n)):
                        dev_arr.copy_(torch.from_numpy(host_arr))
            self._zero_metrics(self.state)
            t_train = time.perf_counter()
            if self.dp is not None:
                for step in range(steps):
                    g = min(batch_size, n - step * batch_size)
                    self._sync_lr()
                    self._step("train_dp:%d" % g, self.dp.slice_of(g)[0])
                    for cb in per_batch:
                        cb.on_batch_end(step, {"batch": step, "size": g})
                self._average_stats()
            elif not per_batch and self.use_graphs and self.timeline is None:
                # nothing on the host looks at single batches: the epoch's full batches are ONE call into the library
                # (floor(n / batch) replays of the captured step), the ragged last batch a second one
                self._sync_lr()
                full = n // batch_size
                if full:
                    self.lib.fnet_graph_launch_n(self._graph(train_kind, batch_size), full, self.sp)
                if n - full * batch_size:
                    self._step(train_kind, n - full * batch_size)
            else:
                for step in range(steps):
                    b = min(batch_size, n - step * batch_size)
                    self._sync_lr()
                    self._step(train_kind, b)
                    for cb in per_batch:
                        cb.on_batch_end(step, {"batch": step, "size": b})
            loss, acc = self._read_metrics(n)
            self.timing["train_s"] += time.perf_counter() - t_train
            self.timing["train_steps"] += steps
            logs = {"loss": loss, "acc": acc}
            if val is not None:
                logs["val_loss"], logs["val_acc"] = self.evaluate(val[0], val[1])
            for cb in cbs:
                cb.on_epoch_end(epoch, logs)
            if verbose:
                print("Epoch %d/%d - %.1fs - %s" % (epoch + 1, epochs, time.time() - t_start,
                                                    " - ".join("%s: %.4f" % kv for kv in logs.items())))
            if self.net.stop_training:
                break
        for cb in cbs:
            cb.on_train_end({})
        result = HistoryResult()
        result.history, result.epoch = history.history, history.epoch
        return result

    def fit_generator(self, generator, steps_per_epoch=None, epochs=1, validation_data=None, callbacks=(), verbose=0, **kw):
        """keras Model.fit_generator (helpers.py:163).  A featurenet_b200 generator (it exposes the arrays it flows
        from) takes the augmentation path of helpers.py:115-167 with the dataset resident on the device; any other
        iterator of (x, y) host batches is streamed through the pipelined feed (_fit_stream)."""
        kw.pop("workers", None)
        if not hasattr(generator, "x") or not hasattr(generator, "y"):
            for name in ("max_queue_size", "use_multiprocessing", "shuffle", "class_weight"):
                kw.pop(name, None)
            return self._fit_stream(generator, steps_per_epoch, epochs, validation_data, callbacks, verbose, **kw)
        return self.fit(generator.x, generator.y, batch_size=generator.batch_size, epochs=epochs,
                        validation_data=validation_data, callbacks=callbacks, verbose=verbose, augment=generator, **kw)

    # ------------------------------------------------------------------ pipelined host feed
    _ring_kinds = ["train_ring%d" % k for k in range(8)]
    _ring_kinds_u8 = ["train_ringu8_%d" % k for k in range(8)]
    RING_SLOTS = 3      # two are enough to overlap upload and compute; the third absorbs host-side jitter

    def _ring_setup(self, batch):
        """Two slots of (device batch, pinned staging batch, pinned result, two events) and a copy stream."""
        self._ensure_capacity(batch, batch)
        ring = getattr(self, "ring", None)
        if ring is not None and ring["capacity"] >= self.capacity:
            return ring
        self.stream.synchronize()
        for key in [k for k in self.programs if k[0].startswith("train_ring")]:
            if key in self.graphs:
                self.lib.fnet_graph_destroy(self.graphs.pop(key))
            self.programs.pop(key)
        cap, n = self.capacity, self.RING_SLOTS
        with self._on_stream():
            ring = {"capacity": cap,
                    "x": torch.empty((n, cap) + tuple(self.shape[0]), dtype=torch.float32, device=self.device),
                    "y": torch.empty((n, cap, self.n_out), dtype=torch.float32, device=self.device),
                    "order": torch.arange(cap, dtype=torch.int32, device=self.device),
                    "x_u8": torch.empty((n, cap) + tuple(self.shape[0]), dtype=torch.uint8, device=self.device),
                    "scale": 255.0,
                    "x_stage": torch.empty((n, cap) + tuple(self.shape[0]), dtype=torch.float32).pin_memory(),
                    "y_stage": torch.empty((n, cap, self.n_out), dtype=torch.float32).pin_memory(),
                    "result": torch.zeros((n, 2), dtype=torch.int64).pin_memory(),
                    "prev": np.zeros(2, np.int64),
                    "copy_stream": self._new_stream(),
                    "slot_event": [torch.cuda.Event() for _ in range(n)],
                    "copied_event": [torch.cuda.Event() for _ in range(n)]}
            for ev in ring["slot_event"] + ring["copied_event"]:
                ev.record(self.stream)                     # materialises the cudaEvent_t handle
        self.stream.synchronize()
        # raw addresses per slot, gathered once: the per-step call must cost the host less than a step costs the GPU
        ring["ptrs"] = [(ring["x"][k].data_ptr(), ring["x_stage"][k].data_ptr(), ring["y"][k].data_ptr(),
                         ring["y_stage"][k].data_ptr(), ring["result"][k].data_ptr(), ring["slot_event"][k].cuda_event,
                         ring["copied_event"][k].cuda_event, ring["x_u8"][k].data_ptr()) for k in range(n)]
        ring["fixed"] = (self.metrics.data_ptr(), ring["prev"].ctypes.data, ring["copy_stream"].cuda_stream,
                         self.stream.cuda_stream)
        ring["prev64"] = ring["prev"].view(np.float64)
        self.ring = ring
        return ring

    def _stream_step(self, step, x, y, scale=None):
        """Enqueue one optimisation step on a host batch without waiting for it.  Returns the running
        (loss_sum, correct) as they stood after the step that used this slot before (RING_SLOTS steps ago).
        A uint8 batch is uploaded as it is (a quarter of the bytes) and divided by `scale` on the device, which is
        what tensorflow_generator.py:310-313 does to the whole dataset on the host."""
        b = int(x.shape[0])
        ring = getattr(self, "ring", None)
        if ring is None or b > ring["capacity"] or ring["capacity"] < self.capacity:
            ring = self._ring_setup(b)
        slot = step % self.RING_SLOTS
        xt = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x))
        yt = y if isinstance(y, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(y, dtype=np.float32))
        as_u8 = xt.dtype == torch.uint8
        if as_u8:
            if scale is None:
                raise ValueError("uint8 batches need the divisor that makes them the model's inputs (generator.scale)")
            if float(scale) != ring["scale"]:
                self.stream.synchronize()
                for key in [k for k in self.programs if k[0].startswith("train_ringu8_")]:
                    if key in self.graphs:
                        self.lib.fnet_graph_destroy(self.graphs.pop(key))
                    self.programs.pop(key)
                ring["scale"] = float(scale)
            if not xt.is_contiguous():
                xt = xt.contiguous()
        elif xt.dtype != torch.float32 or not xt.is_contiguous():
            xt = xt.float().contiguous()
        if yt.dtype != torch.float32 or not yt.is_contiguous():
            yt = yt.float().contiguous()
        if tuple(xt.shape[1:]) != tuple(self.shape[0]) or tuple(yt.shape) != (b, self.n_out):
            raise ValueError("batch of shape %s / %s does not fit the model" % (tuple(xt.shape), tuple(yt.shape)))
        pinned = xt.is_pinned() and yt.is_pinned()
        kinds = self._ring_kinds_u8 if as_u8 else self._ring_kinds
        graph = self.graphs.get((kinds[slot], b))
        if graph is None:
            graph = self._graph(kinds[slot], b)     # one captured step per (slot, batch size)
        self._sync_lr()
        x_slot, x_stage, y_slot, y_stage, result, slot_event, copied_event, x_slot_u8 = ring["ptrs"][slot]
        metrics, prev, copy_stream, compute_stream = ring["fixed"]
        self.lib.fnet_graph_step_streamed(
            graph, x_slot_u8 if as_u8 else x_slot, xt.data_ptr(), None if pinned else x_stage, xt.numel() * xt.element_size(),
            y_slot, yt.data_ptr(), None if pinned else y_stage, yt.numel() * 4,
            metrics, result, prev, 16, copy_stream, compute_stream, slot_event, copied_event)
        if pinned:
            ring.setdefault("alive", {})[slot] = (xt, yt)   # the upload reads the caller's buffer until the slot's next turn
        return float(ring["prev64"][0]), int(ring["prev"][1])

    def _fit_stream(self, generator, steps_per_epoch, epochs, validation_data, callbacks, verbose, initial_epoch=0, **_):
        """fit_generator for host batches.  The upload of batch i+1 (copy stream) overlaps the kernels of batch i
        (compute stream); the 16-byte result of every step comes back asynchronously and is handed to
        on_batch_end RING_SLOTS steps late; epoch metrics are the sample-weighted means keras reports."""
        from .callbacks import History, Callback
        if self.net.optimizer is None:
            raise RuntimeError("compile() the model first")
        if steps_per_epoch is None:
            if not hasattr(generator, "__len__"):
                raise ValueError("steps_per_epoch is needed for a generator without a length")
            steps_per_epoch = len(generator)
        batches = iter(generator)
        val = self._check_xy(validation_data[0], validation_data[1]) if validation_data is not None else None
        history = History()
        cbs = [history] + list(callbacks)
        per_batch = [cb for cb in cbs if getattr(type(cb), "on_batch_end", None) not in (None, Callback.on_batch_end)]
        for cb in cbs:
            cb.set_model(self.net)
            cb.set_params({"epochs": epochs, "steps": steps_per_epoch})
        self.net.stop_training = False
        for cb in cbs:
            cb.on_train_begin({})
        t_start = time.time()
        step_no = getattr(self, "_stream_steps", 0)
        for epoch in range(initial_epoch, epochs):
            for cb in cbs:
                cb.on_epoch_begin(epoch, {})
            self._zero_metrics(self.state)                 # stream-ordered behind the previous epoch's steps; cursor = 0
            seen = 0
            for step in range(steps_per_epoch):
                item = next(batches)
                x, y = item[0], item[1]
                b = int(x.shape[0])
                loss_sum, correct = self._stream_step(step_no, x, y, scale=getattr(generator, "scale", None))
                step_no += 1
                seen += b
                for cb in per_batch:
                    cb.on_batch_end(step, {"batch": step, "size": b, "loss_sum_lagged": loss_sum, "correct_lagged": correct})
            self._stream_steps = step_no
            loss, acc = self._read_metrics(max(seen, 1))
            logs = {"loss": loss, "acc": acc}
            if val is not None:
                logs["val_loss"], logs["val_acc"] = self.evaluate(val[0], val[1])
            for cb in cbs:
                cb.on_epoch_end(epoch, logs)
            if verbose:
                print("Epoch %d/%d - %.1fs - %s" % (epoch + 1, epochs, time.time() - t_start,
                                                    " - ".join("%s: %.4f" % kv for kv in logs.items())))
            if self.net.stop_training:
                break
        for cb in cbs:
            cb.on_train_end({})
        result = HistoryResult()
        result.history, result.epoch = history.history, history.epoch
        return result

    def release(self):
        if not self._streams:
            return
        with torch.cuda.device(self.device):
            for handle in self._streams:
                torch.cuda.ExternalStream(handle.value, device=self.device).synchronize()
            self._drop_graphs()
            for name in ("act", "gact", "aux", "ws", "params", "grads", "adam_m", "adam_v", "stats", "data", "ring", "yb",
                         "_metrics_np", "_metrics_host", "loss_sum", "correct", "metrics", "state", "state_eval", "aug_dev", "ticket"):
                if hasattr(self, name):
                    delattr(self, name)
            self.capacity = self.grad_capacity = 0
        key = self.device.index if self.device.index is not None else torch.cuda.current_device()
        with Engine._idle_lock:
            Engine._idle_streams.setdefault(key, []).extend(self._streams)
        self._streams = []

    def __del__(self):
        if sys is None or sys.is_finalizing():
            return                 # interpreter shutdown: the CUDA context may be gone; the process frees everything
        try:
            if getattr(Engine._capturing, "depth", 0) > 0 and getattr(self, "_streams", None):
                with Engine._parked_lock:              # this thread is recording a graph: release later (see _capture)
                    Engine._parked.append(dict(self.__dict__))
                return
            self.release()
        except Exception:          # noqa: BLE001
            pass
