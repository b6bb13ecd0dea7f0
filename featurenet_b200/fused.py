"""Turn the launch list of one step into ONE launch of the fused step kernel (csrc/fused_step.cu), when every operation
of the step has a fused implementation -- the narrow candidates of main_1block_nas.xml (small-channel convolutions,
concatenation, softmax head), whose step is otherwise 13 launches of ~4 us each for ~7 us of memory traffic.

The engine's program (Engine._build_program: a list of C-ABI calls with fixed pointers) stays the single description
of the step: this module translates each call into an `fnet_fused_op`, works out which operations depend on which
(by the buffers they read and write) and groups them into phases; a grid-wide barrier separates phases, operations
inside a phase follow one another without one.  Anything it does not recognise makes it return None and the engine
keeps the separate launches.
"""
import ctypes as C

import numpy as np
import torch

from . import lib as fl

(FK_BEGIN, FK_GATHER, FK_FEED_U8, FK_CONV_FWD, FK_CONV_DGRAD, FK_CONV_WGRAD, FK_REDUCE, FK_CONCAT_FWD, FK_CONCAT_BWD,
 FK_HEAD_FWD_LOSS, FK_HEAD_DGRAD, FK_HEAD_WGRAD, FK_ACT_BWD, FK_ADAM, FK_SGD) = range(1, 16)


class FusedOp(C.Structure):
    """fnet_fused_op of include/fnet.h."""
    _fields_ = [("kind", C.c_int32), ("barrier_after", C.c_int32), ("flags", C.c_int32), ("act", C.c_int32),
                ("d", C.c_int32 * 16), ("f", C.c_float * 4), ("p", C.c_void_p * 8)]


WORK = {}        # device address of a fused operation list -> (flops, bytes) of the step it replaces (rooflines)


def _ptr(v):
    if v is None:
        return 0
    if isinstance(v, tuple):          # ("scratch", offset): a piece of this program's own scratch, placed later
        return v
    if isinstance(v, C.c_void_p):
        return v.value or 0
    return int(v)


class _Op(object):
    def __init__(self, kind, d=(), f=(), p=(), flags=0, act=0, reads=(), writes=()):
        self.kind, self.d, self.f, self.p, self.flags, self.act = kind, list(d), list(f), [_ptr(x) for x in p], flags, act
        self.reads = {_ptr(x) for x in reads if _ptr(x)}
        self.writes = {_ptr(x) for x in writes if _ptr(x)}
        self.everything = False          # the optimizer: depends on all that came before
        self.phase = 0

    def pack(self, barrier_after):
        op = FusedOp()
        op.kind, op.barrier_after, op.flags, op.act = self.kind, int(barrier_after), int(self.flags), int(self.act)
        for i, v in enumerate(self.d):
            op.d[i] = int(v)
        for i, v in enumerate(self.f):
            op.f[i] = float(v)
        for i, v in enumerate(self.p):
            op.p[i] = v or None
        return op


def _geom(desc):
    return [desc.batch, desc.in_h, desc.in_w, desc.in_c, desc.out_h, desc.out_w, desc.out_c, desc.kh, desc.kw,
            desc.stride_h, desc.stride_w, desc.pad_top, desc.pad_left, desc.dil_h, desc.dil_w]


def try_fuse(engine, ops):
    """ops: the engine's launch list of one step.  Returns (launch tuple, keep-alive objects) or None."""
    L = engine.lib
    grid = L.fnet_fused_step_grid()
    if grid <= 0:
        return None
    out, scratch_floats = [], [0]
    flops = nbytes = 0

    def scratch(n):
        at = scratch_floats[0]
        scratch_floats[0] += -(-int(n) // 64) * 64          # 256-byte aligned pieces
        return ("scratch", at)

    pending_head = None          # a dense_fwd waiting for the softmax_cce that makes it the fused head
    row_buffers = None
    for label, fn, args in ops:
        name = str(label)
        if fn is None:
            if name in ("fork", "join"):
                continue
            return None           # an all-reduce (data-parallel step) or anything else that is not a launch
        f_, b_ = engine.op_work(name, args)
        flops, nbytes = flops + f_, nbytes + b_
        if pending_head is not None and name != "softmax_cce":
            return None
        if name == "step_begin":
            st, b1, b2 = args[0], args[1], args[2]
            out.append(_Op(FK_BEGIN, f=[b1, b2], p=[st], writes=[st]))
        elif name == "gather_batch":
            batch, x_row, y_row, perm, st, xs, ys, xd, yd, begin, b1, b2 = args[:12]
            if x_row >= 2 ** 31 or y_row >= 2 ** 31:
                return None
            out.append(_Op(FK_GATHER, d=[batch, x_row, y_row, begin], f=[b1, b2], p=[perm, st, xs, ys, xd, yd],
                           writes=[xd, yd, st]))
        elif name == "feed_batch_u8":
            batch, x_row, y_row, st, xs, scale, ys, xd, yd, begin, b1, b2 = args[:12]
            out.append(_Op(FK_FEED_U8, d=[batch, x_row, y_row, begin], f=[b1, b2, scale], p=[st, xs, ys, xd, yd],
                           reads=[xs, ys], writes=[xd, yd, st]))
        elif name == "conv2d_fwd":
            d, x, w, b, y = args[:5]
            if d.act == fl.ACT["softmax"] or not L.fnet_fused_conv_supported(C.byref(d), 0, 1):
                return None
            out.append(_Op(FK_CONV_FWD, d=_geom(d), p=[x, w, b, y], act=d.act, reads=[x, w, b], writes=[y]))
        elif name == "conv2d_bwd_data":
            if len(args) == 10:
                d, dz, w, dx, acc, act, yact = args[:7]
            else:
                (d, dz, w, dx, acc), act, yact = args[:5], 0, None
            if not L.fnet_fused_conv_supported(C.byref(d), 1, 0):
                return None
            out.append(_Op(FK_CONV_DGRAD, d=_geom(d), p=[dz, w, dx, yact], flags=1 if acc else 0, act=act,
                           reads=[dz, w, yact], writes=[dx]))
        elif name == "conv2d_bwd_filter":
            d, x, dz, dw, db = args[:5]
            with_bias = 1 if _ptr(db) else 0
            if not L.fnet_fused_conv_supported(C.byref(d), 2, with_bias):
                return None
            rows = d.kh * d.kw * d.in_c + with_bias
            total = rows * d.out_c
            part = scratch(grid * total)
            out.append(_Op(FK_CONV_WGRAD, d=_geom(d), p=[x, dz, part], flags=with_bias, reads=[x, dz], writes=[part]))
            out.append(_Op(FK_REDUCE, d=[grid, total, d.kh * d.kw * d.in_c * d.out_c], p=[part, dw, db], reads=[part],
                           writes=[dw, db]))
        elif name == "dense_fwd":
            m, k, n, act = args[:4]
            x, w, b, y = args[5:9]
            if act != 0 or not L.fnet_fused_head_supported(m, k, n):
                return None       # only the classifier head (logits for the loss) is fused
            pending_head = (m, k, n, x, w, b, y)
        elif name == "head_fwd_loss":
            m, k, n, x, w, b, probs, targets, scale, dz, loss_sum, correct = args[:12]
            if _ptr(correct) != _ptr(loss_sum) + 8:
                return None
            row_buffers = (scratch(m), scratch(m), m, loss_sum)
            out.append(_Op(FK_HEAD_FWD_LOSS, d=[m, k, n], f=[scale], p=[x, w, b, probs, targets, dz, row_buffers[0], row_buffers[1]],
                           reads=[x, w, b, targets], writes=[probs, dz, row_buffers[0], row_buffers[1]]))
        elif name == "softmax_cce":
            batch, classes, logits, targets, scale, from_logits, probs, dz, loss_sum, correct = args[:10]
            if pending_head is None or not from_logits:
                return None
            m, k, n, x, w, b, y = pending_head
            pending_head = None
            if m != batch or n != classes or _ptr(y) != _ptr(logits) or _ptr(probs) != _ptr(logits):
                return None
            if _ptr(correct) != _ptr(loss_sum) + 8:
                return None       # the kernel closes the step on one 16-byte block [loss sum, correct count]
            row_buffers = (scratch(m), scratch(m), m, loss_sum)
            out.append(_Op(FK_HEAD_FWD_LOSS, d=[m, k, n], f=[scale], p=[x, w, b, probs, targets, dz, row_buffers[0], row_buffers[1]],
                           reads=[x, w, b, targets], writes=[probs, dz, row_buffers[0], row_buffers[1]]))
        elif name == "dense_bwd_data":
            if len(args) == 13:
                m, k, n, _, dz, w, dx, acc, act, yact = args[:10]
            else:
                (m, k, n, _, dz, w, dx, acc), act, yact = args[:8], 0, None
            if not L.fnet_fused_head_supported(m, k, n):
                return None
            out.append(_Op(FK_HEAD_DGRAD, d=[m, k, n], p=[dz, w, dx, yact], flags=1 if acc else 0, act=act,
                           reads=[dz, w, yact], writes=[dx]))
        elif name == "dense_bwd_filter":
            m, k, n, _, x, dz, dw, db = args[:8]
            if not L.fnet_fused_head_supported(m, k, n):
                return None
            out.append(_Op(FK_HEAD_WGRAD, d=[m, k, n], p=[x, dz, dw, db], reads=[x, dz], writes=[dw, db]))
        elif name == "concat_fwd":
            outer, ra, rb, inner, a, b, y = args[:7]
            if outer * (ra + rb) * inner >= 2 ** 31:
                return None
            out.append(_Op(FK_CONCAT_FWD, d=[outer, ra * inner, rb * inner], p=[a, b, y], reads=[a, b], writes=[y]))
        elif name == "concat_bwd":
            if len(args) != 10:
                return None       # the variant with a folded activation derivative has no fused counterpart
            outer, ra, rb, inner, dy, ga, gb, fa, fb = args[:9]
            if outer * (ra + rb) * inner >= 2 ** 31:
                return None
            out.append(_Op(FK_CONCAT_BWD, d=[outer, ra * inner, rb * inner], p=[dy, ga, gb],
                           flags=(1 if fa else 0) | (2 if fb else 0), reads=[dy], writes=[ga, gb]))
        elif name == "act_bwd":
            act, rows, cols, y, dy, dx, acc = args[:7]
            if rows * cols >= 2 ** 31 or act not in (1, 2, 3):
                return None
            out.append(_Op(FK_ACT_BWD, d=[rows * cols], p=[y, dy, dx], flags=1 if acc else 0, act=act, reads=[y, dy], writes=[dx]))
        elif name in ("adam_step", "sgd_step", "step_end"):
            rows = row_buffers[2] if row_buffers else 0
            rl, rh, metrics = (row_buffers[0], row_buffers[1], row_buffers[3]) if row_buffers else (None, None, None)
            if name == "adam_step":
                n, p, g, m_, v_, st, b1, b2, eps, advance = args[:10]
                op = _Op(FK_ADAM, d=[n, advance, rows], f=[b1, b2, eps], p=[p, g, m_, v_, st, rl, rh, metrics])
            elif name == "sgd_step":
                n, p, g, vel, st, momentum, nesterov, advance = args[:8]
                op = _Op(FK_SGD, d=[n, advance, rows, nesterov], f=[momentum], p=[p, g, vel, None, st, rl, rh, metrics])
            else:
                st, advance = args[:2]
                op = _Op(FK_ADAM, d=[0, advance, rows], f=[0.9, 0.999, 1e-7], p=[None, None, None, None, st, rl, rh, metrics])
            if op.d[0] >= 2 ** 31:
                return None
            op.everything = True
            out.append(op)
            row_buffers = None
        else:
            return None
    if pending_head is not None or row_buffers is not None or not out or not out[-1].everything:
        return None
    if len(out) > 48:                 # FS_MAX_OPS of csrc/fused_step.cu
        return None

    # ---- phases: an operation comes after everything it conflicts with (read-after-write, write-after-read, write-after-write)
    for j, op in enumerate(out):
        phase = 0
        for prev in out[:j]:
            clash = op.everything or prev.everything or (prev.writes & (op.reads | op.writes)) or (prev.reads & op.writes)
            if clash:
                phase = max(phase, prev.phase + 1)
        op.phase = phase
    order = sorted(range(len(out)), key=lambda i: (out[i].phase, i))
    out = [out[i] for i in order]

    # ---- device memory: scratch (filter-gradient partials, per-row loss), the barrier word, the operation list
    with engine._on_stream():
        scratch_t = torch.empty(max(scratch_floats[0], 64), dtype=torch.float32, device=engine.device)
        sync_t = torch.zeros(2, dtype=torch.int32, device=engine.device)
    base = scratch_t.data_ptr()
    array = (FusedOp * len(out))()
    for i, op in enumerate(out):
        op.p = [base + 4 * v[1] if isinstance(v, tuple) else v for v in op.p]
        array[i] = op.pack(barrier_after=(i + 1 < len(out) and out[i + 1].phase != op.phase))
    host = np.frombuffer(bytes(array), dtype=np.uint8).copy()
    with engine._on_stream():
        ops_t = torch.from_numpy(host).to(engine.device)
    engine.stream.synchronize()                 # the pageable source of the upload may go away
    WORK[ops_t.data_ptr()] = (flops, nbytes)
    launch = ("fused_step", L.fnet_fused_step, (C.c_void_p(ops_t.data_ptr()), len(out), C.c_void_p(sync_t.data_ptr()), engine.sp))
    summary = {"operations": len(out), "phases": out[-1].phase + 1, "replaced_launches": sum(1 for op in ops if op[1] is not None)}
    return launch, (ops_t, sync_t, scratch_t), summary
