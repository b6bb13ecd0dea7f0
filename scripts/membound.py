"""Achieved HBM GB/s of the memory-bound kernels at FeatureNet-sized tensors (batch 128), L2 flushed between repetitions.
Bytes are the algorithmic ones of Engine.op_work; peak = MEASURED_PEAKS.json hbm_gbs.

  python scripts/membound.py [reps]
"""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from featurenet_b200 import lib as fl  # noqa: E402
from featurenet_b200.engine import Engine  # noqa: E402

L = fl.get()
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
rnd = lambda *s: torch.randn(*s, device="cuda")
PEAK = 6537.3
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except OSError:
    pass
# L2 is flushed by READING 256 MB (clean lines).  Flushing by writing leaves the 126 MB L2 full of dirty lines whose
# write-back then competes with the kernel under test for HBM: on 33 MB tensors that halves the figure.
flush = torch.zeros(64 << 20, dtype=torch.int32, device="cuda")
rows = []


def timed(name, label, args, fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best, total = 1e9, 0.0
    for _ in range(reps):
        flush.sum()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best, total = min(best, ms), total + ms
    _, nbytes = Engine.op_work(label, args)
    ms = total / reps
    rows.append((name, 1e3 * ms, nbytes / ms / 1e6, nbytes / ms / 1e6 / PEAK))
    print("%-44s %8.1f us  %7.0f GB/s  %5.1f%% of %.0f" % (name, 1e3 * ms, nbytes / ms / 1e6, 100 * nbytes / ms / 1e6 / PEAK, PEAK), flush=True)


def pool(kind, k, s, shape):
    B, H, W, Cc = shape
    ho, wo = -(-H // s), -(-W // s)
    pt = max((ho - 1) * s + k - H, 0) // 2
    pl = max((wo - 1) * s + k - W, 0) // 2
    d = fl.Pool2dDesc(B, H, W, Cc, ho, wo, k, k, s, s, pt, pl, fl.POOL_MAX if kind == "max" else fl.POOL_AVG)
    x, y, dy, dx = rnd(*shape), torch.empty(B, ho, wo, Cc, device="cuda"), rnd(B, ho, wo, Cc), torch.empty(*shape, device="cuda")
    arg = torch.empty(B, ho, wo, Cc, dtype=torch.uint8, device="cuda") if kind == "max" else None
    timed("%spool %dx%d s%d fwd %s" % (kind, k, k, s, "x".join(map(str, shape))), "pool2d_fwd", (d,), lambda: L.fnet_pool2d_fwd(d, ptr(x), ptr(y), ptr(arg), None))
    timed("%spool %dx%d s%d bwd" % (kind, k, k, s), "pool2d_bwd", (d,), lambda: L.fnet_pool2d_bwd(d, ptr(dy), ptr(arg), ptr(dx), 0, None))


def depthwise(k, s, shape):
    B, H, W, Cc = shape
    ho, wo = -(-H // s), -(-W // s)
    pt = max((ho - 1) * s + k - H, 0) // 2
    pl = max((wo - 1) * s + k - W, 0) // 2
    d = fl.Conv2dDesc(B, H, W, Cc, ho, wo, Cc, k, k, s, s, pt, pl, 1, 1, 0, 0)
    x, w, b = rnd(*shape), rnd(k, k, Cc, 1), rnd(Cc)
    y, dz, dx = torch.empty(B, ho, wo, Cc, device="cuda"), rnd(B, ho, wo, Cc), torch.empty(*shape, device="cuda")
    dw, db = torch.empty_like(w), torch.empty_like(b)
    ws = torch.empty(max(L.fnet_dwconv2d_workspace_bytes(C.byref(d)), 256), dtype=torch.uint8, device="cuda")
    tag = "dw %dx%d s%d %s" % (k, k, s, "x".join(map(str, shape)))
    timed(tag + " fwd", "dwconv2d_fwd", (d,), lambda: L.fnet_dwconv2d_fwd(d, ptr(x), ptr(w), ptr(b), ptr(y), None))
    timed(tag + " dgrad", "dwconv2d_bwd_data", (d,), lambda: L.fnet_dwconv2d_bwd_data(d, ptr(dz), ptr(w), ptr(dx), 0, None))
    timed(tag + " wgrad", "dwconv2d_bwd_filter", (d,), lambda: L.fnet_dwconv2d_bwd_filter(d, ptr(x), ptr(dz), ptr(dw), ptr(db), ptr(ws), ws.numel(), None))


pool("max", 2, 1, (128, 32, 32, 32))
pool("max", 2, 2, (128, 32, 32, 64))
pool("avg", 3, 2, (128, 32, 32, 64))
depthwise(3, 1, (128, 32, 32, 64))
depthwise(3, 2, (128, 32, 32, 24))
depthwise(5, 1, (128, 16, 16, 192))

# BatchNorm over axis 1 of [128, 64, 32, 24]: outer 128, a 64, inner 768
outer, a, inner = 128, 64, 32 * 24
x, y, dy, dx = rnd(outer, a, inner), torch.empty(outer, a, inner, device="cuda"), rnd(outer, a, inner), torch.empty(outer, a, inner, device="cuda")
gamma, beta, mean, rstd, mm, mv, dg, dbt = (rnd(a) for _ in range(8))
mv.abs_()
ws = torch.empty(max(L.fnet_bn_workspace_bytes(a), 256), dtype=torch.uint8, device="cuda")
timed("bn fwd train 128x64x32x24", "bn_fwd_train", (outer, a, inner), lambda: L.fnet_bn_fwd_train(outer, a, inner, 1e-3, 0.99, ptr(x), ptr(gamma), ptr(beta), ptr(y), ptr(mean), ptr(rstd), ptr(mm), ptr(mv), ptr(ws), ws.numel(), None))
timed("bn bwd", "bn_bwd", (outer, a, inner), lambda: L.fnet_bn_bwd(outer, a, inner, ptr(x), ptr(dy), ptr(gamma), ptr(mean), ptr(rstd), ptr(dx), ptr(dg), ptr(dbt), 0, ptr(ws), ws.numel(), None))

n = 128 * 32 * 32 * 64
a1, a2, a3 = rnd(n), rnd(n), torch.empty(n, device="cuda")
timed("relu fwd 128x32x32x64", "act_fwd", (1, n // 64, 64), lambda: L.fnet_act_fwd(1, n // 64, 64, ptr(a1), ptr(a3), None))
timed("relu bwd", "act_bwd", (1, n // 64, 64), lambda: L.fnet_act_bwd(1, n // 64, 64, ptr(a1), ptr(a2), ptr(a3), 0, None))
timed("add fwd", "add_fwd", (n,), lambda: L.fnet_add_fwd(n, ptr(a1), ptr(a2), ptr(a3), None))
state = torch.zeros(8, dtype=torch.int32, device="cuda")
timed("dropout fwd", "dropout_fwd", (n,), lambda: L.fnet_dropout_fwd(n, 0.25, 1234, 3, ptr(state), ptr(a1), ptr(a3), None))
cat = torch.empty(2 * n, device="cuda")
timed("concat rows fwd (2 x 128x32x32x64)", "concat_fwd", (128, 32, 32, 32 * 64), lambda: L.fnet_concat_rows_fwd(128, 32, 32, 32 * 64, ptr(a1), ptr(a2), ptr(cat), None))
for params in (5341738, 20000000):
    p, g, m, v = (rnd(params) for _ in range(4))
    v.abs_()
    state[2] = torch.tensor(1e-3).view(torch.int32) if False else state[2]
    timed("adam %d parameters" % params, "adam_step", (params,), lambda: L.fnet_adam_step(params, ptr(p), ptr(g), ptr(m), ptr(v), ptr(state), 0.9, 0.999, 1e-7, -1, None))
print(json.dumps([{"kernel": r[0], "us": round(r[1], 1), "gbs": round(r[2]), "frac": round(r[3], 3)} for r in rows]))
