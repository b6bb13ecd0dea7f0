"""Time (and give ncu something small to look at) the three products of one Dense / Conv2D shape.

  python scripts/one_op.py dense M K N [fp32|tf32] [reps]
  python scripts/one_op.py conv B H W C K kh kw stride [fp32|tf32] [reps]
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from featurenet_b200 import lib as fl  # noqa: E402

L = fl.get()
kind = sys.argv[1]
ptr = lambda t: C.c_void_p(t.data_ptr())
rnd = lambda *s: torch.randn(*s, device="cuda")


def timed(name, fn, flops, nbytes, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    total = 0.0
    for _ in range(reps):
        flush.zero_()                       # evict L2 between repetitions
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        total += e0.elapsed_time(e1)
    ms = total / reps
    print("%-8s %9.1f us  %7.2f TF/s  %8.1f GB/s" % (name, 1e3 * ms, flops / ms / 1e9, nbytes / ms / 1e6))


if kind == "dense":
    m, k, n = (int(v) for v in sys.argv[2:5])
    math = fl.MATH[sys.argv[5] if len(sys.argv) > 5 else "fp32"]
    reps = int(sys.argv[6]) if len(sys.argv) > 6 else 10
    x, w, b, dz = rnd(m, k), rnd(k, n), rnd(n), rnd(m, n)
    y, dx, dw, db = torch.empty(m, n, device="cuda"), torch.empty(m, k, device="cuda"), torch.empty(k, n, device="cuda"), torch.empty(n, device="cuda")
    need = L.fnet_dense_workspace_bytes(m, k, n)
    ws = torch.empty(max(need, 256), dtype=torch.uint8, device="cuda")
    flops, nbytes = 2.0 * m * k * n, 4.0 * (m * k + k * n + m * n)
    timed("fwd", lambda: L.fnet_dense_fwd(m, k, n, 0, math, ptr(x), ptr(w), ptr(b), ptr(y), ptr(ws), ws.numel(), None), flops, nbytes, reps)
    timed("dgrad", lambda: L.fnet_dense_bwd_data(m, k, n, math, ptr(dz), ptr(w), ptr(dx), 0, ptr(ws), ws.numel(), None), flops, nbytes, reps)
    timed("wgrad", lambda: L.fnet_dense_bwd_filter(m, k, n, math, ptr(x), ptr(dz), ptr(dw), ptr(db), ptr(ws), ws.numel(), None), flops, nbytes, reps)
    timed("dgrad+act", lambda: L.fnet_dense_bwd_data_act(m, k, n, math, ptr(dz), ptr(w), ptr(dx), 0, 1, ptr(x), ptr(ws), ws.numel(), None), flops, nbytes, reps)
else:
    bsz, h, wd, c, ko, kh, kw, st = (int(v) for v in sys.argv[2:10])
    math = fl.MATH[sys.argv[10] if len(sys.argv) > 10 else "fp32"]
    reps = int(sys.argv[11]) if len(sys.argv) > 11 else 10
    ho, wo = -(-h // st), -(-wd // st)
    pad_h = max((ho - 1) * st + kh - h, 0)
    pad_w = max((wo - 1) * st + kw - wd, 0)
    d = fl.Conv2dDesc(bsz, h, wd, c, ho, wo, ko, kh, kw, st, st, pad_h // 2, pad_w // 2, 1, 1, 0, math)
    x, w, b, dz = rnd(bsz, h, wd, c), rnd(kh, kw, c, ko), rnd(ko), rnd(bsz, ho, wo, ko)
    y, dx, dw, db = torch.empty_like(dz), torch.empty_like(x), torch.empty_like(w), torch.empty_like(b)
    need = L.fnet_conv2d_workspace_bytes(C.byref(d))
    ws = torch.empty(max(need, 256), dtype=torch.uint8, device="cuda")
    flops = 2.0 * bsz * ho * wo * ko * kh * kw * c
    nbytes = 4.0 * (x.numel() + dz.numel() + w.numel())
    timed("fwd", lambda: L.fnet_conv2d_fwd(d, ptr(x), ptr(w), ptr(b), ptr(y), ptr(ws), ws.numel(), None), flops, nbytes, reps)
    timed("dgrad", lambda: L.fnet_conv2d_bwd_data(d, ptr(dz), ptr(w), ptr(dx), 0, ptr(ws), ws.numel(), None), flops, nbytes, reps)
    timed("wgrad", lambda: L.fnet_conv2d_bwd_filter(d, ptr(x), ptr(dz), ptr(dw), ptr(db), ptr(ws), ws.numel(), None), flops, nbytes, reps)
