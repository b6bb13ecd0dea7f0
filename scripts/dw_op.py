"""Time the depthwise-convolution kernels on one shape (development aid): python scripts/dw_op.py B H W C k stride"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from featurenet_b200 import lib as fl
L = fl.get()
b, h, w, c, k, st = (int(v) for v in sys.argv[1:7])
ho, wo = -(-h // st), -(-w // st)
ph, pw = max((ho - 1) * st + k - h, 0), max((wo - 1) * st + k - w, 0)
d = fl.Conv2dDesc(b, h, w, c, ho, wo, c, k, k, st, st, ph // 2, pw // 2, 1, 1, 0, 0)
ptr = lambda t: C.c_void_p(t.data_ptr())
x, wt, bias, dz = (torch.randn(*s, device="cuda") for s in ((b, h, w, c), (k, k, c, 1), (c,), (b, ho, wo, c)))
y, dx, dw, db = torch.empty_like(dz), torch.empty_like(x), torch.empty_like(wt), torch.empty_like(bias)
ws = torch.empty(max(int(L.fnet_dwconv2d_workspace_bytes(C.byref(d))), 256), dtype=torch.uint8, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timed(name, fn, nbytes):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = 0.0
    for _ in range(5):
        flush.zero_(); e0.record(); fn(); e1.record(); torch.cuda.synchronize(); tot += e0.elapsed_time(e1)
    ms = tot / 5
    print("%-6s %8.1f us  %7.1f GB/s" % (name, 1e3 * ms, nbytes / ms / 1e6))
nb = 4.0 * (x.numel() + dz.numel())
timed("fwd", lambda: L.fnet_dwconv2d_fwd(d, ptr(x), ptr(wt), ptr(bias), ptr(y), None), nb)
timed("dgrad", lambda: L.fnet_dwconv2d_bwd_data(d, ptr(dz), ptr(wt), ptr(dx), 0, None), nb)
timed("wgrad", lambda: L.fnet_dwconv2d_bwd_filter(d, ptr(x), ptr(dz), ptr(dw), ptr(db), ptr(ws), ws.numel(), None), nb)
