"""Back-to-back time of the three products of the population's most frequent convolution shapes (batch 128, math auto):
40 launches captured in one CUDA graph, so the figure is the kernel's own duration with warm L2 -- the state it runs in
inside a training step -- not the ~6 us floor of an event pair around a single launch.

  python scripts/conv_sweep.py [fp32|tf32|auto]
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from featurenet_b200 import lib as fl  # noqa: E402

L = fl.get()
ptr = lambda t: C.c_void_p(t.data_ptr())
rnd = lambda *s: torch.randn(*s, device="cuda")
SHAPES = [  # h, w, c, k, kh, kw, stride   (profiles/r2_population_by_operation_fewc.txt)
    (32, 16, 24, 192, 1, 1, 2), (32, 32, 3, 24, 1, 1, 2), (16, 16, 24, 192, 1, 1, 2), (16, 8, 192, 1536, 1, 1, 2),
    (16, 8, 192, 24, 1, 1, 2), (8, 8, 192, 24, 1, 1, 2), (16, 16, 3, 24, 1, 1, 1), (32, 32, 3, 24, 2, 2, 2),
    (32, 32, 3, 12, 1, 1, 2), (4, 2, 24, 192, 1, 1, 1), (4, 2, 192, 1536, 1, 1, 2), (8, 4, 24, 192, 1, 1, 2),
    (2, 1, 96, 768, 1, 1, 2), (16, 16, 24, 3, 1, 1, 2), (8, 8, 24, 192, 1, 1, 2), (32, 32, 24, 24, 3, 3, 1),
    (16, 16, 24, 24, 3, 3, 1), (32, 32, 3, 24, 3, 3, 1), (16, 16, 192, 192, 1, 1, 1),
]


def graph_time(fn, stream, n=40):
    with torch.cuda.stream(stream):
        for _ in range(3):
            fn()
        stream.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            for _ in range(n):
                fn()
        g.replay()
        stream.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(5):
            g.replay()
        e1.record(stream)
        stream.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (5 * n)


def main():
    math = fl.MATH[sys.argv[1] if len(sys.argv) > 1 else "auto"]
    bsz = 128
    stream = torch.cuda.Stream()
    sp = C.c_void_p(stream.cuda_stream)
    print("%-34s %8s %8s %8s   (us per launch, warm L2)" % ("shape", "fwd", "dgrad", "wgrad"))
    total = [0.0, 0.0, 0.0]
    for h, wd, c, ko, kh, kw, st in SHAPES:
        ho, wo = -(-h // st), -(-wd // st)
        pad_h, pad_w = max((ho - 1) * st + kh - h, 0), max((wo - 1) * st + kw - wd, 0)
        d = fl.Conv2dDesc(bsz, h, wd, c, ho, wo, ko, kh, kw, st, st, pad_h // 2, pad_w // 2, 1, 1, 0, math)
        with torch.cuda.stream(stream):
            x, w, b, dz = rnd(bsz, h, wd, c), rnd(kh, kw, c, ko), rnd(ko), rnd(bsz, ho, wo, ko)
            y, dx, dw, db = torch.empty_like(dz), torch.empty_like(x), torch.empty_like(w), torch.empty_like(b)
            need = L.fnet_conv2d_workspace_bytes(C.byref(d))
            ws = torch.empty(max(need, 256), dtype=torch.uint8, device="cuda")
        stream.synchronize()
        t = [graph_time(lambda: L.fnet_conv2d_fwd(d, ptr(x), ptr(w), ptr(b), ptr(y), ptr(ws), ws.numel(), sp), stream),
             graph_time(lambda: L.fnet_conv2d_bwd_data(d, ptr(dz), ptr(w), ptr(dx), 0, ptr(ws), ws.numel(), sp), stream),
             graph_time(lambda: L.fnet_conv2d_bwd_filter(d, ptr(x), ptr(dz), ptr(dw), ptr(db), ptr(ws), ws.numel(), sp), stream)]
        total = [a + v for a, v in zip(total, t)]
        mb = 4e-6 * (x.numel() + dz.numel() + w.numel())
        print("%-34s %8.1f %8.1f %8.1f   %.1f MB  %.2f GFLOP" % ("%dx%dx%d->%d k%dx%d s%d" % (h, wd, c, ko, kh, kw, st), t[0], t[1], t[2], mb,
                                                       2e-9 * bsz * ho * wo * ko * kh * kw * c))
    print("%-34s %8.1f %8.1f %8.1f" % ("sum", *total))


if __name__ == "__main__":
    main()
