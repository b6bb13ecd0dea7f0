"""Size-independent properties of the GEMM-shaped kernels at the FULL sizes of BASELINE.json's configs.

The float64 oracle takes minutes on batch-128 layers, so at full size the kernels are checked
through identities that hold for any size and involve all three products of a layer:

* adjointness: <conv(x), dy> == <x, dgrad(dy)> == <w, wgrad(x, dy)>  (bias: sum(dy) == dbias),
  accumulated on the host in float64 from the device results;
* linearity: conv(x1 + x2) == conv(x1) + conv(x2) without bias and activation;
* determinism: two calls give bit-identical results.

The shapes are the layers of configs[0] (lenet5 on MNIST shape) and configs[1]-style CIFAR layers at
batch 128, plus the widest layers of the "keras" template; both math modes.
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from tests.test_gpu_ops import conv_desc, ptr, workspace  # noqa: E402


@pytest.fixture(scope="module")
def L():
    from featurenet_b200 import lib as fl
    return fl.get()


def dot64(a, b):
    return float((a.double() * b.double()).sum().item())


FULL_CONVS = [
    # x shape, kernel, out channels, stride
    ((128, 28, 28, 1), (5, 5), 6, (1, 1)),        # lenet5 conv1 @MNIST: GEMM (100 352, 6, 25)
    ((128, 28, 28, 6), (5, 5), 12, (1, 1)),       # lenet5 conv2: (100 352, 12, 150)
    ((128, 14, 14, 12), (5, 5), 120, (1, 1)),     # lenet5 conv3: (25 088, 120, 300); col2im data gradient
    ((128, 32, 32, 3), (5, 5), 18, (1, 1)),       # lenet5 conv1 @CIFAR: (131 072, 18, 75)
    ((128, 32, 32, 3), (1, 1), 6, (1, 1)),        # main_1block_nas product #4: 1x1 3 -> 6
    ((128, 32, 32, 32), (3, 3), 64, (1, 1)),      # keras template: (131 072, 64, 288)
    ((128, 32, 32, 64), (3, 3), 64, (2, 2)),      # stride-2 block conv, wide
    ((128, 32, 32, 64), (3, 3), 64, (1, 1)),      # keras template's widest: two channel chunks of the resident-tile kernel
]


@pytest.mark.parametrize("math,tol", [("fp32", 2e-5), ("tf32", 4e-3), ("tf32x3", 2e-5), ("auto", 2e-5)])
@pytest.mark.parametrize("case_id", range(len(FULL_CONVS)))
def test_conv_adjoint_linearity_determinism_full_size(L, case_id, math, tol):
    from featurenet_b200 import lib as fl
    xs, k, kout, s = FULL_CONVS[case_id]
    g = torch.Generator(device="cuda").manual_seed(100 + case_id)
    rnd = lambda *shape: torch.randn(*shape, device="cuda", generator=g)
    desc = conv_desc(fl, xs, k, kout, s, "same", (1, 1), None, math)
    ws = workspace(L.fnet_conv2d_workspace_bytes(C.byref(desc)))
    x, x2 = rnd(*xs), rnd(*xs)
    w, b = rnd(k[0], k[1], xs[3], kout) * 0.2, rnd(kout)
    y_shape = (xs[0], desc.out_h, desc.out_w, kout)
    dy = rnd(*y_shape)

    def fwd(inp, bias):
        out = torch.empty(y_shape, device="cuda")
        L.fnet_conv2d_fwd(desc, ptr(inp), ptr(w), ptr(bias) if bias is not None else None, ptr(out), ptr(ws), ws.numel(), None)
        return out

    y = fwd(x, None)
    dx, dw, db = torch.empty_like(x), torch.empty_like(w), torch.empty_like(b)
    L.fnet_conv2d_bwd_data(desc, ptr(dy), ptr(w), ptr(dx), 0, ptr(ws), ws.numel(), None)
    L.fnet_conv2d_bwd_filter(desc, ptr(x), ptr(dy), ptr(dw), ptr(db), ptr(ws), ws.numel(), None)
    torch.cuda.synchronize()

    # adjointness: three ways to compute the same bilinear form
    ref = dot64(y, dy)
    norm = float(y.double().norm() * dy.double().norm())
    assert abs(dot64(x, dx) - ref) <= tol * norm, "data gradient is not the adjoint of the forward"
    assert abs(dot64(w, dw) - ref) <= tol * norm, "filter gradient is not the adjoint of the forward"
    assert torch.allclose(db.double(), dy.double().sum(dim=(0, 1, 2)), rtol=0, atol=tol * float(dy.double().abs().sum(dim=(0, 1, 2)).max()))

    # bias enters additively
    yb = fwd(x, b)
    assert float((yb - y - b).abs().max()) <= 1e-5 * float(y.abs().max())

    # linearity (exact in fp32 up to rounding of the two sums; tf32 truncates x1 + x2 differently)
    ysum = fwd(x + x2, None)
    y2 = fwd(x2, None)
    assert float((ysum - (y + y2)).abs().max()) <= max(tol, 1e-5) * float(ysum.abs().max())

    # determinism
    dw2, db2, dx2 = torch.empty_like(w), torch.empty_like(b), torch.empty_like(x)
    L.fnet_conv2d_bwd_filter(desc, ptr(x), ptr(dy), ptr(dw2), ptr(db2), ptr(ws), ws.numel(), None)
    L.fnet_conv2d_bwd_data(desc, ptr(dy), ptr(w), ptr(dx2), 0, ptr(ws), ws.numel(), None)
    assert torch.equal(dw, dw2) and torch.equal(db, db2) and torch.equal(dx, dx2) and torch.equal(y, fwd(x, None))
    # without the bias row the split plan may differ: same values up to rounding, and the workspace query covers it
    L.fnet_conv2d_bwd_filter(desc, ptr(x), ptr(dy), ptr(dw2), None, ptr(ws), ws.numel(), None)
    assert float((dw - dw2).abs().max()) <= max(tol, 1e-5) * float(dw.abs().max())


FULL_DENSE = [(25088, 120, 84),        # lenet5 Dense(84) on the rank-4 tensor @MNIST
              (128, 16464, 10),        # lenet5 head @MNIST
              (128, 21504, 10),        # lenet5 head @CIFAR
              (131072, 64, 512),       # keras template Dense(512) on 32x32 positions
              (128, 524288, 10)]       # keras template head (streaming kernels)


@pytest.mark.parametrize("math,tol", [("fp32", 2e-5), ("tf32", 4e-3), ("tf32x3", 2e-5), ("auto", 2e-5)])
@pytest.mark.parametrize("m,k,n", FULL_DENSE)
def test_dense_adjoint_full_size(L, m, k, n, math, tol):
    from featurenet_b200 import lib as fl
    mode = fl.MATH[math]
    g = torch.Generator(device="cuda").manual_seed(m + k + n)
    rnd = lambda *shape: torch.randn(*shape, device="cuda", generator=g)
    x, w, dy = rnd(m, k), rnd(k, n) / np.sqrt(k), rnd(m, n)
    ws = workspace(L.fnet_dense_workspace_bytes(m, k, n))
    y, dx, dw, db = torch.empty(m, n, device="cuda"), torch.empty(m, k, device="cuda"), torch.empty(k, n, device="cuda"), torch.empty(n, device="cuda")
    L.fnet_dense_fwd(m, k, n, 0, mode, ptr(x), ptr(w), None, ptr(y), ptr(ws), ws.numel(), None)
    L.fnet_dense_bwd_data(m, k, n, mode, ptr(dy), ptr(w), ptr(dx), 0, ptr(ws), ws.numel(), None)
    L.fnet_dense_bwd_filter(m, k, n, mode, ptr(x), ptr(dy), ptr(dw), ptr(db), ptr(ws), ws.numel(), None)
    torch.cuda.synchronize()
    ref = dot64(y, dy)
    norm = float(y.double().norm() * dy.double().norm())
    assert abs(dot64(x, dx) - ref) <= tol * norm
    assert abs(dot64(w, dw) - ref) <= tol * norm
    assert torch.allclose(db.double(), dy.double().sum(dim=0), rtol=0, atol=tol * float(dy.double().abs().sum(dim=0).max()))
    # against a library GEMM of the same size (fp32 reference of the same op)
    want = x.double() @ w.double()
    assert float((y.double() - want).abs().max()) <= (5e-3 if math == "tf32" else tol) * float(want.abs().max())


GUARD = 4096      # floats on each side of every output


def guarded(shape):
    """An output tensor embedded in a larger allocation whose margins hold a sentinel."""
    n = int(np.prod(shape))
    whole = torch.full((n + 2 * GUARD,), 12345.0, device="cuda")
    return whole, whole[GUARD: GUARD + n].view(*shape)


def intact(whole, n):
    return bool((whole[:GUARD] == 12345.0).all().item()) and bool((whole[GUARD + n:] == 12345.0).all().item())


@pytest.mark.parametrize("math", ["fp32", "tf32", "tf32x3", "auto"])
@pytest.mark.parametrize("xs,k,kout,s", [((3, 9, 7, 5), (3, 3), 7, (2, 2)), ((5, 13, 11, 33), (3, 3), 130, (1, 1)),
                                         ((130, 6, 6, 4), (1, 1), 16, (1, 1)), ((67, 16, 16, 24), (1, 1), 100, (1, 1)),
                                         ((40, 14, 14, 12), (5, 5), 120, (1, 1)), ((2, 5, 5, 2), (7, 7), 4, (1, 1)),
                                         ((19, 16, 16, 36), (1, 1), 5, (1, 1)), ((160, 16, 16, 24), (3, 3), 40, (1, 1))])
def test_no_kernel_writes_outside_its_output(L, xs, k, kout, s, math):
    """compute-sanitizer is closed on this pool: out-of-bounds WRITES are looked for with guard bands around every
    output of the three products, on ragged shapes that exercise every tile edge, in all math modes."""
    from featurenet_b200 import lib as fl
    g = torch.Generator(device="cuda").manual_seed(7)
    rnd = lambda *shape: torch.randn(*shape, device="cuda", generator=g)
    desc = conv_desc(fl, xs, k, kout, s, "same", (1, 1), "relu", math)
    ws_whole, ws = guarded((max(int(L.fnet_conv2d_workspace_bytes(C.byref(desc))) // 4, 64),))
    ws_bytes = ws.numel() * 4
    x, w, b = rnd(*xs), rnd(k[0], k[1], xs[3], kout) * 0.2, rnd(kout)
    y_shape = (xs[0], desc.out_h, desc.out_w, kout)
    dy = rnd(*y_shape)
    yw, y = guarded(y_shape)
    dxw, dx = guarded(xs)
    dww, dw = guarded(tuple(w.shape))
    dbw, db = guarded((kout,))
    L.fnet_conv2d_fwd(desc, ptr(x), ptr(w), ptr(b), ptr(y), ptr(ws), ws_bytes, None)
    L.fnet_conv2d_bwd_data(desc, ptr(dy), ptr(w), ptr(dx), 0, ptr(ws), ws_bytes, None)
    L.fnet_conv2d_bwd_data_act(desc, ptr(dy), ptr(w), ptr(dx), 1, 2, ptr(x), ptr(ws), ws_bytes, None)
    L.fnet_conv2d_bwd_filter(desc, ptr(x), ptr(dy), ptr(dw), ptr(db), ptr(ws), ws_bytes, None)
    torch.cuda.synchronize()
    for whole, t in ((yw, y), (dxw, dx), (dww, dw), (dbw, db), (ws_whole, ws)):
        assert intact(whole, t.numel())
    assert bool(torch.isfinite(y).all()) and bool(torch.isfinite(dx).all()) and bool(torch.isfinite(dw).all())
