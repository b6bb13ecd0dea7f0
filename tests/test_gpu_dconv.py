"""k x k stride-1 convolutions with the input tile resident in shared memory (csrc/dconv.cu) vs the float64 oracle.

Every case is large enough (>= 74 tiles of 128 padded positions) and aligned enough (channels % 4 == 0, <= 256 output
columns) for that kernel to be the one conv_api.cu picks for the forward product and the data gradient; the filter
gradient of the same layer runs elsewhere and is checked too, so a layer is covered end to end.
"""
import ctypes as C

import numpy as np
import pytest

from oracle import ops
from tests.test_gpu_ops import (L, _release_device_tensors, assert_close, conv_desc, dev, host, ptr,  # noqa: F401
                                workspace)

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

CASES = [
    # x shape, kernel, out channels, padding, activation
    ((12, 32, 32, 64), (3, 3), 64, "same", "relu"),       # keras template: two channel chunks, 64-wide accumulator
    ((12, 32, 32, 32), (3, 3), 32, "valid", "tanh"),      # VALID: the data gradient pads by kh - 1 on every side
    ((6, 64, 32, 24), (3, 1), 192, "same", None),         # the x8 expansion: no padded columns, 192-wide accumulator
    ((24, 14, 14, 12), (5, 5), 120, "same", "tanh"),      # lenet5 conv3: 12 channels in a 16-channel chunk (zero columns)
    ((10, 30, 34, 40), (2, 2), 36, "same", "sigmoid"),    # even kernel (SAME pads bottom / right only), 2 ragged chunks of 24
    ((40, 9, 7, 8), (1, 5), 100, "valid", "relu"),        # 1 x k kernel, odd extents, images smaller than a tile
    ((9, 33, 31, 20), (4, 3), 8, "valid", None),          # 8 output columns (16-wide MMA), even x odd kernel
    ((16, 16, 16, 192), (3, 3), 24, "same", "relu"),      # six chunks, narrow output
]
MODES = [("tf32", 5e-3), ("tf32x3", 2e-5), ("auto", 2e-5)]


@pytest.mark.parametrize("math,tol", MODES)
@pytest.mark.parametrize("case_id", range(len(CASES)))
def test_resident_tile_convolution(L, case_id, math, tol):
    from featurenet_b200 import lib as fl
    xs, k, kout, padding, act = CASES[case_id]
    if k[0] * k[1] * max(xs[3], kout) > 2048:
        tol = max(tol, 1e-4)                  # the stated fp32 bar for very long reductions (see test_gpu_tc.py)
    rng = np.random.default_rng(4000 + case_id)
    x = rng.standard_normal(xs)
    w = rng.standard_normal(k + (xs[3], kout)) / np.sqrt(k[0] * k[1] * xs[3])
    b = rng.standard_normal(kout)
    desc = conv_desc(fl, xs, k, kout, (1, 1), padding, act=act, math=math)
    ws = workspace(L.fnet_conv2d_workspace_bytes(C.byref(desc)))
    y = ops.act_fwd(ops.conv2d_fwd(x, w, b, (1, 1), padding), act)
    dz = rng.standard_normal(y.shape)
    dx, dw, db = ops.conv2d_bwd(x, w, dz, (1, 1), padding)
    xd, wd, bd, dzd = dev(x), dev(w), dev(b), dev(dz)
    yd = torch.full(y.shape, 7.0, device="cuda")
    L.fnet_conv2d_fwd(desc, ptr(xd), ptr(wd), ptr(bd), ptr(yd), ptr(ws), ws.numel(), None)
    assert_close(host(yd), y, tol=tol, what="%s fwd" % math)
    L.fnet_conv2d_fwd(desc, ptr(xd), ptr(wd), None, ptr(yd), ptr(ws), ws.numel(), None)
    assert_close(host(yd), ops.act_fwd(ops.conv2d_fwd(x, w, None, (1, 1), padding), act), tol=tol, what="%s fwd, no bias" % math)
    dxd = torch.full(xs, 7.0, device="cuda")
    L.fnet_conv2d_bwd_data(desc, ptr(dzd), ptr(wd), ptr(dxd), 0, ptr(ws), ws.numel(), None)
    assert_close(host(dxd), dx, tol=tol, what="%s dgrad" % math)
    L.fnet_conv2d_bwd_data(desc, ptr(dzd), ptr(wd), ptr(dxd), 1, ptr(ws), ws.numel(), None)
    assert_close(host(dxd), 2 * dx, tol=tol, what="%s dgrad accumulate" % math)
    dx2 = torch.full(xs, 7.0, device="cuda")
    L.fnet_conv2d_bwd_data(desc, ptr(dzd), ptr(wd), ptr(dx2), 0, ptr(ws), ws.numel(), None)
    L.fnet_conv2d_bwd_data(desc, ptr(dzd), ptr(wd), ptr(dxd), 0, ptr(ws), ws.numel(), None)
    assert torch.equal(dx2, dxd)                                       # deterministic
    # the producing layer's activation derivative folded into the epilogue
    before = rng.standard_normal(xs)
    for eact in ("relu", "tanh"):
        yprev = ops.act_fwd(rng.standard_normal(xs), eact)
        for accumulate in (0, 1):
            want = ops.act_bwd(yprev, dx + (before if accumulate else 0.0), eact)
            dxe = dev(before)
            L.fnet_conv2d_bwd_data_act(desc, ptr(dzd), ptr(wd), ptr(dxe), accumulate, fl.ACT[eact], ptr(dev(yprev)), ptr(ws),
                                       ws.numel(), None)
            assert_close(host(dxe), want, tol=tol, what="%s dgrad epilogue %s accumulate=%d" % (math, eact, accumulate))
    dwd, dbd = torch.empty(w.shape, device="cuda"), torch.empty(kout, device="cuda")
    L.fnet_conv2d_bwd_filter(desc, ptr(xd), ptr(dzd), ptr(dwd), ptr(dbd), ptr(ws), ws.numel(), None)
    assert_close(host(dwd), dw, tol=tol, what="%s wgrad" % math)
    assert_close(host(dbd), db, tol=tol, what="%s bias grad" % math)
