"""Per-op parity: every C-ABI entry point of libfnet.so vs the float64 CPU oracle on seeded inputs.

Bars (north_star): fp32 results within 1e-4 relative (max-norm) of the oracle -- the kernels
land around 1e-6, so the tests assert 2e-5 -- and integer work (argmax taps, dropout masks,
correct counts, gathers) bit-exact.
"""
import ctypes as C

import numpy as np
import pytest

from oracle import ops

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

TOL = 2e-5            # asserted; the stated bar is 1e-4


@pytest.fixture(scope="module")
def L():
    from featurenet_b200 import lib as fl
    return fl.get()


_KEEP = []      # device tensors made by dev() stay alive until the test ends: `ptr(dev(a))` passes a raw
                # pointer, and a temporary freed right away would be handed to the next allocation


@pytest.fixture(autouse=True)
def _release_device_tensors():
    yield
    torch.cuda.synchronize()
    del _KEEP[:]


def dev(a, dtype=np.float32):
    t = torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=dtype))).cuda()
    _KEEP.append(t)
    return t


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def host(t):
    torch.cuda.synchronize()
    return t.cpu().numpy().astype(np.float64)


def relerr(got, want):
    want = np.asarray(want, np.float64)
    scale = np.abs(want).max()
    return np.abs(np.asarray(got, np.float64) - want).max() / (scale if scale > 0 else 1.0)


def assert_close(got, want, tol=TOL, what=""):
    err = relerr(got, want)
    assert err <= tol, "%s: relative error %.3e > %.1e" % (what, err, tol)


def conv_desc(fl, x_shape, k, kout, s, padding, d=(1, 1), act=None, math="fp32"):
    B, H, W, Cin = x_shape
    ho, pt, _ = ops.same_or_valid(H, k[0], s[0], padding, d[0])
    wo, pl, _ = ops.same_or_valid(W, k[1], s[1], padding, d[1])
    return fl.Conv2dDesc(B, H, W, Cin, ho, wo, kout, k[0], k[1], s[0], s[1], pt, pl, d[0], d[1], fl.ACT[act], fl.MATH[math])


def workspace(nbytes):
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device="cuda")


CONV_CASES = [
    # x shape, kernel, out channels, stride, padding, dilation, activation
    ((4, 28, 28, 1), (5, 5), 6, (1, 1), "same", (1, 1), "tanh"),        # lenet5 first layer: N=6 tile, K=25
    ((3, 14, 14, 12), (5, 5), 120, (1, 1), "same", (1, 1), "tanh"),     # N=120 -> 64-wide tiles, K=300
    ((2, 32, 32, 3), (3, 3), 24, (2, 2), "same", (1, 1), "relu"),       # stride 2, N=24
    ((2, 9, 7, 5), (1, 5), 12, (1, 1), "same", (1, 1), "sigmoid"),      # 1xk kernel, odd extents
    ((2, 9, 7, 5), (3, 1), 7, (2, 2), "valid", (1, 1), None),           # valid padding, ragged N
    ((2, 16, 16, 18), (1, 1), 3, (2, 2), "valid", (2, 2), None),        # the "Reg_" 1x1 conv with stride/dilation
    ((1, 5, 5, 2), (7, 7), 4, (1, 1), "same", (1, 1), "relu"),          # mutated 7x7 kernel larger than the image
    ((2, 8, 8, 64), (3, 3), 96, (1, 1), "same", (1, 1), "relu"),        # wide layer
    ((130, 6, 6, 4), (3, 3), 16, (1, 1), "same", (1, 1), None),         # many rows, N=16 tile
    ((16, 14, 14, 12), (5, 5), 120, (1, 1), "same", (1, 1), "tanh"),    # few in-channels: dgrad = dense GEMM + col2im
    ((64, 32, 32, 8), (3, 3), 64, (2, 2), "same", (1, 1), "relu"),      # col2im dgrad with stride 2
    ((48, 20, 20, 4), (3, 3), 100, (1, 1), "same", (2, 2), None),       # col2im dgrad with dilation 2
    ((20, 16, 16, 192), (1, 1), 3, (1, 1), "same", (1, 1), None),       # "Reg_" 192 -> 3 at size: pointwise data gradient
    ((17, 16, 16, 36), (1, 1), 8, (1, 1), "valid", (1, 1), "relu"),     # pointwise kernels, 8 output channels, ragged pixel groups
    ((5, 30, 30, 260), (1, 1), 5, (1, 1), "same", (1, 1), "tanh"),      # pointwise kernels, > 256 channels (32 lanes per pixel)
    ((33, 11, 13, 132), (1, 1), 1, (1, 1), "same", (1, 1), "sigmoid"),  # a single output channel, 33 channel groups
    ((24, 32, 16, 24), (1, 1), 192, (2, 2), "same", (1, 1), "relu"),    # 1x1 stride 2, the NAS space's staple: subsample + plain GEMM
    ((16, 16, 8, 192), (1, 1), 264, (2, 2), "same", (1, 1), None),      # the same, wide on both sides
    ((9, 7, 5, 20), (1, 1), 33, (2, 2), "same", (1, 1), "tanh"),        # odd extents: the last row / column is kept (ceil)
    ((6, 9, 10, 7), (1, 1), 40, (1, 3), "valid", (1, 1), None),         # stride along one axis only, channels not a multiple of 4
    ((40, 32, 32, 3), (1, 1), 24, (2, 2), "same", (1, 1), "relu"),      # 3 -> 24 stride 2: first layer of most sampled products
    # very few input channels, 17..64 output channels: direct kernels over groups of eight output channels (fewc.cu)
    ((7, 16, 12, 3), (3, 1), 24, (1, 1), "same", (1, 1), "relu"),       # 9 filter rows: forward / data gradient there, filter gradient elsewhere
    ((5, 9, 9, 3), (2, 2), 24, (2, 2), "same", (1, 1), "tanh"),         # even kernel, stride 2, odd extents
    ((3, 10, 8, 1), (3, 3), 32, (1, 1), "same", (1, 1), None),          # a single input channel (MNIST-shaped first layers)
    ((6, 8, 8, 4), (1, 1), 64, (1, 1), "valid", (1, 1), "sigmoid"),     # the widest case: 4 -> 64
    ((4, 8, 6, 2), (1, 3), 20, (1, 2), "same", (1, 1), None),           # a last group of four channels only (K = 20)
    ((9, 7, 5, 3), (1, 1), 28, (1, 1), "same", (2, 2), "relu"),         # 1x1 with a dilation (ignored), K = 28
]


@pytest.mark.parametrize("case_id", range(len(CONV_CASES)))
def test_conv2d_forward_backward(L, case_id):
    from featurenet_b200 import lib as fl
    xs, k, kout, s, padding, d, act = CONV_CASES[case_id]
    rng = np.random.default_rng(1000 + case_id)
    x = rng.standard_normal(xs)
    w = rng.standard_normal(k + (xs[3], kout)) * 0.3
    b = rng.standard_normal(kout)
    desc = conv_desc(fl, xs, k, kout, s, padding, d, act)
    ws = workspace(L.fnet_conv2d_workspace_bytes(C.byref(desc)))
    z = ops.conv2d_fwd(x, w, b, s, padding, d)
    y = ops.act_fwd(z, act)
    dy = rng.standard_normal(y.shape)
    dz = ops.act_bwd(y, dy, act)
    dx, dw, db = ops.conv2d_bwd(x, w, dz, s, padding, d)

    xd, wd, bd = dev(x), dev(w), dev(b)
    yd = torch.empty(y.shape, device="cuda")
    L.fnet_conv2d_fwd(desc, ptr(xd), ptr(wd), ptr(bd), ptr(yd), ptr(ws), ws.numel(), None)
    assert_close(host(yd), y, what="conv fwd")

    dzd = dev(dz)
    dxd = torch.full(x.shape, 7.0, device="cuda")
    L.fnet_conv2d_bwd_data(desc, ptr(dzd), ptr(wd), ptr(dxd), 0, ptr(ws), ws.numel(), None)
    assert_close(host(dxd), dx, what="conv dgrad")
    L.fnet_conv2d_bwd_data(desc, ptr(dzd), ptr(wd), ptr(dxd), 1, ptr(ws), ws.numel(), None)
    assert_close(host(dxd), 2 * dx, what="conv dgrad accumulate")

    dwd, dbd = torch.empty(w.shape, device="cuda"), torch.empty(kout, device="cuda")
    L.fnet_conv2d_bwd_filter(desc, ptr(xd), ptr(dzd), ptr(dwd), ptr(dbd), ptr(ws), ws.numel(), None)
    assert_close(host(dwd), dw, what="conv wgrad")
    assert_close(host(dbd), db, what="conv bias grad")
    # same call twice -> bit-identical (fixed reduction order, no atomics)
    dw2 = torch.empty(w.shape, device="cuda")
    L.fnet_conv2d_bwd_filter(desc, ptr(xd), ptr(dzd), ptr(dw2), None, ptr(ws), ws.numel(), None)
    assert torch.equal(dwd, dw2)


@pytest.mark.parametrize("math", ["fp32", "auto"])
@pytest.mark.parametrize("case_id", [0, 2, 7, 9, 12, 16, 17, 18, 19, 20, 21, 22, 24, 26])
def test_conv2d_bwd_data_with_the_producers_activation_derivative(L, case_id, math):
    """fnet_conv2d_bwd_data_act: dx = (accumulate ? dx + v : v) * act'(y), y = output of the layer that produced x."""
    from featurenet_b200 import lib as fl
    xs, k, kout, s, padding, d, _ = CONV_CASES[case_id]
    rng = np.random.default_rng(2000 + case_id)
    w = rng.standard_normal(k + (xs[3], kout)) * 0.3
    desc = conv_desc(fl, xs, k, kout, s, padding, d, None, math)
    ws = workspace(L.fnet_conv2d_workspace_bytes(C.byref(desc)))
    ho, wo = desc.out_h, desc.out_w
    dz = rng.standard_normal((xs[0], ho, wo, kout))
    x = rng.standard_normal(xs)
    dx, _, _ = ops.conv2d_bwd(x, w, dz, s, padding, d)
    before = rng.standard_normal(xs)
    for act in ("relu", "tanh", "sigmoid"):
        y = ops.act_fwd(rng.standard_normal(xs), act)                  # what the producing layer put out
        for accumulate in (0, 1):
            want = ops.act_bwd(y, dx + (before if accumulate else 0.0), act)
            dxd = dev(before)
            L.fnet_conv2d_bwd_data_act(desc, ptr(dev(dz)), ptr(dev(w)), ptr(dxd), accumulate, fl.ACT[act], ptr(dev(y)), ptr(ws),
                                       ws.numel(), None)
            assert_close(host(dxd), want, what="dgrad epilogue %s accumulate=%d" % (act, accumulate))


@pytest.mark.parametrize("m,k,n", [(20, 300, 10), (37, 32772, 10), (5, 40000, 3), (130, 32768, 16), (9, 33001, 7)])
def test_dense_bwd_data_with_the_producers_activation_derivative(L, m, k, n):
    """fnet_dense_bwd_data_act on heads: dx = (accumulate ? dx + v : v) * act'(y); K >= 32 768 with K % 4 == 0 takes the
    four-inputs-per-thread kernel, K % 4 != 0 the one-input-per-thread kernel."""
    from featurenet_b200 import lib as fl
    rng = np.random.default_rng(m + k + n)
    w = rng.standard_normal((k, n)) / np.sqrt(n)
    dz = rng.standard_normal((m, n))
    x = rng.standard_normal((m, k))
    dx, _, _ = ops.dense_bwd(x, w, dz)
    before = rng.standard_normal((m, k))
    ws = workspace(L.fnet_dense_workspace_bytes(m, k, n))
    plain = torch.empty((m, k), device="cuda")
    L.fnet_dense_bwd_data(m, k, n, fl.MATH["fp32"], ptr(dev(dz)), ptr(dev(w)), ptr(plain), 0, ptr(ws), ws.numel(), None)
    assert_close(host(plain), dx, what="dense dgrad")
    for act in ("relu", "tanh"):
        y = ops.act_fwd(rng.standard_normal((m, k)), act)
        for accumulate in (0, 1):
            want = ops.act_bwd(y, dx + (before if accumulate else 0.0), act)
            dxd = dev(before)
            L.fnet_dense_bwd_data_act(m, k, n, fl.MATH["fp32"], ptr(dev(dz)), ptr(dev(w)), ptr(dxd), accumulate, fl.ACT[act], ptr(dev(y)),
                                      ptr(ws), ws.numel(), None)
            assert_close(host(dxd), want, what="dense dgrad epilogue %s accumulate=%d" % (act, accumulate))


def test_bf16_mode_is_refused_not_aliased(L):
    """SURVEY.md names a bf16 mode; it is not built, and asking for it must fail loudly rather than run fp32 kernels."""
    from featurenet_b200 import lib as fl
    desc = conv_desc(fl, (2, 6, 6, 4), (3, 3), 5, (1, 1), "same", math="bf16")
    x, w, y = (torch.zeros(n, device="cuda") for n in (2 * 6 * 6 * 4, 3 * 3 * 4 * 5, 2 * 6 * 6 * 5))
    ws = workspace(1 << 20)
    with pytest.raises(fl.FnetError, match="not built"):
        L.fnet_conv2d_fwd(desc, ptr(x), ptr(w), None, ptr(y), ptr(ws), ws.numel(), None)
    with pytest.raises(fl.FnetError, match="not built"):
        L.fnet_dense_fwd(4, 4, 4, 0, fl.MATH["bf16"], ptr(x), ptr(w), None, ptr(y), ptr(ws), ws.numel(), None)
    from featurenet_b200.engine import Engine
    from tests.util import tiny_graphs
    with pytest.raises(NotImplementedError, match="bf16"):
        Engine(tiny_graphs()["conv_chain"](), math_mode="bf16")


def test_conv2d_softmax_activation_and_no_bias(L):
    from featurenet_b200 import lib as fl
    rng = np.random.default_rng(5)
    x, w = rng.standard_normal((2, 6, 6, 3)), rng.standard_normal((3, 3, 3, 5))
    desc = conv_desc(fl, x.shape, (3, 3), 5, (1, 1), "same", act="softmax")
    ws = workspace(L.fnet_conv2d_workspace_bytes(C.byref(desc)))
    yd = torch.empty((2, 6, 6, 5), device="cuda")
    L.fnet_conv2d_fwd(desc, ptr(dev(x)), ptr(dev(w)), None, ptr(yd), ptr(ws), ws.numel(), None)
    assert_close(host(yd), ops.act_fwd(ops.conv2d_fwd(x, w, None, (1, 1), "same"), "softmax"))


DENSE_CASES = [(128, 16464, 10, "softmax"),      # lenet5 head: split-K forward, wide wgrad
               (1568, 120, 84, "tanh"),          # Dense applied to a rank-4 tensor (rows = B*H*W)
               (64, 64, 512, "relu"), (7, 5, 3, None), (1, 1, 1, "sigmoid"), (300, 33, 17, None),
               (70, 40001, 10, "softmax"),       # long head (keras template class): streaming kernels, ragged K and rows
               (5, 33000, 3, None), (130, 32768, 16, "tanh")]


@pytest.mark.parametrize("m,k,n,act", DENSE_CASES)
def test_dense_forward_backward(L, m, k, n, act):
    from featurenet_b200 import lib as fl
    rng = np.random.default_rng(m * 7 + k)
    x, w, b = rng.standard_normal((m, k)), rng.standard_normal((k, n)) / np.sqrt(k), rng.standard_normal(n)
    y = ops.act_fwd(ops.dense_fwd(x, w, b), act)
    dz = rng.standard_normal((m, n))
    dx, dw, db = ops.dense_bwd(x, w, dz)
    ws = workspace(L.fnet_dense_workspace_bytes(m, k, n))
    xd, wd, bd, dzd = dev(x), dev(w), dev(b), dev(dz)
    yd = torch.empty((m, n), device="cuda")
    L.fnet_dense_fwd(m, k, n, fl.ACT[act], 0, ptr(xd), ptr(wd), ptr(bd), ptr(yd), ptr(ws), ws.numel(), None)
    assert_close(host(yd), y, what="dense fwd")
    dxd, dwd, dbd = torch.empty((m, k), device="cuda"), torch.empty((k, n), device="cuda"), torch.empty(n, device="cuda")
    L.fnet_dense_bwd_data(m, k, n, 0, ptr(dzd), ptr(wd), ptr(dxd), 0, ptr(ws), ws.numel(), None)
    L.fnet_dense_bwd_filter(m, k, n, 0, ptr(xd), ptr(dzd), ptr(dwd), ptr(dbd), ptr(ws), ws.numel(), None)
    assert_close(host(dxd), dx, what="dense dgrad")
    assert_close(host(dwd), dw, what="dense wgrad")
    assert_close(host(dbd), db, what="dense bias grad")


def test_workspace_too_small_is_an_error_not_a_crash(L):
    from featurenet_b200 import lib as fl
    m, k, n = 128, 16464, 24          # 24 columns: not the head kernel (<= 16), one output tile, split reduction
    need = L.fnet_dense_workspace_bytes(m, k, n)
    assert need > 0
    x, w, y = dev(np.zeros((m, k))), dev(np.zeros((k, n))), torch.empty((m, n), device="cuda")
    with pytest.raises(fl.FnetError, match="workspace"):
        L.fnet_dense_fwd(m, k, n, 0, 0, ptr(x), ptr(w), None, ptr(y), None, 0, None)
    with pytest.raises(fl.FnetError):
        L.fnet_dense_fwd(m, -1, n, 0, 0, ptr(x), ptr(w), None, ptr(y), None, 0, None)
    with pytest.raises(fl.FnetError, match="math"):
        L.fnet_dense_fwd(m, k, n, 0, 9, ptr(x), ptr(w), None, ptr(y), None, 0, None)


@pytest.mark.parametrize("xs,k,s,act", [((3, 9, 8, 5), (3, 3), (1, 1), "relu"), ((2, 16, 16, 40), (5, 1), (2, 2), None),
                                        ((2, 7, 7, 3), (2, 2), (2, 2), "tanh"), ((1, 6, 6, 33), (7, 7), (1, 1), "sigmoid"),
                                        # channels % 4 == 0: the four-channels-per-thread kernels
                                        ((4, 12, 10, 16), (3, 3), (1, 1), "relu"), ((3, 9, 9, 8), (3, 3), (2, 2), None),
                                        ((2, 5, 6, 64), (5, 5), (1, 1), "tanh"), ((3, 8, 7, 12), (1, 3), (1, 1), None),
                                        ((2, 9, 6, 20), (1, 5), (2, 1), "relu"), ((5, 4, 4, 4), (1, 1), (1, 1), None),
                                        # stride 1, up to nine taps: four destination pixels per thread (ragged strips, all tap shapes)
                                        ((2, 11, 9, 8), (3, 1), (1, 1), None), ((3, 6, 5, 12), (2, 2), (1, 1), "relu"),
                                        ((2, 33, 31, 24), (3, 3), (1, 1), "tanh"), ((1, 3, 2, 4), (3, 3), (1, 1), None),
                                        # 25 taps and 1 x 5 / 5 x 1 in strips (forward, data gradient and filter gradient)
                                        ((3, 14, 13, 16), (5, 5), (1, 1), "relu"), ((2, 9, 11, 8), (1, 5), (1, 1), None),
                                        ((2, 10, 7, 12), (5, 1), (1, 1), "tanh")])
def test_depthwise_forward_backward(L, xs, k, s, act):
    from featurenet_b200 import lib as fl
    rng = np.random.default_rng(xs[1] * 13 + k[0])
    x, w, b = rng.standard_normal(xs), rng.standard_normal(k + (xs[3], 1)), rng.standard_normal(xs[3])
    desc = conv_desc(fl, xs, k, xs[3], s, "same", act=act)
    y = ops.act_fwd(ops.dwconv2d_fwd(x, w, b, s, "same"), act)
    dz = rng.standard_normal(y.shape)
    dx, dw, db = ops.dwconv2d_bwd(x, w, dz, s, "same")
    xd, wd, bd, dzd = dev(x), dev(w), dev(b), dev(dz)
    yd = torch.empty(y.shape, device="cuda")
    L.fnet_dwconv2d_fwd(desc, ptr(xd), ptr(wd), ptr(bd), ptr(yd), None)
    assert_close(host(yd), y, what="dw fwd")
    dxd = torch.empty(xs, device="cuda")
    L.fnet_dwconv2d_bwd_data(desc, ptr(dzd), ptr(wd), ptr(dxd), 0, None)
    assert_close(host(dxd), dx, what="dw dgrad")
    L.fnet_dwconv2d_bwd_data(desc, ptr(dzd), ptr(wd), ptr(dxd), 1, None)
    assert_close(host(dxd), 2 * dx, what="dw dgrad accumulate")
    L.fnet_dwconv2d_bwd_data(desc, ptr(dzd), ptr(wd), ptr(dxd), 0, None)
    ws = workspace(L.fnet_dwconv2d_workspace_bytes(C.byref(desc)))
    dwd, dbd = torch.empty(w.shape, device="cuda"), torch.empty(xs[3], device="cuda")
    L.fnet_dwconv2d_bwd_filter(desc, ptr(xd), ptr(dzd), ptr(dwd), ptr(dbd), ptr(ws), ws.numel(), None)
    assert_close(host(dwd), dw, what="dw wgrad")
    assert_close(host(dbd), db, what="dw bias grad")


@pytest.mark.parametrize("xs,k,s", [((3, 28, 28, 6), (2, 2), (1, 1)), ((2, 28, 28, 12), (2, 2), (2, 2)),
                                    ((2, 9, 11, 5), (3, 3), (2, 2)), ((2, 7, 7, 3), (3, 3), (3, 3)),
                                    ((2, 8, 8, 4), (1, 1), (2, 2)), ((1, 5, 5, 70), (3, 2), (1, 2)),
                                    # channels % 4 == 0: the four-channels-per-thread kernels
                                    ((5, 16, 16, 32), (2, 2), (1, 1)), ((3, 10, 14, 24), (3, 3), (2, 2)),
                                    ((2, 6, 5, 8), (3, 3), (1, 1)),
                                    # the strip kernels (2x2 / 3x3, stride 1 / 2): odd extents, ragged last strips
                                    ((4, 13, 9, 16), (2, 2), (2, 2)), ((2, 17, 19, 4), (3, 3), (1, 1)),
                                    ((3, 12, 10, 20), (2, 2), (1, 1)), ((2, 15, 15, 8), (3, 3), (2, 2)),
                                    ((2, 3, 2, 8), (3, 3), (1, 1)), ((2, 5, 1, 4), (2, 2), (2, 2))])
def test_pooling(L, xs, k, s):
    from featurenet_b200 import lib as fl
    rng = np.random.default_rng(xs[1] + 31 * k[0])
    x = np.round(rng.standard_normal(xs) * 4) / 4          # coarse values -> ties occur and must resolve identically
    ho, pt, _ = ops.same_or_valid(xs[1], k[0], s[0], "same")
    wo, pl, _ = ops.same_or_valid(xs[2], k[1], s[1], "same")
    xd = dev(x)
    for kind, name in ((fl.POOL_MAX, "max"), (fl.POOL_AVG, "avg")):
        desc = fl.Pool2dDesc(xs[0], xs[1], xs[2], xs[3], ho, wo, k[0], k[1], s[0], s[1], pt, pl, kind)
        yd = torch.empty((xs[0], ho, wo, xs[3]), device="cuda")
        argd = torch.empty((xs[0], ho, wo, xs[3]), dtype=torch.uint8, device="cuda")
        L.fnet_pool2d_fwd(desc, ptr(xd), ptr(yd), ptr(argd) if name == "max" else None, None)
        dy = rng.standard_normal((xs[0], ho, wo, xs[3]))
        dxd = torch.empty(xs, device="cuda")
        L.fnet_pool2d_bwd(desc, ptr(dev(dy)), ptr(argd) if name == "max" else None, ptr(dxd), 0, None)
        if name == "max":
            y, arg = ops.maxpool_fwd(x, k, s, "same")
            assert np.array_equal(host(yd), y)
            torch.cuda.synchronize()
            assert np.array_equal(argd.cpu().numpy(), arg.astype(np.uint8))        # bit-exact tap index
            want = ops.maxpool_bwd(xs, arg, dy, k, s, "same")
            assert_close(host(dxd), want, what="maxpool bwd")
            L.fnet_pool2d_bwd(desc, ptr(dev(dy)), ptr(argd), ptr(dxd), 1, None)       # second gradient piece: accumulate
            assert_close(host(dxd), 2 * want, what="maxpool bwd accumulate")
        else:
            assert_close(host(yd), ops.avgpool_fwd(x, k, s, "same"), what="avgpool fwd")
            assert_close(host(dxd), ops.avgpool_bwd(xs, dy, k, s, "same"), what="avgpool bwd")


def test_global_average_pooling(L):
    rng = np.random.default_rng(2)
    x, dy = rng.standard_normal((3, 7, 5, 37)), rng.standard_normal((3, 37))
    yd, dxd = torch.empty((3, 37), device="cuda"), torch.ones((3, 7, 5, 37), device="cuda")
    L.fnet_gap_fwd(3, 35, 37, ptr(dev(x)), ptr(yd), None)
    L.fnet_gap_bwd(3, 35, 37, ptr(dev(dy)), ptr(dxd), 1, None)
    assert_close(host(yd), ops.gap_fwd(x))
    assert_close(host(dxd), 1.0 + ops.gap_bwd(x.shape, dy))


@pytest.mark.parametrize("shape,axis", [((16, 32, 32, 24), 1), ((128, 14, 14, 12), 1), ((3, 5, 2, 2), 1), ((64, 100), 1),
                                        ((4, 6, 5, 3), 1), ((9, 7, 3, 10), 1)])     # inner extents 15 and 30: the scalar kernels
def test_batchnorm_axis1(L, shape, axis):
    rng = np.random.default_rng(shape[0])
    x = rng.standard_normal(shape) * 2.0 + 3.0             # mean >> std: the single-pass variance must survive
    a = shape[axis]
    outer, inner = shape[0], int(np.prod(shape[2:]))
    gamma, beta = rng.standard_normal(a), rng.standard_normal(a)
    mm, mv = rng.standard_normal(a), rng.random(a) + 0.5
    y, mean, var, var_u = ops.bn_train_fwd(x, gamma, beta, axis)
    dy = rng.standard_normal(shape)
    dx, dgamma, dbeta = ops.bn_train_bwd(x, gamma, mean, var, dy, axis)
    ws = workspace(L.fnet_bn_workspace_bytes(a))
    xd, gd, bd, mmd, mvd = dev(x), dev(gamma), dev(beta), dev(mm), dev(mv)
    yd = torch.empty(shape, device="cuda")
    sm, sr = torch.empty(a, device="cuda"), torch.empty(a, device="cuda")
    L.fnet_bn_fwd_train(outer, a, inner, 1e-3, 0.99, ptr(xd), ptr(gd), ptr(bd), ptr(yd), ptr(sm), ptr(sr), ptr(mmd),
                        ptr(mvd), ptr(ws), ws.numel(), None)
    assert_close(host(yd), y, what="bn train fwd")
    assert_close(host(sm), mean, what="bn mean")
    assert_close(host(sr), 1 / np.sqrt(var + 1e-3), what="bn rstd")
    assert_close(host(mmd), ops.bn_moving_update(mm, mean), what="moving mean")
    assert_close(host(mvd), ops.bn_moving_update(mv, var_u), what="moving var")
    dxd, dgd, dbd = torch.empty(shape, device="cuda"), torch.empty(a, device="cuda"), torch.empty(a, device="cuda")
    L.fnet_bn_bwd(outer, a, inner, ptr(xd), ptr(dev(dy)), ptr(gd), ptr(sm), ptr(sr), ptr(dxd), ptr(dgd), ptr(dbd), 0,
                  ptr(ws), ws.numel(), None)
    assert_close(host(dxd), dx, tol=5e-5, what="bn dx")
    assert_close(host(dgd), dgamma, what="bn dgamma")
    assert_close(host(dbd), dbeta, what="bn dbeta")
    # x as the output of an activated layer below: the derivative of that activation (taken from x itself) rides along,
    # also on top of what the buffer already holds
    from featurenet_b200 import lib as fl
    for act in ("relu", "tanh", "sigmoid"):
        for accumulate in (0, 1):
            before = np.full(shape, 0.25)
            dxa = dev(before)
            L.fnet_bn_bwd_act(outer, a, inner, ptr(xd), ptr(dev(dy)), ptr(gd), ptr(sm), ptr(sr), ptr(dxa), ptr(dgd), ptr(dbd),
                              accumulate, fl.ACT[act], ptr(ws), ws.numel(), None)
            want = ops.act_bwd(x.astype(np.float32).astype(np.float64), dx + (before if accumulate else 0.0), act)
            assert_close(host(dxa), want, tol=5e-5, what="bn dx with %s' accumulate=%d" % (act, accumulate))
    yi = torch.empty(shape, device="cuda")
    L.fnet_bn_fwd_infer(outer, a, inner, 1e-3, ptr(xd), ptr(gd), ptr(bd), ptr(dev(mm)), ptr(dev(mv)), ptr(yi), None)
    assert_close(host(yi), ops.bn_infer_fwd(x, gamma, beta, mm, mv, axis), what="bn infer")


@pytest.mark.parametrize("act", ["relu", "tanh", "sigmoid", "softmax", None])
@pytest.mark.parametrize("rows,cols", [(1000, 10), (37, 129), (5, 3)])
def test_activations(L, act, rows, cols):
    from featurenet_b200 import lib as fl
    rng = np.random.default_rng(rows)
    x, dy = rng.standard_normal((rows, cols)) * 3, rng.standard_normal((rows, cols))
    y = ops.act_fwd(x, act)
    yd, dxd = torch.empty((rows, cols), device="cuda"), torch.zeros((rows, cols), device="cuda")
    L.fnet_act_fwd(fl.ACT[act], rows, cols, ptr(dev(x)), ptr(yd), None)
    assert_close(host(yd), y, what="act fwd")
    L.fnet_act_bwd(fl.ACT[act], rows, cols, ptr(dev(y)), ptr(dev(dy)), ptr(dxd), 0, None)
    assert_close(host(dxd), ops.act_bwd(y, dy, act), what="act bwd")


@pytest.mark.parametrize("n,rate", [(100003, 0.05), (4096, 0.5), (7, 0.7), (1 << 20, 0.3)])
def test_dropout_mask_is_bit_exact_and_recomputed(L, n, rate):
    rng = np.random.default_rng(n)
    x = rng.standard_normal(n).astype(np.float32)
    seed, stream_id, step = 0x1234ABCD5678EF01, 11, 42
    state = dev(np.array([step, 0, 0, 0, 0, 0, 0, 0]), np.int32)
    yd = torch.empty(n, device="cuda")
    L.fnet_dropout_fwd(n, rate, seed, stream_id, ptr(state), ptr(dev(x)), ptr(yd), None)
    mask = ops.dropout_keep_mask(n, rate, seed, stream_id, step)
    got = host(yd)
    assert np.array_equal(got != 0, (mask != 0) & (x != 0))                 # the same units survive
    want32 = np.where(mask != 0, x / np.float32(1.0 - np.float32(rate)), np.float32(0)).astype(np.float32)
    assert np.array_equal(got.astype(np.float32), want32)                   # and fp32 division matches bit for bit
    dxd = torch.ones(n, device="cuda")
    L.fnet_dropout_bwd(n, rate, seed, stream_id, ptr(state), ptr(dev(x)), ptr(dxd), 1, None)
    assert np.array_equal(host(dxd).astype(np.float32), (want32 + np.float32(1)).astype(np.float32))
    state2 = dev(np.array([step + 1, 0, 0, 0, 0, 0, 0, 0]), np.int32)
    y2 = torch.empty(n, device="cuda")
    L.fnet_dropout_fwd(n, rate, seed, stream_id, ptr(state2), ptr(dev(x)), ptr(y2), None)
    if n > 100:
        assert not torch.equal(y2, yd)                                       # a new step draws a new mask


def test_merges_copies_and_padding(L):
    rng = np.random.default_rng(9)
    a, b = rng.standard_normal((3, 4, 5, 6)), rng.standard_normal((3, 4, 5, 6))
    n = a.size
    ad, bd = dev(a), dev(b)
    yd = torch.empty(a.shape, device="cuda")
    L.fnet_add_fwd(n, ptr(ad), ptr(bd), ptr(yd), None)
    assert_close(host(yd), a + b)
    L.fnet_mul_fwd(n, ptr(ad), ptr(bd), ptr(yd), None)
    assert_close(host(yd), a * b)
    L.fnet_mul_bwd(n, ptr(ad), ptr(bd), ptr(yd), 1, None)
    assert_close(host(yd), 2 * a * b)
    L.fnet_accumulate(n, ptr(ad), ptr(yd), 0, None)
    assert np.array_equal(host(yd), a.astype(np.float32).astype(np.float64))
    # Concatenate(axis=1) of NHWC = stacking along H, forward and backward
    c = rng.standard_normal((3, 2, 5, 6))
    cat = torch.empty((3, 6, 5, 6), device="cuda")
    L.fnet_copy_rows(3, 4, 0, 6, 0, 4, 30, ptr(ad), ptr(cat), 0, None)
    L.fnet_copy_rows(3, 2, 0, 6, 4, 2, 30, ptr(dev(c)), ptr(cat), 0, None)
    want = ops.concat_fwd(a, c, 1)
    assert np.array_equal(host(cat), want.astype(np.float32).astype(np.float64))
    back = torch.empty((3, 2, 5, 6), device="cuda")
    L.fnet_copy_rows(3, 6, 4, 2, 0, 2, 30, ptr(cat), ptr(back), 0, None)
    assert np.array_equal(host(back), c.astype(np.float32).astype(np.float64))
    # the one-launch versions
    cd = dev(c)
    cat2 = torch.empty((3, 6, 5, 6), device="cuda")
    L.fnet_concat_rows_fwd(3, 4, 2, 30, ptr(ad), ptr(cd), ptr(cat2), None)
    assert torch.equal(cat2, cat)
    ga, gc = torch.ones((3, 4, 5, 6), device="cuda"), torch.full((3, 2, 5, 6), 9.0, device="cuda")
    L.fnet_concat_rows_bwd(3, 4, 2, 30, ptr(cat), ptr(ga), ptr(gc), 1, 0, None)
    assert np.array_equal(host(ga), (a.astype(np.float32) + np.float32(1)).astype(np.float64))
    assert np.array_equal(host(gc), c.astype(np.float32).astype(np.float64))
    L.fnet_concat_rows_bwd(3, 4, 2, 30, ptr(cat), None, ptr(gc), 0, 1, None)        # only one side wants a gradient
    assert np.array_equal(host(gc), (2 * c.astype(np.float32)).astype(np.float64))
    pd = torch.empty((3, 8, 11, 6), device="cuda")
    L.fnet_pad2d_fwd(3, 4, 5, 6, 2, 3, ptr(ad), ptr(pd), None)
    assert np.array_equal(host(pd), ops.pad2d_fwd(a, (2, 3)).astype(np.float32).astype(np.float64))
    un = torch.zeros(a.shape, device="cuda")
    L.fnet_pad2d_bwd(3, 4, 5, 6, 2, 3, ptr(pd), ptr(un), 0, None)
    assert np.array_equal(host(un), a.astype(np.float32).astype(np.float64))
    L.fnet_fill(n, 2.5, ptr(un), None)
    assert np.all(host(un) == 2.5)


@pytest.mark.parametrize("batch,classes", [(128, 10), (96, 10), (500, 100), (3, 2)])
def test_softmax_cross_entropy_and_accuracy(L, batch, classes):
    rng = np.random.default_rng(batch)
    logits = rng.standard_normal((batch, classes)) * 4
    logits[0, :] = 0.0                                      # an exact tie: argmax must pick index 0
    logits[1, 0] = 60.0                                     # a saturated row: the clip is active
    p = ops.act_fwd(logits, "softmax").astype(np.float32).astype(np.float64)
    labels = rng.integers(0, classes, batch)
    labels[1] = 1
    t = np.eye(classes)[labels]
    loss, correct = ops.cce_fwd(p, t)
    dz = ops.act_bwd(p, ops.cce_bwd_probs(p, t, 1.0 / batch), "softmax")
    dzd = torch.empty((batch, classes), device="cuda")
    loss_d = torch.zeros(1, dtype=torch.float64, device="cuda")
    corr_d = torch.zeros(1, dtype=torch.int64, device="cuda")
    for _ in range(2):                                      # accumulates across calls
        L.fnet_softmax_cce(batch, classes, ptr(dev(p)), ptr(dev(t)), 1.0 / batch, 0, None, ptr(dzd), ptr(loss_d),
                           ptr(corr_d), None)
    torch.cuda.synchronize()
    assert int(corr_d.item()) == 2 * int(correct.sum())                      # integer, exact
    assert abs(loss_d.item() - 2 * loss.sum()) <= 1e-5 * abs(2 * loss.sum())
    assert_close(host(dzd), dz, tol=5e-5, what="dlogits")
    # the same from the logits: the kernel runs the softmax itself and leaves the probabilities in place
    zd = dev(logits)
    loss2, corr2 = torch.zeros(1, dtype=torch.float64, device="cuda"), torch.zeros(1, dtype=torch.int64, device="cuda")
    dz2 = torch.empty((batch, classes), device="cuda")
    L.fnet_softmax_cce(batch, classes, ptr(zd), ptr(dev(t)), 1.0 / batch, 1, ptr(zd), ptr(dz2), ptr(loss2), ptr(corr2), None)
    p64 = ops.act_fwd(logits.astype(np.float32).astype(np.float64), "softmax")
    loss64, correct64 = ops.cce_fwd(p64, t)
    assert_close(host(zd), p64, what="softmax in place")
    assert abs(loss2.item() - loss64.sum()) <= 1e-5 * abs(loss64.sum())
    assert abs(int(corr2.item()) - int(correct64.sum())) <= 1            # a near-tie may flip between fp32 and fp64
    assert_close(host(dz2), ops.act_bwd(p64, ops.cce_bwd_probs(p64, t, 1.0 / batch), "softmax"), tol=5e-5, what="dlogits from logits")


@pytest.mark.parametrize("m,k,n", [(128, 6144, 10), (37, 999, 16), (5, 16464, 3), (200, 64, 10), (1, 7, 2)])
def test_head_forward_and_loss_in_one_launch(L, m, k, n):
    """fnet_head_fwd_loss == fnet_dense_fwd (logits) + softmax + categorical cross-entropy of the oracle, twice in a row
    (the accumulators keep running, the ticket re-arms itself)."""
    rng = np.random.default_rng(m + k + n)
    x = rng.standard_normal((m, k))
    w = rng.standard_normal((k, n)) / np.sqrt(k) * 3.0
    b = rng.standard_normal(n)
    t = np.eye(n)[rng.integers(0, n, m)]
    z = ops.dense_fwd(x.astype(np.float32).astype(np.float64), w.astype(np.float32).astype(np.float64), b.astype(np.float32).astype(np.float64))
    p = ops.act_fwd(z, "softmax")
    loss, correct = ops.cce_fwd(p, t)
    dz = ops.act_bwd(p, ops.cce_bwd_probs(p, t, 1.0 / m), "softmax")
    probs, dzd = torch.empty((m, n), device="cuda"), torch.empty((m, n), device="cuda")
    metrics = torch.zeros(2, dtype=torch.int64, device="cuda")
    ticket = torch.zeros(4, dtype=torch.int32, device="cuda")
    ws = workspace(L.fnet_head_fwd_loss_workspace_bytes(m))
    xd, wd, bd, td = dev(x), dev(w), dev(b), dev(t)
    for _ in range(2):
        L.fnet_head_fwd_loss(m, k, n, ptr(xd), ptr(wd), ptr(bd), ptr(probs), ptr(td), 1.0 / m, ptr(dzd),
                             C.c_void_p(metrics.data_ptr()), C.c_void_p(metrics.data_ptr() + 8), ptr(ticket), ptr(ws), ws.numel(), None)
    torch.cuda.synchronize()
    assert_close(host(probs), p, what="probabilities")
    assert_close(host(dzd), dz, tol=5e-5, what="dlogits")
    assert abs(metrics[0:1].view(torch.float64).item() - 2 * loss.sum()) <= 1e-5 * abs(2 * loss.sum())
    assert abs(int(metrics[1].item()) - 2 * int(correct.sum())) <= 2
    assert int(ticket[0].item()) == 0
    # without a gradient (evaluation) and without a bias
    L.fnet_head_fwd_loss(m, k, n, ptr(xd), ptr(wd), None, ptr(probs), ptr(td), 1.0 / m, None,
                         C.c_void_p(metrics.data_ptr()), C.c_void_p(metrics.data_ptr() + 8), ptr(ticket), ptr(ws), ws.numel(), None)
    assert_close(host(probs), ops.act_fwd(z - b.astype(np.float32).astype(np.float64), "softmax"), what="probabilities without bias")


@pytest.mark.parametrize("outer,ra,rb,inner", [(6, 4, 4, 12), (5, 3, 2, 7), (128, 32, 32, 96)])
def test_concat_backward_with_folded_activation_derivatives(L, outer, ra, rb, inner):
    from featurenet_b200 import lib as fl
    rng = np.random.default_rng(outer + inner)
    dy = rng.standard_normal((outer, ra + rb, inner))
    ya, yb = ops.act_fwd(rng.standard_normal((outer, ra, inner)), "relu"), ops.act_fwd(rng.standard_normal((outer, rb, inner)), "tanh")
    da0, db0 = rng.standard_normal(ya.shape), rng.standard_normal(yb.shape)
    for acc_a, acc_b, act_a, act_b in ((0, 0, "relu", "tanh"), (1, 0, "relu", None), (0, 1, None, "tanh"), (1, 1, "relu", "tanh")):
        want_a = dy[:, :ra] + (da0 if acc_a else 0.0)
        want_b = dy[:, ra:] + (db0 if acc_b else 0.0)
        want_a = ops.act_bwd(ya, want_a, act_a) if act_a else want_a
        want_b = ops.act_bwd(yb, want_b, act_b) if act_b else want_b
        da, db = dev(da0), dev(db0)
        L.fnet_concat_rows_bwd_act(outer, ra, rb, inner, ptr(dev(dy)), ptr(da), ptr(db), acc_a, acc_b, fl.ACT[act_a],
                                   ptr(dev(ya)) if act_a else None, fl.ACT[act_b], ptr(dev(yb)) if act_b else None, None)
        assert_close(host(da), want_a, what="da %s" % act_a)
        assert_close(host(db), want_b, what="db %s" % act_b)
    # one operand only (the other one needs no gradient)
    db = dev(db0)
    L.fnet_concat_rows_bwd_act(outer, ra, rb, inner, ptr(dev(dy)), None, ptr(db), 0, 0, 0, None, fl.ACT["tanh"], ptr(dev(yb)), None)
    assert_close(host(db), ops.act_bwd(yb, dy[:, ra:], "tanh"), what="db alone")


def test_adam_and_sgd_follow_keras(L):
    rng = np.random.default_rng(4)
    n = 1003
    p, m, v = rng.standard_normal(n), np.zeros(n), np.zeros(n)
    pd, md, vd = dev(p), dev(m), dev(v)
    state = torch.zeros(8, dtype=torch.int32, device="cuda")
    state[2:3].view(torch.float32).fill_(1e-3)
    p32 = p.astype(np.float32).astype(np.float64)
    for t in range(1, 6):
        g = rng.standard_normal(n) * (10.0 ** rng.integers(-3, 2))
        g32 = g.astype(np.float32).astype(np.float64)
        L.fnet_step_begin(ptr(state), 0.9, 0.999, None)
        if t % 2:       # the optimizer launch can close the step itself (advance >= 0) or leave it to fnet_step_end
            L.fnet_adam_step(n, ptr(pd), ptr(dev(g)), ptr(md), ptr(vd), ptr(state), 0.9, 0.999, 1e-7, 128, None)
        else:
            L.fnet_adam_step(n, ptr(pd), ptr(dev(g)), ptr(md), ptr(vd), ptr(state), 0.9, 0.999, 1e-7, -1, None)
            L.fnet_step_end(ptr(state), 128, None)
        p32, m, v = ops.adam_step(p32, g32, m, v, t, float(np.float32(1e-3)))
        assert_close(host(pd), p32, what="adam step %d" % t)
    torch.cuda.synchronize()
    assert state[0].item() == 5 and state[1].item() == 5 * 128
    vel = np.zeros(n)
    q = p.astype(np.float32).astype(np.float64)
    qd, veld = dev(p), dev(vel)
    for _ in range(3):
        g = rng.standard_normal(n)
        L.fnet_sgd_step(n, ptr(qd), ptr(dev(g)), ptr(veld), ptr(state), 0.9, 0, -1, None)
        q, vel = ops.sgd_step(q, g.astype(np.float32).astype(np.float64), vel, float(np.float32(1e-3)), 0.9)
    assert_close(host(qd), q, what="sgd")


def test_gather_rows_and_u8_scaling_are_exact(L):
    rng = np.random.default_rng(12)
    for row in (784, 3072, 10, 7):
        src = rng.standard_normal((50, row)).astype(np.float32)
        perm = rng.permutation(50).astype(np.int32)
        state = dev(np.array([0, 13, 0, 0, 0, 0, 0, 0]), np.int32)
        dst = torch.empty((32, row), device="cuda")
        L.fnet_gather_rows(32, row, ptr(dev(perm, np.int32)), ptr(state), ptr(dev(src)), ptr(dst), None)
        torch.cuda.synchronize()
        assert np.array_equal(dst.cpu().numpy(), src[perm[13:45]])
    # inputs and targets in one launch, with the per-step lr_t of fnet_step_begin folded in
    xs, ys = rng.standard_normal((50, 3072)).astype(np.float32), rng.standard_normal((50, 10)).astype(np.float32)
    perm = rng.permutation(50).astype(np.int32)
    state = torch.zeros(8, dtype=torch.int32, device="cuda")
    state[0], state[1] = 4, 9
    state[2:3].view(torch.float32).fill_(1e-3)
    xd, yd = torch.empty((40, 3072), device="cuda"), torch.empty((40, 10), device="cuda")
    L.fnet_gather_batch(40, 3072, 10, ptr(dev(perm, np.int32)), ptr(state), ptr(dev(xs)), ptr(dev(ys)), ptr(xd), ptr(yd),
                        1, 0.9, 0.999, None)
    torch.cuda.synchronize()
    assert np.array_equal(xd.cpu().numpy(), xs[perm[9:49]]) and np.array_equal(yd.cpu().numpy(), ys[perm[9:49]])
    lr_t = float(state[3:4].view(torch.float32).item())
    b1, b2 = float(np.float32(0.9)), float(np.float32(0.999))        # the betas cross the ABI as floats
    want_lr_t = float(np.float32(1e-3)) * np.sqrt(1 - b2 ** 5) / (1 - b1 ** 5)
    assert abs(lr_t - want_lr_t) <= 1e-6 * want_lr_t and state[0].item() == 4 and state[1].item() == 9
    u8 = rng.integers(0, 256, 100001, dtype=np.uint8)
    out = torch.empty(u8.size, device="cuda")
    u8d = torch.from_numpy(u8).cuda()
    L.fnet_u8_to_f32(u8.size, ptr(u8d), 255.0, ptr(out), None)
    torch.cuda.synchronize()
    want = u8.astype("float32")
    want /= 255
    assert np.array_equal(out.cpu().numpy(), want)          # same rounding as numpy's fp32 divide


def test_empty_batches_are_accepted(L):
    from featurenet_b200 import lib as fl
    z = torch.empty(16, device="cuda")          # a valid pointer; sizes below are all zero
    L.fnet_act_fwd(1, 0, 10, ptr(z), ptr(z), None)
    L.fnet_add_fwd(0, ptr(z), ptr(z), ptr(z), None)
    L.fnet_dense_fwd(0, 5, 3, 0, 0, ptr(z), ptr(z), None, ptr(z), None, 0, None)
    L.fnet_gather_rows(0, 10, ptr(z), ptr(z), ptr(z), ptr(z), None)
    desc = fl.Pool2dDesc(0, 4, 4, 3, 4, 4, 2, 2, 1, 1, 0, 0, 0)
    L.fnet_pool2d_fwd(desc, ptr(z), ptr(z), None, None)


def test_conv_linearity_at_full_size(L):
    """Size-independent property at the BASELINE shape (B=128, 32x32x3 -> 18): conv(a*x1 + x2) = a*conv(x1) + conv(x2)."""
    from featurenet_b200 import lib as fl
    rng = np.random.default_rng(1)
    xs = (128, 32, 32, 3)
    x1, x2 = dev(rng.standard_normal(xs)), dev(rng.standard_normal(xs))
    w = dev(rng.standard_normal((5, 5, 3, 18)) * 0.1)
    desc = conv_desc(fl, xs, (5, 5), 18, (1, 1), "same")
    ws = workspace(L.fnet_conv2d_workspace_bytes(C.byref(desc)))
    outs = []
    for x in (x1, x2, 2.0 * x1 + x2):
        y = torch.empty((128, 32, 32, 18), device="cuda")
        L.fnet_conv2d_fwd(desc, ptr(x), ptr(w), None, ptr(y), ptr(ws), ws.numel(), None)
        outs.append(y)
    torch.cuda.synchronize()
    assert relerr(outs[2].cpu().numpy(), (2.0 * outs[0] + outs[1]).cpu().numpy()) < 1e-5
