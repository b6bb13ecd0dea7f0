"""The timed CPU baseline (oracle/torch_cpu.py, fp32 torch ops) computes the same step as the float64 oracle."""
import contextlib
import io

import numpy as np
import pytest

from oracle.net import OracleNet
from oracle.torch_cpu import TorchCpuNet
from tests.util import init_weights, tiny_graphs, build_template


def relerr(a, b):
    b = np.asarray(b, np.float64)
    s = np.abs(b).max()
    return np.abs(np.asarray(a, np.float64) - b).max() / (s if s > 0 else 1.0)


def keras_layout(net_t, rows):
    """Weights of the torch port back in Keras layouts, flat list."""
    out = []
    for row, ts in zip(rows, net_t.params):
        for k, t in enumerate(ts):
            a = t.detach()
            if row["class"] == "Conv2D" and k == 0:
                a = a.permute(2, 3, 1, 0)
            elif row["class"] in ("DepthwiseConv2D", "SeparableConv2D") and k == 0:
                a = a.permute(2, 3, 0, 1)
            elif row["class"] == "SeparableConv2D" and k == 1:
                a = a.permute(2, 3, 1, 0)
            out.append(a.numpy())
    return out


@pytest.mark.parametrize("name", sorted(tiny_graphs().keys()) + ["lenet5"])
def test_torch_port_matches_numpy_oracle(name):
    with contextlib.redirect_stdout(io.StringIO()):
        net = build_template("lenet5", (28, 28, 1)) if name == "lenet5" else tiny_graphs()[name]()
    rows = net.describe()
    rng = np.random.default_rng(2)
    weights = init_weights(net, rng, jitter=True)
    w32 = [[w.astype(np.float32).astype(np.float64) for w in ws] for ws in weights]
    a, b = OracleNet(rows, w32), TorchCpuNet(rows, w32, threads=2)
    classes = net.output_shape[-1]
    for step in range(2):
        x = rng.random((5,) + tuple(net.input_shape[1:])).astype(np.float32)
        t = np.eye(classes, dtype=np.float32)[rng.integers(0, classes, 5)]
        masks = {i: (rng.random((5,) + tuple(r["out_shape"][1:])) > r["cfg"]["rate"]).astype(np.float32)
                 for i, r in enumerate(rows) if r["class"] == "Dropout"}
        la, ca = a.train_step(x.astype(np.float64), t.astype(np.float64), 1e-3, masks)
        lb, cb = b.train_step(x, t, 1e-3, masks)
        assert abs(la - lb) <= 1e-4 * max(1.0, abs(la)) and ca == cb
        for wa, wb in zip([w for ws in a.weights for w in ws], keras_layout(b, rows)):
            assert relerr(wb, wa) <= 2e-4
    la, aa = a.evaluate(x.astype(np.float64), t.astype(np.float64))
    lb, ab = b.evaluate(x, t)
    assert abs(la - lb) <= 1e-4 * max(1.0, abs(la)) and aa == ab


@pytest.mark.parametrize("xs,k,kout,s,padding", [((3, 9, 7, 5), (3, 3), 7, (2, 2), "same"), ((2, 6, 6, 4), (1, 1), 5, (1, 1), "same"),
                                                 ((2, 8, 9, 3), (5, 2), 4, (1, 1), "valid"), ((4, 1, 1, 30), (1, 1), 6, (1, 1), "same")])
def test_float64_conv_products_match_numpy_oracle(xs, k, kout, s, padding):
    """conv_products64 is the checker of the full-size GPU tests: it must agree with oracle/ops.py where both run."""
    from oracle import ops
    from oracle.torch_cpu import conv_products64
    rng = np.random.default_rng(9)
    x, w, b = rng.standard_normal(xs), rng.standard_normal(k + (xs[3], kout)), rng.standard_normal(kout)
    y0 = ops.conv2d_fwd(x, w, b, s, padding)
    dy = rng.standard_normal(y0.shape)
    dx0, dw0, db0 = ops.conv2d_bwd(x, w, dy, s, padding)
    y1, dx1, dw1, db1 = conv_products64(x, w, b, dy, s, padding)
    for a, c in ((y1, y0), (dx1, dx0), (dw1, dw0), (db1, db0)):
        assert relerr(a, c) <= 1e-12


def test_float64_port_matches_numpy_oracle_closely():
    with contextlib.redirect_stdout(io.StringIO()):
        net = tiny_graphs()["pools_and_bn"]()
    import torch
    rows = net.describe()
    rng = np.random.default_rng(4)
    weights = init_weights(net, rng, jitter=True)
    a, b = OracleNet(rows, weights), TorchCpuNet(rows, weights, threads=2, dtype=torch.float64)
    x = rng.random((5,) + tuple(net.input_shape[1:]))
    t = np.eye(3)[rng.integers(0, 3, 5)]
    la, ca = a.train_step(x, t, 1e-3, {})
    lb, cb = b.train_step(x, t, 1e-3, {})
    assert abs(la - lb) <= 1e-10 * max(1.0, abs(la)) and ca == cb
    for i in range(len(rows)):
        assert relerr(b.out_nhwc(i), a.outs[i]) <= 1e-10
    for (i, k), g in b.grads_keras().items():
        assert relerr(g, a.last_grads[i][k]) <= 1e-9
    for wa, wb in zip([w for ws in a.weights for w in ws], b.weights_keras()):
        assert relerr(wb, wa) <= 1e-9
