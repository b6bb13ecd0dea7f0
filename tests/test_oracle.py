"""Consistency checks of the CPU oracle (it has no reference vectors to pin it: see oracle/ops.py).

Forward ops are compared with independent torch-CPU float64 implementations; every backward
is compared with central finite differences of the oracle's own forward, through whole
networks built by the product's graph code, so the chain rule wiring is covered too.
"""
import contextlib
import io

import numpy as np
import pytest
import torch
import torch.nn.functional as tf

from oracle import ops
from oracle.net import OracleNet, lr_schedule

from featurenet_b200 import layers as L
from tests.util import init_weights, build_template, tiny_graphs

RNG = np.random.default_rng(7)


def nhwc(t):
    return t.permute(0, 2, 3, 1).numpy()


def nchw(a):
    return torch.from_numpy(np.ascontiguousarray(a)).permute(0, 3, 1, 2)


@pytest.mark.parametrize("k,s,padding,d", [((3, 3), (1, 1), "same", (1, 1)), ((5, 5), (2, 2), "same", (1, 1)),
                                            ((1, 5), (1, 1), "same", (1, 1)), ((2, 2), (2, 2), "same", (1, 1)),
                                            ((3, 1), (2, 2), "valid", (1, 1)), ((1, 1), (2, 2), "valid", (2, 2)),
                                            ((3, 3), (1, 1), "same", (2, 2))])
def test_conv2d_forward_matches_torch(k, s, padding, d):
    x = RNG.standard_normal((2, 9, 7, 3))
    w = RNG.standard_normal(k + (3, 4))
    b = RNG.standard_normal(4)
    got = ops.conv2d_fwd(x, w, b, s, padding, d)
    _, pt, pb = ops.same_or_valid(9, k[0], s[0], padding, d[0])
    _, pl, pr = ops.same_or_valid(7, k[1], s[1], padding, d[1])
    xt = tf.pad(nchw(x), (pl, pr, pt, pb))
    want = tf.conv2d(xt, torch.from_numpy(w).permute(3, 2, 0, 1), torch.from_numpy(b), stride=s, dilation=d)
    np.testing.assert_allclose(got, nhwc(want), rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("k,s", [((3, 3), (1, 1)), ((5, 1), (2, 2)), ((2, 2), (1, 1))])
def test_depthwise_forward_matches_torch(k, s):
    x = RNG.standard_normal((2, 8, 6, 5))
    w = RNG.standard_normal(k + (5, 1))
    b = RNG.standard_normal(5)
    got = ops.dwconv2d_fwd(x, w, b, s, "same")
    _, pt, pb = ops.same_or_valid(8, k[0], s[0], "same")
    _, pl, pr = ops.same_or_valid(6, k[1], s[1], "same")
    want = tf.conv2d(tf.pad(nchw(x), (pl, pr, pt, pb)), torch.from_numpy(w).permute(2, 3, 0, 1),
                     torch.from_numpy(b), stride=s, groups=5)
    np.testing.assert_allclose(got, nhwc(want), rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("k,s", [((2, 2), (1, 1)), ((2, 2), (2, 2)), ((3, 3), (2, 2)), ((3, 3), (3, 3)), ((1, 1), (2, 2))])
def test_pooling_forward(k, s):
    x = RNG.standard_normal((2, 7, 10, 3))
    _, pt, pb = ops.same_or_valid(7, k[0], s[0], "same")
    _, pl, pr = ops.same_or_valid(10, k[1], s[1], "same")
    got, arg = ops.maxpool_fwd(x, k, s, "same")
    want = tf.max_pool2d(tf.pad(nchw(x), (pl, pr, pt, pb), value=-np.inf), k, s)
    np.testing.assert_array_equal(got, nhwc(want))
    assert arg.min() >= 0 and arg.max() < k[0] * k[1]
    # average: in-bounds divisor == sum-pool of x divided by sum-pool of ones
    num = tf.avg_pool2d(tf.pad(nchw(x), (pl, pr, pt, pb)), k, s, divisor_override=1)
    den = tf.avg_pool2d(tf.pad(torch.ones(1, 1, 7, 10, dtype=torch.float64), (pl, pr, pt, pb)), k, s, divisor_override=1)
    np.testing.assert_allclose(ops.avgpool_fwd(x, k, s, "same"), nhwc(num / den), rtol=1e-12, atol=1e-12)


def test_maxpool_ties_go_to_first_tap():
    x = np.zeros((1, 2, 2, 1))
    y, arg = ops.maxpool_fwd(x, (2, 2), (1, 1), "same")
    assert arg[0, 0, 0, 0] == 0                  # all four equal -> row-major first
    dx = ops.maxpool_bwd(x.shape, arg, np.ones_like(y), (2, 2), (1, 1), "same")
    # windows: (0,0) covers 4 cells -> first; (0,1) covers col 1 (+pad) -> x[0,1]; (1,0) -> x[1,0]; (1,1) -> x[1,1]
    np.testing.assert_array_equal(dx[0, :, :, 0], [[1, 1], [1, 1]])


def test_same_padding_rule():
    assert ops.same_or_valid(28, 5, 1, "same") == (28, 2, 2)
    assert ops.same_or_valid(28, 2, 1, "same") == (28, 0, 1)      # extra pad goes after
    assert ops.same_or_valid(32, 3, 2, "same") == (16, 0, 1)
    assert ops.same_or_valid(7, 3, 3, "same") == (3, 1, 1)
    assert ops.same_or_valid(32, 1, 2, "valid") == (16, 0, 0)
    with pytest.raises(ValueError):
        ops.same_or_valid(2, 3, 1, "valid")


def test_bn_axis1_matches_torch_batchnorm_over_rows():
    x = RNG.standard_normal((4, 6, 5, 3))
    gamma, beta = RNG.standard_normal(6), RNG.standard_normal(6)
    y, mean, var, var_u = ops.bn_train_fwd(x, gamma, beta, axis=1)
    # torch normalises over "channels" = dim 1 of an N,C,* tensor; NHWC with axis=1 is exactly that
    want = tf.batch_norm(torch.from_numpy(x), None, None, torch.from_numpy(gamma), torch.from_numpy(beta),
                         training=True, eps=1e-3)
    np.testing.assert_allclose(y, want.numpy(), rtol=1e-10, atol=1e-10)
    n = 4 * 5 * 3
    np.testing.assert_allclose(var_u, var * n / (n - 1.001), rtol=1e-14)
    yi = ops.bn_infer_fwd(x, gamma, beta, mean, var, axis=1)
    np.testing.assert_allclose(yi, y, rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(ops.bn_moving_update(np.ones(6), mean), 0.99 + 0.01 * mean)


def test_loss_accuracy_and_clip():
    p = np.array([[0.7, 0.2, 0.1], [0.5, 0.5, 0.0], [1.0, 0.0, 0.0]])
    t = np.array([[1., 0, 0], [0, 1., 0], [0, 1., 0]])
    loss, correct = ops.cce_fwd(p, t)
    np.testing.assert_allclose(loss, [-np.log(0.7), -np.log(0.5), -np.log(1e-7)])
    np.testing.assert_array_equal(correct, [True, False, False])      # tie -> first index
    g = ops.cce_bwd_probs(p, t, 1.0)
    assert np.all(g[2] == 0)                                          # clipped entries carry no gradient


def test_adam_and_sgd_formulas():
    p, g = np.array([1.0, -2.0]), np.array([0.5, -0.25])
    p1, m1, v1 = ops.adam_step(p, g, np.zeros(2), np.zeros(2), 1, 1e-3)
    lr_t = 1e-3 * np.sqrt(1 - 0.999) / (1 - 0.9)
    np.testing.assert_allclose(m1, 0.1 * g)
    np.testing.assert_allclose(v1, 0.001 * g * g)
    np.testing.assert_allclose(p1, p - lr_t * m1 / (np.sqrt(v1) + 1e-7))
    p2, vel = ops.sgd_step(p, g, np.array([0.1, 0.1]), 0.01, momentum=0.9)
    np.testing.assert_allclose(vel, 0.09 - 0.01 * g)
    np.testing.assert_allclose(p2, p + vel)


def test_lr_schedule_breakpoints():
    assert lr_schedule(0) == 1e-3 and lr_schedule(80) == 1e-3
    assert lr_schedule(81) == pytest.approx(1e-4) and lr_schedule(121) == pytest.approx(1e-5)
    assert lr_schedule(161) == pytest.approx(1e-6) and lr_schedule(181) == pytest.approx(0.5e-6)


def test_philox_known_answers():
    # Random123 known-answer vectors for philox4x32-10
    zero = ops.philox4x32_10(np.zeros((1, 4), np.uint32), (0, 0))[0]
    assert [int(v) for v in zero] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    ones = ops.philox4x32_10(np.full((1, 4), 0xFFFFFFFF, np.uint32), (0xFFFFFFFF, 0xFFFFFFFF))[0]
    assert [int(v) for v in ones] == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    pi = ops.philox4x32_10(np.array([[0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344]], np.uint32),
                           (0xa4093822, 0x299f31d0))[0]
    assert [int(v) for v in pi] == [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]
    mask = ops.dropout_keep_mask(100000, 0.05, seed=1234, stream=3, step=9)
    assert abs(mask.mean() - 0.95) < 0.003


def numeric_grad(f, a, h=1e-6):
    g = np.zeros_like(a)
    flat, gf = a.reshape(-1), g.reshape(-1)
    for i in range(flat.size):
        old = flat[i]
        flat[i] = old + h
        up = f()
        flat[i] = old - h
        dn = f()
        flat[i] = old
        gf[i] = (up - dn) / (2 * h)
    return g


@pytest.mark.parametrize("name", sorted(tiny_graphs().keys()))
def test_backward_matches_finite_differences(name):
    with contextlib.redirect_stdout(io.StringIO()):
        net = tiny_graphs()[name]()
    rows = net.describe()
    rng = np.random.default_rng(11)
    weights = init_weights(net, rng, jitter=True)
    shape = tuple(net.input_shape[1:])
    x = rng.standard_normal((3,) + shape)
    t = np.eye(net.output_shape[-1])[rng.integers(0, net.output_shape[-1], 3)]
    oracle = OracleNet(rows, weights)
    masks = {i: (rng.random((3,) + tuple(r["out_shape"][1:])) > 0.3).astype(float)
             for i, r in enumerate(rows) if r["class"] == "Dropout"}

    def loss():
        outs = oracle.forward(x, training=True, dropout_masks=masks)
        return ops.cce_fwd(outs[-1], t)[0].mean()

    loss()
    grads = oracle.backward(t)
    checked = 0
    for i, ws in enumerate(oracle.weights):
        for k, w in enumerate(ws):
            if rows[i]["class"] == "BatchNormalization" and k >= 2:
                continue
            if w.size > 400:
                continue
            want = numeric_grad(loss, w)
            np.testing.assert_allclose(grads[i][k], want, rtol=2e-5, atol=2e-8,
                                       err_msg="%s layer %d (%s) weight %d" % (name, i, rows[i]["class"], k))
            checked += 1
    assert checked > 0
    # input gradient (needs every data-path backward, including layers without weights)
    loss()
    oracle.backward(t)
    want = numeric_grad(loss, x)
    np.testing.assert_allclose(oracle.douts[0] if oracle.douts[0] is not None else 0 * x, want, rtol=2e-5, atol=2e-8)


def test_fit_semantics_on_a_tiny_problem():
    """Loop bookkeeping: weighted epoch means, best-val_loss restore, early stop bookkeeping."""
    x_in = L.Input((4,))
    out = L.Dense(3, activation="softmax", name="out")(L.Dense(5, activation="tanh")(x_in))
    net = L.Model(inputs=x_in, outputs=out)
    rng = np.random.default_rng(3)
    x = rng.standard_normal((50, 4))
    y = np.eye(3)[(x[:, 0] > 0).astype(int) + (x[:, 1] > 0)]
    xv, yv = x[:20], y[:20]
    oracle = OracleNet(net.describe(), init_weights(net, rng))
    perms = [rng.permutation(50) for _ in range(6)]
    hist = oracle.fit(x, y, xv, yv, epochs=6, batch_size=16, permutations=perms)
    assert len(hist["loss"]) == 6 and hist["loss"][-1] < hist["loss"][0]
    best = int(np.argmin(hist["val_loss"]))
    val_loss, val_acc = oracle.evaluate(xv, yv, batch_size=32)       # best weights are installed
    assert val_loss == pytest.approx(hist["val_loss"][best], rel=1e-12)
    assert val_acc == pytest.approx(hist["val_acc"][best])
    assert oracle.t == 6 * 4                                          # ceil(50/16) steps per epoch, partial kept


def test_templates_build_for_the_oracle():
    net = build_template("lenet5", (28, 28, 1))
    rng = np.random.default_rng(0)
    oracle = OracleNet(net.describe(), init_weights(net, rng))
    out = oracle.forward(rng.random((2, 28, 28, 1)))[-1]
    assert out.shape == (2, 10)
    np.testing.assert_allclose(out.sum(axis=1), 1.0, rtol=1e-12)
