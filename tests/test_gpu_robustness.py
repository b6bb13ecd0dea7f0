"""Robustness path on the GPU: the inference-mode input gradient, the attack-step kernel and the two attacks
against the float64 oracle (oracle/attacks.py), on small graphs that contain BatchNorm, Dropout and merges."""
import ctypes as C

import numpy as np
import pytest

from oracle import attacks
from oracle.net import OracleNet

from tests.util import init_weights, tiny_graphs
from tests.test_gpu_network import engine_for, relerr

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def setup(name, seed=3):
    net = tiny_graphs()[name]()
    rng = np.random.default_rng(seed)
    weights = init_weights(net, rng, jitter=True)
    weights32 = [[w.astype(np.float32).astype(np.float64) for w in ws] for ws in weights]
    eng = engine_for(net, weights)
    return net, eng, OracleNet(net.describe(), weights32)


@pytest.mark.parametrize("name", sorted(tiny_graphs()))
def test_input_gradient_matches_the_oracle(name):
    net, eng, oracle = setup(name)
    rng = np.random.default_rng(11)
    x = rng.random((6,) + tuple(net.input_shape[1:])).astype(np.float32)
    classes = net.output_shape[-1]
    t = np.eye(classes, dtype=np.float32)[rng.integers(0, classes, 6)]
    want = oracle.input_gradient(x.astype(np.float64), t.astype(np.float64))
    got = eng.input_gradient(x, t).cpu().numpy()
    assert relerr(got, want) <= 1e-4
    # a second call with a device tensor (what the attacks pass) gives the same bits
    again = eng.input_gradient(torch.from_numpy(x).cuda(), t).cpu().numpy()
    assert np.array_equal(got, again)
    # and training still works afterwards (the gradient program must not disturb the train program)
    loss = net.train_on_batch(x, t)
    assert np.isfinite(loss[0])


def test_l2_attack_step_kernel():
    from featurenet_b200 import lib as fl
    L = fl.get()
    rng = np.random.default_rng(4)
    b, row = 7, 301
    x0, g, adv = (rng.random((b, row)).astype(np.float32) for _ in range(3))
    active = np.array([1, 0, 1, 1, 0, 1, 1], np.uint8)
    dev = lambda a: torch.from_numpy(a).cuda()
    x0d, gd, advd, actd = dev(x0), dev(g), dev(adv.copy()), dev(active)
    ptr = lambda t: C.c_void_p(t.data_ptr())
    L.fnet_l2_attack_step(b, row, ptr(x0d), ptr(gd), ptr(advd), ptr(advd), 0.1, 0.25, 0.0, 0.9, ptr(actd), None)
    torch.cuda.synchronize()
    delta = (adv - x0).astype(np.float64) + 0.1 * g / (np.linalg.norm(g.astype(np.float64), axis=1, keepdims=True) + 1e-7)
    n = np.linalg.norm(delta, axis=1, keepdims=True)
    delta = delta * np.where(n > 0.25, 0.25 / (n + 1e-7), 1.0)
    want = np.clip(x0 + delta, 0.0, 0.9)
    want[active == 0] = adv[active == 0]
    assert np.abs(advd.cpu().numpy() - want).max() <= 2e-6


@pytest.mark.parametrize("attack", ["fgsm", "pgd"])
def test_attacks_match_the_oracle(attack):
    from featurenet_b200 import robustness
    name = "pools_and_bn" if "pools_and_bn" in tiny_graphs() else sorted(tiny_graphs())[0]
    net, eng, oracle = setup(name)
    rng = np.random.default_rng(17)
    x = rng.random((9,) + tuple(net.input_shape[1:])).astype(np.float32)
    params = {"norm": 2, "max_iter": 8} if attack == "pgd" else {"norm": 2}
    adv = robustness.craft(net, x, attack, params)
    want = attacks.fgsm_minimal(oracle, x) if attack == "fgsm" else attacks.pgd(oracle, x, max_iter=8)
    assert np.abs(adv - want).max() <= 1e-4
    value, _ = robustness.empirical_robustness(net, x, attack, params)
    assert abs(value - attacks.empirical_robustness(oracle, x.astype(np.float64), want)) <= 1e-4
    rob, clean, fooled = robustness.attack_robustness(net, attack, x, np.eye(net.output_shape[-1], dtype=np.float32)[
        np.argmax(net.predict(x), axis=1)], params)
    assert clean == 1.0 and 0.0 <= fooled <= 1.0 and rob >= 0.0


def test_eval_robustness_through_the_reference_named_runner():
    """tensorflow_generator.py:175-218: scores land on the model, the first requested metric becomes robustness_score."""
    import contextlib
    import io
    import featurenet_b200
    featurenet_b200.install_compat()
    from tensorflow_generator import TensorflowGenerator
    import os
    os.environ["FNET_SYNTH_TRAIN"], os.environ["FNET_SYNTH_TEST"] = "512", "256"
    with contextlib.redirect_stdout(io.StringIO()):
        gen = TensorflowGenerator("lenet5", epochs=1, dataset="mnist", data_augmentation=False, batch_size=64)
    model = gen.model
    model.accuracy = 0.9                       # the reference skips candidates below 0.5 (tensorflow_generator.py:178)
    out = io.StringIO()
    with contextlib.redirect_stdout(out):
        TensorflowGenerator.eval_robustness(model, ["fgsm", "pgd", "cw"], robustness_set_size=48)
    assert "model robustness (clever, pgd, cw, fgsm)" in out.getvalue()
    for score in (model.fgsm_score, model.pgd_score, model.cw_score):
        # 0.0 = no row changed its prediction within the attack's budget
        assert len(score) == 3 and score[0] >= 0.0 and 0.0 <= score[1] <= 1.0 and 0.0 <= score[2] <= 1.0
    assert model.robustness_score == model.fgsm_score[0]
    assert model.clever_score == 0                              # not requested
    assert model.fgsm_score[1] == model.pgd_score[1] == model.cw_score[1]      # the same clean accuracy on the same rows
    np.random.seed(3)
    with contextlib.redirect_stdout(out):
        TensorflowGenerator.eval_robustness(model, ["clever"], robustness_set_size=4)
    assert 0.0 < model.clever_score <= 2.0                      # radius r_l2 = 2 caps every per-image score
    assert model.robustness_score == model.clever_score


@pytest.mark.parametrize("name", sorted(tiny_graphs()))
def test_logit_gradient_matches_the_oracle(name):
    net, eng, oracle = setup(name)
    rng = np.random.default_rng(13)
    x = rng.random((6,) + tuple(net.input_shape[1:])).astype(np.float32)
    classes = net.output_shape[-1]
    head = rng.standard_normal((6, classes)).astype(np.float32)
    want = oracle.logit_gradient(x.astype(np.float64), head.astype(np.float64))
    grad, logits = eng.logit_gradient(x, head)
    got, z = grad.cpu().numpy(), logits.cpu().numpy()
    assert relerr(got, want) <= 1e-4
    assert relerr(z, oracle.logits) <= 1e-4
    # linear in head_dz: the CLEVER shortcut (one pass with e_pred - e_target) equals the difference of two rows
    e = np.eye(classes, dtype=np.float32)
    g0 = eng.logit_gradient(x, np.tile(e[0], (6, 1)))[0].cpu().numpy()
    g1 = eng.logit_gradient(x, np.tile(e[1], (6, 1)))[0].cpu().numpy()
    gd = eng.logit_gradient(x, np.tile(e[0] - e[1], (6, 1)))[0].cpu().numpy()
    assert np.abs(gd - (g0 - g1)).max() <= 1e-5 * max(1.0, np.abs(g0).max())
    # the loss-gradient program and training are undisturbed
    t = e[rng.integers(0, classes, 6)]
    assert relerr(eng.input_gradient(x, t).cpu().numpy(), oracle.input_gradient(x.astype(np.float64), t.astype(np.float64))) <= 1e-4
    assert np.isfinite(net.train_on_batch(x, t)[0])


def test_clever_matches_the_oracle():
    """Same numpy seed -> same pools and picks on both sides; the per-batch maxima of the gradient norm agree to
    fp32 accuracy and the score (a Nelder-Mead Weibull fit on those maxima) to 1e-3."""
    from featurenet_b200 import robustness
    name = "pools_and_bn" if "pools_and_bn" in tiny_graphs() else sorted(tiny_graphs())[0]
    net, eng, oracle = setup(name)
    rng = np.random.default_rng(19)
    xs = rng.random((3,) + tuple(net.input_shape[1:])).astype(np.float32)
    for x in xs:
        np.random.seed(77)
        want, want_maxima = attacks.clever_u(oracle, x, nb_batches=6, batch_size=4, radius=2.0, pool_factor=3)
        np.random.seed(77)
        got, maxima = robustness.clever_u(net, x, 6, 4, 2.0, norm=2, pool_factor=3, return_maxima=True)
        assert relerr(maxima, want_maxima) <= 1e-4
        assert abs(got - want) <= 1e-3 * max(1.0, abs(want))
        assert 0.0 <= got <= 2.0
    np.random.seed(5)
    avg = robustness.clever_score(net, xs, nb_batches=3, batch_size=3, radius=2, pool_factor=3)
    assert 0.0 <= avg <= 2.0
    with pytest.raises(NotImplementedError):
        robustness.clever_u(net, xs[0], 3, 3, 0.1, norm=np.inf)


def test_logits_program_matches_predict_and_the_oracle():
    name = "pools_and_bn" if "pools_and_bn" in tiny_graphs() else sorted(tiny_graphs())[0]
    net, eng, oracle = setup(name)
    x = np.random.default_rng(23).random((7,) + tuple(net.input_shape[1:])).astype(np.float32)
    z = net.logits(x)
    oracle.forward(x.astype(np.float64))
    assert relerr(z, oracle.logits) <= 1e-4
    e = np.exp(z - z.max(axis=1, keepdims=True))
    assert np.abs(e / e.sum(axis=1, keepdims=True) - net.predict(x)).max() <= 1e-6


@pytest.mark.parametrize("strong", [False, True])
def test_carlini_l2_matches_the_oracle(strong, monkeypatch):
    from featurenet_b200 import robustness
    name = "pools_and_bn" if "pools_and_bn" in tiny_graphs() else sorted(tiny_graphs())[0]
    net, eng, oracle = setup(name)
    x = np.random.default_rng(8).random((5,) + tuple(net.input_shape[1:])).astype(np.float32)
    if strong:      # the data's own range and a setting that succeeds, so that every branch of the line search runs
        monkeypatch.setattr(robustness, "CLIP_VALUES", (0.0, 1.0))
        params = {"targeted": False, "norm": 2, "initial_const": 10.0, "learning_rate": 0.1}
        want = attacks.carlini_l2(oracle, x, clip=(0.0, 1.0), initial_const=10.0, learning_rate=0.1)
    else:           # exactly what tensorflow_generator.py:151-173 asks for
        params = {"targeted": False, "norm": 2}
        want = attacks.carlini_l2(oracle, x)
    adv = robustness.craft(net, x, "cw", params)
    assert adv.dtype == np.float32 and adv.shape == x.shape
    # fp32 logits against float64 ones can flip a comparison of the line search on a rare row; the rows that took the
    # same path agree tightly, and both sides fool the model on the same rows
    close = np.abs(adv - want).reshape(5, -1).max(axis=1) <= 1e-3
    assert close.sum() >= 4, np.abs(adv - want).reshape(5, -1).max(axis=1)
    labels = np.argmax(net.predict(x), axis=1)
    fooled = np.argmax(net.predict(adv), axis=1) != labels
    fooled_want = np.argmax(oracle.forward(want)[-1], axis=1) != labels
    assert np.array_equal(fooled[close], fooled_want[close])
    if strong:
        assert fooled.any()
        value, _ = robustness.empirical_robustness(net, x, "cw", params)
        assert value > 0.0
    else:
        assert not fooled.any() and robustness.empirical_robustness(net, x, "cw", params)[0] == 0.0
