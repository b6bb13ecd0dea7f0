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
    assert "cw is not built" in out.getvalue()
    for score in (model.fgsm_score, model.pgd_score):
        # 0.0 = no row changed its prediction within the attack's budget
        assert len(score) == 3 and score[0] >= 0.0 and 0.0 <= score[1] <= 1.0 and 0.0 <= score[2] <= 1.0
    assert model.robustness_score == model.fgsm_score[0]
    assert model.cw_score == 0 and model.clever_score == 0
    assert model.fgsm_score[1] == model.pgd_score[1]            # the same clean accuracy on the same rows
