"""Candidates that share one GPU (featurenet_b200/population.py, workers > 1) must score exactly what they score
alone: eight host threads drive libfnet at once (ctypes releases the GIL), each on its own streams and CUDA graphs.

Only an invalid or out-of-memory candidate may score 0 (tensorflow_generator.py:109-111, helpers.py:172-174);
a candidate lost to a race between threads would be a wrong fitness handed to the evolution drivers.
"""
import contextlib
import copy
import io
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_eight_resident_candidates_score_what_they_score_alone():
    from featurenet_b200 import population as pop
    from featurenet_b200 import synth
    from model.keras_model import KerasFeatureModel
    from products_tree import ProductSet
    from tensorflow_generator import TensorflowGenerator
    with contextlib.redirect_stdout(io.StringIO()):
        formatted = ProductSet(os.path.join(ROOT, "featurenet_b200/data/nas_5_5_10_pop32_seed7.pdt")).format_products()
        (x_train, y_train), (x_test, y_test) = synth.load("cifar", 2048, 512)
    TensorflowGenerator.dataset = "cifar"
    TensorflowGenerator.X_train, TensorflowGenerator.Y_train = x_train, y_train
    TensorflowGenerator.X_test, TensorflowGenerator.Y_test = x_test, y_test
    TensorflowGenerator.input_shape = synth.SHAPES["cifar"]
    # four distinct products, each eight times: 32 jobs for 8 threads
    chosen, costs, mem = [], [], []
    for item in formatted:
        with contextlib.redirect_stdout(io.StringIO()):
            net = KerasFeatureModel.parse_feature_model(copy.deepcopy(item[0])).build((32, 32, 3), 10)
        if net and net.count_params() < 2_000_000:
            chosen.append(item)
            costs.append(pop.estimate_cost(net, 2048, 512, 1))
            mem.append(pop.estimate_bytes(net, 128, 512))
        if len(chosen) == 4:
            break
    assert len(chosen) == 4
    jobs, job_costs, job_mem = chosen * 8, costs * 8, mem * 8

    def train_one_shared(index, item):
        product, original = pop.copy_product(item)
        return pop.train_product(product, "cifar", epochs=1, batch_size=128, product_features=original, shared_gpu=True)

    def train_one_alone(index, item):
        product, original = pop.copy_product(item)
        return pop.train_product(product, "cifar", epochs=1, batch_size=128, product_features=original, shared_gpu=False)

    alone = pop.run_population(chosen, train_one_alone, costs, workers=1)
    assert all(r.get("valid") and not r.get("error") for _, r in alone), alone
    budget = pop.MemoryBudget(int(0.8 * torch.cuda.mem_get_info()[0]))
    for attempt in range(3):
        with contextlib.redirect_stdout(io.StringIO()):
            shared = pop.run_population(jobs, train_one_shared, job_costs, workers=8, mem_bytes=job_mem, budget=budget)
        assert [i for i, _ in shared] == list(range(32))
        for index, result in shared:
            want = alone[index % 4][1]
            assert not result.get("error"), (attempt, index, result.get("error"), result.get("traceback"))
            assert result["valid"] and result["params"] == want["params"]
            # FNET_SEED pins weights, permutations and dropout masks, and every kernel is deterministic
            assert result["accuracy"] == want["accuracy"], (attempt, index, result["accuracy"], want["accuracy"])
            assert result["history"] == want["history"]
    TensorflowGenerator.dataset = None
