"""Graph-construction parity: compat ``model/*`` + ``featurenet_b200.layers`` vs the reference.

The expected values in tests/golden/structural.json were produced by the reference's own
model/*.py, model/mutation/*.py and Products_tree.py (tests/golden/make_structural_golden.py),
so these checks pin layer lists, shapes, parameter counts, connectivity, invalid-candidate
behaviour, the mutation RNG stream and breeding against the reference itself.
"""
import contextlib
import copy
import io
import json
import os

import numpy as np
import pytest

from model.keras_model import KerasFeatureModel
from model.mutation.mutable_base import MutableBase, MutationStrategies
from products_tree import ProductSet

SHAPES = {"mnist": (28, 28, 1), "cifar": (32, 32, 3)}
CFG_KEYS = {
    "Conv2D": ("filters", "kernel", "strides", "padding", "activation", "dilation"),
    "SeparableConv2D": ("filters", "kernel", "strides", "padding", "activation"),
    "DepthwiseConv2D": ("kernel", "strides", "padding", "activation"),
    "Dense": ("units", "activation"),
    "MaxPooling2D": ("pool", "strides", "padding"),
    "AveragePooling2D": ("pool", "strides", "padding"),
    "Dropout": ("rate",),
    "BatchNormalization": ("axis",),
    "Activation": ("activation",),
    "ZeroPadding2D": ("pad",),
    "Concatenate": ("axis",),
}


@pytest.fixture(scope="module")
def golden(golden_dir):
    with open(os.path.join(golden_dir, "structural.json")) as handle:
        return json.load(handle)


def record(model, shape):
    with contextlib.redirect_stdout(io.StringIO()):
        net = model.build(shape, 10)
    if not net:
        return {"valid": False}
    return {"valid": True, "params": net.count_params(), "layers": net.describe()}


def assert_same(mine, ref, where):
    assert mine["valid"] == ref["valid"], where
    if not ref["valid"]:
        return
    assert mine["params"] == ref["params"], where
    assert len(mine["layers"]) == len(ref["layers"]), where
    for i, (a, b) in enumerate(zip(mine["layers"], ref["layers"])):
        at = "%s layer %d" % (where, i)
        assert a["class"] == b["class"], at
        assert a["out_shape"] == b["out_shape"], at
        assert a["params"] == b["params"], at
        assert a["inbound"] == b["inbound"], at
        for key in CFG_KEYS.get(a["class"], ()):
            assert a["cfg"][key] == b["cfg"][key], "%s cfg %s" % (at, key)


def parse(product, **kw):
    with contextlib.redirect_stdout(io.StringIO()):
        return KerasFeatureModel.parse_feature_model(product, **kw)


def load_products(path):
    with contextlib.redirect_stdout(io.StringIO()):
        return ProductSet(path).format_products()


def test_templates(golden):
    for key, ref in golden["templates"].items():
        tpl, ds = key.split("@")
        assert_same(record(parse(tpl), SHAPES[ds]), ref, key)
    # anchors quoted in SURVEY.md / BASELINE.md
    assert golden["templates"]["lenet5@mnist"]["params"] == 212902
    assert golden["templates"]["lenet5@cifar"]["params"] == 268114
    assert golden["templates"]["keras@cifar"]["params"] == 5341738


@pytest.mark.parametrize("pdt", ["nas_1_1_10_seed0.pdt", "nas_5_5_10_seed0.pdt",
                                 "main_1block_nas_5cells_seed0.pdt", "main_light_seed0.pdt"])
def test_sampled_products(golden, golden_dir, pdt):
    refs = golden["products"][pdt]
    formatted = load_products(os.path.join(golden_dir, pdt))
    assert len(formatted) == len(refs)
    for i, ((product, original), ref) in enumerate(zip(formatted, refs)):
        model = parse(product, product_features=original)
        assert_same(record(model, SHAPES["cifar"]), ref, "%s[%d]" % (pdt, i))


def test_shipped_10products_are_all_invalid(golden):
    # the shipped population predates the parsers; every candidate fails to build
    rows = golden["products"]["10Products.pdt"]
    assert len(rows) == 10 and not any(r["valid"] for r in rows)


def test_10products_invalid_here_too():
    path = "/root/reference/datasets/10Products.pdt"
    if not os.path.isfile(path):
        pytest.skip("reference tree not present on this box")
    for product, original in load_products(path):
        assert record(parse(product, product_features=original), SHAPES["cifar"]) == {"valid": False}


def test_mutation_walks_follow_the_reference_rng_stream(golden):
    MutableBase.mutation_stategy = MutationStrategies.CHOICE
    for key, steps in golden["mutation_walks"].items():
        tpl, seed = key.split("/seed")
        np.random.seed(int(seed))
        model = parse(tpl)
        for n, ref in enumerate(steps):
            with contextlib.redirect_stdout(io.StringIO()):
                model.mutate(0.1)
            assert_same(record(model, SHAPES["cifar"]), ref, "%s step %d" % (key, n))


def test_breeding(golden):
    MutableBase.mutation_stategy = MutationStrategies.CHOICE
    for key, rows in golden["breeding"].items():
        seed = int(key[len("seed"):])
        np.random.seed(100 + seed)
        a, b = parse("lenet5"), parse("keras")
        for _ in range(3):
            a.mutate(0.1)
            b.mutate(0.1)
        got = []
        for ratio in (1.0, 0.5, 0.0):
            child = a.breed(b, ratio)
            got.append(record(child, SHAPES["cifar"]))
            child.mutate(0.1)
            got.append(record(child, SHAPES["cifar"]))
        for n, (mine, ref) in enumerate(zip(got, rows)):
            assert_same(mine, ref, "%s[%d]" % (key, n))


def test_in_place_reversal_and_depth(golden, golden_dir):
    pdt = "nas_5_5_10_seed0.pdt"
    refs = golden["reparse"][pdt]
    formatted = load_products(os.path.join(golden_dir, pdt))[:4]
    for i, ((product, _), ref) in enumerate(zip(formatted, refs)):
        first = parse(product)
        again = parse(product)
        deep = parse(copy.deepcopy(product), depth=2)
        for tag, model, r in (("first", first, ref[0]), ("again", again, ref[1]), ("depth2", deep, ref[2])):
            assert_same(record(model, SHAPES["cifar"]), r, "%s[%d] %s" % (pdt, i, tag))


def test_product_reader_matches_reference_reader(golden_dir):
    ref_dir = "/root/reference"
    if not os.path.isdir(ref_dir):
        pytest.skip("reference tree not present on this box")
    import importlib.util
    import sys
    import types
    sys.modules.setdefault("validators", types.ModuleType("validators"))
    spec = importlib.util.spec_from_file_location("ref_products_tree", os.path.join(ref_dir, "Products_tree.py"))
    ref = importlib.util.module_from_spec(spec)
    with contextlib.redirect_stdout(io.StringIO()):
        spec.loader.exec_module(ref)
        for pdt in ("nas_1_1_10_seed0.pdt", "nas_5_5_10_seed0.pdt"):
            path = os.path.join(golden_dir, pdt)
            for binary in (False, True):
                mine, theirs = ProductSet(path, binary), ref.ProductSet(path, binary)
                assert mine.features == theirs.features
                assert mine.products == theirs.products
                assert mine.format_products() == theirs.format_products()
                assert ProductSet.filter_leaves(mine.features) == ref.ProductSet.filter_leaves(theirs.features)
