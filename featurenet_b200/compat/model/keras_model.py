# -*- coding: utf-8 -*-
"""Product / template -> architecture description -> executable network.

Behaviour of /root/reference/model/keras_model.py:23-219.  ``build`` chains the blocks,
appends ``Flatten`` + ``Dense(n_classes, softmax, name="out")``, rejects candidates above
20 M parameters and turns any construction error into ``None`` (printed traceback), which
is what makes ``TensorflowGenerator`` report "model is not valid".

The reference's ``multi_gpu_model`` branch (keras_model.py:137-146) becomes data-parallel training over the
ranks of the process group when ``FNET_DP=1`` (``featurenet_b200.dataparallel``); without it whole candidates
are placed on GPUs by ``featurenet_b200.population``.
"""
from __future__ import absolute_import, division, print_function, unicode_literals

import random
import traceback
import uuid

from .mutation.mutable_model import MutableModel
from .block import Block
from .output import Out  # noqa: F401
from .cell import Cell  # noqa: F401

from featurenet_b200 import layers as L


class KerasFeatureVector(object):
    """Genome used by evolution.py: accuracy, attributes, feature bits (keras_model.py:23-56)."""
    features = []
    attributes = [0, 0, 0, 0]
    accuracy = None

    def __init__(self, accuracy, attributes, features):
        self.features = features
        self.attributes = attributes
        self.accuracy = accuracy

    def mutate(self, rate=0.05):
        keep = [random.random() > rate for _ in range(len(self.features))]
        self.features = [bit if k else 1 - bit for bit, k in zip(self.features, keep)]

    def cross_over(self, second_vector, crossover_type="onepoint"):
        if crossover_type == "onepoint":
            cut = random.randint(0, len(self.features))
            return KerasFeatureVector(0, [0, 0, 0, 0], self.features[0:cut] + second_vector.features[cut:])

    def to_vector(self):
        return [self.accuracy] + [self.attributes] + self.features

    @staticmethod
    def from_vector(vect):
        return KerasFeatureVector(vect[0], vect[1], vect[2:])

    def __str__(self):
        return "{}:{}".format(";".join(str(a) for a in self.attributes), self.accuracy)

    @property
    def fitness(self):
        return 0 if self.accuracy is None else self.accuracy


class KerasFeatureModel(MutableModel):
    blocks = []
    outputs = []
    optimizers = []
    features = []
    features_label = []
    nb_flops = 0
    nb_params = 0
    robustness_score = 0
    clever_score = 0
    pgd_score = 0
    cw_score = 0
    fgsm_score = 0
    model = None
    accuracy = 0
    use_multigpu = True
    robustness_scores = ["clever", "cw", "fgsm", "pgd"]
    metrics = []
    layers = {"pool": [], "conv": []}

    max_parameters = 20000000

    def __init__(self, name=""):
        self._name = name if name else str(uuid.uuid1())[:10]
        super(KerasFeatureModel, self).__init__()

    def get_custom_parameters(self):
        params = {}
        for block in self.blocks:
            params.update(block.get_custom_parameters())
        return params

    def to_kerasvector(self):
        nb_layers = len(self.model.layers) if self.model else 0
        shape = [len(self.blocks), nb_layers, self.nb_params, self.nb_flops]
        robustness = [self.robustness_score, [self.clever_score, self.fgsm_score, self.pgd_score, self.cw_score]]
        return KerasFeatureVector(self.accuracy, [self._name, shape, robustness, self.metrics], self.features)

    # the reference also calls this name from evolution.py:69
    def to_vector(self):
        return self.to_kerasvector()

    def build(self, input_shape, output_shape, max_parameters=20000000):
        self.outputs = []
        model = None
        try:
            print("Build Tensorflow model")
            stack = [L.Input(input_shape)]
            network_input = stack[0]
            for block in self.blocks:
                stack, produced = block.build_tensorflow_model(stack)
                self.outputs = self.outputs + produced

            tail = self.outputs[-1] if len(self.outputs) else stack[0]
            tail = tail.content if hasattr(tail, "content") else tail
            if tail.shape.ndims > 2:
                tail = L.Flatten()(tail)
            self.outputs = [L.Dense(output_shape, activation="softmax", name="out")(tail)]

            model = L.Model(outputs=self.outputs, inputs=network_input, name=self._name)
            if model.count_params() > KerasFeatureModel.max_parameters:
                print("#### model is bigger than 20M params. Skipped")
                model.summary()
                return None

            if KerasFeatureModel.use_multigpu:
                # keras_model.py:137-146: with two or more GPUs the model is wrapped for batch-split data parallelism,
                # otherwise "multi gpu not available" is printed and the switch turns itself off.  Here the GPUs are the
                # ranks of the process group and the mode is asked for with FNET_DP=1 (a population run keeps whole
                # candidates per GPU instead: featurenet_b200.population).
                from featurenet_b200 import dataparallel
                if dataparallel.wanted():
                    model = dataparallel.multi_gpu_model(model)
                else:
                    print("multi gpu not available")
                    KerasFeatureModel.use_multigpu = False
        except Exception as e:
            print("error", e)
            print(traceback.format_exc())
            if model:
                model.summary()
            return None

        self.model = model
        return model

    def dump_blocks(self):
        return {"blocks": self.blocks, "accuracy": self.accuracy}

    @staticmethod
    def parse_blocks(block_model, name=None):
        model = KerasFeatureModel(name=name)
        model.blocks = block_model.get("blocks", [])
        model.accuracy = block_model.get("accuracy", 0)
        return model

    @staticmethod
    def parse_feature_model(feature_model, name=None, depth=1, product_features=None, features_label=None):
        print("building keras model from feature model tree")
        model = KerasFeatureModel(name=name)
        if product_features:
            model.features = [1 if str(x).isdigit() and int(x) > 0 else 0 for x in product_features]
        if features_label:
            model.features_label = features_label
        model.blocks = []
        if len(feature_model) == 0:
            return model

        if isinstance(feature_model, str):
            model.blocks = KerasFeatureModel.get_from_template(feature_model)
        else:
            for _ in range(depth):
                for block_dict in feature_model:
                    # reversed in place: a product parsed twice comes out in the other order (keras_model.py:199)
                    block_dict["children"] = list(reversed(block_dict["children"]))
                    model.blocks.append(Block.parse_feature_model(block_dict, model))
            model.blocks.sort(key=lambda b: b.get_name())
        return model

    @staticmethod
    def get_from_template(feature_model):
        if feature_model == "lenet5":
            from .leNet import lenet5_blocks
            return lenet5_blocks()
        if feature_model == "keras":
            from .kerasNet import standard_blocks
            return standard_blocks()
        return []
