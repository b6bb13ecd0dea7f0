# -*- coding: utf-8 -*-
"""``TensorflowGenerator``: build, train and evaluate one candidate -- the call every driver makes.

Same constructor, static methods, class attributes and result fields as
/root/reference/tensorflow_generator.py:73-395, so full.py, full_mnist.py, full_cifar.py,
full_lenet5.py, evolution.py, pledge_evolution.py and full_evolution.py run unchanged with
``PYTHONPATH=featurenet_b200/compat``.  Underneath, ``model.build`` records a layer DAG and
``fit`` / ``evaluate`` run hand-written sm_100a kernels (featurenet_b200.engine).

Differences that cannot be avoided on this box, all loud:
  * datasets: ``keras.datasets`` needs a download; ``init_dataset`` uses ``FNET_DATA_DIR/<name>.npz``
    (keys x_train, y_train, x_test, y_test; uint8 images, integer labels) when present, otherwise a
    synthetic set of the same shape and size (featurenet_b200.synth), and says which;
  * robustness (``eval_robustness``): all four scores (clever, pgd, cw, fgsm) run on the engine
    (featurenet_b200/robustness.py), restated from ART 0.9.0's published algorithms (ART is un-vendored);
  * ``export_png`` needs Graphviz ``dot``; it is a no-op after the first failure, as in the reference.
"""
from __future__ import absolute_import, division, print_function, unicode_literals

import os
import time

import numpy as np

from model.keras_model import KerasFeatureModel
from helpers import train_model
from featurenet_b200 import layers as _layers
from featurenet_b200 import synth as _synth
from featurenet_b200.optimizers import Adam


def get_flops():
    """The reference profiles the TF graph (and never calls it, tensorflow_generator.py:272); analytic here."""
    model = TensorflowGenerator._last_network
    return 0 if model is None else 3 * model.train_flops_per_image()


def lr_schedule(epoch):
    lr = 1e-3
    if epoch > 180:
        lr *= 0.5e-3
    elif epoch > 160:
        lr *= 1e-3
    elif epoch > 120:
        lr *= 1e-2
    elif epoch > 80:
        lr *= 1e-1
    return lr


def reset_keras():
    """Tear down the previous candidate's device state (the reference closes and reopens the TF session)."""
    previous = TensorflowGenerator._last_network
    if previous is not None:
        previous.release()
        TensorflowGenerator._last_network = None
    _layers.reset_uids()


class TensorflowGenerator(object):
    model_graph = ""
    accuracy = 0
    training_time = 0
    params = 0
    flops = 0
    stop_training = False
    X_train = []
    Y_train = []
    X_test = []
    Y_test = []
    history = ([], [])
    dataset = None
    input_shape = (0, 0, 0)
    default_batchsize = 64
    num_classes = 10
    default_robustness_set_size = 500

    model_graph_export = True
    eval_metrics = {}

    datasets_classes = {"mnist": 10, "cifar": 10, "cifar10": 10, "cifar100": 100}

    _last_network = None

    def __init__(self, product, epochs=12, dataset="mnist", data_augmentation=True, depth=1, product_features=None,
                 features_label=None, no_train=False, clear_memory=True, batch_size=128, eval_robustness=None,
                 save_path=None, robustness_set_size=0):
        if batch_size == 0:
            batch_size = TensorflowGenerator.default_batchsize

        if product:
            self.model = KerasFeatureModel.parse_feature_model(product, name="", depth=depth,
                                                               product_features=product_features,
                                                               features_label=features_label)
            print("====> Loading new feature model with {0} blocks".format(len(self.model.blocks)))
            model = TensorflowGenerator.build(self.model, dataset, clear_memory=clear_memory)
            if not model:
                print("#### model is not valid ####")
                return
            if no_train:
                self.print()
                return
            if save_path:
                save_path = "{}{}".format(save_path, self.model._name)

            history, training_time, score, keras_model = TensorflowGenerator.train(
                self.model, epochs, batch_size, dataset, data_augmentation, save_path=save_path)
            if not keras_model:
                print("#### model is not valid ####")
                return
            if eval_robustness:
                TensorflowGenerator.eval_robustness(self.model, eval_robustness, robustness_set_size)
            if TensorflowGenerator.eval_metrics:
                for (m, f) in TensorflowGenerator.eval_metrics:
                    self.model.metrics[m] = f(keras_model)

            self.params = self.model.nb_params
            self.flops = self.model.nb_flops
            self.training_time = training_time
            self.accuracy = self.model.accuracy
            self.history = (history.history["acc"], history.history["val_acc"])

    @property
    def keras_model(self):
        return self.model.model

    @property
    def metrics(self):
        return self.model.metrics

    @staticmethod
    def add_metric(metric, func):
        TensorflowGenerator.eval_metrics[metric] = func

    @staticmethod
    def eval_attack_robustness(keras_model, attack_name, norm, robustness_set_size=0):
        """(empirical robustness, clean accuracy, adversarial accuracy) -- tensorflow_generator.py:151-173."""
        from featurenet_b200 import robustness
        attack_params = {"norm": norm}
        if attack_name == "cw":
            attack_params["targeted"] = False
        elif attack_name == "pgd":
            attack_params["eps_step"] = 0.1
            attack_params["eps"] = 1.
        if robustness_set_size == 0:
            adv_set, y_set = TensorflowGenerator.X_robustness, TensorflowGenerator.Y_robustness
        else:
            adv_set = TensorflowGenerator.X_test[0:min(len(TensorflowGenerator.X_test), robustness_set_size)]
            y_set = TensorflowGenerator.Y_test[0:min(len(TensorflowGenerator.Y_test), robustness_set_size)]
        return robustness.attack_robustness(keras_model, attack_name, adv_set, y_set, attack_params)

    @staticmethod
    def eval_robustness(model, scores=[], robustness_set_size=0):
        """The second fitness of the evolution drivers (tensorflow_generator.py:175-218).  The gradient attacks
        (fgsm, pgd, cw) and CLEVER all run on the engine."""
        keras_model = model.model
        if not keras_model or model.accuracy < 0.5:
            return
        begin_robustness = time.time()
        score_metrics = model.robustness_scores if not scores else scores
        try:
            from featurenet_b200 import robustness
            norm = 2
            r_l1, r_l2, r_li = 40, 2, 0.1
            nb_batches = 10
            batch_size = 5
            radius = r_l1 if norm == 1 else (r_l2 if norm == 2 else r_li)
            if "clever" in score_metrics:
                if robustness_set_size == 0:
                    x_set = TensorflowGenerator.X_robustness
                else:
                    x_set = TensorflowGenerator.X_test[0:robustness_set_size]
                model.clever_score = robustness.clever_score(keras_model, x_set, nb_batches, batch_size, radius, norm=norm,
                                                             pool_factor=3)
            if "pgd" in score_metrics:
                model.pgd_score = TensorflowGenerator.eval_attack_robustness(keras_model, "pgd", norm, robustness_set_size)
            if "cw" in score_metrics:
                model.cw_score = TensorflowGenerator.eval_attack_robustness(keras_model, "cw", norm, robustness_set_size)
            if "fgsm" in score_metrics:
                model.fgsm_score = TensorflowGenerator.eval_attack_robustness(keras_model, "fgsm", norm, robustness_set_size)
        except Exception as e:
            import traceback
            print("error", e)
            print(traceback.format_exc())
        robustness_time = time.time() - begin_robustness
        first = getattr(model, "{}_score".format(scores[0]), [0]) if len(scores) else model.clever_score
        model.robustness_score = first[0] if isinstance(first, (list, tuple)) else first
        print('model robustness (clever, pgd, cw, fgsm): {} time:{}'.format(
            (model.clever_score, model.pgd_score, model.cw_score, model.fgsm_score), robustness_time))

    @staticmethod
    def build(model, dataset, clear_memory=True):
        if clear_memory:
            reset_keras()
        TensorflowGenerator.init_dataset(dataset)
        keras_model = model.build(TensorflowGenerator.input_shape, TensorflowGenerator.datasets_classes.get(dataset))
        if not keras_model:
            return keras_model
        # optimizers = [Adam(lr=lr_schedule(0)), "sgd"]: only the first entry is ever used
        keras_model.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=lr_schedule(0)))
        TensorflowGenerator._last_network = keras_model
        model.nb_params = keras_model.count_params()
        # the reference leaves nb_flops at 0 (its get_flops() is never called, tensorflow_generator.py:272); the vector
        # consumers (ui/back, plots) get the analytic training cost per image: 3 x the forward multiply-adds x 2
        model.nb_flops = 3 * keras_model.train_flops_per_image()
        print("model blocks,layers,params,flops: {} ".format(model.to_kerasvector()))
        return keras_model

    @staticmethod
    def train(model, epochs, batch_size, dataset, data_augmentation=True, save_path=None):
        score = []
        keras_model = model.model if hasattr(model, "model") else model
        begin_training = time.time()
        model_path = "{}.h5".format(save_path) if save_path else None

        keras_model, history = train_model(keras_model, TensorflowGenerator.X_train, TensorflowGenerator.Y_train,
                                           TensorflowGenerator.X_test, TensorflowGenerator.Y_test, epochs, batch_size,
                                           True, data_augmentation, model_path)
        training_time = time.time() - begin_training

        if keras_model:
            if model_path:
                keras_model.save(model_path)
                TensorflowGenerator.export_png(keras_model, save_path)
            score = keras_model.evaluate(TensorflowGenerator.X_test, TensorflowGenerator.Y_test, verbose=0)
            model.accuracy = score[1]
            print("Test loss: {} Test accuracy: {} training_time {}".format(score[0], score[1], training_time))
        return history, training_time, score, keras_model

    @staticmethod
    def load_arrays(dataset):
        """(x_train uint8/float, y_train int), (x_test, y_test) -- local npz if provided, else synthetic."""
        folder = os.environ.get("FNET_DATA_DIR")
        if folder:
            name = {"cifar": "cifar10"}.get(dataset, dataset)
            path = os.path.join(folder, name + ".npz")
            if os.path.isfile(path):
                with np.load(path) as z:
                    print("dataset {} loaded from {}".format(dataset, path))
                    return (z["x_train"], z["y_train"].reshape(-1)), (z["x_test"], z["y_test"].reshape(-1))
        n_train = int(os.environ.get("FNET_SYNTH_TRAIN", _synth.SIZES[dataset][0]))
        n_test = int(os.environ.get("FNET_SYNTH_TEST", _synth.SIZES[dataset][1]))
        print("dataset {}: no local copy (keras.datasets needs a download) -> synthetic {}x{} train / {} test".format(
            dataset, n_train, "x".join(str(s) for s in _synth.SHAPES[dataset]), n_test))
        return _synth.make_synth_uint8(_synth.SHAPES[dataset], n_train, n_test, _synth.CLASSES[dataset])

    @staticmethod
    def init_dataset(dataset, data_augmentation=False):
        TensorflowGenerator.num_classes = 10
        if TensorflowGenerator.dataset != dataset:
            (x_train, y_train), (x_test, y_test) = TensorflowGenerator.load_arrays(dataset)
            if dataset == "cifar100":
                TensorflowGenerator.num_classes = 100
            rows, cols = x_train.shape[1], x_train.shape[2]
            channels = x_train.shape[3] if len(x_train.shape) == 4 else 1
            y_train = _synth.to_categorical(np.asarray(y_train).reshape(-1), TensorflowGenerator.num_classes)
            y_test = _synth.to_categorical(np.asarray(y_test).reshape(-1), TensorflowGenerator.num_classes)
            x_train = x_train.reshape(x_train.shape[0], rows, cols, channels)
            x_test = x_test.reshape(x_test.shape[0], rows, cols, channels)
            TensorflowGenerator.input_shape = (rows, cols, channels)
            x_train = x_train.astype("float32")
            x_test = x_test.astype("float32")
            x_train /= 255
            x_test /= 255

            TensorflowGenerator.X_train, TensorflowGenerator.X_test = x_train, x_test
            TensorflowGenerator.Y_train, TensorflowGenerator.Y_test = y_train, y_test
            size = TensorflowGenerator.default_robustness_set_size
            TensorflowGenerator.X_robustness, TensorflowGenerator.Y_robustness = x_test[0:size], y_test[0:size]
            TensorflowGenerator.dataset = dataset

    @staticmethod
    def export_png(model, path):
        if not TensorflowGenerator.model_graph_export:
            return
        print("saving model png to {}.png".format(path))
        print("error export model image: Graphviz `dot` is not available")
        TensorflowGenerator.model_graph_export = False

    def print(self, include_summary=True, invalid_params=True, export_png=True):
        model = self.model.model
        if include_summary:
            model.summary()
        if invalid_params:
            for name, (node, params) in self.model.get_custom_parameters().items():
                print("{0}:{1}".format(name, params))
        if TensorflowGenerator.model_graph and export_png:
            TensorflowGenerator.export_png(model, TensorflowGenerator.model_graph)

    def load_products(self, product):
        def walk(node, level=0):
            for child in node.get("children"):
                walk(child, level + 1)
        walk(product)
