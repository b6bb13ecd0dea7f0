"""The object ``Model(...)`` returns: a pruned layer DAG plus (lazily) its CUDA executor.

It exposes the part of the Keras ``Model`` surface that FeatureNet's runner and drivers touch
(/root/reference/tensorflow_generator.py:221-277, helpers.py:69-175, model/keras_model.py:97-130):
``layers``, ``count_params``, ``summary``, ``compile``, ``fit``, ``evaluate``, ``predict``,
``get_weights`` / ``set_weights``, ``save``, ``optimizer.lr`` and ``stop_training``.

All arithmetic runs in the hand-written kernels of ``csrc/`` through the C ABI; this file
only plans memory and sequences launches.  There is no CPU execution path: using any
compute method without the CUDA library raises.
"""
from . import layers as L


class Network(object):
    def __init__(self, inputs, output, name=None):
        self.name = name or "network"
        self.input = inputs[0] if isinstance(inputs, (list, tuple)) else inputs
        self.output = output
        self.layers = L.reachable_layers([output])
        if self.input.node not in self.layers:
            raise ValueError("Graph disconnected: the output does not depend on the input")
        self.stop_training = False
        self.optimizer = None
        self.loss = None
        self.metrics_names = ["loss", "acc"]
        self._engine = None
        self._init_seed = None

    # ------------------------------------------------------------------ structure
    @property
    def input_shape(self):
        return tuple(self.input.shape.as_list())

    @property
    def output_shape(self):
        return tuple(self.output.shape.as_list())

    def count_params(self):
        return sum(node.nb_params for node in self.layers)

    def describe(self):
        """Layer list with connectivity by index; the structural-parity fixtures compare this."""
        position = {id(node): i for i, node in enumerate(self.layers)}
        rows = []
        for node in self.layers:
            row = node.describe()
            row["inbound"] = [position[id(t.node)] for t in node.inputs]
            rows.append(row)
        return rows

    def summary(self, print_fn=print):
        print_fn("Model: %s" % self.name)
        print_fn("%-44s %-24s %10s" % ("Layer (type)", "Output Shape", "Param #"))
        for node in self.layers:
            print_fn("%-44s %-24s %10d" % ("%s (%s)" % (node.name[:28], node.class_name),
                                            str(tuple(node.output.shape.as_list())), node.nb_params))
        print_fn("Total params: %d" % self.count_params())

    def train_flops_per_image(self):
        """Analytic multiply-add count x2, forward pass, per image (fills the reference's unused `flops`)."""
        total = 0
        for node in self.layers:
            out = node.output.shape.as_list()[1:]
            n_out = 1
            for d in out:
                n_out *= d
            if node.kind == "conv2d":
                kh, kw = node.cfg["kernel"]
                cin = node.inputs[0].shape.as_list()[-1]
                total += 2 * n_out * kh * kw * cin
            elif node.kind == "dwconv2d":
                kh, kw = node.cfg["kernel"]
                total += 2 * n_out * kh * kw
            elif node.kind == "sepconv2d":
                kh, kw = node.cfg["kernel"]
                cin = node.inputs[0].shape.as_list()[-1]
                total += 2 * (n_out // out[-1]) * cin * kh * kw + 2 * n_out * cin
            elif node.kind == "dense":
                total += 2 * n_out * node.inputs[0].shape.as_list()[-1]
        return total

    # ------------------------------------------------------------------ execution (CUDA only)
    def _need_engine(self):
        if self._engine is None:
            from .engine import Engine
            from . import dataparallel
            self._engine = Engine(self)
            if getattr(self, "data_parallel", False) or dataparallel.wanted():
                import torch.distributed as dist
                self._engine.enable_data_parallel(dist)
        return self._engine

    def enable_data_parallel(self, dist=None):
        """Train this candidate data-parallel over the process group (keras_model.py:137-146's multi_gpu_model)."""
        self.data_parallel = True
        if self._engine is not None and self._engine.dp is None:
            if dist is None:
                import torch.distributed as dist
            self._engine.enable_data_parallel(dist)
        return self

    def compile(self, loss="categorical_crossentropy", metrics=None, optimizer=None, **_):
        from .optimizers import resolve_optimizer
        if loss != "categorical_crossentropy":
            raise ValueError("only categorical_crossentropy is on the FeatureNet path")
        self.loss = loss
        self.optimizer = resolve_optimizer(optimizer)
        return self

    def fit(self, x, y, batch_size=32, epochs=1, validation_data=None, shuffle=True, callbacks=None,
            verbose=0, **kw):
        return self._need_engine().fit(x, y, batch_size=batch_size, epochs=epochs,
                                       validation_data=validation_data, shuffle=shuffle,
                                       callbacks=callbacks or [], verbose=verbose, **kw)

    def fit_generator(self, generator, steps_per_epoch=None, epochs=1, validation_data=None, callbacks=None,
                      verbose=0, **kw):
        return self._need_engine().fit_generator(generator, steps_per_epoch=steps_per_epoch, epochs=epochs,
                                                 validation_data=validation_data, callbacks=callbacks or [],
                                                 verbose=verbose, **kw)

    def evaluate(self, x, y, batch_size=32, verbose=0, **_):
        return self._need_engine().evaluate(x, y, batch_size=batch_size)

    def predict(self, x, batch_size=32, **_):
        return self._need_engine().predict(x, batch_size=batch_size)

    def train_on_batch(self, x, y, **_):
        """keras Model.train_on_batch: one optimisation step on a host batch, returns [loss, accuracy]."""
        return self._need_engine().train_on_batch(x, y)

    def input_gradient(self, x, y):
        """dL/dx in inference mode (device tensor): what the robustness attacks differentiate."""
        return self._need_engine().input_gradient(x, y)

    def logits(self, x):
        """Pre-softmax outputs of the head in inference mode (host array)."""
        return self._need_engine().logits(x)

    def logit_gradient(self, x, head_dz):
        """(sum_j head_dz[:, j] * dlogit_j/dx, logits) in inference mode: class gradients for CLEVER."""
        return self._need_engine().logit_gradient(x, head_dz)

    def get_weights(self):
        return self._need_engine().get_weights()

    def set_weights(self, weights):
        return self._need_engine().set_weights(weights)

    def save(self, path, **_):
        return self._need_engine().save(path)

    def release(self):
        """Free device memory (the reference tears the TF session down between candidates)."""
        if self._engine is not None:
            self._engine.release()
            self._engine = None
