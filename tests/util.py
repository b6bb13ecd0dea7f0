"""Shared helpers for the tests: seeded weights, small graphs covering the whole op vocabulary."""
import contextlib
import io

import numpy as np

from featurenet_b200 import layers as L


def init_weights(net, rng, jitter=False):
    """Keras-order float64 weights per layer: glorot-uniform kernels, zero bias, BN (1,0,0,1).

    jitter=True perturbs biases / BN parameters so that no gradient is trivially zero.
    """
    out = []
    for node in net.layers:
        ws = []
        for spec in node.weights:
            if spec.init == "glorot_uniform":
                w = rng.uniform(-L.glorot_limit(spec.fans), L.glorot_limit(spec.fans), size=spec.shape)
            elif spec.init == "ones":
                w = np.ones(spec.shape)
            else:
                w = np.zeros(spec.shape)
            if jitter and spec.init != "glorot_uniform":
                w = w + 0.3 * rng.standard_normal(spec.shape)
                if spec.name == "moving_variance":
                    w = np.abs(w) + 0.5
            ws.append(w.astype(np.float64))
        out.append(ws)
    return out


def build_template(name, shape, classes=10):
    from model.keras_model import KerasFeatureModel
    with contextlib.redirect_stdout(io.StringIO()):
        model = KerasFeatureModel.parse_feature_model(name)
        return model.build(shape, classes)


def _head(t, classes=3):
    if t.shape.ndims > 2:
        t = L.Flatten()(t)
    return L.Dense(classes, activation="softmax", name="out")(t)


def tiny_graphs():
    """name -> builder of a small Network; together they touch every layer kind and both merge paths."""
    def conv_chain():
        x = L.Input((6, 5, 2))
        t = L.Conv2D(3, (3, 3), strides=(1, 1), padding="same", activation="tanh")(x)
        t = L.Conv2D(4, (1, 5), strides=(2, 2), padding="same", activation="relu")(t)
        t = L.Conv2D(2, 1, strides=2, padding="valid")(t)
        return L.Model(inputs=x, outputs=_head(t))

    def separable_depthwise():
        x = L.Input((5, 6, 3))
        t = L.SeparableConv2D(4, (3, 1), strides=(1, 1), padding="same", activation="sigmoid")(x)
        t = L.DepthwiseConv2D((2, 2), strides=(2, 2), padding="same", activation="tanh")(t)
        return L.Model(inputs=x, outputs=_head(t))

    def pools_and_bn():
        x = L.Input((6, 6, 2))
        t = L.Conv2D(3, (3, 3), padding="same", activation="tanh")(x)
        t = L.BatchNormalization(axis=1)(t)
        t = L.MaxPooling2D((2, 2), strides=(1, 1), padding="same")(t)
        t = L.AveragePooling2D((3, 3), strides=(2, 2), padding="same")(t)
        return L.Model(inputs=x, outputs=_head(t))

    def merges():
        x = L.Input((4, 4, 2))
        a = L.Conv2D(3, (3, 3), padding="same", activation="tanh")(x)
        b = L.Conv2D(3, (1, 1), padding="same", activation="sigmoid")(x)
        s = L.Add()([a, b])
        c = L.Concatenate(axis=1)([s, a])              # stacks along H: [B, 8, 4, 3]
        m = L.Multiply()([c, c])
        return L.Model(inputs=x, outputs=_head(m))

    def dropout_pad_act_gap():
        x = L.Input((4, 5, 3))
        t = L.ZeroPadding2D((1, 1))(x)
        t = L.Conv2D(4, (3, 3), padding="same")(t)
        t = L.Activation("relu")(t)
        t = L.Dropout(0.3)(t)
        t = L.Dense(5, activation="tanh")(t)            # Dense on a rank-4 tensor
        t = L.GlobalAveragePooling2D()(t)
        t = L.BatchNormalization(axis=1)(t)             # rank-2 BN over features
        t = L.Activation("softmax")(t)
        return L.Model(inputs=x, outputs=_head(t))

    def merges_with_the_input():
        """Merges whose FIRST operand is the network input (no gradient owed to it): the second one still is."""
        x = L.Input((4, 4, 2))
        a = L.Conv2D(2, (3, 3), padding="same", activation="relu")(x)
        s = L.Add()([x, a])
        b = L.Conv2D(2, (1, 1), padding="same", activation="tanh")(s)
        c = L.Concatenate(axis=1)([x, b])              # what every FeatureNet cell does with its input
        d = L.Conv2D(2, (1, 1), padding="same")(c)
        m = L.Multiply()([c, d])
        return L.Model(inputs=x, outputs=_head(m))

    return {"merges_with_the_input": merges_with_the_input, "conv_chain": conv_chain, "separable_depthwise": separable_depthwise, "pools_and_bn": pools_and_bn,
            "merges": merges, "dropout_pad_act_gap": dropout_pad_act_gap}
