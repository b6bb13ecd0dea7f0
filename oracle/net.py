"""Whole-network forward / backward / training loop on top of oracle.ops (float64, CPU).

TEST INFRASTRUCTURE ONLY; PARITY UNPINNED (see oracle/ops.py for both statements).

``OracleNet`` consumes the plain layer rows of ``Network.describe()`` (class name, config,
connectivity) -- data, not code, so the oracle shares no arithmetic with the CUDA engine.
``fit`` restates the loop of /root/reference/helpers.py:69-175 as Keras 2.2.4 runs it:
per-epoch LearningRateScheduler -> batches (last partial batch kept) -> validation with
the training batch size -> ReduceLROnPlateau / EarlyStopping(val_acc, +0.01, 20) /
ModelCheckpoint(best val_loss) -> best weights restored at the end.  Everything random
(initial weights, per-epoch permutations, dropout masks) is an input.
"""
import math

import numpy as np

from . import ops

F64 = np.float64

WEIGHT_NAMES = {
    "Conv2D": ("kernel", "bias"), "DepthwiseConv2D": ("depthwise_kernel", "bias"),
    "SeparableConv2D": ("depthwise_kernel", "pointwise_kernel", "bias"), "Dense": ("kernel", "bias"),
    "BatchNormalization": ("gamma", "beta", "moving_mean", "moving_variance"),
}


def lr_schedule(epoch):
    """/root/reference/helpers.py:45-67."""
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


class OracleNet(object):
    def __init__(self, rows, weights):
        """rows: Network.describe(); weights: list (per layer) of lists of arrays in Keras order."""
        self.rows = rows
        self.weights = [[np.asarray(w, F64).copy() for w in ws] for ws in weights]
        self.t = 0                      # optimizer iterations
        self.adam_m = [[np.zeros_like(w) for w in ws] for ws in self.weights]
        self.adam_v = [[np.zeros_like(w) for w in ws] for ws in self.weights]

    # ------------------------------------------------------------------ forward
    def forward(self, x, training=False, dropout_masks=None):
        """Returns the list of layer outputs; keeps what backward needs in self.cache."""
        outs, cache = [], []
        dropout_masks = dropout_masks or {}
        for i, row in enumerate(self.rows):
            cls, cfg, W = row["class"], row["cfg"], self.weights[i]
            ins = [outs[j] for j in row["inbound"]]
            c = None
            if cls == "InputLayer":
                y = np.asarray(x, F64)
            elif cls == "Conv2D":
                z = ops.conv2d_fwd(ins[0], W[0], W[1], cfg["strides"], cfg["padding"], cfg["dilation"])
                y = ops.act_fwd(z, cfg["activation"])
            elif cls == "DepthwiseConv2D":
                z = ops.dwconv2d_fwd(ins[0], W[0], W[1], cfg["strides"], cfg["padding"])
                y = ops.act_fwd(z, cfg["activation"])
            elif cls == "SeparableConv2D":
                mid = ops.dwconv2d_fwd(ins[0], W[0], None, cfg["strides"], cfg["padding"])
                z = ops.conv2d_fwd(mid, W[1], W[2], (1, 1), "valid")
                y = ops.act_fwd(z, cfg["activation"])
                c = mid
            elif cls == "Dense":
                z = ops.dense_fwd(ins[0], W[0], W[1])
                y = ops.act_fwd(z, cfg["activation"])
                if i == len(self.rows) - 1:
                    self.logits = z                 # what ART's predict(logits=True) / class_gradient(logits=True) read
            elif cls == "MaxPooling2D":
                y, c = ops.maxpool_fwd(ins[0], cfg["pool"], cfg["strides"], cfg["padding"])
            elif cls == "AveragePooling2D":
                y = ops.avgpool_fwd(ins[0], cfg["pool"], cfg["strides"], cfg["padding"])
            elif cls == "GlobalAveragePooling2D":
                y = ops.gap_fwd(ins[0])
            elif cls == "Flatten":
                y = ins[0].reshape(ins[0].shape[0], -1)
            elif cls == "Dropout":
                if training:
                    mask = np.asarray(dropout_masks[i], F64).reshape(ins[0].shape)
                    y = ops.dropout_fwd(ins[0], mask, cfg["rate"])
                    c = mask
                else:
                    y = ins[0]
            elif cls == "BatchNormalization":
                if training:
                    y, mean, var, var_unbiased = ops.bn_train_fwd(ins[0], W[0], W[1], cfg["axis"])
                    c = (mean, var, var_unbiased)
                else:
                    y = ops.bn_infer_fwd(ins[0], W[0], W[1], W[2], W[3], cfg["axis"])
            elif cls == "Activation":
                y = ops.act_fwd(ins[0], cfg["activation"])
            elif cls == "Add":
                y = ins[0] + ins[1]
            elif cls == "Multiply":
                y = ins[0] * ins[1]
            elif cls == "Concatenate":
                y = ops.concat_fwd(ins[0], ins[1], cfg["axis"])
            elif cls == "ZeroPadding2D":
                y = ops.pad2d_fwd(ins[0], cfg["pad"])
            else:
                raise NotImplementedError(cls)
            outs.append(y)
            cache.append(c)
        self.outs, self.cache = outs, cache
        return outs

    # ------------------------------------------------------------------ backward
    def input_gradient(self, x, targets):
        """dL/dx of the mean-over-batch CCE in INFERENCE mode (dropout off, BatchNorm on its moving statistics):
        what an adversarial attack differentiates (ART's KerasClassifier.loss_gradient runs with learning_phase 0)."""
        self.forward(x, training=False)
        self.backward(targets, training=False)
        return self.douts[0]

    def logit_gradient(self, x, head_dz):
        """sum_j head_dz[b, j] * dZ_j/dx[b] for the LOGITS Z of the head, in inference mode: one row of ART's
        ``class_gradient(x, logits=True)`` per one-hot head_dz, any linear combination of rows otherwise
        (model/metrics.py:309-314 uses grads[:, pred] - grads[:, target])."""
        self.forward(x, training=False)
        self.backward(None, training=False, head_dz=np.asarray(head_dz, F64))
        return self.douts[0]

    def backward(self, targets, training=True, head_dz=None):
        """Gradients of mean-over-batch CCE for the last forward. Returns per-layer weight grads; self.douts[0] is dL/dx.
        head_dz: start from this gradient with respect to the head's logits instead of from the loss."""
        rows, outs = self.rows, self.outs
        B = outs[0].shape[0]
        douts = [None] * len(rows)
        if head_dz is None:
            douts[-1] = ops.cce_bwd_probs(outs[-1], np.asarray(targets, F64), 1.0 / B)
        else:
            douts[-1] = head_dz
        grads = [[np.zeros_like(w) for w in ws] for ws in self.weights]

        def give(j, g):
            douts[j] = g if douts[j] is None else douts[j] + g

        for i in range(len(rows) - 1, -1, -1):
            row = rows[i]
            cls, cfg, W, dy = row["class"], row["cfg"], self.weights[i], douts[i]
            if dy is None or cls == "InputLayer":
                continue
            if not training and cls == "Dropout":
                give(row["inbound"][0], dy)
                continue
            if not training and cls == "BatchNormalization":
                give(row["inbound"][0], ops.bn_infer_bwd(dy, W[0], W[3], cfg["axis"]))
                continue
            src = row["inbound"]
            x = outs[src[0]] if src else None
            y = outs[i]
            if cls == "Conv2D":
                dz = ops.act_bwd(y, dy, cfg["activation"])
                dx, grads[i][0], grads[i][1] = ops.conv2d_bwd(x, W[0], dz, cfg["strides"], cfg["padding"], cfg["dilation"])
                give(src[0], dx)
            elif cls == "DepthwiseConv2D":
                dz = ops.act_bwd(y, dy, cfg["activation"])
                dx, grads[i][0], grads[i][1] = ops.dwconv2d_bwd(x, W[0], dz, cfg["strides"], cfg["padding"])
                give(src[0], dx)
            elif cls == "SeparableConv2D":
                dz = ops.act_bwd(y, dy, cfg["activation"])
                dmid, grads[i][1], grads[i][2] = ops.conv2d_bwd(self.cache[i], W[1], dz, (1, 1), "valid")
                dx, grads[i][0], _ = ops.dwconv2d_bwd(x, W[0], dmid, cfg["strides"], cfg["padding"])
                give(src[0], dx)
            elif cls == "Dense":
                dz = dy if (head_dz is not None and i == len(rows) - 1) else ops.act_bwd(y, dy, cfg["activation"])
                dx, grads[i][0], grads[i][1] = ops.dense_bwd(x, W[0], dz)
                give(src[0], dx)
            elif cls == "MaxPooling2D":
                give(src[0], ops.maxpool_bwd(x.shape, self.cache[i], dy, cfg["pool"], cfg["strides"], cfg["padding"]))
            elif cls == "AveragePooling2D":
                give(src[0], ops.avgpool_bwd(x.shape, dy, cfg["pool"], cfg["strides"], cfg["padding"]))
            elif cls == "GlobalAveragePooling2D":
                give(src[0], ops.gap_bwd(x.shape, dy))
            elif cls == "Flatten":
                give(src[0], dy.reshape(x.shape))
            elif cls == "Dropout":
                give(src[0], ops.dropout_bwd(dy, self.cache[i], cfg["rate"]))
            elif cls == "BatchNormalization":
                mean, var, _ = self.cache[i]
                dx, grads[i][0], grads[i][1] = ops.bn_train_bwd(x, W[0], mean, var, dy, cfg["axis"])
                give(src[0], dx)
            elif cls == "Activation":
                give(src[0], ops.act_bwd(y, dy, cfg["activation"]))
            elif cls == "Add":
                give(src[0], dy)
                give(src[1], dy)
            elif cls == "Multiply":
                give(src[0], dy * outs[src[1]])
                give(src[1], dy * outs[src[0]])
            elif cls == "Concatenate":
                da, db = ops.concat_bwd(dy, outs[src[0]].shape[cfg["axis"]], cfg["axis"])
                give(src[0], da)
                give(src[1], db)
            elif cls == "ZeroPadding2D":
                give(src[0], ops.pad2d_bwd(dy, cfg["pad"]))
            else:
                raise NotImplementedError(cls)
        self.douts = douts
        return grads

    # ------------------------------------------------------------------ one optimisation step
    def train_step(self, x, targets, lr, dropout_masks=None, optimizer="adam", after_forward=None):
        """Forward (training phase), loss/accuracy of that forward, backward, update. Returns (loss_sum, n_correct).
        after_forward(self) may adjust self.outs before the backward pass (tests align ReLU kinks with it)."""
        outs = self.forward(x, training=True, dropout_masks=dropout_masks)
        loss, correct = ops.cce_fwd(outs[-1], np.asarray(targets, F64))
        if after_forward is not None:
            after_forward(self)
        grads = self.backward(targets)
        self.t += 1
        for i, row in enumerate(self.rows):
            names = WEIGHT_NAMES.get(row["class"], ())
            for k, name in enumerate(names):
                if name.startswith("moving_"):
                    continue
                if optimizer == "adam":
                    self.weights[i][k], self.adam_m[i][k], self.adam_v[i][k] = ops.adam_step(
                        self.weights[i][k], grads[i][k], self.adam_m[i][k], self.adam_v[i][k], self.t, lr)
                else:
                    self.weights[i][k], self.adam_m[i][k] = ops.sgd_step(
                        self.weights[i][k], grads[i][k], self.adam_m[i][k], lr)
            if row["class"] == "BatchNormalization":
                mean, _, var_unbiased = self.cache[i]
                self.weights[i][2] = ops.bn_moving_update(self.weights[i][2], mean)
                self.weights[i][3] = ops.bn_moving_update(self.weights[i][3], var_unbiased)
        self.last_grads = grads
        return float(loss.sum()), int(correct.sum())

    def evaluate(self, x, targets, batch_size=32):
        """keras Model.evaluate: sample-weighted mean loss and accuracy, inference phase."""
        n = x.shape[0]
        loss_sum, n_correct = 0.0, 0
        for lo in range(0, n, batch_size):
            outs = self.forward(x[lo: lo + batch_size], training=False)
            loss, correct = ops.cce_fwd(outs[-1], np.asarray(targets[lo: lo + batch_size], F64))
            loss_sum += float(loss.sum())
            n_correct += int(correct.sum())
        return loss_sum / n, n_correct / n

    def snapshot(self):
        return [[w.copy() for w in ws] for ws in self.weights]

    # ------------------------------------------------------------------ helpers.train_model
    def fit(self, x, y, x_val, y_val, epochs, batch_size, permutations, dropout_mask_fn=None, scheduler=True):
        """helpers.train_model(..., scheduler=True, data_augmentation=False) as Keras 2.2.4 executes it.

        permutations[e] is the sample order of epoch e; dropout_mask_fn(step, layer_index, shape) -> 0/1 mask.
        Returns history dict; leaves the best-val_loss weights installed (helpers.py:169).
        """
        history = {"loss": [], "acc": [], "val_loss": [], "val_acc": [], "lr": []}
        n = x.shape[0]
        lr = lr_schedule(0)
        plateau_best, plateau_wait = np.inf, 0                 # ReduceLROnPlateau(monitor val_loss)
        stop_best, stop_wait = -np.inf, 0                      # EarlyStopping(val_acc, max, 0.01, 20)
        ckpt_best, best_weights = np.inf, self.snapshot()      # ModelCheckpoint(val_loss, min)
        step = 0
        for epoch in range(epochs):
            if scheduler:
                lr = lr_schedule(epoch)                        # LearningRateScheduler.on_epoch_begin
            order = permutations[epoch]
            loss_sum, n_correct = 0.0, 0
            for lo in range(0, n, batch_size):
                idx = order[lo: lo + batch_size]
                masks = None
                if dropout_mask_fn is not None:
                    masks = {i: dropout_mask_fn(step, i, (len(idx),) + tuple(self.rows[i]["out_shape"][1:]))
                             for i, row in enumerate(self.rows) if row["class"] == "Dropout"}
                ls, nc = self.train_step(x[idx], y[idx], lr, masks)
                loss_sum += ls
                n_correct += nc
                step += 1
            val_loss, val_acc = self.evaluate(x_val, y_val, batch_size)
            history["loss"].append(loss_sum / n)
            history["acc"].append(n_correct / n)
            history["val_loss"].append(val_loss)
            history["val_acc"].append(val_acc)
            history["lr"].append(lr)
            if scheduler:
                # ReduceLROnPlateau.on_epoch_end (factor sqrt(0.1), patience 5, min_lr 0.5e-6, min_delta 1e-4)
                if val_loss < plateau_best - 1e-4:
                    plateau_best, plateau_wait = val_loss, 0
                else:
                    plateau_wait += 1
                    if plateau_wait >= 5:
                        if lr > 0.5e-6:
                            lr = max(lr * math.sqrt(0.1), 0.5e-6)
                            plateau_wait = 0
                # EarlyStopping.on_epoch_end
                if val_acc - 0.01 > stop_best:
                    stop_best, stop_wait = val_acc, 0
                else:
                    stop_wait += 1
            if val_loss < ckpt_best:                           # ModelCheckpoint(save_best_only)
                ckpt_best, best_weights = val_loss, self.snapshot()
            if scheduler and stop_wait >= 20:
                break
        self.weights = best_weights
        return history
