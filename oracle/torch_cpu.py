"""fp32, multi-threaded CPU port of the train step, used ONLY as the timed CPU baseline.

TEST INFRASTRUCTURE / BASELINE ONLY; PARITY UNPINNED (see oracle/ops.py).  The reference's CPU
path is TensorFlow 1.12 + Keras 2.2.4, which cannot be installed here.
This module stands in for it: the same layer rows executed by torch's CPU kernels (MKL-DNN /
OpenMP, all host threads), backward by autograd, Keras-formula Adam.  tests/test_oracle_torch.py
checks it against the float64 NumPy oracle, so the thing being timed computes the same step.
"""
import numpy as np
import torch
import torch.nn.functional as F

from .ops import same_or_valid

ACT = {None: lambda z: z, "linear": lambda z: z, "relu": torch.relu, "tanh": torch.tanh,
       "sigmoid": torch.sigmoid, "softmax": lambda z: torch.softmax(z, dim=-1)}


def _pad_same(x, k, s, padding, d=(1, 1)):
    """x is NCHW; returns x padded the TF way."""
    _, pt, pb = same_or_valid(x.shape[2], k[0], s[0], padding, d[0])
    _, pl, pr = same_or_valid(x.shape[3], k[1], s[1], padding, d[1])
    return F.pad(x, (pl, pr, pt, pb)) if (pt or pb or pl or pr) else x


class TorchCpuNet(object):
    """Activations are kept NCHW (torch's fast layout); weights are converted once from Keras layouts."""

    def __init__(self, rows, weights, threads=None, dtype=torch.float32):
        """dtype=torch.float64 turns the port into a fast float64 checker for layers and candidates at BASELINE's
        full sizes (batch 128), where the NumPy oracle of oracle/net.py takes minutes."""
        if threads:
            torch.set_num_threads(int(threads))
        self.dtype = dtype
        self.np_dtype = np.float64 if dtype == torch.float64 else np.float32
        self.rows = rows
        self.params = []          # per layer: list of torch tensors (requires_grad for trainable ones)
        for row, ws in zip(rows, weights):
            cls, ts = row["class"], []
            for k, w in enumerate(ws):
                t = torch.tensor(np.asarray(w, self.np_dtype))
                if cls in ("Conv2D",) and k == 0:
                    t = t.permute(3, 2, 0, 1).contiguous()                       # HWIO -> OIHW
                elif cls in ("DepthwiseConv2D", "SeparableConv2D") and k == 0:
                    t = t.permute(2, 3, 0, 1).contiguous()                       # [kh,kw,C,1] -> [C,1,kh,kw]
                elif cls == "SeparableConv2D" and k == 1:
                    t = t.permute(3, 2, 0, 1).contiguous()
                trainable = not (cls == "BatchNormalization" and k >= 2)
                t.requires_grad_(trainable)
                ts.append(t)
            self.params.append(ts)
        self.trainable = [t for ts in self.params for t in ts if t.requires_grad]
        self.m = [torch.zeros_like(t) for t in self.trainable]
        self.v = [torch.zeros_like(t) for t in self.trainable]
        self.t = 0

    def forward(self, x_nhwc, training=True, dropout_masks=None, relu_masks=None):
        """relu_masks {layer: 0/1 array in NHWC}: use this mask instead of (z > 0) for that layer's ReLU.  A parity test
        passes the masks of the implementation under test: an element that lands on the other side of the kink within
        fp32 rounding is a difference of the backward pass's INPUTS (it switches a whole gradient path on or off)."""
        outs = []
        relu_masks = relu_masks or {}

        def activate(i, z, name, channel_dim=1):
            if name == "relu" and i in relu_masks:
                mask = torch.as_tensor(np.asarray(relu_masks[i], self.np_dtype))
                return z * (mask.permute(0, 3, 1, 2) if z.dim() == 4 else mask)
            if name == "softmax":
                return torch.softmax(z, dim=channel_dim)
            return ACT[name](z)

        rank4 = []              # whether outs[i] is stored NCHW (True) or as a plain [B, F] tensor
        for i, row in enumerate(self.rows):
            cls, cfg, W = row["class"], row["cfg"], self.params[i]
            ins = [outs[j] for j in row["inbound"]]
            is4 = len(row["out_shape"]) == 4
            if cls == "InputLayer":
                y = x_nhwc.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last) if is4 else x_nhwc
            elif cls == "Conv2D":
                xp = _pad_same(ins[0], cfg["kernel"], cfg["strides"], cfg["padding"], cfg["dilation"])
                y = activate(i, F.conv2d(xp, W[0], W[1], stride=tuple(cfg["strides"]), dilation=tuple(cfg["dilation"])),
                             cfg["activation"])
            elif cls == "DepthwiseConv2D":
                xp = _pad_same(ins[0], cfg["kernel"], cfg["strides"], cfg["padding"])
                y = activate(i, F.conv2d(xp, W[0], W[1], stride=tuple(cfg["strides"]), groups=xp.shape[1]), cfg["activation"])
            elif cls == "SeparableConv2D":
                xp = _pad_same(ins[0], cfg["kernel"], cfg["strides"], cfg["padding"])
                mid = F.conv2d(xp, W[0], None, stride=tuple(cfg["strides"]), groups=xp.shape[1])
                y = activate(i, F.conv2d(mid, W[1], W[2]), cfg["activation"])
            elif cls == "Dense":
                if rank4[row["inbound"][0]]:
                    y = activate(i, F.conv2d(ins[0], W[0].t().reshape(W[0].shape[1], W[0].shape[0], 1, 1), W[1]), cfg["activation"])
                else:
                    y = activate(i, ins[0] @ W[0] + W[1], cfg["activation"], channel_dim=-1)
            elif cls == "MaxPooling2D":
                xp = _pad_same(ins[0], cfg["pool"], cfg["strides"], cfg["padding"])
                if xp is not ins[0]:
                    _, pt, pb = same_or_valid(ins[0].shape[2], cfg["pool"][0], cfg["strides"][0], cfg["padding"])
                    _, pl, pr = same_or_valid(ins[0].shape[3], cfg["pool"][1], cfg["strides"][1], cfg["padding"])
                    xp = F.pad(ins[0], (pl, pr, pt, pb), value=float("-inf"))
                y = F.max_pool2d(xp, tuple(cfg["pool"]), tuple(cfg["strides"]))
            elif cls == "AveragePooling2D":
                xp = _pad_same(ins[0], cfg["pool"], cfg["strides"], cfg["padding"])
                ones = _pad_same(torch.ones(1, 1, ins[0].shape[2], ins[0].shape[3], dtype=self.dtype), cfg["pool"], cfg["strides"], cfg["padding"])
                num = F.avg_pool2d(xp, tuple(cfg["pool"]), tuple(cfg["strides"]), divisor_override=1)
                den = F.avg_pool2d(ones, tuple(cfg["pool"]), tuple(cfg["strides"]), divisor_override=1)
                y = num / den
            elif cls == "GlobalAveragePooling2D":
                y = ins[0].mean(dim=(2, 3))
            elif cls == "Flatten":
                y = ins[0].permute(0, 2, 3, 1).reshape(ins[0].shape[0], -1) if rank4[row["inbound"][0]] else ins[0]
            elif cls == "Dropout":
                if training:
                    if dropout_masks is not None:
                        mask = torch.tensor(np.asarray(dropout_masks[i], self.np_dtype)).reshape(
                            (-1,) + tuple(row["out_shape"][1:]))
                        mask = mask.permute(0, 3, 1, 2) if is4 else mask
                        y = ins[0] * mask / (1.0 - cfg["rate"])
                    else:
                        y = F.dropout(ins[0], cfg["rate"], training=True)
                else:
                    y = ins[0]
            elif cls == "BatchNormalization":
                # Keras axis=1 of NHWC = the H axis = dim 2 of NCHW; for rank 2 it is the feature axis
                xin = ins[0].transpose(1, 2) if is4 else ins[0]
                if training:
                    dims = [d for d in range(xin.dim()) if d != 1]
                    n = xin.numel() // xin.shape[1]
                    mean = xin.mean(dim=dims)
                    var = xin.var(dim=dims, unbiased=False)
                    with torch.no_grad():
                        W[2].mul_(0.99).add_(0.01 * mean)
                        W[3].mul_(0.99).add_(0.01 * var * n / (n - 1.001))
                else:
                    mean, var = W[2], W[3]
                shape = [1, -1] + [1] * (xin.dim() - 2)
                inv = W[0] / torch.sqrt(var + 1e-3)
                yb = xin * inv.reshape(shape) + (W[1] - mean * inv).reshape(shape)
                y = yb.transpose(1, 2) if is4 else yb
            elif cls == "Activation":
                y = activate(i, ins[0], cfg["activation"], channel_dim=1 if is4 else -1)
            elif cls == "Add":
                y = ins[0] + ins[1]
            elif cls == "Multiply":
                y = ins[0] * ins[1]
            elif cls == "Concatenate":
                axis = cfg["axis"]
                y = torch.cat(ins, dim={1: 2, 2: 3, 3: 1}[axis] if is4 else axis)
            elif cls == "ZeroPadding2D":
                y = F.pad(ins[0], (cfg["pad"][1], cfg["pad"][1], cfg["pad"][0], cfg["pad"][0]))
            else:
                raise NotImplementedError(cls)
            outs.append(y)
            rank4.append(is4)
        return outs

    @staticmethod
    def loss_and_correct(p, t):
        q = p / p.sum(dim=-1, keepdim=True)
        c = torch.clamp(q, 1e-7, 1.0 - 1e-7)
        loss = -(t * torch.log(c)).sum(dim=-1)
        return loss, (t.argmax(dim=-1) == p.argmax(dim=-1))

    def train_step(self, x, t, lr=1e-3, dropout_masks=None, relu_masks=None):
        x = torch.as_tensor(x, dtype=self.dtype)
        t = torch.as_tensor(t, dtype=self.dtype)
        outs = self.forward(x, True, dropout_masks, relu_masks)
        self.outs = outs
        if getattr(self, "keep_activation_grads", False):       # parity debugging: dL/d(output of every layer)
            for y in outs:
                if y.requires_grad:
                    y.retain_grad()
        loss, correct = self.loss_and_correct(outs[-1], t)
        for p in self.trainable:
            p.grad = None
        loss.mean().backward()
        self.t += 1
        lr_t = lr * np.sqrt(1.0 - 0.999 ** self.t) / (1.0 - 0.9 ** self.t)
        with torch.no_grad():
            for p, m, v in zip(self.trainable, self.m, self.v):
                g = p.grad if p.grad is not None else torch.zeros_like(p)
                m.mul_(0.9).add_(g, alpha=0.1)
                v.mul_(0.999).addcmul_(g, g, value=0.001)
                p.sub_(lr_t * m / (v.sqrt() + 1e-7))
        return float(loss.detach().sum()), int(correct.sum())

    def evaluate(self, x, t, batch_size=256):
        loss_sum, n_correct = 0.0, 0
        with torch.no_grad():
            for lo in range(0, x.shape[0], batch_size):
                xb = torch.as_tensor(x[lo: lo + batch_size], dtype=self.dtype)
                tb = torch.as_tensor(t[lo: lo + batch_size], dtype=self.dtype)
                loss, correct = self.loss_and_correct(self.forward(xb, False)[-1], tb)
                loss_sum += float(loss.sum())
                n_correct += int(correct.sum())
        return loss_sum / x.shape[0], n_correct / x.shape[0]

    # ---- views in the engine's conventions (NHWC activations, Keras weight layouts), for the parity tests
    def out_nhwc(self, i):
        y = self.outs[i].detach()
        return (y.permute(0, 2, 3, 1) if y.dim() == 4 else y).contiguous().numpy()

    def out_grad_nhwc(self, i):
        g = self.outs[i].grad
        if g is None:
            return None
        return (g.permute(0, 2, 3, 1) if g.dim() == 4 else g).contiguous().numpy()

    @staticmethod
    def _to_keras(cls, k, t):
        if cls == "Conv2D" and k == 0:
            return t.permute(2, 3, 1, 0)                                         # OIHW -> HWIO
        if cls in ("DepthwiseConv2D", "SeparableConv2D") and k == 0:
            return t.permute(2, 3, 0, 1)                                         # [C,1,kh,kw] -> [kh,kw,C,1]
        if cls == "SeparableConv2D" and k == 1:
            return t.permute(2, 3, 1, 0)
        return t

    def grads_keras(self):
        """{(layer, weight index): gradient in the Keras layout} of the last train_step."""
        out = {}
        for i, (row, ts) in enumerate(zip(self.rows, self.params)):
            for k, t in enumerate(ts):
                if t.requires_grad:
                    g = t.grad if t.grad is not None else torch.zeros_like(t)
                    out[(i, k)] = self._to_keras(row["class"], k, g.detach()).contiguous().numpy()
        return out

    def weights_keras(self):
        return [self._to_keras(row["class"], k, t.detach()).contiguous().numpy()
                for row, ts in zip(self.rows, self.params) for k, t in enumerate(ts)]

    def fit_epoch(self, x, t, batch_size, permutation, lr=1e-3):
        """One epoch in Keras' order (helpers.py:105): batches cut from the given permutation, ragged last batch kept.
        Returns (mean loss, accuracy) over the epoch's samples."""
        loss_sum, correct = 0.0, 0
        for lo in range(0, x.shape[0], batch_size):
            idx = permutation[lo: lo + batch_size]
            ls, c = self.train_step(x[idx], t[idx], lr)
            loss_sum += ls
            correct += c
        return loss_sum / x.shape[0], correct / x.shape[0]


def conv_products64(x, w, b, dy, strides, padding, dilation=(1, 1)):
    """Forward, data gradient, filter gradient and bias gradient of one Conv2D in float64 (NHWC / HWIO in and out):
    the checker for the three GEMM-shaped products at batch 128.  Dense is the 1x1 case on a [M,1,1,K] tensor."""
    xt = torch.tensor(np.asarray(x, np.float64)).permute(0, 3, 1, 2).requires_grad_(True)
    wt = torch.tensor(np.asarray(w, np.float64)).permute(3, 2, 0, 1).contiguous().requires_grad_(True)
    bt = torch.tensor(np.asarray(b, np.float64)).requires_grad_(True) if b is not None else None
    xp = _pad_same(xt, w.shape[:2], strides, padding, dilation)
    y = F.conv2d(xp, wt, bt, stride=tuple(strides), dilation=tuple(dilation))
    y.backward(torch.tensor(np.asarray(dy, np.float64)).permute(0, 3, 1, 2))
    return (y.detach().permute(0, 2, 3, 1).contiguous().numpy(), xt.grad.permute(0, 2, 3, 1).contiguous().numpy(),
            wt.grad.permute(2, 3, 1, 0).contiguous().numpy(), bt.grad.numpy() if bt is not None else None)
