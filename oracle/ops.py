"""CPU restatement (NumPy, float64, NHWC) of the arithmetic on FeatureNet's train+evaluate path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``featurenet_b200/`` imports this package; only
``tests/``, ``__graft_entry__.smoke()`` and the CPU-baseline legs of ``bench.py`` do, and
only as the checker.

PARITY UNPINNED.  The reference (/root/reference) contains no arithmetic of its own: every
layer is a Keras 2.2.4 call (README:5) executed by TensorFlow 1.12, neither of which is
vendored, installed or installable here (Python 3.12, no network).  The reference has no
numeric tests, golden vectors or fixtures for this path either (tests_mutant.py and
tests_pledge.py assert nothing).  So each function below restates the *published* behaviour
of the Keras/TF op named in its docstring, cites the reference call site that selects it,
and is cross-checked in tests/test_oracle.py against an independent implementation
(torch CPU float64 ops and finite differences) -- a consistency check, not a pin.

Third-party semantics assumed (each is exercised by a test):
  * TF SAME padding: out = ceil(in/stride), total = max((out-1)*stride + k_eff - in, 0),
    the smaller half in front.
  * AvgPool SAME divides by the number of in-bounds taps; MaxPoolGrad routes the gradient to
    the first maximum in row-major window order.
  * Keras BatchNormalization (non-fused path, axis != -1): biased batch variance for the
    normalisation, moving variance fed with var * n / (n - (1 + eps)); momentum 0.99, eps 1e-3.
  * tf.nn.dropout: kept units scaled by 1/keep_prob.
  * categorical_crossentropy on probabilities: renormalise, clip to [1e-7, 1-1e-7], -sum t log p.
  * Keras Adam: lr_t = lr * sqrt(1-b2^t)/(1-b1^t), p -= lr_t * m / (sqrt(v) + 1e-7).
"""
import numpy as np

F64 = np.float64


# ----------------------------------------------------------------------------- geometry
def same_or_valid(n, k, s, padding, d=1):
    """(out, pad_before, pad_after) along one axis; TF rule (see module docstring)."""
    k_eff = (k - 1) * d + 1
    if padding == "same":
        out = -(-n // s)
        total = max((out - 1) * s + k_eff - n, 0)
        return out, total // 2, total - total // 2
    out = (n - k_eff) // s + 1
    if n < k_eff or out <= 0:
        raise ValueError("negative output size")
    return out, 0, 0


def _pad_hw(x, ph, pw, value=0.0):
    if ph == (0, 0) and pw == (0, 0):
        return x
    return np.pad(x, ((0, 0), ph, pw, (0, 0)), mode="constant", constant_values=value)


def _tap(xp, r, s, ho, wo, sh, sw, dh=1, dw=1):
    """View of the padded input seen by kernel tap (r, s): shape [B, ho, wo, C]."""
    return xp[:, r * dh: r * dh + (ho - 1) * sh + 1: sh, s * dw: s * dw + (wo - 1) * sw + 1: sw, :]


# ----------------------------------------------------------------------------- activations
def act_fwd(z, kind):
    """keras.activations.{relu,tanh,sigmoid,softmax}; softmax over the last axis."""
    if kind is None or kind == "linear":
        return z
    if kind == "relu":
        return np.maximum(z, 0.0)
    if kind == "tanh":
        return np.tanh(z)
    if kind == "sigmoid":
        return 1.0 / (1.0 + np.exp(-z))
    if kind == "softmax":
        e = np.exp(z - z.max(axis=-1, keepdims=True))
        return e / e.sum(axis=-1, keepdims=True)
    raise ValueError(kind)


def act_bwd(y, dy, kind):
    """Gradient w.r.t. the pre-activation given the activation's OUTPUT y."""
    if kind is None or kind == "linear":
        return dy
    if kind == "relu":
        return dy * (y > 0)
    if kind == "tanh":
        return dy * (1.0 - y * y)
    if kind == "sigmoid":
        return dy * y * (1.0 - y)
    if kind == "softmax":
        return y * (dy - (dy * y).sum(axis=-1, keepdims=True))
    raise ValueError(kind)


# ----------------------------------------------------------------------------- convolutions
def conv2d_fwd(x, w, b, strides=(1, 1), padding="same", dilation=(1, 1)):
    """keras Conv2D pre-activation (model/input.py:294, model/operation.py:186). w is HWIO."""
    B, H, W, C = x.shape
    kh, kw, _, K = w.shape
    ho, pt, pb = same_or_valid(H, kh, strides[0], padding, dilation[0])
    wo, pl, pr = same_or_valid(W, kw, strides[1], padding, dilation[1])
    xp = _pad_hw(x, (pt, pb), (pl, pr))
    y = np.zeros((B, ho, wo, K), F64)
    for r in range(kh):
        for s in range(kw):
            y += _tap(xp, r, s, ho, wo, strides[0], strides[1], dilation[0], dilation[1]) @ w[r, s]
    if b is not None:
        y += b
    return y


def conv2d_bwd(x, w, dz, strides=(1, 1), padding="same", dilation=(1, 1)):
    """Gradients of conv2d_fwd w.r.t. x, w, b given dz = dL/d(pre-activation)."""
    B, H, W, C = x.shape
    kh, kw, _, K = w.shape
    ho, pt, pb = same_or_valid(H, kh, strides[0], padding, dilation[0])
    wo, pl, pr = same_or_valid(W, kw, strides[1], padding, dilation[1])
    xp = _pad_hw(x, (pt, pb), (pl, pr))
    dxp = np.zeros_like(xp)
    dw = np.zeros_like(w)
    dz2 = dz.reshape(-1, K)
    for r in range(kh):
        for s in range(kw):
            tap = _tap(xp, r, s, ho, wo, strides[0], strides[1], dilation[0], dilation[1])
            dw[r, s] = tap.reshape(-1, C).T @ dz2
            _tap(dxp, r, s, ho, wo, strides[0], strides[1], dilation[0], dilation[1])[...] += dz @ w[r, s].T
    dx = dxp[:, pt: pt + H, pl: pl + W, :]
    return dx, dw, dz2.sum(axis=0)


def dwconv2d_fwd(x, w, b, strides=(1, 1), padding="same"):
    """keras DepthwiseConv2D pre-activation, depth multiplier 1 (model/input.py:298). w is [kh,kw,C,1]."""
    B, H, W, C = x.shape
    kh, kw = w.shape[:2]
    ho, pt, pb = same_or_valid(H, kh, strides[0], padding)
    wo, pl, pr = same_or_valid(W, kw, strides[1], padding)
    xp = _pad_hw(x, (pt, pb), (pl, pr))
    y = np.zeros((B, ho, wo, C), F64)
    for r in range(kh):
        for s in range(kw):
            y += _tap(xp, r, s, ho, wo, strides[0], strides[1]) * w[r, s, :, 0]
    if b is not None:
        y += b
    return y


def dwconv2d_bwd(x, w, dz, strides=(1, 1), padding="same"):
    B, H, W, C = x.shape
    kh, kw = w.shape[:2]
    ho, pt, pb = same_or_valid(H, kh, strides[0], padding)
    wo, pl, pr = same_or_valid(W, kw, strides[1], padding)
    xp = _pad_hw(x, (pt, pb), (pl, pr))
    dxp = np.zeros_like(xp)
    dw = np.zeros_like(w)
    for r in range(kh):
        for s in range(kw):
            tap = _tap(xp, r, s, ho, wo, strides[0], strides[1])
            dw[r, s, :, 0] = (tap * dz).sum(axis=(0, 1, 2))
            _tap(dxp, r, s, ho, wo, strides[0], strides[1])[...] += dz * w[r, s, :, 0]
    return dxp[:, pt: pt + H, pl: pl + W, :], dw, dz.sum(axis=(0, 1, 2))


def dense_fwd(x, w, b):
    """keras Dense pre-activation over the last axis of a rank-2 or rank-4 tensor (model/input.py:190)."""
    return x @ w + b


def dense_bwd(x, w, dz):
    k, n = w.shape
    x2, dz2 = x.reshape(-1, k), dz.reshape(-1, n)
    return dz @ w.T, x2.T @ dz2, dz2.sum(axis=0)


# ----------------------------------------------------------------------------- pooling
def maxpool_fwd(x, pool, strides, padding="same"):
    """keras MaxPooling2D (model/input.py:233). Returns (y, argmax) with argmax = flat tap index r*kw+s."""
    B, H, W, C = x.shape
    ho, pt, pb = same_or_valid(H, pool[0], strides[0], padding)
    wo, pl, pr = same_or_valid(W, pool[1], strides[1], padding)
    xp = _pad_hw(x, (pt, pb), (pl, pr), value=-np.inf)
    best = np.full((B, ho, wo, C), -np.inf, F64)
    arg = np.zeros((B, ho, wo, C), np.int32)
    for r in range(pool[0]):
        for s in range(pool[1]):
            tap = _tap(xp, r, s, ho, wo, strides[0], strides[1])
            better = tap > best            # strict: the first maximum in row-major order wins
            best = np.where(better, tap, best)
            arg = np.where(better, r * pool[1] + s, arg)
    return best, arg


def maxpool_bwd(x_shape, arg, dy, pool, strides, padding="same"):
    B, H, W, C = x_shape
    ho, pt, pb = same_or_valid(H, pool[0], strides[0], padding)
    wo, pl, pr = same_or_valid(W, pool[1], strides[1], padding)
    dxp = np.zeros((B, H + pt + pb, W + pl + pr, C), F64)
    for r in range(pool[0]):
        for s in range(pool[1]):
            _tap(dxp, r, s, ho, wo, strides[0], strides[1])[...] += dy * (arg == r * pool[1] + s)
    return dxp[:, pt: pt + H, pl: pl + W, :]


def _avg_counts(H, W, pool, strides, padding):
    ho, pt, pb = same_or_valid(H, pool[0], strides[0], padding)
    wo, pl, pr = same_or_valid(W, pool[1], strides[1], padding)
    ones = _pad_hw(np.ones((1, H, W, 1), F64), (pt, pb), (pl, pr))
    cnt = np.zeros((1, ho, wo, 1), F64)
    for r in range(pool[0]):
        for s in range(pool[1]):
            cnt += _tap(ones, r, s, ho, wo, strides[0], strides[1])
    return cnt, (ho, wo, pt, pb, pl, pr)


def avgpool_fwd(x, pool, strides, padding="same"):
    """keras AveragePooling2D (model/input.py:237); the divisor counts in-bounds taps only."""
    B, H, W, C = x.shape
    cnt, (ho, wo, pt, pb, pl, pr) = _avg_counts(H, W, pool, strides, padding)
    xp = _pad_hw(x, (pt, pb), (pl, pr))
    y = np.zeros((B, ho, wo, C), F64)
    for r in range(pool[0]):
        for s in range(pool[1]):
            y += _tap(xp, r, s, ho, wo, strides[0], strides[1])
    return y / cnt


def avgpool_bwd(x_shape, dy, pool, strides, padding="same"):
    B, H, W, C = x_shape
    cnt, (ho, wo, pt, pb, pl, pr) = _avg_counts(H, W, pool, strides, padding)
    dxp = np.zeros((B, H + pt + pb, W + pl + pr, C), F64)
    g = dy / cnt
    for r in range(pool[0]):
        for s in range(pool[1]):
            _tap(dxp, r, s, ho, wo, strides[0], strides[1])[...] += g
    return dxp[:, pt: pt + H, pl: pl + W, :]


def gap_fwd(x):
    """keras GlobalAveragePooling2D (model/input.py:240)."""
    return x.mean(axis=(1, 2))


def gap_bwd(x_shape, dy):
    B, H, W, C = x_shape
    return np.broadcast_to(dy[:, None, None, :] / (H * W), x_shape).copy()


# ----------------------------------------------------------------------------- batch norm
def _bn_axes(ndim, axis):
    return tuple(a for a in range(ndim) if a != axis)


def _bn_bcast(v, ndim, axis):
    shape = [1] * ndim
    shape[axis] = -1
    return v.reshape(shape)


def bn_train_fwd(x, gamma, beta, axis=1, eps=1e-3):
    """keras BatchNormalization, training phase (model/operation.py:142,150: axis is always 1).

    Returns y, batch mean, biased batch variance, and the value fed to the moving variance.
    """
    axes = _bn_axes(x.ndim, axis)
    n = 1
    for a in axes:
        n *= x.shape[a]
    mean = x.mean(axis=axes)
    var = ((x - _bn_bcast(mean, x.ndim, axis)) ** 2).mean(axis=axes)
    inv = gamma / np.sqrt(var + eps)
    y = x * _bn_bcast(inv, x.ndim, axis) + _bn_bcast(beta - mean * inv, x.ndim, axis)
    return y, mean, var, var * n / (n - (1.0 + eps))


def bn_train_bwd(x, gamma, mean, var, dy, axis=1, eps=1e-3):
    axes = _bn_axes(x.ndim, axis)
    n = 1
    for a in axes:
        n *= x.shape[a]
    rstd = 1.0 / np.sqrt(var + eps)
    xhat = (x - _bn_bcast(mean, x.ndim, axis)) * _bn_bcast(rstd, x.ndim, axis)
    dbeta = dy.sum(axis=axes)
    dgamma = (dy * xhat).sum(axis=axes)
    dx = _bn_bcast(gamma * rstd / n, x.ndim, axis) * (
        n * dy - _bn_bcast(dbeta, x.ndim, axis) - xhat * _bn_bcast(dgamma, x.ndim, axis))
    return dx, dgamma, dbeta


def bn_infer_fwd(x, gamma, beta, moving_mean, moving_var, axis=1, eps=1e-3):
    inv = gamma / np.sqrt(moving_var + eps)
    return x * _bn_bcast(inv, x.ndim, axis) + _bn_bcast(beta - moving_mean * inv, x.ndim, axis)


def bn_infer_bwd(dy, gamma, moving_var, axis=1, eps=1e-3):
    """Input gradient of the inference-mode affine map (the statistics are constants)."""
    return dy * _bn_bcast(gamma / np.sqrt(moving_var + eps), dy.ndim, axis)


def bn_moving_update(moving, value, momentum=0.99):
    """K.moving_average_update: moving <- moving*momentum + value*(1-momentum)."""
    return moving * momentum + value * (1.0 - momentum)


# ----------------------------------------------------------------------------- dropout
def dropout_fwd(x, keep_mask, rate):
    """tf.nn.dropout given the 0/1 keep mask (model/operation.py:114): kept units are scaled by 1/(1-rate)."""
    return x * keep_mask / (1.0 - rate)


dropout_bwd = dropout_fwd      # same linear map applied to dy


_PHILOX_M0, _PHILOX_M1 = 0xD2511F53, 0xCD9E8D57
_PHILOX_W0, _PHILOX_W1 = 0x9E3779B9, 0xBB67AE85


def philox4x32_10(counter, key):
    """Philox-4x32-10 on uint32 arrays: counter [n,4], key (k0,k1). The engine's dropout stream."""
    c = [counter[:, i].astype(np.uint64) for i in range(4)]
    k0, k1 = np.uint64(key[0]), np.uint64(key[1])
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0 = np.uint64(_PHILOX_M0) * c[0]
        p1 = np.uint64(_PHILOX_M1) * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & mask
        hi1, lo1 = p1 >> np.uint64(32), p1 & mask
        c = [(hi1 ^ c[1] ^ k0) & mask, lo1, (hi0 ^ c[3] ^ k1) & mask, lo0]
        k0 = (k0 + np.uint64(_PHILOX_W0)) & mask
        k1 = (k1 + np.uint64(_PHILOX_W1)) & mask
    return np.stack([v.astype(np.uint32) for v in c], axis=1)


def dropout_keep_mask(n, rate, seed, stream, step):
    """The engine's mask: element i uses word i%4 of philox(counter=(i//4, step, stream, 0), key=(seed lo, hi)).

    keep <=> uniform >= rate with uniform = word * 2^-32 (integer compare: word >= ceil(rate * 2^32)).
    """
    blocks = (n + 3) // 4
    counter = np.zeros((blocks, 4), np.uint32)
    counter[:, 0] = np.arange(blocks, dtype=np.uint64) & 0xFFFFFFFF
    counter[:, 1] = np.uint32(step & 0xFFFFFFFF)
    counter[:, 2] = np.uint32(stream & 0xFFFFFFFF)
    words = philox4x32_10(counter, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)).reshape(-1)[:n]
    threshold = dropout_threshold(rate)
    return (words.astype(np.uint64) >= np.uint64(threshold)).astype(F64)


def dropout_threshold(rate):
    import math
    # the rate crosses the C ABI as a float: the threshold is taken from its fp32 value
    return min(int(math.ceil(float(np.float32(rate)) * 4294967296.0)), 4294967296)


# ----------------------------------------------------------------------------- merges / reshapes
def concat_fwd(a, b, axis=1):
    """keras Concatenate (model/operation.py:238; axis is always 1: along H for NHWC)."""
    return np.concatenate([a, b], axis=axis)


def concat_bwd(dy, size_a, axis=1):
    idx_a = [slice(None)] * dy.ndim
    idx_b = [slice(None)] * dy.ndim
    idx_a[axis] = slice(0, size_a)
    idx_b[axis] = slice(size_a, None)
    return dy[tuple(idx_a)], dy[tuple(idx_b)]


def pad2d_fwd(x, pad):
    """keras ZeroPadding2D((ph, pw)) (model/operation.py:136)."""
    return np.pad(x, ((0, 0), (pad[0], pad[0]), (pad[1], pad[1]), (0, 0)))


def pad2d_bwd(dy, pad):
    H, W = dy.shape[1] - 2 * pad[0], dy.shape[2] - 2 * pad[1]
    return dy[:, pad[0]: pad[0] + H, pad[1]: pad[1] + W, :]


# ----------------------------------------------------------------------------- loss / metric
EPS = 1e-7


def cce_fwd(p, t):
    """keras categorical_crossentropy on probabilities + categorical_accuracy (tensorflow_generator.py:233-235).

    Returns per-sample loss and per-sample correctness (argmax ties -> first index).
    """
    q = p / p.sum(axis=-1, keepdims=True)
    c = np.clip(q, EPS, 1.0 - EPS)
    loss = -(t * np.log(c)).sum(axis=-1)
    correct = (np.argmax(t, axis=-1) == np.argmax(p, axis=-1))
    return loss, correct


def cce_bwd_probs(p, t, scale):
    """dL/dp of scale * sum_samples cce (through the renormalisation and the clip)."""
    s = p.sum(axis=-1, keepdims=True)
    q = p / s
    inside = (q >= EPS) & (q <= 1.0 - EPS)
    c = np.clip(q, EPS, 1.0 - EPS)
    gq = -(t / c) * inside
    return scale * (gq - (gq * q).sum(axis=-1, keepdims=True)) / s


# ----------------------------------------------------------------------------- optimizers
def adam_step(p, g, m, v, t, lr, beta1=0.9, beta2=0.999, eps=1e-7):
    """keras.optimizers.Adam.get_updates, iteration t (1-based) (tensorflow_generator.py:232)."""
    lr_t = lr * np.sqrt(1.0 - beta2 ** t) / (1.0 - beta1 ** t)
    m = beta1 * m + (1.0 - beta1) * g
    v = beta2 * v + (1.0 - beta2) * g * g
    return p - lr_t * m / (np.sqrt(v) + eps), m, v


def sgd_step(p, g, mom, lr, momentum=0.0, nesterov=False):
    """keras.optimizers.SGD.get_updates (the "sgd" of lenet5.py:21 and of the unused list entry)."""
    vel = momentum * mom - lr * g
    new_p = p + momentum * vel - lr * g if nesterov else p + vel
    return new_p, vel


def glorot_uniform(rng, shape, fans):
    limit = np.sqrt(6.0 / (fans[0] + fans[1]))
    return rng.uniform(-limit, limit, size=shape)
