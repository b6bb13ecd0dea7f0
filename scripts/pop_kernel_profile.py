"""Which (operation, shape) pairs the benchmark population spends its training steps in: every buildable product of the
32-candidate sample, 5 profiled steps each (CUDA events around every launch), aggregated over the population.
Columns: share of the population's step time, launches, average us, achieved GB/s and TFLOP/s (algorithmic)."""
import collections
import contextlib
import copy
import io
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import featurenet_b200  # noqa: E402

featurenet_b200.install_compat()
import numpy as np  # noqa: E402
import torch  # noqa: E402
from featurenet_b200 import synth  # noqa: E402
from featurenet_b200.optimizers import Adam  # noqa: E402
from products_tree import ProductSet  # noqa: E402
from model.keras_model import KerasFeatureModel  # noqa: E402


def shape_key(label, args):
    if label.startswith("conv2d") or label.startswith("dwconv2d"):
        d = args[0]
        return "%dx%dx%d->%d k%dx%d s%d out %dx%d" % (d.in_h, d.in_w, d.in_c, d.out_c, d.kh, d.kw, d.stride_h, d.out_h, d.out_w)
    if label.startswith("dense"):
        return "m%d k%d n%d" % (args[0], args[1], args[2])
    if label.startswith("pool2d"):
        d = args[0]
        return "%dx%dx%d k%dx%d s%d" % (d.in_h, d.in_w, d.channels, d.kh, d.kw, d.stride_h)
    if label.startswith("bn_"):
        return "outer%d a%d inner%d" % (args[0], args[1], args[2])
    return ""


def main():
    steps = 5
    with contextlib.redirect_stdout(io.StringIO()):
        formatted = ProductSet(os.path.join(ROOT, "featurenet_b200/data/nas_5_5_10_pop32_seed7.pdt")).format_products()
        (xtr, ytr), _ = synth.load("cifar", 2048, 16)
    agg = collections.defaultdict(lambda: [0, 0.0, 0, 0])
    per_candidate = []
    for idx, (product, original) in enumerate(formatted):
        with contextlib.redirect_stdout(io.StringIO()):
            net = KerasFeatureModel.parse_feature_model(copy.deepcopy(product)).build((32, 32, 3), 10)
        if not net:
            continue
        net.compile(loss="categorical_crossentropy", optimizer=Adam(1e-3))
        eng = net._need_engine()
        eng._ensure_capacity(128, 128)
        eng._set_data("train", xtr, ytr)
        eng._sync_lr()
        for _ in range(3):
            eng._step("train", 128)
        eng.stream.synchronize()
        eng.timeline = []
        with torch.cuda.stream(eng.stream):
            torch.cuda._sleep(int(5e7))
        for _ in range(steps):
            eng._step("train", 128)
        eng.stream.synchronize()
        ops = [op for op in eng.programs[("train", 128)] if op[1] is not None]
        total = 0.0
        for n, (label, e0, e1) in enumerate(eng.timeline):
            args = ops[n % len(ops)][2]
            ms = e0.elapsed_time(e1)
            f, b = eng.op_work(str(label), args)
            row = agg[(str(label), shape_key(str(label), args))]
            row[0] += 1
            row[1] += ms
            row[2] += f
            row[3] += b
            total += ms
        eng.timeline = None
        per_candidate.append((idx, net.count_params(), total / steps))
        net.release()
    whole = sum(r[1] for r in agg.values())
    print("population step time (sum over candidates, eager, serialised launches): %.2f ms" % (whole / steps))
    for idx, params, ms in sorted(per_candidate, key=lambda t: -t[2])[:10]:
        print("  candidate %2d  %9d params  %.3f ms/step" % (idx, params, ms))
    print("%-20s %-42s %6s %6s %9s %8s %7s" % ("operation", "shape", "share", "n", "avg us", "GB/s", "TF/s"))
    for (label, key), (n, ms, f, b) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
        print("%-20s %-42s %5.1f%% %6d %9.1f %8.0f %7.2f" % (label, key, 100 * ms / whole, n, 1e3 * ms / n, b / (ms / 1e3) / 1e9, f / (ms / 1e3) / 1e12))
    by_label = collections.defaultdict(float)
    for (label, _), row in agg.items():
        by_label[label] += row[1]
    print("by operation:", ", ".join("%s %.1f%%" % (k, 100 * v / whole) for k, v in sorted(by_label.items(), key=lambda kv: -kv[1])))


if __name__ == "__main__":
    main()
