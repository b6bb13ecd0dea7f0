"""Per-kernel CUDA-event profile of one candidate of the benchmark population (development aid)."""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import featurenet_b200
featurenet_b200.install_compat()
import numpy as np, torch
from featurenet_b200 import synth
from featurenet_b200.optimizers import Adam
from products_tree import ProductSet
from model.keras_model import KerasFeatureModel

idx = int(sys.argv[1])
with contextlib.redirect_stdout(io.StringIO()):
    product, original = ProductSet(os.path.join(ROOT, "featurenet_b200/data/nas_5_5_10_pop32_seed7.pdt")).format_products()[idx]
    net = KerasFeatureModel.parse_feature_model(product).build((32, 32, 3), 10)
    (xtr, ytr), _ = synth.load("cifar", 2048, 16)
for r in net.describe():
    print("  ", r["class"], r["in_shapes"], "->", r["out_shape"], {k: v for k, v in r["cfg"].items() if k in ("kernel", "strides", "activation", "filters", "units", "rate")}, r["params"])
net.compile(loss="categorical_crossentropy", optimizer=Adam(1e-3))
eng = net._need_engine()
eng._ensure_capacity(128, 128)
eng._set_data("train", xtr, ytr)
eng._sync_lr()
table = eng.profile_steps("train", 128, steps=10)
tot = sum(v["ms_total"] for v in table.values())
for k, v in sorted(table.items(), key=lambda kv: -kv[1]["ms_total"]):
    print("%-22s n=%3d avg %8.1f us  share %5.1f%%  GB/s %7.1f  TF/s %6.2f" % (k, v["launches"], 1e3 * v["ms_total"] / v["launches"], 100 * v["ms_total"] / tot,
          v["bytes"] / (v["ms_total"] / 1e3) / 1e9, v["flops"] / (v["ms_total"] / 1e3) / 1e12))
print("eager total per step %.3f ms" % (tot / 10))
