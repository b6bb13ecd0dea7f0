"""Per-candidate timing of the benchmark population (development aid): where does a population's time go?"""
import contextlib, io, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import featurenet_b200
featurenet_b200.install_compat()
import numpy as np, torch
from featurenet_b200 import synth
from products_tree import ProductSet
from tensorflow_generator import TensorflowGenerator

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
with contextlib.redirect_stdout(io.StringIO()):
    formatted = ProductSet(os.path.join(ROOT, "featurenet_b200/data/nas_5_5_10_pop32_seed7.pdt")).format_products()[:n]
    (xtr, ytr), (xte, yte) = synth.load("cifar", 10000, 2000)
TensorflowGenerator.dataset = "cifar"
TensorflowGenerator.X_train, TensorflowGenerator.Y_train, TensorflowGenerator.X_test, TensorflowGenerator.Y_test = xtr, ytr, xte, yte
TensorflowGenerator.input_shape = (32, 32, 3)
for i, (product, original) in enumerate(formatted):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        gen = TensorflowGenerator(product, epochs=1, dataset="cifar", data_augmentation=False, batch_size=128, product_features=original)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    net = gen.model.model
    print("%2d params %9d layers %3d  total %.3fs train %.3fs  acc %.3f  mflop/img %.1f" % (
        i, gen.params, len(net.layers) if net else 0, dt, gen.training_time, gen.accuracy,
        3 * net.train_flops_per_image() / 1e6 if net else 0))
