"""cProfile of one sequential pass over the benchmark population on a small set: where the HOST time of a candidate goes."""
import cProfile
import contextlib
import copy
import io
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import featurenet_b200  # noqa: E402

featurenet_b200.install_compat()
import torch  # noqa: E402
from featurenet_b200 import population as pop, synth  # noqa: E402
from products_tree import ProductSet  # noqa: E402
from tensorflow_generator import TensorflowGenerator  # noqa: E402

with contextlib.redirect_stdout(io.StringIO()):
    formatted = ProductSet(os.path.join(ROOT, "featurenet_b200/data/nas_5_5_10_pop32_seed7.pdt")).format_products()
    (xtr, ytr), (xte, yte) = synth.load("cifar", 2048, 512)
TensorflowGenerator.dataset = "cifar"
TensorflowGenerator.X_train, TensorflowGenerator.Y_train, TensorflowGenerator.X_test, TensorflowGenerator.Y_test = xtr, ytr, xte, yte
TensorflowGenerator.input_shape = (32, 32, 3)


def one_pass():
    for product, original in formatted:
        item = copy.deepcopy((product, original))
        with contextlib.redirect_stdout(io.StringIO()):
            pop.train_product(item[0], "cifar", epochs=1, batch_size=128, product_features=item[1], shared_gpu=True)


one_pass()
torch.cuda.synchronize()
prof = cProfile.Profile()
prof.enable()
one_pass()
torch.cuda.synchronize()
prof.disable()
out = io.StringIO()
pstats.Stats(prof, stream=out).sort_stats("cumulative").print_stats(45)
print(out.getvalue()[:9000])
