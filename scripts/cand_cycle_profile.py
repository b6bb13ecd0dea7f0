"""cProfile of one whole candidate cycle (build + 1 epoch + validation + evaluate) after a warm-up cycle (development aid)."""
import contextlib, cProfile, io, os, pstats, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import featurenet_b200
featurenet_b200.install_compat()
import torch
from featurenet_b200 import synth
from products_tree import ProductSet
from tensorflow_generator import TensorflowGenerator

idx = int(sys.argv[1]) if len(sys.argv) > 1 else 1
with contextlib.redirect_stdout(io.StringIO()):
    formatted = ProductSet(os.path.join(ROOT, "featurenet_b200/data/nas_5_5_10_pop32_seed7.pdt")).format_products()
    (xtr, ytr), (xte, yte) = synth.load("cifar", 10000, 2000)
TensorflowGenerator.dataset = "cifar"
TensorflowGenerator.X_train, TensorflowGenerator.Y_train, TensorflowGenerator.X_test, TensorflowGenerator.Y_test = xtr, ytr, xte, yte
TensorflowGenerator.input_shape = (32, 32, 3)
import copy
def cycle():
    product, original = copy.deepcopy(formatted[idx])
    with contextlib.redirect_stdout(io.StringIO()):
        return TensorflowGenerator(product, epochs=1, dataset="cifar", data_augmentation=False, batch_size=128, product_features=original)
for _ in range(2):
    cycle()
torch.cuda.synchronize()
t0 = time.perf_counter(); g = cycle(); torch.cuda.synchronize(); print("cycle %.2f ms, params %d, training_time %.2f ms" % (1e3 * (time.perf_counter() - t0), g.params, 1e3 * g.training_time))
pr = cProfile.Profile(); pr.enable(); cycle(); torch.cuda.synchronize(); pr.disable()
out = io.StringIO(); pstats.Stats(pr, stream=out).sort_stats("cumulative").print_stats(28); print(out.getvalue()[:6000])
