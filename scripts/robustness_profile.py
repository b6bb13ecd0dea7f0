"""Wall time of the robustness fitness (tensorflow_generator.py:175-218) on a trained lenet5: fgsm, pgd, cw, clever."""
import contextlib, io, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("FNET_SYNTH_TRAIN", "20000")
os.environ.setdefault("FNET_SYNTH_TEST", "2000")
import featurenet_b200
featurenet_b200.install_compat()
import numpy as np, torch
from tensorflow_generator import TensorflowGenerator

n = int(sys.argv[1]) if len(sys.argv) > 1 else 500
with contextlib.redirect_stdout(io.StringIO()):
    gen = TensorflowGenerator("lenet5", epochs=2, dataset="mnist", data_augmentation=False, batch_size=128)
model = gen.model
print("lenet5 trained: accuracy %.3f in %.2f s" % (gen.accuracy, gen.training_time))
model.accuracy = max(model.accuracy, 0.9)
for metric, count in (("fgsm", n), ("pgd", n), ("cw", n), ("clever", min(n, 50))):
    np.random.seed(1)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        TensorflowGenerator.eval_robustness(model, [metric], robustness_set_size=count)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print("%-6s on %3d images: %.2f s  score %s" % (metric, count, dt, getattr(model, metric + "_score")))
