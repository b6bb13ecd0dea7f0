"""Where the wall time of one candidate goes (development aid): build, engine set-up, capture, steps, validation."""
import contextlib, io, os, sys, time
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
    formatted = ProductSet(os.path.join(ROOT, "featurenet_b200/data/nas_5_5_10_pop32_seed7.pdt")).format_products()
    (xtr, ytr), (xte, yte) = synth.load("cifar", 10000, 2000)
torch.zeros(1, device="cuda"); torch.cuda.synchronize()
from featurenet_b200.engine import DeviceDataCache
DeviceDataCache.get(xtr, torch.device("cuda:0")); DeviceDataCache.get(ytr, torch.device("cuda:0"))
DeviceDataCache.get(xte, torch.device("cuda:0")); DeviceDataCache.get(yte, torch.device("cuda:0"))
for rep in range(2):
    import copy
    product = copy.deepcopy(formatted[idx][0])
    marks = [("start", time.perf_counter())]
    def mark(name):
        torch.cuda.synchronize(); marks.append((name, time.perf_counter()))
    with contextlib.redirect_stdout(io.StringIO()):
        net = KerasFeatureModel.parse_feature_model(product).build((32, 32, 3), 10)
    mark("parse+build graph")
    net.compile(loss="categorical_crossentropy", optimizer=Adam(1e-3))
    eng = net._need_engine(); mark("engine init (arenas, weights)")
    eng._ensure_capacity(1000, 128); mark("activation buffers")
    eng._set_data("train", xtr, ytr); eng._sync_lr(); mark("set data")
    eng._zero_metrics(eng.state); eng._step("train", 128); mark("first step (program + capture + run)")
    for _ in range(77): eng._step("train", 128)
    mark("77 steps")
    eng._step("train", 16); mark("remainder step (second capture)")
    eng.evaluate(xte, yte); mark("validation (2 captures + 2 steps)")
    eng.evaluate(xte, yte); mark("evaluate again")
    net.release(); mark("release")
    print("rep", rep, " | ".join("%s %.1f ms" % (n, 1e3 * (t - marks[i][1])) for i, (n, t) in enumerate(marks[1:])))
