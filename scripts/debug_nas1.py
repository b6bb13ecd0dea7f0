import os, sys, contextlib, io
sys.path.insert(0, '/root/repo')
import numpy as np, torch
import tests.conftest
from tests.test_gpu_candidates import CANDIDATES
from tests.util import init_weights
from featurenet_b200.optimizers import Adam
for jitter in (False, True):
  for nograph in ("0", "1"):
    os.environ["FNET_NO_GRAPH"] = nograph
    net = CANDIDATES["nas1#4"]()
    rng = np.random.default_rng(77)
    weights = init_weights(net, rng, jitter=jitter)
    net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
    eng = net._need_engine()
    eng.set_weights([w.astype(np.float32) for ws in weights for w in ws])
    x = rng.random((128, 32, 32, 3)).astype(np.float32)
    t = np.eye(10, dtype=np.float32)[rng.integers(0, 10, 128)]
    print("jitter", jitter, "nograph", nograph, eng.train_on_batch(x, t))
    torch.cuda.synchronize()
    for i, n in enumerate(eng.nodes):
        print(i, n.kind, n.cfg.get("activation"), eng.src[i], eng.shape[i],
              "act", float(eng.act[eng._resolve(i, True)][:128].abs().max()) ,
              "gact", float(eng.gact[eng._resolve(i, True)][:128].abs().max()) if eng._resolve(i, True) in eng.gact else None)
    g = eng.grads.cpu().numpy()
    for (i, k), (off, size) in eng.w_off.items():
        print("grad", i, k, off, size, np.abs(g[off:off+size]).max())
    print([op[0] for op in eng.programs[("train_fed", 128)]])
