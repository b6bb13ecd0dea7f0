"""Per-layer gradient error of one training step against the float64 oracle (development aid).
  FNET_PRECISION=auto python scripts/grad_debug.py keras@cifar 4"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import featurenet_b200
featurenet_b200.install_compat()
import numpy as np, torch
from oracle.net import OracleNet
from tests.util import init_weights
from tests.test_gpu_network import make_net, engine_for, oracle_masks, grads_from_engine, relerr

name, batch = sys.argv[1], int(sys.argv[2])
net = make_net(name)
rows = net.describe()
rng = np.random.default_rng(21)
weights = init_weights(net, rng, jitter=True)
weights32 = [[w.astype(np.float32).astype(np.float64) for w in ws] for ws in weights]
classes = net.output_shape[-1]
eng = engine_for(net, weights)
oracle = OracleNet(rows, weights32)
x = rng.random((batch,) + tuple(net.input_shape[1:])).astype(np.float32)
t = np.eye(classes, dtype=np.float32)[rng.integers(0, classes, batch)]
masks = oracle_masks(eng, rows, 0, batch)
oracle.train_step(x.astype(np.float64), t.astype(np.float64), 1e-3, masks)
eng.train_on_batch(x, t)
torch.cuda.synchronize()
for i, row in enumerate(rows):
    got = eng.act[eng._resolve(i, True)][:batch].cpu().numpy().reshape(oracle.outs[i].shape)
    print("fwd  %2d %-22s %.2e" % (i, row["class"], relerr(got, oracle.outs[i])))
for i in range(len(rows) - 1, 0, -1):
    r = eng._resolve(i, True)
    if r in eng.gact and oracle.douts[i] is not None and r == i:
        got = eng.gact[r][:batch].cpu().numpy().reshape(oracle.douts[i].shape)
        print("dout %2d %-22s %.2e" % (i, rows[i]["class"], relerr(got, oracle.douts[i])))
for (i, k), g in grads_from_engine(eng).items():
    print("grad %2d %-22s w%d %.2e" % (i, rows[i]["class"], k, relerr(g, oracle.last_grads[i][k])))
