"""Per-weight gradient error of the `wide` candidate at batch 128 against the float64 port, per math mode."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tests.conftest  # noqa: F401,E402
from tests.test_gpu_candidates import CANDIDATES, relerr  # noqa: E402
from tests.util import init_weights  # noqa: E402
from oracle.torch_cpu import TorchCpuNet  # noqa: E402
from featurenet_b200.optimizers import Adam  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "wide#9"
port_out = None
for mode in ("fp32",):
    os.environ["FNET_PRECISION"] = mode
    net = CANDIDATES[name]()
    rows = net.describe()
    rng = np.random.default_rng(77)
    weights = init_weights(net, rng, jitter=True)
    w32 = [[w.astype(np.float32).astype(np.float64) for w in ws] for ws in weights]
    net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
    eng = net._need_engine()
    eng.set_weights([w.astype(np.float32) for ws in weights for w in ws])
    x = rng.random((128, 32, 32, 3)).astype(np.float32)
    t = np.eye(10, dtype=np.float32)[rng.integers(0, 10, 128)]
    if port_out is None:
        port = TorchCpuNet(rows, w32, dtype=torch.float64)
        port.keep_activation_grads = True
        port.train_step(x, t, float(np.float32(1e-3)), {})
        port_out = ([port.out_nhwc(i) for i in range(len(rows))], port.grads_keras(), [port.out_grad_nhwc(i) for i in range(len(rows))])
        for i, row in enumerate(rows):
            print(i, row["class"], row["cfg"], row["inbound"], row["out_shape"])
    eng.train_on_batch(x, t)
    torch.cuda.synchronize()
    g = eng.grads.cpu().numpy()
    print("mode", mode)
    for i, row in enumerate(rows):
        got = eng.act[eng._resolve(i, True)][:128].cpu().numpy().reshape(port_out[0][i].shape)
        flips = ""
        if row["cfg"].get("activation") == "relu":
            flips = " relu sign flips %d of %d" % (int(((got > 0) != (port_out[0][i] > 0)).sum()), got.size)
        print("  fwd %2d %-22s %.2e%s" % (i, row["class"], relerr(got, port_out[0][i]), flips))
    for i, row in enumerate(rows):
        r = eng._resolve(i, True)
        if r in eng.gact and port_out[2][i] is not None and row["class"] != "Flatten":
            got = eng.gact[r][:128].cpu().numpy().reshape(port_out[2][i].shape)
            # the engine's buffer holds dL/d(pre-activation) for activated GEMM layers once act_bwd ran in place: compare both ways
            want = port_out[2][i]
            act = row["cfg"].get("activation")
            if act == "relu":
                want_pre = want * (port_out[0][i] > 0)
                print("  dact %2d %-22s vs dL/dy %.2e   vs dL/dz %.2e" % (i, row["class"], relerr(got, want), relerr(got, want_pre)))
            else:
                print("  dact %2d %-22s vs dL/dy %.2e" % (i, row["class"], relerr(got, want)))
    for (i, k), (off, size) in sorted(eng.w_off.items()):
        want = port_out[1][(i, k)]
        got = g[off: off + size].reshape(want.shape)
        print("  grad layer %2d %-20s w%d %-18s %.2e" % (i, rows[i]["class"], k, str(want.shape), relerr(got, want)))
