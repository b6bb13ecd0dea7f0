"""Run under torchrun with 2+ GPUs: data-parallel training of one candidate through the ORDINARY entry points
(helpers.train_model -> Network.fit / evaluate with FNET_DP=1) == single-GPU training on the same global batches
(a network without BatchNorm, whose batch statistics are per-rank by design).  Two epochs with validation after each
and an evaluate() in between the two fits, so that buffers captured graphs point at are re-planned under the trainer's
feet (graphs vs. reallocation).  Prints DP_CHECK_OK."""
import contextlib
import io
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("FNET_SEED", "20261106")
import featurenet_b200  # noqa: E402

featurenet_b200.install_compat()
from featurenet_b200.optimizers import Adam  # noqa: E402
from tests.util import build_template, init_weights  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    rng = np.random.default_rng(5)
    x = rng.random((200, 28, 28, 1)).astype(np.float32)
    y = np.eye(10, dtype=np.float32)[rng.integers(0, 10, 200)]
    xv = rng.random((1101, 28, 28, 1)).astype(np.float32)         # larger than the training batch: evaluate re-plans buffers
    yv = np.eye(10, dtype=np.float32)[rng.integers(0, 10, 1101)]
    orders = [rng.permutation(200) for _ in range(3)]

    def make(dp):
        os.environ["FNET_DP"] = "1" if dp else "0"
        with contextlib.redirect_stdout(io.StringIO()):
            net = build_template("lenet5", (28, 28, 1))
        net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
        return net

    net_dp = make(True)
    weights = [w.astype(np.float32) for ws in init_weights(net_dp, np.random.default_rng(9)) for w in ws]
    eng = net_dp._need_engine()
    assert eng.dp is not None and eng.dp.world == world
    eng.set_weights(weights)
    h1 = net_dp.fit(x, y, batch_size=48, epochs=1, permutations=orders, validation_data=(xv[:300], yv[:300]))   # 200 = 4 x 48 + 8
    mid_dp = net_dp.evaluate(xv, yv)                              # 1101 rows: sharded, ragged, grows the buffers
    h2 = net_dp.fit(x, y, batch_size=48, epochs=3, initial_epoch=1, permutations=orders, validation_data=(xv[:300], yv[:300]))
    w_dp = eng.get_weights()
    end_dp = net_dp.evaluate(xv, yv)

    net_1 = make(False)
    eng1 = net_1._need_engine()
    assert eng1.dp is None
    eng1.set_weights(weights)
    g1 = net_1.fit(x, y, batch_size=48, epochs=1, permutations=orders, validation_data=(xv[:300], yv[:300]))
    mid_1 = net_1.evaluate(xv, yv)
    g2 = net_1.fit(x, y, batch_size=48, epochs=3, initial_epoch=1, permutations=orders, validation_data=(xv[:300], yv[:300]))
    w_1 = eng1.get_weights()
    end_1 = net_1.evaluate(xv, yv)

    worst = max(float(np.abs(a - b).max() / (2e-4 * np.abs(b).max() + 0.05 * 1e-3)) for a, b in zip(w_dp, w_1))
    close = lambda a, b, tol=2e-4: abs(a - b) <= tol * max(1.0, abs(b))
    hist_ok = all(close(a, b) for key in ("loss", "val_loss") for a, b in
                  zip(h1.history[key] + h2.history[key], g1.history[key] + g2.history[key]))
    acc_ok = all(abs(a - b) <= 1.0 / 200 + 1e-9 for key in ("acc", "val_acc") for a, b in
                 zip(h1.history[key] + h2.history[key], g1.history[key] + g2.history[key]))
    eval_ok = close(mid_dp[0], mid_1[0]) and close(end_dp[0], end_1[0]) and abs(mid_dp[1] - mid_1[1]) <= 1.5 / 1101 \
        and abs(end_dp[1] - end_1[1]) <= 1.5 / 1101
    digest = torch.tensor([float(np.abs(w).sum()) for w in w_dp], device="cuda", dtype=torch.float64)
    gathered = [torch.zeros_like(digest) for _ in range(world)]
    dist.all_gather(gathered, digest)
    same = all(bool(torch.equal(g, gathered[0])) for g in gathered)
    ok = worst <= 1.0 and hist_ok and acc_ok and eval_ok and same
    if rank == 0:
        print("world %d, single graph %s: losses dp %s / single %s; evaluate dp %s / single %s; weights within %.3f of the "
              "bound; replicas identical: %s; all-reduce %d bytes x %d calls"
              % (world, eng.dp_single_graph, [round(v, 5) for v in h1.history["loss"] + h2.history["loss"]],
                 [round(v, 5) for v in g1.history["loss"] + g2.history["loss"]], [round(v, 5) for v in end_dp],
                 [round(v, 5) for v in end_1], worst, same, eng.dp.allreduce_bytes, eng.dp.allreduce_calls))
        print("DP_CHECK_OK" if ok else "DP_CHECK_FAILED hist %s acc %s eval %s" % (hist_ok, acc_ok, eval_ok))
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
