"""Run under torchrun with 2+ GPUs: data-parallel training of one candidate == single-GPU training on the same
global batches (a network without BatchNorm, whose statistics are per-rank by design).  Prints DP_CHECK_OK."""
import contextlib
import io
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import featurenet_b200  # noqa: E402

featurenet_b200.install_compat()
from featurenet_b200.dataparallel import DataParallelTrainer  # noqa: E402
from featurenet_b200.optimizers import Adam  # noqa: E402
from tests.util import build_template, init_weights  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    rng = np.random.default_rng(5)
    x = rng.random((200, 28, 28, 1)).astype(np.float32)
    y = np.eye(10, dtype=np.float32)[rng.integers(0, 10, 200)]
    order = rng.permutation(200)
    weights = None

    def make():
        with contextlib.redirect_stdout(io.StringIO()):
            net = build_template("lenet5", (28, 28, 1))
        net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
        return net

    net_dp = make()
    weights = [w.astype(np.float32) for ws in init_weights(net_dp, np.random.default_rng(9)) for w in ws]
    net_dp._need_engine().set_weights(weights)
    trainer = DataParallelTrainer(net_dp, dist, rank, world)
    loss_dp, acc_dp = trainer.fit_epoch(x, y, 48, order)          # 200 = 4 x 48 + 8: ragged last batch, 8 rows over the ranks
    w_dp = net_dp._need_engine().get_weights()

    net_1 = make()
    net_1._need_engine().set_weights(weights)
    hist = net_1.fit(x, y, batch_size=48, epochs=1, permutations=[order])
    w_1 = net_1._need_engine().get_weights()

    worst = max(float(np.abs(a - b).max() / (1e-4 * np.abs(b).max() + 0.05 * 1e-3)) for a, b in zip(w_dp, w_1))
    same_across_ranks = torch.tensor([float(np.abs(w_dp[-2]).sum())], device="cuda", dtype=torch.float64)
    gathered = [torch.zeros_like(same_across_ranks) for _ in range(world)]
    dist.all_gather(gathered, same_across_ranks)
    ok = worst <= 1.0 and abs(loss_dp - hist.history["loss"][0]) <= 1e-4 * abs(loss_dp) \
        and abs(acc_dp - hist.history["acc"][0]) <= 1.0 / 200 + 1e-9 \
        and all(float(g.item()) == float(gathered[0].item()) for g in gathered)
    if rank == 0:
        print("dp loss %.6f single %.6f; acc %.4f / %.4f; weights within %.3f of the bound; replicas identical: %s"
              % (loss_dp, hist.history["loss"][0], acc_dp, hist.history["acc"][0], worst,
                 all(float(g.item()) == float(gathered[0].item()) for g in gathered)))
        print("DP_CHECK_OK" if ok else "DP_CHECK_FAILED")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
