"""Where the fused step's time goes: per-operation GPU-timer stamps of the bench candidate (nas1 #4), averaged over
launches, plus the whole-launch time by CUDA events with and without a CUDA graph."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from featurenet_b200 import fused, synth  # noqa: E402
from featurenet_b200.optimizers import Adam  # noqa: E402

NAMES = {v: k for k, v in vars(fused).items() if k.startswith("FK_")}


def main():
    workload = sys.argv[1] if len(sys.argv) > 1 else "nas1"
    model, net, dataset = bench.build_candidate(workload)
    (x, y), _ = synth.load(dataset, 12800, 128)
    net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
    eng = net._need_engine()
    eng._ensure_capacity(128, 128)
    eng._set_data("train", x, y)
    eng._sync_lr()
    ops = eng._program("train", 128)
    print("program:", [str(op[0]) for op in ops], eng.fused_summary)
    if str(ops[0][0]) != "fused_step":
        return
    ops_ptr, n_ops, sync_ptr, sp = ops[0][2]
    ops_t = eng._fused_keep[("train", 128)][0]
    host = ops_t.cpu().numpy().tobytes()
    arr = (fused.FusedOp * n_ops).from_buffer_copy(host)
    stamps = torch.zeros(n_ops + 1, dtype=torch.int64, device="cuda")
    total = np.zeros(n_ops + 1)
    reps = 200
    for r in range(reps + 20):
        if r % 90 == 0:
            eng._zero_metrics(eng.state)
        eng.lib.fnet_fused_step_timed(ops_ptr, n_ops, sync_ptr, C.c_void_p(stamps.data_ptr()), sp)
        eng.stream.synchronize()
        if r >= 20:
            s = stamps.cpu().numpy().astype(np.float64)
            total += s - s[0]
    total /= reps
    prev = 0.0
    for i in range(n_ops):
        print("%2d %-16s barrier_after=%d  %7.2f us (ends at %7.2f)" % (i, NAMES.get(arr[i].kind, arr[i].kind), arr[i].barrier_after,
                                                                       (total[i + 1] - prev) / 1e3, total[i + 1] / 1e3))
        prev = total[i + 1]
    # whole launch: eager back-to-back and graph replay
    for mode in ("eager", "graph"):
        g = eng._graph("train", 128)
        eng._zero_metrics(eng.state)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        eng.stream.synchronize()
        e0.record(eng.stream)
        for _ in range(90):
            if mode == "eager":
                eng.lib.fnet_fused_step(ops_ptr, n_ops, sync_ptr, sp)
            else:
                eng.lib.fnet_graph_launch(g, sp)
        e1.record(eng.stream)
        eng.stream.synchronize()
        print(mode, "%.2f us per step" % (1e3 * e0.elapsed_time(e1) / 90))


if __name__ == "__main__":
    main()
