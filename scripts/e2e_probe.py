"""Where the end-to-end (host-fed) step of the bench candidate spends its time (development aid)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import bench  # noqa: E402
from featurenet_b200 import synth  # noqa: E402
from featurenet_b200.optimizers import Adam  # noqa: E402

model, net, dataset = bench.build_candidate("nas1")
with bench.quiet():
    (xtr, ytr), _ = synth.load(dataset, 4096, 16)
net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
eng = net._need_engine()
B = 128
xs = torch.from_numpy(xtr[:B * 8]).pin_memory()
ys = torch.from_numpy(ytr[:B * 8]).pin_memory()
for i in range(5):
    eng.train_on_batch(xs[:B], ys[:B])
torch.cuda.synchronize()


def wall(fn, n=300):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(n):
        fn(i)
    torch.cuda.synchronize()
    return 1e6 * (time.perf_counter() - t0) / n


graph = eng.graphs[("train_fed", B)]
dst = eng.act[0]
print("train_on_batch            %7.1f us" % wall(lambda i: eng.train_on_batch(xs[(i % 8) * B:(i % 8 + 1) * B], ys[(i % 8) * B:(i % 8 + 1) * B])))
print("graph launch + sync       %7.1f us" % wall(lambda i: (eng.lib.fnet_graph_launch(graph, eng.sp), eng.stream.synchronize())))
print("graph launch, no sync     %7.1f us" % wall(lambda i: eng.lib.fnet_graph_launch(graph, eng.sp)))
with torch.cuda.stream(eng.stream):
    print("H2D x (1.57 MB) + sync    %7.1f us" % wall(lambda i: (dst[:B].copy_(xs[:B], non_blocking=True), eng.stream.synchronize())))
    print("H2D x, no sync (pipelined)%7.1f us" % wall(lambda i: dst[:B].copy_(xs[:B], non_blocking=True)))
    big = torch.empty(64 << 20, dtype=torch.uint8).pin_memory()
    dbig = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")
    us = wall(lambda i: dbig.copy_(big, non_blocking=True), n=20)
    print("H2D 64 MB                 %7.1f us  = %.1f GB/s" % (us, (64 << 20) / us / 1e3))
print("step_fed C call           %7.1f us" % wall(lambda i: eng.lib.fnet_graph_step_fed(
    graph, eng.act[0].data_ptr(), xs.data_ptr(), B * 3072 * 4, eng.yb.data_ptr(), ys.data_ptr(), B * 10 * 4,
    eng.metrics.data_ptr(), eng._metrics_host.data_ptr(), 16, eng.sp)))
print("step_fed, no x copy       %7.1f us" % wall(lambda i: eng.lib.fnet_graph_step_fed(
    graph, None, None, 0, eng.yb.data_ptr(), ys.data_ptr(), B * 10 * 4,
    eng.metrics.data_ptr(), eng._metrics_host.data_ptr(), 16, eng.sp)))
