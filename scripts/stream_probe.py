"""Where the time of a streamed (fit_generator) step goes: host Python, the upload, the kernels (development aid)."""
import contextlib, cProfile, io, os, pstats, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import featurenet_b200
featurenet_b200.install_compat()
import numpy as np, torch
import bench
from featurenet_b200 import synth
from featurenet_b200.optimizers import Adam

B = 128
with contextlib.redirect_stdout(io.StringIO()):
    _, net, ds = bench.build_candidate("nas1")
    (xtr, ytr), _ = synth.load(ds, 4096, 16)
net.compile(loss="categorical_crossentropy", optimizer=Adam(1e-3))
eng = net._need_engine()
xs, ys = torch.from_numpy(xtr[: B * 8]).pin_memory(), torch.from_numpy(ytr[: B * 8]).pin_memory()
pairs = [(xs[i * B: (i + 1) * B], ys[i * B: (i + 1) * B]) for i in range(8)]


def feed():
    i = 0
    while True:
        yield pairs[i & 7]
        i += 1


g = feed()
net.fit_generator(g, steps_per_epoch=20, epochs=1)
K = 2000
torch.cuda.synchronize()
t0 = time.perf_counter()
net.fit_generator(g, steps_per_epoch=K, epochs=1)
torch.cuda.synchronize()
print("fit_generator      %7.1f us/step" % (1e6 * (time.perf_counter() - t0) / K))

# upload alone
dst = torch.empty_like(pairs[0][0], device="cuda")
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    for _ in range(20):
        dst.copy_(pairs[0][0], non_blocking=True)
    s.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s)
    for i in range(200):
        dst.copy_(pairs[i & 7][0], non_blocking=True)
    e1.record(s)
    s.synchronize()
print("upload alone       %7.1f us/batch (%.1f GB/s)" % (1e3 * e0.elapsed_time(e1) / 200, dst.numel() * 4 / (e0.elapsed_time(e1) / 200) / 1e6))

# the captured step alone
graph = eng._graph("train_ring0", B)
for _ in range(20):
    eng.lib.fnet_graph_launch(graph, eng.sp)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(eng.stream)
for _ in range(500):
    eng.lib.fnet_graph_launch(graph, eng.sp)
e1.record(eng.stream)
eng.stream.synchronize()
print("captured step      %7.1f us" % (1e3 * e0.elapsed_time(e1) / 500))

# host side only: the same calls with the GPU far behind is not separable, so profile the Python
pr = cProfile.Profile()
pr.enable()
net.fit_generator(g, steps_per_epoch=K, epochs=1)
pr.disable()
out = io.StringIO()
pstats.Stats(pr, stream=out).sort_stats("cumulative").print_stats(14)
print("\n".join(out.getvalue().splitlines()[:40]))
