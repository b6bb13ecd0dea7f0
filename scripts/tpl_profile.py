"""Per-kernel CUDA-event profile of a template ("lenet5" / "keras") at batch 128 (development aid)."""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import featurenet_b200
featurenet_b200.install_compat()
import numpy as np, torch
from featurenet_b200 import synth
from featurenet_b200.optimizers import Adam
from model.keras_model import KerasFeatureModel

tpl, ds = sys.argv[1], sys.argv[2]
with contextlib.redirect_stdout(io.StringIO()):
    if tpl in ("nas1", "wide"):                     # bench.py workloads that are products of a .pdt, not templates
        import bench
        _, net, ds = bench.build_candidate(tpl)
    else:
        net = KerasFeatureModel.parse_feature_model(tpl).build(synth.SHAPES[ds], 10)
    (xtr, ytr), _ = synth.load(ds, 16384, 16)
net.compile(loss="categorical_crossentropy", optimizer=Adam(1e-3))
eng = net._need_engine()
eng._ensure_capacity(128, 128)
eng._set_data("train", xtr, ytr)
eng._sync_lr()
table = eng.profile_steps("train", 128, steps=10)
ops = [op for op in eng.programs[("train", 128)] if op[1] is not None]
# per-launch detail: label + shape hint
per = {}
for n, (label, e0, e1) in enumerate([]):
    pass
tot = sum(v["ms_total"] for v in table.values())
for k, v in sorted(table.items(), key=lambda kv: -kv[1]["ms_total"]):
    print("%-22s n=%3d avg %8.1f us  share %5.1f%%  GB/s %7.1f  TF/s %6.2f" % (k, v["launches"], 1e3 * v["ms_total"] / v["launches"], 100 * v["ms_total"] / tot,
          v["bytes"] / (v["ms_total"] / 1e3) / 1e9, v["flops"] / (v["ms_total"] / 1e3) / 1e12))
print("eager total per step %.3f ms" % (tot / 10))
# individual launches of the GEMM-shaped ops, in program order
eng.timeline = []
with torch.cuda.stream(eng.stream):
    torch.cuda._sleep(int(5e7))
eng._step("train", 128); eng.stream.synchronize()
for (label, e0, e1), (_, _, args) in zip(eng.timeline, ops):
    if label.startswith(("conv2d", "dense", "dwconv")):
        d = args[0]
        shape = ("B%d %dx%dx%d -> %dx%dx%d k%dx%d s%d" % (d.batch, d.in_h, d.in_w, d.in_c, d.out_h, d.out_w, d.out_c, d.kh, d.kw, d.stride_h)
                 if hasattr(d, "in_c") else "m%d k%d n%d" % (args[0], args[1], args[2]))
        f, b = eng.op_work(label, args)
        ms = e0.elapsed_time(e1)
        print("  %-20s %-40s %8.1f us  %6.2f TF/s %7.1f GB/s" % (label, shape, 1e3 * ms, f / ms / 1e9, b / ms / 1e6))
eng.timeline = None
# the captured step (graph launches back to back, filter gradients overlapped on the side stream)
for _ in range(5):
    eng._step("train", 128)
g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
g0.record(eng.stream)
for _ in range(50):
    eng._step("train", 128)
g1.record(eng.stream)
eng.stream.synchronize()
print("graph step %.3f ms (overlap %s)" % (g0.elapsed_time(g1) / 50, eng.overlap_wgrad))
