#!/usr/bin/env python
"""Benchmark of the per-candidate train+evaluate path (contract: see the task statement / DESIGN.md).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload nas1|lenet5|wide]

A "step" is one optimisation step (forward, loss, backward, Adam) over one batch of 128
synthetic images of the candidate named by BASELINE.json configs[1]: product #4 of the
seed-0 sample of main_1block_nas.xml (5 cells, 1 block) on CIFAR-shaped 32x32x3 data.
`value` is device-timed whole-job train images/s with the dataset resident in HBM; `e2e` is
the same step driven through the public API (`Network.fit_generator`, and `train_on_batch` beside it) from pinned HOST
buffers, H2D and D2H inside the timed region.  With N > 1 every rank trains its own copy of
the candidate (population sharding: no data-path collective), weak scaling.
`--impl reference` times the CPU port of the same step (oracle/torch_cpu.py, all host
threads): the reference's own TF 1.12 / Keras 2.2.4 cannot be installed on this box.
"""
import argparse
import contextlib
import copy
import io
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import featurenet_b200  # noqa: E402

featurenet_b200.install_compat()

BATCH = 128
METRIC = "per-candidate train images/sec (candidates trained+evaluated/hour reported beside it)"
WORKLOADS = {
    # name: (product file or template, product index, dataset, description)
    "nas1": ("featurenet_b200/data/main_1block_nas_5cells_seed0.pdt", 4, "cifar",
             "main_1block_nas.xml (5 cells, 1 block) seed-0 sample, product #4, CIFAR-shape 32x32x3, batch 128"),
    "lenet5": ("lenet5", None, "mnist", "lenet5 template, MNIST-shape 28x28x1, batch 128"),
    "wide": ("tests/golden/nas_5_5_10_seed0.pdt", 9, "cifar",
             "nas_5_5_10.xml seed-0 sample, product #9 (98.6 MFLOP/img fwd), CIFAR-shape, batch 128"),
    "keras": ("keras", None, "cifar", "keras template (5.34 M parameters, 634 MFLOP/img train), CIFAR-shape, batch 128"),
}


def quiet():
    return contextlib.redirect_stdout(io.StringIO())


# The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner), so
# file descriptor 1 is pointed at stderr for the whole run and the JSON line goes to a private copy of the real one.
_JSON_OUT = None


def claim_stdout():
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def config_for(workload, gpus):
    """The `config` object of the JSON line: identical, key for key, in the product arm and the reference arm."""
    return {"workload": WORKLOADS[workload][3], "batch": BATCH, "optimizer": "adam(1e-3)",
            "parallelism": "candidate-sharded x%d (no collective)" % gpus,
            "l2": "inputs larger than L2: the resident dataset (614 MB CIFAR-shape / 188 MB MNIST-shape) is streamed "
                  "batch by batch through the epoch permutation"}


def build_candidate(workload):
    from model.keras_model import KerasFeatureModel
    from products_tree import ProductSet
    from featurenet_b200 import synth
    source, index, dataset, _ = WORKLOADS[workload]
    with quiet():
        if index is None:
            model = KerasFeatureModel.parse_feature_model(source)
        else:
            product, original = ProductSet(os.path.join(ROOT, source)).format_products()[index]
            model = KerasFeatureModel.parse_feature_model(product, product_features=original)
        net = model.build(synth.SHAPES[dataset], synth.CLASSES[dataset])
    if not net:
        raise SystemExit("benchmark candidate did not build")
    return model, net, dataset


def peaks(math="auto"):
    """Roofline denominators: HBM copy rate and the dense rate of the arithmetic the math mode runs on.

    MEASURED_PEAKS.json (driver-written) has HBM and bf16; profiles/r2_peaks_tf32_ffma.json (scripts/measure_peaks.py,
    same method: 8192^3 matmul, best of 10) adds cuBLAS TF32 and fp32-FFMA.  fp32 mode -> FFMA peak; tf32 -> TF32 peak;
    auto / tf32x3 -> TF32 peak / 3 on the tensor-core kernels (three MMAs per algorithmic product); bf16 -> bf16 peak."""
    out = {"hbm_gbs": 6650.0, "tflops": 1590.0, "source": "fallback (B200_PROFILING.md)"}
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as handle:
            p = json.load(handle)
        out = {"hbm_gbs": p["hbm_gbs"], "tflops": p["bf16_tflops"], "source": "measured (MEASURED_PEAKS.json, burst)"}
    extra = os.path.join(ROOT, "profiles", "r2_peaks_tf32_ffma.json")
    if math != "bf16" and os.path.isfile(extra):
        with open(extra) as handle:
            q = json.load(handle)
        if math == "fp32":
            out["tflops"], what = q["fp32_ffma_tflops"], "fp32 FFMA"
        elif math == "tf32":
            out["tflops"], what = q["tf32_tflops"], "TF32"
        else:
            out["tflops"], what = q["tf32_tflops"] / 3.0, "TF32 / 3 (three MMAs per product)"
        out["source"] += "; tensor peak = %s from profiles/r2_peaks_tf32_ffma.json" % what
    return out


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.FIELDS,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def wait_for_samples(self, count=2, limit_s=3.0):
        """Block until nvidia-smi is up and sampling: its start-up (NVML init) must not fall into the timed region."""
        t0 = time.perf_counter()
        while self.proc is not None and len(self.rows) < count and time.perf_counter() - t0 < limit_s:
            time.sleep(0.05)

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) == 7:
                self.rows.append(parts)

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        self.proc.wait()          # our own child, by handle -- never by pattern
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = sorted({names[i] for r in self.rows for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def cpu_port_rate(net, dataset, seconds, steps=None, threads=None, warm_steps=2, warm_seconds=0.0):
    """Train images/s of the CPU port (oracle/torch_cpu.py) on batches of the same workload.

    The port is warmed (thread pool, oneDNN primitive cache) for at least `warm_steps` steps and
    `warm_seconds` seconds before the timed part, so that a short run reports its steady state."""
    import torch
    from oracle.torch_cpu import TorchCpuNet
    from featurenet_b200 import synth
    from tests.util import init_weights
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    try:
        # keep freed activation buffers in the heap: a fresh process otherwise maps and faults in every temporary of
        # every step (2x slower than the same port inside a long-running process, which hid this from the first runs)
        import ctypes
        libc = ctypes.CDLL("libc.so.6")
        libc.mallopt(-1, 2 ** 31 - 1)       # M_TRIM_THRESHOLD
        libc.mallopt(-3, 2 ** 30)           # M_MMAP_THRESHOLD
    except OSError:
        pass
    rng = np.random.default_rng(1236)
    port = TorchCpuNet(net.describe(), init_weights(net, rng), threads=threads)
    shape, classes = synth.SHAPES[dataset], synth.CLASSES[dataset]
    x = rng.random((BATCH,) + shape).astype(np.float32)
    t = np.eye(classes, dtype=np.float32)[rng.integers(0, classes, BATCH)]
    warmed, w0 = 0, time.perf_counter()
    while warmed < warm_steps or time.perf_counter() - w0 < warm_seconds:
        port.train_step(x, t)
        warmed += 1
    done, t0 = 0, time.perf_counter()
    while True:
        port.train_step(x, t)
        done += 1
        elapsed = time.perf_counter() - t0
        if (steps is not None and done >= steps) or (steps is None and elapsed >= seconds):
            break
    return done * BATCH / elapsed, done, elapsed, threads


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    _, net, dataset = build_candidate(args.workload)
    port_steps = args.steps
    t0 = time.perf_counter()
    rate, done, elapsed, threads = cpu_port_rate(net, dataset, seconds=0, steps=port_steps, warm_steps=max(args.warmup, 3),
                                                 warm_seconds=3.0)
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": "images/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * elapsed / done, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_for(args.workload, args.gpus),
        "cpu_baseline": {"value": rate, "unit": "images/s", "cores": threads, "kind": "port",
                         "sample": "%d train steps of batch %d through oracle/torch_cpu.py (TF1.12/Keras2.2.4 "
                                   "is not installable here)" % (done, BATCH)},
        "e2e": {"value": rate, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": time.perf_counter() - t0,
    }
    emit(line)


MIN_TIMED_SECONDS = 0.5


def measure_workload(workload, args, torch, dist, world, device_index, full=True):
    """Device-timed steps (+ optionally e2e, per-kernel profile, whole-candidate cycle) for one workload."""
    from featurenet_b200 import synth
    from featurenet_b200.optimizers import Adam
    model, net, dataset = build_candidate(workload)
    with quiet():
        (x_train, y_train), (x_test, y_test) = synth.load(dataset)
    n = x_train.shape[0]
    net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
    eng = net._need_engine()
    eng._ensure_capacity(BATCH, BATCH)
    eng._set_data("train", x_train, y_train)
    perm = torch.from_numpy(np.random.default_rng(1237).permutation(n).astype(np.int32))
    with torch.cuda.stream(eng.stream):
        eng.data["train"][2].copy_(perm)
    eng._sync_lr()
    steps_per_epoch = n // BATCH
    done_in_epoch = [0]
    graph = eng._graph("train", BATCH)

    def run(k):
        """k optimisation steps; the host side is one library call per stretch of an epoch (as Engine.fit does it)."""
        while k > 0:
            if done_in_epoch[0] == 0:
                eng._zero_metrics(eng.state)            # epoch boundary: cursor back to 0
            stretch = min(k, steps_per_epoch - done_in_epoch[0])
            eng.lib.fnet_graph_launch_n(graph, stretch, eng.sp)
            done_in_epoch[0] = (done_in_epoch[0] + stretch) % steps_per_epoch
            k -= stretch

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # warm-up, and a first estimate of the step time: the timed region is the driver's `--steps` block repeated until
    # it lasts at least MIN_TIMED_SECONDS (a 20-step block of a 40 us step is 0.8 ms: shorter than one clock sample)
    run(max(args.warmup, 3))
    barrier()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record(eng.stream)
    run(args.steps)
    p1.record(eng.stream)
    barrier()
    est_block_ms = max(p0.elapsed_time(p1), 1e-3)
    repeats = int(min(max(1, math.ceil(1e3 * MIN_TIMED_SECONDS / est_block_ms)), 20000))
    if world > 1:                                       # every rank runs the same number of blocks
        t = torch.tensor([repeats], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        repeats = int(t.item())
    sampler = ClockSampler(device_index)
    sampler.wait_for_samples()
    time.sleep(1.0)                                     # the sampler has been up for >= 1 s before the first event
    barrier()
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(repeats + 1)]
    marks[0].record(eng.stream)
    for r in range(repeats):
        run(args.steps)
        marks[r + 1].record(eng.stream)
    barrier()
    blocks = [marks[r].elapsed_time(marks[r + 1]) for r in range(repeats)]
    total_ms = marks[0].elapsed_time(marks[-1])
    clocks = sampler.stop()
    ms = total_ms / repeats                              # mean block: exactly `--steps` steps
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    out = {"ms": ms, "value": world * args.steps * BATCH / (ms / 1e3), "clocks": clocks,
           "repeats": repeats, "timed_s": total_ms / 1e3, "block_ms_median": float(np.median(blocks)),
           "block_ms_min": float(min(blocks)), "block_ms_max": float(max(blocks)),
           "launch_calls_per_step": eng.launches_per_step[("train", BATCH)],
           "params": net.count_params(), "layers": len(net.layers),
           "train_mflop_per_image": 3 * net.train_flops_per_image() / 1e6}
    if not full:
        net.release()
        return out, net

    # ---- end to end through the public API with host buffers
    k = max(args.steps, int(math.ceil(MIN_TIMED_SECONDS / max(ms / args.steps / 1e3, 1e-6))))
    k = min(k, 20000)
    (xu8, _), _ = synth.make_synth_uint8(synth.SHAPES[dataset], BATCH * 8, 0, synth.CLASSES[dataset])
    xs_u8 = torch.from_numpy(np.ascontiguousarray(xu8)).pin_memory()          # the same pixels, before `/ 255`
    xs = torch.from_numpy(x_train[: BATCH * 8]).pin_memory()
    ys = torch.from_numpy(y_train[: BATCH * 8]).pin_memory()

    class HostBatches(object):
        """An endless generator of pinned host batches; with `scale` set they are the uint8 pixels."""
        def __init__(self, x, scale=None):
            self.xh, self.i = x, 0
            if scale is not None:
                self.scale = scale

        def __iter__(self):
            return self

        def __next__(self):
            lo = (self.i % 8) * BATCH
            self.i += 1
            return self.xh[lo: lo + BATCH], ys[lo: lo + BATCH]

    def wall_max(seconds):
        if world > 1:
            t = torch.tensor([seconds], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            seconds = float(t.item())
        return seconds

    def streamed(feed):
        net.fit_generator(feed, steps_per_epoch=max(args.warmup, 100), epochs=1)     # captures the slots' graphs, warms the host loop
        barrier()
        t0 = time.perf_counter()
        hist = net.fit_generator(feed, steps_per_epoch=k, epochs=1)   # ends with the epoch's loss/accuracy read
        barrier()
        return wall_max(time.perf_counter() - t0), float(hist.history["loss"][-1])

    # (1) fit_generator: the reference's streaming call (helpers.py:163).  Every step uploads its own batch from
    # pinned host memory and returns its 16-byte result; the upload of step i+1 overlaps the kernels of step i.
    # (a) the batch travels as the dataset's uint8 pixels and is divided by 255 on the device (what
    # tensorflow_generator.py:310-313 does on the host); (b) the batch travels as float32, already scaled.
    u8_s, u8_loss = streamed(HostBatches(xs_u8, scale=255.0))
    f32_s, f32_loss = streamed(HostBatches(xs))
    # (2) train_on_batch: one blocking call per step (upload, step, result, wait)
    kb = min(k, 2000)
    for i in range(3):
        eng.train_on_batch(xs[:BATCH], ys[:BATCH])
    barrier()
    t0 = time.perf_counter()
    for i in range(kb):
        lo = (i % 8) * BATCH
        eng.train_on_batch(xs[lo: lo + BATCH], ys[lo: lo + BATCH])
    barrier()
    blocking_s = wall_max(time.perf_counter() - t0)
    row = int(np.prod(x_train.shape[1:]))
    out["e2e"] = {"value": world * k * BATCH / u8_s, "unit": "images/s",
                  "h2d_bytes_per_step": int(BATCH * (row + y_train.shape[1] * 4)),
                  "d2h_bytes_per_step": 16, "steps": k, "last_loss": u8_loss,
                  "api": "Network.fit_generator on pinned host batches of uint8 pixels (generator.scale = 255: the "
                         "`astype(float32) / 255` of tensorflow_generator.py:310-313 runs on the device), 3-slot pipelined feed",
                  "fp32_feed_value": world * k * BATCH / f32_s,
                  "fp32_feed_h2d_bytes_per_step": int(BATCH * (row + y_train.shape[1]) * 4),
                  "fp32_feed_note": "the same call on float32 batches scaled on the host (4x the upload)",
                  "train_on_batch_value": world * kb * BATCH / blocking_s,
                  "train_on_batch_note": "float32 batches through the blocking per-step call (no overlap of upload and compute)"}

    # ---- per-kernel profile (CUDA events around every launch on the launching stream)
    table = eng.profile_steps("train", BATCH, steps=50)
    out["profile"] = table
    out["profiled_steps"] = 50

    # ---- one whole candidate: build + compile + 1 epoch + validation + evaluate (rank-local)
    from tensorflow_generator import TensorflowGenerator
    from featurenet_b200.population import copy_product
    TensorflowGenerator.dataset = dataset
    TensorflowGenerator.X_train, TensorflowGenerator.Y_train = x_train, y_train
    TensorflowGenerator.X_test, TensorflowGenerator.Y_test = x_test, y_test
    TensorflowGenerator.input_shape = synth.SHAPES[dataset]
    source, index, _, _ = WORKLOADS[workload]
    if index is None:
        product = source
    else:
        from products_tree import ProductSet
        with quiet():
            product = ProductSet(os.path.join(ROOT, source)).format_products()[index][0]
    net.release()
    cycles = []
    for _ in range(3):                                   # later cycles = steady state (dataset resident)
        barrier()
        t0 = time.perf_counter()
        with quiet():
            gen = TensorflowGenerator(copy_product(product), epochs=1, dataset=dataset, data_augmentation=False, batch_size=BATCH)
        barrier()
        cycles.append(time.perf_counter() - t0)
    timing = dict(gen.model.model._engine.timing)
    out["candidate_cycle_s"] = cycles[-1]
    out["candidate_accuracy"] = float(gen.accuracy)
    out["candidate_phases"] = {
        "train_steps": timing["train_steps"], "train_s": timing["train_s"], "capture_s": timing["capture_s"],
        "alloc_s": timing["alloc_s"], "validation_and_evaluate_s": timing["eval_s"], "eval_steps": timing["eval_steps"],
        "build_and_host_s": cycles[-1] - timing["train_s"] - timing["eval_s"] - timing["alloc_s"],
        "train_us_per_step": 1e6 * (timing["train_s"] - timing["capture_s"]) / max(timing["train_steps"], 1),
        "note": "host seconds; train_s spans the enqueue of the whole epoch to the read of its loss (it contains "
                "capture_s, the one-off capture of the train graphs); eval = validation after the epoch + final evaluate"}
    return out, net


POPULATION_FILE = "featurenet_b200/data/nas_5_5_10_pop32_seed7.pdt"
# BASELINE configs[3] names a 100-product population of the nas_5_5_10 space; the reference does not ship it, so it is
# drawn by featurenet_b200/data/make_populations.py (95 of its 100 products build): `bench.py --population 100`
POPULATION_FILE_100 = "featurenet_b200/data/nas_5_5_10_100Products_seed0.pdt"


def run_population_bench(args, dist, world, rank, n_train, n_test, epochs=1):
    """A fixed population (nas_5_5_10 sample) sharded by candidate over the ranks: build + train + evaluate each."""
    from featurenet_b200 import population as pop
    from featurenet_b200 import synth
    from model.keras_model import KerasFeatureModel
    from products_tree import ProductSet
    from tensorflow_generator import TensorflowGenerator
    import torch
    with quiet():
        source = POPULATION_FILE_100 if args.population > 32 else POPULATION_FILE
        formatted = ProductSet(os.path.join(ROOT, source)).format_products()
        # weak scaling: the same `--population` candidates PER GPU (the sample repeated), so that the number of
        # candidates grows with the box and the makespan is not bound by the single heaviest candidate
        formatted = formatted[: args.population] * (1 if args.population_total else world)
        (x_train, y_train), (x_test, y_test) = synth.load("cifar", n_train, n_test)
    TensorflowGenerator.dataset = "cifar"
    TensorflowGenerator.X_train, TensorflowGenerator.Y_train = x_train, y_train
    TensorflowGenerator.X_test, TensorflowGenerator.Y_test = x_test, y_test
    TensorflowGenerator.input_shape = synth.SHAPES["cifar"]
    costs, mem, buildable = [], [], 0
    for product, _ in formatted:
        with quiet():
            net = KerasFeatureModel.parse_feature_model(copy.deepcopy(product)).build(synth.SHAPES["cifar"], 10)
        costs.append(pop.estimate_cost(net, n_train, n_test, epochs) if net else 1.0)
        mem.append(pop.estimate_bytes(net, BATCH, n_test) if net else 0)
        buildable += 1 if net else 0

    workers = max(1, args.resident)
    free_bytes, _ = torch.cuda.mem_get_info()
    budget = pop.MemoryBudget(int(0.8 * free_bytes)) if workers > 1 else None

    def train_one(index, item):
        product, original = pop.copy_product(item)        # parsing reverses the product's children in place
        return pop.train_product(product, "cifar", epochs=epochs, batch_size=BATCH, product_features=original,
                                 shared_gpu=workers > 1)

    def train_giant(index, item):
        product, original = pop.copy_product(item)
        return pop.train_product(product, "cifar", epochs=epochs, batch_size=BATCH, product_features=original, data_parallel=True)

    giants = pop.pick_giants(costs, world, args.giant_share) if (args.population_total and world > 1 and not args.no_dp_giants) else []

    def one_pass():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        with quiet():
            res = pop.run_population(formatted, train_one, costs, rank=rank, world_size=world,
                                     dist=dist if world > 1 else None, workers=workers, mem_bytes=mem, budget=budget,
                                     train_giant=train_giant if giants else None, giant_share=args.giant_share)
        torch.cuda.synchronize()
        return res, pop.max_over_ranks(time.perf_counter() - t0, dist if world > 1 else None, device="cuda")

    _, first_seconds = one_pass()      # first pass also pays one-off costs: lazy kernel loading, first allocations
    results, seconds = one_pass()      # steady state, what a long evolution run sees
    valid = [r for _, r in results if r.get("valid")]
    errors = [{"candidate": i, "rank": r.get("rank"), "error": str(r.get("error"))[:300]}
              for i, r in results if r.get("error") and not r.get("valid")]
    retried = sum(1 for _, r in results if r.get("retried_alone_after"))
    return {"candidates": len(results), "valid": len(valid), "expected_valid": buildable, "errors": errors,
            "retried_alone": retried, "data_parallel_giants": giants, "seconds": seconds, "first_pass_seconds": first_seconds,
            "resident_per_gpu": workers, "epochs": epochs,
            "candidates_per_hour": 3600.0 * len(results) / seconds,
            "scaling": ("strong (%d candidates in all, LPT-sharded over all ranks)" if args.population_total else
                        "weak (%d candidates per GPU, LPT-sharded over all ranks)") % args.population,
            "what": "%s (%d products, %d buildable): build + %d epoch(s) on %d synthetic CIFAR-shape images + "
                    "validation + evaluate on %d" % (os.path.basename(source), len(results), buildable, epochs, n_train, n_test),
            "mean_accuracy": float(np.mean([r["accuracy"] for r in valid])) if valid else 0.0,
            "per_rank_candidates": [sum(1 for _, r in results if r.get("rank") == k) for k in range(world)]}


def roofline_from(table, pk):
    label = max(table, key=lambda k: table[k]["ms_total"])
    row = table[label]
    total_ms = sum(r["ms_total"] for r in table.values())
    secs = row["ms_total"] / 1e3
    intensity = row["flops"] / max(row["bytes"], 1)
    ridge = pk["tflops"] * 1e12 / (pk["hbm_gbs"] * 1e9)
    if row["flops"] and intensity >= ridge:
        achieved, peak, unit, bound = row["flops"] / secs / 1e12, pk["tflops"], "TFLOP/s", "tensor"
    else:
        achieved, peak, unit, bound = row["bytes"] / secs / 1e9, pk["hbm_gbs"], "GB/s", "hbm"
    # dram bytes per launch of this kernel from a committed `ncu --set full` capture of this program, when there is
    # one for the kernel that is dominant today (profiles/r2_traffic.json names the capture); otherwise null
    traffic, traffic_source = None, None
    tpath = os.path.join(ROOT, "profiles", "r2_traffic.json")
    if os.path.isfile(tpath):
        with open(tpath) as handle:
            entry = json.load(handle).get(label)
        if entry:
            traffic, traffic_source = entry.get("dram_bytes_per_launch"), entry.get("source")
    return {"bound": bound, "achieved": achieved, "peak": peak, "unit": unit, "frac": achieved / peak,
            "traffic": traffic, "traffic_source": traffic_source, "kernel": label, "launches": row["launches"],
            "avg_us": 1e3 * row["ms_total"] / row["launches"], "share_of_step": row["ms_total"] / total_ms,
            "algorithmic_bytes_per_launch": row["bytes"] / row["launches"],
            "algorithmic_flops_per_launch": row["flops"] / row["launches"], "peak_source": pk["source"],
            "tflops_achieved": row["flops"] / secs / 1e12}


def run_dp(args, torch, dist, world, rank, local):
    """BASELINE configs[4], the single large candidate: the `keras` template (5.34 M parameters) trained data-parallel
    over all ranks, global batch 128, through Engine's ordinary data-parallel step."""
    from featurenet_b200 import synth
    from featurenet_b200.optimizers import Adam
    os.environ["FNET_DP"] = "1"
    workload = args.workload if args.workload != "nas1" else "keras"
    model, net, dataset = build_candidate(workload)
    with quiet():
        (x_train, y_train), _ = synth.load(dataset, 12800, 128)
    n = x_train.shape[0]
    net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
    eng = net._need_engine()
    if world > 1 and eng.dp is None:
        raise SystemExit("--dp needs the process group (launch with torchrun)")
    local_rows = eng.dp.slice_of(BATCH)[0] if eng.dp else BATCH
    eng._ensure_capacity(max(local_rows, 1), max(local_rows, 1))
    eng._set_data("train", x_train, y_train)
    eng._sync_lr()
    kind = "train_dp:%d" % BATCH if eng.dp else "train"
    steps_per_epoch = n // BATCH
    done = [0]

    def run(k):
        for _ in range(k):
            if done[0] % steps_per_epoch == 0:
                eng._zero_metrics(eng.state)
            eng._step(kind, local_rows)
            done[0] += 1

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    run(max(args.warmup, 3))
    barrier()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record(eng.stream)
    run(args.steps)
    p1.record(eng.stream)
    barrier()
    repeats = int(min(max(1, math.ceil(1e3 * MIN_TIMED_SECONDS / max(p0.elapsed_time(p1), 1e-3))), 2000))
    if world > 1:
        t = torch.tensor([repeats], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        repeats = int(t.item())
    sampler = ClockSampler(local)
    sampler.wait_for_samples()
    time.sleep(1.0)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(eng.stream)
    for _ in range(repeats):
        run(args.steps)
    e1.record(eng.stream)
    barrier()
    ms = e0.elapsed_time(e1) / repeats
    clocks = sampler.stop()
    # the collective alone, same message, same stream
    ar_ms = None
    if eng.dp:
        for _ in range(5):
            eng._allreduce_grads()
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record(eng.stream)
        for _ in range(50):
            eng._allreduce_grads()
        a1.record(eng.stream)
        barrier()
        ar_ms = a0.elapsed_time(a1) / 50
    if world > 1:
        t = torch.tensor([ms, ar_ms or 0.0], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ar_ms = float(t[0].item()), float(t[1].item())
    if rank == 0:
        msg = 4 * eng.n_params
        line = {"metric": METRIC, "value": args.steps * BATCH / (ms / 1e3), "unit": "images/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": dict(config_for(workload, world), parallelism="data-parallel x%d, global batch %d (%d rows per rank), "
                               "one NCCL all-reduce of the flat gradient arena between the captured backward and optimizer graphs" % (world, BATCH, local_rows)),
                "clocks": clocks, "gpu_launches": eng.launches_per_step[(kind, local_rows)] * args.steps,
                "timing": {"repeats": repeats},
                "data_parallel": {"params": net.count_params(), "allreduce_bytes": msg if eng.dp else 0,
                                  "allreduce_ms_alone": ar_ms, "single_graph": bool(eng.dp and eng.dp_single_graph),
                                  "allreduce_algbw_gbs": (msg / (ar_ms * 1e-3) / 1e9) if ar_ms else None,
                                  "allreduce_busbw_gbs": (msg / (ar_ms * 1e-3) / 1e9 * 2 * (world - 1) / world) if ar_ms else None,
                                  "allreduce_share_of_step": (ar_ms / (ms / args.steps)) if ar_ms else None}}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_ours(args):
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for the product path)")
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if args.dp:
        return run_dp(args, torch, dist, world, rank, local)
    t_wall = time.perf_counter()
    math_mode = os.environ.get("FNET_PRECISION", "auto").lower()
    main, net = measure_workload(args.workload, args, torch, dist, world, local, full=True)
    extra = {}
    if rank == 0 and world == 1 and not args.quick:
        for other in ("lenet5", "wide"):
            if other != args.workload:
                res, _ = measure_workload(other, args, torch, dist, 1, local, full=False)
                extra[other] = {"images_per_s": res["value"], "ms_per_step": res["ms"] / args.steps,
                                "params": res["params"], "train_mflop_per_image": res["train_mflop_per_image"],
                                "launch_calls_per_step": res["launch_calls_per_step"],
                                "workload": WORKLOADS[other][3], "dtype": "f32"}
                # the same step in the other math modes (reported beside, never as the headline): fp32 = FFMA everywhere,
                # tf32x3 = split operands on the tensor core wherever it runs (both at the fp32 tolerance; the default
                # "auto" picks between them per product); tf32 = plain tf32 operands, 5e-3 per-op tolerance
                saved = os.environ.get("FNET_PRECISION")
                by_mode = {}
                for mode in ("fp32", "tf32x3", "tf32"):
                    os.environ["FNET_PRECISION"] = mode
                    alt, _ = measure_workload(other, args, torch, dist, 1, local, full=False)
                    by_mode[mode] = alt["ms"] / args.steps
                if saved is None:
                    os.environ.pop("FNET_PRECISION", None)
                else:
                    os.environ["FNET_PRECISION"] = saved
                extra[other]["ms_per_step_other_modes"] = by_mode
    population = population_small = None
    if args.population > 0:
        # BASELINE configs[2..3]: the population on the full-size synthetic set (50 000 / 10 000 images)
        population = run_population_bench(args, dist, world, rank, 50000, 10000, epochs=args.population_epochs)
        if not args.quick:
            population_small = run_population_bench(args, dist, world, rank, 10000, 2000, epochs=1)
    # every rank holds the same gathered results, so every rank takes the same decision
    for name, p in (("population", population), ("population_small", population_small)):
        if p is not None and p["valid"] != p["expected_valid"]:
            if rank == 0:
                sys.stderr.write("bench.py: %s trained %d of %d buildable candidates; errors: %s\n"
                                 % (name, p["valid"], p["expected_valid"], json.dumps(p["errors"])))
                sys.stderr.write("bench.py: refusing to report a throughput that lost candidates\n")
            if world > 1:
                dist.barrier()
                dist.destroy_process_group()
            raise SystemExit(3)
    if rank == 0:
        pk = peaks(math_mode)
        line = {
            "metric": METRIC, "value": main["value"], "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": main["ms"] / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": {"fp32": "f32", "tf32": "tf32", "tf32x3": "f32 (tf32x3 on tensor cores)", "bf16": "bf16"}.get(math_mode, "f32"), "data": "synthetic",
            "config": config_for(args.workload, world),
            "math": math_mode + " (auto = fp32 accuracy: FFMA or three-MMA tf32 split per product, both asserted <= 2e-5 of the float64 oracle)",
            "candidate": {"params": main["params"], "layers": main["layers"], "train_mflop_per_image": main["train_mflop_per_image"]},
            "timing": {"repeats": main["repeats"], "timed_s": main["timed_s"], "block_ms_median": main["block_ms_median"],
                       "block_ms_min": main["block_ms_min"], "block_ms_max": main["block_ms_max"],
                       "what": "the `--steps` block repeated `repeats` times back to back between CUDA events (>= %.1f s in all); "
                               "ms_per_step = mean block / steps, max over ranks; the clock sampler runs >= 1 s before the first event"
                               % MIN_TIMED_SECONDS},
            "clocks": main["clocks"], "e2e": main["e2e"],
            "gpu_launches": main["launch_calls_per_step"] * args.steps,
            "roofline": roofline_from(main["profile"], pk),
            "candidates_per_hour": world * 3600.0 / main["candidate_cycle_s"],
            "candidate_cycle": {"seconds": main["candidate_cycle_s"], "accuracy": main["candidate_accuracy"],
                                "what": "build + compile + 1 epoch (full-size synthetic set) + validation + evaluate",
                                "phases": main["candidate_phases"]},
            "profile": {k: {"launches": v["launches"], "ms_total": round(v["ms_total"], 4)}
                        for k, v in sorted(main["profile"].items(), key=lambda kv: -kv[1]["ms_total"])},
            "other_workloads": extra,
            "population": population,
            "population_small": population_small,
        }
        if world == 1:
            _, net2, dataset = build_candidate(args.workload)
            rate, done, elapsed, threads = cpu_port_rate(net2, dataset, seconds=args.cpu_seconds)
            line["cpu_baseline"] = {"value": rate, "unit": "images/s", "cores": threads, "kind": "port",
                                    "sample": "%d train steps of batch %d in %.1f s through oracle/torch_cpu.py"
                                              % (done, BATCH, elapsed)}
        # whole-step roofline: every launch at the better of its two ceilings vs the measured step
        ideal_s = sum(max(r["flops"] / (pk["tflops"] * 1e12), r["bytes"] / (pk["hbm_gbs"] * 1e9))
                      for r in main["profile"].values()) / main["profiled_steps"]
        line["roofline"]["whole_step"] = {
            "ideal_us": 1e6 * ideal_s, "measured_us": 1e3 * line["ms_per_step"], "frac": 1e3 * ideal_s / line["ms_per_step"],
            "formula": "sum over launches of max(flops / peak_tensor, bytes / peak_hbm), over the measured step time"}
        line["wall_s"] = time.perf_counter() - t_wall
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="nas1", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--quick", action="store_true", help="skip the secondary workloads")
    ap.add_argument("--population", type=int, default=32, help="candidates of the sharded population run (0 = skip)")
    ap.add_argument("--resident", type=int, default=8, help="candidates kept resident per GPU in the population run")
    ap.add_argument("--population-total", action="store_true", help="--population counts the candidates of the whole job "
                    "(strong scaling: BASELINE configs[3], one 100-product population over 8 GPUs) instead of per GPU")
    ap.add_argument("--no-dp-giants", action="store_true", help="with --population-total: do not train the candidates that "
                    "exceed a rank's fair share data-parallel over all ranks (plain longest-first placement)")
    ap.add_argument("--giant-share", type=float, default=0.6, help="a candidate whose estimated cost exceeds this part of a "
                    "rank's fair share of the population is trained data-parallel over all ranks")
    ap.add_argument("--population-epochs", type=int, default=1, help="epochs per candidate of the full-size population run")
    ap.add_argument("--dp", action="store_true", help="data-parallel training of ONE large candidate (the `keras` "
                    "template) over all ranks: NCCL all-reduce of the gradient arena every step")
    args = ap.parse_args()
    claim_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
