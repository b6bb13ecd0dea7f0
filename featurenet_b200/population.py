"""Population scheduling: whole candidates are placed on GPUs, one process per GPU, no data-path collective.

The reference trains the candidates of a population one after the other in a single process
(/root/reference/full.py:28-36, pledge_evolution.py:117-125, full_evolution.py:244-256; its
multiprocessing attempts are commented out, evolution.py:93-98).  Candidates are independent
(build, train, evaluate) jobs, so here they are sharded by candidate:

* every rank (launched by torchrun, one per GPU) computes the SAME assignment from the same
  cost estimates -- longest-processing-time-first onto the least loaded rank -- so no work
  has to be communicated;
* each rank trains its share with the ordinary single-GPU engine;
* results travel back as Python objects over the control plane (``all_gather_object``), in
  candidate order, so report files are identical to a sequential run.

A generation barrier exists only because the evolution drivers select on all fitnesses.
"""
import contextlib
import io
import threading
import time


def estimate_cost(network, n_train, n_eval, epochs):
    """Relative cost of one candidate: analytic train flops plus a per-launch latency term.

    Small candidates are launch-bound (a step costs ~ #layers x a few microseconds), large
    ones FLOP-bound; the sum of the two orders them well enough for LPT placement.
    """
    flops_per_image = 3.0 * network.train_flops_per_image()
    steps_images = float(n_train) * epochs
    launch_term = 4.0e6 * len(network.layers)          # ~ flops a B200 could do in one launch latency, per image-batch
    return steps_images * (flops_per_image + launch_term / 128.0) + n_eval * flops_per_image / 3.0


def estimate_bytes(network, batch_size, n_eval):
    """Device memory one resident candidate needs, from the layer shapes alone (what Engine allocates: activations for
    the inference batch, gradients for the training batch, four parameter arenas, workspace allowance)."""
    per_image = 0
    for node in network.layers:
        if node.kind == "flatten":
            continue
        n = 1
        for d in node.output.shape.as_list()[1:]:
            n *= int(d)
        per_image += 4 * n * (2 if node.kind == "sepconv2d" else 1)
    from .engine import Engine
    eval_batch = int(min(n_eval, max(32, Engine.EVAL_ACT_BYTES // max(per_image, 1)), 1000))
    rows = max(eval_batch, batch_size) + batch_size               # activations + gradients
    return int(1.25 * per_image * rows + 24 * network.count_params() + (64 << 20))


class MemoryBudget(object):
    """Admission control for candidates that share one GPU: a candidate starts once its estimated footprint fits in
    what is left of the budget.  Without it, which candidates happen to be resident together decides whether one of
    them runs out of memory -- and the reference's convention turns that into accuracy 0 (helpers.py:172-174)."""

    def __init__(self, total_bytes):
        self.total = int(total_bytes)
        self.free = int(total_bytes)
        self.cv = threading.Condition()

    def acquire(self, need):
        need = min(int(need), self.total)                         # a candidate larger than the budget runs alone
        with self.cv:
            while self.free < need:
                self.cv.wait()
            self.free -= need
        return need

    def release(self, held):
        with self.cv:
            self.free += held
            self.cv.notify_all()


def is_oom(result):
    return isinstance(result, dict) and (result.get("oom") or "OutOfMemory" in str(result.get("error", "")) or
                                         "out of memory" in str(result.get("error", "")))


def assign_lpt(costs, world_size):
    """Longest-processing-time-first: returns a list (per rank) of candidate indices, deterministic."""
    order = sorted(range(len(costs)), key=lambda i: (-costs[i], i))
    loads = [0.0] * world_size
    shards = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda j: (loads[j], j))
        shards[r].append(i)
        loads[r] += costs[i]
    return shards


def assign_lpt_subset(indices, costs, world_size):
    """assign_lpt over a subset of the candidates (the others are trained some other way)."""
    order = sorted(indices, key=lambda i: (-costs[i], i))
    loads = [0.0] * world_size
    shards = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda j: (loads[j], j))
        shards[r].append(i)
        loads[r] += costs[i]
    return shards


def pick_giants(costs, world_size, share=0.6):
    """Candidates too large to be one GPU's job: each would take more than `share` of a rank's fair part of the whole
    population, so whatever rank gets it decides the makespan.  They are trained data-parallel over ALL ranks first."""
    if world_size < 2 or not costs:
        return []
    fair = sum(costs) / world_size
    return sorted(i for i in range(len(costs)) if costs[i] > share * fair)


def max_over_ranks(value, dist=None, device=None):
    """The slowest rank's value (a multi-GPU time is the max over ranks, never a wall clock of one)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_population(products, train_one, costs, rank=0, world_size=1, dist=None, workers=1, mem_bytes=None, budget=None,
                   train_giant=None, giant_share=0.6):
    """Train this rank's share of ``products``; gather (index, result) pairs on every rank, in index order.

    train_one(index, product) -> picklable result (e.g. dict(accuracy=..., history=..., params=...)).
    Failures are reported, not raised: an invalid or out-of-memory candidate yields ``{"error": ...}``,
    as the reference turns them into accuracy 0 (tensorflow_generator.py:109-111, helpers.py:172-174).

    workers > 1 keeps that many candidates resident on this rank's GPU at once, each driven by its own host
    thread on its own CUDA stream.  FeatureNet's candidates are mostly launch-latency-bound chains of tiny
    kernels (a step of a 6-feature network is ~15 launches of a few microseconds), so several chains
    interleave on the 148 SMs almost for free; the work queue is longest-first, pulled dynamically.

    mem_bytes[i] (estimate_bytes) with a MemoryBudget makes the co-residency memory-aware; a candidate that still runs
    out of memory beside its neighbours is trained again ALONE after the others are done, so that only a candidate
    that does not fit the GPU by itself is scored 0.

    train_giant(index, product), when given with an initialised process group, is called by ALL ranks together for
    every candidate that pick_giants() names, before the rest is sharded: the one-large-candidate mode of the engine
    (featurenet_b200.dataparallel) applied to the candidates that would otherwise be the tail of the population.
    """
    giants = pick_giants(costs, world_size, giant_share) if (train_giant is not None and dist is not None and dist.is_initialized()) else []
    giant_results = []
    for index in giants:
        started = time.time()
        try:
            result = train_giant(index, products[index])
        except Exception as e:                       # noqa: BLE001
            result = {"error": "%s: %s" % (type(e).__name__, e), "accuracy": 0}
        if isinstance(result, dict):
            result.setdefault("seconds", time.time() - started)
            result["rank"], result["data_parallel"] = -1, world_size
        giant_results.append((index, result))
    if giants:
        box = [giant_results]
        dist.broadcast_object_list(box, src=0)       # every rank keeps rank 0's record: identical result lists everywhere
        giant_results = box[0]
    rest = [i for i in range(len(costs)) if i not in set(giants)]
    mine = assign_lpt_subset(rest, costs, world_size)[rank]

    def one(index):
        started = time.time()
        held = budget.acquire(mem_bytes[index]) if (budget is not None and mem_bytes is not None) else 0
        try:
            result = train_one(index, products[index])
        except Exception as e:                       # noqa: BLE001 -- one bad candidate must not kill the population
            import traceback
            result = {"error": "%s: %s" % (type(e).__name__, e), "accuracy": 0,
                      "traceback": traceback.format_exc(limit=6)}
        finally:
            if held:
                budget.release(held)
        if isinstance(result, dict):
            result.setdefault("seconds", time.time() - started)
            result.setdefault("rank", rank)
        return (index, result)

    if workers > 1 and len(mine) > 1:
        from concurrent.futures import ThreadPoolExecutor
        initializer = None
        try:
            import torch
            if torch.cuda.is_available():
                device = torch.cuda.current_device()     # the current device is per thread: hand it to the workers
                initializer = lambda: torch.cuda.set_device(device)
        except ImportError:
            pass
        with ThreadPoolExecutor(max_workers=workers, initializer=initializer) as pool:
            local = list(pool.map(one, mine))            # `mine` is already longest-first
    else:
        local = [one(index) for index in mine]
    retry = [k for k, (_, result) in enumerate(local) if is_oom(result)]
    if retry and workers > 1:
        try:
            import torch
            torch.cuda.synchronize()
            torch.cuda.empty_cache()
        except ImportError:
            pass
        for k in retry:
            index, first = local[k]
            index, again = one(index)
            if isinstance(again, dict):
                again["retried_alone_after"] = str(first.get("error", "out of memory"))[:200]
            local[k] = (index, again)
    if dist is not None and dist.is_initialized() and world_size > 1:
        gathered = [None] * world_size
        dist.all_gather_object(gathered, local)
        merged = [pair for part in gathered for pair in part]
    else:
        merged = local
    merged = merged + giant_results
    return sorted(merged, key=lambda pair: pair[0])


def copy_product(item):
    """A product (list of block dicts, or a (product, original features) pair) that can be parsed without changing the
    caller's copy.  Parsing reverses `children` IN PLACE on the block dicts (keras_model.py:199) and on the cell dicts two
    levels below (cell.py:87 of the reference), so a population that is trained twice must hand out copies; a deepcopy
    of the whole feature tree costs 4.5 ms per candidate (a quarter of the host time of a small candidate), this one
    copies the three levels that are written to and shares the rest."""
    if isinstance(item, tuple):
        return (copy_product(item[0]),) + tuple(item[1:])
    if isinstance(item, str):
        return item

    def level(node, depth):
        fresh = dict(node)
        kids = node.get("children")
        if kids is not None:
            fresh["children"] = [level(k, depth - 1) if depth > 0 and isinstance(k, dict) else k for k in kids]
        return fresh

    return [level(block, 2) for block in item]


def train_product(product, dataset, epochs, batch_size, product_features=None, data_augmentation=False, shared_gpu=False,
                  data_parallel=False):
    """One candidate through the reference-named runner; returns the fields the drivers log.

    shared_gpu=True is for candidates that run side by side on one GPU (run_population(workers > 1)): the runner
    must not tear down "the previous candidate" (it may be a neighbour that is still training), so this function
    releases its own network when it is done, and it does not touch the process-wide stdout.
    """
    from tensorflow_generator import TensorflowGenerator
    import helpers
    import os
    helpers.clear_failure()
    phases = None
    if data_parallel:
        # every rank of the process group makes this call together: the engine cuts each batch into per-rank slices
        previous = os.environ.get("FNET_DP")
        os.environ["FNET_DP"] = "1"
        try:
            with contextlib.redirect_stdout(io.StringIO()):
                gen = TensorflowGenerator(product, epochs=epochs, dataset=dataset, data_augmentation=data_augmentation,
                                          batch_size=batch_size, product_features=product_features, clear_memory=False)
            network = getattr(getattr(gen, "model", None), "model", None)
            if network is not None:
                network.release()
        finally:
            if previous is None:
                os.environ.pop("FNET_DP", None)
            else:
                os.environ["FNET_DP"] = previous
    elif shared_gpu:
        gen = TensorflowGenerator(product, epochs=epochs, dataset=dataset, data_augmentation=data_augmentation,
                                  batch_size=batch_size, product_features=product_features, clear_memory=False)
        network = getattr(getattr(gen, "model", None), "model", None)
        if network is not None:
            engine = getattr(network, "_engine", None)
            phases = dict(engine.timing) if engine is not None else None
            network.release()
    else:
        with contextlib.redirect_stdout(io.StringIO()):
            gen = TensorflowGenerator(product, epochs=epochs, dataset=dataset, data_augmentation=data_augmentation,
                                      batch_size=batch_size, product_features=product_features)
    out = {"accuracy": float(gen.accuracy), "params": int(gen.params), "training_time": float(gen.training_time),
           "history": (list(gen.history[0]), list(gen.history[1])), "valid": bool(gen.params)}
    if phases:                        # host seconds of this candidate's engine (alloc, capture, train, eval)
        out["phases"] = phases
    failure = helpers.last_failure()
    if failure:                       # the runner swallowed it (reference convention); the scheduler wants to know
        out["oom"], out["error"] = True, failure
    return out
