"""Candidate sharding on CPU: LPT placement, and the N>1 control plane under gloo with world_size 2."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from featurenet_b200 import population as pop


def test_lpt_assignment_is_balanced_and_deterministic():
    costs = [9.0, 1.0, 8.0, 2.0, 7.0, 3.0, 6.0, 4.0, 5.0]
    shards = pop.assign_lpt(costs, 2)
    assert sorted(i for s in shards for i in s) == list(range(9))
    loads = [sum(costs[i] for i in s) for s in shards]
    assert abs(loads[0] - loads[1]) <= max(costs) * 0.15
    assert pop.assign_lpt(costs, 2) == shards
    assert pop.assign_lpt(costs, 1) == [sorted(range(9), key=lambda i: (-costs[i], i))]
    assert pop.assign_lpt([], 4) == [[], [], [], []]


def test_single_rank_run_collects_results_and_survives_a_bad_candidate():
    def train_one(index, product):
        if product == "bad":
            raise ValueError("cannot build")
        return {"accuracy": 0.1 * index}
    out = pop.run_population(["a", "bad", "c"], train_one, [1.0, 1.0, 1.0])
    assert [i for i, _ in out] == [0, 1, 2]
    assert out[1][1]["accuracy"] == 0 and "cannot build" in out[1][1]["error"]
    assert out[2][1]["accuracy"] == pytest.approx(0.2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, queue):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        products = ["p%d" % i for i in range(7)]
        costs = [7.0, 1.0, 6.0, 2.0, 5.0, 3.0, 4.0]
        trained_here = []

        def train_one(index, product):
            trained_here.append(index)
            return {"accuracy": index / 10.0, "name": product}

        results = pop.run_population(products, train_one, costs, rank=rank, world_size=world, dist=dist)
        slowest = pop.max_over_ranks(10.0 * (rank + 1), dist)
        queue.put((rank, trained_here, results, slowest))
    finally:
        dist.destroy_process_group()


def test_two_ranks_under_gloo_share_the_population_without_overlap():
    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, queue)) for r in range(2)]
    for p in procs:
        p.start()
    got = [queue.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    got.sort()
    (r0, mine0, res0, slow0), (r1, mine1, res1, slow1) = got
    assert sorted(mine0 + mine1) == list(range(7)) and not set(mine0) & set(mine1)
    assert res0 == res1                                            # every rank ends with the full, ordered result list
    assert [i for i, _ in res0] == list(range(7))
    assert [r["name"] for _, r in res0] == ["p%d" % i for i in range(7)]
    assert {r["rank"] for _, r in res0} == {0, 1}
    assert slow0 == slow1 == 20.0                                  # max over ranks


def test_cost_estimate_orders_small_and_large_candidates():
    from tests.util import build_template
    small = build_template("lenet5", (28, 28, 1))
    large = build_template("keras", (32, 32, 3))
    assert pop.estimate_cost(large, 50000, 10000, 1) > 3 * pop.estimate_cost(small, 50000, 10000, 1)


def test_memory_budget_serialises_what_does_not_fit():
    import threading
    import time
    from featurenet_b200.population import MemoryBudget
    budget = MemoryBudget(100)
    active, peak = [0], [0]
    lock = threading.Lock()

    def job(need):
        held = budget.acquire(need)
        with lock:
            active[0] += held
            peak[0] = max(peak[0], active[0])
        time.sleep(0.02)
        with lock:
            active[0] -= held
        budget.release(held)

    threads = [threading.Thread(target=job, args=(n,)) for n in (60, 60, 30, 500, 10, 40)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=10)
        assert not t.is_alive()
    assert peak[0] <= 100 and budget.free == 100


def test_copy_product_protects_the_callers_tree_and_parses_the_same():
    """Parsing mutates the product tree in place; the scheduler's cheap copy must protect the caller exactly as a deepcopy."""
    import contextlib
    import copy
    import io
    import os
    from model.keras_model import KerasFeatureModel
    from products_tree import ProductSet
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with contextlib.redirect_stdout(io.StringIO()):
        formatted = ProductSet(os.path.join(root, "featurenet_b200/data/nas_5_5_10_pop32_seed7.pdt")).format_products()[:8]
    pristine = copy.deepcopy(formatted)
    for item, reference in zip(formatted, pristine):
        with contextlib.redirect_stdout(io.StringIO()):
            a = KerasFeatureModel.parse_feature_model(pop.copy_product(item)[0]).build((32, 32, 3), 10)
            b = KerasFeatureModel.parse_feature_model(copy.deepcopy(reference)[0]).build((32, 32, 3), 10)
            again = KerasFeatureModel.parse_feature_model(pop.copy_product(item)[0]).build((32, 32, 3), 10)
        assert item == reference                                     # the caller's tree is untouched
        assert bool(a) == bool(b)
        if a:
            strip = lambda rows: [{k: v for k, v in r.items() if k != "name"} for r in rows]
            assert strip(a.describe()) == strip(b.describe()) == strip(again.describe())


def _giant_worker(rank, world, port, queue):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        products = ["p%d" % i for i in range(6)]
        costs = [1.0, 30.0, 1.0, 2.0, 1.0, 1.0]            # candidate 1 alone exceeds a rank's fair share (18)
        alone, together = [], []

        def train_one(index, product):
            alone.append(index)
            return {"accuracy": index / 10.0}

        def train_giant(index, product):
            together.append(index)
            t = torch.tensor([1.0])
            dist.all_reduce(t)                              # every rank is here: a collective completes
            return {"accuracy": 0.5, "ranks_present": int(t.item())}

        results = pop.run_population(products, train_one, costs, rank=rank, world_size=world, dist=dist, train_giant=train_giant)
        queue.put((rank, alone, together, results))
    finally:
        dist.destroy_process_group()


def test_giants_are_trained_by_all_ranks_together_and_the_rest_is_sharded():
    assert pop.pick_giants([1.0, 30.0, 1.0, 2.0, 1.0, 1.0], 2) == [1]
    assert pop.pick_giants([1.0] * 8, 2) == [] and pop.pick_giants([5.0, 1.0], 1) == []
    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_giant_worker, args=(r, 2, port, queue)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(queue.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, alone0, together0, res0), (_, alone1, together1, res1) = got
    assert together0 == together1 == [1]                            # both ranks trained the giant, together
    assert sorted(alone0 + alone1) == [0, 2, 3, 4, 5] and not set(alone0) & set(alone1)
    assert res0 == res1 and [i for i, _ in res0] == list(range(6))
    assert res0[1][1]["ranks_present"] == 2 and res0[1][1]["data_parallel"] == 2
