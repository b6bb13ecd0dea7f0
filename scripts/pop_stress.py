"""Repeat the co-resident population pass and report, per pass, how many buildable candidates were trained and every
error text: the reproduction loop for "candidates vanish under --resident 8" (VERDICT r1, weak 4).

  python scripts/pop_stress.py --passes 50 --resident 8 [--train 10000 --test 2000] [--no-budget]
"""
import argparse
import contextlib
import copy
import io
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("FNET_SEED", "20261106")     # same weights and permutations every pass
import featurenet_b200  # noqa: E402

featurenet_b200.install_compat()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--passes", type=int, default=50)
    ap.add_argument("--resident", type=int, default=8)
    ap.add_argument("--train", type=int, default=10000)
    ap.add_argument("--test", type=int, default=2000)
    ap.add_argument("--no-budget", action="store_true")
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--per-candidate", action="store_true")
    args = ap.parse_args()
    import torch
    from featurenet_b200 import population as pop, synth
    from model.keras_model import KerasFeatureModel
    from products_tree import ProductSet
    from tensorflow_generator import TensorflowGenerator
    quiet = lambda: contextlib.redirect_stdout(io.StringIO())
    with quiet():
        formatted = ProductSet(os.path.join(ROOT, "featurenet_b200/data/nas_5_5_10_pop32_seed7.pdt")).format_products()
        (x_train, y_train), (x_test, y_test) = synth.load("cifar", args.train, args.test)
    TensorflowGenerator.dataset = "cifar"
    TensorflowGenerator.X_train, TensorflowGenerator.Y_train = x_train, y_train
    TensorflowGenerator.X_test, TensorflowGenerator.Y_test = x_test, y_test
    TensorflowGenerator.input_shape = synth.SHAPES["cifar"]
    costs, mem, buildable = [], [], 0
    for product, _ in formatted:
        with quiet():
            net = KerasFeatureModel.parse_feature_model(copy.deepcopy(product)).build((32, 32, 3), 10)
        costs.append(pop.estimate_cost(net, args.train, args.test, 1) if net else 1.0)
        mem.append(pop.estimate_bytes(net, args.batch, args.test) if net else 0)
        buildable += 1 if net else 0
    budget = None if args.no_budget else pop.MemoryBudget(int(0.8 * torch.cuda.mem_get_info()[0]))
    print("buildable %d of %d, estimated bytes %.1f GB total, largest %.1f GB" % (buildable, len(formatted), sum(mem) / 2**30, max(mem) / 2**30))

    def train_one(index, item):
        product, original = pop.copy_product(item)
        return pop.train_product(product, "cifar", epochs=1, batch_size=args.batch, product_features=original, shared_gpu=args.resident > 1 or args.per_candidate)

    reference = None
    bad = 0
    for n in range(args.passes):
        t0 = time.perf_counter()
        with quiet():
            res = pop.run_population(formatted, train_one, costs, workers=args.resident, mem_bytes=mem, budget=budget)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        valid = sum(1 for _, r in res if r.get("valid"))
        errors = [(i, r.get("error")) for i, r in res if r.get("error")]
        retried = sum(1 for _, r in res if r.get("retried_alone_after"))
        accs = [r.get("accuracy") for _, r in res]
        if reference is None:
            reference = accs
        same = accs == reference
        differ = [(i, a, b) for i, (a, b) in enumerate(zip(accs, reference)) if a != b and not (a != a and b != b)]
        if valid != buildable or not same:
            bad += 1
        print("pass %2d: valid %d/%d retried %d same_accuracies %s %.2fs peak_mem %.1f GB %s" % (
            n, valid, buildable, retried, same, dt, torch.cuda.max_memory_allocated() / 2**30, json.dumps(errors)[:600]), flush=True)
        ph = [r.get("phases") for _, r in res if r.get("phases")]
        if ph:
            keys = sorted(ph[0])
            print("   engine seconds summed over candidates: " + ", ".join("%s %.2f" % (k, sum(p[k] for p in ph)) for k in keys)
                  + ", training_time %.2f" % sum(r.get("training_time", 0) for _, r in res), flush=True)
        if args.per_candidate and n == args.passes - 1:
            for i, r in res:
                p = r.get("phases")
                if p:
                    print("   candidate %2d params %8d train %.3fs (%d steps, %.0f us/step) eval %.3fs (%d steps, %.0f us/step)" % (
                        i, r["params"], p["train_s"], p["train_steps"], 1e6 * (p["train_s"] - p["capture_s"]) / max(p["train_steps"], 1),
                        p["eval_s"], p["eval_steps"], 1e6 * p["eval_s"] / max(p["eval_steps"], 1)), flush=True)
        if not same:
            print("   differ (index, now, pass 0):", differ[:12], flush=True)
    print("RESULT bad_passes=%d of %d" % (bad, args.passes))
    return 1 if bad else 0


if __name__ == "__main__":
    raise SystemExit(main())
