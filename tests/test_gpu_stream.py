"""The pipelined host feed behind ``fit_generator`` (helpers.py:163 is its reference call site): batches uploaded on
a copy stream while the previous step runs must train EXACTLY like the same batches fed one at a time through
``train_on_batch`` -- the kernels are the same, only the way the batch reaches the device differs."""
import numpy as np
import pytest

from tests.util import init_weights, tiny_graphs
from tests.test_gpu_network import engine_for

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def make(name, seed=3):
    net = tiny_graphs()[name]()
    weights = init_weights(net, np.random.default_rng(seed), jitter=True)
    return net, engine_for(net, weights)


def batches_for(net, sizes, seed=29):
    rng = np.random.default_rng(seed)
    classes = net.output_shape[-1]
    out = []
    for b in sizes:
        x = rng.random((b,) + tuple(net.input_shape[1:])).astype(np.float32)
        y = np.eye(classes, dtype=np.float32)[rng.integers(0, classes, b)]
        out.append((x, y))
    return out


@pytest.mark.parametrize("pinned", [False, True])
@pytest.mark.parametrize("name", sorted(tiny_graphs()))
def test_streamed_steps_equal_train_on_batch(name, pinned):
    sizes = [8, 8, 8, 8, 5, 8, 8, 8, 8, 5]          # two epochs of five steps, ragged last batch
    net_a, eng_a = make(name)
    net_b, eng_b = make(name)
    data = batches_for(net_a, sizes)
    per_step = [net_a.train_on_batch(x, y) for x, y in data]
    feed = data
    if pinned:
        feed = [(torch.from_numpy(x).pin_memory(), torch.from_numpy(y).pin_memory()) for x, y in data]

    seen = []

    from featurenet_b200.callbacks import Callback

    class Spy(Callback):
        def on_batch_end(self, batch, logs=None):
            seen.append(dict(logs))

    hist = net_b.fit_generator(iter(feed), steps_per_epoch=5, epochs=2, callbacks=[Spy()])
    for wa, wb in zip(eng_a.get_weights(), eng_b.get_weights()):
        assert np.array_equal(wa, wb)
    # epoch logs are the sample-weighted means of the per-step results
    for e in range(2):
        chunk, n = per_step[5 * e: 5 * e + 5], sizes[5 * e: 5 * e + 5]
        loss = sum(l * b for (l, _), b in zip(chunk, n)) / sum(n)
        acc = sum(a * b for (_, a), b in zip(chunk, n)) / sum(n)
        assert abs(hist.history["loss"][e] - loss) <= 1e-5 * max(1.0, abs(loss))
        assert abs(hist.history["acc"][e] - acc) <= 1e-6
    # every step reported; the lagged running sums are those of the step that used the same slot before
    assert [s["batch"] for s in seen] == [0, 1, 2, 3, 4] * 2 and [s["size"] for s in seen] == sizes
    running = np.cumsum([l * b for (l, _), b in zip(per_step[:5], sizes[:5])])
    earlier = running[4 - eng_b.RING_SLOTS]
    assert abs(seen[4]["loss_sum_lagged"] - earlier) <= 1e-4 * max(1.0, earlier)


def test_stream_then_other_paths_still_work():
    name = sorted(tiny_graphs())[0]
    net, eng = make(name)
    data = batches_for(net, [16] * 6)
    net.fit_generator(iter(data), steps_per_epoch=6, epochs=1, validation_data=data[0])
    x, y = data[0]
    assert np.isfinite(net.train_on_batch(x, y)[0])
    assert np.isfinite(net.evaluate(x, y)[0])
    h = net.fit(np.concatenate([d[0] for d in data]), np.concatenate([d[1] for d in data]), batch_size=16, epochs=1)
    assert np.isfinite(h.history["loss"][0])
    # a generator with a length needs no steps_per_epoch; a larger batch than any seen so far regrows the ring
    class Seq(object):
        def __init__(self, items):
            self.items = items
        def __len__(self):
            return len(self.items)
        def __iter__(self):
            while True:
                for it in self.items:
                    yield it
    big = batches_for(net, [40, 40, 40])
    h = net.fit_generator(Seq(big), epochs=2)
    assert len(h.history["loss"]) == 2 and np.isfinite(h.history["loss"][-1])
    with pytest.raises(ValueError):
        net.fit_generator(iter(big), epochs=1)                       # no length, no steps_per_epoch
    with pytest.raises(ValueError):
        net.fit_generator(iter([(big[0][0][:, :1], big[0][1])]), steps_per_epoch=1)    # wrong input shape
