"""fnet_dense_fwd_static: the dense forward for hosts that cannot allocate device memory.

A plug-in loaded into the reference's TensorFlow 1.x process finds the GPU already reserved by
the session, so it cannot cudaMalloc a workspace; this entry point keeps the split-reduction
scratch in a fixed array inside libfnet.so instead.  The test drives it through the C ABI twice
in a row (the second call finds the first call's leftovers in the shared array) and compares
with fnet_dense_fwd, which gets an ordinary caller-provided workspace.

Collected last on purpose (see tests/conftest.py): a run that stops at the first failure has
then already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

TOL = 2e-5      # the tolerance of tests/test_gpu_ops.py (fnet_dense_fwd may pick a different fp32 kernel for the same shape)


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def run_both(L, fl, m, k, n, act, seed):
    rng = np.random.default_rng(seed)
    x = torch.from_numpy(rng.standard_normal((m, k)).astype(np.float32)).cuda()
    w = torch.from_numpy((rng.standard_normal((k, n)) / np.sqrt(k)).astype(np.float32)).cuda()
    b = torch.from_numpy(rng.standard_normal(n).astype(np.float32)).cuda()
    need = L.fnet_dense_workspace_bytes(m, k, n)
    ws = torch.empty(max(int(need), 256), dtype=torch.uint8, device="cuda")
    want = torch.empty((m, n), device="cuda")
    got = torch.full((m, n), float("nan"), device="cuda")
    L.fnet_dense_fwd(m, k, n, fl.ACT[act], fl.MATH["fp32"], ptr(x), ptr(w), ptr(b), ptr(want), ptr(ws), ws.numel(), None)
    L.fnet_dense_fwd_static(m, k, n, fl.ACT[act], ptr(x), ptr(w), ptr(b), ptr(got), None)
    torch.cuda.synchronize()
    # and against the float64 CPU oracle of the same layer, so that the pair cannot agree on a wrong answer
    from oracle import ops
    z = ops.dense_fwd(x.cpu().numpy().astype(np.float64), w.cpu().numpy().astype(np.float64), b.cpu().numpy().astype(np.float64))
    oracle = ops.act_fwd(z, act)
    assert np.abs(got.cpu().numpy() - oracle).max() <= TOL * np.abs(oracle).max()
    return got.cpu().numpy(), want.cpu().numpy(), need


def test_two_calls_in_a_row_match_the_workspace_entry_point():
    from featurenet_b200 import lib as fl
    L = fl.get()
    # call 1: the lenet5 classifier head -- one output tile, the reduction over 16464 is split, scratch is used
    got, want, need = run_both(L, fl, 128, 16464, 10, "softmax", seed=1)
    assert 0 < need <= fl.STATIC_SCRATCH_BYTES
    assert np.abs(got - want).max() <= TOL * np.abs(want).max()
    # call 2, straight after: another shape whose partials cover only part of what call 1 left in the array
    got, want, need = run_both(L, fl, 48, 4096, 24, "tanh", seed=2)
    assert 0 < need <= fl.STATIC_SCRATCH_BYTES
    assert np.abs(got - want).max() <= TOL * np.abs(want).max()
    # and a shape with enough output tiles that its forward is not split (no scratch touched) still goes through
    got, want, _ = run_both(L, fl, 40000, 24, 12, "relu", seed=3)
    assert np.abs(got - want).max() <= TOL * np.abs(want).max()


def test_shapes_beyond_the_built_in_scratch_are_refused():
    from featurenet_b200 import lib as fl
    L = fl.get()
    m, k, n = 1024, 65536, 64
    assert L.fnet_dense_workspace_bytes(m, k, n) > fl.STATIC_SCRATCH_BYTES
    x = torch.zeros((m, k), device="cuda")
    w = torch.zeros((k, n), device="cuda")
    y = torch.zeros((m, n), device="cuda")
    with pytest.raises(fl.FnetError, match="built-in array holds"):
        L.fnet_dense_fwd_static(m, k, n, 0, ptr(x), ptr(w), None, ptr(y), None)
    with pytest.raises(fl.FnetError, match="bad GEMM shape"):
        L.fnet_dense_fwd_static(4, 0, 4, 0, ptr(x), ptr(w), None, ptr(y), None)
