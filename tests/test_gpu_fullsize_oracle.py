"""The GEMM-shaped kernels against a float64 oracle AT the sizes BASELINE.json's configs run them (batch 128).

tests/test_gpu_properties.py checks these shapes through identities (adjointness, linearity, determinism);
a forward kernel that is wrong but linear passes those.  Here every product of every full-size layer -- forward,
data gradient, filter gradient, bias gradient -- is compared element by element with
oracle/torch_cpu.py::conv_products64 (torch CPU kernels in float64, itself checked against the NumPy oracle at small
sizes in tests/test_oracle_torch.py), in all four math modes.  These are the shapes that reach the big-tile variants
of igemm_simt.cu (256x{8,16}, 128x{32,64}), the TMA-fed kernels and the streaming head kernels (K >= 32 768).

Bars: fp32, tf32x3 and auto within the stated 1e-4 relative (max-norm); tf32 within 5e-3 (10-bit operands).
"""
import ctypes as C

import numpy as np
import pytest

from oracle.torch_cpu import conv_products64

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from tests.test_gpu_ops import conv_desc, ptr, workspace, relerr  # noqa: E402
from tests.test_gpu_properties import FULL_CONVS, FULL_DENSE  # noqa: E402

MODES = [("fp32", 1e-4), ("tf32", 5e-3), ("tf32x3", 1e-4), ("auto", 1e-4)]


@pytest.fixture(scope="module")
def L():
    from featurenet_b200 import lib as fl
    return fl.get()


_ORACLE = {}


def conv_oracle(case_id):
    """Inputs and float64 results of one full-size layer, computed once and shared by the four math modes."""
    if case_id not in _ORACLE:
        xs, k, kout, s = FULL_CONVS[case_id]
        rng = np.random.default_rng(500 + case_id)
        x = rng.standard_normal(xs).astype(np.float32)
        w = (0.2 * rng.standard_normal((k[0], k[1], xs[3], kout))).astype(np.float32)
        b = rng.standard_normal(kout).astype(np.float32)
        from oracle import ops
        ho = ops.same_or_valid(xs[1], k[0], s[0], "same")[0]
        wo = ops.same_or_valid(xs[2], k[1], s[1], "same")[0]
        dy = rng.standard_normal((xs[0], ho, wo, kout)).astype(np.float32)
        _ORACLE.clear()                       # one layer's float64 tensors at a time (up to 0.5 GB each)
        _ORACLE[case_id] = (x, w, b, dy) + conv_products64(x, w, b, dy, s, "same")
    return _ORACLE[case_id]


@pytest.mark.parametrize("math,tol", MODES)
@pytest.mark.parametrize("case_id", range(len(FULL_CONVS)))
def test_conv_products_match_float64_oracle_at_batch_128(L, case_id, math, tol):
    from featurenet_b200 import lib as fl
    xs, k, kout, s = FULL_CONVS[case_id]
    x, w, b, dy, y64, dx64, dw64, db64 = conv_oracle(case_id)
    desc = conv_desc(fl, xs, k, kout, s, "same", (1, 1), None, math)
    ws = workspace(L.fnet_conv2d_workspace_bytes(C.byref(desc)))
    xd, wd, bd, dyd = (torch.from_numpy(a).cuda() for a in (x, w, b, dy))
    y = torch.empty(dy.shape, device="cuda")
    dx, dw, db = torch.empty_like(xd), torch.empty_like(wd), torch.empty_like(bd)
    L.fnet_conv2d_fwd(desc, ptr(xd), ptr(wd), ptr(bd), ptr(y), ptr(ws), ws.numel(), None)
    L.fnet_conv2d_bwd_data(desc, ptr(dyd), ptr(wd), ptr(dx), 0, ptr(ws), ws.numel(), None)
    L.fnet_conv2d_bwd_filter(desc, ptr(xd), ptr(dyd), ptr(dw), ptr(db), ptr(ws), ws.numel(), None)
    torch.cuda.synchronize()
    assert relerr(y.cpu().numpy(), y64) <= tol, "forward"
    assert relerr(dx.cpu().numpy(), dx64) <= tol, "data gradient"
    assert relerr(dw.cpu().numpy(), dw64) <= tol, "filter gradient"
    assert relerr(db.cpu().numpy(), db64) <= tol, "bias gradient"


@pytest.mark.parametrize("math,tol", MODES)
@pytest.mark.parametrize("m,k,n", FULL_DENSE)
def test_dense_products_match_float64_oracle_at_full_size(L, m, k, n, math, tol):
    from featurenet_b200 import lib as fl
    mode = fl.MATH[math]
    rng = np.random.default_rng(m + k + n)
    x = rng.standard_normal((m, k)).astype(np.float32)
    w = (rng.standard_normal((k, n)) / np.sqrt(k)).astype(np.float32)
    b = rng.standard_normal(n).astype(np.float32)
    dy = rng.standard_normal((m, n)).astype(np.float32)
    xt, wt, dyt = (torch.from_numpy(a).double() for a in (x, w, dy))
    y64 = (xt @ wt + torch.from_numpy(b).double()).numpy()
    dx64, dw64, db64 = (dyt @ wt.t()).numpy(), (xt.t() @ dyt).numpy(), dyt.sum(dim=0).numpy()
    xd, wd, bd, dyd = (torch.from_numpy(a).cuda() for a in (x, w, b, dy))
    ws = workspace(L.fnet_dense_workspace_bytes(m, k, n))
    y, dx, dw, db = (torch.empty(*shape, device="cuda") for shape in ((m, n), (m, k), (k, n), (n,)))
    L.fnet_dense_fwd(m, k, n, 0, mode, ptr(xd), ptr(wd), ptr(bd), ptr(y), ptr(ws), ws.numel(), None)
    L.fnet_dense_bwd_data(m, k, n, mode, ptr(dyd), ptr(wd), ptr(dx), 0, ptr(ws), ws.numel(), None)
    L.fnet_dense_bwd_filter(m, k, n, mode, ptr(xd), ptr(dyd), ptr(dw), ptr(db), ptr(ws), ws.numel(), None)
    torch.cuda.synchronize()
    assert relerr(y.cpu().numpy(), y64) <= tol, "forward"
    assert relerr(dx.cpu().numpy(), dx64) <= tol, "data gradient"
    assert relerr(dw.cpu().numpy(), dw64) <= tol, "filter gradient"
    assert relerr(db.cpu().numpy(), db64) <= tol, "bias gradient"
