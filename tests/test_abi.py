"""The drop-in boundary without a GPU: libfnet.so builds, loads and exports exactly what include/fnet.h declares,
argument checking works (no kernel is launched here), and the ctypes table matches the header."""
import ctypes as C
import os
import re

import pytest

from featurenet_b200 import lib as fl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    text = open(os.path.join(ROOT, "include", "fnet.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fnet_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def L():
    if not os.path.isfile(fl.LIB_PATH):
        fl.build()
    return fl.get()


def test_every_declared_symbol_is_exported_and_bound(L):
    declared = header_functions()
    assert len(declared) >= 45
    dll = C.CDLL(fl.LIB_PATH)
    for name in declared:
        assert hasattr(dll, name), "%s is declared in include/fnet.h but not exported" % name
    assert sorted(fl.SIGNATURES) == declared


def test_version_and_error_reporting(L):
    assert L.fnet_version() == 100
    with pytest.raises(fl.FnetError, match="bad GEMM shape"):
        L.fnet_dense_fwd(4, -1, 3, 0, 0, None, None, None, None, None, 0, None)
    assert b"bad GEMM shape" in L.fnet_last_error()
    with pytest.raises(fl.FnetError, match="null descriptor"):
        L.fnet_conv2d_fwd(None, None, None, None, None, None, 0, None)
    bad = fl.Conv2dDesc(1, 8, 8, 3, 8, 8, 4, 3, 3, 0, 1, 1, 1, 1, 1, 0, 0)          # stride 0
    with pytest.raises(fl.FnetError, match="geometry"):
        L.fnet_conv2d_fwd(bad, None, None, None, None, None, 0, None)
    with pytest.raises(fl.FnetError, match="unknown activation"):
        L.fnet_act_fwd(17, 4, 4, None, None, None)
    with pytest.raises(fl.FnetError):
        L.fnet_copy_rows(1, 4, 3, 4, 0, 2, 8, None, None, 0, None)                  # rows 3..5 of a 4-row tensor


def test_workspace_queries_are_host_only(L):
    # lenet5 head forward: one output tile, K = 16464 -> the reduction is split, partials need room
    assert L.fnet_dense_workspace_bytes(128, 16464, 10) >= 2 * 128 * 10 * 4
    assert L.fnet_dense_workspace_bytes(0, 16, 16) == 0
    d = fl.Conv2dDesc(128, 28, 28, 1, 28, 28, 6, 5, 5, 1, 1, 2, 2, 1, 1, 2, 0)
    assert L.fnet_conv2d_workspace_bytes(C.byref(d)) >= 26 * 6 * 4                   # split filter gradient
    assert L.fnet_bn_workspace_bytes(32) == 32 * 64 * 2 * 4
    assert L.fnet_dwconv2d_workspace_bytes(C.byref(fl.Conv2dDesc(8, 16, 16, 40, 16, 16, 40, 3, 3, 1, 1, 1, 1, 1, 1, 0, 0))) > 0


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(fl.FnetError, match="no CPU fallback"):
        fl._Lib(str(tmp_path / "libfnet.so"))
