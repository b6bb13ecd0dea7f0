"""ctypes binding of libfnet.so (the C ABI declared in include/fnet.h).

There is no fallback: if the shared library is missing or a call fails, an exception is
raised.  ``build()`` compiles the library in-tree with nvcc for sm_100a.
"""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(CSRC, "libfnet.so")

ACT = {None: 0, "linear": 0, "relu": 1, "tanh": 2, "sigmoid": 3, "softmax": 4}
POOL_MAX, POOL_AVG = 0, 1
MATH = {"fp32": 0, "tf32": 1, "bf16": 2, "tf32x3": 3, "auto": 4}
STATIC_SCRATCH_BYTES = 4 << 20      # FNET_STATIC_SCRATCH_BYTES in include/fnet.h


class FnetError(RuntimeError):
    pass


class Conv2dDesc(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "batch", "in_h", "in_w", "in_c", "out_h", "out_w", "out_c", "kh", "kw", "stride_h", "stride_w",
        "pad_top", "pad_left", "dil_h", "dil_w", "act", "math")]


class Pool2dDesc(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "batch", "in_h", "in_w", "channels", "out_h", "out_w", "kh", "kw", "stride_h", "stride_w",
        "pad_top", "pad_left", "kind")]


_P = C.c_void_p
_I32, _I64, _U32, _U64, _F, _SZ = C.c_int32, C.c_int64, C.c_uint32, C.c_uint64, C.c_float, C.c_size_t
_CD, _PD = C.POINTER(Conv2dDesc), C.POINTER(Pool2dDesc)

# name -> (restype, argtypes); every symbol of include/fnet.h
SIGNATURES = {
    "fnet_version": (C.c_int, []),
    "fnet_last_error": (C.c_char_p, []),
    "fnet_conv2d_workspace_bytes": (_SZ, [_CD]),
    "fnet_conv2d_fwd": (C.c_int, [_CD, _P, _P, _P, _P, _P, _SZ, _P]),
    "fnet_conv2d_bwd_data": (C.c_int, [_CD, _P, _P, _P, C.c_int, _P, _SZ, _P]),
    "fnet_conv2d_bwd_data_act": (C.c_int, [_CD, _P, _P, _P, C.c_int, C.c_int, _P, _P, _SZ, _P]),
    "fnet_conv2d_bwd_filter": (C.c_int, [_CD, _P, _P, _P, _P, _P, _SZ, _P]),
    "fnet_dense_workspace_bytes": (_SZ, [_I64, _I64, _I64]),
    "fnet_dense_fwd": (C.c_int, [_I64, _I64, _I64, C.c_int, C.c_int, _P, _P, _P, _P, _P, _SZ, _P]),
    "fnet_dense_fwd_static": (C.c_int, [_I64, _I64, _I64, C.c_int, _P, _P, _P, _P, _P]),
    "fnet_dense_bwd_data": (C.c_int, [_I64, _I64, _I64, C.c_int, _P, _P, _P, C.c_int, _P, _SZ, _P]),
    "fnet_dense_bwd_data_act": (C.c_int, [_I64, _I64, _I64, C.c_int, _P, _P, _P, C.c_int, C.c_int, _P, _P, _SZ, _P]),
    "fnet_dense_bwd_filter": (C.c_int, [_I64, _I64, _I64, C.c_int, _P, _P, _P, _P, _P, _SZ, _P]),
    "fnet_dwconv2d_workspace_bytes": (_SZ, [_CD]),
    "fnet_dwconv2d_fwd": (C.c_int, [_CD, _P, _P, _P, _P, _P]),
    "fnet_dwconv2d_bwd_data": (C.c_int, [_CD, _P, _P, _P, C.c_int, _P]),
    "fnet_dwconv2d_bwd_filter": (C.c_int, [_CD, _P, _P, _P, _P, _P, _SZ, _P]),
    "fnet_pool2d_fwd": (C.c_int, [_PD, _P, _P, _P, _P]),
    "fnet_pool2d_bwd": (C.c_int, [_PD, _P, _P, _P, C.c_int, _P]),
    "fnet_gap_fwd": (C.c_int, [_I32, _I32, _I32, _P, _P, _P]),
    "fnet_gap_bwd": (C.c_int, [_I32, _I32, _I32, _P, _P, C.c_int, _P]),
    "fnet_bn_workspace_bytes": (_SZ, [_I32]),
    "fnet_bn_fwd_train": (C.c_int, [_I32, _I32, _I32, _F, _F, _P, _P, _P, _P, _P, _P, _P, _P, _P, _SZ, _P]),
    "fnet_bn_fwd_infer": (C.c_int, [_I32, _I32, _I32, _F, _P, _P, _P, _P, _P, _P, _P]),
    "fnet_bn_bwd_infer": (C.c_int, [_I32, _I32, _I32, _F, _P, _P, _P, _P, C.c_int, _P]),
    "fnet_l2_attack_step": (C.c_int, [_I32, _I64, _P, _P, _P, _P, _F, _F, _F, _F, _P, _P]),
    "fnet_bn_bwd": (C.c_int, [_I32, _I32, _I32, _P, _P, _P, _P, _P, _P, _P, _P, C.c_int, _P, _SZ, _P]),
    "fnet_bn_bwd_act": (C.c_int, [_I32, _I32, _I32, _P, _P, _P, _P, _P, _P, _P, _P, C.c_int, C.c_int, _P, _SZ, _P]),
    "fnet_act_fwd": (C.c_int, [C.c_int, _I64, _I64, _P, _P, _P]),
    "fnet_act_bwd": (C.c_int, [C.c_int, _I64, _I64, _P, _P, _P, C.c_int, _P]),
    "fnet_dropout_fwd": (C.c_int, [_I64, _F, _U64, _U32, _P, _P, _P, _P]),
    "fnet_dropout_bwd": (C.c_int, [_I64, _F, _U64, _U32, _P, _P, _P, C.c_int, _P]),
    "fnet_add_fwd": (C.c_int, [_I64, _P, _P, _P, _P]),
    "fnet_mul_fwd": (C.c_int, [_I64, _P, _P, _P, _P]),
    "fnet_mul_bwd": (C.c_int, [_I64, _P, _P, _P, C.c_int, _P]),
    "fnet_accumulate": (C.c_int, [_I64, _P, _P, C.c_int, _P]),
    "fnet_copy_rows": (C.c_int, [_I32, _I32, _I32, _I32, _I32, _I32, _I32, _P, _P, C.c_int, _P]),
    "fnet_pad2d_fwd": (C.c_int, [_I32, _I32, _I32, _I32, _I32, _I32, _P, _P, _P]),
    "fnet_pad2d_bwd": (C.c_int, [_I32, _I32, _I32, _I32, _I32, _I32, _P, _P, C.c_int, _P]),
    "fnet_fill": (C.c_int, [_I64, _F, _P, _P]),
    "fnet_softmax_cce": (C.c_int, [_I32, _I32, _P, _P, _F, C.c_int, _P, _P, _P, _P, _P]),
    "fnet_step_begin": (C.c_int, [_P, _F, _F, _P]),
    "fnet_step_end": (C.c_int, [_P, _I32, _P]),
    "fnet_adam_step": (C.c_int, [_I64, _P, _P, _P, _P, _P, _F, _F, _F, _I32, _P]),
    "fnet_sgd_step": (C.c_int, [_I64, _P, _P, _P, _P, _F, C.c_int, _I32, _P]),
    "fnet_gather_batch": (C.c_int, [_I32, _I64, _I64, _P, _P, _P, _P, _P, _P, C.c_int, _F, _F, _P]),
    "fnet_concat_rows_fwd": (C.c_int, [_I32, _I32, _I32, _I32, _P, _P, _P, _P]),
    "fnet_concat_rows_bwd": (C.c_int, [_I32, _I32, _I32, _I32, _P, _P, _P, C.c_int, C.c_int, _P]),
    "fnet_concat_rows_bwd_act": (C.c_int, [_I32, _I32, _I32, _I32, _P, _P, _P, C.c_int, C.c_int, C.c_int, _P, C.c_int, _P, _P]),
    "fnet_gather_rows": (C.c_int, [_I32, _I64, _P, _P, _P, _P, _P]),
    "fnet_gather_shift_flip": (C.c_int, [_I32, _I32, _I32, _I32, _P, _P, _P, _P, _P, _P, _P, _P]),
    "fnet_u8_to_f32": (C.c_int, [_I64, _P, _F, _P, _P]),
    "fnet_head_fwd_loss_workspace_bytes": (_SZ, [_I64]),
    "fnet_head_fwd_loss": (C.c_int, [_I64, _I64, _I64, _P, _P, _P, _P, _P, _F, _P, _P, _P, _P, _P, _SZ, _P]),
    "fnet_fused_step_grid": (C.c_int, []),
    "fnet_fused_conv_supported": (C.c_int, [_CD, C.c_int, C.c_int]),
    "fnet_fused_head_supported": (C.c_int, [_I64, _I64, _I64]),
    "fnet_fused_step": (C.c_int, [_P, _I32, _P, _P]),
    "fnet_fused_step_timed": (C.c_int, [_P, _I32, _P, _P, _P]),
    "fnet_stream_create": (C.c_int, [C.POINTER(_P)]),
    "fnet_stream_destroy": (C.c_int, [_P]),
    "fnet_graph_begin": (C.c_int, [_P]),
    "fnet_graph_end": (C.c_int, [_P, C.POINTER(_P)]),
    "fnet_graph_launch": (C.c_int, [_P, _P]),
    "fnet_graph_launch_n": (C.c_int, [_P, _I32, _P]),
    "fnet_state_set_lr": (C.c_int, [_P, _F, _P]),
    "fnet_feed_batch_u8": (C.c_int, [_I32, _I64, _I64, _P, _P, _F, _P, _P, _P, C.c_int, _F, _F, _P]),
    "fnet_graph_destroy": (C.c_int, [_P]),
    "fnet_graph_step_fed": (C.c_int, [_P, _P, _P, C.c_size_t, _P, _P, C.c_size_t, _P, _P, C.c_size_t, _P]),
    "fnet_conv2d_bwd_data_on_tensor_core": (C.c_int, [_CD]),
    "fnet_dense_bwd_data_on_tensor_core": (C.c_int, [C.c_int64, C.c_int64, C.c_int64, C.c_int]),
    "fnet_graph_step_streamed": (C.c_int, [_P, _P, _P, _P, C.c_size_t, _P, _P, _P, C.c_size_t, _P, _P, _P, C.c_size_t,
                                           _P, _P, _P, _P]),
}

_QUERIES = {"fnet_version", "fnet_conv2d_bwd_data_on_tensor_core", "fnet_dense_bwd_data_on_tensor_core",
            "fnet_fused_step_grid", "fnet_fused_conv_supported", "fnet_fused_head_supported"}    # return a value, not a status
_STATUS_RETURNING = {n for n, (r, _) in SIGNATURES.items() if r is C.c_int and n not in _QUERIES}


def build(verbose=False):
    """Compile libfnet.so in tree (nvcc, -gencode arch=compute_100a,code=sm_100a, -lineinfo)."""
    out = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
    if verbose or out.returncode != 0:
        print(out.stdout)
        print(out.stderr)
    if out.returncode != 0:
        raise FnetError("building libfnet.so failed (see output above)")
    return LIB_PATH


class _Lib(object):
    """Loaded library; attribute access returns a checked wrapper that raises on a non-zero status."""

    def __init__(self, path):
        if not os.path.isfile(path):
            raise FnetError("%s not found: run `python -c 'import __graft_entry__ as g; g.build()'` "
                            "(there is no CPU fallback)" % path)
        self._dll = C.CDLL(path)
        self.path = path
        self.launches = 0            # number of C-ABI calls that enqueue GPU work (a claim for bench.py)
        for name, (restype, argtypes) in SIGNATURES.items():
            fn = getattr(self._dll, name)          # AttributeError here = header / library mismatch
            fn.restype = restype
            fn.argtypes = argtypes
            setattr(self, name, self._checked(name, fn) if name in _STATUS_RETURNING else fn)

    def _checked(self, name, fn):
        def call(*args):
            rc = fn(*args)
            if rc != 0:
                raise FnetError("%s failed (%d): %s" % (name, rc, self._dll.fnet_last_error().decode()))
            self.launches += 1
            return rc
        call.__name__ = name
        return call


_lib = None


def get():
    global _lib
    if _lib is None:
        _lib = _Lib(LIB_PATH)
        if _lib.fnet_version() != 100:
            raise FnetError("libfnet.so version mismatch")
    return _lib
