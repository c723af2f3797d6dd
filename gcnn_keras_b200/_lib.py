"""ctypes binding of libkgcnn_b200.so (the C-ABI declared in include/kgcnn_b200.h).

There is no CPU fallback: if the shared library is missing, importing any compute path raises.
Build it with ``python -m gcnn_keras_b200.build`` (or ``__graft_entry__.build()``).
"""
import ctypes
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libkgcnn_b200.so")
HEADER_PATH = os.path.join(_HERE, "..", "include", "kgcnn_b200.h")

_p = ctypes.c_void_p
_i = ctypes.c_int
_l = ctypes.c_int64
_f = ctypes.c_float
_d = ctypes.c_double
_z = ctypes.c_size_t

# name -> (restype, argtypes); mirrors include/kgcnn_b200.h one to one (tests/test_abi.py checks it)
SIGNATURES = {
    "kgcnn_version": (_i, []),
    "kgcnn_error_string": (ctypes.c_char_p, [_i]),
    "kgcnn_launch_count": (ctypes.c_longlong, []),
    "kgcnn_launch_count_reset": (None, []),
    "kgcnn_shift_index": (_i, [_p, _l, _i, _p, _p, _l, _p, _p, _p]),
    "kgcnn_value_rowids": (_i, [_p, _l, _l, _p, _p]),
    "kgcnn_cast_i64_i32": (_i, [_p, _l, _p, _p]),
    "kgcnn_pad": (_i, [_p, _p, _l, _l, _l, ctypes.c_uint32, ctypes.c_uint32, _i, _p, _p, _i, _p]),
    "kgcnn_unpad": (_i, [_p, _p, _l, _l, _l, _l, _p, _p]),
    "kgcnn_check_sorted": (_i, [_p, _i, _l, _l, _p, _i, _p]),
    "kgcnn_plan_workspace_bytes": (_z, [_l, _l]),
    "kgcnn_plan_build": (_i, [_p, _i, _l, _l, _i, _p, _p, _p, _z, _p]),
    "kgcnn_permute_i32": (_i, [_p, _i, _p, _l, _p, _p]),
    "kgcnn_gather_rows": (_i, [_p, _l, _i, _p, _i, _l, _p, _l, _p]),
    "kgcnn_gather_rows_i64": (_i, [_p, _l, _i, _p, _i, _l, _p, _l, _p]),
    "kgcnn_gather_mul_segment_reduce": (_i, [_p, _i, _p, _i, _p, _p, _p, _i, _i, _l, _l, _i, _p, _p, _p, _i, _i, _p]),
    "kgcnn_row_stats": (_i, [_p, _l, _p, _p]),
    "kgcnn_long_rows_build": (_i, [_p, _l, _i, _i, _p, _p, _p]),
    "kgcnn_gather2_mul": (_i, [_p, _p, _i, _p, _p, _i, _i, _l, _i, _p, _p]),
    "kgcnn_bias_act": (_i, [_p, _p, _p, _p, _i, _i, _l, _i, _p, _p]),
    "kgcnn_bias_act_pair": (_i, [_p, _p, _p, _p, _i, _i, _l, _i, _p, _p, _p]),
    "kgcnn_colsum_workspace_bytes": (_z, [_l, _i]),
    "kgcnn_colsum": (_i, [_p, _l, _i, _p, _p, _z, _p]),
    "kgcnn_colsum_static": (_i, [_p, _l, _i, _p, _p]),
    "kgcnn_dense_tc_workspace_bytes": (_z, [_i, _l, _i, _i]),
    "kgcnn_dense_tc": (_i, [_i, _p, _p, _l, _i, _i, _p, _p, _p, _z, _p]),
    "kgcnn_dense_act": (_i, [_p, _p, _l, _i, _i, _p, _i, _p, _p, _p]),
    "kgcnn_dense_tn_grouped_workspace_bytes": (_z, [_i, _p, _p, _p]),
    "kgcnn_dense_tn_grouped": (_i, [_i, _p, _p, _p, _p, _p, _p, _p, _p, _z, _p]),
    "kgcnn_dense_rel_workspace_bytes": (_z, [_i, _i, _i, _i]),
    "kgcnn_dense_rel": (_i, [_i, _p, _p, _p, _i, _l, _i, _i, _p, _p, _p, _z, _p]),
    "kgcnn_edge_geometry": (_i, [_p, _p, _p, _l, _i, _i, _i, _p, _p, _f, _i, _f, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "kgcnn_bessel_frequency_grad_workspace_bytes": (_z, [_i]),
    "kgcnn_bessel_frequency_grad": (_i, [_p, _p, _p, _l, _i, _p, _p, _f, _p, _p, _p, _p, _z, _p]),
    "kgcnn_acsf_g2": (_i, [_p, _p, _p, _p, _p, _p, _i, _p, _i, _i, _f, _l, _p, _i, _p, _p, _p, _p]),
    "kgcnn_acsf_g4": (_i, [_p, _p, _p, _p, _i, _p, _p, _p, _p, _p, _i, _i, _f, _f, _i, _l, _p, _p, _p, _p, _l, _i, _p, _p, _p, _p]),
    "kgcnn_painn_msg_nodes": (_i, [_p, _p, _p, _p, _p, _p, _p, _i, _l, _l, _i, _i, _p, _p, _p]),
    "kgcnn_painn_msg_senders": (_i, [_p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _l, _l, _i, _i, _p, _p, _p]),
    "kgcnn_painn_msg_edges": (_i, [_p, _p, _p, _p, _p, _p, _p, _p, _l, _i, _i, _p, _p, _p, _p]),
    "kgcnn_offsets_workspace_bytes": (_z, [_l]),
    "kgcnn_offsets_from_counts": (_i, [_p, _l, _p, _p, _z, _p]),
    "kgcnn_angle_count": (_i, [_p, _l, _l, _p, _p, _i, _i, _i, _i, _i, _p, _p]),
    "kgcnn_angle_emit": (_i, [_p, _p, _l, _l, _p, _p, _i, _i, _i, _i, _i, _p, _p, _l, _p, _p, _p, _p]),
    "kgcnn_range_count": (_i, [_p, _i, _p, _l, _l, _i, _i, _d, _i, _i, _i, _p, _p]),
    "kgcnn_range_emit": (_i, [_p, _i, _p, _l, _l, _i, _i, _d, _i, _i, _i, _p, _p, _p, _p, _p]),
    "kgcnn_segment_softmax": (_i, [_p, _p, _p, _l, _l, _i, _i, _p, _p]),
    "kgcnn_cfconv_fused_workspace_bytes": (_z, [_l]),
    "kgcnn_gauss_basis": (_i, [_p, _l, _i, _f, _f, _f, _i, _p, _p, _p]),
    "kgcnn_cfconv_fused": (_i, [_i, _p, _l, _i, _f, _f, _f, _p, _p, _p, _p, _i, _i, _i, _p, _p, _p, _p, _l, _p, _p, _p, _p, _p, _p, _p, _p, _z, _p]),
    "kgcnn_allreduce_peer": (_i, [_p, _i, _l, _f, _p, _p]),
    "kgcnn_guard_fold": (_i, [_p, _p, _i, _p, _p]),
    "kgcnn_adam_clipnorm": (_i, [_p, _p, _p, _p, _p, _p, _l, _f, _p, _f, _f, _f, _f, _p, _p, _p, _i, _p]),
}

OP_SUM, OP_MEAN, OP_MAX, OP_MIN = 0, 1, 2, 3
BASIS_NONE, BASIS_GAUSS, BASIS_BESSEL = 0, 1, 2
ACT_CODES = {"linear": 0, "shifted_softplus": 1, "swish": 2, "tanh": 3, "relu": 4, "softplus2": 5, "sigmoid": 6, "leaky_relu": 7}

_lib = None


class KgcnnError(RuntimeError):
    pass


def header_symbols():
    """Function names declared in include/kgcnn_b200.h."""
    with open(HEADER_PATH) as fh:
        text = fh.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(kgcnn_[a-z0-9_]+)\s*\(", text)))


def lib():
    """Load (once) and return the ctypes handle; raises if the CUDA library was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise KgcnnError(
                "libkgcnn_b200.so not found at %s. The CUDA extension is required (there is no CPU "
                "fallback); build it with `python -m gcnn_keras_b200.build`." % LIB_PATH)
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(code, what=""):
    if code != 0:
        msg = lib().kgcnn_error_string(code).decode()
        raise KgcnnError("%s failed: %s (%d)" % (what or "kgcnn call", msg, code))


def launch_count():
    return int(lib().kgcnn_launch_count())


def launch_count_reset():
    lib().kgcnn_launch_count_reset()
