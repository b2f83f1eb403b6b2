"""ctypes binding of libtdssf.so (include/tdssf.h).  There is no CPU or PyTorch fallback: if the
library is missing, importing this module raises."""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("TDSSF_LIB") or os.path.join(_PKG, "libtdssf.so")  # TDSSF_LIB: experiment variants

OK, EINVAL, ECUDA, ENOSUP, EDEVICE = 0, -1, -2, -3, -4
F32, F16, BF16 = 0, 1, 2
MAX_CAMS = 6
REC_BYTES = 16


class MlpWeights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("W1", "b1", "W2", "b2", "W3", "b3")]


class MlpGrads(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("dW1", "db1", "dW2", "db2", "dW3", "db3")]


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `python 3d-semantic-segmentation-fusion_b200/build.py` "
        "(nvcc, sm_100a).  tdssf has no CPU fallback.")

lib = C.CDLL(LIB_PATH)

_p, _i, _i64, _sz, _f, _d = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float, C.c_double

# name -> (restype, argtypes); kept in one table so tests can check it against include/tdssf.h
SIGNATURES = {
    "tdssf_version": (_i, []),
    "tdssf_last_error": (C.c_char_p, []),
    "tdssf_check_device": (_i, []),
    "tdssf_project": (_i, [_p, _i64, _i64, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p]),
    "tdssf_scale_pixels": (_i, [_p, _p, _i, _i, _i64, _f, _f, _p, _p]),
    "tdssf_gather_mean": (_i, [_p, _i, _p, _i, _i, _i, _i, _i, _p, _p]),
    "tdssf_mlp_fwd_f32": (_i, [_p, _p, _i64, C.POINTER(MlpWeights), _p, _p, _p, _p]),
    "tdssf_mlp_bwd_workspace_bytes": (_sz, [_i64]),
    "tdssf_mlp_bwd_f32": (_i, [_p, _p, _p, _p, _p, _i64, C.POINTER(MlpWeights), C.POINTER(MlpGrads), _p, _p, _sz, _p]),
    "tdssf_mlp_pack_bytes": (_sz, []),
    "tdssf_mlp_pack_weights": (_i, [C.POINTER(MlpWeights), _p, _p]),
    "tdssf_fusion_workspace_bytes": (_sz, [_i, _i, _i]),
    "tdssf_fusion_fwd": (_i, [_p, _i64, _i64, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "tdssf_mlp_fwd_tc": (_i, [_p, _p, _i64, _p, _p, _p, _p, _p]),
    "tdssf_mlp_bwd_tc_workspace_bytes": (_sz, [_i64]),
    "tdssf_mlp_bwd_tc": (_i, [_p, _p, _p, _p, _p, _p, _i64, C.POINTER(MlpWeights), C.POINTER(MlpGrads), _p, _p, _sz, _p]),
    "tdssf_voxelize_workspace_bytes": (_sz, [_i, _i]),
    "tdssf_voxelize": (_i, [_p, _i64, _i, _i, _i, _d, _d, _p, _p, _p, _p, _p, _sz, _p]),
    "tdssf_devox_nn_workspace_bytes": (_sz, [_i, _i, _i]),
    "tdssf_devox_nn": (_i, [_p, _i64, _p, _i64, _i, _i, _i, _p, _f, _p, _sz, _p]),
    "tdssf_devox_nn_static": (_i, [_p, _i64, _p, _i64, _i, _i, _i, _p, _f, _p]),
    "tdssf_devox_gather": (_i, [_p, _p, _i, _i, _i, _i, _p, _p]),
    "tdssf_devox_scatter_add": (_i, [_p, _p, _i, _i, _i, _i, _p, _p]),
    "tdssf_cels_workspace_bytes": (_sz, [_i64]),
    "tdssf_cels_fwd_bwd": (_i, [_p, _p, _p, _i64, _p, _f, _i, _p, C.POINTER(C.c_int64), _p, _p, _p, _p, _sz, _p]),
    "tdssf_cels_count": (_i, [_p, _p, _i64, _i, _p, _p, _sz, _p]),
    "tdssf_cels_fwd_bwd_counted": (_i, [_p, _p, _p, _i64, _p, _f, _i, _p, _p, _p, _p, _p, _sz, _p]),
    "tdssf_cels_graph_stats": (None, [C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "tdssf_cels_backward": (_i, [_p, _sz, _i64, _i64, _p, _p, _p, _p]),
    "tdssf_confusion": (_i, [_p, _p, _i64, _i, _p, _p]),
    "tdssf_p2p_buffer_bytes": (_sz, [_i64]),
    "tdssf_p2p_create": (_i, [_i64, C.POINTER(C.c_void_p), _p]),
    "tdssf_p2p_open": (_i, [_p, C.POINTER(C.c_void_p)]),
    "tdssf_p2p_close": (_i, [_p]),
    "tdssf_p2p_destroy": (_i, [_p]),
    "tdssf_p2p_allreduce_mean": (_i, [C.POINTER(C.c_void_p), _i, _i, _p, _i64, _i64, C.c_uint32, _p]),
    "tdssf_compose_lidar2cam": (_i, [_p, _p, _p, _p, _i, _i, _p, _p, _p]),
    "tdssf_compose_sweep2ref": (_i, [_p, _p, _p, _p, _i, _p, _p]),
    "tdssf_accumulate_sweeps": (_i, [_p, _i64, _p, _p, _i, _i, _p, _p, _p]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here = header and library out of sync
    _fn.restype = _res
    _fn.argtypes = _args


def last_error() -> str:
    return (lib.tdssf_last_error() or b"").decode("utf-8", "replace")


def check(rc: int, what: str = "tdssf") -> None:
    """Error convention of the boundary: EINVAL -> ValueError (the reference raises ValueError for shape
    mismatches, utils/fusion_model.py:45,53,115); everything else -> RuntimeError."""
    if rc == OK:
        return
    msg = f"{what}: {last_error()} (rc={rc})"
    if rc == EINVAL:
        raise ValueError(msg)
    raise RuntimeError(msg)
