"""ctypes binding of the native sm_100a library (include/dvc_b200.h).

The library is built in-tree by ``ops.build`` (nvcc, -gencode arch=compute_100a,code=sm_100a).
On a machine with a GPU a missing/unloadable library is a hard error: GPU paths never fall
back silently to PyTorch ops.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from . import build as _build

_LIB: Optional[C.CDLL] = None


class NativeError(RuntimeError):
    pass


class HeadOut(C.Structure):
    _fields_ = [("loc", C.c_void_p), ("conf", C.c_void_p), ("loc_pitch", C.c_longlong), ("conf_pitch", C.c_longlong),
                ("hw", C.c_int), ("n_loc", C.c_int), ("n_conf", C.c_int), ("loc_off", C.c_longlong),
                ("conf_off", C.c_longlong)]


class DetParams(C.Structure):
    _fields_ = [("num_priors", C.c_int), ("num_classes", C.c_int), ("background", C.c_int),
                ("conf_thresh", C.c_float), ("nms_thresh", C.c_float), ("top_k", C.c_int), ("keep_top_k", C.c_int),
                ("app_conf", C.c_float), ("consider_mask", C.c_uint32), ("out_w", C.c_int), ("out_h", C.c_int)]


class MbxDesc(C.Structure):
    """dvc_mbx_desc (include/dvc_b200.h): one dispatch, written by the client, polled by the worker lane."""
    _fields_ = [("seq", C.c_uint32), ("epoch", C.c_uint32), ("n_valid", C.c_uint32), ("kind", C.c_uint32),
                ("src_off", C.c_uint64), ("batch_id", C.c_uint32), ("reserved", C.c_uint32)]


class MbxLaneCtl(C.Structure):
    """dvc_mbx_lane_ctl: worker-local control block of one lane (device memory)."""
    _fields_ = [("src", C.c_void_p), ("gate_now", C.c_uint32), ("gate_expect", C.c_uint32), ("flag_value", C.c_uint32),
                ("status", C.c_uint32), ("batch_id", C.c_uint32), ("n_valid", C.c_uint32), ("next_seq", C.c_uint32),
                ("reserved", C.c_uint32)]


class MbxHostStatus(C.Structure):
    """dvc_mbx_host_status: what a wait kernel saw, in pinned host memory."""
    _fields_ = [("status", C.c_uint32), ("seq", C.c_uint32), ("batch_id", C.c_uint32), ("n_valid", C.c_uint32),
                ("wait_ns", C.c_uint64)]


MBX_NOTHING, MBX_GO, MBX_STOP, MBX_SEQ_STOP, MBX_MAX_GROUP = 0, 1, 2, 0xFFFFFFFF, 16

_P, _I, _LL, _F, _U32, _SZ = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_uint32, C.c_size_t

_SIGS = {
    "dvc_version": ([], _I),
    "dvc_last_error": ([], C.c_char_p),
    "dvc_launch_count": ([], _LL),
    "dvc_set_device": ([_I], _I),
    "dvc_resize_u8": ([_P, _LL, _I, _I, _I, _P, _LL, _I, _I, _I, _P, _P, _P, _I, _P, _P, _P, _I, _P, _P, _I, _I, _P], _I),
    "dvc_resize_u8_gated": ([_P, _LL, _I, _I, _I, _P, _LL, _I, _I, _I, _P, _P, _P, _I, _P, _P, _P, _I, _P, _P, _I, _I, _P, _P, _P], _I),
    "dvc_preproc_conv0": ([_P, _LL, _P, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _F, _P, _P], _I),
    "dvc_dwconv3x3": ([_P, _I, _I, _I, _I, _I, _P, _P, _I, _P, _P], _I),
    "dvc_gemm_f16": ([_P, _P, _P, _I, _I, _I, _I, _I, _P, _I, _I, _P, _I, _P], _I),
    "dvc_im2col3x3s2": ([_P, _I, _I, _I, _I, _P, _P], _I),
    "dvc_conv3x3s2_f16": ([_P, _I, _I, _I, _I, _P, _P, _I, _P, _I, _I, _I, _P], _I),
    "dvc_debug_im2col_tile": ([_P, _I, _I, _I, _I, _I, _I, _P, _P, _P], _I),
    "dvc_priorbox": ([_P, _P, _I, _I, _I, _I, _I, _P, _I, _F, _P, _I, _P], _I),
    "dvc_detection_scratch_bytes": ([_I, _I, _I], _SZ),
    "dvc_detection_output": ([_P, _P, _P, _P, _I, _P, _P, _P, _P, _P, _P, _P, _U32, _P, _SZ, _P], _I),
    "dvc_detection_output_static": ([_P, _P, _P, _P, _I, _P, _P, _P, _P, _P, _P, _P, _U32, _P], _I),
    "dvc_annotate": ([_P, _LL, _P, _I, _I, _I, _P, _P, _P, _P, _I, _I, _I, _I, _P, _I, _I, _I, _I, _I, _P], _I),
    "dvc_plan_create": ([], _P),
    "dvc_plan_destroy": ([_P], None),
    "dvc_plan_num_ops": ([_P], _I),
    "dvc_plan_add_preproc_conv0": ([_P, _P, _LL, _P, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _F, _P, _I], _I),
    "dvc_plan_add_dwconv": ([_P, _P, _I, _I, _I, _I, _I, _P, _P, _I, _P], _I),
    "dvc_plan_add_gemm": ([_P, _P, _P, _P, _I, _I, _I, _I, _I, _P, _I, _I, _P, _I], _I),
    "dvc_plan_add_im2col": ([_P, _P, _I, _I, _I, _I, _P], _I),
    "dvc_plan_add_conv3x3s2": ([_P, _P, _I, _I, _I, _I, _P, _P, _I, _P, _I, _I, _I], _I),
    "dvc_plan_add_detection": ([_P, _P, _P, _P, _P, _I, _P, _P, _P, _P, _P, _P, _P, _U32, _P, _SZ], _I),
    "dvc_plan_set_source": ([_P, _P, _LL, _P], _I),
    "dvc_plan_set_sinks": ([_P, _P, _P, _P, _P, _P, _P], _I),
    "dvc_plan_set_flag_value": ([_P, _U32], _I),
    "dvc_plan_set_epoch_gate": ([_P, _P, _P, _P], _I),
    "dvc_plan_set_timestamps": ([_P, _P, _P], _I),
    "dvc_plan_add_dwpw_fused": ([_P, _P, _I, _I, _I, _I, _P, _P, _P, _P, _I, _P], _I),
    "dvc_dwpw_fused_s1": ([_P, _I, _I, _I, _I, _P, _P, _P, _P, _I, _P, _P], _I),
    "dvc_dwpw_fused": ([_P, _I, _I, _I, _I, _I, _P, _P, _P, _P, _I, _P, _P], _I),
    "dvc_plan_run": ([_P, _P], _I),
    "dvc_plan_set_head_branch": ([_P, _I], _I),
    "dvc_p2p_alloc": ([_SZ, C.POINTER(C.c_void_p)], _I),
    "dvc_p2p_free": ([_P], _I),
    "dvc_p2p_get_handle": ([_P, _P], _I),
    "dvc_p2p_open": ([_P, C.POINTER(C.c_void_p)], _I),
    "dvc_p2p_close": ([_P], _I),
    "dvc_p2p_enable_peer": ([_I], _I),
    "dvc_epoch_broadcast": ([_P, _I, _U32, _P], _I),
    "dvc_push_multicast": ([_P, _P, _I, _SZ, _P], _I),
    "dvc_peer_read": ([_P, _SZ, _P, _P], _I),
    "dvc_mbx_post": ([_P, _P, _I, _P], _I),
    "dvc_mbx_wait_desc": ([_P, _U32, _P, _P, _P, _P, C.c_ulonglong, _P], _I),
    "dvc_mbx_wait_done": ([_P, _P, _I, _P, C.c_ulonglong, _P], _I),
    "dvc_mbx_dispatch": ([_P, _P, _SZ, _P, _P, _P, _P, C.c_ulonglong, _P, _P, _SZ, _P, _P], _I),
    "dvc_mbx_signal": ([_P, _U32, _P], _I),
    "dvc_mbx_poll": ([_P, _P, _I, _P], _I),
    "dvc_mbx_worker_enqueue": ([_P, _U32, _P, _P, _P, _P, C.c_ulonglong, _P, _P, _P], _I),
    "dvc_mbx_local_dispatch": ([_P, _P, _P, _P, _P, _P, _P, _P, C.c_ulonglong, _P, _P, _SZ, _P, _P], _I),
    "dvc_mbx_copy_out": ([_P, _P, _SZ, _P, _P], _I),
    "dvc_plan_set_mailbox": ([_P, _P, _P, _P, _P, _P], _I),
    "dvc_resize_u8_indirect": ([_P, _LL, _I, _I, _I, _P, _LL, _I, _I, _I, _P, _P, _P, _I, _P, _P, _P, _I, _P, _P, _I, _I, _P, _P, _P], _I),
}


def lib_path() -> str:
    return _build.LIB_PATH


def load(build_if_missing: bool = True) -> C.CDLL:
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        if not build_if_missing:
            raise NativeError(f"{path} is missing: run python -m distributedvolunteercomputing_b200.ops.build")
        _build.build()
    lib = C.CDLL(path)
    for name, (args, res) in _SIGS.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.argtypes = args
        fn.restype = res
    _LIB = lib
    return lib


def available() -> bool:
    """True when the native kernels can run here (library loadable AND a CUDA device present)."""
    if not torch.cuda.is_available():
        return False
    load()
    return True


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().dvc_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what or 'dvc'} failed (rc={rc}): {msg}")


def ptr(t) -> Optional[int]:
    if t is None:
        return None
    if isinstance(t, int):
        return t
    return t.data_ptr()


def set_device(device) -> None:
    """Make `device` current in the library's own CUDA runtime (see dvc_set_device)."""
    idx = torch.device(device).index
    check(load().dvc_set_device(0 if idx is None else idx), "set_device")


def stream_ptr(stream=None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def launch_count() -> int:
    return int(load().dvc_launch_count()) if _LIB is not None or os.path.exists(lib_path()) else 0
