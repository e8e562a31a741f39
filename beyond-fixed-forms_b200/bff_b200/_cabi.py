"""ctypes binding of libbff.so (include/bff.h). No pybind, no torch extension: the library is
a plain C ABI and is called with raw device pointers (`tensor.data_ptr()`) and the current
CUDA stream. There is NO CPU fallback: if the library is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BFF_LIB_PATH") or os.path.join(_HERE, "libbff.so")   # override: experiment builds only

BFF_DEPTH_F32, BFF_DEPTH_U16, BFF_DEPTH_U16_BILINEAR = 0, 1, 2
BFF_XYZ_F32X4, BFF_XYZ_F64X3 = 0, 1


class BffError(RuntimeError):
    pass


class LiftDesc(ctypes.Structure):
    """struct bff_lift_desc (include/bff.h)."""
    _fields_ = [
        ("xyz", c_void_p), ("xyz_kind", c_int32), ("n_points", c_int64),
        ("pose_inv", c_void_p), ("intrinsics", c_void_p), ("n_frames", c_int32),
        ("depth", c_void_p), ("depth_kind", c_int32), ("depth_scale", c_float),
        ("height", c_int32), ("width", c_int32), ("depth_thresh", c_double),
        ("depth_min", c_double), ("depth_max", c_double),
        ("masks2d", c_void_p), ("mask_offset", c_void_p), ("mask_query", c_void_p),
        ("n_masks", c_int32), ("n_queries", c_int32), ("frame_viewed", c_void_p),
        ("rows", c_void_p), ("masked_counts", c_void_p), ("viewed_counts", c_void_p),
        ("debug_pix", c_void_p), ("debug_vis", c_void_p),
        ("frames_per_chunk", c_int32), ("reserved", c_int32),
        ("src_height", c_int32), ("src_width", c_int32), ("resize_x", c_void_p), ("resize_y", c_void_p),
        ("ev_start", c_void_p), ("ev_stop", c_void_p),
        ("n_row_slices", c_int32), ("reserved2", c_int32), ("row_slice_words", c_int64),
        ("peer_rows", c_void_p * 8), ("peer_row0", c_int64),
    ]


class Stage2Desc(ctypes.Structure):
    """struct bff_stage2_desc (include/bff.h)."""
    _fields_ = [
        ("rows", c_void_p), ("n_rows", c_int32), ("n_points", c_int64),
        ("conf", c_void_p), ("label_id", c_void_p), ("mc", c_void_p), ("vc", c_void_p),
        ("iou_thres", c_float), ("min_aggregated", c_int32), ("filter_mode", c_int32), ("filter_fraction", c_double),
        ("agg_rows", c_void_p), ("cap_groups", c_int32), ("n_groups", c_void_p), ("group_first", c_void_p),
        ("group_size", c_void_p), ("group_conf", c_void_p), ("before", c_void_p), ("after", c_void_p),
        ("members", c_void_p), ("filter_value", c_void_p), ("inter_pre", c_void_p), ("pop_pre", c_void_p),
    ]


class RefineDesc(ctypes.Structure):
    """struct bff_refine_desc (include/bff.h)."""
    _fields_ = [
        ("s1_rows", c_void_p), ("n1", c_int32), ("s2_rows", c_void_p), ("n2", c_int32), ("n_points", c_int64),
        ("s2_conf", c_void_p), ("stage1_iou_thres", c_float),
        ("s2_out", c_void_p), ("n2_out", c_void_p), ("s2_conf_out", c_void_p), ("match", c_void_p),
        ("match_iou", c_void_p), ("best_match_log", c_void_p),
    ]


class Stage2DevDesc(ctypes.Structure):
    """struct bff_stage2_dev_desc (include/bff.h)."""
    _fields_ = [
        ("rows", c_void_p), ("n_rows", c_int32), ("n_points", c_int64),
        ("conf", c_void_p), ("label_id", c_void_p), ("mc", c_void_p), ("vc", c_void_p),
        ("n_vote_points", c_int64), ("vote_point0", c_int64),
        ("iou_thres", c_float), ("min_aggregated", c_int32), ("filter_mode", c_int32), ("filter_fraction", c_double),
        ("small_is_float", c_int32), ("ratio_is_float", c_int32), ("small_i", c_int64), ("ratio_i", c_int64),
        ("small_f", c_float), ("ratio_f", c_float), ("mc_cap", c_int32), ("vc_cap", c_int32),
        ("out_rows", c_void_p), ("cap_groups", c_int32), ("result", c_void_p), ("comm", c_void_p),
    ]


class RefineDevDesc(ctypes.Structure):
    """struct bff_refine_dev_desc (include/bff.h)."""
    _fields_ = [
        ("s1_rows", c_void_p), ("n1", c_int32), ("s2_rows", c_void_p), ("n2_dev", c_void_p), ("n2_host", c_int32),
        ("cap2", c_int32), ("n_points", c_int64), ("s2_conf", c_void_p), ("stage1_iou_thres", c_float),
        ("is_query", c_void_p), ("sim", c_void_p), ("select_mode", c_int32), ("sim_percentile", c_double),
        ("sim_thres", c_double), ("refine_iou_thres", c_float),
        ("s2_out", c_void_p), ("final_rows", c_void_p), ("result", c_void_p), ("comm", c_void_p),
    ]


# word offsets of the device result blocks (include/bff.h enums)
S2R_N_GROUPS, S2R_N_SELECTED, S2R_N_OVERLAP, S2R_N_UNIQUE, S2R_ERROR, S2R_FILTER_VALUE, S2R_MC_MAX, S2R_VC_MAX, S2R_HEADER = 0, 1, 2, 3, 4, 5, 6, 7, 16
RFR_N2_IN, RFR_N2_OUT, RFR_N_OR, RFR_N_OTHER, RFR_N_FINAL, RFR_ERROR, RFR_SIM_THRES, RFR_HEADER = 0, 1, 2, 3, 4, 5, 6, 16


def stage2_dev_result_words(n_rows: int, cap: int) -> int:
    return S2R_HEADER + 7 * cap + (cap + 1) + n_rows


def refine_dev_result_words(n1: int, cap2: int) -> int:
    return RFR_HEADER + 8 * cap2 + (cap2 + 1) + n1 + 2 * (n1 + cap2)


# name -> (restype, argtypes); every symbol include/bff.h declares
VOTE_UNIQUE_STATIC_MAX_BINS = 131072       # include/bff.h BFF_VOTE_UNIQUE_STATIC_MAX_BINS

SYMBOLS = {
    "bff_version": (c_int, []),
    "bff_last_error": (c_char_p, []),
    "bff_lift_scene": (c_int, [POINTER(LiftDesc), c_void_p]),
    "bff_lift_scene_launches": (c_int, [POINTER(LiftDesc)]),
    "bff_resize_table": (c_int, [c_int32, c_int32, c_void_p]),
    "bff_depth_resize": (c_int, [c_void_p, c_int32, c_int32, c_int32, c_float, c_void_p, c_void_p, c_int32, c_int32,
                                 c_void_p, c_void_p]),
    "bff_visibility": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int32, c_int32, c_double, c_void_p, c_void_p]),
    "bff_mask_gather": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int32, c_int32, c_int32, c_void_p, c_void_p]),
    "bff_mask_iou": (c_int, [c_void_p, c_int32, c_void_p, c_int32, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "bff_set_iou_path": (c_int, [c_int32]),
    "bff_rows_or_reduce": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int32, c_void_p, c_void_p]),
    "bff_counts_max": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "bff_vote_hist": (c_int, [c_void_p, c_void_p, c_int64, c_int32, c_int32, c_void_p, c_void_p]),
    "bff_vote_unique_table_size": (c_int64, [c_int32, c_int32]),
    "bff_vote_unique": (c_int, [c_void_p, c_int32, c_int32, c_int32, c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
    "bff_vote_unique_static": (c_int, [c_void_p, c_int32, c_int32, c_int32, c_void_p, c_void_p, c_void_p]),
    "bff_point_keep": (c_int, [c_void_p, c_void_p, c_int64, c_int32, c_float, c_void_p, c_void_p]),
    "bff_labels_to_rows": (c_int, [c_void_p, c_int64, c_void_p, c_int32, c_void_p, c_void_p]),
    "bff_rows_permute": (c_int, [c_void_p, c_int32, c_int64, c_void_p, c_void_p, c_void_p]),
    "bff_overlap_resolve": (c_int, [c_void_p, c_int32, c_int64, c_void_p, c_int32, c_void_p]),
    "bff_rows_or_into": (c_int, [c_void_p, c_int32, c_int64, c_void_p, c_int32, c_void_p]),
    "bff_rows_and_popcount": (c_int, [c_void_p, c_int32, c_int64, c_void_p, c_void_p, c_void_p]),
    "bff_pack_bits": (c_int, [c_void_p, c_int32, c_int64, c_void_p, c_void_p]),
    "bff_unpack_bits": (c_int, [c_void_p, c_int32, c_int64, c_void_p, c_void_p]),
    "bff_rle_to_bits": (c_int, [c_void_p, c_int64, c_int32, c_int64, c_void_p, c_void_p]),
    "bff_rle_pairs_to_bits": (c_int, [c_void_p, c_void_p, c_int32, c_int64, c_int64, c_void_p, c_void_p]),
    "bff_stage2_aggregate": (c_int, [c_void_p, POINTER(Stage2Desc), c_void_p]),
    "bff_stage2_aggregate_batch": (c_int, [c_void_p, POINTER(Stage2Desc), c_int32, c_void_p]),
    "bff_refine_match": (c_int, [c_void_p, POINTER(RefineDesc), c_void_p]),
    "bff_refine_match_batch": (c_int, [c_void_p, POINTER(RefineDesc), c_int32, c_void_p]),
    "bff_upload": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "bff_rows_gather": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int32, c_void_p, c_void_p]),
    "bff_stage2_device": (c_int, [c_void_p, POINTER(Stage2DevDesc), c_void_p]),
    "bff_refine_device": (c_int, [c_void_p, POINTER(RefineDevDesc), c_void_p]),
    "bff_refine_select_device": (c_int, [c_void_p, POINTER(RefineDevDesc), c_double, c_void_p]),
    "bff_rows_gather_dev": (c_int, [c_void_p, c_int64, c_void_p, c_int32, c_void_p, c_void_p]),
    "bff_unpack_bits_dc": (c_int, [c_void_p, c_void_p, c_int32, c_int64, c_void_p, c_void_p]),
    "bff_comm_unique_id": (c_int, [c_void_p]),
    "bff_comm_create": (c_int, [c_int, c_int, c_int, c_void_p, POINTER(c_void_p)]),
    "bff_comm_destroy": (None, [c_void_p]),
    "bff_comm_rank": (c_int, [c_void_p]),
    "bff_comm_world": (c_int, [c_void_p]),
    "bff_comm_allreduce_i32": (c_int, [c_void_p, c_void_p, c_int64, c_int32, c_void_p]),
    "bff_comm_allgather_bytes": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "bff_comm_exchange_rows": (c_int, [c_void_p, c_void_p, c_int32, c_void_p, c_int64, c_void_p, c_void_p]),
    "bff_ipc_alloc": (c_int, [c_int64, POINTER(c_void_p), c_void_p]),
    "bff_ipc_open": (c_int, [c_void_p, POINTER(c_void_p)]),
    "bff_ipc_close": (c_int, [c_void_p]),
    "bff_ipc_free": (c_int, [c_void_p]),
    "bff_rows_slice": (c_int, [c_void_p, c_int32, c_int64, c_int64, c_int64, c_void_p, c_void_p]),
    "bff_rows_unslice": (c_int, [c_void_p, c_int32, c_int32, c_int64, c_int64, c_void_p, c_void_p]),
    "bff_ctx_create": (c_int, [c_int, POINTER(c_void_p)]),
    "bff_ctx_destroy": (None, [c_void_p]),
    "bff_lift_scene_host": (c_int, [c_void_p, POINTER(LiftDesc), POINTER(c_int64), POINTER(c_int64)]),
    "bff_mask_iou_host": (c_int, [c_void_p, c_void_p, c_int32, c_void_p, c_int32, c_int64, c_void_p, c_void_p,
                                  c_void_p, c_void_p, POINTER(c_int64), POINTER(c_int64)]),
}

_lib = None
launch_count = 0   # kernels launched through this binding (bench.py's gpu_launches)


def lib():
    """Load libbff.so or raise. There is deliberately no fallback implementation."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise BffError(
                f"{LIB_PATH} not found: build it with `python beyond-fixed-forms_b200/build.py` "
                "(nvcc, sm_100a). bff_b200 has no CPU or PyTorch fallback.")
        h = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(h, name)
            fn.restype = res
            fn.argtypes = args
        _lib = h
    return _lib


def check(rc: int, what: str):
    if rc != 0:
        msg = lib().bff_last_error().decode(errors="replace")
        raise BffError(f"{what} failed with status {rc}: {msg}")


def ptr(t):
    """Device (or host) address of a torch tensor / numpy array, or None."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return c_void_p(t.data_ptr())
    return c_void_p(t.ctypes.data)


def current_stream():
    """Raw handle of torch's current CUDA stream on the current device (the fast private accessor when torch has it:
    torch.cuda.current_stream() builds a Stream object and costs ~15 us per call, which adds up over a chain of launches)."""
    import torch
    try:
        return c_void_p(torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice()))
    except AttributeError:
        return c_void_p(torch.cuda.current_stream().cuda_stream)


_ctxs = {}


def ctx(device_index: int, lane: int = 0):
    """One bff_ctx (stream + scratch arenas) per device, host thread AND lane, created on first use: the pipelines carve
    their scratch from the context's arenas, so chains that run concurrently (queries of one scene on their own streams,
    scene_step.py) each need their own."""
    import threading
    key = (int(device_index), threading.get_ident(), int(lane))
    if key not in _ctxs:
        h = c_void_p()
        check(lib().bff_ctx_create(int(device_index), ctypes.byref(h)), "bff_ctx_create")
        _ctxs[key] = h
    return _ctxs[key]
