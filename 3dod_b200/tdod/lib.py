"""ctypes binding of libtdod.so (include/tdod.h).  No torch types cross the boundary: tensors are passed as raw
device pointers + sizes, the stream as a void*.  There is no CPU fallback: a missing library or a non-CUDA tensor
raises."""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TDOD_LIB", os.path.join(_HERE, "libtdod.so"))   # TDOD_LIB: A/B builds while tuning

RNG_HOST, RNG_PHILOX = 0, 1
BACKPROJECT_NUMPY, BACKPROJECT_TORCH = 0, 1
CENTRE_DEPTH, CENTRE_UNIFORM, CENTRE_PATCH, CENTRE_PATCH_Z = 0, 1, 2, 3
DIMS_PRIOR, DIMS_UNIFORM, DIMS_ASPECT = 0, 1, 2
STD_EXACT, STD_TORCH_CUDA = 0, 1
MODEB_SEG, MODEB_PC = 1, 2

_lib = None


class PssArgs(C.Structure):
    """mirror of `tdod_pss_args` (include/tdod.h)"""
    _fields_ = [
        ("depth", C.c_void_p), ("K", C.c_void_p), ("yaw_table", C.c_void_p),
        ("F", C.c_int32), ("H", C.c_int32), ("W", C.c_int32),
        ("n_obj", C.c_int32), ("n_prop", C.c_int32),
        ("obj_frame", C.c_void_p), ("boxes", C.c_void_p), ("iou_boxes", C.c_void_p),
        ("prior_mean", C.c_void_p), ("prior_std", C.c_void_p), ("rects", C.c_void_p), ("has_rect", C.c_void_p),
        ("rng_mode", C.c_int32),
        ("noise", C.c_void_p), ("yaw_idx", C.c_void_p), ("rot", C.c_void_p),
        ("seed", C.c_uint64), ("obj_id_base", C.c_int64),
        ("best_idx", C.c_void_p), ("best_score", C.c_void_p), ("best_cube", C.c_void_p), ("best_box", C.c_void_p),
        ("all_cubes", C.c_void_p), ("all_boxes", C.c_void_p), ("all_scores", C.c_void_p), ("stats", C.c_void_p),
        ("cluster_size", C.c_int32), ("propose_only", C.c_int32),
        ("centre_mode", C.c_int32), ("dims_mode", C.c_int32), ("im_w", C.c_int32), ("im_h", C.c_int32),
        ("aux", C.c_void_p),
        ("std_mode", C.c_int32),
        ("mode_flags", C.c_int32), ("score_channels", C.c_int32),
        ("masks", C.c_void_p), ("points", C.c_void_p), ("n_points", C.c_void_p), ("P_cap", C.c_int64),
        ("points_f64", C.c_int32), ("all_pc", C.c_void_p),
        ("out_noise", C.c_void_p), ("out_yaw_idx", C.c_void_p), ("out_seeds", C.c_void_p),
    ]


_P, _I, _L, _D, _U64 = C.c_void_p, C.c_int, C.c_int64, C.c_double, C.c_uint64

# name -> (restype, argtypes); must list every function include/tdod.h declares (tests check this)
SIGNATURES = {
    "tdod_version": (_I, []),
    "tdod_strerror": (C.c_char_p, [_I]),
    "tdod_last_error_string": (C.c_char_p, []),
    "tdod_cube_corners": (_I, [_P, _L, _P, _P]),
    "tdod_project_cubes": (_I, [_P, _L, _P, _I, _I, _P, _P, _P]),
    "tdod_project_cubes_indexed": (_I, [_P, _L, _P, _P, _I, _I, _P, _P, _P]),
    "tdod_score_iou": (_I, [_P, _P, _L, _P, _P]),
    "tdod_score_dimensions": (_I, [_P, _P, _P, _L, _P, _P, _L, _P, _P]),
    "tdod_mask_rect_workspace_bytes": (_L, [_L, _I, _I]),
    "tdod_mask_rect": (_I, [_P, _L, _I, _I, _P, _P, _P, _L, _P]),
    "tdod_mask_rect_bitmap": (_I, [_P, _L, _I, _I, _P, _P, _P, _L, _P]),
    "tdod_mask_pack": (_I, [_P, _L, _I, _I, _P, _L, _P]),
    "tdod_mask_rect_packed": (_I, [_L, _I, _I, _P, _P, _P, _L, _P]),
    "tdod_score_corners": (_I, [_P, _I, _P, _L, _P, _P]),
    "tdod_score_segmentation": (_I, [_P, _I, _I, _P, _L, _I, _P, _P]),
    "tdod_score_point_cloud": (_I, [_P, _I, _L, _P, _L, _P, _P]),
    "tdod_box3d_iou": (_I, [_P, _P, _L, _L, _P, _P, _P]),
    "tdod_segmented_order": (_I, [_P, _L, _I, _I, _P, _P, _P]),
    "tdod_accumulate_scores": (_I, [_P, _P, _L, _I, _I, _I, _P, _P]),
    "tdod_backproject": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _P, _P, _P, _P]),
    "tdod_ransac_workspace_bytes": (_L, [_I, _I, _I]),
    "tdod_ransac_plane": (_I, [_P, _I, _P, _I, _L, _P, _U64, _L, _I, _D, _P, _P, _P, _P, _P, _P, _L, _P]),
    "tdod_ransac_plane_static": (_I, [_P, _I, _P, _I, _L, _P, _U64, _L, _I, _D, _P, _P, _P, _P, _P, _P]),
    "tdod_host_pack_masks": (_I, [_P, _L, _I, _I, _P, _I]),
    "tdod_yaw_table": (_I, [_P, _I, _I, _P, _P]),
    "tdod_resize_nearest": (_I, [_P, _I, _I, _I, _I, _P, _P, _P, _I, _I, _P]),
    "tdod_torch_std": (_I, [_P, _L, _I, _P, _P]),
    "tdod_propose_score_select": (_I, [C.POINTER(PssArgs), _P]),
    "tdod_cube_box_fwd": (_I, [_P, _P, _P, _P, _L, _P, _P, _P, _P, _P, _P, _P]),
    "tdod_cube_box_bwd": (_I, [_P, _P, _P, _P, _L, _P, _P, _P, _P, _P, _P, _P]),
    "tdod_z_sweep": (_I, [_P, _P, _P, _P, _P, _P, _I, _L, _P, _P]),
    "tdod_hull": (_I, [_P, _L, _P, _P, _P]),
    "tdod_hull_fill_moments": (_I, [_P, _P, _P, _P, _L, _I, _I, _I, _P, _P]),
    "tdod_hull_fill_grad": (_I, [_P, _P, _P, _P, _L, _I, _I, _P, _P, _P, _P]),
    "tdod_hull_fill": (_I, [_P, _P, _L, _I, _I, _P, _P]),
    "tdod_hull_fill_dense_grad": (_I, [_P, _P, _L, _I, _I, _P, _P, _P, _P]),
    "tdod_depth_select": (_I, [_P, _L, _I, _I, _P, _P, _P, _P, _I, _L, _P, _P, _P]),
}


def lib():
    """Loads libtdod.so once.  Raises (never falls back) when the CUDA extension has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libtdod.so is missing (%s): run `python 3dod_b200/build.py`; there is no CPU path"
                               % LIB_PATH)
        handle = C.CDLL(LIB_PATH)
        missing = [name for name in SIGNATURES if not hasattr(handle, name)]
        if missing:
            raise RuntimeError("libtdod.so is stale: missing symbols %s; rebuild with 3dod_b200/build.py" % missing)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc):
    if rc == 0:
        return
    h = lib()
    msg = h.tdod_strerror(rc).decode()
    if rc < 0:
        raise ValueError("libtdod: %s: %s" % (msg, h.tdod_last_error_string().decode()))
    raise RuntimeError("libtdod: CUDA error %d: %s" % (rc, msg))


def ptr(t, dtype=None, name="tensor"):
    """device pointer of a contiguous CUDA tensor (None -> NULL); anything else raises (no silent copies)."""
    if t is None:
        return None
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor" % name)
    if not t.is_cuda:
        raise ValueError("%s must be a CUDA tensor (libtdod has no CPU path)" % name)
    if t.device.index != torch.cuda.current_device():
        # the launch goes to the CURRENT device's stream: a foreign pointer would fault there (tdod.api switches the device)
        raise ValueError("%s lives on %s while cuda:%d is current: call under torch.cuda.device(%s)"
                         % (name, t.device, torch.cuda.current_device(), t.device))
    if not t.is_contiguous():
        raise ValueError("%s must be contiguous" % name)
    if dtype is not None and t.dtype != dtype:
        raise ValueError("%s must be %s, got %s" % (name, dtype, t.dtype))
    return t.data_ptr()


def stream_ptr(device=None):
    return torch.cuda.current_stream(device).cuda_stream
