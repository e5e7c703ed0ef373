"""ctypes binding of libmdb200.so (the C ABI declared in include/mdb200.h).

The library is built in-tree by ``make -C monodepth2.jl_b200`` (or
``__graft_entry__.build()``).  There is no fallback: if the shared object is missing the
import fails, and if no sm_100 device is present ``mdb_create`` fails.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# MDB200_LIB: development hook (tools/sweep_variants.sh builds the library with other -D settings and times them)
LIB_PATH = os.environ.get("MDB200_LIB") or os.path.join(_HERE, "libmdb200.so")

MAX_SCALES = 8
MAX_SOURCES = 2

c_f32p = C.POINTER(C.c_float)


class LossParams(C.Structure):
    _fields_ = [
        ("min_depth", C.c_double), ("max_depth", C.c_double), ("disparity_smoothness", C.c_double),
        ("alpha", C.c_float), ("ssim_c1", C.c_float), ("ssim_c2", C.c_float),
        ("automasking", C.c_int32), ("normalize_disparity", C.c_int32),
        ("smooth_weight_mode", C.c_int32), ("upsample_align_corners", C.c_int32),
        ("n_global", C.c_int32), ("abs_subgradient", C.c_int32),
    ]


class LossIO(C.Structure):
    _fields_ = [
        ("x", C.c_void_p),
        ("W", C.c_int32), ("H", C.c_int32), ("C", C.c_int32), ("L", C.c_int32), ("N", C.c_int32),
        ("target_id", C.c_int32), ("S", C.c_int32),
        ("source_ids", C.c_int32 * MAX_SOURCES), ("invert", C.c_int32 * MAX_SOURCES),
        ("rvec", C.c_void_p * MAX_SOURCES), ("tvec", C.c_void_p * MAX_SOURCES),
        ("Ls", C.c_int32),
        ("disparity", C.c_void_p * MAX_SCALES),
        ("dw", C.c_int32 * MAX_SCALES), ("dh", C.c_int32 * MAX_SCALES),
        ("scale", C.c_float * MAX_SCALES),
        ("auto_loss", C.c_void_p),
        ("K", C.c_float * 9), ("invK", C.c_float * 9),
        ("loss", C.c_void_p),
        ("d_disparity", C.c_void_p * MAX_SCALES),
        ("d_rvec", C.c_void_p * MAX_SOURCES), ("d_tvec", C.c_void_p * MAX_SOURCES),
        ("vis_warp_loss", C.c_void_p),
        ("x_format", C.c_int32),
    ]


class HostIO(C.Structure):
    """mdb_host_io: HOST pointers of one mdb_train_loss_host call."""
    _fields_ = [
        ("x", C.c_void_p), ("flip", C.c_void_p),
        ("disparity", C.c_void_p * MAX_SCALES),
        ("rvec", C.c_void_p * MAX_SOURCES), ("tvec", C.c_void_p * MAX_SOURCES),
        ("auto_loss", C.c_void_p),
        ("loss", C.c_void_p),
        ("d_disparity", C.c_void_p * MAX_SCALES),
        ("d_rvec", C.c_void_p * MAX_SOURCES), ("d_tvec", C.c_void_p * MAX_SOURCES),
    ]


class AdamParams(C.Structure):
    """mdb_adam_params (Flux ADAM defaults; eta from src/simple_depth.jl:16)."""
    _fields_ = [("eta", C.c_double), ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double)]


class RenderOpts(C.Structure):
    """mdb_render_opts; all zero = src/render.jl as written."""
    _fields_ = [("mine_grid", C.c_int32), ("mine_mask", C.c_int32), ("depth_eps", C.c_float), ("reserved", C.c_int32)]


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `make -C monodepth2.jl_b200` "
            "(there is no CPU fallback for the mdb200 kernels)")
    return C.CDLL(LIB_PATH)


lib = _load()

_vp, _i32, _i64, _f32, _f64, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_double, C.c_size_t

# name -> (restype, argtypes); mirrors include/mdb200.h one to one
PROTOTYPES = {
    "mdb_version": (_i32, []),
    "mdb_create": (_i32, [C.POINTER(_vp), _i32]),
    "mdb_destroy": (_i32, [_vp]),
    "mdb_last_error": (C.c_char_p, [_vp]),
    "mdb_launch_count": (_i64, [_vp]),
    "mdb_profile_begin": (_i32, [_vp, _i32]),
    "mdb_profile_read": (_i32, [_vp, _i32, c_f32p]),
    "mdb_loss_params_default": (None, [C.POINTER(LossParams)]),
    "mdb_hat": (_i32, [_vp, _vp, _i32, _vp, _vp]),
    "mdb_hat_pullback": (_i32, [_vp, _vp, _i32, _vp, _vp]),
    "mdb_so3_exp_map": (_i32, [_vp, _vp, _i32, _vp, _vp]),
    "mdb_compose_T": (_i32, [_vp, _vp, _vp, _i32, _i32, _vp, _vp, _vp]),
    "mdb_compose_T_bwd": (_i32, [_vp, _vp, _vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp]),
    "mdb_disparity_to_depth": (_i32, [_vp, _vp, _i64, _f64, _f64, _vp, _vp]),
    "mdb_backproject": (_i32, [_vp, _vp, c_f32p, _i32, _i32, _i32, _vp, _vp]),
    "mdb_project": (_i32, [_vp, _vp, c_f32p, _vp, _vp, _i32, _i32, _i32, _vp, _vp]),
    "mdb_grid_sample": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _vp]),
    "mdb_warp": (_i32, [_vp, _vp, _vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_i32), _i32, c_f32p,
                        c_f32p, _i32, _i32, _i32, _i32, _i32, _f64, _f64, C.POINTER(_vp), _vp]),
    "mdb_warp_bwd": (_i32, [_vp, _vp, _vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_i32), _i32, c_f32p, c_f32p, _i32, _i32, _i32,
                            _i32, _i32, _f64, _f64, C.POINTER(_vp), _vp, C.POINTER(_vp), C.POINTER(_vp), _vp]),
    "mdb_ssim_bwd_workspace_bytes": (_sz, [_i32, _i32, _i32, _i32]),
    "mdb_ssim_bwd": (_i32, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _f32, _f32, _vp, _vp, _sz, _vp]),
    "mdb_prediction_loss_bwd": (_i32, [_vp, C.POINTER(_vp), _i32, _vp, _vp, _i32, _i32, _i32, _i32, _f32, _f32, _f32, _i32,
                                       C.POINTER(_vp), _vp, _sz, _vp]),
    "mdb_apply_mask_bwd": (_i32, [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp]),
    "mdb_sample_bwd": (_i32, [_vp, _vp, _vp, _vp, _vp, c_f32p, c_f32p, _i32, _i32, _i32, _i32, _i32, _vp, _vp, _sz, _vp]),
    "mdb_plane_volume_rendering_bwd": (_i32, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "mdb_ssim": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _f32, _f32, _vp, _vp]),
    "mdb_photometric_loss": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _f32, _f32, _f32, _vp, _vp]),
    "mdb_prediction_loss": (_i32, [_vp, C.POINTER(_vp), _i32, _vp, _i32, _i32, _i32, _i32, _f32, _f32,
                                   _f32, _vp, _vp]),
    "mdb_automasking_loss": (_i32, [_vp, _vp, _vp, _i64, C.POINTER(_i32), _i32, _i32, _i32, _i32, _i32, _i32,
                                    _f32, _f32, _f32, _vp, _vp]),
    "mdb_apply_mask": (_i32, [_vp, _vp, _vp, _i64, _vp, _vp]),
    "mdb_smooth_loss_workspace_bytes": (_sz, [_i32, _i32, _i32]),
    "mdb_smooth_loss": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _sz, _vp]),
    "mdb_smooth_loss_ex": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _sz, _vp]),
    "mdb_smooth_loss_noalloc": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp, _vp, _vp]),
    "mdb_train_loss_strip_geometry": (_i32, [C.POINTER(LossIO), C.POINTER(_i32)]),
    "mdb_train_loss_workspace_bytes": (_sz, [C.POINTER(LossIO)]),
    "mdb_train_loss_fwdbwd": (_i32, [_vp, C.POINTER(LossIO), C.POINTER(LossParams), _vp, _sz, _vp]),
    "mdb_inv3x3_batched": (_i32, [_vp, _vp, _i32, _vp, _vp]),
    "mdb_create_meshgrid": (_i32, [_vp, _i32, _i32, _vp, _vp]),
    "mdb_plane_xyz_src": (_i32, [_vp, _vp, _vp, c_f32p, _i32, _i32, _i32, _i32, _vp, _vp]),
    "mdb_plane_xyz_tgt": (_i32, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp]),
    "mdb_render_workspace_bytes": (_sz, [_i32, _i32, _i32, _i32]),
    "mdb_sample": (_i32, [_vp, _vp, _vp, _vp, _vp, c_f32p, c_f32p, _i32, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _sz, _vp]),
    "mdb_plane_volume_rendering": (_i32, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _vp]),
    "mdb_render_tgt_rgb_depth_fwd": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, c_f32p, c_f32p, _i32, _i32, _i32, _i32,
                                            _vp, _vp, _vp, _vp, _sz, _vp]),
    "mdb_render_tgt_rgb_depth_bwd": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, c_f32p, c_f32p, _i32, _i32, _i32, _i32,
                                            _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "mdb_render_tgt_rgb_depth_fwd_ex": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, c_f32p, c_f32p, _i32, _i32, _i32, _i32,
                                               C.POINTER(RenderOpts), _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "mdb_render_tgt_rgb_depth_bwd_ex": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, c_f32p, c_f32p, _i32, _i32, _i32, _i32,
                                               C.POINTER(RenderOpts), _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                                               _vp, _vp, _sz, _vp]),
    "mdb_adam_update": (_i32, [_vp, _i32, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp),
                               C.POINTER(_i64), C.POINTER(AdamParams), _vp, _vp, _vp, _vp, _i32, _vp]),
    "mdb_slow_depth_state_bytes": (_sz, [C.POINTER(LossIO)]),
    "mdb_slow_depth_create": (_i32, [_vp, C.POINTER(LossIO), C.POINTER(LossParams), C.POINTER(AdamParams), _vp, _sz,
                                     _vp, _sz, _vp, _i32, C.POINTER(_vp)]),
    "mdb_slow_depth_run": (_i32, [_vp, _i32, _vp]),
    "mdb_slow_depth_destroy": (_i32, [_vp]),
    "mdb_ingest_triplet_u8": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp]),
    "mdb_ingest_kitti_u8": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp, _vp]),
    "mdb_ingest_triplet_u8_planar": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp]),
    "mdb_pose_head": (_i32, [_vp, _vp, _i32, _i32, _i32, _f32, _vp, _vp, _vp]),
    "mdb_pose_head_bwd": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _f32, _vp, _vp]),
    "mdb_disparity_heads": (_i32, [_vp, C.POINTER(_vp), C.POINTER(_i32), C.POINTER(_i32), _i32, _i32, C.POINTER(_vp), _vp]),
    "mdb_disparity_heads_bwd": (_i32, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_i32), C.POINTER(_i32), _i32, _i32, C.POINTER(_vp), _vp]),
    "mdb_item_mean": (_i32, [_vp, _vp, _i64, _i32, _vp, _vp]),
    "mdb_measure_fp32_peak": (_i32, [_vp, _vp, _sz, _i32, C.POINTER(_f64), C.POINTER(_f64), _vp]),
    "mdb_debug_option": (_i32, [C.c_char_p, _i64]),
    "mdb_host_plan_create": (_i32, [_vp, C.POINTER(LossIO), C.POINTER(LossParams), _i32, _i32, _i32, C.POINTER(_vp)]),
    "mdb_train_loss_host": (_i32, [_vp, C.POINTER(HostIO), _vp]),
    "mdb_host_plan_info": (_i32, [_vp, C.POINTER(_i64)]),
    "mdb_host_plan_destroy": (_i32, [_vp]),
    "mdb_packed_size": (_i64, [_i32, _i32]),
    "mdb_packed_bind": (_i32, [C.POINTER(LossIO), _vp, _i32, _i32]),
    "mdb_comm_get_unique_id": (_i32, [_vp, _sz]),
    "mdb_comm_init": (_i32, [_vp, _vp, _i32, _i32]),
    "mdb_comm_peers_export": (_i32, [_vp, _i32, _i64, _vp, _sz]),
    "mdb_comm_peers_attach": (_i32, [_vp, _vp, _i32]),
    "mdb_allreduce_loss_posegrad": (_i32, [_vp, _vp, _vp, _i64, _vp]),
    "mdb_comm_status": (_i32, [_vp, C.POINTER(_i32)]),
    "mdb_comm_destroy": (_i32, [_vp]),
}


def _bind():
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args


_bind()


class MdbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"mdb200 error {code}: {msg}")
        self.code = code


def check(handle, code: int):
    if code != 0:
        msg = lib.mdb_last_error(handle).decode() if handle else "no handle"
        raise MdbError(code, msg)
