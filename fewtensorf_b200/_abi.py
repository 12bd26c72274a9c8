"""ctypes binding of include/fewtensorf.h (libfewtensorf_b200.so).

There is NO fallback: if the shared library is missing or does not export a symbol the
import raises, and every compute entry point needs a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfewtensorf_b200.so")

MODEL_VM, MODEL_CP = 0, 1
SHADE_MLP_FEA, SHADE_MLP_PE, SHADE_MLP, SHADE_SH = 0, 1, 2, 3
ACT_SOFTPLUS, ACT_RELU = 0, 1
SHADING_KINDS = {"MLP_Fea": SHADE_MLP_FEA, "MLP_PE": SHADE_MLP_PE, "MLP": SHADE_MLP, "SH": SHADE_SH}

WS_NAMES = ["COUNTERS", "TMIN", "RAY_OFF", "RAY_CNT", "ACC", "RGB_PRE", "S_RAY", "S_IDX", "S_SIGMA", "S_W", "S_T",
            "S_APP", "S_GSIG", "A_SLOT", "A_FEAT", "A_RGB", "A_H1", "A_H2", "A_GRGB", "A_GH1", "A_GH2", "A_GFEAT", "A_X", "A_GX", "WIMG"]
WS_COUNT = len(WS_NAMES)
WS = {n: i for i, n in enumerate(WS_NAMES)}

_f32p = C.c_void_p   # device pointers travel as integers


class FtrfFactorSet(C.Structure):
    _fields_ = [
        ("n_comp", C.c_int32 * 3),
        ("plane_w", C.c_int32 * 3),
        ("plane_h", C.c_int32 * 3),
        ("line_len", C.c_int32 * 3),
        ("plane", _f32p * 3),
        ("line", _f32p * 3),
        ("g_plane", _f32p * 3),
        ("g_line", _f32p * 3),
        ("mask", _f32p * 3),
        ("mask_stride", C.c_int32 * 3),
    ]


class FtrfField(C.Structure):
    _fields_ = [
        ("model", C.c_int32), ("shading", C.c_int32), ("act", C.c_int32), ("app_dim", C.c_int32),
        ("feature_c", C.c_int32), ("pos_pe", C.c_int32), ("view_pe", C.c_int32), ("fea_pe", C.c_int32),
        ("in_dim", C.c_int32),
        ("aabb", C.c_float * 6), ("inv_aabb", C.c_float * 3), ("step_size", C.c_float),
        ("near_", C.c_float), ("far_", C.c_float),
        ("distance_scale", C.c_float), ("density_shift", C.c_float), ("weight_thres", C.c_float),
        ("has_alpha_mask", C.c_int32), ("am_dim", C.c_int32 * 3), ("am_aabb0", C.c_float * 3),
        ("am_inv", C.c_float * 3), ("am_bits", C.c_void_p), ("am_coarse", C.c_void_p),
        ("den", FtrfFactorSet), ("app", FtrfFactorSet),
        ("basis", _f32p), ("g_basis", _f32p),
        ("w1", _f32p), ("b1", _f32p), ("w2", _f32p), ("b2", _f32p), ("w3", _f32p), ("b3", _f32p),
        ("g_w1", _f32p), ("g_b1", _f32p), ("g_w2", _f32p), ("g_b2", _f32p), ("g_w3", _f32p), ("g_b3", _f32p),
        ("enc_mask", _f32p * 3), ("enc_mask_len", C.c_int32 * 3),
    ]


class FtrfCall(C.Structure):
    _fields_ = [
        ("n_rays", C.c_int64), ("ray_cols", C.c_int32), ("n_samples", C.c_int32), ("is_train", C.c_int32),
        ("ndc", C.c_int32), ("add_bg", C.c_int32), ("save_for_backward", C.c_int32),
        ("rays", _f32p), ("jitter", _f32p), ("rgb_map", _f32p), ("depth_map", _f32p), ("g_rgb_map", _f32p),
        ("cap_samples", C.c_int64), ("cap_app", C.c_int64),
        ("lib_flags", C.c_uint64), ("lib_ctx", C.c_void_p),       # library-owned hand-off state (zero-initialised)
        ("overflow_out", C.c_void_p),
    ]


class FtrfAdamSegment(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("n", C.c_int64), ("lr", C.c_float), ("pad_", C.c_int32)]


class FtrfRegion(C.Structure):
    _fields_ = [("offset", C.c_int64), ("row_len", C.c_int64), ("row_stride", C.c_int64), ("n_rows", C.c_int64)]


MAX_REGIONS = 24
STATIC_SCRATCH_BYTES = 16 << 20    # FTRF_STATIC_SCRATCH_BYTES / FTRF_STATIC_MAX_POINTS (ftrf_shade_points_static)
STATIC_MAX_POINTS = 8192


class FtrfCamera(C.Structure):
    _fields_ = [("H", C.c_int32), ("W", C.c_int32), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float),
                ("cy", C.c_float), ("c2w", C.c_float * 12), ("normalize", C.c_int32), ("blender", C.c_int32)]


# every symbol include/fewtensorf.h declares
EXPORTS = [
    "ftrf_abi_version", "ftrf_last_error", "ftrf_workspace_layout", "ftrf_forward", "ftrf_forward_ex", "ftrf_backward", "ftrf_backward_part",
    "ftrf_stage_march", "ftrf_stage_density", "ftrf_stage_integrate", "ftrf_stage_appearance",
    "ftrf_stage_shade", "ftrf_stage_composite", "ftrf_mse_loss", "ftrf_alpha_mask_words",
    "ftrf_pack_alpha_mask", "ftrf_alpha_mask_coarse_words", "ftrf_build_coarse_mask", "ftrf_density_at", "ftrf_dense_alpha_bits", "ftrf_dilate_alpha_mask",
    "ftrf_unpack_alpha_mask", "ftrf_filter_rays", "ftrf_event_record", "ftrf_stream_wait_event", "ftrf_zero_after", "ftrf_profile_enable", "ftrf_profile_read",
    "ftrf_profile_stage_name", "ftrf_launch_count", "ftrf_launch_count_add", "ftrf_reg_factors", "ftrf_adam_step", "ftrf_adam_step_guarded", "ftrf_generate_rays", "ftrf_multimem_allreduce", "ftrf_p2p_allreduce",
    "ftrf_points_workspace_bytes", "ftrf_density_feature", "ftrf_app_feature", "ftrf_shade_points", "ftrf_shade_points_static",
    "ftrf_resample_factor", "ftrf_crop_factor", "ftrf_ortho_reg",
    "ftrf_p2p_allreduce_regions", "ftrf_multimem_allreduce_regions", "ftrf_allreduce_regions_fused",
]
TEST_LIB_PATH = os.path.join(_HERE, "libfewtensorf_b200_test.so")    # superset build for tests / scripts only
PROF_STAGES = 13


def _load(path=LIB_PATH):
    if not os.path.exists(path):
        raise ImportError(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C fewtensorf_b200/csrc`). fewtensorf_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(path)
    for name in EXPORTS:
        if not hasattr(lib, name):
            raise ImportError(f"{path} does not export {name}")
    lib.ftrf_abi_version.restype = C.c_int
    lib.ftrf_last_error.restype = C.c_char_p
    lib.ftrf_workspace_layout.restype = C.c_size_t
    lib.ftrf_workspace_layout.argtypes = [C.POINTER(FtrfField), C.POINTER(FtrfCall), C.POINTER(C.c_int64)]
    stage_args = [C.POINTER(FtrfField), C.POINTER(FtrfCall), C.c_void_p, C.c_size_t, C.c_void_p]
    lib.ftrf_backward_part.restype = C.c_int
    lib.ftrf_backward_part.argtypes = stage_args + [C.c_int]
    for name in ("ftrf_forward", "ftrf_backward", "ftrf_stage_march", "ftrf_stage_density", "ftrf_stage_integrate",
                 "ftrf_stage_appearance", "ftrf_stage_shade", "ftrf_stage_composite"):
        fn = getattr(lib, name)
        fn.restype = C.c_int
        fn.argtypes = stage_args
    lib.ftrf_forward_ex.restype = C.c_int
    lib.ftrf_forward_ex.argtypes = stage_args + [C.c_void_p, C.c_void_p]
    lib.ftrf_mse_loss.restype = C.c_int
    lib.ftrf_mse_loss.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ftrf_alpha_mask_words.restype = C.c_size_t
    lib.ftrf_alpha_mask_words.argtypes = [C.c_int32, C.c_int32, C.c_int32]
    lib.ftrf_pack_alpha_mask.restype = C.c_int
    lib.ftrf_pack_alpha_mask.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
    lib.ftrf_alpha_mask_coarse_words.restype = C.c_size_t
    lib.ftrf_alpha_mask_coarse_words.argtypes = [C.c_int32, C.c_int32, C.c_int32]
    lib.ftrf_build_coarse_mask.restype = C.c_int
    lib.ftrf_build_coarse_mask.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
    lib.ftrf_density_at.restype = C.c_int
    lib.ftrf_density_at.argtypes = [C.POINTER(FtrfField), C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    i3 = [C.c_int32, C.c_int32, C.c_int32]
    lib.ftrf_dense_alpha_bits.restype = C.c_int
    lib.ftrf_dense_alpha_bits.argtypes = [C.POINTER(FtrfField), C.c_void_p] + i3 + [C.c_float, C.c_float, C.c_void_p,
                                                                                   C.c_void_p, C.c_void_p]
    lib.ftrf_dilate_alpha_mask.restype = C.c_int
    lib.ftrf_dilate_alpha_mask.argtypes = [C.c_void_p] + i3 + [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ftrf_unpack_alpha_mask.restype = C.c_int
    lib.ftrf_unpack_alpha_mask.argtypes = [C.c_void_p] + i3 + [C.c_void_p, C.c_void_p]
    lib.ftrf_filter_rays.restype = C.c_int
    lib.ftrf_filter_rays.argtypes = [C.POINTER(FtrfField), C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_int32,
                                     C.c_void_p, C.c_void_p]
    lib.ftrf_event_record.restype = C.c_int
    lib.ftrf_event_record.argtypes = [C.c_void_p, C.c_void_p]
    lib.ftrf_stream_wait_event.restype = C.c_int
    lib.ftrf_stream_wait_event.argtypes = [C.c_void_p, C.c_void_p]
    lib.ftrf_zero_after.restype = C.c_int
    lib.ftrf_zero_after.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ftrf_profile_enable.restype = C.c_int
    lib.ftrf_profile_enable.argtypes = [C.c_int]
    lib.ftrf_profile_read.restype = C.c_int
    lib.ftrf_profile_read.argtypes = [C.POINTER(C.c_float)]
    lib.ftrf_profile_stage_name.restype = C.c_char_p
    lib.ftrf_profile_stage_name.argtypes = [C.c_int]
    lib.ftrf_multimem_allreduce.restype = C.c_int
    lib.ftrf_multimem_allreduce.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
    lib.ftrf_p2p_allreduce.restype = C.c_int
    lib.ftrf_p2p_allreduce.argtypes = [C.POINTER(C.c_void_p), C.c_size_t, C.c_int, C.c_int, C.c_void_p]
    lib.ftrf_p2p_allreduce_regions.restype = C.c_int
    lib.ftrf_p2p_allreduce_regions.argtypes = [C.POINTER(C.c_void_p), C.POINTER(FtrfRegion), C.c_int, C.c_int, C.c_int, C.c_void_p]
    lib.ftrf_multimem_allreduce_regions.restype = C.c_int
    lib.ftrf_multimem_allreduce_regions.argtypes = [C.c_void_p, C.POINTER(FtrfRegion), C.c_int, C.c_int, C.c_int, C.c_void_p]
    lib.ftrf_allreduce_regions_fused.restype = C.c_int
    lib.ftrf_allreduce_regions_fused.argtypes = [C.POINTER(C.c_void_p), C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(FtrfRegion),
                                                 C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_void_p]
    lib.ftrf_reg_factors.restype = C.c_int
    lib.ftrf_reg_factors.argtypes = [C.POINTER(FtrfFactorSet), C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p,
                                     C.c_int, C.c_void_p]
    lib.ftrf_adam_step.restype = C.c_int
    lib.ftrf_adam_step.argtypes = [C.POINTER(FtrfAdamSegment), C.c_int, C.c_float, C.c_float, C.c_float, C.c_int,
                                   C.c_void_p]
    lib.ftrf_adam_step_guarded.restype = C.c_int
    lib.ftrf_adam_step_guarded.argtypes = [C.POINTER(FtrfAdamSegment), C.c_int, C.c_float, C.c_float, C.c_float, C.c_int,
                                           C.c_void_p, C.c_void_p]
    lib.ftrf_generate_rays.restype = C.c_int
    lib.ftrf_generate_rays.argtypes = [C.POINTER(FtrfCamera), C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
    lib.ftrf_launch_count.restype = C.c_ulonglong
    lib.ftrf_launch_count.argtypes = []
    lib.ftrf_launch_count_add.restype = None
    lib.ftrf_launch_count_add.argtypes = [C.c_longlong]
    lib.ftrf_points_workspace_bytes.restype = C.c_size_t
    lib.ftrf_points_workspace_bytes.argtypes = [C.POINTER(FtrfField), C.c_int64]
    lib.ftrf_density_feature.restype = C.c_int
    lib.ftrf_density_feature.argtypes = [C.POINTER(FtrfField), C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    lib.ftrf_app_feature.restype = C.c_int
    lib.ftrf_app_feature.argtypes = [C.POINTER(FtrfField), C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.ftrf_shade_points.restype = C.c_int
    lib.ftrf_shade_points.argtypes = [C.POINTER(FtrfField), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                      C.c_void_p, C.c_size_t, C.c_void_p]
    lib.ftrf_shade_points_static.restype = C.c_int
    lib.ftrf_shade_points_static.argtypes = [C.POINTER(FtrfField), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                             C.c_void_p]
    lib.ftrf_resample_factor.restype = C.c_int
    lib.ftrf_resample_factor.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
    lib.ftrf_crop_factor.restype = C.c_int
    lib.ftrf_crop_factor.argtypes = [C.c_void_p] + [C.c_int32] * 7 + [C.c_void_p, C.c_void_p]
    lib.ftrf_ortho_reg.restype = C.c_int
    lib.ftrf_ortho_reg.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                   C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    if lib.ftrf_abi_version() != 2:
        raise ImportError(f"{path} ABI version mismatch")
    return lib


def load_test_lib():
    """The superset test build (cross-check MLP back ends, unit-test entry points, debug switches).  For tests/ and
    scripts/ only: the product never loads it."""
    lib = _load(TEST_LIB_PATH)
    lib.ftrf_test_gemm3x.restype = C.c_int
    lib.ftrf_test_gemm3x.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_int,
                                     C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                     C.c_void_p]
    lib.ftrf_test_wgrad3x.restype = C.c_int
    lib.ftrf_test_wgrad3x.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_int,
                                      C.c_void_p, C.c_void_p]
    lib.ftrf_set_shade_path.restype = C.c_int
    lib.ftrf_set_shade_path.argtypes = [C.c_int]
    return lib


lib = _load()


_raw_stream = None


def current_stream_handle() -> int:
    """cudaStream_t of torch's current stream on the current device.  torch.cuda.current_stream() builds a Stream
    object through two device-index helpers (~10 us; the hot path asks five times per step); torch's raw getter
    answers in well under a microsecond."""
    global _raw_stream
    import torch
    if _raw_stream is None:
        _raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", False)
    if _raw_stream:
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


def check(code: int, what: str = "") -> None:
    """Error convention of the boundary: non-zero return code -> RuntimeError(ftrf_last_error())."""
    if code != 0:
        msg = lib.ftrf_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"fewtensorf {what} failed ({code}): {msg}")
