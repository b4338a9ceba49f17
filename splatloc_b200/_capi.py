"""ctypes binding of libsplatloc_b200.so (the C ABI in include/splatloc_b200.h).

This is the only place the shared library is loaded.  There is no fallback: if
the library is missing or a CUDA device is absent, the product path raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# SPLATLOC_B200_LIB: load another build of the same library (kernel tuning experiments)
LIB_PATH = os.environ.get("SPLATLOC_B200_LIB") or os.path.join(_HERE, "libsplatloc_b200.so")

ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_size_t)

_F = C.c_void_p  # every device pointer travels as an integer address

class AdamGroup(C.Structure):          # sl_adam_group
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("width", C.c_int), ("lr", C.c_float), ("freeze_by_marker", C.c_int)]


class DensifyGroup(C.Structure):       # sl_densify_group
    _fields_ = [("param_in", C.c_void_p), ("param_out", C.c_void_p), ("exp_avg_in", C.c_void_p), ("exp_avg_out", C.c_void_p),
                ("exp_avg_sq_in", C.c_void_p), ("exp_avg_sq_out", C.c_void_p), ("width", C.c_int)]


# name -> (restype, argtypes); kept in sync with include/splatloc_b200.h and
# checked against the header's declarations by tests/test_capi_symbols.py.
SIGNATURES = {
    "sl_version": (C.c_char_p, []),
    "sl_last_error": (C.c_char_p, []),
    "sl_geometry_bytes": (C.c_size_t, [C.c_int]),
    "sl_binning_bytes": (C.c_size_t, [C.c_int]),
    "sl_image_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "sl_geometry_bytes_batch": (C.c_size_t, [C.c_int, C.c_int]),
    "sl_image_bytes_batch": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "sl_forward": (C.c_int, [
        ALLOC_FN, C.c_void_p, ALLOC_FN, C.c_void_p, ALLOC_FN, C.c_void_p,
        C.c_int, C.c_int, C.c_int, C.c_int,          # P, D, M, C
        _F, C.c_int, C.c_int,                        # background, width, height
        _F, _F, _F,                                  # means3D, shs, colors_precomp
        _F, _F, C.c_float,                           # opacities, scales, scale_modifier
        _F, _F, _F,                                  # rotations, cov3D_precomp, tile_mask
        _F, _F, _F,                                  # viewmatrix, projmatrix, cam_pos
        C.c_float, C.c_float, C.c_int,               # tan_fovx, tan_fovy, prefiltered
        _F, _F, _F, _F,                              # out_color, out_depth, out_alpha, radii
        C.c_int, C.c_void_p,                         # debug, stream
    ]),
    "sl_backward": (C.c_int, [
        C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,  # P, D, M, R, C
        _F, C.c_int, C.c_int,                        # background, width, height
        _F, _F, _F,                                  # means3D, shs, colors_precomp
        _F, C.c_float, _F,                           # scales, scale_modifier, rotations
        _F, _F, _F,                                  # cov3D_precomp, viewmatrix, projmatrix
        _F, C.c_float, C.c_float, _F,                # campos, tan_fovx, tan_fovy, radii
        _F, _F, _F,                                  # geometry, binning, image chunks
        _F, _F, _F, _F,                              # accum_alphas, dL_dpix, dL_dpix_depth, dL_dpix_alpha
        _F, _F, _F, _F,                              # dL_dmean2D, dL_dabsmean2D, dL_dconic, dL_dopacity
        _F, _F, _F, _F,                              # dL_dcolor, dL_ddepth, dL_dmean3D, dL_dcov3D
        _F, _F, _F,                                  # dL_dsh, dL_dscale, dL_drot
        C.c_int, C.c_void_p,                         # debug, stream
    ]),
    "sl_forward_ext": (C.c_int, [
        ALLOC_FN, C.c_void_p, ALLOC_FN, C.c_void_p, ALLOC_FN, C.c_void_p,
        C.c_int, C.c_int, C.c_int, C.c_int,          # P, D, M, C
        _F, C.c_int, C.c_int,                        # background, width, height
        _F, _F, _F,                                  # means3D, shs, colors_precomp
        _F, _F, C.c_float,                           # opacities, scales, scale_modifier
        _F, _F, _F,                                  # rotations, cov3D_precomp, tile_mask
        _F, _F, _F,                                  # viewmatrix, projmatrix, cam_pos
        C.c_float, C.c_float, C.c_int,               # tan_fovx, tan_fovy, prefiltered
        _F, _F, _F, _F, _F,                          # out_color, out_depth, out_alpha, radii, n_touched
        C.c_int, C.c_void_p,                         # debug, stream
    ]),
    "sl_forward_nosync": (C.c_int, [
        ALLOC_FN, C.c_void_p, ALLOC_FN, C.c_void_p, ALLOC_FN, C.c_void_p,
        C.c_int, C.c_int, C.c_int, C.c_int,          # P, D, M, C
        _F, C.c_int, C.c_int,                        # background, width, height
        _F, _F, _F,                                  # means3D, shs, colors_precomp
        _F, _F, C.c_float,                           # opacities, scales, scale_modifier
        _F, _F, _F,                                  # rotations, cov3D_precomp, tile_mask
        _F, _F, _F,                                  # viewmatrix, projmatrix, cam_pos
        C.c_float, C.c_float, C.c_int,               # tan_fovx, tan_fovy, prefiltered
        _F, _F, _F, _F, _F,                          # out_color, out_depth, out_alpha, radii, n_touched
        C.c_int, _F, C.c_void_p,                     # max_pairs, status, stream
    ]),
    "sl_backward_ext": (C.c_int, [
        C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,  # P, D, M, R, C
        _F, C.c_int, C.c_int,                        # background, width, height
        _F, _F, _F,                                  # means3D, shs, colors_precomp
        _F, C.c_float, _F,                           # scales, scale_modifier, rotations
        _F, _F, _F,                                  # cov3D_precomp, viewmatrix, projmatrix
        _F, C.c_float, C.c_float, _F,                # campos, tan_fovx, tan_fovy, radii
        _F, _F, _F,                                  # geometry, binning, image chunks
        _F, _F, _F, _F,                              # accum_alphas, dL_dpix, dL_dpix_depth, dL_dpix_alpha
        _F, _F, _F, _F,                              # dL_dmean2D, dL_dabsmean2D, dL_dconic, dL_dopacity
        _F, _F, _F, _F,                              # dL_dcolor, dL_ddepth, dL_dmean3D, dL_dcov3D
        _F, _F, _F, _F,                              # dL_dsh, dL_dscale, dL_drot, dL_dtau
        C.c_int, C.c_void_p,                         # debug, stream
    ]),
    "sl_forward_batch": (C.c_int, [
        ALLOC_FN, C.c_void_p, ALLOC_FN, C.c_void_p, ALLOC_FN, C.c_void_p,
        C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,  # n_views, P, D, M, C
        _F, C.c_int, C.c_int,                        # background, width, height
        _F, _F, _F,                                  # means3D, shs, colors_precomp
        _F, _F, C.c_float,                           # opacities, scales, scale_modifier
        _F, _F, _F,                                  # rotations, cov3D_precomp, tile_mask
        _F, _F, _F,                                  # viewmatrices, projmatrices, cam_positions
        C.c_float, C.c_float, C.c_int,               # tan_fovx, tan_fovy, prefiltered
        _F, _F, _F, _F, _F,                          # out_color, out_depth, out_alpha, radii, n_touched
        C.c_int, _F, C.c_int, C.c_void_p,            # max_pairs, status, debug, stream
    ]),
    "sl_backward_batch": (C.c_int, [
        C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,  # n_views, P, D, M, R, C
        _F, C.c_int, C.c_int,                        # background, width, height
        _F, _F, _F,                                  # means3D, shs, colors_precomp
        _F, C.c_float, _F,                           # scales, scale_modifier, rotations
        _F, _F, _F,                                  # cov3D_precomp, viewmatrices, projmatrices
        _F, C.c_float, C.c_float, _F,                # cam_positions, tan_fovx, tan_fovy, radii
        _F, _F, _F,                                  # geometry, binning, image chunks
        _F, _F, _F, _F,                              # accum_alphas, dL_dpix, dL_dpix_depth, dL_dpix_alpha
        _F, _F, _F, _F,                              # dL_dmean2D, dL_dabsmean2D, dL_dconic, dL_dopacity
        _F, _F, _F, _F,                              # dL_dcolor, dL_ddepth, dL_dmean3D, dL_dcov3D
        _F, _F, _F, _F,                              # dL_dsh, dL_dscale, dL_drot, dL_dtau
        C.c_int, C.c_void_p,                         # debug, stream
    ]),
    "sl_ssim_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "sl_ssim_forward": (C.c_int, [C.c_int, C.c_int, C.c_int, _F, _F, C.POINTER(C.c_float), _F, _F, _F, C.c_void_p]),
    "sl_ssim_backward": (C.c_int, [C.c_int, C.c_int, C.c_int, _F, _F, C.POINTER(C.c_float), _F, _F, C.c_float, C.c_float,
                                   _F, C.c_void_p]),
    "sl_exchange_arrive": (C.c_int, [_F, C.c_int, C.c_void_p]),                    # flags_multicast, rank, stream
    "sl_exchange_wait": (C.c_int, [_F, C.c_int, _F, C.c_int, C.c_int, C.c_int, _F, C.c_void_p]),  # flags, world, waits, ahead, advance, timeout_ms, status, stream
    "sl_exchange_push": (C.c_int, [_F, C.c_int, C.c_int, _F, C.c_int, C.c_void_p]),   # source (device), P, C, exchange_multicast, max_ctas, stream
    "sl_backward_batch_push": (C.c_int, [
        C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,  # n_views, P, D, M, R, C
        _F, C.c_int, C.c_int,                        # background, width, height
        _F, _F, _F,                                  # means3D, shs, colors_precomp
        _F, C.c_float, _F,                           # scales, scale_modifier, rotations
        _F, _F, _F,                                  # cov3D_precomp, viewmatrices, projmatrices
        _F, C.c_float, C.c_float, _F,                # cam_positions, tan_fovx, tan_fovy, radii
        _F, _F, _F,                                  # geometry, binning, image chunks
        _F, _F, _F, _F,                              # accum_alphas, dL_dpix, dL_dpix_depth, dL_dpix_alpha
        _F, _F, _F, _F,                              # dL_dmean2D, dL_dabsmean2D, dL_dconic, dL_dopacity
        _F, _F, _F, _F,                              # dL_dcolor, dL_ddepth, dL_dmean3D, dL_dcov3D
        _F, _F, _F, _F,                              # dL_dsh, dL_dscale, dL_drot, dL_dtau
        _F, C.c_int, C.c_void_p,                     # exchange_multicast, debug, stream
    ]),
    "sl_forward_render": (C.c_int, [
        ALLOC_FN, C.c_void_p, ALLOC_FN, C.c_void_p, ALLOC_FN, C.c_void_p,
        C.c_int, C.c_int, _F, C.c_int, C.c_int,      # n_views, P, background, width, height
        _F, _F, _F,                                  # means3D, features_dc, kp_score
        _F, _F, C.c_int, C.c_float,                  # raw_opacity, raw_scaling, scale_dim, scale_modifier
        _F, _F,                                      # raw_rotation, tile_mask
        _F, _F, _F, C.c_float, C.c_float,            # viewmatrices, projmatrices, cam_positions, tan_fovx, tan_fovy
        _F, _F, _F, _F,                              # out_color, out_depth, out_alpha, radii
        C.c_int, _F, C.c_int, C.c_void_p,            # max_pairs, status, debug, stream
    ]),
    "sl_backward_render": (C.c_int, [
        C.c_int, C.c_int, C.c_int, _F, C.c_int, C.c_int,   # n_views, P, R, background, width, height
        _F, _F, _F,                                  # means3D, features_dc, kp_score
        _F, _F, C.c_int, C.c_float, _F,              # raw_opacity, raw_scaling, scale_dim, scale_modifier, raw_rotation
        _F, _F, _F, C.c_float, C.c_float, _F,        # viewmatrices, projmatrices, cam_positions, tan_fovx, tan_fovy, radii
        _F, _F, _F,                                  # geometry, binning, image chunks
        _F, _F, _F, _F,                              # accum_alphas, dL_dpix, dL_dpix_depth, dL_dpix_alpha
        _F, _F, _F, _F, _F, _F,                      # dL_dmean2D, dL_dabsmean2D, dL_dconic, dL_dcolor, dL_ddepth, dL_dcov3D
        _F, _F, _F, _F, _F, _F,                      # dL_dmean3D, dL_dfeatures_dc, dL_dkp_score, dL_draw_opacity/scaling/rotation
        C.c_int, C.c_void_p,                         # debug, stream
    ]),
    "sl_adam_step": (C.c_int, [C.c_int, C.c_int, C.POINTER(AdamGroup), C.c_int, C.c_double, C.c_double, C.c_double,
                               _F, C.c_float, C.c_void_p]),
    "sl_adam_step_counted": (C.c_int, [C.c_int, C.c_int, C.POINTER(AdamGroup), _F, C.c_double, C.c_double, C.c_double,
                                       _F, C.c_float, C.c_void_p]),
    "sl_densify_stats": (C.c_int, [C.c_int, C.c_int, _F, _F, _F, _F, _F, C.c_void_p]),
    "sl_densify_scratch_bytes": (C.c_size_t, [C.c_int]),
    "sl_densify_plan": (C.c_int, [C.c_int, _F, _F, _F, C.c_int, _F, _F, C.c_float, C.c_float, C.c_float, C.c_float,
                                  C.c_int, C.c_int, _F, C.c_size_t, _F, C.c_void_p]),
    "sl_densify_apply": (C.c_int, [C.c_int, C.c_int, C.POINTER(DensifyGroup), C.c_int, C.c_int, C.c_int, _F, _F, _F,
                                   C.c_void_p]),
    "sl_mark_visible": (C.c_int, [C.c_int, _F, _F, _F, _F, C.c_void_p]),
    "sl_knn_scratch_bytes": (C.c_size_t, [C.c_int]),
    "sl_knn": (C.c_int, [C.c_int, _F, _F, _F, C.c_size_t, C.c_void_p]),
    "sl_knn_static": (C.c_int, [C.c_int, _F, _F, C.c_void_p]),
    "sl_mapping_loss_workspace_bytes": (C.c_size_t, []),
    "sl_mapping_loss": (C.c_int, [
        C.c_int, C.c_int, C.c_int,                   # width, height, channels
        _F, _F, _F, _F, _F,                          # image, depth, gt_image, gt_depth, gt_kp
        _F, _F, C.c_float, C.c_int,                  # exposure_a, exposure_b, rgb_boundary_threshold, initialization
        _F, _F, _F, _F, C.c_void_p,                  # dL_dimage, dL_ddepth, out, workspace, stream
    ]),
    "sl_launch_count": (C.c_longlong, []),
    "sl_set_binning_mode": (None, [C.c_int]),
    "sl_set_record_stream": (None, [C.c_int]),
    "sl_set_chained_launches": (None, [C.c_int]),
    "sl_profile_enable": (None, [C.c_int]),
    "sl_profile_stage_count": (C.c_int, []),
    "sl_profile_stage_name": (C.c_char_p, [C.c_int]),
    "sl_profile_read": (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_longlong), C.c_int]),
}

_lib = None


def load() -> C.CDLL:
    """Load the library once; raise (never fall back) when it is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m splatloc_b200.build` "
                "(there is no CPU or PyTorch fallback for the rasterizer)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        if os.environ.get("SPLATLOC_B200_BINNING"):      # tuning experiments: see sl_set_binning_mode
            lib.sl_set_binning_mode(int(os.environ["SPLATLOC_B200_BINNING"]))
        if os.environ.get("SPLATLOC_B200_CHAINED_LAUNCHES"):  # A/B: see sl_set_chained_launches
            lib.sl_set_chained_launches(int(os.environ["SPLATLOC_B200_CHAINED_LAUNCHES"]))
        if os.environ.get("SPLATLOC_B200_RECORD_STREAM"):  # A/B: see sl_set_record_stream
            lib.sl_set_record_stream(int(os.environ["SPLATLOC_B200_RECORD_STREAM"]))
        _lib = lib
    return _lib


def last_error() -> str:
    return load().sl_last_error().decode("utf-8", "replace")


def check(rc: int) -> int:
    """Raise RuntimeError with the library's message on a negative return code."""
    if rc < 0:
        raise RuntimeError(last_error() or f"libsplatloc_b200 error {rc}")
    return rc


def profile_read() -> dict:
    """{stage name: (total ms, timed calls)} accumulated since sl_profile_enable(1)."""
    lib = load()
    n = lib.sl_profile_stage_count()
    ms = (C.c_double * n)()
    calls = (C.c_longlong * n)()
    lib.sl_profile_read(ms, calls, n)
    return {lib.sl_profile_stage_name(i).decode(): (ms[i], calls[i]) for i in range(n)}
