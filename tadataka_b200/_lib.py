"""ctypes binding of libtdk_b200.so (the C-ABI declared in include/tdk.h).

There is no CPU fallback: if the shared library is missing, import fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# TDK_B200_LIB: development override (an experimental build of the same sources)
LIB_PATH = os.environ.get("TDK_B200_LIB") or os.path.join(_HERE, "libtdk_b200.so")

OK = 0
ERR_INVALID, ERR_CUDA, ERR_AGE, ERR_SINGULAR, ERR_WORKSPACE, ERR_SEARCH_LEN, ERR_RANGE = \
    -1, -2, -3, -4, -5, -6, -7

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -m tadataka_b200.build` "
        "(nvcc, sm_100a).  tadataka_b200 has no CPU fallback.")

lib = C.CDLL(LIB_PATH)

P = C.c_void_p
I64 = C.c_int64
D = C.c_double
SZ = C.c_size_t


def _sig(name, restype, *argtypes):
    f = getattr(lib, name)
    f.restype = restype
    f.argtypes = list(argtypes)
    return f


last_error = _sig("tdk_last_error", C.c_char_p)
version = _sig("tdk_version", C.c_char_p)
set_kernel_plan = _sig("tdk_set_kernel_plan", C.c_int, C.c_int)
set_subpixel_refinement = _sig("tdk_set_subpixel_refinement", C.c_int, C.c_int)
DEFAULT_PLAN = -1  # automatic: sorted line records (2) except for single small keyframes (0); update_depth.cu
update_depth_workspace_bytes = _sig("tdk_update_depth_workspace_bytes", SZ, I64, I64, I64, I64)
update_depth_workspace_bytes_min = _sig("tdk_update_depth_workspace_bytes_min", SZ, I64, I64, I64, I64)
update_depth_dev = _sig("tdk_update_depth_dev", C.c_int, P, P, P, P, P, P, P, P, P, P, I64, I64,
                        I64, I64, P, P, P, P, P, C.c_int, P, SZ, P)
update_depth_frames_dev = _sig("tdk_update_depth_frames_dev", C.c_int, P, P, P, P, P, P, P, P, P,
                               P, I64, I64, I64, I64, P, P, P, P, P, C.c_int, P, SZ, P)
update_depth_tables_dev = _sig("tdk_update_depth_tables_dev", C.c_int, P, P, P, P, P, P, P, P, P,
                               P, I64, I64, I64, I64, P, P, P, P, P, C.c_int, P, SZ, P)
update_depth_static_dev = _sig("tdk_update_depth_static_dev", C.c_int, P, P, P, P, P, P, P, P, P,
                               P, I64, I64, I64, I64, P, P, P, P, P, C.c_int, C.c_int, P)
STATIC_MAX_PAIRS = 64
update_depth_status = _sig("tdk_update_depth_status", C.c_int, P, P)
update_depth_maps_host = _sig("tdk_update_depth_maps_host", C.c_int, P, P, P, P, P, P, P, P, P, P, I64, I64, I64,
                              I64, P, P, P, P, P, C.c_int)
selftest_host_wire = _sig("tdk_selftest_host_wire", C.c_int, P, P, I64, P, P)
update_depth_host = _sig("tdk_update_depth_host", C.c_int, P, P, P, P, P, P, P, P, P, P, I64, I64,
                         I64, I64, P, P, P, P, P, C.c_int)
update_depth_frames_host = _sig("tdk_update_depth_frames_host", C.c_int, P, P, P, P, P, P, P, P, P, P, I64, I64,
                                I64, I64, P, P, P, P, P, C.c_int)
estimate_debug_host = _sig("tdk_estimate_debug_host", C.c_int, P, D, D, P, P, P, P, P, P, P, I64,
                           I64, P, P, P)
fusion_dev = _sig("tdk_fusion_dev", C.c_int, P, P, P, P, P, I64, C.c_int, P, P, P)
regularize_dev = _sig("tdk_regularize_dev", C.c_int, P, P, P, I64, I64, I64, P, P)
increment_age_workspace_bytes = _sig("tdk_increment_age_workspace_bytes", SZ, I64, I64, I64)
increment_age_dev = _sig("tdk_increment_age_dev", C.c_int, P, P, P, P, P, I64, I64, I64, P, P, SZ,
                         P)
propagate_workspace_bytes = _sig("tdk_propagate_workspace_bytes", SZ, I64, I64, I64)
propagate_dev = _sig("tdk_propagate_dev", C.c_int, P, P, P, P, P, I64, I64, I64, D, D, D, P, P, P,
                     SZ, P)
propagate_age_dev = _sig("tdk_propagate_age_dev", C.c_int, P, P, P, P, P, P, I64, I64, I64, D, D, D, P, P, P, P,
                         SZ, P)
increment_age_host = _sig("tdk_increment_age_host", C.c_int, P, P, P, P, P, I64, I64, I64, P)
propagate_host = _sig("tdk_propagate_host", C.c_int, P, P, P, P, P, I64, I64, I64, D, D, D, P, P)
fusion_host = _sig("tdk_fusion_host", C.c_int, P, P, P, P, P, I64, C.c_int, P, P)
regularize_host = _sig("tdk_regularize_host", C.c_int, P, P, P, I64, I64, I64, P)
mapper_arena_bytes = _sig("tdk_mapper_arena_bytes", SZ, I64, I64, I64)
mapper_create = _sig("tdk_mapper_create", C.c_int, P, SZ, I64, I64, I64, P, D, D, D, C.c_int, C.c_int, C.c_int, P,
                     C.POINTER(C.c_void_p))
mapper_destroy = _sig("tdk_mapper_destroy", None, P)
mapper_reset = _sig("tdk_mapper_reset", C.c_int, P, P, C.c_int, P, P, P, P)
mapper_track = _sig("tdk_mapper_track", C.c_int, P, P, C.c_int, P, P, P)
mapper_sync = _sig("tdk_mapper_sync", C.c_int, P)
mapper_prune = _sig("tdk_mapper_prune", C.c_int, P)
mapper_state = _sig("tdk_mapper_state", C.c_int, P, C.POINTER(I64), C.POINTER(SZ), C.POINTER(SZ), C.POINTER(SZ),
                    C.POINTER(SZ), C.POINTER(SZ), C.POINTER(I64))
IMAGE_F64, IMAGE_U8, IMAGE_RGB8, IMAGE_RGBA8 = 0, 1, 2, 3
warp_vecs_dev = _sig("tdk_warp_vecs_dev", C.c_int, P, P, P, I64, P, P, P)
interpolate_dev = _sig("tdk_interpolate_dev", C.c_int, P, I64, I64, P, I64, P, P, P)
project_vecs_dev = _sig("tdk_project_vecs_dev", C.c_int, P, I64, P, P)
inv_project_vecs_dev = _sig("tdk_inv_project_vecs_dev", C.c_int, P, P, I64, P, P)
transform_dev = _sig("tdk_transform_dev", C.c_int, P, P, I64, P, P)
calc_depth0_dev = _sig("tdk_calc_depth0_dev", C.c_int, P, P, P, I64, P, P)
warp_vecs_host = _sig("tdk_warp_vecs_host", C.c_int, P, P, P, I64, P, P)
interpolate_host = _sig("tdk_interpolate_host", C.c_int, P, I64, I64, P, I64, P)
project_vecs_host = _sig("tdk_project_vecs_host", C.c_int, P, I64, P)
inv_project_vecs_host = _sig("tdk_inv_project_vecs_host", C.c_int, P, P, I64, P)
transform_host = _sig("tdk_transform_host", C.c_int, P, P, I64, P)
calc_depth0_host = _sig("tdk_calc_depth0_host", C.c_int, P, P, P, I64, P)
age_usage_dev = _sig("tdk_age_usage_dev", C.c_int, P, I64, I64, I64, I64, P, P)
render_scene_dev = _sig("tdk_render_scene_dev", C.c_int, P, P, P, I64, C.c_double, P, P, I64, I64, I64, P, P, P)
sqrt_threshold = _sig("tdk_sqrt_threshold", C.c_double, C.c_double)
selftest_shared_division_dev = _sig("tdk_selftest_shared_division_dev", C.c_int, I64, C.c_uint64, P, P)

image_gradient_dev = _sig("tdk_image_gradient_dev", C.c_int, P, I64, I64, P, P, P)
dvo_normal_equations_dev = _sig("tdk_dvo_normal_equations_dev", C.c_int, P, P, P, P, P, P, C.c_int, D, D,
                                P, P, P, I64, I64, P, P)
dvo_student_t_dev = _sig("tdk_dvo_student_t_dev", C.c_int, P, P, P, D, D, P, P, P, I64, I64, P, P)
dvo_photometric_error_dev = _sig("tdk_dvo_photometric_error_dev", C.c_int, P, P, P, P, P, P, I64, I64, P,
                                 P)

dvo_iterate_dev = _sig("tdk_dvo_iterate_dev", C.c_int, P, P, P, P, P, P, C.c_int, D, P, P, P, I64, I64, C.c_int, C.c_int,
                       P, P)
gaussian_blur_dev = _sig("tdk_gaussian_blur_dev", C.c_int, P, I64, I64, P, C.c_int, P, P, P)
resample_bilinear_dev = _sig("tdk_resample_bilinear_dev", C.c_int, P, I64, I64, D, P, I64, I64, P)
dvo_tukey_workspace_bytes = _sig("tdk_dvo_tukey_workspace_bytes", SZ, I64, I64)
dvo_tukey_scale_dev = _sig("tdk_dvo_tukey_scale_dev", C.c_int, P, P, P, D, P, P, P, I64, I64, P, P, SZ, P)

# every symbol include/tdk.h declares (tests check the header against this list and the .so)
EXPORTS = [
    "tdk_last_error", "tdk_version", "tdk_set_kernel_plan", "tdk_update_depth_workspace_bytes",
    "tdk_update_depth_workspace_bytes_min",
    "tdk_update_depth_dev", "tdk_update_depth_frames_dev", "tdk_update_depth_tables_dev", "tdk_update_depth_status",
    "tdk_update_depth_static_dev", "tdk_selftest_host_wire", "tdk_update_depth_maps_host",
    "tdk_update_depth_host", "tdk_estimate_debug_host", "tdk_fusion_dev", "tdk_regularize_dev",
    "tdk_increment_age_workspace_bytes", "tdk_increment_age_dev", "tdk_propagate_workspace_bytes",
    "tdk_propagate_dev", "tdk_warp_vecs_dev", "tdk_interpolate_dev", "tdk_project_vecs_dev",
    "tdk_inv_project_vecs_dev", "tdk_transform_dev", "tdk_calc_depth0_dev",
    "tdk_selftest_shared_division_dev", "tdk_render_scene_dev", "tdk_age_usage_dev", "tdk_propagate_age_dev", "tdk_sqrt_threshold",
    "tdk_image_gradient_dev", "tdk_dvo_normal_equations_dev", "tdk_dvo_student_t_dev",
    "tdk_dvo_photometric_error_dev", "tdk_dvo_iterate_dev", "tdk_gaussian_blur_dev", "tdk_resample_bilinear_dev",
    "tdk_increment_age_host", "tdk_propagate_host", "tdk_fusion_host", "tdk_regularize_host",
    "tdk_mapper_arena_bytes", "tdk_mapper_create", "tdk_mapper_destroy", "tdk_mapper_reset", "tdk_mapper_track",
    "tdk_mapper_sync", "tdk_mapper_prune", "tdk_mapper_state",
    "tdk_update_depth_frames_host", "tdk_warp_vecs_host", "tdk_interpolate_host", "tdk_project_vecs_host",
    "tdk_inv_project_vecs_host", "tdk_transform_host", "tdk_calc_depth0_host", "tdk_dvo_tukey_workspace_bytes", "tdk_dvo_tukey_scale_dev", "tdk_set_subpixel_refinement",
]


class TdkError(RuntimeError):
    pass


def check(rc):
    """Map a C-ABI status to the exception the reference-facing API documents."""
    if rc == OK:
        return
    msg = (last_error() or b"").decode()
    if rc in (ERR_AGE, ERR_SINGULAR, ERR_SEARCH_LEN, ERR_RANGE, ERR_INVALID):
        raise ValueError(msg)
    raise TdkError(f"libtdk_b200 error {rc}: {msg}")
