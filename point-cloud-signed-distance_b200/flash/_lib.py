"""Loads libflashsdf.so (the CUDA C-ABI library) and declares its prototypes.

Fails loudly: there is no Python/CPU implementation behind these calls.
"""
import ctypes as C
import os

from ._desc import ModelDesc, Options, SolverParams, Timing, c_double_p, c_int32_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FLASHSDF_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "libflashsdf.so")

# every symbol include/flashsdf.h declares
SYMBOLS = [
    "flash_options_default", "flash_abi_version", "flash_last_error", "flash_model_create",
    "flash_destroy", "flash_num_positions", "flash_num_states", "flash_num_surfaces",
    "flash_points_set", "flash_points_set_f32", "flash_points_set_device", "flash_num_points",
    "flash_cost_grad", "flash_skin_eval", "flash_raycast", "flash_kinect_rays",
    "flash_estimate_state", "flash_comm_unique_id", "flash_comm_init", "flash_comm_world",
    "flash_profiling", "flash_last_timing", "flash_launch_count", "flash_device_info",
    "flash_grid_report", "flash_grid_download", "flash_contour_mesh",
    "flash_points_set_async", "flash_points_set_async_f32", "flash_points_swap", "flash_points_sync",
    "flash_frames_set", "flash_frames_set_async", "flash_num_frames", "flash_cost_grad_frames",
    "flash_urdf_load", "flash_urdf_desc", "flash_urdf_body_name", "flash_urdf_surface_mesh", "flash_urdf_set_joint_type", "flash_urdf_free",
    "flash_model_load_urdf", "flash_read_point_cloud", "flash_points_set_file", "flash_points_set_file_async", "flash_num_devices",
    "flash_skin_eval_static",
]

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "libflashsdf.so not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C point-cloud-signed-distance_b200`. There is no CPU fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.flash_options_default.argtypes = [C.POINTER(Options)]; L.flash_options_default.restype = None
    L.flash_abi_version.restype = C.c_int
    L.flash_last_error.argtypes = [vp]; L.flash_last_error.restype = C.c_char_p
    L.flash_model_create.argtypes = [C.POINTER(ModelDesc), C.POINTER(Options), C.POINTER(vp)]
    L.flash_destroy.argtypes = [vp]; L.flash_destroy.restype = None
    for f in ("flash_num_positions", "flash_num_states", "flash_num_surfaces", "flash_comm_world", "flash_num_devices"):
        getattr(L, f).argtypes = [vp]; getattr(L, f).restype = C.c_int
    L.flash_points_set.argtypes = [vp, c_double_p, C.c_int64]
    L.flash_points_set_f32.argtypes = [vp, C.POINTER(C.c_float), C.c_int64]
    L.flash_points_set_device.argtypes = [vp, vp, C.c_int64]
    L.flash_urdf_load.argtypes = [C.c_char_p, C.POINTER(C.c_char_p), C.c_int32, C.POINTER(vp)]
    L.flash_urdf_desc.argtypes = [vp]; L.flash_urdf_desc.restype = C.POINTER(ModelDesc)
    L.flash_urdf_body_name.argtypes = [vp, C.c_int32]; L.flash_urdf_body_name.restype = C.c_char_p
    L.flash_urdf_surface_mesh.argtypes = [vp, C.c_int32]; L.flash_urdf_surface_mesh.restype = C.c_char_p
    L.flash_urdf_set_joint_type.argtypes = [vp, C.c_int32, C.c_int32]
    L.flash_urdf_free.argtypes = [vp]; L.flash_urdf_free.restype = None
    L.flash_model_load_urdf.argtypes = [C.c_char_p, C.POINTER(C.c_char_p), C.c_int32, C.POINTER(Options), C.POINTER(vp)]
    L.flash_read_point_cloud.argtypes = [C.c_char_p, c_double_p, c_double_p, c_double_p, C.c_int64, C.POINTER(C.c_int64)]
    L.flash_points_set_file.argtypes = [vp, C.c_char_p, c_double_p]
    L.flash_points_set_file_async.argtypes = [vp, C.c_char_p, c_double_p]
    L.flash_points_set_async.argtypes = [vp, c_double_p, C.c_int64]
    L.flash_points_set_async_f32.argtypes = [vp, C.POINTER(C.c_float), C.c_int64]
    L.flash_points_swap.argtypes = [vp]
    L.flash_points_sync.argtypes = [vp]
    L.flash_frames_set.argtypes = [vp, C.POINTER(c_double_p), C.POINTER(C.c_int64), C.c_int32]
    L.flash_frames_set_async.argtypes = [vp, C.POINTER(c_double_p), C.POINTER(C.c_int64), C.c_int32]
    L.flash_num_frames.argtypes = [vp]
    L.flash_cost_grad_frames.argtypes = [vp, c_double_p, C.c_int32, C.c_double, c_double_p, c_double_p]
    L.flash_num_points.argtypes = [vp]; L.flash_num_points.restype = C.c_int64
    L.flash_cost_grad.argtypes = [vp, c_double_p, C.c_int32, C.c_double, c_double_p, c_double_p, c_int32_p]
    L.flash_skin_eval.argtypes = [vp, c_double_p, c_double_p, C.c_int64, c_double_p, c_int32_p]
    L.flash_skin_eval_static.argtypes = [vp, vp, vp, C.c_int64, vp, vp]   # host or device addresses
    L.flash_raycast.argtypes = [vp, c_double_p, c_double_p, C.c_int64, c_double_p, c_double_p, c_double_p, C.POINTER(C.c_int64)]
    L.flash_kinect_rays.argtypes = [C.c_int32, C.c_int32, C.c_double, C.c_double, c_double_p]
    L.flash_estimate_state.argtypes = [vp, c_double_p, C.c_int32, C.POINTER(SolverParams), c_double_p, c_double_p, c_int32_p]
    L.flash_comm_unique_id.argtypes = [vp]
    L.flash_comm_init.argtypes = [vp, C.c_int32, C.c_int32, vp]
    L.flash_profiling.argtypes = [vp, C.c_int32]
    L.flash_last_timing.argtypes = [vp, C.POINTER(Timing)]
    L.flash_launch_count.argtypes = [vp]; L.flash_launch_count.restype = C.c_int64
    L.flash_device_info.argtypes = [vp, c_int32_p, c_int32_p, C.POINTER(C.c_int64)]
    L.flash_grid_report.argtypes = [vp, C.c_int32, c_int32_p, c_double_p, c_double_p]
    L.flash_grid_download.argtypes = [vp, C.c_int32, c_double_p, c_double_p, C.POINTER(C.c_float)]
    L.flash_contour_mesh.argtypes = [vp, c_double_p, C.c_int32, c_double_p, c_double_p, C.c_double, C.c_double, C.POINTER(C.c_float), C.c_int64,
                                     C.POINTER(C.c_int64)]
    _lib = L
    return L
