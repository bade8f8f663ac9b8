"""ctypes binding of libfvsrn_b200.so (include/fvsrn_b200.h). Fails loudly when the library is missing."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfvsrn_b200.so")

# every symbol include/fvsrn_b200.h declares (tests/test_abi.py checks the library exports all of them)
SYMBOLS = [
    "fvsrn_abi_version", "fvsrn_last_error", "fvsrn_network_create", "fvsrn_network_free", "fvsrn_network_load_volnet",
    "fvsrn_network_save_volnet", "fvsrn_network_set_input", "fvsrn_network_set_fourier", "fvsrn_network_disable_fourier",
    "fvsrn_network_set_output_mode", "fvsrn_network_add_layer", "fvsrn_network_set_latent_grid_layout",
    "fvsrn_network_set_time_grid", "fvsrn_network_set_ensemble_grid", "fvsrn_network_set_box", "fvsrn_network_get_box",
    "fvsrn_network_valid", "fvsrn_network_num_parameters", "fvsrn_network_compute_max_warps",
    "fvsrn_network_set_time_and_ensemble", "fvsrn_network_set_gradient_mode", "fvsrn_network_get_info", "fvsrn_network_set_kernel_mode", "fvsrn_network_get_kernel_mode", "fvsrn_network_prepare",
    "fvsrn_network_clear_gpu_resources", "fvsrn_evaluate", "fvsrn_evaluate_host", "fvsrn_render_dvr",
    "fvsrn_render_dvr_host", "fvsrn_extract_color", "fvsrn_extract_color_ex", "fvsrn_extract_color_ex_noalloc", "fvsrn_generate_rays", "fvsrn_evaluate_tf", "fvsrn_query",
    "fvsrn_volume_load_cvol", "fvsrn_volume_save_cvol", "fvsrn_volume_from_dense", "fvsrn_volume_get_info", "fvsrn_volume_free",
    "fvsrn_volume_prepare", "fvsrn_render_dvr_grid", "fvsrn_evaluate_grid",
    "fvsrn_peer_buffer_create", "fvsrn_peer_buffer_open", "fvsrn_peer_buffer_close", "fvsrn_peer_buffer_free",
]


class NetworkInfo(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "has_time", "has_direction", "num_fourier", "direction_in_fourier", "output_mode", "num_layers",
        "hidden_channels", "latent_channels", "grid_x", "grid_y", "grid_z", "time_min", "time_num", "time_step",
        "ensemble_min", "ensemble_num", "grid_encoding", "activation")] + [
        ("activation_param", C.c_float), ("channels_out", C.c_int), ("gradient_mode", C.c_int),
        ("finite_differences_stepsize", C.c_float)]


class DvrParams(C.Structure):
    _fields_ = [
        ("width", C.c_int), ("height", C.c_int), ("batch", C.c_int),
        ("tile_x0", C.c_int), ("tile_y0", C.c_int), ("tile_w", C.c_int), ("tile_h", C.c_int),
        ("camera_matrix_dev", C.c_void_p), ("fov_y_radians", C.c_float),
        ("stepsize", C.c_float), ("alpha_early_out", C.c_float), ("density_min", C.c_float), ("density_max", C.c_float),
        ("early_out", C.c_int), ("blend_mode", C.c_int),
        ("tf_kind", C.c_int), ("tf_dev", C.c_void_p), ("tf_resolution", C.c_int), ("tf_batch", C.c_int),
        ("samples_per_pixel", C.c_int), ("view_times_host", C.c_void_p),
        ("out_stride_b", C.c_int64), ("out_stride_c", C.c_int64), ("out_stride_y", C.c_int64), ("out_stride_x", C.c_int64),
        ("step_count_dev", C.c_void_p), ("counters_dev", C.c_void_p),
        ("output_format", C.c_int), ("channel_mode", C.c_int), ("tonemap", C.c_int), ("max_exposure", C.c_float),
        ("brdf_flags", C.c_int), ("light_type", C.c_int), ("light", C.c_float * 3),
        ("brdf_magnitude_scaling", C.c_float), ("brdf_ambient", C.c_float), ("brdf_specular", C.c_float),
        ("brdf_magnitude_center", C.c_float), ("brdf_magnitude_radius", C.c_float), ("brdf_specular_exponent", C.c_int),
    ]


class VolumeInfo(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("num_features", "feature", "size_x", "size_y", "size_z", "channels", "data_type")] + [
        ("world_size", C.c_float * 3), ("box_min", C.c_float * 3), ("box_size", C.c_float * 3)]


class QueryInfo(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "device", "sm_count", "block_threads", "blocks_per_sm", "grid_blocks", "smem_bytes_per_block", "regs_per_thread",
        "rays_per_warp_tile_x", "rays_per_warp_tile_y")] + [("kernel_launches", C.c_ulonglong)]


_lib = None


def lib():
    """Loads libfvsrn_b200.so. There is no Python/CPU fallback: a missing library is an error."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C fv-srn_b200/csrc`). The fV-SRN B200 path has no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, i, f, i64 = C.c_void_p, C.c_int, C.c_float, C.c_int64
    L.fvsrn_last_error.restype = C.c_char_p
    sig = {
        "fvsrn_network_create": [C.POINTER(vp)],
        "fvsrn_network_free": [vp],
        "fvsrn_network_load_volnet": [C.c_char_p, C.c_size_t, C.POINTER(vp)],
        "fvsrn_network_save_volnet": [vp, vp, C.c_size_t, C.POINTER(C.c_size_t)],
        "fvsrn_network_set_input": [vp, i, i],
        "fvsrn_network_set_fourier": [vp, vp, i, i, i],
        "fvsrn_network_disable_fourier": [vp],
        "fvsrn_network_set_output_mode": [vp, C.c_char_p],
        "fvsrn_network_add_layer": [vp, vp, vp, i, i, C.c_char_p, f],
        "fvsrn_network_set_latent_grid_layout": [vp, i, i, i, i, i],
        "fvsrn_network_set_time_grid": [vp, i, vp, i, i, i, i, i, C.POINTER(C.c_double)],
        "fvsrn_network_set_ensemble_grid": [vp, i, vp, i, i, i, i, i, C.POINTER(C.c_double)],
        "fvsrn_network_set_box": [vp, C.POINTER(f * 3), C.POINTER(f * 3)],
        "fvsrn_network_get_box": [vp, C.POINTER(f * 3), C.POINTER(f * 3)],
        "fvsrn_network_valid": [vp],
        "fvsrn_network_num_parameters": [vp],
        "fvsrn_network_compute_max_warps": [vp, i, i],
        "fvsrn_network_set_time_and_ensemble": [vp, f, i],
        "fvsrn_network_set_gradient_mode": [vp, i, f],
        "fvsrn_network_get_info": [vp, C.POINTER(NetworkInfo)],
        "fvsrn_network_set_kernel_mode": [vp, i],
        "fvsrn_network_get_kernel_mode": [vp],
        "fvsrn_network_prepare": [vp, i, vp],
        "fvsrn_network_clear_gpu_resources": [vp],
        "fvsrn_evaluate": [vp, vp, vp, vp, i64, vp],
        "fvsrn_evaluate_host": [vp, vp, vp, vp, i64, vp],
        "fvsrn_render_dvr": [vp, C.POINTER(DvrParams), vp, vp],
        "fvsrn_render_dvr_host": [vp, C.POINTER(DvrParams), vp, vp],
        "fvsrn_extract_color": [vp, vp, i, i, i, vp],
        "fvsrn_extract_color_ex": [vp, vp, i, i, i, i, i, f, i, vp],
        "fvsrn_extract_color_ex_noalloc": [vp, vp, i, i, i, i, i, f, i, vp],
        "fvsrn_generate_rays": [vp, i, f, i, i, vp, vp, vp],
        "fvsrn_volume_load_cvol": [C.c_char_p, C.c_size_t, C.POINTER(vp)],
        "fvsrn_volume_save_cvol": [vp, vp, C.c_size_t, C.POINTER(C.c_size_t)],
        "fvsrn_volume_from_dense": [vp, i, i, i, C.POINTER(f * 3), C.POINTER(vp)],
        "fvsrn_volume_get_info": [vp, i, C.POINTER(VolumeInfo)],
        "fvsrn_volume_free": [vp],
        "fvsrn_volume_prepare": [vp, i, i, vp],
        "fvsrn_render_dvr_grid": [vp, C.POINTER(DvrParams), vp, vp],
        "fvsrn_evaluate_grid": [vp, vp, vp, i64, vp],
        "fvsrn_evaluate_tf": [i, vp, i, vp, vp, i64, f, f, f, vp],
        "fvsrn_query": [vp, i, C.POINTER(QueryInfo)],
        "fvsrn_peer_buffer_create": [C.c_size_t, C.POINTER(vp), C.c_char_p],
        "fvsrn_peer_buffer_open": [C.c_char_p, C.POINTER(vp)],
        "fvsrn_peer_buffer_close": [vp],
        "fvsrn_peer_buffer_free": [vp],
    }
    for name, args in sig.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = None if name in ("fvsrn_network_free", "fvsrn_volume_free") else C.c_int
    _lib = L
    return L


def check(status: int):
    """Non-zero status -> RuntimeError, like the reference's C++ exceptions surfacing through pybind11."""
    if status != 0:
        raise RuntimeError(lib().fvsrn_last_error().decode("utf-8", "replace"))
