"""ctypes loader for libngp_b200.so (the C ABI declared in include/ngp_b200.h).

The product path has NO fallback: if the shared library is missing or a call fails, a RuntimeError is raised
(the reference raises std::runtime_error through pybind11 -> RuntimeError, main.cu:195-209).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NGP_B200_LIB") or os.path.join(_HERE, "libngp_b200.so")   # NGP_B200_LIB: A/B timing builds (csrc/Makefile target `ab`)

_lib = None


class NgpError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NgpError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
        _lib = C.CDLL(LIB_PATH)
        _declare(_lib)
    return _lib


def check(status):
    if status != 0:
        raise NgpError(lib().ngp_last_error().decode())


u32, u64, i32, f32, vp, cp = C.c_uint32, C.c_uint64, C.c_int, C.c_float, C.c_void_p, C.c_char_p


class AdamHyper(C.Structure):
    _fields_ = [("learning_rate", f32), ("beta1", f32), ("beta2", f32), ("epsilon", f32), ("l2_reg", f32), ("loss_scale", f32),
                ("n_matrix_params", u64), ("optimize_matrix_params", C.c_int32), ("optimize_non_matrix_params", C.c_int32),
                ("ema_decay", f32), ("ema_step", u32)]


class TrainingImage(C.Structure):
    _fields_ = [("pixels", vp), ("data_type", C.c_int32), ("res", C.c_int32 * 2), ("focal_length", f32 * 2), ("principal_point", f32 * 2),
                ("lens_mode", C.c_int32), ("lens_params", f32 * 4), ("xform", f32 * 12)]


class TrainStats(C.Structure):
    _fields_ = [("training_step", u32), ("loss", f32), ("rays_per_batch", u32), ("measured_batch_size", u32),
                ("measured_batch_size_before_compaction", u32), ("n_rays_total", u32), ("learning_rate", f32), ("density_grid_mean", f32)]


# name -> (restype, argtypes); every symbol include/ngp_b200.h declares
SYMBOLS = {
    "ngp_last_error": (cp, []),
    "ngp_abi_version": (i32, []),
    "ngp_kernel_launch_count": (u64, []),
    "ngp_profile_enable": (None, [i32]),
    "ngp_profile_report": (i32, [cp, u64, i32]),
    "ngp_profile_timeline": (i32, [cp, u64]),
    "ngp_microbench_l2_gather": (i32, [vp, u64, u32, C.POINTER(C.c_double)]),
    "ngp_microbench_l2_red": (i32, [vp, u64, i32, u32, C.POINTER(C.c_double)]),
    "ngp_host_has_mask_pixels": (i32, [vp, u64, C.POINTER(i32)]),
    "ngp_encoding_create": (i32, [cp, u32, C.POINTER(vp)]),
    "ngp_encoding_destroy": (None, [vp]),
    "ngp_encoding_n_params": (u64, [vp]),
    "ngp_encoding_padded_output_width": (u32, [vp]),
    "ngp_encoding_level_table": (u32, [vp, C.POINTER(u32), C.POINTER(f32), C.POINTER(u32)]),
    "ngp_encoding_initialize_params": (i32, [vp, vp, u64, vp]),
    "ngp_encoding_forward": (i32, [vp, vp, vp, vp, u32, u32, vp, i32]),
    "ngp_encoding_backward": (i32, [vp, vp, vp, u32, u32, vp, i32, vp]),
    "ngp_encoding_backward_input": (i32, [vp, vp, vp, vp, u32, u32, vp, i32, vp]),
    "ngp_encoding_corners": (i32, [vp, vp, vp, u32, u32, u32, vp, vp]),
    "ngp_nerf_network_create": (i32, [cp, C.POINTER(vp)]),
    "ngp_nerf_network_destroy": (None, [vp]),
    "ngp_nerf_network_n_params": (u64, [vp]),
    "ngp_nerf_network_layout": (i32, [vp, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]),
    "ngp_nerf_network_initialize_params": (i32, [vp, vp, u64, vp]),
    "ngp_nerf_network_inference": (i32, [vp, vp, vp, vp, u32, vp]),
    "ngp_nerf_network_inference_noalloc": (i32, [vp, vp, vp, vp, u32, vp]),
    "ngp_noalloc_max_samples": (u32, []),
    "ngp_nerf_network_density": (i32, [vp, vp, vp, vp, u32, u32, vp]),
    "ngp_nerf_network_forward_backward": (i32, [vp, vp, vp, vp, u32, vp, vp]),
    "ngp_nerf_network_forward_backward_input": (i32, [vp, vp, vp, vp, u32, vp, vp, vp]),
    "ngp_encoded_mlp_create": (i32, [cp, u32, u32, C.POINTER(vp)]),
    "ngp_encoded_mlp_destroy": (None, [vp]),
    "ngp_encoded_mlp_n_params": (u64, [vp]),
    "ngp_encoded_mlp_layout": (i32, [vp, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]),
    "ngp_encoded_mlp_initialize_params": (i32, [vp, vp, u64, vp]),
    "ngp_encoded_mlp_inference": (i32, [vp, vp, vp, vp, u32, u32, vp]),
    "ngp_encoded_mlp_forward_backward": (i32, [vp, vp, vp, vp, u32, u32, vp, vp]),
    "ngp_adam_step": (i32, [vp, C.POINTER(AdamHyper), u64, vp, vp, vp, vp, vp, vp, vp, vp]),
    "ngp_encoding_set_max_level_gpu": (i32, [vp, vp]),
    "ngp_testbed_create": (i32, [i32, C.POINTER(vp)]),
    "ngp_testbed_destroy": (None, [vp]),
    "ngp_testbed_reload_network_from_json": (i32, [vp, cp]),
    "ngp_testbed_reload_network_from_json_base": (i32, [vp, cp, cp]),
    "ngp_testbed_reload_network_from_file": (i32, [vp, cp]),
    "ngp_testbed_set_root_dir": (i32, [vp, cp]),
    "ngp_builtin_base_config": (i32, [C.c_int32, C.POINTER(cp)]),
    "ngp_mode_from_scene": (i32, [cp, C.POINTER(C.c_int32)]),
    "ngp_snapshot_mode": (i32, [cp, C.POINTER(C.c_int32)]),
    "ngp_testbed_set_image_rgba8": (i32, [vp, vp, vp, C.c_int32, C.c_int32]),
    "ngp_testbed_set_nerf_dataset": (i32, [vp, vp, u32, vp, C.c_int32]),
    "ngp_nerf_dataset_open": (i32, [cp, C.POINTER(vp)]),
    "ngp_nerf_dataset_close": (None, [vp]),
    "ngp_nerf_dataset_n_images": (u32, [vp]),
    "ngp_nerf_dataset_image_path": (i32, [vp, u32, cp, u64]),
    "ngp_nerf_dataset_set_image": (i32, [vp, u32, vp, C.c_int32, C.c_int32, C.c_int32]),
    "ngp_nerf_dataset_depth_path": (i32, [vp, u32, cp, u64]),
    "ngp_nerf_dataset_set_depth": (i32, [vp, u32, vp, C.c_int32, C.c_int32]),
    "ngp_testbed_training_view_motion": (i32, [vp, u32, vp, vp, vp]),
    "ngp_testbed_get_camera_extrinsics": (i32, [vp, C.c_int32, C.c_int32, vp]),
    "ngp_testbed_set_training_depth": (i32, [vp, vp, C.c_int32, vp, C.c_int32, f32]),
    "ngp_nerf_dataset_info": (i32, [vp, C.POINTER(C.c_int32), C.POINTER(f32), C.POINTER(f32), C.POINTER(C.c_int32)]),
    "ngp_nerf_dataset_view": (i32, [vp, u32, C.POINTER(TrainingImage)]),
    "ngp_nerf_dataset_view_motion": (i32, [vp, u32, vp, vp]),
    "ngp_testbed_load_nerf_dataset": (i32, [vp, vp, vp]),
    "ngp_testbed_set_training_image": (i32, [vp, vp, u32, vp]),
    "ngp_testbed_set_camera_intrinsics": (i32, [vp, vp, C.c_int32, f32, f32, f32, f32, f32, f32, f32, f32]),
    "ngp_testbed_set_camera_extrinsics": (i32, [vp, vp, C.c_int32, vp, C.c_int32]),
    "ngp_testbed_set_camera_extrinsics_rolling_shutter": (i32, [vp, vp, C.c_int32, vp, vp, vp, C.c_int32]),
    "ngp_testbed_reset": (i32, [vp, C.c_int32]),
    "ngp_testbed_set_image": (i32, [vp, vp, vp, C.c_int32, C.c_int32]),
    "ngp_testbed_override_sdf_training_data": (i32, [vp, vp, vp, vp, u32, vp, vp, f32]),
    "ngp_testbed_train": (i32, [vp, vp, u32]),
    "ngp_testbed_train_begin": (i32, [vp, vp, u32]),
    "ngp_testbed_train_end": (i32, [vp, vp, u32, u32]),
    "ngp_testbed_set_shard": (i32, [vp, u32, u32]),
    "ngp_testbed_pending_counters": (i32, [vp, vp, vp]),
    "ngp_comm_unique_id": (i32, [vp, C.c_char_p]),
    "ngp_testbed_sync_inference_params": (i32, [vp, vp]),
    "ngp_testbed_comm_init": (i32, [vp, vp, u32, u32, C.c_char_p]),
    "ngp_testbed_stats": (i32, [vp, C.POINTER(TrainStats)]),
    "ngp_testbed_n_params": (u64, [vp]),
    "ngp_testbed_n_encoding_params": (u64, [vp]),
    "ngp_json_to_msgpack": (i32, [cp, vp, u64, C.POINTER(u64)]),
    "ngp_msgpack_roundtrip": (i32, [vp, u64, vp, u64, C.POINTER(u64)]),
    "ngp_testbed_save_snapshot": (i32, [vp, vp, cp, C.c_int32, C.c_int32]),
    "ngp_testbed_load_snapshot": (i32, [vp, vp, cp]),
    "ngp_testbed_param_buffer": (i32, [vp, i32, C.POINTER(vp), C.POINTER(u64)]),
    "ngp_testbed_density_grid": (i32, [vp, i32, C.POINTER(vp), C.POINTER(u64)]),
    "ngp_testbed_render": (i32, [vp, vp, vp, vp, vp, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, vp, C.c_int32, vp]),
    "ngp_testbed_set_render_lens": (i32, [vp, C.c_int32, vp]),
    "ngp_testbed_render_frame": (i32, [vp, vp, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, vp, C.c_int32]),
    "ngp_testbed_reset_camera": (i32, [vp]),
    "ngp_testbed_set_nerf_camera_matrix": (i32, [vp, vp]),
    "ngp_testbed_set_camera_to_training_view": (i32, [vp, C.c_int32]),
    "ngp_testbed_set_vector": (i32, [vp, cp, vp, u32]),
    "ngp_testbed_get_vector": (i32, [vp, cp, vp, u32]),
    "ngp_testbed_training_view": (i32, [vp, u32, C.POINTER(TrainingImage)]),
    "ngp_testbed_n_training_views": (u32, [vp]),
    "ngp_testbed_compute_image_mse": (i32, [vp, vp, C.c_int32, C.POINTER(f32)]),
    "ngp_testbed_set_option": (i32, [vp, cp, C.c_double]),
    "ngp_testbed_get_option": (i32, [vp, cp, C.POINTER(C.c_double)]),
    "ngp_testbed_debug_buffer": (i32, [vp, cp, C.POINTER(vp), C.POINTER(u64)]),
    "ngp_synth_camera": (i32, [u32, u32, f32, C.POINTER(f32)]),
    "ngp_synth_render_view": (i32, [vp, u32, u32, f32, f32, C.c_int32, C.c_int32, vp]),
}


def _declare(l):
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(l, name)   # AttributeError here == ABI drift: fail loudly
        fn.restype = res
        fn.argtypes = args
