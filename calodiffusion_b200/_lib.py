"""ctypes binding of libcalodiffusion_b200.so (C ABI declared in include/calodiffusion_b200.h)."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcalodiffusion_b200.so")

CD_MAX_LEVELS = 4
CD_NUM_STATE_BUFS = 4

EXPORTS = [
    "cd_last_error", "cd_version", "cd_device_check", "cd_unet_create", "cd_unet_destroy", "cd_unet_set_weight",
    "cd_unet_finalize", "cd_unet_set_option", "cd_unet_workspace_bytes", "cd_unet_launch_count", "cd_unet_forward",
    "cd_denoise", "cd_sample_run", "cd_fill_normal", "cd_reverse_norm", "cd_test_conv", "cd_level_shape", "cd_unet_profile_report",
    "cd_unet_set_geom_map", "cd_unet_voxels", "cd_geom_apply", "cd_layer_model_create", "cd_reverse_norm_ex", "cd_geom_decode_sparse",
    "cd_train_step", "cd_train_loss", "cd_train_adam", "cd_train_refresh", "cd_train_flat_size", "cd_train_num_params",
    "cd_train_param_info", "cd_train_buffer", "cd_conv_noalloc", "cd_conv_noalloc_max_bytes",
]


class LayerDesc(C.Structure):
    _fields_ = [("dim_in", C.c_int), ("hidden", C.c_int), ("num_layers", C.c_int), ("cond_size", C.c_int), ("cond_dim", C.c_int),
                ("time_embed", C.c_int), ("sigma_data", C.c_float), ("max_batch", C.c_int)]


class UnetDesc(C.Structure):
    _fields_ = [
        ("Z", C.c_int), ("P", C.c_int), ("R", C.c_int),
        ("n_sizes", C.c_int), ("layer_sizes", C.c_int * (CD_MAX_LEVELS + 1)),
        ("in_channels", C.c_int), ("rz_input", C.c_int), ("phi_input", C.c_int),
        ("cond_size", C.c_int), ("cond_dim", C.c_int), ("groups", C.c_int), ("dataset_num", C.c_int),
        ("time_embed", C.c_int), ("sigma_data", C.c_float), ("max_batch", C.c_int),
        ("r_centres", C.c_float * 64),
    ]


class Eval(C.Structure):
    _fields_ = [
        ("sigma", C.c_float),
        ("x_in", C.c_int), ("x_out", C.c_int), ("x0_buf", C.c_int),
        ("k", C.c_float * 5),
        ("a1", C.c_int), ("a2", C.c_int), ("a3", C.c_int),
        ("noise_index", C.c_int),
        ("pre_a", C.c_float), ("pre_b", C.c_float), ("pre_noise_index", C.c_int),
        ("rec_x0", C.c_int), ("rec_x", C.c_int), ("rec_x_buf", C.c_int),
        ("pre_src", C.c_int), ("d_out", C.c_int), ("d_scale", C.c_float),
    ]


class ReverseNormDesc(C.Structure):
    _fields_ = [
        ("Z", C.c_int), ("PR", C.c_int), ("logit_mean", C.c_float), ("logit_std", C.c_float),
        ("totalE_mean", C.c_float), ("totalE_std", C.c_float), ("layers_mean", C.c_float), ("layers_std", C.c_float),
        ("emin", C.c_float), ("emax", C.c_float), ("max_deposit", C.c_float), ("ecut", C.c_float),
    ]


class ReverseNormExDesc(C.Structure):
    _fields_ = [
        ("nseg", C.c_int), ("total", C.c_int), ("map_kind", C.c_int), ("a_scale", C.c_float), ("a_shift", C.c_float),
        ("alpha", C.c_float), ("log_cap", C.c_float), ("post_scale", C.c_float), ("post_shift", C.c_float),
        ("totalE_mean", C.c_float), ("totalE_std", C.c_float), ("layers_mean", C.c_float), ("layers_std", C.c_float),
        ("guard_eps", C.c_float), ("energy_mode", C.c_int), ("emin", C.c_float), ("emax", C.c_float),
        ("max_deposit", C.c_float), ("ecut", C.c_float), ("e_stride", C.c_int), ("final_scale", C.c_int),
    ]


class LibraryMissing(RuntimeError):
    pass


_lib = None


def load():
    """Load the CUDA library; raises LibraryMissing (never falls back to another implementation)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LibraryMissing(
            f"{LIB_PATH} not found: build it with calodiffusion_b200/csrc/build.sh (or __graft_entry__.build()). "
            "calodiffusion_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, u64, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_float
    lib.cd_last_error.restype = C.c_char_p
    lib.cd_version.restype = i32
    lib.cd_device_check.restype = i32
    lib.cd_unet_create.restype = vp
    lib.cd_unet_create.argtypes = [C.POINTER(UnetDesc)]
    lib.cd_layer_model_create.restype = vp
    lib.cd_layer_model_create.argtypes = [C.POINTER(LayerDesc)]
    lib.cd_unet_destroy.argtypes = [vp]
    lib.cd_unet_destroy.restype = None
    lib.cd_unet_set_weight.argtypes = [vp, C.c_char_p, vp, C.POINTER(i64), i32]
    lib.cd_unet_finalize.argtypes = [vp]
    lib.cd_unet_set_option.argtypes = [vp, C.c_char_p, i32]
    lib.cd_unet_workspace_bytes.argtypes = [vp]
    lib.cd_unet_workspace_bytes.restype = i64
    lib.cd_unet_launch_count.argtypes = [vp]
    lib.cd_unet_launch_count.restype = i64
    lib.cd_unet_forward.argtypes = [vp, vp, vp, vp, vp, i32, vp]
    lib.cd_denoise.argtypes = [vp, vp, vp, vp, vp, i32, vp]
    lib.cd_sample_run.argtypes = [vp, vp, f32, vp, C.POINTER(Eval), i32, vp, u64, i32, i32, vp, vp, vp, i32, vp, i32]
    lib.cd_fill_normal.argtypes = [vp, i64, i64, u64, i32, i32, vp]
    lib.cd_reverse_norm.argtypes = [C.POINTER(ReverseNormDesc), vp, vp, vp, vp, vp, i32, vp]
    lib.cd_reverse_norm_ex.argtypes = [C.POINTER(ReverseNormExDesc), vp, vp, vp, vp, vp, vp, i32, vp]
    lib.cd_geom_decode_sparse.argtypes = [vp, vp, vp, i32, i32, i32, i32, vp, vp, i32, i32, u64, vp, vp]
    lib.cd_train_step.argtypes = [vp, vp, vp, vp, vp, f32, i32, vp]
    lib.cd_train_loss.argtypes = [vp, C.POINTER(C.c_double), vp]
    lib.cd_train_adam.argtypes = [vp, f32, f32, f32, f32, f32, vp]
    lib.cd_train_refresh.argtypes = [vp, vp]
    lib.cd_train_flat_size.argtypes = [vp]
    lib.cd_train_flat_size.restype = i64
    lib.cd_train_num_params.argtypes = [vp]
    lib.cd_train_param_info.argtypes = [vp, i32, C.c_char_p, i32, C.POINTER(i64), C.POINTER(i64)]
    lib.cd_train_buffer.argtypes = [vp, i32]
    lib.cd_train_buffer.restype = vp
    lib.cd_test_conv.argtypes = [vp, i32, i32, i32, i32, C.POINTER(i32), C.POINTER(i32), i32, vp, vp, vp, vp, i32, i32, vp]
    lib.cd_level_shape.argtypes = [vp, i32, C.POINTER(i32)]
    lib.cd_conv_noalloc.argtypes = [vp, i32, i32, i32, i32, C.POINTER(i32), C.POINTER(i32), i32, vp, vp, vp, vp, i32, vp]
    lib.cd_conv_noalloc_max_bytes.argtypes = []
    lib.cd_conv_noalloc_max_bytes.restype = i64
    lib.cd_unet_profile_report.argtypes = [vp, C.c_char_p, i32]
    lib.cd_unet_set_geom_map.argtypes = [vp, i32, i32, i32, vp, vp, vp]
    lib.cd_unet_voxels.argtypes = [vp]
    lib.cd_unet_voxels.restype = i32
    lib.cd_geom_apply.argtypes = [vp, vp, vp, i32, i32, vp, vp, i32, vp]
    _lib = lib
    return lib


def check(rc, what="call"):
    if rc != 0:
        raise RuntimeError(f"calodiffusion_b200 {what} failed: {load().cd_last_error().decode()}")
