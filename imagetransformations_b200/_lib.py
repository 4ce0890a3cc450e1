"""ctypes binding of libimgtx.so (the C ABI declared in include/imgtx.h).

There is deliberately no fallback: if the shared library is missing or a call fails, we raise.  The
product path never routes through CPU code.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libimgtx.so")

_lib = None


class ImgtxError(RuntimeError):
    pass


class GeomRecord(C.Structure):
    """imgtx_geom_t (64 bytes)."""
    _fields_ = [("kind", C.c_int32), ("dx", C.c_int32), ("dy", C.c_int32), ("pad", C.c_int32),
                ("m", C.c_double * 6)]


# enums (mirrors include/imgtx.h)
PW_BRIGHTNESS, PW_SCALE_ABS, PW_POSTERIZE, PW_SOLARIZE, PW_IDENTITY = range(5)
(GEOM_IDENTITY, GEOM_AFFINE, GEOM_SHIFT, GEOM_FLIP_LR, GEOM_CROP_REFLECT, GEOM_ROT180, GEOM_ROT90,
 GEOM_ROT270) = range(8)
ENH_BRIGHTNESS, ENH_CONTRAST, ENH_COLOR, ENH_SHARPNESS, ENH_EQUALIZE, ENH_HISTEQ_YUV = range(6)
FILTER_BOX, FILTER_BILINEAR, FILTER_BICUBIC, FILTER_LANCZOS = range(4)
NOISE_GAUSS_F32, NOISE_GAUSS_F64, NOISE_IMPULSE, NOISE_SHOT = range(4)
TOU8_DENORMALIZE, TOU8_CLAMP01 = 1, 2
CLAMP_NONE, CLAMP_SCALAR, CLAMP_PER_CHANNEL = range(3)
FC_GAUSSIAN_NOISE, FC_BRIGHTNESS, FC_CONTRAST, FC_ADD_NOISE = range(4)
MAX_EPS = 8
CC_GAUSSIAN, CC_SHOT, CC_IMPULSE, CC_SPECKLE, CC_BRIGHTNESS, CC_CONTRAST, CC_FOG, CC_FROST = range(8)
(VIEW_IDENTITY, VIEW_FLIP, VIEW_CROP_RESIZE, VIEW_BRIGHTNESS, VIEW_CONTRAST, VIEW_CROP_REFLECT,
 VIEW_FLIP_CROP_REFLECT, VIEW_RESIZE_CENTER_CROP) = range(8)
AUG_NONE, AUG_GEOM, AUG_POSTERIZE, AUG_SOLARIZE, AUG_EQUALIZE = range(5)
CC_FAST, CC_REPLAY_ALL = 1, 2   # `flags` of the imgtx_cc_* entry points
CC_SHOT_FLAT = 0x101             # table layout selector of the on-chip shot_noise (include/imgtx.h)
ERR_UNSUPPORTED = -3

_vp = C.c_void_p
_i64 = C.c_int64
_int = C.c_int
_f = C.c_float

# name -> argtypes (restype is always int unless listed in _RESTYPES)
_PROTOS = {
    "imgtx_version": [],
    "imgtx_last_error": [],
    "imgtx_device_info": [C.POINTER(C.c_int), C.POINTER(C.c_int)],
    "imgtx_image_checksums_u8": [_vp, _i64, _i64, _vp, _vp],
    "imgtx_pointwise_u8": [_vp, _vp, _i64, _i64, _int, _f, _vp, _vp],
    "imgtx_geometry_u8": [_vp, _vp, _i64, _int, _int, _int, _int, _int, C.POINTER(GeomRecord), _vp, _vp, _vp],
    "imgtx_enhance_u8": [_vp, _vp, _i64, _int, _int, _int, _int, _f, _vp, _vp],
    "imgtx_resample_table_words": [_int, _int, _int],
    "imgtx_resample_table_host": [_int, _int, _int, _vp],
    "imgtx_resample_u8": [_vp, _vp, _i64, _int, _int, _int, _int, _int, _int, _int, _vp, _int, _int, _vp, _i64, _vp,
                          _i64, _int, _int, _int, _int, _vp],
    "imgtx_box_pixelate_u8": [_vp, _vp, _i64, _int, _int, _int, _int, _int, _int, _vp, _vp],
    "imgtx_resize_chain_u8": [_vp, _vp, _i64, _int, _int, _int, _int, _int, _vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _vp],
    "imgtx_cv_gaussian_kernel_host": [_int, C.c_double, _vp],
    "imgtx_cv_gaussian_blur_u8": [_vp, _vp, _i64, _int, _int, _int, _int, C.c_double, _int, _vp],
    "imgtx_cc_op32_mma_u8": [_vp, _vp, _i64, _vp, _i64, _int, _vp, _vp],
    "imgtx_cc_zoom_levels_u8": [_vp, _vp, _i64, _int, _int, _int, _int, _vp, _vp, _vp, _vp, _vp, _vp],
    "imgtx_pool_chain_f32": [_vp, _i64, _vp, _i64, _int, _int, _vp, _int, _vp, _vp, _vp, _vp, _vp, C.c_uint64, _vp],
    "imgtx_noalloc_scratch_bytes": [],
    "imgtx_cv_gaussian_blur_noalloc_u8": [_vp, _vp, _i64, _int, _int, _int, _int, C.c_double, _vp],
    "imgtx_pil_gaussian_blur_u8": [_vp, _vp, _i64, _int, _int, _int, _f, _vp],
    "imgtx_box_filter_h_u8": [_vp, _vp, _i64, _int, _int, _int, _int, _vp],
    "imgtx_pil_transform_u8": [_vp, _vp, _i64, _int, _int, _int, _int, _int, _int, _int, _vp, _vp, _vp, _vp],
    "imgtx_adaptive_chain_u8": [_vp, _vp, _i64, _int, _int, _int, _vp, _vp],
    "imgtx_perspective_tensor_u8": [_vp, _vp, _i64, _int, _int, _int, _vp, _f, _vp],
    "imgtx_gaussian_blur_f32": [_vp, _vp, _i64, _int, _int, _int, _f, _vp],
    "imgtx_noise_injected_u8": [_vp, _vp, _i64, _int, _int, _int, _int, _vp, C.c_double, _vp],
    "imgtx_noise_philox_u8": [_vp, _vp, _i64, _int, _int, _int, _int, C.c_double, C.c_uint64, _vp],
    "imgtx_cc_pointwise_u8": [_vp, _vp, _i64, _int, _int, _int, _int, C.c_double, C.c_double, _vp, _vp, C.c_uint64, _int,
                              _int, _int, _int, _vp, _vp],
    "imgtx_cc_noise_table_words": [_int, C.c_double],
    "imgtx_cc_noise_table_host": [_int, C.c_double, _vp, _i64],
    "imgtx_cc_noise_onchip_u8": [_vp, _vp, _i64, _int, _int, _int, _int, _vp, _i64, C.c_uint64, _int, _vp, _vp],
    "imgtx_cc_defocus_u8": [_vp, _vp, _i64, _int, _int, _int, _int, _vp, _vp, _int, _vp, _vp],
    "imgtx_cc_csr_u8": [_vp, _vp, _i64, _int, _int, _int, _vp, _vp, _vp, _int, _vp],
    "imgtx_cc_csr_ell_u8": [_vp, _vp, _i64, _int, _int, _int, _vp, _vp, _vp, _int, _int, _int, _vp, _vp],
    "imgtx_cc_glass_u8": [_vp, _vp, _i64, _int, _int, _int, _vp, _int, _int, _vp, C.c_uint64, _int, _vp, _vp],
    "imgtx_cc_motion_u8": [_vp, _vp, _i64, _int, _int, _int, _int, _vp, _vp, C.c_double, C.c_double, C.c_uint64, _int, _vp, _vp],
    "imgtx_cc_snow_u8": [_vp, _vp, _i64, _int, _int, _int, _f, _f, _f, _vp, _vp, _int, _vp, _vp, _vp, C.c_double,
                         C.c_double, C.c_double, C.c_double, C.c_uint64, _int, _vp, _vp],
    "imgtx_cc_elastic_u8": [_vp, _vp, _i64, _int, _int, _int, C.c_double, _vp, _int, _vp, _vp, _vp, C.c_double,
                            C.c_uint64, _int, _vp, _vp],
    "imgtx_cc_jpeg_u8": [_vp, _vp, _i64, _int, _int, _int, _vp, _int, _vp, _vp],
    "imgtx_to_tensor_f32": [_vp, _vp, _i64, _int, _int, _int, _vp, _vp, _vp],
    "imgtx_to_uint8_f32": [_vp, _vp, _i64, _int, _int, _int, _vp, _vp, _int, _vp],
    "imgtx_augmix_f32": [_vp, _vp, _i64, _int, _int, _int, _int, _vp, _vp, _vp],
    "imgtx_views_f32": [_vp, _int, _vp, _i64, _int, _int, _int, _vp, _vp, _vp, _vp, _int, _vp, _vp, _vp, _vp, _vp],
    "imgtx_sign_step_f32": [_vp, _vp, _vp, _vp, _i64, _int, _i64, _vp, _int, _f, _int, _vp, _vp, _vp],
    "imgtx_sign_step_post_f32": [_vp, _vp, _vp, _i64, _int, _i64, _vp, _int, _f, _int, _vp, _vp, _vp, _vp, _vp, _vp],
    "imgtx_pgd_step_f32": [_vp, _vp, _vp, _vp, _i64, _int, _i64, _f, _f, _int, _vp, _vp, _vp],
    "imgtx_channel_affine_f32": [_vp, _vp, _i64, _int, _i64, _vp, _vp, _int, _vp],
    "imgtx_float_corrupt_f32": [_vp, _vp, _vp, _i64, _int, _f, _f, _int, _vp],
    "imgtx_shannon_entropy_f32": [_vp, _i64, _int, _int, _int, _vp, _vp, _vp, _vp, _vp],
    "imgtx_tensor_tta_f32": [_vp, _vp, _i64, _int, _int, _int, _int, _vp, _vp],
    "imgtx_mean_views_f32": [_vp, _vp, _i64, _int, _vp],
}
_RESTYPES = {"imgtx_last_error": C.c_char_p, "imgtx_resample_table_words": C.c_int64,
             "imgtx_cc_noise_table_words": C.c_int64, "imgtx_noalloc_scratch_bytes": C.c_int64}


def load():
    """Load libimgtx.so (once). Raises ImportError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C imagetransformations_b200/csrc`). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in _PROTOS.items():
        fn = getattr(lib, name)  # AttributeError if the .so is stale: fail loudly
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, C.c_int)
    _lib = lib
    return lib


def check(rc: int):
    if rc != 0:
        msg = load().imgtx_last_error()
        raise ImgtxError(f"libimgtx error {rc}: {msg.decode() if msg else '?'}")


def exported_names():
    return list(_PROTOS)
