"""ctypes binding of libaqua_b200.so (include/aqua_b200.h).

The product path has no fallback: if the library is missing or no sm_100 GPU is present, calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

LIB_NAME = "libaqua_b200.so"
LIB_PATH = os.environ.get("AQUA_LIB_PATH") or os.path.join(os.path.dirname(os.path.abspath(__file__)), LIB_NAME)  # override: A/B of two builds

AQ_OK = 0
AQ_ACT_NONE, AQ_ACT_SILU, AQ_ACT_SILU_FAST = 0, 1, 2
AQ_JPEG_TABLE_BYTES = 8640
# tilings measured on a B200 for the plans the CLI / bench build (tools/make_tune_table.sh); AQUA_TUNE_TABLE overrides
TUNE_TABLE = os.environ.get("AQUA_TUNE_TABLE") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "tuned", "b200.txt")


class AquaError(RuntimeError):
    pass


class GeoRow(C.Structure):
    """aq_geo_row (include/aqua_b200.h)."""
    _fields_ = [("cls", C.c_int), ("xmin", C.c_int), ("xmax", C.c_int), ("ymin", C.c_int), ("ymax", C.c_int),
                ("conf", C.c_double), ("xmin_m", C.c_double), ("ymin_m", C.c_double), ("xmax_m", C.c_double),
                ("ymax_m", C.c_double)]


class GeoDet(C.Structure):
    """aq_geo_det (include/aqua_b200.h)."""
    _fields_ = [("image", C.c_int), ("cls", C.c_int), ("year", C.c_int), ("xmin", C.c_int), ("xmax", C.c_int), ("ymin", C.c_int),
                ("ymax", C.c_int), ("conf", C.c_double), ("xmin_m", C.c_double), ("ymin_m", C.c_double), ("xmax_m", C.c_double),
                ("ymax_m", C.c_double), ("xmin_a", C.c_double), ("ymin_a", C.c_double), ("xmax_a", C.c_double), ("ymax_a", C.c_double)]


class JpegInfo(C.Structure):
    """aq_jpeg_info (include/aqua_b200.h)."""
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("ncomp", C.c_int), ("hs", C.c_int), ("vs", C.c_int),
                ("mcus_x", C.c_int), ("mcus_y", C.c_int), ("restart_interval", C.c_int), ("coef_count", C.c_size_t)]


class ConvDesc(C.Structure):
    _fields_ = [("src", C.c_int), ("src_c_off", C.c_int), ("c_in", C.c_int),
                ("dst", C.c_int), ("dst_c_off", C.c_int), ("c_out", C.c_int),
                ("k", C.c_int), ("stride", C.c_int), ("act", C.c_int),
                ("res", C.c_int), ("res_c_off", C.c_int),
                ("weight", C.c_void_p), ("bias", C.c_void_p),
                ("tile_n", C.c_int), ("mblk", C.c_int), ("mode", C.c_int), ("cluster", C.c_int),
                ("up_src", C.c_int), ("up_c_off", C.c_int), ("up_c", C.c_int)]


class DetectDesc(C.Structure):
    _fields_ = [("src", C.c_int), ("src_c_off", C.c_int), ("c_in", C.c_int),
                ("nc", C.c_int), ("level_row_off", C.c_int), ("stride", C.c_float),
                ("anchors_px", C.c_float * 6),
                ("weight", C.c_void_p), ("bias", C.c_void_p)]


# name -> (restype, argtypes); every symbol include/aqua_b200.h declares
SIGNATURES = {
    "aq_abi_version": (C.c_int, []),
    "aq_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "aq_destroy": (None, [C.c_void_p]),
    "aq_last_error": (C.c_char_p, [C.c_void_p]),
    "aq_tune_load": (C.c_int, [C.c_void_p, C.c_char_p]),
    "aq_tune_misses": (C.c_int, [C.c_void_p]),
    "aq_plan_begin": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]),
    "aq_plan_buffer": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "aq_plan_stage_begin": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "aq_plan_stage_end": (C.c_int, [C.c_void_p]),
    "aq_plan_conv": (C.c_int, [C.c_void_p, C.POINTER(ConvDesc)]),
    "aq_plan_conv_chain": (C.c_int, [C.c_void_p, C.POINTER(ConvDesc), C.POINTER(ConvDesc), C.c_int]),
    "aq_plan_sppf_pool": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "aq_plan_upsample2x": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "aq_plan_detect": (C.c_int, [C.c_void_p, C.POINTER(DetectDesc)]),
    "aq_plan_end": (C.c_int, [C.c_void_p]),
    "aq_buffer_info": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                 C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "aq_pred_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "aq_detect_windows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_float, C.c_double, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    "aq_detect_window_list": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_int,
                                        C.c_float, C.c_double, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_void_p, C.c_void_p,
                                        C.c_void_p]),
    "aq_preprocess_windows_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                           C.c_void_p, C.c_void_p]),
    "aq_tile_stats_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_void_p, C.c_void_p]),
    "aq_tile_stats": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                C.c_void_p, C.c_void_p]),
    "aq_geocode_label_text": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_double), C.c_int, C.c_int, C.c_int,
                                        C.c_void_p, C.c_int]),
    "aq_geocode_dets": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                  C.POINTER(C.c_char_p), C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]),
    "aq_geo_class_name": (C.c_char_p, [C.c_int]),
    "aq_transform_3857_to_3035": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "aq_lonlat_to_3035": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "aq_cage_areas": (C.c_int, [C.c_int, C.c_double, C.c_double, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "aq_dbscan": (C.c_int, [C.c_void_p, C.c_int, C.c_double, C.c_int, C.c_void_p]),
    "aq_plan_num_launches": (C.c_int, [C.c_void_p]),
    "aq_profile_ops": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_float), C.c_int]),
    "aq_profile_ops_insitu": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_float), C.c_int]),
    "aq_op_describe": (C.c_int, [C.c_void_p, C.c_int, C.c_char_p, C.c_size_t]),
    "aq_profile_op_waves": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float)]),
    "aq_trace_op": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]),
    "aq_debug_plan_conv": (C.c_int, [C.c_int] * 10 + [C.c_char_p, C.c_size_t]),
    "aq_preprocess": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "aq_preprocess_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "aq_forward": (C.c_int, [C.c_void_p, C.c_void_p]),
    "aq_nms": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_double, C.c_int, C.c_int,
                         C.POINTER(C.c_int), C.c_int, C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                         C.c_void_p]),
    "aq_detect": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_double,
                            C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "aq_detect_submit": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                   C.c_double, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_void_p, C.c_void_p,
                                   C.c_void_p]),
    "aq_detect_wait": (C.c_int, [C.c_void_p, C.c_int]),
    "aq_detect_stage_ms": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_float)]),
    "aq_format_labels": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                   C.POINTER(C.c_char_p), C.c_int, C.c_int]),
    "aq_format_label_text": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p,
                                       C.c_size_t]),
    "aq_jpeg_probe": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(JpegInfo)]),
    "aq_jpeg_entropy_decode": (C.c_int, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p]),
    "aq_jpeg_entropy_decode_packed": (C.c_int, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                                C.POINTER(C.c_size_t), C.c_void_p]),
    "aq_jpeg_reconstruct_packed": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_int,
                                             C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "aq_jpeg_scan_prepare": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(JpegInfo), C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t),
                                       C.c_void_p, C.c_void_p]),
    "aq_jpeg_scan_prepare_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p,
                                             C.POINTER(C.c_size_t), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "aq_jpeg_decode_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int,
                                        C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "aq_jpeg_debug_parallel_decode": (C.c_int, [C.c_void_p, C.c_size_t, C.c_void_p, C.POINTER(JpegInfo), C.c_int, C.c_void_p,
                                                C.c_size_t, C.POINTER(C.c_int)]),
    "aq_jpeg_reconstruct": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                      C.c_int, C.c_void_p, C.c_void_p]),
    "aq_memcpy": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]),
    "aq_sync": (C.c_int, [C.c_void_p, C.c_void_p]),
}

_lib = None


def load() -> C.CDLL:
    """Load the shared library and bind every declared symbol; raise if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise AquaError(f"{LIB_PATH} is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                        f"(or `make -C aquaculture_b200/csrc`). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing: loud by design
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(lib, ctx, rc: int, what: str) -> int:
    if rc < 0:
        msg = lib.aq_last_error(ctx)
        raise AquaError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")
    return rc
