"""ctypes binding of libdcvc_b200.so (include/dcvc_b200.h).

The library is the product's only compute path: importing works anywhere (so that host logic can
be tested on a CPU box), but every compute entry fails loudly when the shared object is missing or
when there is no CUDA device -- there is no CPU or PyTorch fallback.
"""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "libdcvc_b200.so")

c_int, c_float, c_void_p, c_size_t = ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_size_t
c_int32, c_int64 = ctypes.c_int32, ctypes.c_int64


class View(ctypes.Structure):
    _fields_ = [("data", c_void_p), ("h", c_int32), ("w", c_int32), ("c", c_int32),
                ("cstride", c_int32), ("coff", c_int32)]


class Conv(ctypes.Structure):
    _fields_ = [("inp", View), ("out", View), ("res1", View), ("res2", View), ("out2", View),
                ("weight", c_void_p), ("weight_tc", c_void_p), ("bias", c_void_p),
                ("scale", c_void_p),
                ("ksize", c_int32), ("stride", c_int32), ("pad", c_int32), ("cout_w", c_int32),
                ("gate", c_int32), ("shuffle", c_int32),
                ("in_slope", c_float), ("out_slope", c_float), ("gate_slope", c_float),
                ("out2_slope", c_float), ("mode", c_int32), ("clamp01", c_int32),
                ("weight_tc16", c_void_p), ("in_f16", c_int32), ("out_f16", c_int32),
                ("res1_f16", c_int32), ("res2_f16", c_int32)]


class Ffn(ctypes.Structure):
    _fields_ = [("inp", View), ("out", View), ("w1", c_void_p), ("w2", c_void_p),
                ("b1", c_void_p), ("b2", c_void_p), ("scale", c_void_p),
                ("slope1", c_float), ("slope2", c_float), ("out_f16", c_int32)]


class Quad(ctypes.Structure):
    _fields_ = [("y", View), ("params", View), ("prior", View), ("y_hat", View), ("y_out", View),
                ("edges", c_void_p), ("step", c_int32), ("distribution", c_int32)]


class Dual(ctypes.Structure):
    _fields_ = [("y", View), ("fused", View), ("prior", View), ("cat", View), ("y_out", View),
                ("edges", c_void_p), ("post_scale", c_void_p), ("step", c_int32),
                ("distribution", c_int32)]


class Tables(ctypes.Structure):
    _fields_ = [("cdf", c_void_p), ("sizes", c_void_p), ("offsets", c_void_p),
                ("count", c_int32), ("stride", c_int32), ("lut", c_void_p)]


VP = ctypes.POINTER(View)

# name -> (restype, argtypes); tests/test_host_logic.py holds this table, the header and the .so exports equal
SIGNATURES = {
    "dcvc_last_error": (ctypes.c_char_p, []),
    "dcvc_version": (c_int, []),
    "dcvc_launch_count": (c_int64, []),
    "dcvc_launch_count_add": (None, [c_int64]),
    "dcvc_conv2d": (c_int, [ctypes.POINTER(Conv), c_void_p]),
    "dcvc_conv_tc_eligible": (c_int, [ctypes.POINTER(Conv)]),
    "dcvc_conv_tc_geometry": (c_int, [c_int, c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_int),
                                      ctypes.POINTER(c_int)]),
    "dcvc_conv_tc_geometry16": (c_int, [c_int, c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_int),
                                        ctypes.POINTER(c_int)]),
    "dcvc_ffn": (c_int, [ctypes.POINTER(Ffn), c_void_p]),
    "dcvc_ffn_eligible": (c_int, [c_int, c_int64]),
    "dcvc_dwconv3x3": (c_int, [VP, VP, c_void_p, c_void_p, c_void_p]),
    "dcvc_dwconv3x3_f16": (c_int, [VP, VP, c_void_p, c_void_p, c_void_p]),
    "dcvc_nchw_to_nhwc": (c_int, [c_void_p, VP, c_void_p]),
    "dcvc_nhwc_to_nchw": (c_int, [VP, c_void_p, c_int, c_void_p]),
    "dcvc_copy": (c_int, [VP, VP, c_float, c_float, c_void_p]),
    "dcvc_copy_f16": (c_int, [VP, VP, c_float, c_float, c_void_p]),
    "dcvc_copy_f16_f16": (c_int, [VP, VP, c_float, c_float, c_void_p]),
    "dcvc_copy_f16_narrow": (c_int, [VP, VP, c_void_p]),
    "dcvc_fill": (c_int, [VP, c_float, c_void_p]),
    "dcvc_pool2": (c_int, [VP, VP, c_int, c_void_p]),
    "dcvc_pool2_f16": (c_int, [VP, VP, c_int, c_void_p]),
    "dcvc_resize_bilinear": (c_int, [VP, VP, c_float, c_void_p]),
    "dcvc_flow_warp": (c_int, [VP, VP, VP, c_void_p]),
    "dcvc_offset_diversity": (c_int, [VP, VP, VP, c_void_p, c_void_p, VP, c_void_p]),
    "dcvc_offset_diversity_f16": (c_int, [VP, VP, VP, c_void_p, c_void_p, VP, c_void_p]),
    "dcvc_offset_diversity_f16_f16": (c_int, [VP, VP, VP, c_void_p, c_void_p, VP, c_void_p]),
    "dcvc_flow_warp_f16": (c_int, [VP, VP, VP, c_void_p]),
    "dcvc_flow_warp_f16_f16": (c_int, [VP, VP, VP, c_void_p]),
    "dcvc_frame_bytes": (c_size_t, [c_int, c_int, c_int]),
    "dcvc_frame_u8_to_view": (c_int, [c_void_p, c_int, c_int, c_int, VP, c_void_p]),
    "dcvc_plane_sse_ws_bytes": (c_size_t, [c_int, c_int]),
    "dcvc_view_plane_sse": (c_int, [VP, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                    c_void_p, c_void_p]),
    "dcvc_quad_bits_ws_bytes": (c_size_t, [c_int64]),
    "dcvc_quad_step_enc": (c_int, [ctypes.POINTER(Quad), c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_void_p]),
    "dcvc_quad_step_enc_static_max": (c_int64, []),
    "dcvc_quad_step_enc_static": (c_int, [ctypes.POINTER(Quad), c_void_p, c_void_p, c_void_p,
                                          c_void_p]),
    "dcvc_quad_step_indexes": (c_int, [ctypes.POINTER(Quad), c_void_p, c_void_p]),
    "dcvc_quad_step_dec": (c_int, [ctypes.POINTER(Quad), c_void_p, c_void_p]),
    "dcvc_dual_step_enc": (c_int, [ctypes.POINTER(Dual), c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_void_p]),
    "dcvc_dual_step_indexes": (c_int, [ctypes.POINTER(Dual), c_void_p, c_void_p]),
    "dcvc_dual_step_dec": (c_int, [ctypes.POINTER(Dual), c_void_p, c_void_p]),
    "dcvc_channel_scale": (c_int, [VP, VP, c_void_p, c_int, c_void_p]),
    "dcvc_channel_scale_f16": (c_int, [VP, VP, c_void_p, c_int, c_void_p]),
    "dcvc_se_ws_bytes": (c_size_t, [c_int, c_int, c_int]),
    "dcvc_se_gate": (c_int, [VP, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "dcvc_se_gate_f16": (c_int, [VP, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "dcvc_z_quant": (c_int, [VP, VP, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "dcvc_z_dequant": (c_int, [c_void_p, VP, c_void_p]),
    "dcvc_pmf_to_quantized_cdf": (c_int, [c_void_p, c_int, c_int, c_void_p]),
    "dcvc_rans_workspace_bytes": (c_size_t, [c_int, c_int, c_int64]),
    "dcvc_rans_create": (c_int, [ctypes.POINTER(c_void_p), c_int, c_int, c_int64, c_void_p,
                                 c_size_t, c_void_p]),
    "dcvc_rans_destroy": (None, [c_void_p]),
    "dcvc_rans_enc_reset": (c_int, [c_void_p]),
    "dcvc_rans_enc_push": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64,
                                   ctypes.POINTER(Tables)]),
    "dcvc_rans_enc_flush": (c_int, [c_void_p, c_void_p, c_size_t, ctypes.POINTER(c_size_t),
                                    c_void_p]),
    "dcvc_rans_enc_flush_begin": (c_int, [c_void_p, c_size_t, c_void_p]),
    "dcvc_rans_enc_flush_enqueue": (c_int, [c_void_p, c_size_t, c_void_p]),
    "dcvc_rans_enc_flush_mark": (c_int, [c_void_p, c_void_p]),
    "dcvc_rans_enc_flush_mark_captured": (c_int, [c_void_p, c_void_p]),
    "dcvc_rans_enc_flush_set_pending": (c_int, [c_void_p]),
    "dcvc_rans_enc_flush_end": (c_int, [c_void_p, c_void_p, c_size_t, ctypes.POINTER(c_size_t)]),
    "dcvc_rans_dec_set_stream": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p]),
    "dcvc_rans_dec_pull": (c_int, [c_void_p, c_void_p, c_int64, c_int64, ctypes.POINTER(Tables),
                                   c_void_p, c_void_p]),
    "dcvc_rans_dec_finish": (c_int, [c_void_p, c_void_p]),
    "dcvc_rans_dec_check": (c_int, [c_void_p]),
    "dcvc_rans_status": (c_int, [c_void_p, c_void_p]),
}

_lib = None


class DcvcError(RuntimeError):
    pass


def lib():
    """The loaded library.  Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.isfile(SO_PATH):
            raise DcvcError(
                f"{SO_PATH} is missing: run `python -m dcvc_b200.build` (or __graft_entry__.build()); "
                "dcvc_b200 has no CPU/PyTorch fallback")
        L = ctypes.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        msg = lib().dcvc_last_error()
        raise DcvcError(f"libdcvc_b200 error {rc}: {msg.decode() if msg else ''}")


def launch_count():
    return int(lib().dcvc_launch_count())
