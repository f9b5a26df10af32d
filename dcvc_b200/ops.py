"""Python front end of the kernels in libdcvc_b200.so: one function per C entry point, taking
`Act`s and packed device weights.  Thin by design -- it builds the C descriptor and launches on
torch's current stream.  No function here computes anything in Python or torch.
"""
import ctypes

import torch

from . import _lib
from .tensor import Act, NULL_VIEW, _stream

# conv precision mode: 0 = fp32 FMA on CUDA cores (exact mode), 1 = tcgen05 TF32 where eligible
CONV_MODE = 0


# bench.py's per-launch profile: when a list, every wrapper appends
# (kernel name, algorithmic flops, algorithmic bytes, start event, end event)
PROFILE = None


class _Prof:
    __slots__ = ("name", "flops", "nbytes", "e0")

    def __init__(self, name, flops=0, nbytes=0):
        self.name, self.flops, self.nbytes = name, flops, nbytes

    def __enter__(self):
        if PROFILE is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *exc):
        if PROFILE is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            PROFILE.append((self.name, self.flops, self.nbytes, self.e0, e1))
        return False


def _nb(*acts):
    return sum((2 if a.f16 else 4) * a.h * a.w * a.c for a in acts if a is not None)


# binary16 storage of private conv -> conv intermediates on the tensor-core path (DCVC_F16=0
# keeps every activation fp32).  Same 11-bit significand the TF32 MMA reads from fp32 storage.
import os as _os
F16 = _os.environ.get("DCVC_F16", "1") != "0"


# maps with fewer output pixels stay on the CUDA-core kernel (conv_tc_ok in conv_tc.cu): 1/64
# resolution hyperprior tensors at 1080p are 17x30 = 510 pixels and do go to the tensor cores
TC_MIN_PIXELS = 256


def f16_ok(h, w, c):
    """May an h x w x c activation that only convolutions read be stored as binary16?  Both its
    producer and its consumers must be on the tensor-core path (>= TC_MIN_PIXELS output pixels) and the
    pixel's channel vector must be whole 16-byte groups."""
    return F16 and CONV_MODE == 1 and h * w >= TC_MIN_PIXELS and c % 8 == 0


def set_conv_mode(mode):
    global CONV_MODE
    assert mode in (0, 1)
    CONV_MODE = int(mode)


class ConvW:
    """Packed device weights of one convolution (built by layers.pack_conv)."""
    __slots__ = ("w", "w_tc", "b", "cin", "cout_w", "k", "gate", "tc_planar", "w_tc16", "pack16")

    def __init__(self, w, b, cin, cout_w, k, gate=False, w_tc=None, tc_planar=False):
        self.w, self.b, self.cin, self.cout_w, self.k, self.gate, self.w_tc = \
            w, b, cin, cout_w, k, gate, w_tc
        self.tc_planar = tc_planar    # w_tc columns ordered (subpixel, channel): shuffle = 2
        self.w_tc16 = None            # binary16 NK packing, made on first use by pack16()
        self.pack16 = None


def conv(x, cw, stride=1, pad=None, out=None, in_slope=1.0, act=1.0, res=None, res2=None,
         scale=None, shuffle=False, out2=None, out2_slope=1.0, gate_slope=0.1, clamp01=False,
         stream=None, f16_out=False, act_copy=None):
    """Fused convolution; see dcvc_conv2d in include/dcvc_b200.h for the exact semantics.
    f16_out=True asks for a binary16 result (granted when f16_ok() holds for it).
    act_copy=slope: also produce lrelu(out, slope) as a second tensor of out's element type (the
    activated input of the ResBlock that follows) and return (out, copy)."""
    assert x.c == cw.cin, (x, cw.cin)
    k = cw.k
    pad = k // 2 if pad is None else pad
    ho = (x.h + 2 * pad - k) // stride + 1
    wo = (x.w + 2 * pad - k) // stride + 1
    ec, eh, ew = cw.cout_w, ho, wo
    if shuffle:
        ec, eh, ew = ec // 4, ho * 2, wo * 2
    if cw.gate:
        ec //= 2
    if out is None:
        # (depth-to-space: only the planar-packed epilogue writes binary16, without residuals)
        f16_out = (f16_out and f16_ok(eh, ew, ec) and cw.w_tc is not None
                   and (not shuffle or (cw.tc_planar and res is None and ho * wo >= TC_MIN_PIXELS))
                   and not cw.gate and res2 is None and out2 is None)
        out = Act.new(eh, ew, ec, x.device, f16=f16_out)
    if act_copy is not None:
        assert out2 is None and res is None and res2 is None and not shuffle and not cw.gate
        out2, out2_slope = Act.new(eh, ew, ec, x.device, f16=out.f16), act_copy
    if CONV_MODE == 1 and in_slope != 1.0 and cw.w_tc is not None and ho * wo >= TC_MIN_PIXELS:
        # the TMA-fed tensor-core kernel has no input prologue: materialise lrelu(x) once
        # (as binary16 when the copy is private to this conv)
        x = copy(x, slope=in_slope, f16=x.f16 or (f16_ok(ho, wo, x.c) and cw.pack16 is not None))
        in_slope = 1.0
    d = _lib.Conv()
    d.inp = x.view()
    d.out = out.view()
    d.res1 = res.view() if res is not None else NULL_VIEW
    d.res2 = res2.view() if res2 is not None else NULL_VIEW
    d.out2 = out2.view() if out2 is not None else NULL_VIEW
    d.weight = cw.w.data_ptr()
    d.weight_tc = cw.w_tc.data_ptr() if cw.w_tc is not None else None
    d.in_f16, d.out_f16 = int(x.f16), int(out.f16)
    d.res1_f16 = int(res is not None and res.f16)
    d.res2_f16 = int(res2 is not None and res2.f16)
    if x.f16:
        if cw.w_tc16 is None:
            cw.w_tc16 = cw.pack16()
        d.weight_tc16 = cw.w_tc16.data_ptr()
    d.bias = cw.b.data_ptr() if cw.b is not None else None
    d.scale = scale.data_ptr() if scale is not None else None
    d.ksize, d.stride, d.pad, d.cout_w = k, stride, pad, cw.cout_w
    assert shuffle or not cw.tc_planar, "planar-packed weights are for depth-to-space convs"
    d.gate, d.shuffle = int(cw.gate), (2 if cw.tc_planar else 1) if shuffle else 0
    d.in_slope, d.out_slope, d.gate_slope, d.out2_slope = in_slope, act, gate_slope, out2_slope
    d.mode = CONV_MODE
    d.clamp01 = int(clamp01)
    if PROFILE is None:
        _lib.check(_lib.lib().dcvc_conv2d(ctypes.byref(d), _stream(stream)))
        return (out, out2) if act_copy is not None else out
    path = "tcgen05" if _lib.lib().dcvc_conv_tc_eligible(ctypes.byref(d)) else "fma"
    tags = ("i" if x.f16 else "I") + ("o" if out.f16 else "O") + \
        ("" if res is None else "r" if res.f16 else "R") + \
        ("" if res2 is None else "2h" if res2.f16 else "2")
    name = f"conv{k}x{k}s{stride} {cw.cin}->{cw.cout_w} @{ho}x{wo} [{path}] {tags}"
    flops = 2 * ho * wo * cw.cout_w * cw.cin * k * k
    nbytes = _nb(x, out, res, res2, out2) + 4 * cw.w.numel()
    with _Prof(name + ("+a" if act_copy is not None else ""), flops, nbytes):
        _lib.check(_lib.lib().dcvc_conv2d(ctypes.byref(d), _stream(stream)))
    return (out, out2) if act_copy is not None else out


# DCVC_ACT_COPY=0: ResBlock inputs are activated by a separate copy pass again
ACT_COPY = _os.environ.get("DCVC_ACT_COPY", "1") != "0"

# fused ConvFFN (dcvc_ffn): DCVC_FFN=0 keeps the two separate 1x1 convolutions
FFN = _os.environ.get("DCVC_FFN", "1") != "0"


def ffn_ok(x):
    """May DepthConvBlock's ConvFFN on the binary16 activation x run as the fused kernel?"""
    return (FFN and CONV_MODE == 1 and x.f16 and
            bool(_lib.lib().dcvc_ffn_eligible(x.c, x.h * x.w)))


def ffn(x, w1, b1, w2, b2, slope1=0.1, slope2=0.1, scale=None, out=None, f16_out=False,
        stream=None):
    """out = (x + lrelu(W2 . lrelu(W1 . x + b1) + b2)) * scale in one kernel (dcvc_ffn)."""
    if out is None:
        out = Act.new(x.h, x.w, x.c, x.device, f16=f16_out and f16_ok(x.h, x.w, x.c))
    d = _lib.Ffn()
    d.inp, d.out = x.view(), out.view()
    d.w1, d.w2, d.b1, d.b2 = w1.data_ptr(), w2.data_ptr(), b1.data_ptr(), b2.data_ptr()
    d.scale = scale.data_ptr() if scale is not None else None
    d.slope1, d.slope2, d.out_f16 = slope1, slope2, int(out.f16)
    flops = 2 * x.h * x.w * x.c * 4 * x.c * 2
    with _Prof(f"ffn {x.c}->{4 * x.c}->{x.c} @{x.h}x{x.w} [tcgen05] i" + ("o" if out.f16 else "O"),
               flops, _nb(x, out) + 2 * (w1.numel() + w2.numel())):
        _lib.check(_lib.lib().dcvc_ffn(ctypes.byref(d), _stream(stream)))
    return out


def dwconv3x3(x, w, b, out=None, stream=None):
    if out is None:
        out = Act.new(x.h, x.w, x.c, x.device, f16=x.f16)
    assert out.f16 == x.f16
    fn = _lib.lib().dcvc_dwconv3x3_f16 if x.f16 else _lib.lib().dcvc_dwconv3x3
    with _Prof("dwconv3x3" + (" f16" if x.f16 else ""), 18 * x.h * x.w * x.c, _nb(x, out)):
        _lib.check(fn(x.view(), out.view(), w.data_ptr(), b.data_ptr() if b is not None else None,
                      _stream(stream)))
    return out


def copy(src, out=None, h=None, w=None, mul=1.0, slope=1.0, stream=None, f16=False):
    """Channel-slice copy; with (h, w) != src size: replicate pad / crop.  f16=True: a new
    binary16 destination (same size only)."""
    if out is None:
        out = Act.new(src.h if h is None else h, src.w if w is None else w, src.c, src.device,
                      f16=f16)
    if src.f16 and not out.f16:
        raise ValueError("copy: binary16 -> fp32 is not provided (binary16 tensors feed convolutions)")
    if out.f16 and not src.f16 and src.c < 8:
        assert mul == 1.0 and slope == 1.0
        with _Prof(f"copy {out.c}ch @{out.h}x{out.w} f16", 0, _nb(src, out)):
            _lib.check(_lib.lib().dcvc_copy_f16_narrow(src.view(), out.view(), _stream(stream)))
        return out
    fn = (_lib.lib().dcvc_copy_f16_f16 if src.f16 else _lib.lib().dcvc_copy_f16) if out.f16 \
        else _lib.lib().dcvc_copy
    with _Prof(f"copy {out.c}ch @{out.h}x{out.w}" + (" lrelu" if slope != 1.0 else "") +
               (" f16" if out.f16 else ""), 0, _nb(src, out)):
        _lib.check(fn(src.view(), out.view(), mul, slope, _stream(stream)))
    return out


def fill(dst, value, stream=None):
    _lib.check(_lib.lib().dcvc_fill(dst.view(), value, _stream(stream)))
    return dst


def pool2(x, mode, out=None, stream=None):
    if out is None:
        out = Act.new(x.h // 2, x.w // 2, x.c, x.device, f16=x.f16)
    assert out.f16 == x.f16
    fn = _lib.lib().dcvc_pool2_f16 if x.f16 else _lib.lib().dcvc_pool2
    with _Prof("pool2" + (" f16" if x.f16 else ""), 0, _nb(x, out)):
        _lib.check(fn(x.view(), out.view(), mode, _stream(stream)))
    return out


def avg_pool2(x, out=None):
    return pool2(x, 0, out)


def max_pool2(x, out=None):
    return pool2(x, 1, out)


def resize_bilinear(x, h, w, mul=1.0, out=None, stream=None):
    if out is None:
        out = Act.new(h, w, x.c, x.device)
    with _Prof("resize_bilinear", 0, _nb(x, out)):
        _lib.check(_lib.lib().dcvc_resize_bilinear(x.view(), out.view(), mul, _stream(stream)))
    return out


def flow_warp(feat, flow, out=None, stream=None):
    """out may be a binary16 Act (a context only tensor-core convolutions read)."""
    if out is None:
        out = Act.new(flow.h, flow.w, feat.c, feat.device)
    if feat.f16 and not out.f16:
        raise ValueError("flow_warp: a binary16 source needs a binary16 destination")
    fn = (_lib.lib().dcvc_flow_warp_f16_f16 if feat.f16 else _lib.lib().dcvc_flow_warp_f16) \
        if out.f16 else _lib.lib().dcvc_flow_warp
    with _Prof("flow_warp" + (" f16" if out.f16 else "") + (" f16src" if feat.f16 else ""), 0,
               _nb(feat, flow, out)):
        _lib.check(fn(feat.view(), flow.view(), out.view(), _stream(stream)))
    return out


def offset_diversity(x, offs_half, flow, fusion_w, fusion_b, out=None, stream=None):
    if out is None:
        out = Act.new(x.h, x.w, 48, x.device)
    if x.f16 and not out.f16:
        raise ValueError("offset_diversity: a binary16 source needs a binary16 destination")
    fn = (_lib.lib().dcvc_offset_diversity_f16_f16 if x.f16 else _lib.lib().dcvc_offset_diversity_f16) \
        if out.f16 else _lib.lib().dcvc_offset_diversity
    with _Prof("offset_diversity" + (" f16" if out.f16 else ""), 0, _nb(x, offs_half, flow, out)):
        _lib.check(fn(x.view(), offs_half.view(), flow.view(), fusion_w.data_ptr(),
                      fusion_b.data_ptr(), out.view(), _stream(stream)))
    return out


def to_nchw(x, clamp01=False, stream=None):
    """Contiguous NCHW copy [1,C,H,W] of an Act."""
    out = torch.empty((1, x.c, x.h, x.w), dtype=torch.float32, device=x.device)
    _lib.check(_lib.lib().dcvc_nhwc_to_nchw(x.view(), out.data_ptr(), int(clamp01),
                                            _stream(stream)))
    return out


# ---------------------------------------------------------------------------- entropy model
def _quad(y, params, prior, y_hat, y_out, edges, step, distribution):
    q = _lib.Quad()
    q.y = y.view() if y is not None else NULL_VIEW
    q.params = params.view()
    q.prior = prior.view() if prior is not None else NULL_VIEW
    q.y_hat = y_hat.view()
    q.y_out = y_out.view()
    q.edges = edges.data_ptr()
    q.step = step
    q.distribution = distribution
    return q


def quad_step_enc(y, params, prior, y_hat, y_out, edges, step, distribution, symbols=None,
                  indexes=None, bits=None, bits_ws=None, stream=None):
    q = _quad(y, params, prior, y_hat, y_out, edges, step, distribution)
    with _Prof("quad_step_enc"):
        _lib.check(_lib.lib().dcvc_quad_step_enc(
            ctypes.byref(q), symbols.data_ptr() if symbols is not None else None,
            indexes.data_ptr() if indexes is not None else None,
            bits.data_ptr() if bits is not None else None,
            bits_ws.data_ptr() if bits_ws is not None else None, _stream(stream)))


def quad_step_indexes(params, prior, y_hat, y_out, edges, step, indexes, stream=None):
    q = _quad(None, params, prior, y_hat, y_out, edges, step, 0)
    with _Prof("quad_step_indexes"):
        _lib.check(_lib.lib().dcvc_quad_step_indexes(ctypes.byref(q), indexes.data_ptr(),
                                                     _stream(stream)))


def quad_step_dec(params, prior, y_hat, y_out, edges, step, symbols, stream=None):
    q = _quad(None, params, prior, y_hat, y_out, edges, step, 0)
    with _Prof("quad_step_dec"):
        _lib.check(_lib.lib().dcvc_quad_step_dec(ctypes.byref(q), symbols.data_ptr(), _stream(stream)))


def z_quant(z, z_hat, symbols=None, fparams=None, bits=None, bits_ws=None, stream=None):
    with _Prof("z_quant"):
        _lib.check(_lib.lib().dcvc_z_quant(
            z.view(), z_hat.view(), symbols.data_ptr() if symbols is not None else None,
            fparams.data_ptr() if fparams is not None else None,
            bits.data_ptr() if bits is not None else None,
            bits_ws.data_ptr() if bits_ws is not None else None, _stream(stream)))


def z_dequant(symbols, z_hat, stream=None):
    with _Prof("z_dequant"):
        _lib.check(_lib.lib().dcvc_z_dequant(symbols.data_ptr(), z_hat.view(), _stream(stream)))


# ---------------------------------------------------------------------------- DCVC-HEM pieces
def _dual(y, fused, prior, cat, y_out, edges, step, distribution, post_scale):
    q = _lib.Dual()
    q.y = y.view() if y is not None else NULL_VIEW
    q.fused = fused.view() if fused is not None else NULL_VIEW
    q.prior = prior.view() if prior is not None else NULL_VIEW
    q.cat = cat.view()
    q.y_out = y_out.view()
    q.edges = edges.data_ptr()
    q.post_scale = post_scale.data_ptr() if post_scale is not None else None
    q.step, q.distribution = step, distribution
    return q


def dual_step_enc(y, fused, prior, cat, y_out, edges, step, distribution, post_scale=None,
                  symbols=None, indexes=None, bits=None, bits_ws=None, stream=None):
    q = _dual(y, fused, prior, cat, y_out, edges, step, distribution, post_scale)
    with _Prof("dual_step_enc"):
        _lib.check(_lib.lib().dcvc_dual_step_enc(
            ctypes.byref(q), symbols.data_ptr() if symbols is not None else None,
            indexes.data_ptr() if indexes is not None else None,
            bits.data_ptr() if bits is not None else None,
            bits_ws.data_ptr() if bits_ws is not None else None, _stream(stream)))


def dual_step_indexes(fused, prior, cat, y_out, edges, step, indexes, stream=None):
    q = _dual(None, fused, prior, cat, y_out, edges, step, 0, None)
    with _Prof("dual_step_indexes"):
        _lib.check(_lib.lib().dcvc_dual_step_indexes(ctypes.byref(q), indexes.data_ptr(),
                                                     _stream(stream)))


def dual_step_dec(fused, prior, cat, y_out, edges, step, symbols, post_scale=None, stream=None):
    q = _dual(None, fused, prior, cat, y_out, edges, step, 0, post_scale)
    with _Prof("dual_step_dec"):
        _lib.check(_lib.lib().dcvc_dual_step_dec(ctypes.byref(q), symbols.data_ptr(),
                                                 _stream(stream)))


def channel_scale(x, vec, divide=False, out=None, stream=None):
    """out = x * vec[c] (or x / vec[c]); fp32 views, or binary16 in and out."""
    if out is None:
        out = Act.new(x.h, x.w, x.c, x.device, f16=x.f16)
    if out.f16 != x.f16:
        raise ValueError("channel_scale: source and destination must hold the same element type")
    fn = _lib.lib().dcvc_channel_scale_f16 if x.f16 else _lib.lib().dcvc_channel_scale
    with _Prof("channel_scale" + (" f16" if x.f16 else ""), 0, _nb(x, out)):
        _lib.check(fn(x.view(), out.view(), vec.data_ptr(), int(divide), _stream(stream)))
    return out


def se_gate(x, w1, w2, stream=None):
    """SELayer gate [C] of the activation x, fp32 or binary16 (w1 [C/r][C], w2 [C][C/r])."""
    hidden = w1.shape[0]
    gate = torch.empty(x.c, dtype=torch.float32, device=x.device)
    ws = torch.empty(int(_lib.lib().dcvc_se_ws_bytes(x.h, x.w, x.c)) // 8 + 1, dtype=torch.float64,
                     device=x.device)
    fn = _lib.lib().dcvc_se_gate_f16 if x.f16 else _lib.lib().dcvc_se_gate
    with _Prof("se_gate" + (" f16" if x.f16 else ""), 0, _nb(x)):
        _lib.check(fn(x.view(), w1.data_ptr(), w2.data_ptr(), hidden, gate.data_ptr(), ws.data_ptr(),
                      _stream(stream)))
    return gate
