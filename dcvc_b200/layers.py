"""Building blocks of the DCVC-DC networks as sequences of fused kernel launches.

Each function mirrors one module of the reference (cited per function, paths relative to
/root/reference/DCVC-DC/src/models) and states which elementwise work it folds into which
convolution's prologue/epilogue.  `Net` owns the packed device weights of one model.
"""
import torch

from . import ops
from .ops import ConvW
from .tensor import Act


def _round_tf32(w):
    """fp32 -> nearest TF32 (round half away, like cvt.rna.tf32): the tensor core reads only the
    top 19 bits, so pre-rounding the weights removes the truncation bias on the B operand."""
    bits = w.contiguous().view(torch.int32)
    return ((bits + 0x1000) & ~0x1FFF).view(torch.float32)


def _pack_tc(w, planar=False):
    """NK-packed weights for the tcgen05 path: [n_pad][kh*kw][cin_pad], zero padded, TF32-rounded
    (geometry from dcvc_conv_tc_geometry).  planar=True (depth-to-space convs, dcvc_conv.shuffle
    = 2) orders the GEMM columns (subpixel, channel) instead of the reference's (channel,
    subpixel), so that one 32-column chunk of the epilogue is 32 consecutive output channels of
    one output pixel."""
    import ctypes
    from . import _lib
    cout, cin, kh, kw = w.shape
    if planar:
        c = cout // 4
        w = w.reshape(c, 4, cin, kh, kw).permute(1, 0, 2, 3, 4).reshape(cout, cin, kh, kw)
    n_pad, cin_pad, kc = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    _lib.check(_lib.lib().dcvc_conv_tc_geometry(cout, cin, ctypes.byref(n_pad),
                                                ctypes.byref(cin_pad), ctypes.byref(kc)))
    out = torch.zeros((n_pad.value, kh * kw, cin_pad.value), dtype=torch.float32, device=w.device)
    out[:cout, :, :cin] = _round_tf32(w.permute(0, 2, 3, 1).reshape(cout, kh * kw, cin))
    return out.contiguous()


def _pack_tc16(w, planar=False):
    """binary16 NK packing for convs whose input is a binary16 activation (dcvc_conv.in_f16):
    [n_pad][kh*kw][cin_pad16], geometry from dcvc_conv_tc_geometry16; planar as in _pack_tc."""
    import ctypes
    from . import _lib
    cout, cin, kh, kw = w.shape
    if planar:
        c = cout // 4
        w = w.reshape(c, 4, cin, kh, kw).permute(1, 0, 2, 3, 4).reshape(cout, cin, kh, kw)
    n_pad, cin_pad, kc = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    _lib.check(_lib.lib().dcvc_conv_tc_geometry16(cout, cin, ctypes.byref(n_pad),
                                                  ctypes.byref(cin_pad), ctypes.byref(kc)))
    out = torch.zeros((n_pad.value, kh * kw, cin_pad.value), dtype=torch.float16, device=w.device)
    out[:cout, :, :cin] = w.permute(0, 2, 3, 1).reshape(cout, kh * kw, cin).to(torch.float16)
    return out.contiguous()


class Net:
    """Packed weights by reference parameter name (packing happens once per load/.to())."""

    def __init__(self, state_dict, device):
        self.device = torch.device(device)
        self.sd = {k: v.detach().to(self.device, torch.float32).contiguous()
                   for k, v in state_dict.items()}
        self._conv = {}
        self._dw = {}
        self._misc = {}

    def has(self, name):
        return (name + ".weight") in self.sd

    def cw(self, name, gate=False, cin_perm=None, shuffle=False, cout_pad=None):
        """KN-packed conv weights: [kh*kw][cin][cout]; gate=True interleaves the two halves of the
        output channels (ConvFFN2's chunk(2), layers.py:193-195) so the gate pairs are adjacent;
        cin_perm reorders the input channels (new channel i reads reference channel cin_perm[i],
        -1 = a padding channel with zero weights) so a concatenated input can be laid out
        alignment-friendly; shuffle=True marks a
        depth-to-space conv: its tensor-core copy is packed planar (see _pack_tc); cout_pad
        appends zero filters (and zero biases) up to that many output channels, so that a
        3-channel image convolution qualifies for the vectorised epilogues."""
        key = (name, gate, cin_perm, shuffle, cout_pad)
        r = self._conv.get(key)
        if r is None:
            w = self.sd[name + ".weight"]
            b = self.sd.get(name + ".bias")
            cout, cin, kh, kw = w.shape
            assert kh == kw
            if cin_perm is not None:       # an entry of -1 is a zero (padding) input channel
                idx = torch.tensor(cin_perm, device=w.device)
                w = w[:, idx.clamp(min=0)] * (idx >= 0).to(w.dtype).view(1, -1, 1, 1)
                cin = len(cin_perm)
            if cout_pad is not None and cout_pad > cout:
                w = torch.cat((w, w.new_zeros((cout_pad - cout, cin, kh, kw))), 0)
                if b is not None:
                    b = torch.cat((b, b.new_zeros(cout_pad - cout)), 0)
                cout = cout_pad
            if gate:
                half = cout // 2
                perm = torch.stack((torch.arange(half), torch.arange(half) + half), 1).reshape(-1)
                perm = perm.to(w.device)
                w = w[perm]
                b = b[perm] if b is not None else None
            wp = w.permute(2, 3, 1, 0).reshape(kh * kw * cin, cout).contiguous()
            r = ConvW(wp, b.contiguous() if b is not None else None, cin, cout, kh, gate,
                      w_tc=_pack_tc(w, planar=shuffle and cout % 16 == 0),
                      tc_planar=shuffle and cout % 16 == 0)
            if not shuffle or r.tc_planar:
                r.pack16 = lambda w=w, pl=r.tc_planar: _pack_tc16(w, pl)
            self._conv[key] = r
        return r

    def ffn(self, p):
        """Packed weights of the fused ConvFFN (dcvc_ffn) of block prefix p: W1 binary16
        [ceil(C/64)][4C][64] (K in atoms of 64, zero-padded), W2 binary16 [4C/64][C][64], fp32
        biases."""
        r = self._misc.get(("ffn", p))
        if r is None:
            w1 = self.sd[p + ".conv.0.weight"]          # [4C, C, 1, 1]
            w2 = self.sd[p + ".conv.2.weight"]          # [C, 4C, 1, 1]
            n1, c = w1.shape[0], w1.shape[1]
            ka = (c + 63) // 64                          # K atoms of 64 (C = 128: two)
            p1 = torch.zeros((ka, n1, 64), dtype=torch.float16, device=w1.device)
            w1f = w1.reshape(n1, c).to(torch.float16)
            for a in range(ka):
                k = min(64, c - 64 * a)
                p1[a, :, :k] = w1f[:, 64 * a:64 * a + k]
            p2 = w2.reshape(c, n1 // 64, 64).permute(1, 0, 2).to(torch.float16).contiguous()
            r = (p1.contiguous(), self.sd[p + ".conv.0.bias"].contiguous(), p2,
                 self.sd[p + ".conv.2.bias"].contiguous())
            self._misc[("ffn", p)] = r
        return r

    def dw(self, name):
        r = self._dw.get(name)
        if r is None:
            w = self.sd[name + ".weight"]            # [C,1,3,3]
            c = w.shape[0]
            r = (w.reshape(c, 9).t().contiguous(), self.sd[name + ".bias"].contiguous())
            self._dw[name] = r
        return r

    def dwpw(self, p):
        """DepthConv's depthwise 3x3 followed by its 1x1 `conv2` (no activation in between,
        layers.py:150-163) as ONE dense 3x3 convolution: W[co, ci, ky, kx] = conv2[co, ci] *
        depth_conv[ci, ky, kx], bias = conv2.bias + conv2 @ depth_conv.bias.  Zero padding of the
        depthwise input is the dense convolution's own padding, so the two are the same function
        (up to rounding); nine times the 1x1's MMAs, but no depthwise kernel and no intermediate
        tensor."""
        name = p + ".__dwpw"
        if name + ".weight" not in self.sd:
            dw = self.sd[p + ".depth_conv.weight"]              # [C, 1, 3, 3]
            w2 = self.sd[p + ".conv2.weight"][:, :, 0, 0]       # [Co, C]
            self.sd[name + ".weight"] = (w2[:, :, None, None] * dw[:, 0][None]).contiguous()
            self.sd[name + ".bias"] = self.sd[p + ".conv2.bias"] + w2 @ self.sd[p + ".depth_conv.bias"]
        return self.cw(name)

    def vec(self, name):
        r = self._misc.get(name)
        if r is None:
            r = self.sd[name].reshape(-1).contiguous()
            self._misc[name] = r
        return r


# ------------------------------------------------------------------------------- blocks
def subpel(net, p, x, act=1.0, out=None, scale=None, res=None, f16_out=False):
    """subpel_conv1x1/3x3 (layers.py:23-39): conv to 4C with the depth-to-space store fused."""
    return ops.conv(x, net.cw(p + ".0", shuffle=True), shuffle=True, act=act, out=out, scale=scale,
                    res=res, f16_out=f16_out)


def conv_for_res_block(net, name, x, slope, **kw):
    """A convolution whose only reader is a ResBlock: on the tensor-core path (no input prologue
    there) it also emits the block's activated input, saving the LeakyReLU copy pass.
    Returns (out, lrelu(out) or None)."""
    cw = net.cw(name)
    k, stride = cw.k, kw.get("stride", 1)
    ho, wo = (x.h + 2 * (k // 2) - k) // stride + 1, (x.w + 2 * (k // 2) - k) // stride + 1
    if ops.CONV_MODE == 1 and cw.w_tc is not None and ho * wo >= ops.TC_MIN_PIXELS and ops.ACT_COPY:
        return ops.conv(x, cw, act_copy=slope, **kw)
    return ops.conv(x, cw, **kw), None


def res_block(net, p, x, slope=0.01, end_with_relu=False, out=None, res2=None, scale=None,
              f16_out=False, x_act=None):
    """ResBlock (video_net.py:58-76): lrelu(x) is conv1's input prologue, the second activation
    conv1's epilogue, `x +` (and an optional outer residual/scale) conv2's epilogue."""
    if x_act is not None:          # lrelu(x) already made by x's producer (conv_for_res_block)
        t = ops.conv(x_act, net.cw(p + ".conv1"), act=slope, f16_out=True)
    else:
        t = ops.conv(x, net.cw(p + ".conv1"), in_slope=slope, act=slope, f16_out=True)
    return ops.conv(t, net.cw(p + ".conv2"), act=slope if end_with_relu else 1.0, res=x,
                    res2=res2, scale=scale, out=out, f16_out=f16_out)


import os as _os
# "cmin,cmax,pxmin": DepthConv blocks whose depthwise + 1x1 pair runs as one dense 3x3 (Net.dwpw)
_DWPW = [int(v) for v in _os.environ.get("DCVC_DWPW", "64,256,100000").split(",")]


def _fuse_dwpw(t):
    return (ops.CONV_MODE == 1 and t.f16 and _DWPW[0] <= t.c <= _DWPW[1]
            and t.h * t.w >= _DWPW[2])


def depth_conv(net, p, x, out=None, f16_out=False):
    """DepthConv (layers.py:135-163), stride 1.  Everything only the next kernel of the block
    reads (the adaptor's identity, conv1 and depth_conv outputs) is binary16 in tensor-core mode."""
    identity = x
    if net.has(p + ".adaptor"):
        identity = ops.conv(x, net.cw(p + ".adaptor"), f16_out=True)
    t = ops.conv(x, net.cw(p + ".conv1.0"), act=0.01, f16_out=True)
    if _fuse_dwpw(t):
        return ops.conv(t, net.dwpw(p), res=identity, out=out, f16_out=f16_out)
    w, b = net.dw(p + ".depth_conv")
    t = ops.dwconv3x3(t, w, b)
    return ops.conv(t, net.cw(p + ".conv2"), res=identity, out=out, f16_out=f16_out)


def dcb(net, p, x, out=None, scale=None, f16_out=False):
    """DepthConvBlock (layers.py:199-209) = DepthConv + ConvFFN (layers.py:166-179); an outer
    `* q` rides on the last epilogue.  f16_out: the caller is the only reader of the result and
    reads it with a convolution (block chains)."""
    t = depth_conv(net, p + ".block.0", x, f16_out=True)
    if ops.ffn_ok(t):
        w1, b1, w2, b2 = net.ffn(p + ".block.1")
        return ops.ffn(t, w1, b1, w2, b2, 0.1, 0.1, scale=scale, out=out, f16_out=f16_out)
    u = ops.conv(t, net.cw(p + ".block.1.conv.0"), act=0.1, f16_out=True)
    return ops.conv(u, net.cw(p + ".block.1.conv.2"), act=0.1, res=t, scale=scale, out=out,
                    f16_out=f16_out)


def dcb2(net, p, x, out=None, scale=None, f16_out=False):
    """DepthConvBlock2 (layers.py:212-222) = DepthConv + ConvFFN2 (layers.py:182-196); the
    x1 * lrelu(x2) gate is the epilogue of the up-projection."""
    t = depth_conv(net, p + ".block.0", x, f16_out=True)
    u = ops.conv(t, net.cw(p + ".block.1.conv", gate=True), gate_slope=0.1)
    return ops.conv(u, net.cw(p + ".block.1.conv_out"), res=t, scale=scale, out=out,
                    f16_out=f16_out)


def rbws(net, p, x, out=None, f16_out=False):
    """ResidualBlockWithStride (layers.py:42-73), stride 2."""
    t = ops.conv(x, net.cw(p + ".conv1"), stride=2, act=0.01, f16_out=True)
    idn = ops.conv(x, net.cw(p + ".downsample"), stride=2, pad=0, f16_out=True)
    return ops.conv(t, net.cw(p + ".conv2"), act=0.1, res=idn, out=out, f16_out=f16_out)


def rbu(net, p, x, out=None, scale=None, f16_out=False):
    """ResidualBlockUpsample (layers.py:76-101)."""
    up = subpel(net, p + ".upsample", x, f16_out=f16_out)
    t = subpel(net, p + ".subpel_conv", x, act=0.01, f16_out=True)
    return ops.conv(t, net.cw(p + ".conv"), act=0.1, res=up, scale=scale, out=out, f16_out=f16_out)


def unet(net, p, x, block, out=None, f16_out=False):
    """UNet / UNet2 (video_net.py:129-214); the two skip concatenations are shared buffers, read
    and written by convolutions and the max pool only: binary16 when every level down to the
    quarter-resolution one is tensor-core sized."""
    dev = x.device
    f16 = ops.f16_ok(x.h // 4, x.w // 4, 64)
    cat2 = Act.new(x.h, x.w, 64, dev, f16=f16)
    cat3 = Act.new(x.h // 2, x.w // 2, 128, dev, f16=f16)
    x1 = block(net, p + ".conv1", x, out=cat2.slice(0, 32))
    x2 = block(net, p + ".conv2", ops.max_pool2(x1), out=cat3.slice(0, 64))
    x3 = block(net, p + ".conv3", ops.max_pool2(x2), f16_out=True)
    for i in range(4):
        x3 = block(net, f"{p}.context_refine.{i}", x3, f16_out=True)
    subpel(net, p + ".up3", x3, out=cat3.slice(64, 64))
    d3 = block(net, p + ".up_conv3", cat3, f16_out=f16)
    subpel(net, p + ".up2", d3, out=cat2.slice(32, 32))
    return block(net, p + ".up_conv2", cat2, out=out, f16_out=f16_out)


def hyper_enc5(net, p, x):
    """get_hyper_enc_dec_models(reduce_enc_layer=False) encoder (video_net.py:226-237)."""
    x = ops.conv(x, net.cw(p + ".0"), act=0.01)
    x = ops.conv(x, net.cw(p + ".2"), act=0.01)
    x = ops.conv(x, net.cw(p + ".4"), stride=2, act=0.01)
    x = ops.conv(x, net.cw(p + ".6"), act=0.01)
    return ops.conv(x, net.cw(p + ".8"), stride=2)


def hyper_enc3(net, p, x):
    """reduce_enc_layer=True encoder (video_net.py:218-225)."""
    x = ops.conv(x, net.cw(p + ".0"), act=0.01)
    x = ops.conv(x, net.cw(p + ".2"), stride=2, act=0.01)
    return ops.conv(x, net.cw(p + ".4"), stride=2)


def hyper_dec(net, p, x):
    """hyper synthesis (video_net.py:239-249)."""
    x = ops.conv(x, net.cw(p + ".0"), act=0.01)
    x = subpel(net, p + ".2", x, act=0.01)
    x = ops.conv(x, net.cw(p + ".4"), act=0.01)
    x = subpel(net, p + ".6", x, act=0.01)
    return ops.conv(x, net.cw(p + ".8"))
