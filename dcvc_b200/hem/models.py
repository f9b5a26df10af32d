"""DCVC-HEM `IntraNoAR` and `DMC` behind the reference's module API
(DCVC-HEM/src/models/image_model.py:15-171, video_model.py:131-519, common_model.py:14-188),
executed on the same sm_100a kernel set as the DCVC-DC models plus the dual-prior, channel-scale
and SE-gate kernels (csrc/hem.cu).

Differences from DCVC-DC that shape this file: the rate is a CONTINUOUS q_scale
(curr_q = LowerBound(q_basic, 0.5) * q_scale; real streams carry it rounded to 1/100 in the frame
header, stream_helper.py:41-45,102-143); latents are divided by curr_q after the analysis
transform and multiplied back before synthesis; the spatial prior is the two-step checkerboard
("dual") prior; conv stacks are plain ResidualBlocks and UNets with SE layers; the bitstream is one
rANS stream without a part header, in the order mv_z, mv_y 0-1, z, y 0-1.
"""
import time

import numpy as np
import torch

from .. import entropy_models as em
from .. import layers, ops
from ..common_model import CompressionModel, get_downsampled_shape
from ..tensor import Act
from . import arch as harch
from . import stream_helper as sh


# ------------------------------------------------------------------------------ blocks
def residual_block(net, p, x, out=None, f16_out=False):
    """ResidualBlock (layers.py:103-128): conv -> lrelu(.01) -> conv -> lrelu(.01) -> + x"""
    t = ops.conv(x, net.cw(p + ".conv1"), act=0.01, f16_out=True)
    return ops.conv(t, net.cw(p + ".conv2"), act=0.01, res=x, out=out, f16_out=f16_out)


def conv_block_residual(net, p, x, out=None, f16_out=False):
    """ConvBlockResidual (video_net.py:166-181): up_dim(x) + SE(conv(lrelu(conv(x))))"""
    t = ops.conv(x, net.cw(p + ".conv.0"), act=0.01, f16_out=True)
    t = ops.conv(t, net.cw(p + ".conv.2"), f16_out=True)       # the SE branch: binary16 in tensor-core mode
    gate = ops.se_gate(t, net.sd[p + ".conv.3.fc.0.weight"], net.sd[p + ".conv.3.fc.2.weight"])
    ops.channel_scale(t, gate, out=t)
    return ops.conv(x, net.cw(p + ".up_dim"), res=t, out=out, f16_out=f16_out)


def hem_unet(net, p, x, out=None, f16_out=False):
    """UNet (video_net.py:184-223): the two skip concatenations are shared buffers, binary16 in
    tensor-core mode like every tensor here that only convolutions and the max pool read (the SE
    branch too: the global mean is accumulated in float64 from the binary16 convolution result)."""
    dev = x.device
    f16 = ops.f16_ok(x.h // 4, x.w // 4, 64)
    cat2 = Act.new(x.h, x.w, 64, dev, f16=f16)                 # [x1 32 | up2 32]
    cat3 = Act.new(x.h // 2, x.w // 2, 128, dev, f16=f16)      # [x2 64 | up3 64]
    x1 = conv_block_residual(net, p + ".conv1", x, out=cat2.slice(0, 32))
    x2 = conv_block_residual(net, p + ".conv2", ops.max_pool2(x1), out=cat3.slice(0, 64))
    x3 = conv_block_residual(net, p + ".conv3", ops.max_pool2(x2), f16_out=True)
    for i in range(4):
        x3 = layers.res_block(net, f"{p}.context_refine.{i}", x3, slope=0.0, f16_out=True)
    layers.subpel(net, p + ".up3", x3, out=cat3.slice(64, 64))
    d3 = conv_block_residual(net, p + ".up_conv3", cat3, f16_out=True)
    layers.subpel(net, p + ".up2", d3, out=cat2.slice(32, 32))
    return conv_block_residual(net, p + ".up_conv2", cat2, out=out, f16_out=f16_out)


def analysis(net, p, x):
    """get_enc_dec_models encoder (video_net.py:226-235)"""
    # the chain's tensors are read by convolutions only: binary16 in tensor-core mode, as long as
    # the stride-2 reader's OUTPUT is tensor-core sized too (a binary16 view on the FMA path is an
    # error, never a reinterpretation)
    def ok(div):
        return ops.f16_ok(x.h // div, x.w // div, 8)
    t = layers.rbws(net, p + ".0", x, f16_out=True)
    t = residual_block(net, p + ".1", t, f16_out=ok(4))
    t = layers.rbws(net, p + ".2", t, f16_out=True)
    t = residual_block(net, p + ".3", t, f16_out=ok(8))
    t = layers.rbws(net, p + ".4", t, f16_out=True)
    t = residual_block(net, p + ".5", t, f16_out=ok(16))
    return ops.conv(t, net.cw(p + ".6"), stride=2)


def synthesis(net, p, x, f16_out=False):
    """get_enc_dec_models decoder (video_net.py:237-247)"""
    t = residual_block(net, p + ".0", x, f16_out=True)
    t = layers.rbu(net, p + ".1", t, f16_out=True)
    t = residual_block(net, p + ".2", t, f16_out=True)
    t = layers.rbu(net, p + ".3", t, f16_out=True)
    t = residual_block(net, p + ".4", t, f16_out=True)
    t = layers.rbu(net, p + ".5", t, f16_out=True)
    t = residual_block(net, p + ".6", t, f16_out=True)
    cout = net.sd[p + ".7.0.weight"].shape[0]
    if cout % 16:      # 8 filters (2-channel flow): zero filters up to 16 -> planar packing, which
        # is what a binary16 input needs (as the DCVC-DC mv decoder does)
        cw = net.cw(p + ".7.0", shuffle=True, cout_pad=16)
        return ops.conv(t, cw, shuffle=True).slice(0, cout // 4)
    return layers.subpel(net, p + ".7", t, f16_out=f16_out)


def three_convs(net, p, x, out=None):
    """prior fusion / spatial prior stacks: conv3x3, LeakyReLU(0.2), conv3x3, LeakyReLU(0.2), conv3x3"""
    t = ops.conv(x, net.cw(p + ".0"), act=0.2, f16_out=True)
    t = ops.conv(t, net.cw(p + ".2"), act=0.2, f16_out=True)
    return ops.conv(t, net.cw(p + ".4"), out=out)


# ------------------------------------------------------------------------------ base
class HemCompressionModel(CompressionModel):
    """CompressionModel (DCVC-HEM/src/models/common_model.py:14-188) on the device."""

    def __init__(self, spec, y_distribution, z_names, anchor_num, coder_mode, stream_part):
        super().__init__(spec, y_distribution, z_names, False, stream_part, anchor_num, coder_mode)

    def curr_q(self, basic_name, q_scale):
        """get_curr_q: LowerBound(q_basic, 0.5) * q_scale (image_model.py:61-63,
        video_model.py:255-261); q_scale is a Python float or a 0-d tensor."""
        basic = torch.clamp_min(self.net.sd[basic_name], 0.5)
        return (basic * float(q_scale)).reshape(-1).contiguous()

    def forward_dual_prior(self, y, fused, prior_name, post_scale, write=False, bits=None):
        """forward_dual_prior / compress_dual_prior (common_model.py:102-159).  y is already
        divided by curr_q.  Returns (y_hat * quant_step * curr_q, [(symbols, indexes)] * 2)."""
        C, h, w, dev = y.c, y.h, y.w, y.device
        n = C // 2 * h * w
        cat = Act.new(h, w, 4 * C, dev)
        y_out = Act.new(h, w, C, dev)
        edges = self._ensure_edges()
        dist = em.DIST_ID[self.y_distribution]
        ws = torch.empty((n + 255) // 256, dtype=torch.float32, device=dev) if bits is not None \
            else None
        coded = []
        for step in range(2):
            prior = three_convs(self.net, prior_name, cat) if step else None
            sym = idx = None
            if write:
                sym = torch.empty(n, dtype=torch.int16, device=dev)
                idx = torch.empty(n, dtype=torch.uint8, device=dev)
                coded.append((sym, idx))
            ops.dual_step_enc(y, fused if step == 0 else None, prior, cat, y_out, edges, step, dist,
                              post_scale, sym, idx, bits, ws)
        return y_out, coded

    def decompress_dual_prior(self, fused, prior_name, post_scale):
        """decompress_dual_prior (common_model.py:161-188)"""
        C, h, w, dev = fused.c // 3, fused.h, fused.w, fused.device
        n = C // 2 * h * w
        cat = Act.new(h, w, 4 * C, dev)
        y_out = Act.new(h, w, C, dev)
        edges = self._ensure_edges()
        idx = torch.empty(n, dtype=torch.uint8, device=dev)
        sym = torch.empty(n, dtype=torch.int16, device=dev)
        for step in range(2):
            prior = three_convs(self.net, prior_name, cat) if step else None
            f = fused if step == 0 else None
            ops.dual_step_indexes(f, prior, cat, y_out, edges, step, idx)
            self.entropy_coder.decode(idx, n, 1, self._tables["y"], sym)
            ops.dual_step_dec(f, prior, cat, y_out, edges, step, sym, post_scale)
        return y_out

    def _tap_coded(self, name, coded, lat):
        if self.taps is not None:
            for st, (sym, idx) in enumerate(coded):
                shp = (1, lat.c // 2, lat.h, lat.w)
                self.taps[f"{name}_q_w_{st}"] = sym.view(shp).clone()
                self.taps[f"{name}_idx_w_{st}"] = idx.view(shp).clone()

    def _tap(self, name, act):
        if self.taps is not None:
            self.taps[name] = act.nchw().detach().float().clone()


# ------------------------------------------------------------------------------ intra
class IntraNoAR(HemCompressionModel):
    def __init__(self, N=192, anchor_num=4, coder_mode="hem", stream_part=1):
        super().__init__(harch.hem_intra_spec(N, anchor_num), "gaussian", ("bit_estimator_z",),
                         anchor_num, coder_mode, stream_part)
        self.N = int(N)

    @staticmethod
    def get_q_scales_from_ckpt(ckpt_path):
        """image_model.py:110-114"""
        from ..stream_helper import get_state_dict
        return get_state_dict(ckpt_path)["q_scale"].reshape(-1)

    def _encode_side(self, x, q_scale, write, bits_y=None, bits_z=None):
        net = self.net
        cq = self.curr_q("q_basic", q_scale)
        y = ops.channel_scale(analysis(net, "enc", Act.wrap(x)), cq, divide=True)
        z = layers.hyper_enc5(net, "hyper_enc", y)
        z_hat, z_sym = self.quant_z(z, "bit_estimator_z", write, bits_z)
        fused = three_convs(net, "y_prior_fusion", layers.hyper_dec(net, "hyper_dec", z_hat))
        y_hat, coded = self.forward_dual_prior(y, fused, "y_spatial_prior", cq, write, bits_y)
        self._tap("i_y", y)
        self._tap("i_z_hat", z_hat)
        self._tap("i_fused", fused)
        self._tap("i_y_hat", y_hat)
        self._tap_coded("i_y", coded, y)
        return y_hat, coded, z_hat, z_sym

    def _synthesis(self, y_hat, clamp):
        net = self.net
        t = hem_unet(net, "refine.0", synthesis(net, "dec", y_hat, f16_out=True), f16_out=True)
        return ops.conv(t, net.cw("refine.1", cout_pad=4), clamp01=clamp).slice(0, 3)

    def forward(self, x, q_scale=None):
        """image_model.py:65-108 (estimate mode; mse/ssim of the training loss are not produced)"""
        bits = torch.zeros(2, dtype=torch.float32, device=x.device)
        y_hat, _, _, _ = self._encode_side(x, q_scale, False, bits[0:1], bits[1:2])
        x_hat = self._synthesis(y_hat, clamp=False)
        pixel_num = x.shape[2] * x.shape[3]
        bpp_y, bpp_z = bits[0:1] / pixel_num, bits[1:2] / pixel_num
        return {"x_hat": x_hat.nchw(), "bit": float((torch.sum(bpp_y + bpp_z) * pixel_num).item()),
                "bpp": bpp_y + bpp_z, "bpp_y": bpp_y, "bpp_z": bpp_z}

    def compress(self, x, q_scale):
        """image_model.py:139-160; stream order z, y step 0, y step 1"""
        zs = get_downsampled_shape(x.shape[2], x.shape[3], 64)
        ec = self.entropy_coder
        ec.reserve(self.N * (x.shape[2] // 16 * (x.shape[3] // 16) + zs[0] * zs[1]))
        _, coded, z_hat, z_sym = self._encode_side(x, q_scale, True)
        ec.reset()
        self.encode_z(z_sym, z_hat, "bit_estimator_z")
        self.encode_y(coded)
        ec.flush()
        return {"bit_stream": ec.get_encoded_stream()}

    def decompress(self, bit_stream, height, width, q_scale):
        """image_model.py:162-171"""
        net = self.net
        ec = self.entropy_coder
        cq = self.curr_q("q_basic", q_scale)
        z_size = get_downsampled_shape(height, width, 64)
        yh, yw = get_downsampled_shape(height, width, 16)
        ec.reserve(self.N * (yh * yw + z_size[0] * z_size[1]))
        ec.set_stream(bit_stream)
        z_hat = self.decode_z(self.N, z_size, "bit_estimator_z")
        fused = three_convs(net, "y_prior_fusion", layers.hyper_dec(net, "hyper_dec", z_hat))
        y_hat = self.decompress_dual_prior(fused, "y_spatial_prior", cq)
        ec.finish_decode()
        x_hat = self._synthesis(y_hat, clamp=True)
        if not ec.lazy_check:
            ec.check()
        return {"x_hat": x_hat.nchw()}

    def encode_decode(self, x, q_scale, output_path=None, pic_width=None, pic_height=None):
        """image_model.py:116-137"""
        if output_path is None:
            return self.forward(x, q_scale)
        assert pic_height is not None
        assert pic_width is not None
        q_scale, q_index = sh.get_rounded_q(q_scale)
        compressed = self.compress(x, q_scale)
        sh.encode_i(pic_height, pic_width, q_index, compressed["bit_stream"], output_path)
        bit = sh.filesize(output_path) * 8
        height, width, q_index, bit_stream = sh.decode_i(output_path)
        decompressed = self.decompress(bit_stream, height, width, q_index / 100)
        return {"bit": bit, "x_hat": decompressed["x_hat"]}


# ------------------------------------------------------------------------------ inter
class DMC(HemCompressionModel):
    def __init__(self, anchor_num=4, coder_mode="hem", stream_part=1):
        super().__init__(harch.hem_dmc_spec(anchor_num), "laplace",
                         ("bit_estimator_z", "bit_estimator_z_mv"), anchor_num, coder_mode,
                         stream_part)

    @staticmethod
    def get_q_scales_from_ckpt(ckpt_path):
        """video_model.py:248-253 -> (y_q_scales, mv_y_q_scales)"""
        from ..stream_helper import get_state_dict
        ckpt = get_state_dict(ckpt_path)
        return ckpt["y_q_scale"].reshape(-1), ckpt["mv_y_q_scale"].reshape(-1)

    def _spynet(self, x, ref):
        """ME_Spynet (video_net.py:117-146), identical to DCVC-DC's"""
        net, dev, L = self.net, x.device, 4
        p1, p2 = [x], [ref]
        for lv in range(L - 1):
            p1.append(ops.avg_pool2(p1[lv]))
            p2.append(ops.avg_pool2(p2[lv]))
        flow = Act.new(p2[L - 1].h // 2, p2[L - 1].w // 2, 2, dev, zero=True)
        perm = (0, 1, 2, 3, 4, 5, -1, -1, 6, 7, -1, -1, -1, -1, -1, -1)
        for lv in range(L):
            k = L - 1 - lv
            h, w = p1[k].h, p1[k].w
            inp = Act.new(h, w, 16, dev, zero=True)
            ops.copy(p1[k], out=inp.slice(0, 3))
            up = ops.resize_bilinear(flow, h, w, mul=2.0, out=inp.slice(8, 2))
            ops.flow_warp(p2[k], up, out=inp.slice(3, 3))
            p = f"optic_flow.moduleBasic.{lv}"
            res = inp.slice(8, 4)
            if ops.f16_ok(h, w, 16):
                inp = ops.copy(inp, f16=True)
            t = ops.conv(inp, net.cw(p + ".conv1", cin_perm=perm), act=0.0, f16_out=True)
            t = ops.conv(t, net.cw(p + ".conv2"), act=0.0, f16_out=True)
            t = ops.conv(t, net.cw(p + ".conv3"), act=0.0, f16_out=True)
            t = ops.conv(t, net.cw(p + ".conv4"), act=0.0, f16_out=True)
            flow = ops.conv(t, net.cw(p + ".conv5", cout_pad=4), res=res).slice(0, 2)
        return flow

    def _mv_fused(self, mv_z_hat, dpb):
        """mv_hyper_prior_decoder + cat(ref_mv_y or zeros) + mv_y_prior_fusion (:273-279)"""
        net = self.net
        t = layers.hyper_dec(net, "mv_hyper_prior_decoder", mv_z_hat)      # 128 channels
        cat = Act.new(t.h, t.w, 192, t.device)
        ops.copy(t, out=cat.slice(0, 128))
        if dpb["ref_mv_y"] is None:
            ops.fill(cat.slice(128, 64), 0.0)
        else:
            ops.copy(Act.wrap(dpb["ref_mv_y"]), out=cat.slice(128, 64))
        return three_convs(net, "mv_y_prior_fusion", cat)

    def _y_fused(self, z_hat, dpb, ctx3):
        """temporal_prior_encoder + contextual_hyper_prior_decoder + cat(ref_y or zeros) +
        y_prior_fusion (:291-299)"""
        net = self.net
        h, w = ctx3.h // 4, ctx3.w // 4
        cat = Act.new(h, w, 480, ctx3.device)
        t = ops.conv(ctx3, net.cw("temporal_prior_encoder.0"), stride=2, act=0.1)
        ops.conv(t, net.cw("temporal_prior_encoder.2"), stride=2, out=cat.slice(0, 192))
        hier = layers.hyper_dec(net, "contextual_hyper_prior_decoder", z_hat)
        ops.copy(hier, out=cat.slice(192, 192))
        if dpb["ref_y"] is None:
            ops.fill(cat.slice(384, 96), 0.0)
        else:
            ops.copy(Act.wrap(dpb["ref_y"]), out=cat.slice(384, 96))
        return three_convs(net, "y_prior_fusion", cat)

    def _motion_compensation(self, dpb, mv):
        """motion_compensation + multi_scale_feature_extractor + MultiScaleContextFusion
        (:233-246, :17-68)"""
        net = self.net
        H, W, dev = mv.h, mv.w, mv.device
        mv2 = ops.resize_bilinear(mv, H // 2, W // 2, mul=0.5)
        mv3 = ops.resize_bilinear(mv2, H // 4, W // 4, mul=0.5)
        # binary16 for everything only tensor-core convolutions, the warps and copies read (the
        # same rule as dcvc_b200/video_model.py: feature levels, fusion concatenations, contexts)
        f16 = ops.f16_ok(H // 8, W // 8, 64)
        if dpb["ref_feature"] is None:
            feat = ops.conv(Act.wrap(dpb["ref_frame"]), net.cw("feature_adaptor_I"), f16_out=True)
        else:
            feat = ops.conv(Act.wrap(dpb["ref_feature"]), net.cw("feature_adaptor_P"), f16_out=True)
        fe = "feature_extractor"
        t, ta = layers.conv_for_res_block(net, fe + ".conv1", feat, 0.01, f16_out=True)
        l1 = layers.res_block(net, fe + ".res_block1", t, f16_out=f16, x_act=ta)
        t, ta = layers.conv_for_res_block(net, fe + ".conv2", l1, 0.01, stride=2, f16_out=True)
        l2 = layers.res_block(net, fe + ".res_block2", t, f16_out=f16, x_act=ta)
        t, ta = layers.conv_for_res_block(net, fe + ".conv3", l2, 0.01, stride=2, f16_out=True)
        l3 = layers.res_block(net, fe + ".res_block3", t, f16_out=f16, x_act=ta)
        cat1 = Act.new(H, W, 128, dev, f16=f16)                  # [context2_up | context1]
        cat2 = Act.new(H // 2, W // 2, 128, dev, f16=f16)        # [context3_up | context2]
        c1 = ops.flow_warp(l1, mv, out=cat1.slice(64, 64))
        c2 = ops.flow_warp(l2, mv2, out=cat2.slice(64, 64))
        c3 = ops.flow_warp(l3, mv3, out=Act.new(H // 4, W // 4, 64, dev, f16=f16))
        cf = "context_fusion_net"
        t = layers.subpel(net, cf + ".conv3_up", c3, f16_out=True)
        layers.res_block(net, cf + ".res_block3_up", t, out=cat2.slice(0, 64))
        t, ta = layers.conv_for_res_block(net, cf + ".conv3_out", c3, 0.01, f16_out=True)
        # the contexts are written straight into the decoder-side concatenation buffers
        # ([up 64 | ctx3 64], [up 64 | ctx2 64], [feature 32 | ctx1 64]; _recon recognises them):
        # no copy there, and binary16 like every tensor only convolutions and copies read
        D3 = Act.new(H // 4, W // 4, 128, dev, f16=f16)
        D2 = Act.new(H // 2, W // 2, 128, dev, f16=f16)
        R = Act.new(H, W, 96, dev, f16=f16)
        ctx3 = layers.res_block(net, cf + ".res_block3_out", t, res2=c3, out=D3.slice(64, 64),
                                x_act=ta)
        t = layers.subpel(net, cf + ".conv2_up", cat2, f16_out=True)
        layers.res_block(net, cf + ".res_block2_up", t, out=cat1.slice(0, 64))
        t, ta = layers.conv_for_res_block(net, cf + ".conv2_out", cat2, 0.01, f16_out=True)
        ctx2 = layers.res_block(net, cf + ".res_block2_out", t, res2=c2, out=D2.slice(64, 64),
                                x_act=ta)
        t, ta = layers.conv_for_res_block(net, cf + ".conv1_out", cat1, 0.01, f16_out=True)
        ctx1 = layers.res_block(net, cf + ".res_block1_out", t, res2=c1, out=R.slice(32, 64),
                                x_act=ta)
        return ctx1, ctx2, ctx3

    def _contextual_encoder(self, x, ctx1, ctx2, ctx3):
        """ContextualEncoder (:71-91); conv1's input is laid out [ctx1 64 | x 3]"""
        net, dev = self.net, x.device
        e1 = Act.new(x.h, x.w, 67, dev, f16=ctx1.f16)
        ops.copy(ctx1, out=e1.slice(0, 64))
        ops.copy(x, out=e1.slice(64, 3))
        perm = tuple(range(3, 67)) + (0, 1, 2)
        e2 = Act.new(x.h // 2, x.w // 2, 128, dev, f16=ctx2.f16)
        ops.conv(e1, net.cw("contextual_encoder.conv1", cin_perm=perm), stride=2, out=e2.slice(0, 64))
        ops.copy(ctx2, out=e2.slice(64, 64))
        t = layers.res_block(net, "contextual_encoder.res1", e2, 0.1, True, f16_out=ctx2.f16)
        e3 = Act.new(x.h // 4, x.w // 4, 128, dev, f16=ctx3.f16)
        ops.conv(t, net.cw("contextual_encoder.conv2"), stride=2, out=e3.slice(0, 64))
        ops.copy(ctx3, out=e3.slice(64, 64))
        t = layers.res_block(net, "contextual_encoder.res2", e3, 0.1, True, f16_out=ctx3.f16)
        t = ops.conv(t, net.cw("contextual_encoder.conv3"), stride=2,
                     f16_out=ops.f16_ok(x.h // 16, x.w // 16, 8))
        return ops.conv(t, net.cw("contextual_encoder.conv4"), stride=2)

    def _recon(self, y_hat, ctx1, ctx2, ctx3):
        """ContextualDecoder + ReconGeneration (:94-128, :301-302)"""
        net, dev = self.net, y_hat.device
        cd = "contextual_decoder"
        t = layers.subpel(net, cd + ".up1", y_hat, f16_out=True)
        def cat_buffer(ctx, total, at):
            """[new `at` channels | ctx]: the buffer ctx already lives in (_motion_compensation),
            else a new one with ctx copied in"""
            if ctx.cs == total and ctx.c0 == at and ctx.c == total - at:
                return Act(ctx.buf, 0, total)
            b = Act.new(ctx.h, ctx.w, total, dev, f16=ctx.f16)
            ops.copy(ctx, out=b.slice(at, total - at))
            return b
        d3 = cat_buffer(ctx3, 128, 64)
        layers.subpel(net, cd + ".up2", t, out=d3.slice(0, 64))
        t = layers.res_block(net, cd + ".res1", d3, 0.1, True, f16_out=True)
        d2 = cat_buffer(ctx2, 128, 64)
        layers.subpel(net, cd + ".up3", t, out=d2.slice(0, 64))
        t = layers.res_block(net, cd + ".res2", d2, 0.1, True, f16_out=True)
        # recon_generation_net(recon_image_feature, context1) (:301): first_conv reads
        # [feature 32 | context1 64]
        r = cat_buffer(ctx1, 96, 32)
        layers.subpel(net, cd + ".up4", t, out=r.slice(0, 32))
        rg = "recon_generation_net"
        t = ops.conv(r, net.cw(rg + ".first_conv"), f16_out=True)
        t = hem_unet(net, rg + ".unet_1", t, f16_out=True)
        feature = hem_unet(net, rg + ".unet_2", t)
        x_hat4 = ops.conv(feature, net.cw(rg + ".recon_conv", cout_pad=4), clamp01=False)
        return x_hat4.slice(0, 3), feature

    def _encode_side(self, x, dpb, mv_y_q_scale, y_q_scale, write, bits=None):
        net = self.net
        b = (lambda i: bits[i:i + 1]) if bits is not None else (lambda i: None)
        mv_q = self.curr_q("mv_y_q_basic", mv_y_q_scale)
        y_q = self.curr_q("y_q_basic", y_q_scale)
        xa, ref = Act.wrap(x), Act.wrap(dpb["ref_frame"])
        est_mv = self._spynet(xa, ref)
        mv_y = ops.channel_scale(analysis(net, "mv_encoder", est_mv), mv_q, divide=True)
        mv_z = layers.hyper_enc5(net, "mv_hyper_prior_encoder", mv_y)
        mv_z_hat, mv_z_sym = self.quant_z(mv_z, "bit_estimator_z_mv", write, b(3))
        mv_fused = self._mv_fused(mv_z_hat, dpb)
        mv_y_hat, mv_coded = self.forward_dual_prior(mv_y, mv_fused, "mv_y_spatial_prior", mv_q,
                                                     write, b(2))
        mv_hat = synthesis(net, "mv_decoder", mv_y_hat)
        ctx1, ctx2, ctx3 = self._motion_compensation(dpb, mv_hat)
        y = ops.channel_scale(self._contextual_encoder(xa, ctx1, ctx2, ctx3), y_q, divide=True)
        z = layers.hyper_enc3(net, "contextual_hyper_prior_encoder", y)
        z_hat, z_sym = self.quant_z(z, "bit_estimator_z", write, b(1))
        fused = self._y_fused(z_hat, dpb, ctx3)
        y_hat, coded = self.forward_dual_prior(y, fused, "y_spatial_prior", y_q, write, b(0))
        x_hat, feature = self._recon(y_hat, ctx1, ctx2, ctx3)
        for k, v in (("est_mv", est_mv), ("mv_y", mv_y), ("mv_z_hat", mv_z_hat),
                     ("mv_fused", mv_fused), ("mv_y_hat", mv_y_hat), ("mv_hat", mv_hat),
                     ("ctx1", ctx1), ("ctx2", ctx2), ("ctx3", ctx3), ("y", y), ("z_hat", z_hat),
                     ("fused", fused), ("y_hat", y_hat), ("x_hat", x_hat), ("feature", feature)):
            self._tap("p_" + k, v)
        self._tap_coded("p_mv_y", mv_coded, mv_y)
        self._tap_coded("p_y", coded, y)
        new_dpb = {"ref_frame": x_hat.nchw(), "ref_feature": feature.nchw(),
                   "ref_y": y_hat.nchw(), "ref_mv_y": mv_y_hat.nchw()}
        return new_dpb, (mv_z_hat, mv_z_sym), (z_hat, z_sym), mv_coded, coded

    def _reserve(self, h, w):
        zs = get_downsampled_shape(h, w, 64)
        ys = get_downsampled_shape(h, w, 16)
        self.entropy_coder.reserve((64 + 96) * ys[0] * ys[1] + 128 * zs[0] * zs[1])

    def compress(self, x, dpb, mv_y_q_scale, y_q_scale):
        """video_model.py:263-324; stream order mv_z, mv_y 0-1, z, y 0-1.  As in the reference the
        returned frame is NOT clamped here and the key is spelt "dbp" (:315)."""
        ec = self.entropy_coder
        self._reserve(x.shape[2], x.shape[3])
        new_dpb, mv_zs, zs, mv_coded, coded = self._encode_side(x, dpb, mv_y_q_scale, y_q_scale, True)
        ec.reset()
        self.encode_z(mv_zs[1], mv_zs[0], "bit_estimator_z_mv")
        self.encode_y(mv_coded)
        self.encode_z(zs[1], zs[0], "bit_estimator_z")
        self.encode_y(coded)
        ec.flush()
        return {"dbp": new_dpb, "bit_stream": ec.get_encoded_stream()}

    def decompress(self, dpb, string, height, width, mv_y_q_scale, y_q_scale):
        """video_model.py:326-375"""
        net = self.net
        ec = self.entropy_coder
        mv_q = self.curr_q("mv_y_q_basic", mv_y_q_scale)
        y_q = self.curr_q("y_q_basic", y_q_scale)
        self._reserve(height, width)
        ec.set_stream(string)
        z_size = get_downsampled_shape(height, width, 64)
        mv_z_hat = self.decode_z(64, z_size, "bit_estimator_z_mv")
        mv_fused = self._mv_fused(mv_z_hat, dpb)
        mv_y_hat = self.decompress_dual_prior(mv_fused, "mv_y_spatial_prior", mv_q)
        mv_hat = synthesis(net, "mv_decoder", mv_y_hat)
        ctx1, ctx2, ctx3 = self._motion_compensation(dpb, mv_hat)
        z_hat = self.decode_z(64, z_size, "bit_estimator_z")
        fused = self._y_fused(z_hat, dpb, ctx3)
        y_hat = self.decompress_dual_prior(fused, "y_spatial_prior", y_q)
        ec.finish_decode()
        x_hat, feature = self._recon(y_hat, ctx1, ctx2, ctx3)
        recon = x_hat.nchw().clamp(0, 1)
        if not ec.lazy_check:
            ec.check()
        return {"dpb": {"ref_frame": recon, "ref_feature": feature.nchw(), "ref_y": y_hat.nchw(),
                        "ref_mv_y": mv_y_hat.nchw()}}

    def encode_decode(self, x, dpb, output_path=None, pic_width=None, pic_height=None,
                      mv_y_q_scale=None, y_q_scale=None):
        """video_model.py:377-415"""
        if output_path is not None:
            mv_y_q_scale, mv_y_q_index = sh.get_rounded_q(mv_y_q_scale)
            y_q_scale, y_q_index = sh.get_rounded_q(y_q_scale)
            encoded = self.compress(x, dpb, mv_y_q_scale, y_q_scale)
            sh.encode_p(encoded["bit_stream"], mv_y_q_index, y_q_index, output_path)
            bits = sh.filesize(output_path) * 8
            mv_y_q_index, y_q_index, string = sh.decode_p(output_path)
            start = time.time()
            decoded = self.decompress(dpb, string, pic_height, pic_width, mv_y_q_index / 100,
                                      y_q_index / 100)
            return {"dpb": decoded["dpb"], "bit": bits, "decoding_time": time.time() - start}
        encoded = self.forward_one_frame(x, dpb, mv_y_q_scale=mv_y_q_scale, y_q_scale=y_q_scale)
        return {"dpb": encoded["dpb"], "bit_y": encoded["bit_y"].item(),
                "bit_z": encoded["bit_z"].item(), "bit_mv_y": encoded["bit_mv_y"].item(),
                "bit_mv_z": encoded["bit_mv_z"].item(), "bit": encoded["bit"].item(),
                "decoding_time": 0}

    def forward_one_frame(self, x, dpb, mv_y_q_scale=None, y_q_scale=None):
        """video_model.py:417-513 (estimate mode; the training-loss entries mse / ssim / me_mse /
        BDQ are not produced)"""
        bits = torch.zeros(4, dtype=torch.float32, device=x.device)    # y, z, mv_y, mv_z
        new_dpb, _, _, _, _ = self._encode_side(x, dpb, mv_y_q_scale, y_q_scale, False, bits)
        pixel_num = x.shape[2] * x.shape[3]
        bpp_y, bpp_z = bits[0:1] / pixel_num, bits[1:2] / pixel_num
        bpp_mv_y, bpp_mv_z = bits[2:3] / pixel_num, bits[3:4] / pixel_num
        bpp = bpp_y + bpp_z + bpp_mv_y + bpp_mv_z
        return {"bpp_mv_y": bpp_mv_y, "bpp_mv_z": bpp_mv_z, "bpp_y": bpp_y, "bpp_z": bpp_z,
                "bpp": bpp, "dpb": new_dpb, "bit": torch.sum(bpp) * pixel_num,
                "bit_y": torch.sum(bpp_y) * pixel_num, "bit_z": torch.sum(bpp_z) * pixel_num,
                "bit_mv_y": torch.sum(bpp_mv_y) * pixel_num,
                "bit_mv_z": torch.sum(bpp_mv_z) * pixel_num}

    def forward(self, x, dpb, mv_y_q_scale=None, y_q_scale=None):
        return self.forward_one_frame(x, dpb, mv_y_q_scale=mv_y_q_scale, y_q_scale=y_q_scale)
