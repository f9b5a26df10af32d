"""`DMC` -- the DCVC-DC inter (P-frame) model behind the reference's module API
(DCVC-DC/src/models/video_model.py:235-628), executed as fused sm_100a kernel launches.

Same constructor signature, state_dict keys, method signatures, dpb contract and return dicts as
the reference.  dpb entries are [1,C,H,W] fp32 CUDA tensors (strided views of NHWC buffers);
they round-trip through the caller without copies.
"""
import time

import torch

from . import arch, layers, ops, tensor
from .common_model import CompressionModel, get_downsampled_shape
from .stream_helper import encode_p, decode_p, filesize
from .tensor import Act

CH = arch  # channel constants g_ch_* (video_model.py:19-23)


class DMC(CompressionModel):
    def __init__(self, anchor_num=4, ec_thread=False, stream_part=1, inplace=False,
                 coder_mode="compat"):
        del inplace
        super().__init__(arch.dmc_spec(anchor_num), "laplace",
                         ("bit_estimator_z", "bit_estimator_z_mv"), ec_thread, stream_part,
                         anchor_num, coder_mode)

    def _tap(self, name, act):
        if self.taps is not None:
            self.taps[name] = act.nchw().detach().float().clone()

    # ------------------------------------------------------------------ q
    @staticmethod
    def get_q_scales_from_ckpt(ckpt_path):
        """video_model.py:(static) -> y_q_scale_enc, y_q_scale_dec, mv_y_q_scale_enc, mv_y_q_scale_dec"""
        from .stream_helper import get_state_dict
        ckpt = get_state_dict(ckpt_path)
        return (ckpt["y_q_scale_enc"].reshape(-1), ckpt["y_q_scale_dec"].reshape(-1),
                ckpt["mv_y_q_scale_enc"].reshape(-1), ckpt["mv_y_q_scale_dec"].reshape(-1))

    def get_q_for_inference(self, q_in_ckpt, q_index):
        """video_model.py:413-423 -> (mv_y_q_enc, mv_y_q_dec, y_q_enc, y_q_dec)"""
        return tuple(self.get_curr_q(f"{n}_q_scale_{d}", f"{n}_q_basic_{d}", q_in_ckpt, q_index)
                     for n, d in (("mv_y", "enc"), ("mv_y", "dec"), ("y", "enc"), ("y", "dec")))

    # ------------------------------------------------------------------ motion estimation
    def _spynet(self, x, ref):
        """ME_Spynet.forward + MEBasic (video_net.py:79-126): avg-pool pyramids, per level
        bilinear x2 (*2), warp, 5 x conv7x7 with ReLU; the level input cat(im1, warp(im2), flow)
        is one 8-channel buffer, `flow_up +` is the last conv's residual."""
        net = self.net
        dev = x.device
        L = 4
        p1, p2 = [x], [ref]
        for lv in range(L - 1):
            p1.append(ops.avg_pool2(p1[lv]))
            p2.append(ops.avg_pool2(p2[lv]))
        flow = Act.new(p2[L - 1].h // 2, p2[L - 1].w // 2, 2, dev, zero=True)
        # level input, 16 channels [im1 3 | warp 3 | 0 0 | flow_up 2 | 0 0 | 0 0 0 0]: a 7x7 tap is
        # one K=16 MMA whatever the channel count, and [flow_up | 0 0] is a whole 16-byte group,
        # so conv5 (padded to 4 filters) takes the vectorised residual epilogue
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

    def _mv_enc(self, est_mv, ref_mv_feature, q):
        """MvEnc.forward (video_model.py:138-146)"""
        net = self.net
        m = "mv_encoder"
        # block outputs that only convolutions read are binary16 (f16_out)
        # (their stride-2 consumers must still be tensor-core sized: half() checks that)
        def half(a):
            return ops.f16_ok(a.h // 2, a.w // 2, a.c)
        t = layers.rbws(net, m + ".enc_1.0", est_mv)
        t = layers.dcb(net, m + ".enc_1.1", t, scale=q, f16_out=half(t))
        if ref_mv_feature is None:
            t = layers.rbws(net, m + ".enc_2", t)
            t = layers.dcb(net, m + ".adaptor_0", t, f16_out=half(t))
        else:
            cat = Act.new(t.h // 2, t.w // 2, 128, t.device)
            layers.rbws(net, m + ".enc_2", t, out=cat.slice(0, 64))
            ops.copy(ref_mv_feature, out=cat.slice(64, 64))
            t = layers.dcb(net, m + ".adaptor_1", cat, f16_out=half(cat))
        t = layers.rbws(net, m + ".enc_3.0", t)
        t = layers.dcb(net, m + ".enc_3.1", t, f16_out=half(t))
        return ops.conv(t, net.cw(m + ".enc_3.2"), stride=2)

    def _mv_dec(self, mv_y_hat, q):
        """MvDec.forward (video_model.py:165-170) -> (mv [2], mv_feature [64 @1/4])"""
        net = self.net
        m = "mv_decoder"
        t = layers.dcb(net, m + ".dec_1.0", mv_y_hat, f16_out=True)
        t = layers.rbu(net, m + ".dec_1.1", t)
        t = layers.dcb(net, m + ".dec_1.2", t, f16_out=True)
        t = layers.rbu(net, m + ".dec_1.3", t)
        feature = layers.dcb(net, m + ".dec_1.4", t,
                             out=tensor.placed("ref_mv_feature", t.h, t.w, 64))
        t = layers.rbu(net, m + ".dec_2", feature, scale=q)
        t = layers.dcb(net, m + ".dec_3.0", t, f16_out=True)
        # 8 -> 16 filters (two zero channels per pixel): planar packing, vectorised epilogue
        mv4 = ops.conv(t, net.cw(m + ".dec_3.1.0", shuffle=True, cout_pad=16), shuffle=True)
        return mv4.slice(0, 2), feature

    # ------------------------------------------------------------------ priors
    def _mv_params(self, mv_z_hat, dpb, catb, c):
        """mv_prior_param_decoder (video_model.py:375-385) -> params slice of catb"""
        net = self.net
        t = layers.hyper_dec(net, "mv_hyper_prior_decoder", mv_z_hat)
        ref = dpb["ref_mv_y"]
        if ref is None:
            if t.h != catb.h or t.w != catb.w:
                t = ops.copy(t, h=catb.h, w=catb.w)
            t = layers.dcb(net, "mv_y_prior_fusion_adaptor_0", t, f16_out=True)
        else:
            cat = Act.new(catb.h, catb.w, 128, catb.device)
            ops.copy(t, out=cat.slice(0, 64))
            ops.copy(Act.wrap(ref), out=cat.slice(64, 64))
            t = layers.dcb(net, "mv_y_prior_fusion_adaptor_1", cat, f16_out=True)
        t = layers.dcb(net, "mv_y_prior_fusion.0", t, f16_out=True)
        layers.dcb(net, "mv_y_prior_fusion.1", t, out=catb.slice(c, 3 * c))

    def _res_params(self, z_hat, dpb, ctx3, catb, c):
        """res_prior_param_decoder (video_model.py:387-399) incl. temporal_prior_encoder"""
        net = self.net
        ref = dpb["ref_y"]
        cat = Act.new(catb.h, catb.w, 256 if ref is None else 384, catb.device)
        t = ops.conv(ctx3, net.cw("temporal_prior_encoder.0"), stride=2, act=0.1)
        ops.conv(t, net.cw("temporal_prior_encoder.2"), stride=2, out=cat.slice(0, 128))
        hier = layers.hyper_dec(net, "contextual_hyper_prior_decoder", z_hat)
        ops.copy(hier, out=cat.slice(128, 128))
        if ref is None:
            t = layers.dcb(net, "y_prior_fusion_adaptor_0", cat, f16_out=True)
        else:
            ops.copy(Act.wrap(ref), out=cat.slice(256, 128))
            t = layers.dcb(net, "y_prior_fusion_adaptor_1", cat, f16_out=True)
        t = layers.dcb(net, "y_prior_fusion.0", t, f16_out=True)
        layers.dcb(net, "y_prior_fusion.1", t, out=catb.slice(c, 3 * c))

    def _prior_fn(self, prefix):
        net = self.net

        def fn(step, catb):
            t = ops.conv(catb, net.cw(f"{prefix}_spatial_prior_adaptor_{step}"), f16_out=True)
            for i in range(3):
                t = layers.dcb(net, f"{prefix}_spatial_prior.{i}", t, f16_out=i < 2)
            return t
        return fn

    # ------------------------------------------------------------------ motion compensation
    def _motion_compensation(self, dpb, mv, frame_idx, bufs):
        """motion_compensation + multi_scale_feature_extractor + OffsetDiversity +
        MultiScaleContextFusion (video_model.py:26-118,343-364).  Final contexts are written into
        the decoder-side concatenation buffers `bufs` (R: [ctx1|res], D2: [f|ctx2], D3: [f|ctx3])."""
        net = self.net
        dev = mv.device
        ref_frame = Act.wrap(dpb["ref_frame"])
        H, W = mv.h, mv.w
        # [context1_init 48 | warpframe 3 | mv 2]: only align.conv_offset.0 (stride 2) reads it
        aux = Act.new(H, W, 53, dev, f16=ops.f16_ok(H // 2, W // 2, 56))
        if aux.f16:      # the five narrow channels go through one fp32 scratch pixel group
            tmp = Act.new(H, W, 5, dev, cstride=8)
            ops.flow_warp(ref_frame, mv, out=tmp.slice(0, 3))
            ops.copy(mv, out=tmp.slice(3, 2))
            ops.copy(tmp, out=aux.slice(48, 5))
        else:
            ops.flow_warp(ref_frame, mv, out=aux.slice(48, 3))
            ops.copy(mv, out=aux.slice(51, 2))
        mv2 = ops.resize_bilinear(mv, H // 2, W // 2, mul=0.5)
        mv3 = ops.resize_bilinear(mv2, H // 4, W // 4, mul=0.5)
        if dpb["ref_feature"] is None:
            feat = ops.conv(ref_frame, net.cw("feature_adaptor_I"), f16_out=True)
        else:
            k = [0, 1, 0, 2][frame_idx % 4]
            feat = ops.conv(Act.wrap(dpb["ref_feature"]), net.cw(f"feature_adaptor.{k}"),
                            f16_out=True)
        fe = "feature_extractor"
        # a ResBlock's input is private to the block (LeakyReLU copy + residual): binary16
        # the three feature levels are read by the warps / offset diversity (binary16 -> binary16
        # gathers) and by stride-2 convolutions only: binary16 when every consumer's buffer is
        fe16 = aux.f16 and ops.f16_ok(H // 8, W // 8, 96)
        t, ta = layers.conv_for_res_block(net, fe + ".conv1", feat, 0.01, f16_out=True)
        l1 = layers.res_block(net, fe + ".res_block1", t, f16_out=fe16, x_act=ta)
        t, ta = layers.conv_for_res_block(net, fe + ".conv2", l1, 0.01, stride=2, f16_out=True)
        l2 = layers.res_block(net, fe + ".res_block2", t, f16_out=fe16, x_act=ta)
        t, ta = layers.conv_for_res_block(net, fe + ".conv3", l2, 0.01, stride=2, f16_out=True)
        l3 = layers.res_block(net, fe + ".res_block3", t, f16_out=fe16, x_act=ta)
        ops.flow_warp(l1, mv, out=aux.slice(0, 48))
        # OffsetDiversity (align)
        t = ops.conv(aux, net.cw("align.conv_offset.0"), stride=2, act=0.1, f16_out=True)
        t = ops.conv(t, net.cw("align.conv_offset.2"), act=0.1, f16_out=True)
        offs = ops.conv(t, net.cw("align.conv_offset.4"))
        # the fusion net's concatenations are read by tensor-core convolutions only (as input
        # and as the second residual): binary16 in tensor-core mode
        cat1 = Act.new(H, W, 96, dev,                      # [context2_up 48 | context1 48]
                       f16=ops.f16_ok(H, W, 96))
        cat2 = Act.new(H // 2, W // 2, 128, dev,           # [context3_up 64 | context2 64]
                       f16=ops.f16_ok(H // 2, W // 2, 128))
        fw = net.sd["align.fusion.weight"].reshape(48, 6)
        c1 = ops.offset_diversity(l1, offs, mv, fw, net.sd["align.fusion.bias"],
                                  out=cat1.slice(48, 48))
        c2 = ops.flow_warp(l2, mv2, out=cat2.slice(64, 64))
        # c3 is read by two convolutions and as a second residual: binary16
        c3 = ops.flow_warp(l3, mv3, out=Act.new(H // 4, W // 4, l3.c, dev,
                                                f16=ops.f16_ok(H // 4, W // 4, l3.c)))
        # MultiScaleContextFusion
        cf = "context_fusion_net"
        t = layers.subpel(net, cf + ".conv3_up", c3, f16_out=True)
        layers.res_block(net, cf + ".res_block3_up", t, out=cat2.slice(0, 64))
        t, ta = layers.conv_for_res_block(net, cf + ".conv3_out", c3, 0.01, f16_out=True)
        ctx3 = layers.res_block(net, cf + ".res_block3_out", t, res2=c3, out=bufs["ctx3"],
                                x_act=ta)
        t = layers.subpel(net, cf + ".conv2_up", cat2, f16_out=True)
        layers.res_block(net, cf + ".res_block2_up", t, out=cat1.slice(0, 48))
        t, ta = layers.conv_for_res_block(net, cf + ".conv2_out", cat2, 0.01, f16_out=True)
        ctx2 = layers.res_block(net, cf + ".res_block2_out", t, res2=c2, out=bufs["ctx2"],
                                x_act=ta)
        t, ta = layers.conv_for_res_block(net, cf + ".conv1_out", cat1, 0.01, f16_out=True)
        ctx1 = layers.res_block(net, cf + ".res_block1_out", t, res2=c1, out=bufs["ctx1"],
                                x_act=ta)
        return ctx1, ctx2, ctx3

    def _new_bufs(self, H, W, dev):
        # recon input [ctx1 48 | res 32]: two convolutions write it, two read it -> binary16
        # (the encoder's stride-2 conv1 reads ctx1 too: its output must be tensor-core sized)
        R = Act.new(H, W, 80, dev, f16=ops.f16_ok(H // 2, W // 2, 80))
        # decoder res2 / res1 inputs [f 64 | ctx2 64], [f 96 | ctx3 96]: convolutions, LeakyReLU
        # copies and the encoder's own concatenations read them; ctx3 also feeds a stride-2 conv
        f16 = ops.f16_ok(H // 8, W // 8, 64)
        D2 = Act.new(H // 2, W // 2, 128, dev, f16=f16)
        D3 = Act.new(H // 4, W // 4, 192, dev, f16=f16)
        return {"R": R, "D2": D2, "D3": D3, "ctx1": R.slice(0, 48), "ctx2": D2.slice(64, 64),
                "ctx3": D3.slice(96, 96)}

    # ------------------------------------------------------------------ contextual coder
    def _contextual_encoder(self, x, ctx1, ctx2, ctx3, q):
        """ContextualEncoder.forward (video_model.py:185-193); conv1's input is laid out
        [ctx1 48 | x 3] (weights permuted accordingly) to keep the 48-channel slice aligned."""
        net = self.net
        dev = x.device
        e1 = Act.new(x.h, x.w, 51, dev, f16=ctx1.f16)
        ops.copy(ctx1, out=e1.slice(0, 48))
        ops.copy(x, out=e1.slice(48, 3))
        perm = tuple(range(3, 51)) + (0, 1, 2)
        e2 = Act.new(x.h // 2, x.w // 2, 128, dev, f16=ctx2.f16)
        ops.conv(e1, net.cw("contextual_encoder.conv1", cin_perm=perm), stride=2,
                 out=e2.slice(0, 64))
        ops.copy(ctx2, out=e2.slice(64, 64))
        t = layers.res_block(net, "contextual_encoder.res1", e2, 0.1, True, scale=q,
                             f16_out=ctx3.f16)
        e3 = Act.new(x.h // 4, x.w // 4, 192, dev, f16=ctx3.f16)
        ops.conv(t, net.cw("contextual_encoder.conv2"), stride=2, out=e3.slice(0, 96))
        ops.copy(ctx3, out=e3.slice(96, 96))
        t = layers.res_block(net, "contextual_encoder.res2", e3, 0.1, True, f16_out=ctx3.f16)
        t = ops.conv(t, net.cw("contextual_encoder.conv3"), stride=2)
        return ops.conv(t, net.cw("contextual_encoder.conv4"), stride=2)

    def _recon(self, y_hat, bufs, q):
        """get_recon_and_feature = ContextualDecoder + ReconGeneration (video_model.py:208-232,
        401-405); x_hat.clamp_(0,1) is the last conv's epilogue."""
        net = self.net
        cd = "contextual_decoder"
        t = layers.subpel(net, cd + ".up1", y_hat)
        layers.subpel(net, cd + ".up2", t, out=bufs["D3"].slice(0, 96))
        t = layers.res_block(net, cd + ".res1", bufs["D3"], 0.1, True, f16_out=True)
        layers.subpel(net, cd + ".up3", t, scale=q, out=bufs["D2"].slice(0, 64))
        t = layers.res_block(net, cd + ".res2", bufs["D2"], 0.1, True, f16_out=True)
        layers.subpel(net, cd + ".up4", t, out=bufs["R"].slice(48, 32))
        rg = "recon_generation_net"
        # the reference calls recon_generation_net(recon_image_feature, context1) (:403), so
        # first_conv's input channels are [res 32 | ctx1 48]; the buffer is [ctx1 | res]
        perm = tuple(range(32, 80)) + tuple(range(0, 32))
        t = ops.conv(bufs["R"], net.cw(rg + ".first_conv", cin_perm=perm), f16_out=True)
        t = layers.unet(net, rg + ".unet_1", t, layers.dcb, f16_out=True)
        feature = layers.unet(net, rg + ".unet_2", t, layers.dcb,
                              out=tensor.placed("ref_feature", t.h, t.w, 48))
        # 3 -> 4 output channels (a zero filter): the image pixel is a 16-byte group either way,
        # and whole groups take the TMA-store epilogue instead of the scalar one
        xp = tensor.placed("ref_frame", t.h, t.w, 3)
        x_hat4 = ops.conv(feature, net.cw(rg + ".recon_conv", cout_pad=4), clamp01=True,
                          out=Act(xp.buf, 0, 4) if xp is not None and xp.cs == 4 else None)
        return x_hat4.slice(0, 3), feature

    # ------------------------------------------------------------------ shared encoder side
    def _encode_side(self, x, dpb, q_in_ckpt, q_index, frame_idx, write, bits=None, coder=None,
                     reconstruct=True):
        """compress (:425-453) / forward_one_frame (:559-585) up to the entropy coder.
        `coder(mv_z, z, mv_coded, coded)` is called once every symbol of the frame is known and
        before the reconstruction is enqueued."""
        mv_q_enc, mv_q_dec, y_q_enc, y_q_dec = self.get_q_for_inference(q_in_ckpt, q_index)
        xa = Act.wrap(x)
        ref = Act.wrap(dpb["ref_frame"])
        H, W = xa.h, xa.w
        b = (lambda i: bits[i:i + 1]) if bits is not None else (lambda i: None)
        est_mv = self._spynet(xa, ref)
        rmf = dpb["ref_mv_feature"]
        mv_y = self._mv_enc(est_mv, Act.wrap(rmf) if rmf is not None else None, mv_q_enc)
        mv_z = layers.hyper_enc5(self.net, "mv_hyper_prior_encoder", self.pad_for_y(mv_y))
        mv_z_hat, mv_z_sym = self.quant_z(mv_z, "bit_estimator_z_mv", write, b(3))
        mv_catb = self.new_prior_buffer(mv_y.h, mv_y.w, mv_y.c)
        self._mv_params(mv_z_hat, dpb, mv_catb, mv_y.c)
        mv_y_hat, mv_coded = self.forward_four_part_prior(mv_y, mv_catb, self._prior_fn("mv_y"),
                                                          write, b(2), tag="ref_mv_y")
        mv_hat, mv_feature = self._mv_dec(mv_y_hat, mv_q_dec)
        bufs = self._new_bufs(H, W, xa.device)
        ctx1, ctx2, ctx3 = self._motion_compensation(dpb, mv_hat, frame_idx, bufs)
        y = self._contextual_encoder(xa, ctx1, ctx2, ctx3, y_q_enc)
        z = layers.hyper_enc3(self.net, "contextual_hyper_prior_encoder", self.pad_for_y(y))
        z_hat, z_sym = self.quant_z(z, "bit_estimator_z", write, b(1))
        catb = self.new_prior_buffer(y.h, y.w, y.c)
        self._res_params(z_hat, dpb, ctx3, catb, y.c)
        y_hat, coded = self.forward_four_part_prior(y, catb, self._prior_fn("y"), write, b(0),
                                                    tag="ref_y")
        if coder is not None:
            coder((mv_z_hat, mv_z_sym), (z_hat, z_sym), mv_coded, coded)
        if not reconstruct:     # encode_decode: the decoder produces the (bit-identical) dpb
            return None, (mv_z_hat, mv_z_sym), (z_hat, z_sym), mv_coded, coded
        x_hat, feature = self._recon(y_hat, bufs, y_q_dec)
        if self.taps is not None:
            for k, v in (("est_mv", est_mv), ("mv_y", mv_y), ("mv_z_hat", mv_z_hat),
                         ("mv_params", mv_catb.slice(mv_y.c, 3 * mv_y.c)), ("mv_y_hat", mv_y_hat),
                         ("mv_hat", mv_hat), ("ctx1", ctx1), ("ctx2", ctx2), ("ctx3", ctx3),
                         ("y", y), ("z_hat", z_hat), ("params", catb.slice(y.c, 3 * y.c)),
                         ("y_hat", y_hat), ("x_hat", x_hat), ("feature", feature),
                         ("recon_res", bufs["R"].slice(48, 32))):
                self._tap("p_" + k, v)
            if write:       # coded symbols / CDF indexes per step, the reference's [1,C/4,h,w] order
                for name, cd, lat in (("mv_y", mv_coded, mv_y), ("y", coded, y)):
                    for st, (sym, idx) in enumerate(cd):
                        shp = (1, lat.c // 4, lat.h, lat.w)
                        self.taps[f"p_{name}_q_w_{st}"] = sym.view(shp).clone()
                        self.taps[f"p_{name}_idx_w_{st}"] = idx.view(shp).clone()
        new_dpb = {"ref_frame": x_hat.nchw(), "ref_feature": feature.nchw(),
                   "ref_mv_feature": mv_feature.nchw(), "ref_y": y_hat.nchw(),
                   "ref_mv_y": mv_y_hat.nchw()}
        return new_dpb, (mv_z_hat, mv_z_sym), (z_hat, z_sym), mv_coded, coded

    # ------------------------------------------------------------------ reference API
    def compress(self, x, dpb, q_in_ckpt, q_index, frame_idx, defer=False, reconstruct=True):
        """video_model.py:425-481; stream order mv_z, z, mv_y parts 0..3, y parts 0..3.
        The coder kernels are enqueued BEFORE the frame's reconstruction (the reference hands the
        symbols to its encoder thread at the same point, :449-476, and reconstructs after); the
        host then waits for the coder's event only, so the reconstruction is still running on the
        GPU when this returns -- whoever reads the dpb next does so in stream order.
        reconstruct=False (used by encode_decode, which takes the dpb from its own decompress
        call as the reference does, :535-543) skips the encoder-side reconstruction, whose only
        product is the returned dpb; the stream is unaffected."""
        ec = self.entropy_coder
        ec.reserve(self._frame_symbols(x.shape[2], x.shape[3]))

        def coder(mv_zs, zs, mv_coded, coded):
            ec.reset()
            self.encode_z(mv_zs[1], mv_zs[0], "bit_estimator_z_mv")
            self.encode_z(zs[1], zs[0], "bit_estimator_z")
            self.encode_y(mv_coded)
            self.encode_y(coded)
            ec.flush()

        new_dpb, _, _, _, _ = self._encode_side(x, dpb, q_in_ckpt, q_index, frame_idx, True,
                                                coder=coder, reconstruct=reconstruct)
        if defer:      # (pipeline.py) "bit_stream" is a callable: the host wait happens when called
            return {"dpb": new_dpb, "bit_stream": ec.get_encoded_stream}
        return {"dpb": new_dpb, "bit_stream": ec.get_encoded_stream()}

    def decompress(self, dpb, string, height, width, q_in_ckpt, q_index, frame_idx):
        """video_model.py:483-520"""
        ec = self.entropy_coder
        self._decode_begin(string, height, width)
        out = self._decompress_body(dpb, height, width, q_in_ckpt, q_index, frame_idx, True)
        if not ec.lazy_check:       # a truncated / corrupt / wrong-table frame raises here
            ec.check()
        return out

    @staticmethod
    def _frame_symbols(height, width):
        """symbols one P frame hands to the coder: mv_y + y latents and their hyper latents"""
        zs = get_downsampled_shape(height, width, 64)
        ys = get_downsampled_shape(height, width, 16)
        return (64 + arch.CH_16X) * (ys[0] * ys[1] + zs[0] * zs[1])

    def _decode_begin(self, string, height, width):
        """host part of decompress: stream parsing, upload, decoder state (not graph-capturable)"""
        ec = self.entropy_coder
        ec.reserve(self._frame_symbols(height, width))
        ec.set_stream(string)

    def _decompress_body(self, dpb, height, width, q_in_ckpt, q_index, frame_idx, finish):
        """every kernel of decompress, in stream order (this is what dcvc_b200.graphs records);
        finish: enqueue the decoder's error-word copy right after the last entropy decode"""
        _, mv_q_dec, _, y_q_dec = self.get_q_for_inference(q_in_ckpt, q_index)
        dev = self.device
        ec = self.entropy_coder
        z_size = get_downsampled_shape(height, width, 64)
        yh, yw = get_downsampled_shape(height, width, 16)
        mv_z_hat = self.decode_z(64, z_size, "bit_estimator_z_mv")
        z_hat = self.decode_z(arch.CH_16X, z_size, "bit_estimator_z")
        mv_catb = self.new_prior_buffer(yh, yw, 64)
        self._mv_params(mv_z_hat, dpb, mv_catb, 64)
        mv_y_hat = self.decompress_four_part_prior(mv_catb, 64, self._prior_fn("mv_y"),
                                                   tag="ref_mv_y")
        mv_hat, mv_feature = self._mv_dec(mv_y_hat, mv_q_dec)
        bufs = self._new_bufs(mv_hat.h, mv_hat.w, dev)
        _, _, ctx3 = self._motion_compensation(dpb, mv_hat, frame_idx, bufs)
        catb = self.new_prior_buffer(yh, yw, arch.CH_16X)
        self._res_params(z_hat, dpb, ctx3, catb, arch.CH_16X)
        y_hat = self.decompress_four_part_prior(catb, arch.CH_16X, self._prior_fn("y"),
                                                tag="ref_y")
        if finish:
            ec.finish_decode()
        x_hat, feature = self._recon(y_hat, bufs, y_q_dec)
        return {"dpb": {"ref_frame": x_hat.nchw(), "ref_feature": feature.nchw(),
                        "ref_mv_feature": mv_feature.nchw(), "ref_y": y_hat.nchw(),
                        "ref_mv_y": mv_y_hat.nchw()}}

    def encode_decode(self, x, dpb, q_in_ckpt, q_index, output_path=None, pic_width=None,
                      pic_height=None, frame_idx=0):
        """video_model.py:522-557"""
        if output_path is not None:
            device = x.device
            torch.cuda.synchronize(device=device)
            t0 = time.time()
            # the reference discards compress()'s dpb here and continues from the decoder's
            # (:535-543); both are bit-identical, so only the decoder reconstructs the frame
            run = self.__dict__.get("_graph_runner") or self      # dcvc_b200.graphs, when enabled
            encoded = run.compress(x, dpb, q_in_ckpt, q_index, frame_idx,
                                   reconstruct=self.taps is not None)
            encode_p(encoded["bit_stream"], q_in_ckpt, q_index, frame_idx, output_path)
            bits = filesize(output_path) * 8
            torch.cuda.synchronize(device=device)
            t1 = time.time()
            q_in_ckpt, q_index, frame_idx, string = decode_p(output_path)
            decoded = run.decompress(dpb, string, pic_height, pic_width, q_in_ckpt, q_index,
                                     frame_idx)
            torch.cuda.synchronize(device=device)
            t2 = time.time()
            return {"dpb": decoded["dpb"], "bit": bits, "encoding_time": t1 - t0,
                    "decoding_time": t2 - t1}
        encoded = self.forward_one_frame(x, dpb, q_in_ckpt=q_in_ckpt, q_index=q_index,
                                         frame_idx=frame_idx)
        return {"dpb": encoded["dpb"], "bit": encoded["bit"].item(), "encoding_time": 0,
                "decoding_time": 0}

    def forward_one_frame(self, x, dpb, q_in_ckpt=False, q_index=None, frame_idx=0):
        """video_model.py:559-628 (estimate mode)"""
        bits = torch.zeros(4, dtype=torch.float32, device=x.device)   # y, z, mv_y, mv_z
        new_dpb, _, _, _, _ = self._encode_side(x, dpb, q_in_ckpt, q_index, frame_idx, False, bits)
        pixel_num = x.shape[2] * x.shape[3]
        bpp_y, bpp_z = bits[0:1] / pixel_num, bits[1:2] / pixel_num
        bpp_mv_y, bpp_mv_z = bits[2:3] / pixel_num, bits[3:4] / pixel_num
        bpp = bpp_y + bpp_z + bpp_mv_y + bpp_mv_z
        return {"bpp_mv_y": bpp_mv_y, "bpp_mv_z": bpp_mv_z, "bpp_y": bpp_y, "bpp_z": bpp_z,
                "bpp": bpp, "dpb": new_dpb, "bit": torch.sum(bpp) * pixel_num,
                "bit_y": torch.sum(bpp_y) * pixel_num, "bit_z": torch.sum(bpp_z) * pixel_num,
                "bit_mv_y": torch.sum(bpp_mv_y) * pixel_num,
                "bit_mv_z": torch.sum(bpp_mv_z) * pixel_num}
