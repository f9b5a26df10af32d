"""TEST INFRASTRUCTURE -- CPU oracle of the DCVC-DC per-frame coding path.

`OracleIntra` / `OracleDMC` restate IntraNoAR / DMC (reference:
DCVC-DC/src/models/image_model.py:61-252, video_model.py:235-628) on top of
oracle/nets_oracle.py (network arithmetic), oracle/entropy_oracle.py (entropy model) and
oracle/rans_oracle.c (coder).  torch fp32 on CPU.  They are the checker for tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline; the product never imports them.

`taps` (a dict) receives named intermediates so that tests can teacher-force individual
kernels with the oracle's inputs.
"""
import numpy as np
import torch

from . import entropy_oracle as eo
from . import nets_oracle as no
from . import rans_oracle


def fine_q_table(anchors):
    """load_state_dict's 64-point table, video_model.py:328-341 / image_model.py:161-167:
    np.exp(np.linspace(np.log(a0), np.log(a3), 64)) on fp32 scalars -> float64 table."""
    a = anchors.reshape(-1).detach().cpu().numpy().astype(np.float32)
    return np.exp(np.linspace(np.log(a[0]), np.log(a[3]), 64))


def downsampled_shape(height, width, p):
    """utils/stream_helper.py:34-37"""
    nh = (height + p - 1) // p * p
    nw = (width + p - 1) // p * p
    return int(nh / p + 0.5), int(nw / p + 0.5)


class _Coder:
    """EntropyCoder facade, entropy_models.py:9-55, on the oracle's C coder."""

    def __init__(self, stream_part):
        self.enc = rans_oracle.RansEncoder(False, stream_part)
        self.dec = rans_oracle.RansDecoder(stream_part)

    def reset(self):
        self.enc.reset()

    def encode(self, symbols, indexes, tables):
        s = symbols.reshape(-1).clamp(-30000, 30000).to(torch.int16).numpy()
        i = indexes.reshape(-1).to(torch.int16).numpy()
        self.enc.encode_with_indexes(s, i, *tables)

    def flush(self):
        self.enc.flush()
        return self.enc.get_encoded_stream().tobytes()

    def set_stream(self, stream):
        self.dec.set_stream(stream)

    def decode(self, indexes, tables):
        i = indexes.reshape(-1).to(torch.int16).numpy()
        out = self.dec.decode_stream(i, *tables)
        return torch.from_numpy(out.astype(np.float32)).reshape(indexes.shape)


class _Base:
    distribution = "laplace"

    def __init__(self, state_dict, stream_part=1):
        self.sd = {k: v.detach().float().cpu() for k, v in state_dict.items()}
        self.stream_part = stream_part
        self.coder = None
        self.y_tables = None
        self.taps = None

    def tap(self, name, t):
        if self.taps is not None:
            self.taps[name] = t.detach().clone() if torch.is_tensor(t) else t

    def update(self):
        """CompressionModel.update(force=True), common_model.py:63-68"""
        self.coder = _Coder(self.stream_part)
        self.y_tables = eo.gaussian_tables(self.distribution)

    def q(self, scale_name, basic_name, q_in_ckpt, q_index):
        """get_curr_q, common_model.py:35-37 (fine table is float64 -> result stays fp32)."""
        basic = self.sd[basic_name]
        if q_in_ckpt:
            return basic * self.sd[scale_name].reshape(-1)[q_index]
        return basic * fine_q_table(self.sd[scale_name])[q_index]

    def _z_indexes(self, z):
        c = z.shape[1]
        return torch.arange(c, dtype=torch.int).view(1, -1, 1, 1).repeat(1, 1, z.shape[2], z.shape[3])

    def _encode_z(self, z_hat, tables):
        self.coder.encode(z_hat, self._z_indexes(z_hat), tables)

    def _decode_z(self, channel, size, tables):
        idx = self._z_indexes(torch.empty(1, channel, size[0], size[1]))
        return self.coder.decode(idx, tables)

    def _encode_y(self, quad):
        for s in range(4):
            self.coder.encode(quad["y_q_w"][s],
                              eo.build_indexes(quad["scales_w"][s], self.distribution),
                              self.y_tables)

    def _decode_y_fn(self):
        def fn(scales_r):
            return self.coder.decode(eo.build_indexes(scales_r, self.distribution), self.y_tables)
        return fn


class OracleIntra(_Base):
    """IntraNoAR, image_model.py:61-252"""
    distribution = "gaussian"

    def update(self):
        super().update()
        self.z_tables = eo.factorized_tables(self.sd, "bit_estimator_z")

    def get_q(self, q_in_ckpt, q_index):
        return (self.q("q_scale_enc", "q_basic_enc", q_in_ckpt, q_index),
                self.q("q_scale_dec", "q_basic_dec", q_in_ckpt, q_index))

    def _prior_fn(self):
        sd = self.sd

        def fn(step, params):
            t = no.conv(sd, f"y_spatial_prior_adaptor_{step}", params)
            for i in range(3):
                t = no.dcb2(sd, f"y_spatial_prior.{i}", t)
            return t
        return fn

    def _params(self, z_hat, pads):
        params = no.intra_hyper_dec(self.sd, z_hat)
        params = no.dcb2(self.sd, "y_prior_fusion.1", no.dcb2(self.sd, "y_prior_fusion.0", params))
        return eo.slice_to_y(params, pads)

    def _analysis(self, x, q_enc):
        y = no.intra_encoder(self.sd, x, q_enc)
        y_pad, pads = eo.pad_for_y(y)
        z = no.intra_hyper_enc(self.sd, y_pad)
        return y, torch.round(z), pads

    def _synthesis(self, y_hat, q_dec):
        return no.intra_refine(self.sd, no.intra_decoder(self.sd, y_hat, q_dec))

    def forward(self, x, q_in_ckpt=False, q_index=None):
        """image_model.py:114-149 (estimate mode; x_hat not clamped)."""
        q_enc, q_dec = self.get_q(q_in_ckpt, q_index)
        y, z_hat, pads = self._analysis(x, q_enc)
        params = self._params(z_hat, pads)
        quad = eo.quad_forward(y, params, self._prior_fn())
        x_hat = self._synthesis(quad["y_hat"], q_dec)
        bits_y = eo.y_bits(quad["y_q"], quad["scales_hat"], "gaussian")
        bits_z = eo.z_bits(self.sd, "bit_estimator_z", z_hat)
        pixel_num = x.shape[2] * x.shape[3]
        bpp_y = torch.sum(bits_y, dim=(1, 2, 3)) / pixel_num
        bpp_z = torch.sum(bits_z, dim=(1, 2, 3)) / pixel_num
        for k, v in (("y", y), ("z_hat", z_hat), ("params", params), ("y_q", quad["y_q"]),
                     ("y_hat", quad["y_hat"]), ("scales_hat", quad["scales_hat"])):
            self.tap("i_" + k, v)
        return {"x_hat": x_hat, "bit": torch.sum(bpp_y + bpp_z) * pixel_num,
                "bpp": bpp_y + bpp_z, "bpp_y": bpp_y, "bpp_z": bpp_z}

    def compress(self, x, q_in_ckpt, q_index):
        """image_model.py:198-229; stream order z, y parts 0..3."""
        q_enc, q_dec = self.get_q(q_in_ckpt, q_index)
        y, z_hat, pads = self._analysis(x, q_enc)
        params = self._params(z_hat, pads)
        quad = eo.quad_forward(y, params, self._prior_fn())
        self.coder.reset()
        self._encode_z(z_hat, self.z_tables)
        self._encode_y(quad)
        stream = self.coder.flush()
        x_hat = self._synthesis(quad["y_hat"], q_dec).clamp_(0, 1)
        for k, v in (("y", y), ("z_hat", z_hat), ("params", params), ("y_q", quad["y_q"]),
                     ("y_hat", quad["y_hat"])):
            self.tap("i_" + k, v)
        for s in range(4):
            self.tap(f"i_y_q_w_{s}", quad["y_q_w"][s])
            self.tap(f"i_scales_w_{s}", quad["scales_w"][s])
        return {"bit_stream": stream, "x_hat": x_hat}

    def decompress(self, bit_stream, height, width, q_in_ckpt, q_index):
        """image_model.py:231-252"""
        _, q_dec = self.get_q(q_in_ckpt, q_index)
        self.coder.set_stream(bit_stream)
        z_size = downsampled_shape(height, width, 64)
        yh, yw = downsampled_shape(height, width, 16)
        pads = ((yw + 3) // 4 * 4 - yw, (yh + 3) // 4 * 4 - yh)
        z_hat = self._decode_z(self.sd["bit_estimator_z.f1.h"].numel(), z_size, self.z_tables)
        params = self._params(z_hat, pads)
        y_hat = eo.quad_decompress(params, self._prior_fn(), self._decode_y_fn())
        return {"x_hat": self._synthesis(y_hat, q_dec).clamp_(0, 1)}


class OracleDMC(_Base):
    """DMC, video_model.py:235-628"""
    distribution = "laplace"

    def update(self):
        super().update()
        self.z_tables = eo.factorized_tables(self.sd, "bit_estimator_z")
        self.mv_z_tables = eo.factorized_tables(self.sd, "bit_estimator_z_mv")

    def get_q(self, q_in_ckpt, q_index):
        """get_q_for_inference, video_model.py:413-423"""
        return tuple(self.q(f"{n}_q_scale_{d}", f"{n}_q_basic_{d}", q_in_ckpt, q_index)
                     for n, d in (("mv_y", "enc"), ("mv_y", "dec"), ("y", "enc"), ("y", "dec")))

    def _prior_fn(self, prefix):
        sd = self.sd

        def fn(step, params):
            t = no.conv(sd, f"{prefix}_spatial_prior_adaptor_{step}", params)
            for i in range(3):
                t = no.dcb(sd, f"{prefix}_spatial_prior.{i}", t)
            return t
        return fn

    def _mv_params(self, mv_z_hat, dpb, pads):
        """mv_prior_param_decoder, video_model.py:375-385"""
        sd = self.sd
        p = eo.slice_to_y(no.hyper_dec(sd, "mv_hyper_prior_decoder", mv_z_hat), pads)
        if dpb["ref_mv_y"] is None:
            p = no.dcb(sd, "mv_y_prior_fusion_adaptor_0", p)
        else:
            p = no.dcb(sd, "mv_y_prior_fusion_adaptor_1", torch.cat((p, dpb["ref_mv_y"]), 1))
        return no.dcb(sd, "mv_y_prior_fusion.1", no.dcb(sd, "mv_y_prior_fusion.0", p))

    def _res_params(self, z_hat, dpb, context3, pads):
        """res_prior_param_decoder, video_model.py:387-399"""
        sd = self.sd
        hier = eo.slice_to_y(no.hyper_dec(sd, "contextual_hyper_prior_decoder", z_hat), pads)
        t = no.conv(sd, "temporal_prior_encoder.0", context3, 2, 1)
        t = no.conv(sd, "temporal_prior_encoder.2", no.lrelu(t, 0.1), 2, 1)
        if dpb["ref_y"] is None:
            p = no.dcb(sd, "y_prior_fusion_adaptor_0", torch.cat((t, hier), 1))
        else:
            p = no.dcb(sd, "y_prior_fusion_adaptor_1", torch.cat((t, hier, dpb["ref_y"]), 1))
        return no.dcb(sd, "y_prior_fusion.1", no.dcb(sd, "y_prior_fusion.0", p))

    def _motion_compensation(self, dpb, mv, frame_idx):
        """motion_compensation + multi_scale_feature_extractor, video_model.py:343-364"""
        sd = self.sd
        warpframe = no.flow_warp(dpb["ref_frame"], mv)
        mv2 = no.bilinear_down2(mv) / 2
        mv3 = no.bilinear_down2(mv2) / 2
        if dpb["ref_feature"] is None:
            feat = no.conv(sd, "feature_adaptor_I", dpb["ref_frame"], 1, 1)
        else:
            k = [0, 1, 0, 2][frame_idx % 4]
            feat = no.conv(sd, f"feature_adaptor.{k}", dpb["ref_feature"])
        f1, f2, f3 = no.feature_extractor(sd, feat)
        c1_init = no.flow_warp(f1, mv)
        c1 = no.offset_diversity(sd, f1, torch.cat((c1_init, warpframe, mv), 1), mv)
        c2 = no.flow_warp(f2, mv2)
        c3 = no.flow_warp(f3, mv3)
        self.tap("p_ctx1_pre", c1)
        return no.context_fusion(sd, c1, c2, c3)

    def _recon(self, y_hat, c1, c2, c3, q_dec):
        """get_recon_and_feature, video_model.py:401-405 (note the argument order of
        recon_generation_net(recon_image_feature, context1), :403)"""
        r = no.contextual_decoder(self.sd, y_hat, c2, c3, q_dec)
        feature, x_hat = no.recon_generation(self.sd, r, c1)
        return x_hat.clamp_(0, 1), feature

    def _encode_side(self, x, dpb, q_in_ckpt, q_index, frame_idx):
        """shared by compress (:425-453) and forward_one_frame (:559-585)"""
        sd = self.sd
        mv_q_enc, mv_q_dec, y_q_enc, y_q_dec = self.get_q(q_in_ckpt, q_index)
        est_mv = no.spynet(sd, x, dpb["ref_frame"])
        mv_y = no.mv_enc(sd, est_mv, dpb["ref_mv_feature"], mv_q_enc)
        mv_y_pad, pads = eo.pad_for_y(mv_y)
        mv_z_hat = torch.round(no.hyper_enc5(sd, "mv_hyper_prior_encoder", mv_y_pad))
        mv_params = self._mv_params(mv_z_hat, dpb, pads)
        mv_quad = eo.quad_forward(mv_y, mv_params, self._prior_fn("mv_y"))
        mv_hat, mv_feature = no.mv_dec(sd, mv_quad["y_hat"], mv_q_dec)
        c1, c2, c3 = self._motion_compensation(dpb, mv_hat, frame_idx)
        y = no.contextual_encoder(sd, x, c1, c2, c3, y_q_enc)
        y_pad, pads = eo.pad_for_y(y)
        z_hat = torch.round(no.hyper_enc3(sd, "contextual_hyper_prior_encoder", y_pad))
        params = self._res_params(z_hat, dpb, c3, pads)
        quad = eo.quad_forward(y, params, self._prior_fn("y"))
        x_hat, feature = self._recon(quad["y_hat"], c1, c2, c3, y_q_dec)
        for k, v in (("est_mv", est_mv), ("mv_y", mv_y), ("mv_z_hat", mv_z_hat),
                     ("mv_params", mv_params), ("mv_y_q", mv_quad["y_q"]),
                     ("mv_y_hat", mv_quad["y_hat"]), ("mv_hat", mv_hat), ("ctx1", c1),
                     ("ctx2", c2), ("ctx3", c3), ("y", y), ("z_hat", z_hat), ("params", params),
                     ("y_q", quad["y_q"]), ("y_hat", quad["y_hat"]), ("x_hat", x_hat)):
            self.tap("p_" + k, v)
        new_dpb = {"ref_frame": x_hat, "ref_feature": feature, "ref_mv_feature": mv_feature,
                   "ref_y": quad["y_hat"], "ref_mv_y": mv_quad["y_hat"]}
        return new_dpb, mv_z_hat, z_hat, mv_quad, quad

    def compress(self, x, dpb, q_in_ckpt, q_index, frame_idx):
        """video_model.py:425-481; stream order mv_z, z, mv_y parts 0..3, y parts 0..3."""
        new_dpb, mv_z_hat, z_hat, mv_quad, quad = self._encode_side(x, dpb, q_in_ckpt, q_index,
                                                                    frame_idx)
        self.coder.reset()
        self._encode_z(mv_z_hat, self.mv_z_tables)
        self._encode_z(z_hat, self.z_tables)
        self._encode_y(mv_quad)
        self._encode_y(quad)
        for s in range(4):
            self.tap(f"p_mv_y_q_w_{s}", mv_quad["y_q_w"][s])
            self.tap(f"p_mv_scales_w_{s}", mv_quad["scales_w"][s])
            self.tap(f"p_y_q_w_{s}", quad["y_q_w"][s])
            self.tap(f"p_scales_w_{s}", quad["scales_w"][s])
        return {"dpb": new_dpb, "bit_stream": self.coder.flush()}

    def decompress(self, dpb, string, height, width, q_in_ckpt, q_index, frame_idx):
        """video_model.py:483-520"""
        sd = self.sd
        _, mv_q_dec, _, y_q_dec = self.get_q(q_in_ckpt, q_index)
        self.coder.set_stream(string)
        z_size = downsampled_shape(height, width, 64)
        yh, yw = downsampled_shape(height, width, 16)
        pads = ((yw + 3) // 4 * 4 - yw, (yh + 3) // 4 * 4 - yh)
        mv_z_hat = self._decode_z(sd["bit_estimator_z_mv.f1.h"].numel(), z_size, self.mv_z_tables)
        z_hat = self._decode_z(sd["bit_estimator_z.f1.h"].numel(), z_size, self.z_tables)
        mv_params = self._mv_params(mv_z_hat, dpb, pads)
        mv_y_hat = eo.quad_decompress(mv_params, self._prior_fn("mv_y"), self._decode_y_fn())
        mv_hat, mv_feature = no.mv_dec(sd, mv_y_hat, mv_q_dec)
        c1, c2, c3 = self._motion_compensation(dpb, mv_hat, frame_idx)
        params = self._res_params(z_hat, dpb, c3, pads)
        y_hat = eo.quad_decompress(params, self._prior_fn("y"), self._decode_y_fn())
        x_hat, feature = self._recon(y_hat, c1, c2, c3, y_q_dec)
        return {"dpb": {"ref_frame": x_hat, "ref_feature": feature,
                        "ref_mv_feature": mv_feature, "ref_y": y_hat, "ref_mv_y": mv_y_hat}}

    def forward_one_frame(self, x, dpb, q_in_ckpt=False, q_index=None, frame_idx=0):
        """video_model.py:559-628 (estimate mode)"""
        new_dpb, mv_z_hat, z_hat, mv_quad, quad = self._encode_side(x, dpb, q_in_ckpt, q_index,
                                                                    frame_idx)
        pixel_num = x.shape[2] * x.shape[3]
        b_y = torch.sum(eo.y_bits(quad["y_q"], quad["scales_hat"], "laplace"), dim=(1, 2, 3))
        b_mv = torch.sum(eo.y_bits(mv_quad["y_q"], mv_quad["scales_hat"], "laplace"), dim=(1, 2, 3))
        b_z = torch.sum(eo.z_bits(self.sd, "bit_estimator_z", z_hat), dim=(1, 2, 3))
        b_mvz = torch.sum(eo.z_bits(self.sd, "bit_estimator_z_mv", mv_z_hat), dim=(1, 2, 3))
        bpp_y, bpp_z = b_y / pixel_num, b_z / pixel_num
        bpp_mv_y, bpp_mv_z = b_mv / pixel_num, b_mvz / pixel_num
        bpp = bpp_y + bpp_z + bpp_mv_y + bpp_mv_z
        return {"bpp_mv_y": bpp_mv_y, "bpp_mv_z": bpp_mv_z, "bpp_y": bpp_y, "bpp_z": bpp_z,
                "bpp": bpp, "dpb": new_dpb, "bit": torch.sum(bpp) * pixel_num,
                "bit_y": torch.sum(bpp_y) * pixel_num, "bit_z": torch.sum(bpp_z) * pixel_num,
                "bit_mv_y": torch.sum(bpp_mv_y) * pixel_num,
                "bit_mv_z": torch.sum(bpp_mv_z) * pixel_num}
