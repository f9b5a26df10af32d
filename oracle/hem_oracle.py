"""TEST INFRASTRUCTURE -- CPU oracle of the DCVC-HEM per-frame coding path: `OracleHemIntra` /
`OracleHemDMC` restate IntraNoAR / DMC of DCVC-HEM (reference DCVC-HEM/src/models/image_model.py:15-171,
video_model.py:131-519, common_model.py:84-188, video_net.py:63-277, src/layers/layers.py:19-128)
in plain torch fp32 on top of oracle/nets_oracle.py (shared conv blocks), oracle/entropy_oracle.py
(tables, indexes, bit estimates) and oracle/rans_oracle.c (the coder: HEM's
rans_interface.cpp:85-244 is the same rANS arithmetic as DCVC-DC's single-part stream without the
part-header byte, so the HEM stream is the DC port's parts=1 stream minus its first byte).

Pinned by tests/test_hem_oracle.py against tests/golden/hem_*.npz, which
tests/golden/make_golden_hem.py produced by running the UNMODIFIED reference in the build container.
The product never imports this module.
"""
import numpy as np
import torch

from . import entropy_oracle as eo
from . import nets_oracle as no
from . import rans_oracle


def get_rounded_q(q_scale):
    """utils/stream_helper.py:41-45"""
    q_scale = np.clip(float(q_scale), 0.01, 655.)
    q_index = int(np.round(q_scale * 100))
    return q_index / 100, q_index


def downsampled_shape(height, width, p):
    nh = (height + p - 1) // p * p
    nw = (width + p - 1) // p * p
    return int(nh / p + 0.5), int(nw / p + 0.5)


# ------------------------------------------------------------------------------ nets
def residual_block(sd, p, x):
    """ResidualBlock, layers.py:103-128"""
    out = no.lrelu(no.conv(sd, p + ".conv1", x, 1, 1), 0.01)
    out = no.lrelu(no.conv(sd, p + ".conv2", out, 1, 1), 0.01)
    return out + x


def se_layer(sd, p, x):
    """SELayer, video_net.py:149-163"""
    y = torch.mean(x, dim=(-1, -2))
    y = torch.relu(y @ sd[p + ".fc.0.weight"].t())
    y = torch.sigmoid(y @ sd[p + ".fc.2.weight"].t())
    return x * y[:, :, None, None]


def conv_block_residual(sd, p, x):
    """ConvBlockResidual, video_net.py:166-181"""
    x1 = no.conv(sd, p + ".conv.2", no.lrelu(no.conv(sd, p + ".conv.0", x, 1, 1), 0.01), 1, 1)
    return no.conv(sd, p + ".up_dim", x) + se_layer(sd, p + ".conv.3", x1)


def unet(sd, p, x):
    """UNet, video_net.py:184-223"""
    x1 = conv_block_residual(sd, p + ".conv1", x)
    x2 = conv_block_residual(sd, p + ".conv2", torch.max_pool2d(x1, 2, 2))
    x3 = conv_block_residual(sd, p + ".conv3", torch.max_pool2d(x2, 2, 2))
    for i in range(4):
        x3 = no.res_block(sd, f"{p}.context_refine.{i}", x3, 0.0)
    d3 = conv_block_residual(sd, p + ".up_conv3", torch.cat((x2, no.subpel(sd, p + ".up3", x3, 1)), 1))
    return conv_block_residual(sd, p + ".up_conv2", torch.cat((x1, no.subpel(sd, p + ".up2", d3, 1)), 1))


def analysis(sd, p, x):
    """get_enc_dec_models encoder, video_net.py:226-235"""
    for i in (0, 2, 4):
        x = residual_block(sd, f"{p}.{i + 1}", no.rbws(sd, f"{p}.{i}", x))
    return no.conv(sd, p + ".6", x, 2, 1)


def synthesis(sd, p, x):
    """get_enc_dec_models decoder, video_net.py:237-247"""
    for i in (0, 2, 4):
        x = no.rbu(sd, f"{p}.{i + 1}", residual_block(sd, f"{p}.{i}", x))
    return no.subpel(sd, p + ".7", residual_block(sd, p + ".6", x), 1)


def three_convs(sd, p, x):
    x = no.lrelu(no.conv(sd, p + ".0", x, 1, 1), 0.2)
    x = no.lrelu(no.conv(sd, p + ".2", x, 1, 1), 0.2)
    return no.conv(sd, p + ".4", x, 1, 1)


# ------------------------------------------------------------------------------ dual prior
def get_mask(h, w):
    """common_model.py:84-91"""
    micro = torch.tensor(((1, 0), (0, 1)), dtype=torch.float32)
    m0 = micro.repeat(h // 2, w // 2)[None, None]
    return m0, torch.ones_like(m0) - m0


def _pwm(y, scales, means, mask):
    """process_with_mask, common_model.py:93-100"""
    scales_hat, means_hat = scales * mask, means * mask
    y_q = torch.round((y - means_hat) * mask)
    return y_q, y_q + means_hat, scales_hat


def dual_forward(y, means, scales, quant_step, prior_fn):
    """forward_dual_prior, common_model.py:102-156; returns every tensor either branch returns"""
    m0, m1 = get_mask(y.shape[2], y.shape[3])
    quant_step = torch.clamp_min(quant_step, 0.5)                 # LowerBound.apply(q, 0.5)
    y = y / quant_step
    y_0, y_1 = y.chunk(2, 1)
    s0, s1 = scales.chunk(2, 1)
    mu0, mu1 = means.chunk(2, 1)
    q00, h00, sh00 = _pwm(y_0, s0, mu0, m0)
    q11, h11, sh11 = _pwm(y_1, s1, mu1, m1)
    params = torch.cat((h00, h11, means, scales, quant_step), 1)
    s0, mu0, s1, mu1 = prior_fn(params).chunk(4, 1)
    q01, h01, sh01 = _pwm(y_0, s0, mu0, m1)
    q10, h10, sh10 = _pwm(y_1, s1, mu1, m0)
    y_q = torch.cat((q00 + q01, q11 + q10), 1)
    y_hat = torch.cat((h00 + h01, h11 + h10), 1) * quant_step
    scales_hat = torch.cat((sh00 + sh01, sh11 + sh10), 1)
    return {"y_q": y_q, "y_hat": y_hat, "scales_hat": scales_hat,
            "y_q_w": (q00 + q11, q01 + q10), "scales_w": (sh00 + sh11, sh01 + sh10)}


def dual_decompress(means, scales, quant_step, prior_fn, decode_fn):
    """decompress_dual_prior, common_model.py:161-188"""
    m0, m1 = get_mask(means.shape[2], means.shape[3])
    quant_step = torch.clamp_min(quant_step, 0.5)
    s0, s1 = scales.chunk(2, 1)
    mu0, mu1 = means.chunk(2, 1)
    r0 = decode_fn(s0 * m0 + s1 * m1)
    h00, h11 = (r0 + mu0) * m0, (r0 + mu1) * m1
    params = torch.cat((h00, h11, means, scales, quant_step), 1)
    s0, mu0, s1, mu1 = prior_fn(params).chunk(4, 1)
    r1 = decode_fn(s0 * m1 + s1 * m0)
    h01, h10 = (r1 + mu0) * m1, (r1 + mu1) * m0
    return torch.cat((h00 + h01, h11 + h10), 1) * quant_step


def y_bits(y_q, sigma, distribution):
    """get_y_gaussian_bits / get_y_laplace_bits, common_model.py:58-70 (HEM floors the Gaussian
    sigma at 0.11, the Laplace one at 1e-5)"""
    sigma = sigma.clamp(0.11 if distribution == "gaussian" else 1e-5, 1e10)
    d = eo._cdf_fn(distribution)(torch.zeros_like(sigma), sigma)
    return eo.probs_to_bits(d.cdf(y_q + 0.5) - d.cdf(y_q - 0.5))


# ------------------------------------------------------------------------------ coder
class _Coder:
    """EntropyCoder (src/entropy_models/entropy_models.py:9-50) on the oracle's C coder: one
    part, the part-header byte dropped / re-added (see the module docstring)."""

    def __init__(self):
        self.enc = rans_oracle.RansEncoder(False, 1)
        self.dec = rans_oracle.RansDecoder(1)

    def reset(self):
        self.enc.reset()

    def encode(self, symbols, indexes, tables):
        s = symbols.reshape(-1).clamp(-30000, 30000).to(torch.int16).numpy()
        self.enc.encode_with_indexes(s, indexes.reshape(-1).to(torch.int16).numpy(), *tables)

    def flush(self):
        self.enc.flush()
        full = self.enc.get_encoded_stream().tobytes()
        assert full[0] == 1          # one part, 2-byte size field flag: constant
        return full[1:]

    def set_stream(self, stream):
        self.dec.set_stream(b"\x01" + bytes(stream))

    def decode(self, indexes, tables):
        out = self.dec.decode_stream(indexes.reshape(-1).to(torch.int16).numpy(), *tables)
        return torch.from_numpy(out.astype(np.float32)).reshape(indexes.shape)


class _Base:
    distribution = "laplace"

    def __init__(self, state_dict):
        self.sd = {k: v.detach().float().cpu() for k, v in state_dict.items()}
        self.coder = None
        self.taps = None

    def tap(self, name, t):
        if self.taps is not None:
            self.taps[name] = t.detach().clone() if torch.is_tensor(t) else t

    def update(self):
        self.coder = _Coder()
        self.y_tables = eo.gaussian_tables(self.distribution)

    def curr_q(self, basic, q_scale):
        return torch.clamp_min(self.sd[basic], 0.5) * float(q_scale)

    def _z_indexes(self, c, h, w):
        return torch.arange(c, dtype=torch.int).view(1, -1, 1, 1).repeat(1, 1, h, w)

    def _encode_z(self, z_hat, tables):
        self.coder.encode(z_hat, self._z_indexes(*z_hat.shape[1:]), tables)

    def _decode_z(self, c, size, tables):
        return self.coder.decode(self._z_indexes(c, size[0], size[1]), tables)

    def _encode_y(self, d):
        for s in range(2):
            self.coder.encode(d["y_q_w"][s], eo.build_indexes(d["scales_w"][s], self.distribution),
                              self.y_tables)

    def _decode_fn(self):
        return lambda scales_r: self.coder.decode(eo.build_indexes(scales_r, self.distribution),
                                                  self.y_tables)


class OracleHemIntra(_Base):
    """IntraNoAR(N=192), image_model.py:15-171"""
    distribution = "gaussian"

    def update(self):
        super().update()
        self.z_tables = eo.factorized_tables(self.sd, "bit_estimator_z")

    def _fused(self, z_hat):
        return three_convs(self.sd, "y_prior_fusion", no.hyper_dec(self.sd, "hyper_dec", z_hat)).chunk(3, 1)

    def _prior_fn(self):
        return lambda p: three_convs(self.sd, "y_spatial_prior", p)

    def _synthesis(self, y_hat):
        t = unet(self.sd, "refine.0", synthesis(self.sd, "dec", y_hat))
        return no.conv(self.sd, "refine.1", t, 1, 1)

    def _encode_side(self, x, q_scale):
        cq = self.curr_q("q_basic", q_scale)
        y = analysis(self.sd, "enc", x) / cq
        z_hat = torch.round(no.hyper_enc5(self.sd, "hyper_enc", y))
        q_step, scales, means = self._fused(z_hat)
        d = dual_forward(y, means, scales, q_step, self._prior_fn())
        self.tap("i_y", y)
        self.tap("i_z_hat", z_hat)
        for s in range(2):
            self.tap(f"i_y_q_w_{s}", d["y_q_w"][s])
            self.tap(f"i_scales_w_{s}", d["scales_w"][s])
        return d, z_hat, cq

    def forward(self, x, q_scale):
        """image_model.py:65-108 (estimate mode)"""
        d, z_hat, cq = self._encode_side(x, q_scale)
        x_hat = self._synthesis(d["y_hat"] * cq)
        n = x.shape[2] * x.shape[3]
        bpp_y = torch.sum(y_bits(d["y_q"], d["scales_hat"], "gaussian"), dim=(1, 2, 3)) / n
        bpp_z = torch.sum(eo.z_bits(self.sd, "bit_estimator_z", z_hat), dim=(1, 2, 3)) / n
        return {"x_hat": x_hat, "bit": float(torch.sum(bpp_y + bpp_z) * n), "bpp_y": bpp_y, "bpp_z": bpp_z}

    def compress(self, x, q_scale):
        """image_model.py:139-160"""
        d, z_hat, _ = self._encode_side(x, q_scale)
        self.coder.reset()
        self._encode_z(z_hat, self.z_tables)
        self._encode_y(d)
        return {"bit_stream": self.coder.flush()}

    def decompress(self, bit_stream, height, width, q_scale):
        """image_model.py:162-171"""
        cq = self.curr_q("q_basic", q_scale)
        self.coder.set_stream(bit_stream)
        z_hat = self._decode_z(self.sd["q_basic"].numel(), downsampled_shape(height, width, 64),
                               self.z_tables)
        q_step, scales, means = self._fused(z_hat)
        y_hat = dual_decompress(means, scales, q_step, self._prior_fn(), self._decode_fn()) * cq
        return {"x_hat": self._synthesis(y_hat).clamp_(0, 1)}


class OracleHemDMC(_Base):
    """DMC, video_model.py:131-519"""
    distribution = "laplace"

    def update(self):
        super().update()
        self.z_tables = eo.factorized_tables(self.sd, "bit_estimator_z")
        self.mv_z_tables = eo.factorized_tables(self.sd, "bit_estimator_z_mv")

    def _mv_fused(self, mv_z_hat, dpb):
        p = no.hyper_dec(self.sd, "mv_hyper_prior_decoder", mv_z_hat)
        ref = dpb["ref_mv_y"]
        if ref is None:
            ref = torch.zeros(1, p.shape[1] // 2, p.shape[2], p.shape[3])
        return three_convs(self.sd, "mv_y_prior_fusion", torch.cat((p, ref), 1)).chunk(3, 1)

    def _y_fused(self, z_hat, dpb, c3):
        sd = self.sd
        hier = no.hyper_dec(sd, "contextual_hyper_prior_decoder", z_hat)
        t = no.conv(sd, "temporal_prior_encoder.0", c3, 2, 1)
        t = no.conv(sd, "temporal_prior_encoder.2", no.lrelu(t, 0.1), 2, 1)
        ref = dpb["ref_y"]
        if ref is None:
            ref = torch.zeros(1, t.shape[1] // 2, t.shape[2], t.shape[3])
        return three_convs(sd, "y_prior_fusion", torch.cat((t, hier, ref), 1)).chunk(3, 1)

    def _motion_compensation(self, dpb, mv):
        """video_model.py:227-246"""
        sd = self.sd
        mv2 = no.bilinear_down2(mv) / 2
        mv3 = no.bilinear_down2(mv2) / 2
        if dpb["ref_feature"] is None:
            feat = no.conv(sd, "feature_adaptor_I", dpb["ref_frame"], 1, 1)
        else:
            feat = no.conv(sd, "feature_adaptor_P", dpb["ref_feature"])
        f1, f2, f3 = no.feature_extractor(sd, feat)
        return no.context_fusion(sd, no.flow_warp(f1, mv), no.flow_warp(f2, mv2), no.flow_warp(f3, mv3))

    def _contextual_encoder(self, x, c1, c2, c3):
        sd, p = self.sd, "contextual_encoder"
        f = no.conv(sd, p + ".conv1", torch.cat((x, c1), 1), 2, 1)
        f = no.res_block(sd, p + ".res1", torch.cat((f, c2), 1), 0.1, True)
        f = no.conv(sd, p + ".conv2", f, 2, 1)
        f = no.res_block(sd, p + ".res2", torch.cat((f, c3), 1), 0.1, True)
        return no.conv(sd, p + ".conv4", no.conv(sd, p + ".conv3", f, 2, 1), 2, 1)

    def _recon(self, y_hat, c1, c2, c3):
        sd, p = self.sd, "contextual_decoder"
        f = no.subpel(sd, p + ".up2", no.subpel(sd, p + ".up1", y_hat, 3), 3)
        f = no.res_block(sd, p + ".res1", torch.cat((f, c3), 1), 0.1, True)
        f = no.subpel(sd, p + ".up3", f, 3)
        f = no.res_block(sd, p + ".res2", torch.cat((f, c2), 1), 0.1, True)
        res = no.subpel(sd, p + ".up4", f, 3)
        rg = "recon_generation_net"
        # recon_generation_net(recon_image_feature, context1) (:301): the 32-channel feature
        # comes FIRST in first_conv's input, whatever the parameter names of :124 suggest
        f = no.conv(sd, rg + ".first_conv", torch.cat((res, c1), 1), 1, 1)
        f = unet(sd, rg + ".unet_2", unet(sd, rg + ".unet_1", f))
        return no.conv(sd, rg + ".recon_conv", f, 1, 1), f

    def _encode_side(self, x, dpb, mv_y_q_scale, y_q_scale):
        sd = self.sd
        mv_q, y_q = self.curr_q("mv_y_q_basic", mv_y_q_scale), self.curr_q("y_q_basic", y_q_scale)
        est_mv = no.spynet(sd, x, dpb["ref_frame"])
        mv_y = analysis(sd, "mv_encoder", est_mv) / mv_q
        mv_z_hat = torch.round(no.hyper_enc5(sd, "mv_hyper_prior_encoder", mv_y))
        mq, ms, mm = self._mv_fused(mv_z_hat, dpb)
        mvd = dual_forward(mv_y, mm, ms, mq, lambda p: three_convs(sd, "mv_y_spatial_prior", p))
        mv_y_hat = mvd["y_hat"] * mv_q
        mv_hat = synthesis(sd, "mv_decoder", mv_y_hat)
        c1, c2, c3 = self._motion_compensation(dpb, mv_hat)
        y = self._contextual_encoder(x, c1, c2, c3) / y_q
        z_hat = torch.round(no.hyper_enc3(sd, "contextual_hyper_prior_encoder", y))
        q, s, m = self._y_fused(z_hat, dpb, c3)
        d = dual_forward(y, m, s, q, lambda p: three_convs(sd, "y_spatial_prior", p))
        y_hat = d["y_hat"] * y_q
        x_hat, feature = self._recon(y_hat, c1, c2, c3)
        for k, v in (("est_mv", est_mv), ("mv_y", mv_y), ("mv_z_hat", mv_z_hat),
                     ("mv_y_hat", mv_y_hat), ("mv_hat", mv_hat), ("ctx1", c1), ("ctx2", c2),
                     ("ctx3", c3), ("y", y), ("z_hat", z_hat), ("y_hat", y_hat), ("x_hat", x_hat)):
            self.tap("p_" + k, v)
        for st in range(2):
            self.tap(f"p_mv_y_q_w_{st}", mvd["y_q_w"][st])
            self.tap(f"p_mv_scales_w_{st}", mvd["scales_w"][st])
            self.tap(f"p_y_q_w_{st}", d["y_q_w"][st])
            self.tap(f"p_scales_w_{st}", d["scales_w"][st])
        dpb2 = {"ref_frame": x_hat, "ref_feature": feature, "ref_y": y_hat, "ref_mv_y": mv_y_hat}
        return dpb2, mv_z_hat, z_hat, mvd, d

    def compress(self, x, dpb, mv_y_q_scale, y_q_scale):
        """video_model.py:263-324: order mv_z, mv_y 0-1, z, y 0-1; frame not clamped; key "dbp"."""
        dpb2, mv_z_hat, z_hat, mvd, d = self._encode_side(x, dpb, mv_y_q_scale, y_q_scale)
        self.coder.reset()
        self._encode_z(mv_z_hat, self.mv_z_tables)
        self._encode_y(mvd)
        self._encode_z(z_hat, self.z_tables)
        self._encode_y(d)
        return {"dbp": dpb2, "bit_stream": self.coder.flush()}

    def decompress(self, dpb, string, height, width, mv_y_q_scale, y_q_scale):
        """video_model.py:326-375"""
        sd = self.sd
        mv_q, y_q = self.curr_q("mv_y_q_basic", mv_y_q_scale), self.curr_q("y_q_basic", y_q_scale)
        self.coder.set_stream(string)
        zs = downsampled_shape(height, width, 64)
        mv_z_hat = self._decode_z(64, zs, self.mv_z_tables)
        mq, ms, mm = self._mv_fused(mv_z_hat, dpb)
        mv_y_hat = dual_decompress(mm, ms, mq, lambda p: three_convs(sd, "mv_y_spatial_prior", p),
                                   self._decode_fn()) * mv_q
        mv_hat = synthesis(sd, "mv_decoder", mv_y_hat)
        c1, c2, c3 = self._motion_compensation(dpb, mv_hat)
        z_hat = self._decode_z(64, zs, self.z_tables)
        q, s, m = self._y_fused(z_hat, dpb, c3)
        y_hat = dual_decompress(m, s, q, lambda p: three_convs(sd, "y_spatial_prior", p),
                                self._decode_fn()) * y_q
        x_hat, feature = self._recon(y_hat, c1, c2, c3)
        return {"dpb": {"ref_frame": x_hat.clamp(0, 1), "ref_feature": feature, "ref_y": y_hat,
                        "ref_mv_y": mv_y_hat}}

    def forward_one_frame(self, x, dpb, mv_y_q_scale, y_q_scale):
        """video_model.py:417-513 (estimate mode): bits only"""
        dpb2, mv_z_hat, z_hat, mvd, d = self._encode_side(x, dpb, mv_y_q_scale, y_q_scale)
        bits = {"bit_y": y_bits(d["y_q"], d["scales_hat"], "laplace").sum(),
                "bit_mv_y": y_bits(mvd["y_q"], mvd["scales_hat"], "laplace").sum(),
                "bit_z": eo.z_bits(self.sd, "bit_estimator_z", z_hat).sum(),
                "bit_mv_z": eo.z_bits(self.sd, "bit_estimator_z_mv", mv_z_hat).sum()}
        bits["bit"] = sum(bits.values())
        bits["dpb"] = dpb2
        return bits
