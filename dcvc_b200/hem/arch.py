"""Parameter inventory of the two DCVC-HEM inference models (the ``state_dict`` contract):
`IntraNoAR(N=192)` (reference DCVC-HEM/src/models/image_model.py:15-56) and `DMC`
(DCVC-HEM/src/models/video_model.py:131-231), built from video_net.py:63-277 and
src/layers/layers.py:19-128.  Names and shapes follow the reference's module tree so that the
released acmmm2022_*.pth.tar checkpoints load unchanged (tests/test_hem.py holds the inventory to
the reference's own state_dict when the reference is present)."""
from ..arch import Spec

CH_MV, CH_N, CH_M = 64, 64, 96


class HemSpec(Spec):
    def residual_block(self, p, cin, cout):                  # layers.py:103-128
        self.conv(p + ".conv1", cin, cout, 3)
        self.conv(p + ".conv2", cout, cout, 3)

    def conv_block_residual(self, p, cin, cout):             # video_net.py:166-181 (+ SELayer :149-163)
        self.conv(p + ".conv.0", cin, cout, 3)
        self.conv(p + ".conv.2", cout, cout, 3)
        self[p + ".conv.3.fc.0.weight"] = (cout // 16, cout)
        self[p + ".conv.3.fc.2.weight"] = (cout, cout // 16)
        self.conv(p + ".up_dim", cin, cout, 1)

    def hem_unet(self, p, cin, cout):                        # video_net.py:184-223
        self.conv_block_residual(p + ".conv1", cin, 32)
        self.conv_block_residual(p + ".conv2", 32, 64)
        self.conv_block_residual(p + ".conv3", 64, 128)
        for i in range(4):
            self.res_block(f"{p}.context_refine.{i}", 128)
        self.conv(p + ".up3.0", 128, 64 * 4, 1)
        self.conv_block_residual(p + ".up_conv3", 128, 64)
        self.conv(p + ".up2.0", 64, 32 * 4, 1)
        self.conv_block_residual(p + ".up_conv2", 64, cout)

    def enc_dec(self, penc, pdec, cin, cout, ch):            # video_net.py:226-249
        self.rbws(penc + ".0", cin, ch)
        self.residual_block(penc + ".1", ch, ch)
        self.rbws(penc + ".2", ch, ch)
        self.residual_block(penc + ".3", ch, ch)
        self.rbws(penc + ".4", ch, ch)
        self.residual_block(penc + ".5", ch, ch)
        self.conv(penc + ".6", ch, ch, 3)
        self.residual_block(pdec + ".0", ch, ch)
        self.rbu(pdec + ".1", ch, ch)
        self.residual_block(pdec + ".2", ch, ch)
        self.rbu(pdec + ".3", ch, ch)
        self.residual_block(pdec + ".4", ch, ch)
        self.rbu(pdec + ".5", ch, ch)
        self.residual_block(pdec + ".6", ch, ch)
        self.conv(pdec + ".7.0", ch, cout * 4, 1)

    def hem_hyper_pair(self, penc, pdec, y_ch, z_ch):        # video_net.py:252-277
        self.conv(penc + ".0", y_ch, z_ch, 3)
        for i in (2, 4, 6, 8):
            self.conv(f"{penc}.{i}", z_ch, z_ch, 3)
        self.hem_hyper_dec(pdec, y_ch, z_ch)

    def hem_hyper_dec(self, pdec, y_ch, z_ch):
        self.conv(pdec + ".0", z_ch, y_ch, 3)
        self.conv(pdec + ".2.0", y_ch, y_ch * 4, 1)
        self.conv(pdec + ".4", y_ch, y_ch * 3 // 2, 3)
        self.conv(pdec + ".6.0", y_ch * 3 // 2, y_ch * 3 // 2 * 4, 1)
        self.conv(pdec + ".8", y_ch * 3 // 2, y_ch * 2, 3)

    def three_convs(self, p, c0, c1, c2, c3):                # prior fusion / spatial prior stacks
        self.conv(p + ".0", c0, c1, 3)
        self.conv(p + ".2", c1, c2, 3)
        self.conv(p + ".4", c2, c3, 3)


def hem_intra_spec(N=192, anchor_num=4):
    s = HemSpec()
    s.bit_estimator("bit_estimator_z", N)
    s.enc_dec("enc", "dec", 3, 16, N)
    s.hem_unet("refine.0", 16, 16)
    s.conv("refine.1", 16, 3, 3)
    s.hem_hyper_pair("hyper_enc", "hyper_dec", N, N)
    s.three_convs("y_prior_fusion", N * 2, N * 3, N * 3, N * 3)
    s.three_convs("y_spatial_prior", N * 4, N * 3, N * 3, N * 2)
    s["q_basic"] = (1, N, 1, 1)
    s["q_scale"] = (anchor_num, 1, 1, 1)
    return s


def hem_dmc_spec(anchor_num=4):
    s = HemSpec()
    s.bit_estimator("bit_estimator_z", CH_N)
    s.bit_estimator("bit_estimator_z_mv", CH_N)
    for lv in range(4):
        p = f"optic_flow.moduleBasic.{lv}"
        for i, (ci, co) in enumerate(((8, 32), (32, 64), (64, 32), (32, 16), (16, 2))):
            s.conv(f"{p}.conv{i + 1}", ci, co, 7)
    s.enc_dec("mv_encoder", "mv_decoder", 2, 2, CH_MV)
    s.hem_hyper_pair("mv_hyper_prior_encoder", "mv_hyper_prior_decoder", CH_MV, CH_N)
    s.three_convs("mv_y_prior_fusion", CH_MV * 3, CH_MV * 3, CH_MV * 3, CH_MV * 3)
    s.three_convs("mv_y_spatial_prior", CH_MV * 4, CH_MV * 3, CH_MV * 3, CH_MV * 2)
    s.conv("feature_adaptor_I", 3, CH_N, 3)
    s.conv("feature_adaptor_P", CH_N, CH_N, 1)
    fe = "feature_extractor"                                  # video_model.py:17-38
    for i in (1, 2, 3):
        s.conv(f"{fe}.conv{i}", CH_N, CH_N, 3)
        s.res_block(f"{fe}.res_block{i}", CH_N)
    cf = "context_fusion_net"                                 # video_model.py:41-68
    s.conv(cf + ".conv3_up.0", CH_N, CH_N * 4, 3)
    s.res_block(cf + ".res_block3_up", CH_N)
    s.conv(cf + ".conv3_out", CH_N, CH_N, 3)
    s.res_block(cf + ".res_block3_out", CH_N)
    s.conv(cf + ".conv2_up.0", CH_N * 2, CH_N * 4, 3)
    s.res_block(cf + ".res_block2_up", CH_N)
    s.conv(cf + ".conv2_out", CH_N * 2, CH_N, 3)
    s.res_block(cf + ".res_block2_out", CH_N)
    s.conv(cf + ".conv1_out", CH_N * 2, CH_N, 3)
    s.res_block(cf + ".res_block1_out", CH_N)
    ce = "contextual_encoder"                                 # video_model.py:71-91
    s.conv(ce + ".conv1", CH_N + 3, CH_N, 3)
    s.res_block(ce + ".res1", CH_N * 2, bottleneck=True)
    s.conv(ce + ".conv2", CH_N * 2, CH_N, 3)
    s.res_block(ce + ".res2", CH_N * 2, bottleneck=True)
    s.conv(ce + ".conv3", CH_N * 2, CH_N, 3)
    s.conv(ce + ".conv4", CH_N, CH_M, 3)
    s.three_convs("contextual_hyper_prior_encoder", CH_M, CH_N, CH_N, CH_N)
    s.hem_hyper_dec("contextual_hyper_prior_decoder", CH_M, CH_N)
    s.conv("temporal_prior_encoder.0", CH_N, CH_M * 3 // 2, 3)
    s.conv("temporal_prior_encoder.2", CH_M * 3 // 2, CH_M * 2, 3)
    s.three_convs("y_prior_fusion", CH_M * 5, CH_M * 4, CH_M * 3, CH_M * 3)
    s.three_convs("y_spatial_prior", CH_M * 4, CH_M * 3, CH_M * 3, CH_M * 2)
    cd = "contextual_decoder"                                 # video_model.py:94-114
    s.conv(cd + ".up1.0", CH_M, CH_N * 4, 3)
    s.conv(cd + ".up2.0", CH_N, CH_N * 4, 3)
    s.res_block(cd + ".res1", CH_N * 2, bottleneck=True)
    s.conv(cd + ".up3.0", CH_N * 2, CH_N * 4, 3)
    s.res_block(cd + ".res2", CH_N * 2, bottleneck=True)
    s.conv(cd + ".up4.0", CH_N * 2, 32 * 4, 3)
    rg = "recon_generation_net"                               # video_model.py:117-128
    s.conv(rg + ".first_conv", CH_N + 32, CH_N, 3)
    s.hem_unet(rg + ".unet_1", CH_N, CH_N)
    s.hem_unet(rg + ".unet_2", CH_N, CH_N)
    s.conv(rg + ".recon_conv", CH_N, 3, 3)
    s["mv_y_q_basic"] = (1, CH_MV, 1, 1)
    s["y_q_basic"] = (1, CH_M, 1, 1)
    s["mv_y_q_scale"] = (anchor_num, 1, 1, 1)
    s["y_q_scale"] = (anchor_num, 1, 1, 1)
    return s


def hem_synth_state_dict(spec, seed, gain=1.0):
    """Deterministic random-init weights (SURVEY 8(d) "Weights": HEM's own init produces large
    latents, so tests also run gain=0.5).  Conv weights U(-b, b), b = gain / sqrt(fan_in); SE
    linear layers the same with fan_in = in_features; q_basic = 1 + 0.1 U(-1, 1) (some below the
    0.5 LowerBound are NOT produced; the bound is exercised by tests/test_hem.py directly);
    q_scale anchors distinct and decreasing."""
    from collections import OrderedDict
    import torch
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    sd = OrderedDict()
    for name, shape in spec.items():
        if name.endswith(".weight"):
            fan_in = shape[1] * (shape[2] * shape[3] if len(shape) == 4 else 1)
            sd[name] = (torch.rand(shape, generator=g) * 2 - 1) * (gain / fan_in ** 0.5)
        elif name.endswith(".bias"):
            w = spec[name[:-5] + ".weight"]
            sd[name] = (torch.rand(shape, generator=g) * 2 - 1) * (gain / (w[1] * w[2] * w[3]) ** 0.5)
        elif "bit_estimator" in name:
            sd[name] = torch.randn(shape, generator=g) * 0.01
        elif "q_scale" in name:
            sd[name] = torch.tensor((1.54, 1.08, 0.73, 0.5), dtype=torch.float32).reshape(shape)
        elif "q_basic" in name:
            sd[name] = 1.0 + 0.1 * (torch.rand(shape, generator=g) * 2 - 1)
        else:
            raise KeyError(name)
    return sd
