"""Parameter inventory of the two DCVC-DC models (the ``state_dict`` contract).

The released checkpoints (cvpr2023_image_*.pth.tar / cvpr2023_video_*.pth.tar) must load
unchanged, so names and shapes follow the reference's module tree exactly
(reference: DCVC-DC/src/models/video_model.py:235-323, image_model.py:61-105,
video_net.py:58-251, layers.py:18-222, entropy_models.py:58-109).  The inventory is
declarative: a ``Spec`` is an ordered dict name -> shape, built by small block helpers; the
host models (video_model.py / image_model.py) walk the same helpers when they compile a frame
plan, so a block's parameters and its kernel sequence are defined side by side.

tests/test_arch.py checks the inventory against the reference's own ``state_dict()`` when
/root/reference is present and against a committed name/shape manifest otherwise.
"""
from collections import OrderedDict

CH_1X, CH_2X, CH_4X, CH_8X, CH_16X = 48, 64, 96, 96, 128   # video_model.py:19-23
CH_MV = 64
CH_N = 64


class Spec(OrderedDict):
    def conv(self, name, cin, cout, k, groups=1):
        self[name + ".weight"] = (cout, cin // groups, k, k)
        self[name + ".bias"] = (cout,)

    # ---- layers.py building blocks -------------------------------------------------
    def depth_conv(self, p, cin, cout):                      # layers.py:135-152
        self.conv(p + ".conv1.0", cin, cin, 1)
        self.conv(p + ".depth_conv", cin, cin, 3, groups=cin)
        self.conv(p + ".conv2", cin, cout, 1)
        if cin != cout:
            self.conv(p + ".adaptor", cin, cout, 1)

    def dcb(self, p, cin, cout):                             # layers.py:199-209, 166-175
        self.depth_conv(p + ".block.0", cin, cout)
        inner = max(min(cout * 4, 1024), cout * 2)
        self.conv(p + ".block.1.conv.0", cout, inner, 1)
        self.conv(p + ".block.1.conv.2", inner, cout, 1)

    def dcb2(self, p, cin, cout):                            # layers.py:212-222, 182-190
        self.depth_conv(p + ".block.0", cin, cout)
        self.conv(p + ".block.1.conv", cout, cout * 4, 1)
        self.conv(p + ".block.1.conv_out", cout * 2, cout, 1)

    def rbws(self, p, cin, cout):                            # layers.py:51-60 (stride 2)
        self.conv(p + ".conv1", cin, cout, 3)
        self.conv(p + ".conv2", cout, cout, 3)
        self.conv(p + ".downsample", cin, cout, 1)

    def rbu(self, p, cin, cout):                             # layers.py:85-91
        self.conv(p + ".subpel_conv.0", cin, cout * 4, 1)
        self.conv(p + ".conv", cout, cout, 3)
        self.conv(p + ".upsample.0", cin, cout * 4, 1)

    def res_block(self, p, ch, bottleneck=False):            # video_net.py:59-67
        mid = ch // 2 if bottleneck else ch
        self.conv(p + ".conv1", ch, mid, 3)
        self.conv(p + ".conv2", mid, ch, 3)

    def unet(self, p, cin, cout, block):                     # video_net.py:129-149 / 173-193
        blk = getattr(self, block)
        blk(p + ".conv1", cin, 32)
        blk(p + ".conv2", 32, 64)
        blk(p + ".conv3", 64, 128)
        for i in range(4):
            blk(f"{p}.context_refine.{i}", 128, 128)
        self.conv(p + ".up3.0", 128, 64 * 4, 1)
        blk(p + ".up_conv3", 128, 64)
        self.conv(p + ".up2.0", 64, 32 * 4, 1)
        blk(p + ".up_conv2", 64, cout)

    def bit_estimator(self, p, ch):                          # entropy_models.py:58-109
        for k in ("f1", "f2", "f3"):
            self[f"{p}.{k}.h"] = (1, ch, 1, 1)
            self[f"{p}.{k}.b"] = (1, ch, 1, 1)
            self[f"{p}.{k}.a"] = (1, ch, 1, 1)
        self[f"{p}.f4.h"] = (1, ch, 1, 1)
        self[f"{p}.f4.b"] = (1, ch, 1, 1)

    def hyper_pair(self, penc, pdec, y_ch, z_ch, reduced):   # video_net.py:217-251
        if reduced:
            self.conv(penc + ".0", y_ch, z_ch, 3)
            self.conv(penc + ".2", z_ch, z_ch, 3)
            self.conv(penc + ".4", z_ch, z_ch, 3)
        else:
            self.conv(penc + ".0", y_ch, z_ch, 3)
            for i in (2, 4, 6, 8):
                self.conv(f"{penc}.{i}", z_ch, z_ch, 3)
        self.conv(pdec + ".0", z_ch, y_ch, 3)
        self.conv(pdec + ".2.0", y_ch, y_ch * 4, 1)
        self.conv(pdec + ".4", y_ch, y_ch, 3)
        self.conv(pdec + ".6.0", y_ch, y_ch * 4, 1)
        self.conv(pdec + ".8", y_ch, y_ch, 3)


def dmc_spec(anchor_num=4):
    """Registration order == the reference's (parameters of CompressionModel first)."""
    s = Spec()
    s.bit_estimator("bit_estimator_z", CH_16X)
    s.bit_estimator("bit_estimator_z_mv", 64)
    for lv in range(4):                                      # video_net.py:79-102
        p = f"optic_flow.moduleBasic.{lv}"
        for i, (ci, co) in enumerate(((8, 32), (32, 64), (64, 32), (32, 16), (16, 2))):
            s.conv(f"{p}.conv{i + 1}", ci, co, 7)
    s.conv("align.conv_offset.0", CH_1X + 5, CH_2X, 3)       # video_model.py:26-41
    s.conv("align.conv_offset.2", CH_2X, CH_2X, 3)
    s.conv("align.conv_offset.4", CH_2X, 3 * 16 * 2, 3)
    s.conv("align.fusion", CH_1X * 2, CH_1X, 1, groups=16)
    m = "mv_encoder"                                         # video_model.py:121-136
    s.rbws(m + ".enc_1.0", 2, CH_MV)
    s.dcb(m + ".enc_1.1", CH_MV, CH_MV)
    s.rbws(m + ".enc_2", CH_MV, CH_MV)
    s.dcb(m + ".adaptor_0", CH_MV, CH_MV)
    s.dcb(m + ".adaptor_1", CH_MV * 2, CH_MV)
    s.rbws(m + ".enc_3.0", CH_MV, CH_MV)
    s.dcb(m + ".enc_3.1", CH_MV, CH_MV)
    s.conv(m + ".enc_3.2", CH_MV, CH_MV, 3)
    s.hyper_pair("mv_hyper_prior_encoder", "mv_hyper_prior_decoder", CH_MV, CH_N, False)
    s.dcb("mv_y_prior_fusion_adaptor_0", CH_MV, CH_MV * 2)
    s.dcb("mv_y_prior_fusion_adaptor_1", CH_MV * 2, CH_MV * 2)
    s.dcb("mv_y_prior_fusion.0", CH_MV * 2, CH_MV * 3)
    s.dcb("mv_y_prior_fusion.1", CH_MV * 3, CH_MV * 3)
    for i in (1, 2, 3):
        s.conv(f"mv_y_spatial_prior_adaptor_{i}", CH_MV * 4, CH_MV * 3, 1)
    s.dcb("mv_y_spatial_prior.0", CH_MV * 3, CH_MV * 3)
    s.dcb("mv_y_spatial_prior.1", CH_MV * 3, CH_MV * 3)
    s.dcb("mv_y_spatial_prior.2", CH_MV * 3, CH_MV * 2)
    m = "mv_decoder"                                         # video_model.py:149-163
    s.dcb(m + ".dec_1.0", CH_MV, CH_MV)
    s.rbu(m + ".dec_1.1", CH_MV, CH_MV)
    s.dcb(m + ".dec_1.2", CH_MV, CH_MV)
    s.rbu(m + ".dec_1.3", CH_MV, CH_MV)
    s.dcb(m + ".dec_1.4", CH_MV, CH_MV)
    s.rbu(m + ".dec_2", CH_MV, CH_MV)
    s.dcb(m + ".dec_3.0", CH_MV, CH_MV)
    s.conv(m + ".dec_3.1.0", CH_MV, 2 * 4, 1)
    s.conv("feature_adaptor_I", 3, CH_1X, 3)
    for i in range(3):
        s.conv(f"feature_adaptor.{i}", CH_1X, CH_1X, 1)
    m = "feature_extractor"                                  # video_model.py:66-74
    s.conv(m + ".conv1", CH_1X, CH_1X, 3)
    s.res_block(m + ".res_block1", CH_1X)
    s.conv(m + ".conv2", CH_1X, CH_2X, 3)
    s.res_block(m + ".res_block2", CH_2X)
    s.conv(m + ".conv3", CH_2X, CH_4X, 3)
    s.res_block(m + ".res_block3", CH_4X)
    m = "context_fusion_net"                                 # video_model.py:89-101
    s.conv(m + ".conv3_up.0", CH_4X, CH_2X * 4, 3)
    s.res_block(m + ".res_block3_up", CH_2X)
    s.conv(m + ".conv3_out", CH_4X, CH_4X, 3)
    s.res_block(m + ".res_block3_out", CH_4X)
    s.conv(m + ".conv2_up.0", CH_2X * 2, CH_1X * 4, 3)
    s.res_block(m + ".res_block2_up", CH_1X)
    s.conv(m + ".conv2_out", CH_2X * 2, CH_2X, 3)
    s.res_block(m + ".res_block2_out", CH_2X)
    s.conv(m + ".conv1_out", CH_1X * 2, CH_1X, 3)
    s.res_block(m + ".res_block1_out", CH_1X)
    m = "contextual_encoder"                                 # video_model.py:173-183
    s.conv(m + ".conv1", CH_1X + 3, CH_2X, 3)
    s.res_block(m + ".res1", CH_2X * 2, bottleneck=True)
    s.conv(m + ".conv2", CH_2X * 2, CH_4X, 3)
    s.res_block(m + ".res2", CH_4X * 2, bottleneck=True)
    s.conv(m + ".conv3", CH_4X * 2, CH_8X, 3)
    s.conv(m + ".conv4", CH_8X, CH_16X, 3)
    s.hyper_pair("contextual_hyper_prior_encoder", "contextual_hyper_prior_decoder",
                 CH_16X, CH_16X, True)
    s.conv("temporal_prior_encoder.0", CH_4X, CH_8X, 3)
    s.conv("temporal_prior_encoder.2", CH_8X, CH_16X, 3)
    s.dcb("y_prior_fusion_adaptor_0", CH_16X * 2, CH_16X * 3)
    s.dcb("y_prior_fusion_adaptor_1", CH_16X * 3, CH_16X * 3)
    s.dcb("y_prior_fusion.0", CH_16X * 3, CH_16X * 3)
    s.dcb("y_prior_fusion.1", CH_16X * 3, CH_16X * 3)
    for i in (1, 2, 3):
        s.conv(f"y_spatial_prior_adaptor_{i}", CH_16X * 4, CH_16X * 3, 1)
    s.dcb("y_spatial_prior.0", CH_16X * 3, CH_16X * 3)
    s.dcb("y_spatial_prior.1", CH_16X * 3, CH_16X * 3)
    s.dcb("y_spatial_prior.2", CH_16X * 3, CH_16X * 2)
    m = "contextual_decoder"                                 # video_model.py:196-206
    s.conv(m + ".up1.0", CH_16X, CH_8X * 4, 3)
    s.conv(m + ".up2.0", CH_8X, CH_4X * 4, 3)
    s.res_block(m + ".res1", CH_4X * 2, bottleneck=True)
    s.conv(m + ".up3.0", CH_4X * 2, CH_2X * 4, 3)
    s.res_block(m + ".res2", CH_2X * 2, bottleneck=True)
    s.conv(m + ".up4.0", CH_2X * 2, 32 * 4, 3)
    m = "recon_generation_net"                               # video_model.py:219-225
    s.conv(m + ".first_conv", CH_1X + 32, CH_1X, 3)
    s.unet(m + ".unet_1", CH_1X, CH_1X, "dcb")
    s.unet(m + ".unet_2", CH_1X, CH_1X, "dcb")
    s.conv(m + ".recon_conv", CH_1X, 3, 3)
    for n, ce, cd in (("mv_y", CH_MV, CH_MV), ("y", CH_2X * 2, CH_2X)):   # :311-323
        s[f"{n}_q_basic_enc"] = (1, ce, 1, 1)
        s[f"{n}_q_scale_enc"] = (anchor_num, 1, 1, 1)
        s[f"{n}_q_basic_dec"] = (1, cd, 1, 1)
        s[f"{n}_q_scale_dec"] = (anchor_num, 1, 1, 1)
    return s


def intra_spec(N=256, anchor_num=4):
    s = Spec()
    s.bit_estimator("bit_estimator_z", N)
    s.rbws("enc.enc_1.0", 3, 128)                            # image_model.py:16-30
    s.dcb2("enc.enc_1.1", 128, 128)
    s.rbws("enc.enc_2.0", 128, 192)
    s.dcb2("enc.enc_2.1", 192, 192)
    s.rbws("enc.enc_2.2", 192, N)
    s.dcb2("enc.enc_2.3", N, N)
    s.conv("enc.enc_2.4", N, N, 3)
    s.dcb2("hyper_enc.0", N, N)                              # image_model.py:68-78
    s.conv("hyper_enc.1", N, N, 3)
    s.conv("hyper_enc.3", N, N, 3)
    s.rbu("hyper_dec.0", N, N)
    s.rbu("hyper_dec.1", N, N)
    s.dcb2("hyper_dec.2", N, N)
    s.dcb2("y_prior_fusion.0", N, N * 2)                     # image_model.py:80-92
    s.dcb2("y_prior_fusion.1", N * 2, N * 3)
    for i in (1, 2, 3):
        s.conv(f"y_spatial_prior_adaptor_{i}", N * 4, N * 3, 1)
    s.dcb2("y_spatial_prior.0", N * 3, N * 3)
    s.dcb2("y_spatial_prior.1", N * 3, N * 2)
    s.dcb2("y_spatial_prior.2", N * 2, N * 2)
    s.dcb2("dec.dec_1.0", N, N)                              # image_model.py:38-53
    s.rbu("dec.dec_1.1", N, N)
    s.dcb2("dec.dec_1.2", N, N)
    s.rbu("dec.dec_1.3", N, 192)
    s.dcb2("dec.dec_1.4", 192, 192)
    s.rbu("dec.dec_1.5", 192, 128)
    s.dcb2("dec.dec_2.0", 128, 128)
    s.rbu("dec.dec_2.1", 128, 16)
    s.unet("refine.0", 16, 16, "dcb2")                       # image_model.py:95-98
    s.conv("refine.1", 16, 3, 3)
    s["q_basic_enc"] = (1, 128, 1, 1)                        # image_model.py:100-105
    s["q_scale_enc"] = (anchor_num, 1, 1, 1)
    s["q_basic_dec"] = (1, 128, 1, 1)
    s["q_scale_dec"] = (anchor_num, 1, 1, 1)
    return s


def synth_state_dict(spec, seed, q_anchors=(1.5, 1.15, 0.85, 0.6)):
    """Deterministic random-init weights for tests and benchmarks (no checkpoints offline).

    Conv weights/biases ~ U(-b, b), b = 1/sqrt(fan_in) (PyTorch's default Conv2d init scale);
    factorised-prior parameters ~ N(0, 0.01) like the reference (entropy_models.py:62-68);
    q_scale anchors distinct and decreasing, q_basic = 1 + 0.1*U(-1,1), so that q handling
    is exercised (a fresh reference model has every q == 1).
    CPU torch.Generator => identical on every machine with the same torch build.
    """
    import torch
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    sd = OrderedDict()
    for name, shape in spec.items():
        if name.endswith(".weight"):
            fan_in = shape[1] * shape[2] * shape[3]
            b = 1.0 / fan_in ** 0.5
            sd[name] = (torch.rand(shape, generator=g) * 2 - 1) * b
        elif name.endswith(".bias"):
            w = spec[name[:-5] + ".weight"]
            b = 1.0 / (w[1] * w[2] * w[3]) ** 0.5
            sd[name] = (torch.rand(shape, generator=g) * 2 - 1) * b
        elif "bit_estimator" in name:
            sd[name] = torch.randn(shape, generator=g) * 0.01
        elif "q_scale" in name:
            sd[name] = torch.tensor(q_anchors, dtype=torch.float32).reshape(shape) * \
                (1.0 + 0.05 * (len(sd) % 3))
        elif "q_basic" in name:
            sd[name] = 1.0 + 0.1 * (torch.rand(shape, generator=g) * 2 - 1)
        else:
            raise KeyError(name)
    return sd
