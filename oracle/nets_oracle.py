"""TEST INFRASTRUCTURE -- CPU restatement of the reference's network arithmetic.

Plain torch fp32 functional code over a flat ``state_dict`` (name -> tensor).  It restates
what the reference's nn.Module graph computes for the DCVC-DC per-frame coding path; it is
NOT used by the product (dcvc_b200/), only by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference leg.

Pinned against the unmodified reference (imported from /root/reference) by
tests/golden/make_golden.py -> tests/test_oracle_golden.py.

Every function cites the reference file:line (relative to /root/reference/DCVC-DC) it follows.
"""
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------- primitives
def conv(sd, name, x, stride=1, pad=0, groups=1):
    return F.conv2d(x, sd[name + ".weight"], sd[name + ".bias"], stride, pad, 1, groups)


def lrelu(x, slope=0.01):
    return F.leaky_relu(x, slope)


def shuffle2(x):
    return F.pixel_shuffle(x, 2)


def subpel(sd, name, x, k):
    """src/models/layers.py:23-34 -- conv to 4*C then depth-to-space(2). `name` is the Sequential."""
    return shuffle2(conv(sd, name + ".0", x, 1, k // 2))


def flow_warp(feature, flow):
    """src/models/video_net.py:11-38 -- bilinear backward warp, border padding,
    align_corners=True: sample `feature` at pixel (x + flow_x, y + flow_y)."""
    n, _, h, w = flow.shape
    gx = torch.linspace(-1.0, 1.0, w, dtype=torch.float32).view(1, 1, 1, w).expand(n, -1, h, -1)
    gy = torch.linspace(-1.0, 1.0, h, dtype=torch.float32).view(1, 1, h, 1).expand(n, -1, -1, w)
    grid = torch.cat([gx, gy], 1)
    fl = torch.cat([flow[:, 0:1] / ((feature.size(3) - 1.0) / 2.0),
                    flow[:, 1:2] / ((feature.size(2) - 1.0) / 2.0)], 1)
    grid = (grid + fl).permute(0, 2, 3, 1)
    return F.grid_sample(feature, grid, mode="bilinear", padding_mode="border",
                         align_corners=True)


def bilinear_up2(x):
    """src/models/video_net.py:41-47"""
    return F.interpolate(x, (x.size(2) * 2, x.size(3) * 2), mode="bilinear", align_corners=False)


def bilinear_down2(x):
    """src/models/video_net.py:50-55"""
    return F.interpolate(x, (x.size(2) // 2, x.size(3) // 2), mode="bilinear",
                         align_corners=False)


# --------------------------------------------------------------------------- blocks
def res_block(sd, p, x, slope=0.01, end_with_relu=False):
    """src/models/video_net.py:58-76 (bottleneck only changes the weight shapes)."""
    out = lrelu(x, slope)
    out = conv(sd, p + ".conv1", out, 1, 1)
    out = lrelu(out, slope)
    out = conv(sd, p + ".conv2", out, 1, 1)
    if end_with_relu:
        out = lrelu(out, slope)
    return x + out


def depth_conv(sd, p, x):
    """src/models/layers.py:135-163 (stride 1 only: the DC models never use stride 2)."""
    identity = x
    if (p + ".adaptor.weight") in sd:
        identity = conv(sd, p + ".adaptor", x)
    out = lrelu(conv(sd, p + ".conv1.0", x), 0.01)
    out = conv(sd, p + ".depth_conv", out, 1, 1, groups=out.size(1))
    out = conv(sd, p + ".conv2", out)
    return out + identity


def conv_ffn(sd, p, x):
    """src/models/layers.py:166-179"""
    out = lrelu(conv(sd, p + ".conv.0", x), 0.1)
    out = lrelu(conv(sd, p + ".conv.2", out), 0.1)
    return x + out


def conv_ffn2(sd, p, x):
    """src/models/layers.py:182-196"""
    x1, x2 = conv(sd, p + ".conv", x).chunk(2, 1)
    return x + conv(sd, p + ".conv_out", x1 * lrelu(x2, 0.1))


def dcb(sd, p, x):
    """DepthConvBlock, src/models/layers.py:199-209"""
    return conv_ffn(sd, p + ".block.1", depth_conv(sd, p + ".block.0", x))


def dcb2(sd, p, x):
    """DepthConvBlock2, src/models/layers.py:212-222"""
    return conv_ffn2(sd, p + ".block.1", depth_conv(sd, p + ".block.0", x))


def rbws(sd, p, x, stride=2):
    """ResidualBlockWithStride, src/models/layers.py:42-73"""
    out = lrelu(conv(sd, p + ".conv1", x, stride, 1), 0.01)
    out = lrelu(conv(sd, p + ".conv2", out, 1, 1), 0.1)
    identity = conv(sd, p + ".downsample", x, stride) if stride != 1 else x
    return out + identity


def rbu(sd, p, x):
    """ResidualBlockUpsample, src/models/layers.py:76-101"""
    out = lrelu(subpel(sd, p + ".subpel_conv", x, 1), 0.01)
    out = lrelu(conv(sd, p + ".conv", out, 1, 1), 0.1)
    return out + subpel(sd, p + ".upsample", x, 1)


def unet(sd, p, x, block):
    """UNet / UNet2, src/models/video_net.py:129-214 (`block` = dcb or dcb2)."""
    x1 = block(sd, p + ".conv1", x)
    x2 = block(sd, p + ".conv2", F.max_pool2d(x1, 2, 2))
    x3 = block(sd, p + ".conv3", F.max_pool2d(x2, 2, 2))
    for i in range(4):
        x3 = block(sd, f"{p}.context_refine.{i}", x3)
    d3 = subpel(sd, p + ".up3", x3, 1)
    d3 = block(sd, p + ".up_conv3", torch.cat((x2, d3), 1))
    d2 = subpel(sd, p + ".up2", d3, 1)
    return block(sd, p + ".up_conv2", torch.cat((x1, d2), 1))


def hyper_enc5(sd, p, x):
    """get_hyper_enc_dec_models(reduce_enc_layer=False) encoder, src/models/video_net.py:226-237"""
    x = lrelu(conv(sd, p + ".0", x, 1, 1))
    x = lrelu(conv(sd, p + ".2", x, 1, 1))
    x = lrelu(conv(sd, p + ".4", x, 2, 1))
    x = lrelu(conv(sd, p + ".6", x, 1, 1))
    return conv(sd, p + ".8", x, 2, 1)


def hyper_enc3(sd, p, x):
    """reduce_enc_layer=True encoder, src/models/video_net.py:218-225"""
    x = lrelu(conv(sd, p + ".0", x, 1, 1))
    x = lrelu(conv(sd, p + ".2", x, 2, 1))
    return conv(sd, p + ".4", x, 2, 1)


def hyper_dec(sd, p, x):
    """src/models/video_net.py:239-249"""
    x = lrelu(conv(sd, p + ".0", x, 1, 1))
    x = lrelu(subpel(sd, p + ".2", x, 1))
    x = lrelu(conv(sd, p + ".4", x, 1, 1))
    x = lrelu(subpel(sd, p + ".6", x, 1))
    return conv(sd, p + ".8", x, 1, 1)


# --------------------------------------------------------------------------- DMC sub-nets
def me_basic(sd, p, x):
    """MEBasic, src/models/video_net.py:79-95"""
    x = F.relu(conv(sd, p + ".conv1", x, 1, 3))
    x = F.relu(conv(sd, p + ".conv2", x, 1, 3))
    x = F.relu(conv(sd, p + ".conv3", x, 1, 3))
    x = F.relu(conv(sd, p + ".conv4", x, 1, 3))
    return conv(sd, p + ".conv5", x, 1, 3)


def spynet(sd, im1, im2, p="optic_flow"):
    """ME_Spynet.forward, src/models/video_net.py:104-126"""
    L = 4
    l1, l2 = [im1], [im2]
    for lv in range(L - 1):
        l1.append(F.avg_pool2d(l1[lv], 2, 2))
        l2.append(F.avg_pool2d(l2[lv], 2, 2))
    sh = l2[L - 1].shape
    flow = torch.zeros(im1.size(0), 2, sh[2] // 2, sh[3] // 2, dtype=im1.dtype)
    for lv in range(L):
        up = bilinear_up2(flow) * 2.0
        k = L - 1 - lv
        flow = up + me_basic(sd, f"{p}.moduleBasic.{lv}",
                             torch.cat([l1[k], flow_warp(l2[k], up), up], 1))
    return flow


def mv_enc(sd, x, context, q, p="mv_encoder"):
    """MvEnc.forward, src/models/video_model.py:138-146"""
    out = dcb(sd, p + ".enc_1.1", rbws(sd, p + ".enc_1.0", x))
    out = out * q
    out = rbws(sd, p + ".enc_2", out)
    if context is None:
        out = dcb(sd, p + ".adaptor_0", out)
    else:
        out = dcb(sd, p + ".adaptor_1", torch.cat((out, context), 1))
    out = dcb(sd, p + ".enc_3.1", rbws(sd, p + ".enc_3.0", out))
    return conv(sd, p + ".enc_3.2", out, 2, 1)


def mv_dec(sd, x, q, p="mv_decoder"):
    """MvDec.forward, src/models/video_model.py:165-170"""
    f = dcb(sd, p + ".dec_1.0", x)
    f = rbu(sd, p + ".dec_1.1", f)
    f = dcb(sd, p + ".dec_1.2", f)
    f = rbu(sd, p + ".dec_1.3", f)
    feature = dcb(sd, p + ".dec_1.4", f)
    out = rbu(sd, p + ".dec_2", feature) * q
    mv = subpel(sd, p + ".dec_3.1", dcb(sd, p + ".dec_3.0", out), 1)
    return mv, feature


def feature_extractor(sd, f, p="feature_extractor"):
    """FeatureExtractor.forward, src/models/video_model.py:76-86"""
    l1 = res_block(sd, p + ".res_block1", conv(sd, p + ".conv1", f, 1, 1))
    l2 = res_block(sd, p + ".res_block2", conv(sd, p + ".conv2", l1, 2, 1))
    l3 = res_block(sd, p + ".res_block3", conv(sd, p + ".conv3", l2, 2, 1))
    return l1, l2, l3


def offset_diversity(sd, x, aux, flow, p="align"):
    """OffsetDiversity.forward, src/models/video_model.py:43-63 (16 groups, 2 offsets, 40 px)."""
    B, C, H, W = x.shape
    G, K = 16, 2
    out = lrelu(conv(sd, p + ".conv_offset.0", aux, 2, 1), 0.1)
    out = lrelu(conv(sd, p + ".conv_offset.2", out, 1, 1), 0.1)
    out = conv(sd, p + ".conv_offset.4", out, 1, 1)
    out = bilinear_up2(out)
    o1, o2, mask = torch.chunk(out, 3, dim=1)
    mask = torch.sigmoid(mask)
    offset = 40 * torch.tanh(torch.cat((o1, o2), dim=1))
    offset = offset + flow.repeat(1, G * K, 1, 1)
    offset = offset.view(B * G * K, 2, H, W)
    mask = mask.view(B * G * K, 1, H, W)
    xs = x.view(B * G, C // G, H, W).repeat(K, 1, 1, 1)
    xs = flow_warp(xs, offset) * mask
    xs = xs.view(B, C * K, H, W)
    return conv(sd, p + ".fusion", xs, groups=G)


def context_fusion(sd, c1, c2, c3, p="context_fusion_net"):
    """MultiScaleContextFusion.forward, src/models/video_model.py:103-118"""
    c3_up = res_block(sd, p + ".res_block3_up", subpel(sd, p + ".conv3_up", c3, 3))
    c3_out = res_block(sd, p + ".res_block3_out", conv(sd, p + ".conv3_out", c3, 1, 1))
    cat2 = torch.cat((c3_up, c2), 1)
    c2_up = res_block(sd, p + ".res_block2_up", subpel(sd, p + ".conv2_up", cat2, 3))
    c2_out = res_block(sd, p + ".res_block2_out", conv(sd, p + ".conv2_out", cat2, 1, 1))
    c1_out = res_block(sd, p + ".res_block1_out",
                       conv(sd, p + ".conv1_out", torch.cat((c2_up, c1), 1), 1, 1))
    return c1 + c1_out, c2 + c2_out, c3 + c3_out


def contextual_encoder(sd, x, c1, c2, c3, q, p="contextual_encoder"):
    """ContextualEncoder.forward, src/models/video_model.py:185-193"""
    f = conv(sd, p + ".conv1", torch.cat([x, c1], 1), 2, 1)
    f = res_block(sd, p + ".res1", torch.cat([f, c2], 1), 0.1, True)
    f = f * q
    f = conv(sd, p + ".conv2", f, 2, 1)
    f = res_block(sd, p + ".res2", torch.cat([f, c3], 1), 0.1, True)
    f = conv(sd, p + ".conv3", f, 2, 1)
    return conv(sd, p + ".conv4", f, 2, 1)


def contextual_decoder(sd, y_hat, c2, c3, q, p="contextual_decoder"):
    """ContextualDecoder.forward, src/models/video_model.py:208-216"""
    f = subpel(sd, p + ".up1", y_hat, 3)
    f = subpel(sd, p + ".up2", f, 3)
    f = res_block(sd, p + ".res1", torch.cat([f, c3], 1), 0.1, True)
    f = subpel(sd, p + ".up3", f, 3)
    f = f * q
    f = res_block(sd, p + ".res2", torch.cat([f, c2], 1), 0.1, True)
    return subpel(sd, p + ".up4", f, 3)


def recon_generation(sd, ctx, res, p="recon_generation_net"):
    """ReconGeneration.forward, src/models/video_model.py:227-232"""
    f = conv(sd, p + ".first_conv", torch.cat((ctx, res), 1), 1, 1)
    f = unet(sd, p + ".unet_1", f, dcb)
    f = unet(sd, p + ".unet_2", f, dcb)
    return f, conv(sd, p + ".recon_conv", f, 1, 1)


# --------------------------------------------------------------------------- intra sub-nets
def intra_encoder(sd, x, q, p="enc"):
    """IntraEncoder.forward, src/models/image_model.py:32-35"""
    out = dcb2(sd, p + ".enc_1.1", rbws(sd, p + ".enc_1.0", x))
    out = out * q
    out = dcb2(sd, p + ".enc_2.1", rbws(sd, p + ".enc_2.0", out))
    out = dcb2(sd, p + ".enc_2.3", rbws(sd, p + ".enc_2.2", out))
    return conv(sd, p + ".enc_2.4", out, 2, 1)


def intra_decoder(sd, x, q, p="dec"):
    """IntraDecoder.forward, src/models/image_model.py:55-58"""
    out = rbu(sd, p + ".dec_1.1", dcb2(sd, p + ".dec_1.0", x))
    out = rbu(sd, p + ".dec_1.3", dcb2(sd, p + ".dec_1.2", out))
    out = rbu(sd, p + ".dec_1.5", dcb2(sd, p + ".dec_1.4", out))
    out = out * q
    return rbu(sd, p + ".dec_2.1", dcb2(sd, p + ".dec_2.0", out))


def intra_hyper_enc(sd, y, p="hyper_enc"):
    """IntraNoAR.hyper_enc, src/models/image_model.py:68-73"""
    out = dcb2(sd, p + ".0", y)
    out = lrelu(conv(sd, p + ".1", out, 2, 1))
    return conv(sd, p + ".3", out, 2, 1)


def intra_hyper_dec(sd, z, p="hyper_dec"):
    """IntraNoAR.hyper_dec, src/models/image_model.py:74-78"""
    return dcb2(sd, p + ".2", rbu(sd, p + ".1", rbu(sd, p + ".0", z)))


def intra_refine(sd, x, p="refine"):
    """IntraNoAR.refine, src/models/image_model.py:95-98"""
    return conv(sd, p + ".1", unet(sd, p + ".0", x, dcb2), 1, 1)
