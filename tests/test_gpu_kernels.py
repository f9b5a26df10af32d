"""-m gpu parity tests, kernel by kernel: the CUDA path (through the C ABI) against the CPU
oracle on the same seeded inputs.  Float kernels: tolerance stated per test; integer/byte
kernels: bit-exact."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from dcvc_b200 import arch, layers, ops
from dcvc_b200 import entropy_models as em
from dcvc_b200.tensor import Act
from oracle import entropy_oracle as eo
from oracle import nets_oracle as no
from oracle import rans_oracle

from gpu_util import dev, from_act, maxdiff, to_act

pytestmark = pytest.mark.gpu


def _rand(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.rand(*shape, generator=g) * 2 - 1) * scale


def _conv_sd(name, cin, cout, k, seed=0, groups=1):
    g = torch.Generator().manual_seed(seed)
    b = 1.0 / (cin // groups * k * k) ** 0.5
    return {name + ".weight": (torch.rand(cout, cin // groups, k, k, generator=g) * 2 - 1) * b,
            name + ".bias": (torch.rand(cout, generator=g) * 2 - 1) * b}


# ------------------------------------------------------------------------------ convolution
@pytest.mark.parametrize("cin,cout,k,stride,h,w", [
    (8, 32, 7, 1, 24, 40), (3, 48, 3, 1, 17, 23), (53, 64, 3, 2, 32, 48), (48, 48, 3, 1, 32, 32),
    (64, 256, 1, 1, 16, 24), (16, 2, 7, 1, 20, 20), (128, 96, 3, 2, 16, 16), (48, 3, 3, 1, 16, 16),
    (2, 64, 3, 2, 32, 32), (64, 64, 1, 2, 10, 14), (384, 384, 1, 1, 6, 8), (20, 20, 3, 1, 9, 131),
])
def test_conv_plain(cin, cout, k, stride, h, w):
    sd = _conv_sd("c", cin, cout, k, seed=cin + cout)
    x = _rand(1, cin, h, w, seed=1)
    pad = k // 2 if k > 1 else 0
    ref = F.conv2d(x, sd["c.weight"], sd["c.bias"], stride, pad)
    net = layers.Net(sd, dev())
    out = ops.conv(to_act(x), net.cw("c"), stride=stride, pad=pad)
    torch.cuda.synchronize()
    assert maxdiff(from_act(out), ref) < 2e-5          # fp32 FMA vs fp32 reference


def test_conv_fused_epilogue_and_slices():
    """bias + lrelu + residuals + scale + out2, reading/writing channel slices of wider buffers,
    with an input prologue activation."""
    cin, cout, h, w = 24, 40, 12, 20
    sd = _conv_sd("c", cin, cout, 3, seed=5)
    x, r1, r2 = _rand(1, cin, h, w, seed=1), _rand(1, cout, h, w, seed=2), _rand(1, cout, h, w, seed=3)
    q = torch.rand(cout, generator=torch.Generator().manual_seed(11)) + 0.5
    ref = F.leaky_relu(F.conv2d(F.leaky_relu(x, 0.1), sd["c.weight"], sd["c.bias"], 1, 1), 0.01)
    ref = (ref + r1 + r2) * q.view(1, -1, 1, 1)
    net = layers.Net(sd, dev())
    xa = to_act(x, cstride=32, coff=4)
    out = to_act(torch.zeros(1, cout, h, w), cstride=48, coff=8)
    out2 = Act.new(h, w, cout, dev())
    ops.conv(xa, net.cw("c"), in_slope=0.1, act=0.01, res=to_act(r1), res2=to_act(r2, 44, 1),
             scale=q.to(dev()), out=out, out2=out2, out2_slope=0.2)
    torch.cuda.synchronize()
    assert maxdiff(from_act(out), ref) < 2e-5
    assert maxdiff(from_act(out2), F.leaky_relu(ref, 0.2)) < 2e-5


def test_conv_shuffle_gate_clamp():
    sd = _conv_sd("c", 32, 64, 3, seed=9)
    x = _rand(1, 32, 10, 14, seed=4)
    net = layers.Net(sd, dev())
    ref = F.pixel_shuffle(F.conv2d(x, sd["c.weight"], sd["c.bias"], 1, 1), 2)
    out = ops.conv(to_act(x), net.cw("c"), shuffle=True)
    assert maxdiff(from_act(out), ref) < 2e-5
    x1, x2 = F.conv2d(x, sd["c.weight"], sd["c.bias"], 1, 1).chunk(2, 1)
    out = ops.conv(to_act(x), net.cw("c", gate=True), gate_slope=0.1)
    assert maxdiff(from_act(out), x1 * F.leaky_relu(x2, 0.1)) < 2e-5
    out = ops.conv(to_act(x * 4), net.cw("c"), clamp01=True)
    assert maxdiff(from_act(out), F.conv2d(x * 4, sd["c.weight"], sd["c.bias"], 1, 1).clamp(0, 1)) < 5e-5


def test_dwconv():
    c, h, w = 70, 13, 19
    sd = _conv_sd("d", c, c, 3, groups=c, seed=3)
    x = _rand(1, c, h, w, seed=8)
    ref = F.conv2d(x, sd["d.weight"], sd["d.bias"], 1, 1, 1, c)
    net = layers.Net(sd, dev())
    wt, b = net.dw("d")
    out = ops.dwconv3x3(to_act(x), wt, b)
    assert maxdiff(from_act(out), ref) < 1e-5


# ------------------------------------------------------------------------------ blocks
def _block_sd(builder, *args, seed=0):
    spec = arch.Spec()
    getattr(spec, builder)("b", *args)
    return arch.synth_state_dict(spec, seed)


@pytest.mark.parametrize("block,oracle_fn,cin,cout", [
    ("dcb", no.dcb, 64, 64), ("dcb", no.dcb, 48, 32), ("dcb2", no.dcb2, 32, 32),
    ("dcb2", no.dcb2, 64, 96)])
def test_depth_conv_blocks(block, oracle_fn, cin, cout):
    sd = _block_sd(block, cin, cout, seed=cin)
    x = _rand(1, cin, 14, 18, seed=2)
    ref = oracle_fn(sd, "b", x)
    out = getattr(layers, block)(layers.Net(sd, dev()), "b", to_act(x))
    assert maxdiff(from_act(out), ref) < 5e-5


def test_residual_blocks():
    x = _rand(1, 32, 16, 24, seed=3)
    sd = _block_sd("rbws", 32, 48)
    assert maxdiff(from_act(layers.rbws(layers.Net(sd, dev()), "b", to_act(x))), no.rbws(sd, "b", x)) < 5e-5
    sd = _block_sd("rbu", 32, 24)
    assert maxdiff(from_act(layers.rbu(layers.Net(sd, dev()), "b", to_act(x))), no.rbu(sd, "b", x)) < 5e-5
    sd = _block_sd("res_block", 32)
    for slope, end in ((0.01, False), (0.1, True)):
        ref = no.res_block(sd, "b", x, slope, end)
        out = layers.res_block(layers.Net(sd, dev()), "b", to_act(x), slope, end)
        assert maxdiff(from_act(out), ref) < 5e-5


@pytest.mark.parametrize("block,ob", [("dcb", no.dcb), ("dcb2", no.dcb2)])
def test_unet(block, ob):
    sd = _block_sd("unet", 16, 16, block, seed=4)
    x = _rand(1, 16, 16, 24, seed=5)
    ref = no.unet(sd, "b", x, ob)
    out = layers.unet(layers.Net(sd, dev()), "b", to_act(x), getattr(layers, block))
    assert maxdiff(from_act(out), ref) < 2e-4


def test_hyper_nets():
    spec = arch.Spec()
    spec.hyper_pair("e5", "d5", 64, 64, False)
    spec.hyper_pair("e3", "d3", 128, 128, True)
    sd = arch.synth_state_dict(spec, 6)
    net = layers.Net(sd, dev())
    y = _rand(1, 64, 8, 12, seed=1)
    assert maxdiff(from_act(layers.hyper_enc5(net, "e5", to_act(y))), no.hyper_enc5(sd, "e5", y)) < 5e-5
    z = _rand(1, 64, 2, 3, seed=2)
    assert maxdiff(from_act(layers.hyper_dec(net, "d5", to_act(z))), no.hyper_dec(sd, "d5", z)) < 5e-5
    y = _rand(1, 128, 8, 12, seed=3)
    assert maxdiff(from_act(layers.hyper_enc3(net, "e3", to_act(y))), no.hyper_enc3(sd, "e3", y)) < 5e-5


# ------------------------------------------------------------------------------ warp group
def test_flow_warp_and_resample():
    h, w = 36, 52
    feat = _rand(1, 48, h, w, seed=1)
    flow = _rand(1, 2, h, w, seed=2, scale=9.0)
    ref = no.flow_warp(feat, flow)
    out = ops.flow_warp(to_act(feat), to_act(flow))
    assert maxdiff(from_act(out), ref) < 2e-5
    up = ops.resize_bilinear(to_act(flow), 2 * h, 2 * w, mul=2.0)
    assert maxdiff(from_act(up), no.bilinear_up2(flow) * 2.0) < 1e-5
    dn = ops.resize_bilinear(to_act(flow), h // 2, w // 2, mul=0.5)
    assert maxdiff(from_act(dn), no.bilinear_down2(flow) / 2) < 1e-6
    odd = _rand(1, 2, 17, 31, seed=3)
    dn = ops.resize_bilinear(to_act(odd), 8, 15)
    assert maxdiff(from_act(dn), no.bilinear_down2(odd)) < 1e-5
    assert maxdiff(from_act(ops.avg_pool2(to_act(feat))), F.avg_pool2d(feat, 2, 2)) < 1e-6
    assert maxdiff(from_act(ops.max_pool2(to_act(feat))), F.max_pool2d(feat, 2, 2)) == 0.0


def test_gathers_from_binary16_sources():
    """flow_warp / offset_diversity with binary16 source AND destination views equal the fp32-
    source kernels run on the same rounded values, rounded once at the store."""
    h, w = 36, 52
    feat = _rand(1, 48, h, w, seed=1)
    flow = _rand(1, 2, h, w, seed=2, scale=9.0)
    f16 = ops.copy(to_act(feat), f16=True)
    feat_r = to_act(feat.half().float())
    want = ops.flow_warp(feat_r, to_act(flow), out=Act.new(h, w, 48, dev(), f16=True))
    got = ops.flow_warp(f16, to_act(flow), out=Act.new(h, w, 48, dev(), f16=True))
    assert torch.equal(want.buf, got.buf)
    offs = to_act(_rand(1, 96, h // 2, w // 2, seed=4, scale=0.7))
    fw = _rand(1, 1, 48, 6, seed=5).reshape(48, 6).to(dev()).contiguous()
    fb = _rand(1, 1, 1, 48, seed=6).reshape(48).to(dev()).contiguous()
    wide = torch.full((h, w, 96), 7.0, device=dev(), dtype=torch.float16)
    want = ops.offset_diversity(feat_r, offs, to_act(flow), fw, fb, out=Act.new(h, w, 48, dev(), f16=True))
    got = ops.offset_diversity(f16, offs, to_act(flow), fw, fb, out=Act(wide, 48, 48))
    torch.cuda.synchronize()
    assert torch.equal(want.buf[:, :, :48], wide[:, :, 48:])
    assert bool((wide[:, :, :48] == 7.0).all())
    with pytest.raises(ValueError):
        ops.flow_warp(f16, to_act(flow))                   # binary16 source -> fp32: not provided


def test_copy_pad_crop_and_layout():
    y = _rand(1, 8, 6, 7, seed=1)
    padded = ops.copy(to_act(y), h=8, w=8)
    assert maxdiff(from_act(padded), F.pad(y, (0, 1, 0, 2), mode="replicate")) == 0.0
    cropped = ops.copy(padded, h=6, w=7)
    assert maxdiff(from_act(cropped), y) == 0.0
    x = _rand(1, 3, 10, 12, seed=2).to(dev())
    a = Act.wrap(x)                                        # NCHW -> NHWC kernel
    assert maxdiff(from_act(a), x.cpu()) == 0.0
    b = Act.wrap(a.nchw())                                 # zero-copy round trip
    assert b.buf.data_ptr() == a.buf.data_ptr() and b.c0 == a.c0 and b.c == a.c
    assert maxdiff(ops.to_nchw(a).cpu(), x.cpu()) == 0.0


def test_offset_diversity():
    spec = arch.Spec()
    spec.conv("align.conv_offset.0", 53, 64, 3)
    spec.conv("align.conv_offset.2", 64, 64, 3)
    spec.conv("align.conv_offset.4", 64, 96, 3)
    spec.conv("align.fusion", 96, 48, 1, groups=16)
    sd = arch.synth_state_dict(spec, 2)
    sd["align.conv_offset.4.weight"] *= 8              # make offsets/masks non-trivial
    h, w = 32, 48
    x = _rand(1, 48, h, w, seed=1)
    aux = _rand(1, 53, h, w, seed=2)
    flow = _rand(1, 2, h, w, seed=3, scale=4.0)
    ref = no.offset_diversity(sd, x, aux, flow)
    net = layers.Net(sd, dev())
    t = ops.conv(to_act(aux), net.cw("align.conv_offset.0"), stride=2, act=0.1)
    t = ops.conv(t, net.cw("align.conv_offset.2"), act=0.1)
    offs = ops.conv(t, net.cw("align.conv_offset.4"))
    out = ops.offset_diversity(to_act(x), offs, to_act(flow),
                               net.sd["align.fusion.weight"].reshape(48, 6),
                               net.sd["align.fusion.bias"])
    assert maxdiff(from_act(out), ref) < 2e-4


# ------------------------------------------------------------------------------ entropy model
@pytest.mark.parametrize("dist,C,h,w", [("laplace", 64, 6, 8), ("gaussian", 128, 5, 7)])
def test_quad_steps_teacher_forced(dist, C, h, w):
    """Symbols, CDF indexes and y_hat bit-exact given the oracle's float inputs (the spatial
    prior's outputs are fed from the oracle so no conv rounding enters)."""
    g = torch.Generator().manual_seed(11)
    y = torch.randn(1, C, h, w, generator=g) * 6
    params = torch.cat((torch.rand(1, C, h, w, generator=g) * 2,            # quant_step (some < .5)
                        torch.exp(torch.rand(1, C, h, w, generator=g) * 9 - 5),   # scales
                        torch.randn(1, C, h, w, generator=g)), 1)
    priors = {s: torch.cat((torch.exp(torch.rand(1, C, h, w, generator=g) * 9 - 5),
                            torch.randn(1, C, h, w, generator=g)), 1) for s in (1, 2, 3)}
    ref = eo.quad_forward(y, params, lambda step, _: priors[step])
    edges = em.index_edges(dist).to(dev())
    catb = Act.new(h, w, 4 * C, dev())
    ops.fill(catb.slice(0, C), 0.0)
    ops.copy(to_act(params), out=catb.slice(C, 3 * C))
    y_out = Act.new(h, w, C, dev())
    ya = to_act(y)
    n = C // 4 * h * w
    bits = torch.zeros(1, device=dev())
    ws = torch.empty((n + 255) // 256, device=dev())
    for step in range(4):
        prior = to_act(priors[step]) if step else None
        sym = torch.empty(n, dtype=torch.int16, device=dev())
        idx = torch.empty(n, dtype=torch.uint8, device=dev())
        ops.quad_step_enc(ya, catb.slice(C, 3 * C), prior, catb.slice(0, C), y_out, edges, step,
                          em.DIST_ID[dist], sym, idx, bits, ws)
        ref_sym = ref["y_q_w"][step].reshape(-1).clamp(-30000, 30000).to(torch.int16)
        ref_idx = eo.build_indexes(ref["scales_w"][step], dist).reshape(-1)
        assert torch.equal(sym.cpu(), ref_sym), f"symbols differ at step {step}"
        assert torch.equal(idx.cpu().int(), ref_idx), f"indexes differ at step {step}"
        # decoder side: indexes from the same params, scatter reproduces y_hat bit-exactly
        idx2 = torch.empty_like(idx)
        d_hat, d_out = Act.new(h, w, C, dev(), zero=True), Act.new(h, w, C, dev(), zero=True)
        ops.quad_step_indexes(catb.slice(C, 3 * C), prior, d_hat, d_out, edges, step, idx2)
        assert torch.equal(idx2, idx)
    assert torch.equal(from_act(y_out), ref["y_hat"])
    ref_bits = eo.y_bits(ref["y_q"], ref["scales_hat"], dist).sum().item()
    assert abs(bits.item() - ref_bits) <= 1e-4 * ref_bits


def test_scale_index_all_edges():
    """Every bin edge +-1 ulp maps to the reference's index."""
    for dist in ("laplace", "gaussian"):
        edges = em.index_edges(dist)
        eb = edges[1:].view(torch.int32)
        s = torch.cat([(eb + d).view(torch.float32) for d in (-2, -1, 0, 1, 2)] +
                      [torch.tensor([0.0, 1e-6, 1e-5, 63.9, 64.0, 65.0, 1e4])])
        C = 4
        n = s.numel()
        w = (n + 3) // 4 * 4
        sp = torch.ones(w)
        sp[:n] = s
        # one-row latent: channel group 0 position (0, even x) only codes group 0 at step 0
        params = torch.zeros(1, 3 * C, 2, 2 * w)
        params[0, C, 0, 0::2] = sp                     # scales of channel 0 at m == 0 positions
        catb = to_act(params)
        y_hat, y_out = Act.new(2, 2 * w, C, dev(), zero=True), Act.new(2, 2 * w, C, dev(), zero=True)
        idx = torch.zeros(2 * 2 * w, dtype=torch.uint8, device=dev())
        ops.quad_step_indexes(catb, None, y_hat, y_out, edges.to(dev()), 0, idx)
        got = idx.cpu().view(2, 2 * w)[0, 0::2][:n].int()
        assert torch.equal(got, eo.build_indexes(s, dist))


def test_z_quant_and_bits():
    C, h, w = 64, 5, 7
    sd = arch.synth_state_dict(arch.dmc_spec(), 100)
    z = _rand(1, C, h, w, seed=4, scale=5.0)
    p = "bit_estimator_z_mv"
    rows = []
    for k in ("f1", "f2", "f3"):
        rows += [sd[f"{p}.{k}.h"], sd[f"{p}.{k}.b"], sd[f"{p}.{k}.a"]]
    rows += [sd[f"{p}.f4.h"], sd[f"{p}.f4.b"]]
    fparams = torch.stack([t.reshape(-1) for t in rows], 0).contiguous().to(dev())
    z_hat = Act.new(h, w, C, dev())
    sym = torch.empty(C * h * w, dtype=torch.int16, device=dev())
    bits = torch.zeros(1, device=dev())
    ws = torch.empty((C * h * w + 255) // 256, device=dev())
    ops.z_quant(to_act(z), z_hat, sym, fparams, bits, ws)
    ref = torch.round(z)
    assert torch.equal(from_act(z_hat), ref)
    assert torch.equal(sym.cpu(), ref.reshape(-1).to(torch.int16))
    ref_bits = eo.z_bits(sd, p, ref).sum().item()
    assert abs(bits.item() - ref_bits) <= 1e-4 * ref_bits
    back = Act.new(h, w, C, dev())
    ops.z_dequant(sym, back)
    assert torch.equal(from_act(back), ref)
