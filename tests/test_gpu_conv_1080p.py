"""-m gpu: the heaviest convolution launches of a 1080p P frame (profiles/r01_pframe_kernel_table.json
rows, SURVEY Appendix A shapes) at their REAL sizes and with their real epilogue variants, against
cuDNN fp32 (`F.conv2d` on the GPU with TF32 disabled) used purely as the checker.

The small-shape tests (test_gpu_conv_tc.py) cannot reach what these do: persistent CTAs looping
over ~16 000 tiles, halo windows with multi-chunk resident weights at 1088x1920, the TMA-store
depth-to-space epilogue with its four strided maps, tensors beyond 2^31 bytes.

Operand model: a binary16 activation is exact input; weights of a binary16-input conv are rounded
to binary16 (kind::f16), so the reference uses the same rounded operands and only fp32 summation
order differs.  fp32 activations go through kind::tf32 (operands cut to 10 mantissa bits):
bounded like test_gpu_conv_tc.py.  A binary16 OUTPUT adds one rounding: 2^-11 relative.
"""
import pytest
import torch
import torch.nn.functional as F

from dcvc_b200 import layers, ops
from dcvc_b200.tensor import Act

from gpu_util import dev

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _tc_mode():
    prev = ops.CONV_MODE
    ops.set_conv_mode(1)
    a, b = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = a, b
    ops.set_conv_mode(prev)


def _rand(shape, seed, f16=False):
    g = torch.Generator(device=dev()).manual_seed(seed)
    t = torch.rand(shape, generator=g, device=dev()) * 2 - 1
    return t.half() if f16 else t


def _act(t_nchw, cstride=None, coff=0):
    """GPU [1,C,H,W] (fp32 or half) -> Act holding the same values (inside a wider buffer when
    cstride/coff are given)."""
    _, c, h, w = t_nchw.shape
    f16 = t_nchw.dtype == torch.float16
    cs = cstride if cstride is not None else ((c + 7) // 8 * 8 if f16 else (c + 3) // 4 * 4)
    buf = torch.zeros((h, w, cs), dtype=t_nchw.dtype, device=dev())
    buf[:, :, coff:coff + c] = t_nchw[0].permute(1, 2, 0)
    return Act(buf, coff, c)


def _sd(cin, cout, k, seed):
    g = torch.Generator().manual_seed(seed)
    b = 1.0 / (cin * k * k) ** 0.5
    return {"c.weight": (torch.rand(cout, cin, k, k, generator=g) * 2 - 1) * b,
            "c.bias": (torch.rand(cout, generator=g) * 2 - 1) * b}


# name, cin, cout, k, stride, H_in, W_in, in_f16, out_f16, res ('', 'r' f16, 'R' fp32),
# res2 ('', '2h'), extra in {'', 'act_copy', 'shuffle', 'clamp4', 'gate', 'scale'}
CASES = [
    ("ffn_up_128_512_q", 128, 512, 1, 1, 272, 480, True, True, "", "", ""),
    ("ffn_down_512_128_q", 512, 128, 1, 1, 272, 480, True, True, "r", "", "scale"),
    ("c3_48_48_io", 48, 48, 3, 1, 1088, 1920, True, True, "", "", ""),
    ("c3_48_48_ior", 48, 48, 3, 1, 1088, 1920, True, True, "r", "", ""),
    ("c3_48_48_ior2h", 48, 48, 3, 1, 1088, 1920, True, True, "r", "2h", ""),
    ("c3_48_48_io_actcopy", 48, 48, 3, 1, 1088, 1920, True, True, "", "", "act_copy"),
    ("spynet_7x7_64_32", 64, 32, 7, 1, 1088, 1920, True, True, "", "", ""),
    ("spynet_7x7_32_64", 32, 64, 7, 1, 1088, 1920, True, True, "", "", ""),
    ("spynet_7x7_32_16", 32, 16, 7, 1, 1088, 1920, True, True, "", "", ""),
    ("spynet_7x7_16_32_first", 16, 32, 7, 1, 1088, 1920, True, True, "", "", ""),
    ("spynet_7x7_16_4_flow_res", 16, 4, 7, 1, 1088, 1920, True, False, "R", "", ""),
    ("adaptor_I_3_48", 3, 48, 3, 1, 1088, 1920, False, True, "", "", ""),
    ("dc_conv2_128_128_r", 128, 128, 1, 1, 272, 480, True, True, "r", "", ""),
    ("dc_conv2_64_48_r", 64, 48, 1, 1, 1088, 1920, True, True, "r", "", ""),
    ("c3_64_128_r", 64, 128, 3, 1, 544, 960, True, True, "r", "", ""),
    ("c1_64_64", 64, 64, 1, 1, 1088, 1920, True, True, "", "", ""),
    ("c3_96_48_actcopy", 96, 48, 3, 1, 1088, 1920, True, True, "", "", "act_copy"),
    ("recon_first_80_48", 80, 48, 3, 1, 1088, 1920, True, True, "", "", ""),
    ("c3_128_192", 128, 192, 3, 1, 544, 960, True, True, "", "", ""),
    ("c1_48_32_r", 48, 32, 1, 1, 1088, 1920, True, True, "r", "", ""),
    ("prior_384_1024", 384, 1024, 1, 1, 68, 120, True, True, "", "", ""),
    ("prior_1024_384_r", 1024, 384, 1, 1, 68, 120, True, True, "r", "", ""),
    ("recon_conv_48_4_clamp", 48, 4, 3, 1, 1088, 1920, False, False, "", "", "clamp4"),
    ("c3_64_64_half", 64, 64, 3, 1, 544, 960, True, True, "", "", ""),
    ("c3_128_128_half", 128, 128, 3, 1, 544, 960, True, True, "", "", ""),
    ("c3_128_64_half", 128, 64, 3, 1, 544, 960, True, True, "", "", ""),
    ("subpel1_64_128_tma_store", 64, 128, 1, 1, 544, 960, True, True, "", "", "shuffle"),
    ("subpel3_128_192_d2s", 128, 192, 3, 1, 544, 960, True, True, "", "", "shuffle"),
    ("subpel3_192_256_d2s_fp32out", 192, 256, 3, 1, 272, 480, True, False, "", "", "shuffle"),
    ("align_offset_56_64_s2", 56, 64, 3, 2, 1088, 1920, True, True, "", "", ""),
    ("ctx_enc_conv1_56_64_s2", 56, 64, 3, 2, 1088, 1920, True, True, "", "", ""),
    ("feat_conv2_48_64_s2", 48, 64, 3, 2, 1088, 1920, True, True, "", "", "act_copy"),
    ("intra_ffn2_gate_128_512", 128, 512, 1, 1, 544, 960, True, False, "", "", "gate"),
    ("rbws_down_1x1_s2", 64, 64, 1, 2, 544, 960, True, True, "", "", ""),
    ("fp32_48_48_tf32_R2", 48, 48, 3, 1, 1088, 1920, False, False, "R", "2", "scale"),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_conv_at_1080p_shape(case):
    name, cin, cout, k, stride, h, w, in16, out16, res, res2, extra = case
    sd = _sd(cin, cout, k, seed=cin * 31 + cout + k)
    pad = k // 2 if stride == 1 or k > 1 else 0
    x = _rand((1, cin, h, w), 1, in16)
    wt = sd["c.weight"].to(dev())
    if in16:
        wt = wt.half().float()
    ref = F.conv2d(x.float(), wt, sd["c.bias"].to(dev()), stride, pad)
    ho, wo = ref.shape[2], ref.shape[3]
    net = layers.Net(sd, dev())
    shuffle = extra == "shuffle"
    cw = net.cw("c", gate=extra == "gate", shuffle=shuffle)
    kw = {}
    slope = 1.0
    if extra == "gate":
        x1, x2 = ref.chunk(2, 1)
        ref = x1 * F.leaky_relu(x2, 0.1)
        kw["gate_slope"] = 0.1
    elif shuffle:
        ref = F.pixel_shuffle(ref, 2)
        kw["shuffle"] = True
    else:
        slope = 0.1
        ref = F.leaky_relu(ref, slope)
    ec, eh, ew = ref.shape[1], ref.shape[2], ref.shape[3]
    if res:
        r1 = _rand((1, ec, eh, ew), 2, res == "r")
        ref = ref + r1.float()
        kw["res"] = _act(r1)
    if res2:
        r2 = _rand((1, ec, eh, ew), 3, res2 == "2h")
        ref = ref + r2.float()
        kw["res2"] = _act(r2)
    if extra == "scale":
        q = (torch.rand(ec, generator=torch.Generator().manual_seed(11)) + 0.5).to(dev())
        ref = ref * q.view(1, -1, 1, 1)
        kw["scale"] = q
    if extra == "clamp4":
        ref = (ref * 3).clamp(0, 1)                  # exercise both clamp sides
        kw["scale"] = torch.full((ec,), 3.0, device=dev())
        kw["clamp01"] = True
    if extra == "act_copy":
        kw["act_copy"] = 0.01
    out = Act.new(eh, ew, ec, dev(), f16=out16)
    if extra == "act_copy":
        got, got2 = ops.conv(_act(x), cw, stride=stride, pad=pad, act=slope, f16_out=out16, **kw)
    else:
        got = ops.conv(_act(x), cw, stride=stride, pad=pad, act=slope, out=out, **kw)
        got2 = None
    torch.cuda.synchronize()
    assert got.f16 == out16, (got.f16, out16)
    K = cin * k * k
    if in16:
        atol = 3e-5 * max(1.0, K ** 0.5 / 8)         # fp32 summation order only
    else:
        atol = 4 * (4e-3 * 0.58 * 0.58 + 1e-5)       # TF32 operands (test_gpu_conv_tc._check)
    if extra in ("scale", "clamp4"):
        atol *= 3
    rtol = 2.0 ** -10 if out16 else 0.0

    def check(t, want, what):
        d = (t.nchw().float() - want).abs()
        lim = atol + rtol * want.abs()
        bad = d > lim
        assert not bool(bad.any()), (name, what, float(d.max()), int(bad.sum()))
        if not in16 and extra != "clamp4":
            assert abs(float((t.nchw().float() - want).mean())) < atol / 40, (name, what, "bias")

    check(got, ref, "out")
    if got2 is not None:
        check(got2, F.leaky_relu(ref, 0.01), "activated copy")


def test_fused_ffn_at_1080p_shapes():
    """dcvc_ffn (two tcgen05 GEMMs, hidden activation on chip) at the P frame's shapes (C = 128:
    the chunked kernel with streamed weights; 67x119 = a ragged last tile) against the two-convolution definition on the same binary16 operands; the hidden activation is
    rounded to binary16 between the GEMMs, which the reference mirrors."""
    for c, h, w, out16 in ((48, 1088, 1920, True), (32, 1088, 1920, False), (64, 544, 960, True),
                           (128, 272, 480, True), (128, 272, 480, False), (128, 67, 119, False)):
        g = torch.Generator().manual_seed(c)
        sd = {"b.conv.0.weight": (torch.rand(4 * c, c, 1, 1, generator=g) * 2 - 1) / c ** 0.5,
              "b.conv.0.bias": torch.rand(4 * c, generator=g) * 0.2 - 0.1,
              "b.conv.2.weight": (torch.rand(c, 4 * c, 1, 1, generator=g) * 2 - 1) / (4 * c) ** 0.5,
              "b.conv.2.bias": torch.rand(c, generator=g) * 0.2 - 0.1}
        net = layers.Net(sd, dev())
        x = _rand((1, c, h, w), 5, True)
        q = (torch.rand(c, generator=g) + 0.5).to(dev())       # seeded: the bounds below scale with it
        xa = _act(x)
        assert ops.ffn_ok(xa)
        w1, b1, w2, b2 = net.ffn("b")
        got = ops.ffn(xa, w1, b1, w2, b2, 0.1, 0.1, scale=q, f16_out=out16)
        torch.cuda.synchronize()
        hid = F.leaky_relu(F.conv2d(x.float(), sd["b.conv.0.weight"].to(dev()).half().float(),
                                    sd["b.conv.0.bias"].to(dev())), 0.1).half().float()
        ref = F.leaky_relu(F.conv2d(hid, sd["b.conv.2.weight"].to(dev()).half().float(),
                                    sd["b.conv.2.bias"].to(dev())), 0.1)
        ref = (x.float() + ref) * q.view(1, -1, 1, 1)
        d = (got.nchw().float() - ref).abs()
        # a hidden value on a binary16 rounding boundary may round the other way (fp32 summation
        # order): one half-ulp of the hidden tensor times |w2| <= 2^-11 * 4 * (1/sqrt(4c)) ...
        lim = 2e-3 + (2.0 ** -10 if out16 else 0.0) * ref.abs()
        assert not bool((d > lim).any()), (c, float(d.max()))
        # mean error: binary16 rounding of the result (2^-12 relative on average) when out16, plus
        # the hidden tensor's rounding; the unseeded scale vector used to put the 48-channel case
        # within 2 % of a fixed 1e-4 bound (2 failures in 40 runs)
        mean_lim = 5e-5 + (2.0 ** -12 if out16 else 0.0) * float(ref.abs().mean())
        assert float(d.mean()) < mean_lim, (c, float(d.mean()), mean_lim)
