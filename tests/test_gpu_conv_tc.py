"""-m gpu: the tcgen05/TMEM/TMA convolution (mode 1, kind::tf32, fp32 accumulate) against the
fp32 oracle convolution.  Tolerance: TF32 keeps 10 mantissa bits per operand, so each product
carries <= 2^-10 relative error; with K-term sums of random-sign terms the error of an output
is ~ 2^-10 * |x||w| * sqrt(K).  The test bounds max|diff| by 4e-3 * sqrt(K) * rms(x) * rms(w)
and the MEAN signed error (bias) by 10x less."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

from dcvc_b200 import _lib, layers, ops
from dcvc_b200.tensor import Act

from gpu_util import dev, from_act, to_act

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _tc_mode():
    ops.set_conv_mode(1)
    yield
    ops.set_conv_mode(0)


def _rand(*shape, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.rand(*shape, generator=g) * 2 - 1


def _sd(cin, cout, k, seed):
    g = torch.Generator().manual_seed(seed)
    b = 1.0 / (cin * k * k) ** 0.5
    return {"c.weight": (torch.rand(cout, cin, k, k, generator=g) * 2 - 1) * b,
            "c.bias": (torch.rand(cout, generator=g) * 2 - 1) * b}


def _check(out, ref, cin, k):
    bound = 4e-3 * (cin * k * k) ** 0.5 * 0.58 * 0.58 / (cin * k * k) ** 0.5 + 1e-5
    d = out - ref
    assert float(d.abs().max()) < 4 * bound, float(d.abs().max())
    assert abs(float(d.mean())) < bound / 10, float(d.mean())


SHAPES = [
    # cin, cout, k, stride, h, w
    (48, 48, 3, 1, 48, 64), (64, 64, 3, 1, 32, 48), (32, 64, 7, 1, 40, 40), (8, 32, 7, 1, 40, 56),
    (128, 192, 3, 1, 24, 40), (64, 256, 1, 1, 32, 32), (512, 128, 1, 1, 24, 24),
    (16, 2, 7, 1, 32, 32), (48, 3, 3, 1, 32, 48), (80, 48, 3, 1, 30, 50), (96, 96, 3, 1, 17, 67),
    (3, 48, 3, 1, 32, 32), (384, 384, 1, 1, 20, 28), (20, 24, 3, 1, 33, 47),
    (128, 512, 1, 1, 24, 32), (384, 1024, 1, 1, 17, 31), (64, 320, 3, 1, 24, 24),
    (32, 128, 3, 1, 40, 40),    # halo window mode with one channel chunk and streamed weights
]


@pytest.mark.parametrize("cin,cout,k,stride,h,w", SHAPES)
def test_tc_conv_stride1(cin, cout, k, stride, h, w):
    sd = _sd(cin, cout, k, seed=cin * 7 + cout)
    x = _rand(1, cin, h, w, seed=1)
    pad = k // 2
    ref = F.conv2d(x, sd["c.weight"], sd["c.bias"], stride, pad)
    net = layers.Net(sd, dev())
    xa = to_act(x)
    cw = net.cw("c")
    d = _lib.Conv()
    out = ops.conv(xa, cw, stride=stride, pad=pad)
    torch.cuda.synchronize()
    _check(from_act(out), ref, cin, k)


@pytest.mark.parametrize("cin,cout,k,h,w", [(53, 64, 3, 64, 96), (128, 96, 3, 48, 48),
                                            (2, 64, 3, 64, 64), (64, 64, 1, 40, 56),
                                            (3, 128, 3, 64, 64)])
def test_tc_conv_stride2(cin, cout, k, h, w):
    sd = _sd(cin, cout, k, seed=cin + cout)
    x = _rand(1, cin, h, w, seed=2)
    pad = k // 2
    ref = F.conv2d(x, sd["c.weight"], sd["c.bias"], 2, pad)
    net = layers.Net(sd, dev())
    out = ops.conv(to_act(x), net.cw("c"), stride=2, pad=pad)
    torch.cuda.synchronize()
    _check(from_act(out), ref, cin, k)


def test_tc_conv_epilogues_and_slices():
    cin, cout, h, w = 32, 64, 40, 48
    sd = _sd(cin, cout, 3, seed=5)
    x, r1, r2 = _rand(1, cin, h, w, seed=1), _rand(1, cout, h, w, seed=2), _rand(1, cout, h, w, seed=3)
    q = torch.rand(cout, generator=torch.Generator().manual_seed(11)) + 0.5
    conv = F.conv2d(x, sd["c.weight"], sd["c.bias"], 1, 1)
    net = layers.Net(sd, dev())
    xa = to_act(x, cstride=48, coff=8)
    # lrelu + two residuals + scale + second output, into a channel slice
    out = to_act(torch.zeros(1, cout, h, w), cstride=80, coff=12)
    out2 = to_act(torch.zeros(1, cout, h, w))
    ops.conv(xa, net.cw("c"), act=0.01, res=to_act(r1), res2=to_act(r2, 68, 4), scale=q.to(dev()),
             out=out, out2=out2, out2_slope=0.2)
    ref = (F.leaky_relu(conv, 0.01) + r1 + r2) * q.view(1, -1, 1, 1)
    _check(from_act(out), ref, cin, 3)
    _check(from_act(out2), F.leaky_relu(ref, 0.2), cin, 3)
    # depth-to-space
    out = ops.conv(xa, net.cw("c"), shuffle=True, act=0.01)
    _check(from_act(out), F.leaky_relu(F.pixel_shuffle(conv, 2), 0.01), cin, 3)
    # gate
    x1, x2 = conv.chunk(2, 1)
    out = ops.conv(xa, net.cw("c", gate=True), gate_slope=0.1)
    _check(from_act(out), x1 * F.leaky_relu(x2, 0.1), cin, 3)
    # clamp + input prologue (materialised copy)
    out = ops.conv(xa, net.cw("c"), in_slope=0.1, clamp01=True)
    _check(from_act(out), F.conv2d(F.leaky_relu(x, 0.1), sd["c.weight"], sd["c.bias"], 1, 1).clamp(0, 1), cin, 3)


def test_tc_conv_n_tiled_gate_and_shuffle():
    """cout > 256 is split over grid.y; gate pairs / shuffle quads never straddle a tile."""
    cin, cout, h, w = 64, 768, 24, 32
    sd = _sd(cin, cout, 1, seed=11)
    x = _rand(1, cin, h, w, seed=4)
    conv = F.conv2d(x, sd["c.weight"], sd["c.bias"])
    net = layers.Net(sd, dev())
    xa = to_act(x)
    x1, x2 = conv.chunk(2, 1)
    _check(from_act(ops.conv(xa, net.cw("c", gate=True), pad=0, gate_slope=0.1)),
           x1 * F.leaky_relu(x2, 0.1), cin, 1)
    _check(from_act(ops.conv(xa, net.cw("c"), pad=0, shuffle=True)), F.pixel_shuffle(conv, 2), cin, 1)


def test_tc_path_is_taken_and_deterministic():
    sd = _sd(64, 64, 3, seed=1)
    net = layers.Net(sd, dev())
    xa = to_act(_rand(1, 64, 64, 64, seed=3))
    cw = net.cw("c")
    out = ops.Act.new(64, 64, 64, dev())
    d = _lib.Conv()
    d.inp, d.out = xa.view(), out.view()
    d.weight, d.weight_tc, d.bias = cw.w.data_ptr(), cw.w_tc.data_ptr(), cw.b.data_ptr()
    d.ksize, d.stride, d.pad, d.cout_w = 3, 1, 1, 64
    d.in_slope = d.out_slope = d.gate_slope = d.out2_slope = 1.0
    d.mode = 1
    assert _lib.lib().dcvc_conv_tc_eligible(ctypes.byref(d)) == 1
    a = from_act(ops.conv(xa, cw))
    b = from_act(ops.conv(xa, cw))
    assert torch.equal(a, b)            # run-to-run bit-identical (encoder/decoder agreement)


@pytest.mark.parametrize("cin,cout,k,h,w", [(32, 64, 3, 40, 48), (64, 768, 1, 24, 32),
                                            (48, 128, 3, 37, 53), (64, 256, 1, 68, 120)])
def test_tc_conv_planar_shuffle(cin, cout, k, h, w):
    """Depth-to-space with weight_tc packed (subpixel, channel) (dcvc_conv.shuffle = 2, the lean
    epilogue): same result as the reference-order packing, with residual / scale / activation in
    the shuffled output space, into a channel slice, ragged tiles included."""
    sd = _sd(cin, cout, k, seed=21)
    x = _rand(1, cin, h, w, seed=6)
    c4 = cout // 4
    r1 = _rand(1, c4, 2 * h, 2 * w, seed=7)
    q = torch.rand(c4, generator=torch.Generator().manual_seed(11)) + 0.5
    conv = F.conv2d(x, sd["c.weight"], sd["c.bias"], 1, k // 2)
    net = layers.Net(sd, dev())
    xa = to_act(x)
    cw = net.cw("c", shuffle=True)
    assert cw.tc_planar
    out = to_act(torch.zeros(1, c4, 2 * h, 2 * w), cstride=c4 + 8, coff=4)
    ops.conv(xa, cw, shuffle=True, act=0.01, res=to_act(r1), scale=q.to(dev()), out=out)
    ref = (F.leaky_relu(F.pixel_shuffle(conv, 2), 0.01) + r1) * q.view(1, -1, 1, 1)
    _check(from_act(out), ref, cin, k)
    # and it agrees with the generic (reference-order) epilogue
    out_ref = ops.conv(xa, net.cw("c"), shuffle=True)
    out_pl = ops.conv(xa, cw, shuffle=True)
    assert torch.equal(from_act(out_ref), from_act(out_pl))


@pytest.mark.parametrize("cin,cout,k,h,w,in16", [(32, 128, 3, 40, 48, False), (64, 128, 1, 37, 53, True),
                                                 (64, 256, 1, 24, 32, True)])
def test_tc_conv_planar_shuffle_f16_output(cin, cout, k, h, w, in16):
    """Depth-to-space into a binary16 channel slice (the contextual decoder's last upsampling
    writes the binary16 recon buffer): the fp32 result rounded once, neighbours untouched."""
    sd = _sd(cin, cout, k, seed=23)
    x = _rand(1, cin, h, w, seed=9)
    c4 = cout // 4
    net = layers.Net(sd, dev())
    cw = net.cw("c", shuffle=True)
    xa = to_act(x)
    if in16:
        xa = ops.copy(xa, f16=True)
    ref32 = ops.conv(xa, cw, shuffle=True, act=0.01)
    wide = torch.full((2 * h, 2 * w, c4 + 24), 7.0, device=dev(), dtype=torch.float16)
    out = Act(wide, 8, c4)
    ops.conv(xa, cw, shuffle=True, act=0.01, out=out)
    torch.cuda.synchronize()
    assert torch.equal(wide[:, :, 8:8 + c4], ref32.buf[:, :, :c4].half())
    assert bool((wide[:, :, :8] == 7.0).all()) and bool((wide[:, :, 8 + c4:] == 7.0).all())
    with pytest.raises(_lib.DcvcError):       # no residual on this path
        ops.conv(xa, cw, shuffle=True, res=Act.new(2 * h, 2 * w, c4, dev()), out=out)


@pytest.mark.parametrize("mode", [0, 1])
def test_pool2_f16(mode):
    c, h, w = 32, 22, 38
    x = _rand(1, c, h, w, seed=17)
    wide = ops.copy(to_act(torch.cat((x, x), 1)), f16=True)          # 64 channels, pool a slice
    out = Act(torch.full((h // 2, w // 2, 48), 7.0, device=dev(), dtype=torch.float16), 8, c)
    ops.pool2(Act(wide.buf, 32, c), mode, out=out)
    torch.cuda.synchronize()
    xh = x.half().float()
    ref = (F.avg_pool2d(xh, 2) if mode == 0 else F.max_pool2d(xh, 2)).half()
    assert torch.equal(out.buf[:, :, 8:8 + c].cpu(), ref[0].permute(1, 2, 0))
    assert bool((out.buf[:, :, :8] == 7.0).all()) and bool((out.buf[:, :, 8 + c:] == 7.0).all())


@pytest.mark.parametrize("cin,cout,k,stride,h,w,f16", [(48, 48, 3, 1, 40, 56, True), (48, 64, 3, 2, 64, 96, True),
                                                        (96, 48, 3, 1, 33, 47, False), (64, 96, 1, 1, 37, 53, True)])
def test_tc_conv_activated_copy(cin, cout, k, stride, h, w, f16):
    """act_copy: the TMA-store epilogue writes out and lrelu(out) (the following ResBlock's
    activated input) from its two tile buffers; out equals the plain convolution's bits."""
    sd = _sd(cin, cout, k, seed=61)
    x = _rand(1, cin, h, w, seed=19)
    net = layers.Net(sd, dev())
    xa = ops.copy(to_act(x), f16=True) if f16 else to_act(x)
    plain = ops.conv(xa, net.cw("c"), stride=stride, act=0.1, f16_out=f16)
    out, act = ops.conv(xa, net.cw("c"), stride=stride, act=0.1, f16_out=f16, act_copy=0.01)
    torch.cuda.synchronize()
    assert out.f16 == plain.f16 == act.f16 == f16
    assert torch.equal(out.buf, plain.buf)
    o32 = out.buf[:, :, :cout].float()
    want = torch.where(o32 >= 0, o32, o32 * 0.01)
    got = act.buf[:, :, :cout].float()
    # binary16: lrelu is applied before the one rounding (the copy pass rounds twice)
    tol = (want.abs() * 2 ** -10 + 1e-7) if f16 else torch.zeros_like(want)
    assert bool(((got - want).abs() <= tol).all())


def test_copy_f16_narrow():
    x = _rand(1, 3, 21, 35, seed=5)
    wide = torch.full((21, 35, 56), 7.0, device=dev(), dtype=torch.float16)
    ops.copy(to_act(x), out=Act(wide, 48, 3))
    torch.cuda.synchronize()
    assert torch.equal(wide[:, :, 48:51].cpu(), x[0].permute(1, 2, 0).half())
    assert bool((wide[:, :, 51:] == 0).all()) and bool((wide[:, :, :48] == 7.0).all())


@pytest.mark.parametrize("cin,cout,k,h,w", [(64, 48, 1, 40, 48), (48, 192, 1, 37, 53),
                                            (32, 64, 3, 33, 70), (64, 768, 1, 24, 32),
                                            (16, 40, 3, 24, 200), (128, 128, 1, 68, 120)])
def test_tc_conv_tma_store_epilogue(cin, cout, k, h, w):
    """Plain convs leave through the tensor-map store epilogue: bias + LeakyReLU + one residual
    (fetched by TMA into the outgoing tile) + per-channel scale + clamp, into a channel slice of
    a wider buffer whose neighbours must stay untouched; ragged tiles and N-tiled outputs."""
    sd = _sd(cin, cout, k, seed=31)
    x, r1 = _rand(1, cin, h, w, seed=8), _rand(1, cout, h, w, seed=9)
    q = torch.rand(cout, generator=torch.Generator().manual_seed(11)) + 0.5
    conv = F.conv2d(x, sd["c.weight"], sd["c.bias"], 1, k // 2)
    net = layers.Net(sd, dev())
    xa = to_act(x)
    wide = torch.full((h, w, cout + 16), 7.0, device=dev())
    out = Act(wide, 8, cout)
    ops.conv(xa, net.cw("c"), act=0.1, res=to_act(r1, cout + 4, 4), scale=q.to(dev()), out=out)
    ref = (F.leaky_relu(conv, 0.1) + r1) * q.view(1, -1, 1, 1)
    _check(from_act(out), ref, cin, k)
    assert bool((wide[:, :, :8] == 7.0).all()) and bool((wide[:, :, 8 + cout:] == 7.0).all())
    # no residual, clamp
    out = ops.conv(xa, net.cw("c"), clamp01=True)
    _check(from_act(out), conv.clamp(0, 1), cin, k)


F16_SHAPES = [
    # cin, cout, k, stride, h, w     (binary16 K chunks: 64 -> 128 B rows, 32 -> 64 B, 16 -> 32 B)
    (64, 64, 3, 1, 32, 48), (32, 128, 3, 1, 40, 40), (48, 48, 3, 1, 48, 64), (8, 32, 7, 1, 40, 56),
    (128, 192, 3, 1, 24, 40), (192, 48, 1, 1, 37, 53), (512, 128, 1, 1, 24, 24),
    (32, 64, 7, 1, 40, 40), (16, 2, 7, 1, 32, 32), (64, 64, 3, 2, 64, 96), (256, 64, 1, 1, 33, 70),
]


@pytest.mark.parametrize("cin,cout,k,stride,h,w", F16_SHAPES)
def test_tc_conv_f16_input(cin, cout, k, stride, h, w):
    """binary16 activations + binary16 weights (kind::f16, fp32 accumulate) against the fp32
    convolution of the same rounded operands: only the summation order differs, so the bound is
    fp32 accumulation noise, not an operand-precision bound.  Residual / scale epilogue rides
    along (fp32 views)."""
    sd = _sd(cin, cout, k, seed=cin * 3 + cout)
    x, r1 = _rand(1, cin, h, w, seed=11), None
    pad = k // 2
    xa = ops.copy(to_act(x, cstride=(cin + 7) // 8 * 8), f16=True)
    assert xa.f16
    w16 = sd["c.weight"].half().float()
    ref = F.conv2d(x.half().float(), w16, sd["c.bias"], stride, pad)
    ho, wo = ref.shape[2], ref.shape[3]
    r1 = _rand(1, cout, ho, wo, seed=12)
    q = torch.rand(cout, generator=torch.Generator().manual_seed(11)) + 0.5
    net = layers.Net(sd, dev())
    out = ops.conv(xa, net.cw("c"), stride=stride, pad=pad, act=0.1,
                   res=to_act(r1) if cout % 4 == 0 else None,
                   scale=q.to(dev()) if cout % 4 == 0 else None)
    assert not out.f16
    want = F.leaky_relu(ref, 0.1)
    if cout % 4 == 0:
        want = (want + r1) * q.view(1, -1, 1, 1)
    d = from_act(out) - want
    assert float(d.abs().max()) < 2e-5 * max(1.0, (cin * k * k) ** 0.5 / 8), float(d.abs().max())


@pytest.mark.parametrize("cin,cout,k,stride,h,w", [(48, 192, 1, 1, 37, 53), (32, 64, 3, 1, 33, 70),
                                                  (64, 48, 3, 2, 64, 96), (8, 32, 7, 1, 40, 56),
                                                  (64, 768, 1, 1, 24, 32), (48, 40, 3, 1, 24, 200)])
def test_tc_conv_f16_output_and_chain(cin, cout, k, stride, h, w):
    """binary16 output through the tensor-map store (64-byte swizzled rows): equals the fp32
    result rounded to binary16 up to the TF32 bound, leaves neighbouring channels of a wider
    buffer untouched, and feeds a second convolution as a binary16 input."""
    sd = _sd(cin, cout, k, seed=41)
    x = _rand(1, cin, h, w, seed=13)
    pad = k // 2
    conv = F.leaky_relu(F.conv2d(x, sd["c.weight"], sd["c.bias"], stride, pad), 0.1)
    ho, wo = conv.shape[2], conv.shape[3]
    net = layers.Net(sd, dev())
    wide = torch.full((ho, wo, cout + 16), 7.0, device=dev(), dtype=torch.float16)
    out = Act(wide, 8, cout)
    ops.conv(to_act(x), net.cw("c"), stride=stride, pad=pad, act=0.1, out=out)
    got = from_act(Act(wide.float(), 8, cout))
    _check(got, conv, cin, k)
    assert float((got - conv).abs().max()) < 4 * (4e-3 * 0.58 * 0.58 + 1e-5) + float(conv.abs().max()) * 2 ** -11
    assert bool((wide[:, :, :8] == 7.0).all()) and bool((wide[:, :, 8 + cout:] == 7.0).all())
    # requested through the layer-level switch: granted only when f16_ok() holds
    t = ops.conv(to_act(x), net.cw("c"), stride=stride, pad=pad, act=0.1, f16_out=True)
    assert t.f16 == (ops.F16 and ho * wo >= 512 and cout % 8 == 0)
    if not t.f16:
        return
    sd2 = {"c.weight": _sd(cout, 32, 3, seed=43)["c.weight"], "c.bias": _sd(cout, 32, 3, seed=43)["c.bias"]}
    net2 = layers.Net(sd2, dev())
    y = ops.conv(t, net2.cw("c"))
    t32 = t.buf[:, :, :cout].float().permute(2, 0, 1).unsqueeze(0).cpu()
    ref2 = F.conv2d(t32, sd2["c.weight"].half().float(), sd2["c.bias"], 1, 1)
    assert float((from_act(y) - ref2).abs().max()) < 2e-5 * max(1.0, (cout * 9) ** 0.5 / 8)


def test_f16_views_never_fall_back():
    """A binary16 view on a descriptor the tensor-core path cannot take is an error, not a
    silent fp32 reinterpretation."""
    sd = _sd(16, 16, 3, seed=2)
    net = layers.Net(sd, dev())
    small = Act.new(8, 8, 16, dev(), f16=True)            # 64 pixels: below the tensor-core floor
    with pytest.raises(_lib.DcvcError):
        ops.conv(small, net.cw("c"))


def test_dwconv_f16():
    c, h, w = 72, 21, 35
    g = torch.Generator().manual_seed(3)
    sd = {"d.weight": torch.rand(c, 1, 3, 3, generator=g) - 0.5, "d.bias": torch.rand(c, generator=g) - 0.5}
    x = _rand(1, c, h, w, seed=8)
    net = layers.Net(sd, dev())
    wt, b = net.dw("d")
    x16 = ops.copy(to_act(x), f16=True)
    out = ops.dwconv3x3(x16, wt, b)
    assert out.f16
    ref = F.conv2d(x.half().float(), sd["d.weight"], sd["d.bias"], 1, 1, 1, c)
    got = out.buf[:, :, :c].float().permute(2, 0, 1).unsqueeze(0).cpu()
    assert float((got - ref.half().float()).abs().max()) <= float(ref.abs().max()) * 2 ** -10


@pytest.mark.parametrize("block", ["dcb", "res_block", "rbws"])
def test_blocks_with_f16_intermediates(block, monkeypatch):
    """The reference's blocks with their private intermediates stored as binary16 (the default in
    tensor-core mode) against the fp32 oracle and against the all-fp32-storage run: storage
    rounding of an intermediate is the same 11-bit significand the TF32 MMA reads, so both
    comparisons share one bound."""
    from dcvc_b200 import arch
    from oracle import nets_oracle as no
    spec = arch.Spec()
    if block == "dcb":
        spec.dcb("b", 48, 48)
    elif block == "res_block":
        spec.res_block("b", 48)
    else:
        spec.rbws("b", 48, 64)
    sd = arch.synth_state_dict(spec, 7)
    x = _rand(1, 48, 40, 56, seed=5)
    ref = getattr(no, block)(sd, "b", x)
    fn = getattr(layers, block)
    monkeypatch.setattr(ops, "F16", True)
    out16 = from_act(fn(layers.Net(sd, dev()), "b", to_act(x)))
    monkeypatch.setattr(ops, "F16", False)
    out32 = from_act(fn(layers.Net(sd, dev()), "b", to_act(x)))
    scale = float(ref.abs().max())
    assert float((out16 - ref).abs().max()) < 6e-3 * scale
    assert float((out32 - ref).abs().max()) < 6e-3 * scale
    assert float((out16 - out32).abs().max()) < 6e-3 * scale
    assert not torch.equal(out16, out32)          # the binary16 path really ran


@pytest.mark.parametrize("res16,out16", [(True, False), (False, True), (True, True)])
@pytest.mark.parametrize("cin,cout,k,h,w", [(64, 48, 1, 37, 53), (32, 64, 3, 33, 70), (64, 320, 1, 24, 32)])
def test_tc_conv_f16_residual(cin, cout, k, h, w, res16, out16):
    """Residual and result of either element size through the one smem tile of the TMA-store
    epilogue (rows of 64 and 128 bytes overlap, so the warp reads before anyone writes)."""
    sd = _sd(cin, cout, k, seed=51)
    x, r1 = _rand(1, cin, h, w, seed=14), _rand(1, cout, h, w, seed=15)
    q = torch.rand(cout, generator=torch.Generator().manual_seed(11)) + 0.5
    net = layers.Net(sd, dev())
    res = ops.copy(to_act(r1), f16=True) if res16 else to_act(r1)
    rr = r1.half().float() if res16 else r1
    ref = (F.leaky_relu(F.conv2d(x, sd["c.weight"], sd["c.bias"], 1, k // 2), 0.1) + rr) * q.view(1, -1, 1, 1)
    out = Act.new(h, w, cout, dev(), f16=out16)
    ops.conv(to_act(x), net.cw("c"), act=0.1, res=res, scale=q.to(dev()), out=out)
    got = out.buf[:, :, :cout].float().permute(2, 0, 1).unsqueeze(0).cpu()
    bound = 4 * (4e-3 * 0.58 * 0.58 + 1e-5) * 1.5 + (float(ref.abs().max()) * 2 ** -11 if out16 else 0.0)
    assert float((got - ref).abs().max()) < bound, float((got - ref).abs().max())


@pytest.mark.parametrize("cin,cout,h,w,f16", [(48, 48, 40, 56, True), (48, 48, 37, 50, False),
                                               (64, 96, 33, 47, True), (96, 48, 24, 40, False)])
def test_tc_conv_two_residuals_tma_epilogue(cin, cout, h, w, f16):
    """res_block(..., res2=...) of the context fusion net: conv + res1 + res2 in the TMA-store
    epilogue (second residual by direct row loads), ragged tiles, channel-slice views."""
    sd = _sd(cin, cout, 3, seed=21)
    x, r1, r2 = _rand(1, cin, h, w, seed=1), _rand(1, cout, h, w, seed=2), _rand(1, cout, h, w, seed=3)
    net = layers.Net(sd, dev())
    xa = to_act(x)
    xr = x
    if f16:
        xa = ops.copy(xa, f16=True)
        xr = x.half().float()
    wr = sd["c.weight"].half().float() if f16 else sd["c.weight"]
    conv = F.conv2d(xr, wr, sd["c.bias"], 1, 1)
    out = to_act(torch.zeros(1, cout, h, w), cstride=cout + 16, coff=8)
    ops.conv(xa, net.cw("c"), act=0.1, res=to_act(r1), res2=to_act(r2, cout + 20, 4), out=out)
    torch.cuda.synchronize()
    _check(from_act(out), F.leaky_relu(conv, 0.1) + r1 + r2, cin, 3)


@pytest.mark.parametrize("c,h,w,out_f16,scaled", [(48, 40, 56, False, True), (32, 33, 47, True, False),
                                                  (64, 64, 96, False, False), (48, 128, 200, True, True)])
def test_fused_ffn_matches_reference(c, h, w, out_f16, scaled):
    """dcvc_ffn (ConvFFN of DepthConvBlock, layers.py:166-179 + the block's residual) against
    torch fp32 on the binary16-rounded operands the kernel multiplies; the hidden activation is
    rounded to binary16 like the unfused path's private intermediate."""
    g = torch.Generator().manual_seed(c + h)
    sd = {"f.conv.0.weight": torch.randn(4 * c, c, 1, 1, generator=g) / c ** 0.5,
          "f.conv.0.bias": torch.randn(4 * c, generator=g) * 0.1,
          "f.conv.2.weight": torch.randn(c, 4 * c, 1, 1, generator=g) / (4 * c) ** 0.5,
          "f.conv.2.bias": torch.randn(c, generator=g) * 0.1}
    x = _rand(1, c, h, w, seed=7)
    q = (torch.rand(c, generator=g) + 0.5) if scaled else None
    net = layers.Net(sd, dev())
    xa = ops.copy(to_act(x), f16=True)
    old = ops.CONV_MODE
    ops.set_conv_mode(1)
    try:
        assert ops.ffn_ok(xa)
        w1, b1, w2, b2 = net.ffn("f")
        out = None
        if not out_f16:
            out = to_act(torch.zeros(1, c, h, w), cstride=c + 16, coff=8)
        out = ops.ffn(xa, w1, b1, w2, b2, 0.1, 0.1, scale=q.to(dev()) if scaled else None, out=out,
                      f16_out=out_f16)
        torch.cuda.synchronize()
        assert out.f16 == out_f16
    finally:
        ops.set_conv_mode(old)
    xh = x.half().float()
    hid = F.leaky_relu(F.conv2d(xh, sd["f.conv.0.weight"].half().float(), sd["f.conv.0.bias"]), 0.1)
    hid = hid.half().float()
    ref = xh + F.leaky_relu(F.conv2d(hid, sd["f.conv.2.weight"].half().float(), sd["f.conv.2.bias"]), 0.1)
    if scaled:
        ref = ref * q.view(1, -1, 1, 1)
    got = from_act(out)
    tol = 2e-3 if out_f16 else 2e-4
    err = float((got - ref).abs().max() / ref.abs().max())
    assert err < tol, err
