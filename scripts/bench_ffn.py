"""Micro-benchmark of the fused ConvFFN kernel (dcvc_ffn) against the two 1x1 convolutions it
replaces, at the 1080p shapes of the DepthConvBlocks.  usage: python scripts/bench_ffn.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dcvc_b200 import layers, ops
from dcvc_b200.tensor import Act

ops.set_conv_mode(1)
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, iters=10):
    fn()
    torch.cuda.synchronize()
    ms = 0.0
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms += e0.elapsed_time(e1)
    return ms / iters


for c, h, w in ((48, 1088, 1920), (32, 1088, 1920), (64, 544, 960), (128, 272, 480)):
    g = torch.Generator().manual_seed(0)
    sd = {"f.conv.0.weight": torch.randn(4 * c, c, 1, 1, generator=g) * 0.05,
          "f.conv.0.bias": torch.randn(4 * c, generator=g),
          "f.conv.2.weight": torch.randn(c, 4 * c, 1, 1, generator=g) * 0.05,
          "f.conv.2.bias": torch.randn(c, generator=g)}
    net = layers.Net(sd, dev)
    x = ops.copy(Act(torch.randn(h, w, c, device=dev), 0, c), f16=True)
    for f16_out in (True, False):
        out = Act.new(h, w, c, dev, f16=f16_out)
        w1, b1, w2, b2 = net.ffn("f")
        t_f = timeit(lambda: ops.ffn(x, w1, b1, w2, b2, out=out))

        def unfused():
            u = ops.conv(x, net.cw("f.conv.0"), act=0.1, f16_out=True)
            ops.conv(u, net.cw("f.conv.2"), act=0.1, res=x, out=out)
        t_u = timeit(unfused)
        ref = out.buf.float().clone()
        out.buf.zero_()
        ops.ffn(x, w1, b1, w2, b2, out=out)
        torch.cuda.synchronize()
        d = (out.buf.float() - ref).abs()
        print(f"   fused vs unfused: max |d| {float(d.max()):.3e}  mean {float(d.mean()):.3e}  "
              f"(ref rms {float(ref.pow(2).mean().sqrt()):.3f})")
        nbytes = h * w * c * (2 + (2 if f16_out else 4))
        print(f"ffn {c}->{4*c}->{c} @{h}x{w} out {'f16' if f16_out else 'f32'}: fused {t_f:.3f} ms "
              f"({nbytes / t_f / 1e6:.0f} GB/s algorithmic), unfused {t_u:.3f} ms")
