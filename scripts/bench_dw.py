"""Micro-benchmark of the binary16 depthwise 3x3 kernel and of offset_diversity at the 1080p
shapes.  usage: python scripts/bench_dw.py [ncu]   (`ncu`: one launch of each, no timing loop)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dcvc_b200 import ops
from dcvc_b200.tensor import Act

ops.set_conv_mode(1)
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
once = len(sys.argv) > 1 and sys.argv[1] == "ncu"


def timeit(fn, iters=10):
    fn()
    torch.cuda.synchronize()
    if once:
        return float("nan")
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


for c, h, w in ((48, 1088, 1920), (64, 1088, 1920), (32, 1088, 1920), (128, 544, 960)):
    x = ops.copy(Act(torch.randn(h, w, c, device=dev), 0, c), f16=True)
    wt = torch.randn(9, c, device=dev)
    b = torch.randn(c, device=dev)
    out = Act.new(h, w, c, dev, f16=True)
    t = timeit(lambda: ops.dwconv3x3(x, wt, b, out=out))
    print(f"dwconv3x3 f16 {c}ch @{h}x{w}: {t:.3f} ms ({h * w * c * 4 / t / 1e6:.0f} GB/s algorithmic)")

h, w = 1088, 1920
x = Act(torch.randn(h, w, 48, device=dev), 0, 48)
offs = Act(torch.randn(h // 2, w // 2, 96, device=dev) * 0.2, 0, 96)
flow = Act(torch.randn(h, w, 2, device=dev) * 3, 0, 2)
fw, fb = torch.randn(48 * 6, device=dev), torch.randn(48, device=dev)
out = Act.new(h, w, 48, dev, f16=True)
t = timeit(lambda: ops.offset_diversity(x, offs, flow, fw, fb, out=out))
print(f"offset_diversity f16 @{h}x{w}: {t:.3f} ms")
