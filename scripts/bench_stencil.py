"""Micro-benchmark of the non-GEMM kernels (depthwise conv, copy, pool, warp, offset diversity)
at P-frame shapes.  CUDA-event timing, L2 flushed between iterations.
usage: python scripts/bench_stencil.py [filter]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dcvc_b200 import ops
from dcvc_b200.tensor import Act

flt = sys.argv[1] if len(sys.argv) > 1 else ""
dev = torch.device("cuda:0")
ops.set_conv_mode(1)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(name, fn, nbytes, iters=10):
    if flt and flt not in name:
        return
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
    ms /= iters
    print(f"{name:50s} {ms:8.4f} ms {nbytes / ms / 1e6:8.0f} GB/s", flush=True)


for h, w, c in [(1088, 1920, 48), (1088, 1920, 32), (544, 960, 64), (272, 480, 128), (68, 120, 384)]:
    wt = torch.randn(9, c, device=dev)
    b = torch.randn(c, device=dev)
    for f16 in (False, True):
        x = Act.new(h, w, c, dev, f16=f16)
        x.buf.normal_()
        out = Act.new(h, w, c, dev, f16=f16)
        timeit(f"dwconv {h}x{w}x{c} {'f16' if f16 else 'f32'}", lambda: ops.dwconv3x3(x, wt, b, out=out),
               (2 if f16 else 4) * 2 * h * w * c)
    x = Act.new(h, w, c, dev)
    x.buf.normal_()
    o = Act.new(h, w, c, dev)
    timeit(f"copy {h}x{w}x{c} f32", lambda: ops.copy(x, out=o, slope=0.1), 8 * h * w * c)
    o16 = Act.new(h, w, c, dev, f16=True)
    timeit(f"copy {h}x{w}x{c} ->f16", lambda: ops.copy(x, out=o16, slope=0.1), 6 * h * w * c)
    p = Act.new(h // 2, w // 2, c, dev)
    timeit(f"max_pool2 {h}x{w}x{c}", lambda: ops.max_pool2(x, out=p), 5 * h * w * c)

h, w = 1088, 1920
x = Act.new(h, w, 48, dev); x.buf.normal_()
offs = Act.new(h // 2, w // 2, 96, dev); offs.buf.normal_()
flow = Act.new(h, w, 2, dev); flow.buf.normal_()
fw, fb = torch.randn(48, 6, device=dev), torch.randn(48, device=dev)
o = Act.new(h, w, 48, dev)
timeit("offset_diversity 1088x1920", lambda: ops.offset_diversity(x, offs, flow, fw, fb, out=o),
       4 * (2 * h * w * 48 + h * w * 24 + h * w * 2))
