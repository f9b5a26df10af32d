"""Micro-benchmark of the convolution kernels at the 1080p shapes that dominate a P frame
(SURVEY.md appendix A).  CUDA-event timing, L2 flushed between iterations.
usage: python scripts/bench_conv.py [mode] [filter-substring]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dcvc_b200 import layers, ops
from dcvc_b200.tensor import Act

mode = int(sys.argv[1]) if len(sys.argv) > 1 else 1
flt = sys.argv[2] if len(sys.argv) > 2 else ""
iters = int(os.environ.get("ITERS", "10"))
ops.set_conv_mode(mode)
dev = torch.device("cuda:0")
SHAPES = [  # cin, cout, k, stride, h, w, kwargs
    (48, 48, 3, 1, 1088, 1920, {}), (64, 64, 3, 1, 544, 960, {}), (64, 32, 7, 1, 1088, 1920, {}),
    (32, 64, 7, 1, 1088, 1920, {}), (8, 32, 7, 1, 1088, 1920, {}), (128, 192, 3, 1, 544, 960, {}),
    (48, 192, 1, 1, 1088, 1920, {}), (192, 48, 1, 1, 1088, 1920, {}), (128, 512, 1, 1, 272, 480, {}),
    (512, 128, 1, 1, 272, 480, {}), (64, 256, 1, 1, 544, 960, {}), (96, 96, 3, 1, 272, 480, {}),
    (384, 384, 1, 1, 68, 120, {}), (384, 1024, 1, 1, 68, 120, {}), (1024, 384, 1, 1, 68, 120, {}),
    (128, 128, 3, 1, 544, 960, {"shuffle": True}), (53, 64, 3, 2, 1088, 1920, {}),
    (64, 48, 1, 1, 1088, 1920, {}), (48, 32, 1, 1, 1088, 1920, {}), (48, 48, 1, 1, 1088, 1920, {}),
]
if os.environ.get("SHAPES"):   # "cin,cout,k,stride,h,w;..."
    SHAPES = [tuple(int(v) for v in t.split(",")) + ({},) for t in os.environ["SHAPES"].split(";")]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for cin, cout, k, s, h, w, kw in SHAPES:
    name = f"conv{k}x{k}s{s} {cin}->{cout} @{h}x{w}" + (" shuffle" if kw.get("shuffle") else "")
    if flt and flt not in name:
        continue
    g = torch.Generator().manual_seed(0)
    sd = {"c.weight": torch.randn(cout, cin, k, k, generator=g) * 0.05, "c.bias": torch.randn(cout, generator=g)}
    net = layers.Net(sd, dev)
    x = Act(torch.randn(h, w, (cin + 3) // 4 * 4, device=dev), 0, cin)
    f16 = os.environ.get("BENCH_F16", "")          # "in", "out" or "io"
    if "i" in f16 and cin % 8 == 0 and not kw.get("shuffle"):
        x = ops.copy(x, f16=True)
    if "o" in f16 and cout % 8 == 0 and not kw.get("shuffle"):
        kw = dict(kw, f16_out=True)
    cw = net.cw("c", shuffle=bool(kw.get("shuffle")))
    if os.environ.get("BENCH_RES") and not kw.get("shuffle"):      # fp32 residual (res_block conv2)
        ho_, wo_ = (h + 2 * (k // 2) - k) // s + 1, (w + 2 * (k // 2) - k) // s + 1
        kw = dict(kw, res=Act(torch.randn(ho_, wo_, (cout + 3) // 4 * 4, device=dev), 0, cout))
        if os.environ.get("BENCH_RES") == "h":                      # binary16 residual
            kw["res"] = ops.copy(Act(torch.randn(ho_, wo_, (cout + 7) // 8 * 8, device=dev), 0, cout), f16=True)
        if os.environ.get("BENCH_RES") == "2":
            kw["res2"] = Act(torch.randn(ho_, wo_, (cout + 3) // 4 * 4, device=dev), 0, cout)
    out = ops.conv(x, cw, stride=s, **kw)
    torch.cuda.synchronize()
    ms = 0.0
    for _ in range(iters):
        if not os.environ.get("BENCH_NOFLUSH"):
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.conv(x, cw, stride=s, out=out, **kw)
        e1.record()
        torch.cuda.synchronize()
        ms += e0.elapsed_time(e1)
    ms /= iters
    ho, wo = (h + 2 * (k // 2) - k) // s + 1, (w + 2 * (k // 2) - k) // s + 1
    fl = 2.0 * ho * wo * cout * cin * k * k
    by = (2.0 if x.f16 else 4.0) * h * w * cin + (2.0 if out.f16 else 4.0) * ho * wo * cout + 4.0 * cout * cin * k * k
    name += (" f16in" if x.f16 else "") + (" f16out" if out.f16 else "") + (" +res" if "res" in kw else "")
    if "res" in kw:
        by += (2.0 if kw["res"].f16 else 4.0) * ho * wo * cout
    print(f"{name:72s} {ms:8.3f} ms {fl / ms / 1e9:8.1f} TFLOP/s {by / ms / 1e6:8.0f} GB/s", flush=True)
