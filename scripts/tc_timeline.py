"""Role timeline of conv_tc_kernel on CTA 0 (debug build: DCVC_EXTRA_NVCC=-DDCVC_TC_TIMELINE):
per tile, where the MMA warp and the first epilogue warp spend their cycles.
usage: python scripts/tc_timeline.py cin cout k h w [res]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from dcvc_b200 import _lib, layers, ops
from dcvc_b200.tensor import Act

cin, cout, k, h, w = [int(v) for v in sys.argv[1:6]]
res = len(sys.argv) > 6 and sys.argv[6] == "res"
ops.set_conv_mode(1)
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
sd = {"c.weight": torch.randn(cout, cin, k, k, generator=g) * 0.05, "c.bias": torch.randn(cout, generator=g)}
net = layers.Net(sd, dev)
x = ops.copy(Act(torch.randn(h, w, (cin + 7) // 8 * 8, device=dev), 0, cin), f16=True)
kw = {"f16_out": True}
if res:
    kw["res"] = ops.copy(Act(torch.randn(h, w, (cout + 7) // 8 * 8, device=dev), 0, cout), f16=True)
out = ops.conv(x, net.cw("c"), **kw)
for _ in range(3):
    ops.conv(x, net.cw("c"), out=out, **kw)
torch.cuda.synchronize()
_lib.lib(); lib = ctypes.CDLL(_lib.SO_PATH)
buf = np.zeros((2, 4096, 8), dtype=np.uint64)
rc = lib.dcvc_debug_timeline(buf.ctypes.data_as(ctypes.c_void_p), buf.nbytes)
assert rc == 0, rc
m, e = buf[0].astype(np.int64), buf[1].astype(np.int64)
n = int((m[:, 3] > 0).sum())
print(f"tiles on CTA 0: {n}")
m, e = m[:n], e[:n]
e = e[e[:, 3] > 0]          # two epilogue sets: warp 0 sees every other tile
t0 = m[0, 0]
per_tile = (m[-1, 3] - m[0, 0]) / n
print(f"cycles per tile (MMA warp, whole run): {per_tile:.0f}")
print("MMA warp, mean cycles per tile: wait acc free %.0f | wait TMA data %.0f | issue + commit %.0f | loop rest %.0f" % (
    (m[:, 1] - m[:, 0]).mean(), (m[:, 2] - m[:, 1]).mean(), (m[:, 3] - m[:, 2]).mean(),
    (m[1:, 0] - m[:-1, 3]).mean()))
print("EPI warp 0, mean cycles per tile: wait accumulator %.0f | tmem load + hand back %.0f | math + park + store %.0f | to next %.0f" % (
    (e[:, 1] - e[:, 0]).mean(), (e[:, 2] - e[:, 1]).mean(), (e[:, 3] - e[:, 2]).mean(),
    (e[1:, 0] - e[:-1, 3]).mean()))
print("EPI detail: e2->bias/act %.0f | ->residual/scale/clamp %.0f | ->park+fence+syncwarp %.0f | ->store issue %.0f | ->wait_read+next residual+syncwarp %.0f" % (
    (e[:, 4] - e[:, 2]).mean(), (e[:, 5] - e[:, 4]).mean(), (e[:, 6] - e[:, 5]).mean(),
    (e[:, 7] - e[:, 6]).mean(), (e[:, 3] - e[:, 7]).mean()))
d = m[:, 1] - m[:, 0]
print("wait acc free per tile:", d[4:40].tolist())
for i in (20, 21, 22, 23):
    print("tile", i, "mma", (m[i, :4] - t0).tolist(), "epi(warp 0, its tile %d)" % i, (e[i // (2 if len(e) < n else 1)] - t0).tolist())
