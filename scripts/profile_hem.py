"""Per-kernel table of one DCVC-HEM P frame at 1080p (CUDA events per launch, ops.PROFILE)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from dcvc_b200 import hem, ops, synth
from dcvc_b200.hem import arch as harch

dev = torch.device("cuda:0")
ops.set_conv_mode(1)
H, W = 1088, 1920
i_net = hem.IntraNoAR(coder_mode="interleaved", stream_part=8)
p_net = hem.DMC(coder_mode="interleaved", stream_part=8)
i_net.load_state_dict(harch.hem_synth_state_dict(harch.hem_intra_spec(), 0, 1.0))
p_net.load_state_dict(harch.hem_synth_state_dict(harch.hem_dmc_spec(), 100, 1.0))
i_net, p_net = i_net.to(dev).eval(), p_net.to(dev).eval()
i_net.update(); p_net.update()
fr = F.pad(synth.synth_frames(3, 1080, W, seed=0), (0, 0, 0, 8)).to(dev)
with torch.no_grad():
    s = i_net.compress(fr[0:1], 1.0)["bit_stream"]
    dpb = {"ref_frame": i_net.decompress(s, H, W, 1.0)["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    for k in (1, 2):
        if k == 2:
            ops.PROFILE = []
        e = p_net.compress(fr[k:k+1], dpb, 1.0, 1.0)
        dpb = p_net.decompress(dpb, e["bit_stream"], H, W, 1.0, 1.0)["dpb"]
torch.cuda.synchronize()
rows = {}
for name, fl, nb, e0, e1 in ops.PROFILE:
    r = rows.setdefault(name, [0, 0.0, 0, 0])
    r[0] += 1; r[1] += e0.elapsed_time(e1); r[2] += fl; r[3] += nb
tot = sum(r[1] for r in rows.values())
print(f"total {tot:.2f} ms, {sum(r[0] for r in rows.values())} launches")
for name, r in sorted(rows.items(), key=lambda kv: -kv[1][1])[:45]:
    print(f"{name:62s} n={r[0]:3d} ms={r[1]:7.3f} tf={r[2]/(r[1]*1e9+1e-9):6.0f} gbs={r[3]/(r[1]*1e6+1e-9):6.0f}")
