"""Debug aid (GPU box): per-stage differences product vs oracle, teacher-forced per frame."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dcvc_b200
from dcvc_b200 import arch, synth, ops
from oracle import codec_oracle as co

mode = int(sys.argv[1]) if len(sys.argv) > 1 else 0
ops.set_conv_mode(mode)
dev = torch.device("cuda:0")
h = w = 64
sd_i = arch.synth_state_dict(arch.intra_spec(), 0)
sd_p = arch.synth_state_dict(arch.dmc_spec(), 100)
p_net = dcvc_b200.DMC(); p_net.load_state_dict(sd_p); p_net = p_net.to(dev).eval(); p_net.update()
o_i, o_p = co.OracleIntra(sd_i, 1), co.OracleDMC(sd_p, 1)
o_i.update(); o_p.update()
frames = synth.synth_frames(4, h, w, seed=0)
torch.set_num_threads(8)
with torch.no_grad():
    ref = o_i.compress(frames[0:1], False, 10)
    dpbo = {"ref_frame": ref["x_hat"], "ref_feature": None, "ref_mv_feature": None, "ref_y": None, "ref_mv_y": None}
    for k in range(1, 4):
        o_p.taps = {}
        refp = o_p.compress(frames[k:k+1], dpbo, False, 10, k % 4)
        p_net.taps = {}
        d = {kk: (v.to(dev) if v is not None else None) for kk, v in dpbo.items()}
        encp = p_net.compress(frames[k:k+1].to(dev), d, False, 10, k % 4)
        print(f"--- frame {k}: bytes {len(encp['bit_stream'])} vs {len(refp['bit_stream'])}")
        for name in ("est_mv", "mv_y", "mv_z_hat", "mv_params", "mv_y_hat", "mv_hat", "ctx1", "ctx2", "ctx3", "y", "z_hat", "params", "y_hat", "x_hat"):
            a = p_net.taps["p_" + name].cpu(); b = o_p.taps["p_" + name]
            diff = (a - b).abs()
            print(f"{name:10s} max|d| {diff.max():.3e} mean|d| {diff.mean():.3e} ref max {b.abs().max():.3e} n(|d|>0.4) {(diff > 0.4).sum().item()}")
        dpbo = refp["dpb"]
