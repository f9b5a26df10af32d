import sys, os
sys.path.insert(0, "/root/repo")
import torch, dcvc_b200
from dcvc_b200 import arch, ops, synth
from dcvc_b200.graphs import GraphedDMC
dev = torch.device("cuda:0")
ops.set_conv_mode(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
h, w = 128, 192
i_net = dcvc_b200.IntraNoAR(stream_part=2, coder_mode="interleaved"); p_net = dcvc_b200.DMC(stream_part=2, coder_mode="interleaved")
i_net.load_state_dict(arch.synth_state_dict(arch.intra_spec(), 5)); p_net.load_state_dict(arch.synth_state_dict(arch.dmc_spec(), 105))
i_net, p_net = i_net.to(dev).eval(), p_net.to(dev).eval(); i_net.update(); p_net.update()
fr = synth.synth_frames(4, h, w, seed=7).to(dev)
KEYS = ("ref_frame", "ref_feature", "ref_mv_feature", "ref_y", "ref_mv_y")
with torch.no_grad():
    e = i_net.compress(fr[0:1], True, 1); d = i_net.decompress(e["bit_stream"], h, w, True, 1)
    dpb0 = {"ref_frame": d["x_hat"], "ref_feature": None, "ref_mv_feature": None, "ref_y": None, "ref_mv_y": None}
    g = GraphedDMC(p_net)
    for rep in range(3):
        ee = p_net.compress(fr[1:2], dpb0, True, 1, 1)
        ref_s = ee["bit_stream"]; ref_dpb = {k: ee["dpb"][k].clone() for k in KEYS}
        ge = g.compress(fr[1:2], dpb0, True, 1, 1)
        print(rep, "enc stream equal", ge["bit_stream"] == ref_s, len(ge["bit_stream"]), len(ref_s),
              {k: bool(torch.equal(ge["dpb"][k], ref_dpb[k])) for k in KEYS})
        dd = p_net.decompress(dpb0, ref_s, h, w, True, 1, 1)
        try:
            gd = g.decompress(dpb0, ref_s, h, w, True, 1, 1)
            print(rep, "dec", {k: bool(torch.equal(gd["dpb"][k], dd["dpb"][k])) for k in KEYS})
        except Exception as ex:
            print(rep, "dec error", ex)
