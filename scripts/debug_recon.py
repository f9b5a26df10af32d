"""Debug aid (GPU box): localise product-vs-oracle differences inside get_recon_and_feature."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dcvc_b200
from dcvc_b200 import arch, synth, ops, layers
from dcvc_b200.tensor import Act
from oracle import codec_oracle as co, nets_oracle as no
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gpu_util import to_act, from_act

dev = torch.device("cuda:0")
h = w = 64
sd_i = arch.synth_state_dict(arch.intra_spec(), 0)
sd_p = arch.synth_state_dict(arch.dmc_spec(), 100)
p_net = dcvc_b200.DMC(); p_net.load_state_dict(sd_p); p_net = p_net.to(dev).eval(); p_net.update()
o_i, o_p = co.OracleIntra(sd_i, 1), co.OracleDMC(sd_p, 1)
o_i.update(); o_p.update()
frames = synth.synth_frames(2, h, w, seed=0)
torch.set_num_threads(8)
def md(a, b): return f"max {float((a-b).abs().max()):.3e} mean {float((a-b).abs().mean()):.3e} | ref absmax {float(b.abs().max()):.3e}"
with torch.no_grad():
    ref = o_i.compress(frames[0:1], False, 10)
    dpbo = {"ref_frame": ref["x_hat"], "ref_feature": None, "ref_mv_feature": None, "ref_y": None, "ref_mv_y": None}
    p_net.taps = {}
    d = {kk: (v.to(dev) if v is not None else None) for kk, v in dpbo.items()}
    p_net.compress(frames[1:2].to(dev), d, False, 10, 1)
    T = {k: v.cpu() for k, v in p_net.taps.items()}
    q = o_p.get_q(False, 10)[3]
    sd = o_p.sd
    r = no.contextual_decoder(sd, T["p_y_hat"], T["p_ctx2"], T["p_ctx3"], q)
    print("contextual_decoder:", md(T["p_recon_res"], r))
    f0 = no.conv(sd, "recon_generation_net.first_conv", torch.cat((T["p_ctx1"], T["p_recon_res"]), 1), 1, 1)
    u1 = no.unet(sd, "recon_generation_net.unet_1", f0, no.dcb)
    u2 = no.unet(sd, "recon_generation_net.unet_2", u1, no.dcb)
    print("feature (oracle on product inputs):", md(T["p_feature"], u2))
    xh = no.conv(sd, "recon_generation_net.recon_conv", u2, 1, 1).clamp(0, 1)
    print("x_hat:", md(T["p_x_hat"], xh))
    # piecewise on the GPU
    net = p_net.net
    f0g = ops.conv(to_act(torch.cat((T["p_ctx1"], T["p_recon_res"]), 1)), net.cw("recon_generation_net.first_conv"))
    print("first_conv:", md(from_act(f0g), f0))
    u1g = layers.unet(net, "recon_generation_net.unet_1", to_act(f0), layers.dcb)
    print("unet_1:", md(from_act(u1g), u1))
    p = "recon_generation_net.unet_1"
    x1 = no.dcb(sd, p + ".conv1", f0)
    x1g = layers.dcb(net, p + ".conv1", to_act(f0))
    print(" conv1 dcb:", md(from_act(x1g), x1))
    x2 = no.dcb(sd, p + ".conv2", torch.nn.functional.max_pool2d(x1, 2, 2))
    x2g = layers.dcb(net, p + ".conv2", ops.max_pool2(to_act(x1)))
    print(" conv2 dcb:", md(from_act(x2g), x2))
    x3 = no.dcb(sd, p + ".conv3", torch.nn.functional.max_pool2d(x2, 2, 2))
    x3g = layers.dcb(net, p + ".conv3", ops.max_pool2(to_act(x2)))
    print(" conv3 dcb:", md(from_act(x3g), x3))
    t, tg = x3, to_act(x3)
    for i in range(4):
        t = no.dcb(sd, f"{p}.context_refine.{i}", t); tg = layers.dcb(net, f"{p}.context_refine.{i}", tg)
        print(f" refine {i}:", md(from_act(tg), t))
    d3 = no.subpel(sd, p + ".up3", t, 1); d3g = layers.subpel(net, p + ".up3", to_act(t))
    print(" up3:", md(from_act(d3g), d3))
    c3 = torch.cat((x2, d3), 1)
    e3 = no.dcb(sd, p + ".up_conv3", c3); e3g = layers.dcb(net, p + ".up_conv3", to_act(c3))
    print(" up_conv3:", md(from_act(e3g), e3))
