"""-m gpu: DCVC-HEM on the B200 kernel set against the HEM oracle and the golden GOPs the
UNMODIFIED reference produced (tests/golden/hem_gop_*.npz).

  * dual-prior kernels teacher-forced on oracle floats: symbols, CDF indexes, y_hat bit-exact;
  * SE gate / channel scale against their torch definitions;
  * mode 0 (fp32 FMA): every frame's stream BYTE-IDENTICAL to the reference's, recon RMS < 1e-6;
  * mode 1 (tcgen05): recon RMS and stream size against the reference's, bounds stated;
  * encoder recon == decoder recon bit for bit in both modes; estimate-mode bits vs the oracle."""
import os

import numpy as np
import pytest
import torch

from dcvc_b200 import entropy_models as em
from dcvc_b200 import hem, ops, synth
from dcvc_b200.hem import arch as harch
from dcvc_b200.tensor import Act
from oracle import entropy_oracle as eo
from oracle import hem_oracle as ho

from gpu_util import dev, from_act, to_act

pytestmark = pytest.mark.gpu


@pytest.fixture
def conv_mode():
    prev = ops.CONV_MODE
    yield ops.set_conv_mode
    ops.set_conv_mode(prev)


@pytest.mark.parametrize("dist,C,h,w", [("laplace", 64, 6, 8), ("gaussian", 192, 4, 6), ("laplace", 96, 68, 120)])
def test_dual_steps_teacher_forced(dist, C, h, w):
    g = torch.Generator().manual_seed(C + h)
    y = torch.randn(1, C, h, w, generator=g) * 6
    q_step = torch.rand(1, C, h, w, generator=g) * 2                      # some below 0.5
    scales = torch.exp(torch.rand(1, C, h, w, generator=g) * 9 - 5)
    means = torch.randn(1, C, h, w, generator=g)
    prior = torch.cat([torch.exp(torch.rand(1, C // 2, h, w, generator=g) * 9 - 5) if i % 2 == 0
                       else torch.randn(1, C // 2, h, w, generator=g) for i in range(4)], 1)
    post = torch.rand(C, generator=g) + 0.5
    ref = ho.dual_forward(y, means, scales, q_step, lambda p: prior)
    fused = to_act(torch.cat((q_step, scales, means), 1))
    edges = em.index_edges(dist).to(dev())
    cat = Act.new(h, w, 4 * C, dev())
    y_out = Act.new(h, w, C, dev())
    n = C // 2 * h * w
    bits = torch.zeros(1, device=dev())
    ws = torch.empty((n + 255) // 256, device=dev())
    ya, pa = to_act(y), to_act(prior)
    d_cat, d_out = Act.new(h, w, 4 * C, dev()), Act.new(h, w, C, dev())
    for step in range(2):
        sym = torch.empty(n, dtype=torch.int16, device=dev())
        idx = torch.empty(n, dtype=torch.uint8, device=dev())
        ops.dual_step_enc(ya, fused if step == 0 else None, pa if step else None, cat, y_out,
                          edges, step, em.DIST_ID[dist], post.to(dev()), sym, idx, bits, ws)
        assert torch.equal(sym.cpu(), ref["y_q_w"][step].reshape(-1).to(torch.int16)), step
        assert torch.equal(idx.cpu().int(), eo.build_indexes(ref["scales_w"][step], dist).reshape(-1)), step
        # decoder mirror: same indexes, scatter reproduces the encoder's buffers bit for bit
        idx2 = torch.empty_like(idx)
        ops.dual_step_indexes(fused if step == 0 else None, pa if step else None, d_cat, d_out,
                              edges, step, idx2)
        assert torch.equal(idx2, idx)
        ops.dual_step_dec(fused if step == 0 else None, pa if step else None, d_cat, d_out, edges,
                          step, sym, post.to(dev()))
        if step == 0:
            want = torch.cat((torch.cat(ref_h00_h11(ref, y, means, scales, q_step), 1), means, scales,
                              q_step.clamp_min(0.5)), 1)
            assert torch.equal(from_act(cat), want)
            assert torch.equal(d_cat.buf, cat.buf)
    assert torch.equal(from_act(y_out), ref["y_hat"] * post.view(1, -1, 1, 1))
    assert torch.equal(d_out.buf, y_out.buf)
    ref_bits = ho.y_bits(ref["y_q"], ref["scales_hat"], dist).sum().item()
    assert abs(bits.item() - ref_bits) <= 1e-4 * ref_bits


def ref_h00_h11(ref, y, means, scales, q_step):
    """y_hat_0_0, y_hat_1_1 of common_model.py:117-120 recomputed from the oracle's pieces"""
    m0, m1 = ho.get_mask(y.shape[2], y.shape[3])
    yq = y / q_step.clamp_min(0.5)
    y0, y1 = yq.chunk(2, 1)
    mu0, mu1 = means.chunk(2, 1)
    s0, s1 = scales.chunk(2, 1)
    return ho._pwm(y0, s0, mu0, m0)[1], ho._pwm(y1, s1, mu1, m1)[1]


def test_se_gate_and_channel_scale():
    g = torch.Generator().manual_seed(3)
    for c, h, w in ((32, 40, 56), (128, 17, 31)):
        x = torch.randn(1, c, h, w, generator=g)
        w1 = torch.randn(c // 16, c, generator=g) * 0.3
        w2 = torch.randn(c, c // 16, generator=g) * 0.3
        gate = ops.se_gate(to_act(x), w1.to(dev()), w2.to(dev())).cpu()
        want = torch.sigmoid(torch.relu(x.double().mean(dim=(-1, -2)) @ w1.double().t()) @ w2.double().t())[0]
        assert float((gate.double() - want).abs().max()) < 1e-6
        v = torch.rand(c, generator=g) + 0.5
        assert torch.equal(from_act(ops.channel_scale(to_act(x), v.to(dev()))), x * v.view(1, -1, 1, 1))
        assert torch.equal(from_act(ops.channel_scale(to_act(x), v.to(dev()), divide=True)),
                           x / v.view(1, -1, 1, 1))


def _models(seed, gain, coder_mode="hem", parts=1):
    sd_i = harch.hem_synth_state_dict(harch.hem_intra_spec(), seed, gain)
    sd_p = harch.hem_synth_state_dict(harch.hem_dmc_spec(), seed + 100, gain)
    i_net, p_net = hem.IntraNoAR(coder_mode=coder_mode, stream_part=parts), \
        hem.DMC(coder_mode=coder_mode, stream_part=parts)
    i_net.load_state_dict(sd_i)
    p_net.load_state_dict(sd_p)
    i_net, p_net = i_net.to(dev()).eval(), p_net.to(dev()).eval()
    i_net.update(force=True)
    p_net.update(force=True)
    return i_net, p_net, sd_i, sd_p


def _run_gop(i_net, p_net, frames, h, w, q_i, q_mv, q_y):
    """the reference harness's calls with rounded q (encode_decode semantics, test_video.py of HEM)"""
    out = []
    with torch.no_grad():
        qi, _ = ho.get_rounded_q(q_i)
        enc = i_net.compress(frames[0:1].to(dev()), qi)
        dec = i_net.decompress(enc["bit_stream"], h, w, qi)
        out.append((enc["bit_stream"], dec["x_hat"].cpu().contiguous()))
        dpb = {"ref_frame": dec["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
        qm, _ = ho.get_rounded_q(q_mv)
        qy, _ = ho.get_rounded_q(q_y)
        for k in range(1, frames.shape[0]):
            enc = p_net.compress(frames[k:k + 1].to(dev()), dpb, qm, qy)
            dec = p_net.decompress(dpb, enc["bit_stream"], h, w, qm, qy)
            assert torch.equal(enc["dbp"]["ref_frame"].clamp(0, 1), dec["dpb"]["ref_frame"]), k
            for key in ("ref_feature", "ref_y", "ref_mv_y"):
                assert torch.equal(enc["dbp"][key], dec["dpb"][key]), (k, key)
            dpb = dec["dpb"]
            out.append((enc["bit_stream"], dpb["ref_frame"].cpu().contiguous()))
    return out


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_hem_gop_vs_reference_golden(golden_dir, name, mode, conv_mode):
    conv_mode(mode)
    g = np.load(os.path.join(golden_dir, f"hem_gop_{name}.npz"))
    h, w, n, seed = [int(v) for v in g["meta"]]
    gain, q_i, q_mv, q_y = [float(v) for v in g["gain_q"]]
    i_net, p_net, _, _ = _models(seed, gain)
    frames = synth.synth_frames(n, h, w, seed=seed)
    res = _run_gop(i_net, p_net, frames, h, w, q_i, q_mv, q_y)
    same = [res[k][0] == g[f"stream_{k}"].tobytes() for k in range(n)]
    rms = [float((res[k][1] - torch.from_numpy(g[f"x_hat_{k}"])).pow(2).mean().sqrt()) for k in range(n)]
    bits = sum(len(s) for s, _ in res)
    ref_bits = sum(g[f"stream_{k}"].size for k in range(n))
    print(f"hem {name} mode {mode}: byte-identical {same}, x_hat RMS {rms}, bytes {bits} vs {ref_bits}")
    if mode == 0:
        assert all(same), same
        assert max(rms) < 1e-6, rms
    else:
        assert max(rms) < 2e-3, rms                       # 0.02 dB at 38 dB
        assert abs(bits - ref_bits) <= max(0.005 * ref_bits, 16), (bits, ref_bits)


def test_hem_estimate_mode_and_api(tmp_path, conv_mode):
    """forward / forward_one_frame bits against the oracle with UNROUNDED q (estimate mode does not
    round), and encode_decode with frame files (uint16 q indexes in the header)."""
    conv_mode(0)
    h, w, seed = 64, 128, 1
    i_net, p_net, sd_i, sd_p = _models(seed, 0.5)
    o_i, o_p = ho.OracleHemIntra(sd_i), ho.OracleHemDMC(sd_p)
    frames = synth.synth_frames(2, h, w, seed=seed)
    with torch.no_grad():
        r = i_net.forward(frames[0:1].to(dev()), 0.7312)
        ro = o_i.forward(frames[0:1], 0.7312)
        assert abs(r["bit"] - ro["bit"]) <= 1e-3 * ro["bit"]
        assert float((r["x_hat"].cpu() - ro["x_hat"]).abs().max()) < 1e-4
        path = str(tmp_path / "i.bin")
        ri = i_net.encode_decode(frames[0:1].to(dev()), 0.7312, path, pic_width=w, pic_height=h)
        from dcvc_b200.hem import stream_helper as sh
        assert sh.decode_i(path)[:3] == (h, w, 73) and ri["bit"] == sh.filesize(path) * 8
        dpb = {"ref_frame": ri["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
        dpbo = {k: (v.cpu() if v is not None else None) for k, v in dpb.items()}
        p = p_net.forward_one_frame(frames[1:2].to(dev()), dpb, mv_y_q_scale=1.0149, y_q_scale=0.7134)
        po = o_p.forward_one_frame(frames[1:2], dpbo, 1.0149, 0.7134)
        for key in ("bit", "bit_y", "bit_z", "bit_mv_y", "bit_mv_z"):
            assert abs(float(p[key]) - float(po[key])) <= 2e-3 * float(po[key]) + 0.5, key
        ppath = str(tmp_path / "p.bin")
        rp = p_net.encode_decode(frames[1:2].to(dev()), dpb, ppath, pic_width=w, pic_height=h,
                                 mv_y_q_scale=1.0149, y_q_scale=0.7134)
        assert sh.decode_p(ppath)[:2] == (101, 71) and rp["bit"] == sh.filesize(ppath) * 8
        assert set(rp["dpb"]) == {"ref_frame", "ref_feature", "ref_y", "ref_mv_y"}


def test_hem_interleaved_coder_same_recon():
    h, w = 64, 64
    frames = synth.synth_frames(2, h, w, seed=0)
    a = _run_gop(*_models(0, 1.0)[:2], frames, h, w, 1.0, 1.0, 1.0)
    b = _run_gop(*_models(0, 1.0, "interleaved", 2)[:2], frames, h, w, 1.0, 1.0, 1.0)
    for k in range(2):
        assert torch.equal(a[k][1], b[k][1])


def test_hem_1080p_modes_agree_and_enc_equals_dec(conv_mode):
    """The size the hem_sweep workload runs (1920x1080 zero-padded to 1088): there the tensor-core
    mode stores most conv-to-conv activations as binary16 (the golden GOPs are too small for
    that).  I + 2 P, interleaved coder: encoder reconstruction == decoder reconstruction bit for
    bit in both modes (asserted inside _run_gop), and mode 1 against mode 0 (fp32 FMA, held
    byte-identical to the unmodified reference by the golden tests): reconstruction RMS and
    stream size."""
    import torch.nn.functional as F
    h, w, n = 1088, 1920, 3
    frames = F.pad(synth.synth_frames(n, 1080, w, seed=3), (0, 0, 0, 8))
    res = {}
    for mode in (0, 1):
        conv_mode(mode)
        i_net, p_net, _, _ = _models(0, 1.0, coder_mode="interleaved", parts=8)
        res[mode] = _run_gop(i_net, p_net, frames, h, w, 1.0, 1.0, 1.0)
        del i_net, p_net
        torch.cuda.empty_cache()
    for k in range(n):
        rms = float((res[1][k][1] - res[0][k][1]).pow(2).mean().sqrt())
        b0, b1 = len(res[0][k][0]), len(res[1][k][0])
        print(f"hem 1080p frame {k}: x_hat RMS mode1 vs mode0 {rms:.3e}, bytes {b1} vs {b0}")
        assert rms < 2e-3, (k, rms)                       # 0.02 dB at 38 dB
        assert abs(b1 - b0) <= 0.005 * b0, (k, b0, b1)
