"""-m gpu end-to-end parity: IntraNoAR + DMC through the reference-facing module API against the
CPU oracle (and the committed golden vectors produced by the unmodified reference).

Tolerances (BASELINE.json north_star): reconstructions within 0.02 dB PSNR of the reference's,
bpp within 0.5 %; encoder and decoder reconstructions of the product bit-identical; compat-mode
streams decodable by the oracle's coder."""
import os

import numpy as np
import pytest
import torch

import dcvc_b200
from dcvc_b200 import arch, synth
from oracle import codec_oracle as co

from gpu_util import dev, psnr

pytestmark = pytest.mark.gpu


def _models(seed, part_i, part_p, mode="compat"):
    sd_i = arch.synth_state_dict(arch.intra_spec(), seed)
    sd_p = arch.synth_state_dict(arch.dmc_spec(), seed + 100)
    i_net = dcvc_b200.IntraNoAR(stream_part=part_i, coder_mode=mode)
    p_net = dcvc_b200.DMC(stream_part=part_p, coder_mode=mode)
    i_net.load_state_dict(sd_i)
    p_net.load_state_dict(sd_p)
    i_net = i_net.to(dev()).eval()
    p_net = p_net.to(dev()).eval()
    i_net.update(force=True)
    p_net.update(force=True)
    return i_net, p_net, sd_i, sd_p


def _run_gop(i_net, p_net, frames, h, w, q_in_ckpt, q_index):
    out = []
    dpb = None
    with torch.no_grad():
        for k in range(frames.shape[0]):
            x = frames[k:k + 1].to(dev())
            if k == 0:
                enc = i_net.compress(x, q_in_ckpt, q_index)
                dec = i_net.decompress(enc["bit_stream"], h, w, q_in_ckpt, q_index)
                assert torch.equal(enc["x_hat"], dec["x_hat"]), "I frame: enc/dec recon differ"
                dpb = {"ref_frame": dec["x_hat"], "ref_feature": None, "ref_mv_feature": None,
                       "ref_y": None, "ref_mv_y": None}
            else:
                enc = p_net.compress(x, dpb, q_in_ckpt, q_index, k % 4)
                dec = p_net.decompress(dpb, enc["bit_stream"], h, w, q_in_ckpt, q_index, k % 4)
                for key in ("ref_frame", "ref_feature", "ref_mv_feature", "ref_y", "ref_mv_y"):
                    assert torch.equal(enc["dpb"][key], dec["dpb"][key]), f"P frame {k}: {key}"
                dpb = dec["dpb"]
            out.append((enc["bit_stream"], dpb["ref_frame"].detach().cpu().contiguous()))
    return out


@pytest.fixture
def conv_mode():
    """Sets ops.CONV_MODE for one test (0 = fp32 FMA exact mode, 1 = tcgen05 TF32)."""
    from dcvc_b200 import ops
    prev = ops.CONV_MODE
    yield ops.set_conv_mode
    ops.set_conv_mode(prev)


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("name", ["a64", "b96x128"])
def test_gop_vs_golden(golden_dir, name, mode, conv_mode):
    conv_mode(mode)
    g = np.load(os.path.join(golden_dir, f"dc_gop_{name}.npz"))
    h, w, n, q_in_ckpt, q_index, part_i, part_p, seed = [int(v) for v in g["meta"]]
    i_net, p_net, _, _ = _models(seed, part_i, part_p)
    frames = synth.synth_frames(n, h, w, seed=seed)
    res = _run_gop(i_net, p_net, frames, h, w, bool(q_in_ckpt), q_index)
    # Free-running GOP: symbol flips at .5 rounding boundaries (different fp32 summation order
    # than oneDNN) feed back through the dpb, so per-frame drift is allowed 0.05 dB while the
    # sequence averages -- what an RD point reports -- must meet the 0.02 dB / 0.5 % tolerance.
    # test_frames_teacher_forced checks every frame at 0.02 dB without the feedback.
    d_psnrs, bits, ref_bits = [], 0, 0
    for k, (stream, x_hat) in enumerate(res):
        ref = torch.from_numpy(g[f"x_hat_{k}"])
        src = frames[k:k + 1]
        d_psnrs.append(psnr(x_hat, src) - psnr(ref, src))
        bits += len(stream) * 8
        ref_bits += g[f"stream_{k}"].size * 8
        assert abs(d_psnrs[-1]) < 0.05, f"frame {k}: PSNR differs by {d_psnrs[-1]:.4f} dB"
        rb = g[f"stream_{k}"].size * 8
        assert abs(len(stream) * 8 - rb) <= max(0.01 * rb, 64), f"frame {k}: bits {len(stream) * 8} vs {rb}"
    assert abs(sum(d_psnrs) / len(d_psnrs)) < 0.02, d_psnrs
    assert abs(bits - ref_bits) <= 0.005 * ref_bits, (bits, ref_bits)


def _to_dev(dpb):
    return {k: (v.to(dev()) if v is not None else None) for k, v in dpb.items()}


def test_frames_teacher_forced():
    """Every frame coded from the ORACLE's decoded-picture buffer (no drift feedback): recon
    within 0.02 dB PSNR and stream size within 0.5 % (or 8 bytes) of the oracle, per frame."""
    h, w, n, seed = 64, 64, 4, 0
    i_net, p_net, sd_i, sd_p = _models(seed, 1, 1)
    o_i, o_p = co.OracleIntra(sd_i, 1), co.OracleDMC(sd_p, 1)
    o_i.update()
    o_p.update()
    frames = synth.synth_frames(n, h, w, seed=seed)
    torch.set_num_threads(8)
    with torch.no_grad():
        ref = o_i.compress(frames[0:1], False, 10)
        enc = i_net.compress(frames[0:1].to(dev()), False, 10)
        pairs = [(enc["bit_stream"], enc["x_hat"].cpu(), ref["bit_stream"], ref["x_hat"])]
        dpbo = {"ref_frame": ref["x_hat"], "ref_feature": None, "ref_mv_feature": None,
                "ref_y": None, "ref_mv_y": None}
        for k in range(1, n):
            refp = o_p.compress(frames[k:k + 1], dpbo, False, 10, k % 4)
            encp = p_net.compress(frames[k:k + 1].to(dev()), _to_dev(dpbo), False, 10, k % 4)
            decp = p_net.decompress(_to_dev(dpbo), encp["bit_stream"], h, w, False, 10, k % 4)
            assert torch.equal(encp["dpb"]["ref_frame"], decp["dpb"]["ref_frame"])
            pairs.append((encp["bit_stream"], decp["dpb"]["ref_frame"].cpu(), refp["bit_stream"],
                          refp["dpb"]["ref_frame"]))
            dpbo = refp["dpb"]
    for k, (s, xh, rs, rxh) in enumerate(pairs):
        d = abs(psnr(xh, frames[k:k + 1]) - psnr(rxh, frames[k:k + 1]))
        assert d < 0.02, f"frame {k}: dPSNR {d:.4f}"
        assert abs(len(s) - len(rs)) <= max(8, 0.005 * len(rs)), (k, len(s), len(rs))


def test_compat_stream_decodes_with_oracle():
    """The product's compat-mode frames are decodable by the oracle decoder (same container,
    same tables): interchange with the reference."""
    h = w = 64
    i_net, p_net, sd_i, sd_p = _models(0, 1, 1)
    frames = synth.synth_frames(2, h, w, seed=0)
    res = _run_gop(i_net, p_net, frames, h, w, True, 1)
    o_i = co.OracleIntra(sd_i, 1)
    o_i.update()
    with torch.no_grad():
        dec = o_i.decompress(res[0][0], h, w, True, 1)
    assert abs(psnr(dec["x_hat"], frames[0:1]) - psnr(res[0][1], frames[0:1])) < 0.02
    assert float((dec["x_hat"] - res[0][1]).abs().max()) < 2e-2
    # the P frame's stream too, from the oracle's own decoded I frame
    o_p = co.OracleDMC(sd_p, 1)
    o_p.update()
    dpb = {"ref_frame": dec["x_hat"], "ref_feature": None, "ref_mv_feature": None,
           "ref_y": None, "ref_mv_y": None}
    with torch.no_grad():
        decp = o_p.decompress(dpb, res[1][0], h, w, True, 1, 1)
    xp = decp["dpb"]["ref_frame"]
    assert abs(psnr(xp, frames[1:2]) - psnr(res[1][1], frames[1:2])) < 0.02
    assert float((xp - res[1][1]).abs().max()) < 2e-2


def test_estimate_mode_matches_oracle():
    h = w = 64
    i_net, p_net, sd_i, sd_p = _models(0, 1, 1)
    o_i, o_p = co.OracleIntra(sd_i, 1), co.OracleDMC(sd_p, 1)
    frames = synth.synth_frames(2, h, w, seed=0)
    with torch.no_grad():
        r = i_net.forward(frames[0:1].to(dev()), False, 20)
        ro = o_i.forward(frames[0:1], False, 20)
        assert abs(float(r["bit"]) - float(ro["bit"])) < 5e-3 * float(ro["bit"])
        assert float((r["x_hat"].cpu() - ro["x_hat"]).abs().max()) < 1e-2
        dpb = {"ref_frame": r["x_hat"].clamp_(0, 1), "ref_feature": None, "ref_mv_feature": None,
               "ref_y": None, "ref_mv_y": None}
        dpbo = {"ref_frame": ro["x_hat"].clamp_(0, 1), "ref_feature": None,
                "ref_mv_feature": None, "ref_y": None, "ref_mv_y": None}
        p = p_net.forward_one_frame(frames[1:2].to(dev()), dpb, False, 20, 1)
        po = o_p.forward_one_frame(frames[1:2], dpbo, False, 20, 1)
        assert abs(float(p["bit"]) - float(po["bit"])) < 5e-3 * float(po["bit"])
        for key in ("bpp_y", "bpp_z", "bpp_mv_y", "bpp_mv_z"):
            assert abs(float(p[key]) - float(po[key])) <= 5e-3 * float(po[key]) + 1e-6, key


def test_interleaved_mode_round_trip_and_size():
    h, w = 96, 128
    i_c, p_c, _, _ = _models(1, 1, 1, "compat")
    i_v, p_v, _, _ = _models(1, 2, 2, "interleaved")
    frames = synth.synth_frames(3, h, w, seed=1)
    rc = _run_gop(i_c, p_c, frames, h, w, False, 37)
    rv = _run_gop(i_v, p_v, frames, h, w, False, 37)
    for k in range(3):
        assert torch.equal(rc[k][1], rv[k][1]), "reconstruction must not depend on the coder mode"
        assert len(rv[k][0]) <= len(rc[k][0]) * 1.01 + 2 * 132 + 16


@pytest.mark.parametrize("mode", [0, 1])
def test_gop_256_vs_oracle(mode, conv_mode):
    """BASELINE.json configs[0] shape (one 4-frame RGB GOP at 256x256): the product free-running
    in both conv modes against the oracle free-running; sequence-average PSNR within 0.02 dB and
    total bits within 0.5 %.  At this size every conv of mode 1 except the 1/64 hyperprior maps
    runs on the tcgen05 kernel."""
    conv_mode(mode)
    h = w = 256
    n, seed, q = 4, 2, 21
    i_net, p_net, sd_i, sd_p = _models(seed, 1, 1)
    o_i, o_p = co.OracleIntra(sd_i, 1), co.OracleDMC(sd_p, 1)
    o_i.update()
    o_p.update()
    frames = synth.synth_frames(n, h, w, seed=seed)
    res = _run_gop(i_net, p_net, frames, h, w, False, q)
    torch.set_num_threads(8)
    ref = []
    with torch.no_grad():
        r = o_i.compress(frames[0:1], False, q)
        ref.append((r["bit_stream"], r["x_hat"]))
        dpbo = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_mv_feature": None,
                "ref_y": None, "ref_mv_y": None}
        for k in range(1, n):
            rp = o_p.compress(frames[k:k + 1], dpbo, False, q, k % 4)
            dpbo = rp["dpb"]
            ref.append((rp["bit_stream"], dpbo["ref_frame"]))
    dp = [psnr(res[k][1], frames[k:k + 1]) - psnr(ref[k][1], frames[k:k + 1]) for k in range(n)]
    bits = sum(len(s) for s, _ in res)
    rbits = sum(len(s) for s, _ in ref)
    assert max(abs(d) for d in dp) < 0.05, dp
    assert abs(sum(dp) / n) < 0.02, dp
    assert abs(bits - rbits) <= 0.005 * rbits, (bits, rbits)


@pytest.mark.parametrize("mode", [0, 1])
def test_pipeline_equals_sequential(mode):
    """dcvc_b200.pipeline.EncDecPipeline (encoder and decoder instances on two CUDA streams, the
    decoder one frame behind) yields byte-identical streams and bit-identical decoded frames to
    the sequential encode -> decode loop of the reference harness (test_video.py:138-165)."""
    from dcvc_b200 import ops
    from dcvc_b200.pipeline import EncDecPipeline
    old = ops.CONV_MODE
    ops.set_conv_mode(mode)
    try:
        h, w, n = 64, 96, 6
        frames = synth.synth_frames(n, h, w, seed=3)
        i_net, p_net, _, _ = _models(5, 2, 2, "interleaved")
        i_dec, p_dec, _, _ = _models(5, 2, 2, "interleaved")
        # intra period 4: I P P P I P
        ref = []
        for g0 in (0, 4):
            ref += _run_gop(i_net, p_net, frames[g0:g0 + 4], h, w, True, 1)
        pipe = EncDecPipeline(i_net, p_net, i_dec, p_dec, 4, True, 1, dev())
        got_streams, got_frames = [], []
        with torch.no_grad():
            for k in range(n):
                with torch.cuda.stream(pipe.enc_stream):
                    x = frames[k:k + 1].to(dev())
                stream, decoded = pipe.push(x)
                got_streams.append(stream)
                if decoded is not None:
                    got_frames.append(decoded)
            got_frames.append(pipe.flush())
            pipe.join()
            torch.cuda.synchronize()
        assert len(got_frames) == n
        for k in range(n):
            assert got_streams[k] == ref[k][0], f"frame {k}: stream differs"
            assert torch.equal(got_frames[k].detach().cpu(), ref[k][1]), f"frame {k}: recon differs"
        assert torch.equal(pipe.enc_dpb["ref_frame"], pipe.dec_dpb["ref_frame"])
    finally:
        ops.set_conv_mode(old)


@pytest.mark.parametrize("coder_mode", ["compat", "interleaved"])
def test_corrupt_frame_raises_in_decompress(coder_mode):
    """A truncated frame must not decode to garbage silently: IntraNoAR.decompress and
    DMC.decompress raise (the decode kernels' error word is checked once per frame)."""
    from dcvc_b200._lib import DcvcError
    h = w = 64
    i_net, p_net, _, _ = _models(0, 2, 2, coder_mode)
    frames = synth.synth_frames(2, h, w, seed=0)
    with torch.no_grad():
        enc = i_net.compress(frames[0:1].to(dev()), True, 1)
        s = enc["bit_stream"]
        with pytest.raises(DcvcError):
            i_net.decompress(s[:len(s) - len(s) // 3], h, w, True, 1)
        dec = i_net.decompress(s, h, w, True, 1)          # the coder recovers for the next frame
        assert torch.equal(dec["x_hat"], enc["x_hat"])
        dpb = {"ref_frame": dec["x_hat"], "ref_feature": None, "ref_mv_feature": None,
               "ref_y": None, "ref_mv_y": None}
        encp = p_net.compress(frames[1:2].to(dev()), dpb, True, 1, 1)
        sp = encp["bit_stream"]
        with pytest.raises(DcvcError):
            p_net.decompress(dpb, sp[:len(sp) - len(sp) // 3], h, w, True, 1, 1)
        decp = p_net.decompress(dpb, sp, h, w, True, 1, 1)
        assert torch.equal(decp["dpb"]["ref_frame"], encp["dpb"]["ref_frame"])


@pytest.mark.parametrize("name", ["a64", "b96x128"])
def test_mode0_compat_stream_bytes_vs_reference_golden(golden_dir, name, conv_mode):
    """fp32-accumulate mode, compat coder, free-running GOP: every frame's stream against the
    bytes the UNMODIFIED reference wrote for the same weights and frames (tests/golden).  Equal
    bytes mean every quantised symbol and every CDF index of the frame equal the reference's."""
    conv_mode(0)
    g = np.load(os.path.join(golden_dir, f"dc_gop_{name}.npz"))
    h, w, n, q_in_ckpt, q_index, part_i, part_p, seed = [int(v) for v in g["meta"]]
    i_net, p_net, _, _ = _models(seed, part_i, part_p)
    frames = synth.synth_frames(n, h, w, seed=seed)
    res = _run_gop(i_net, p_net, frames, h, w, bool(q_in_ckpt), q_index)
    same = [res[k][0] == g[f"stream_{k}"].tobytes() for k in range(n)]
    rms = [float((res[k][1] - torch.from_numpy(g[f"x_hat_{k}"])).pow(2).mean().sqrt())
           for k in range(n)]
    print(f"golden {name}: streams byte-identical per frame {same}, x_hat RMS vs reference {rms}")
    assert all(same), (same, [len(res[k][0]) - g[f"stream_{k}"].size for k in range(n)])
    assert max(rms) < 1e-6, rms


STAGES = ("est_mv", "mv_y", "mv_z_hat", "mv_params", "mv_y_hat", "mv_hat", "ctx1", "ctx2", "ctx3",
          "y", "z_hat", "params", "y_hat", "x_hat")
# per-stage bound on RMS(product - oracle) / RMS(oracle), teacher-forced on the oracle's dpb.
# mode 0: fp32 reassociation only.  mode 1: TF32 / binary16 operands; the quantised stages
# (*_hat, z_hat) inherit symbol flips (one flip changes one latent by 1), hence looser.
# (measured on a B200: mode 0 <= 9.5e-7 everywhere; mode 1 est_mv 1.3e-4, mv_y 6e-4, contexts
# 3-5e-4, y 5e-4, x_hat 8e-4, and 2 flipped mv_y symbols of 18 432 = 5e-2 of mv_y_hat's RMS)
STAGE_BOUND = {0: dict.fromkeys(STAGES, 2e-6),
               1: {"est_mv": 5e-4, "mv_y": 2e-3, "mv_z_hat": 0.1, "mv_params": 1e-4,
                   "mv_y_hat": 0.1, "mv_hat": 4e-2, "ctx1": 2e-3, "ctx2": 2e-3, "ctx3": 2e-3,
                   "y": 2e-3, "z_hat": 0.1, "params": 1e-4, "y_hat": 0.1, "x_hat": 3e-3}}


@pytest.mark.parametrize("mode", [0, 1])
def test_per_stage_taps_vs_oracle(mode, conv_mode):
    """scripts/debug_taps.py as a test: every named stage of a P frame (motion estimate, mv
    latents, prior parameters, contexts, latents, reconstruction) against the oracle's, frames
    teacher-forced on the oracle's dpb.  In mode 0 the coded symbols and CDF indexes of the whole
    frame must be bit-equal to the oracle's (north_star: bit-exact in the fp32-accumulate mode)."""
    from oracle import entropy_oracle as eo
    conv_mode(mode)
    h, w, n, seed, q = 128, 192, 3, 4, 17
    i_net, p_net, sd_i, sd_p = _models(seed, 1, 1)
    o_i, o_p = co.OracleIntra(sd_i, 1), co.OracleDMC(sd_p, 1)
    o_i.update()
    o_p.update()
    frames = synth.synth_frames(n, h, w, seed=seed)
    torch.set_num_threads(8)
    worst = {}
    with torch.no_grad():
        ref = o_i.compress(frames[0:1], False, q)
        dpbo = {"ref_frame": ref["x_hat"], "ref_feature": None, "ref_mv_feature": None,
                "ref_y": None, "ref_mv_y": None}
        for k in range(1, n):
            o_p.taps, p_net.taps = {}, {}
            refp = o_p.compress(frames[k:k + 1], dpbo, False, q, k % 4)
            p_net.compress(frames[k:k + 1].to(dev()), _to_dev(dpbo), False, q, k % 4)
            for name in STAGES:
                a, b = p_net.taps["p_" + name].cpu(), o_p.taps["p_" + name]
                rel = float((a - b).pow(2).mean().sqrt() / b.pow(2).mean().sqrt().clamp_min(1e-12))
                worst[name] = max(worst.get(name, 0.0), rel)
            flips = idx_flips = total = 0
            for lat, sc in (("p_y", "p_scales"), ("p_mv_y", "p_mv_scales")):
                for s in range(4):
                    a = p_net.taps[f"{lat}_q_w_{s}"].cpu().int()
                    b = o_p.taps[f"{lat}_q_w_{s}"].int()
                    ia = p_net.taps[f"{lat}_idx_w_{s}"].cpu().int()
                    ib = eo.build_indexes(o_p.taps[f"{sc}_w_{s}"], "laplace")
                    flips += int((a != b).sum())
                    idx_flips += int((ia != ib).sum())
                    total += a.numel()
            print(f"mode {mode} frame {k}: symbol flips {flips}/{total}, index flips {idx_flips}")
            if mode == 0:
                assert flips == 0 and idx_flips == 0, (k, flips, idx_flips)
            else:
                assert flips <= 1e-3 * total, (k, flips, total)
            dpbo = refp["dpb"]
            p_net.taps = None
    print(f"mode {mode} per-stage rel RMS:", {k_: f"{v:.2e}" for k_, v in worst.items()})
    for name in STAGES:
        assert worst[name] <= STAGE_BOUND[mode][name], (name, worst[name])
