"""Generates tests/golden/*.npz by running the UNMODIFIED reference (imported in place from
/root/reference, native parts from oracle/_ref) on seeded synthetic weights and frames.

Run in the build container only:   python tests/golden/make_golden.py
The reference has no tests or known-answer vectors of its own, so these
files are what pins oracle/ -- and through it the CUDA path -- to the reference.

Contents
  dc_tables.npz      CDF tables of both GaussianEncoders (laplace / gaussian), the factorised
                     prior tables for the seed-0 weights, the 64-point fine q tables.
  rans_kat.npz       (symbols, indexes, tables) -> container bytes from the reference's
                     compiled coder for stream_part in {1,2,4,16}, with forced escapes.
  dc_gop_<cfg>.npz   per frame of one GOP: container bytes, x_hat, bits (real stream and
                     estimate mode) for IntraNoAR + DMC.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402
from dcvc_b200 import arch, synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))

CONFIGS = {
    # name: (height, width, frames, q_in_ckpt, q_index, stream_part_i, stream_part_p, seed)
    "a64": (64, 64, 4, True, 1, 1, 1, 0),
    "b96x128": (96, 128, 3, False, 37, 2, 2, 1),
}


def build_models(ns, seed, part_i, part_p):
    sd_i = arch.synth_state_dict(arch.intra_spec(), seed)
    sd_p = arch.synth_state_dict(arch.dmc_spec(), seed + 100)
    i_net = ns.IntraNoAR(stream_part=part_i)
    p_net = ns.DMC(stream_part=part_p)
    i_net.load_state_dict(sd_i, strict=True)
    p_net.load_state_dict(sd_p, strict=True)
    i_net.eval()
    p_net.eval()
    i_net.update(force=True)
    p_net.update(force=True)
    return i_net, p_net


def tables(ns):
    i_net, p_net = build_models(ns, 0, 1, 1)
    out = {}
    for tag, helper in (("gaussian", i_net.gaussian_encoder), ("laplace", p_net.gaussian_encoder),
                        ("i_z", i_net.bit_estimator_z), ("p_z", p_net.bit_estimator_z),
                        ("p_z_mv", p_net.bit_estimator_z_mv)):
        cdf, length, offset = helper.get_cdf_info()
        out[tag + "_cdf"] = cdf.astype(np.int32)
        out[tag + "_len"] = length.astype(np.int32)
        out[tag + "_off"] = offset.astype(np.int32)
    out["i_q_enc_fine"] = np.asarray(i_net.q_scale_enc_fine, dtype=np.float64)
    out["i_q_dec_fine"] = np.asarray(i_net.q_scale_dec_fine, dtype=np.float64)
    out["p_y_q_enc_fine"] = np.asarray(p_net.y_q_scale_enc_fine, dtype=np.float64)
    out["p_mv_y_q_dec_fine"] = np.asarray(p_net.mv_y_q_scale_dec_fine, dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "dc_tables.npz"), **out)
    return out


def rans_kat(ns, tabs):
    rans, _ = ref_loader.load_native()
    rng = np.random.default_rng(7)
    cdf, length, offset = tabs["laplace_cdf"], tabs["laplace_len"], tabs["laplace_off"]
    out = {}
    for parts in (1, 2, 4, 16):
        n = 4800                                # divisible by 16: the reference overflows otherwise
        idx = rng.integers(0, 256, n).astype(np.int16)
        scale = 0.01 * (64.0 / 0.01) ** (idx / 255.0)
        sym = np.round(rng.laplace(0, scale)).clip(-30000, 30000).astype(np.int16)
        sym[rng.random(n) < 0.02] *= 37         # force escapes (bypass nibbles), both signs
        enc = rans.RansEncoder(False, parts)
        enc.reset()
        enc.encode_with_indexes(sym, idx, cdf, length, offset)
        enc.encode_with_indexes(sym[::-1].copy(), idx[::-1].copy(), cdf, length, offset)
        enc.flush()
        stream = np.asarray(enc.get_encoded_stream(), dtype=np.uint8).copy()
        del enc
        dec = rans.RansDecoder(parts)
        dec.set_stream(stream)
        back = dec.decode_stream(idx, cdf, length, offset)
        assert (back == sym).all()
        out[f"sym_{parts}"] = sym
        out[f"idx_{parts}"] = idx
        out[f"stream_{parts}"] = stream
    np.savez_compressed(os.path.join(OUT, "rans_kat.npz"), **out)


def gop(ns, name):
    h, w, n, q_in_ckpt, q_index, part_i, part_p, seed = CONFIGS[name]
    i_net, p_net = build_models(ns, seed, part_i, part_p)
    frames = synth.synth_frames(n, h, w, seed=seed)
    out = {"meta": np.array([h, w, n, int(q_in_ckpt), q_index, part_i, part_p, seed])}
    with torch.no_grad():
        # real-stream path, as encode_decode does it (video_model.py:526-547, image_model.py:183-190)
        dpb = None
        for k in range(n):
            x = frames[k:k + 1]
            if k == 0:
                enc = i_net.compress(x, q_in_ckpt, q_index)
                dec = i_net.decompress(enc["bit_stream"], h, w, q_in_ckpt, q_index)
                assert torch.equal(enc["x_hat"], dec["x_hat"])
                dpb = {"ref_frame": dec["x_hat"], "ref_feature": None, "ref_mv_feature": None,
                       "ref_y": None, "ref_mv_y": None}
            else:
                enc = p_net.compress(x, dpb, q_in_ckpt, q_index, k % 4)
                dec = p_net.decompress(dpb, enc["bit_stream"], h, w, q_in_ckpt, q_index, k % 4)
                assert torch.equal(enc["dpb"]["ref_frame"], dec["dpb"]["ref_frame"])
                dpb = dec["dpb"]
                out[f"ref_y_{k}"] = dpb["ref_y"].numpy()
                out[f"ref_mv_y_{k}"] = dpb["ref_mv_y"].numpy()
            out[f"stream_{k}"] = np.frombuffer(enc["bit_stream"], dtype=np.uint8).copy()
            out[f"x_hat_{k}"] = dpb["ref_frame"].numpy().copy()
        # estimate path (forward / forward_one_frame), harness aliasing included
        # (test_video.py:141-167: recon clamped in place, and it IS dpb["ref_frame"])
        dpb = None
        for k in range(n):
            x = frames[k:k + 1]
            if k == 0:
                r = i_net.forward(x, q_in_ckpt, q_index)
                out["est_bit_0"] = np.float64(r["bit"].item())
                dpb = {"ref_frame": r["x_hat"].clamp_(0, 1), "ref_feature": None,
                       "ref_mv_feature": None, "ref_y": None, "ref_mv_y": None}
            else:
                r = p_net.forward_one_frame(x, dpb, q_in_ckpt=q_in_ckpt, q_index=q_index,
                                            frame_idx=k % 4)
                out[f"est_bit_{k}"] = np.float64(r["bit"].item())
                out[f"est_bpp_parts_{k}"] = np.array([r["bpp_y"].item(), r["bpp_z"].item(),
                                                      r["bpp_mv_y"].item(), r["bpp_mv_z"].item()])
                dpb = r["dpb"]
            out[f"est_x_hat_{k}"] = dpb["ref_frame"].numpy().copy()
    np.savez_compressed(os.path.join(OUT, f"dc_gop_{name}.npz"), **out)
    print(name, {k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items()
                 if k.startswith("stream") or k.startswith("est_bit")})


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    ns = ref_loader.import_reference("DCVC-DC")
    tabs = tables(ns)
    rans_kat(ns, tabs)
    for name in CONFIGS:
        gop(ns, name)


if __name__ == "__main__":
    main()
