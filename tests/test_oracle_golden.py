"""Pins oracle/ against golden vectors produced by the UNMODIFIED reference
(tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from dcvc_b200 import arch, synth
from oracle import codec_oracle as co
from oracle import entropy_oracle as eo
from oracle import rans_oracle


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_gaussian_tables_match_reference(golden_dir):
    g = _load(golden_dir, "dc_tables.npz")
    for dist in ("laplace", "gaussian"):
        cdf, sizes, offs = eo.gaussian_tables(dist)
        assert np.array_equal(cdf, g[dist + "_cdf"])
        assert np.array_equal(sizes, g[dist + "_len"])
        assert np.array_equal(offs, g[dist + "_off"])


def test_factorized_tables_match_reference(golden_dir):
    g = _load(golden_dir, "dc_tables.npz")
    sd_i = arch.synth_state_dict(arch.intra_spec(), 0)
    sd_p = arch.synth_state_dict(arch.dmc_spec(), 100)
    for tag, sd, p in (("i_z", sd_i, "bit_estimator_z"), ("p_z", sd_p, "bit_estimator_z"),
                       ("p_z_mv", sd_p, "bit_estimator_z_mv")):
        cdf, sizes, offs = eo.factorized_tables(sd, p)
        assert np.array_equal(cdf, g[tag + "_cdf"]), tag
        assert np.array_equal(sizes, g[tag + "_len"]), tag
        assert np.array_equal(offs, g[tag + "_off"]), tag


def test_fine_q_tables(golden_dir):
    g = _load(golden_dir, "dc_tables.npz")
    sd_i = arch.synth_state_dict(arch.intra_spec(), 0)
    sd_p = arch.synth_state_dict(arch.dmc_spec(), 100)
    assert np.array_equal(co.fine_q_table(sd_i["q_scale_enc"]), g["i_q_enc_fine"])
    assert np.array_equal(co.fine_q_table(sd_i["q_scale_dec"]), g["i_q_dec_fine"])
    assert np.array_equal(co.fine_q_table(sd_p["y_q_scale_enc"]), g["p_y_q_enc_fine"])
    assert np.array_equal(co.fine_q_table(sd_p["mv_y_q_scale_dec"]), g["p_mv_y_q_dec_fine"])


@pytest.mark.parametrize("parts", [1, 2, 4, 16])
def test_rans_known_answer(golden_dir, parts):
    g = _load(golden_dir, "rans_kat.npz")
    t = _load(golden_dir, "dc_tables.npz")
    tabs = (t["laplace_cdf"], t["laplace_len"], t["laplace_off"])
    sym, idx, stream = g[f"sym_{parts}"], g[f"idx_{parts}"], g[f"stream_{parts}"]
    enc = rans_oracle.RansEncoder(False, parts)
    enc.reset()
    enc.encode_with_indexes(sym, idx, *tabs)
    enc.encode_with_indexes(sym[::-1].copy(), idx[::-1].copy(), *tabs)
    enc.flush()
    out = enc.get_encoded_stream()
    assert np.array_equal(out, stream)
    dec = rans_oracle.RansDecoder(parts)
    dec.set_stream(stream.tobytes())
    assert np.array_equal(dec.decode_stream(idx, *tabs), sym)
    assert np.array_equal(dec.decode_stream(idx[::-1].copy(), *tabs), sym[::-1])


def test_rans_ragged_and_empty():
    """n % parts != 0 (the reference overflows a heap buffer there; the format is still
    defined), n < parts, and n == 0."""
    tabs = eo.gaussian_tables("laplace")
    rng = np.random.default_rng(3)
    for parts in (1, 3, 16):
        for n in (0, 1, 5, 1001):
            idx = rng.integers(0, 256, n).astype(np.int16)
            sym = rng.integers(-60, 60, n).astype(np.int16)
            enc = rans_oracle.RansEncoder(False, parts)
            enc.reset()
            enc.encode_with_indexes(sym, idx, *tabs)
            enc.flush()
            s = enc.get_encoded_stream()
            assert s.size >= 1 + 8 * parts
            dec = rans_oracle.RansDecoder(parts)
            dec.set_stream(s.tobytes())
            assert np.array_equal(dec.decode_stream(idx, *tabs), sym)


@pytest.mark.parametrize("name", ["a64", "b96x128"])
def test_gop_matches_reference(golden_dir, name):
    """Oracle compress/decompress/forward vs the reference: streams byte-identical,
    reconstructions equal (same torch build, same op order => same floats)."""
    g = _load(golden_dir, f"dc_gop_{name}.npz")
    h, w, n, q_in_ckpt, q_index, part_i, part_p, seed = [int(v) for v in g["meta"]]
    q_in_ckpt = bool(q_in_ckpt)
    i_net = co.OracleIntra(arch.synth_state_dict(arch.intra_spec(), seed), part_i)
    p_net = co.OracleDMC(arch.synth_state_dict(arch.dmc_spec(), seed + 100), part_p)
    i_net.update()
    p_net.update()
    frames = synth.synth_frames(n, h, w, seed=seed)
    torch.set_num_threads(8)
    with torch.no_grad():
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
            assert np.frombuffer(enc["bit_stream"], np.uint8).tolist() == \
                g[f"stream_{k}"].tolist(), f"frame {k} stream differs"
            np.testing.assert_array_equal(dpb["ref_frame"].numpy(), g[f"x_hat_{k}"])
        dpb = None
        for k in range(n):
            x = frames[k:k + 1]
            if k == 0:
                r = i_net.forward(x, q_in_ckpt, q_index)
                dpb = {"ref_frame": r["x_hat"].clamp_(0, 1), "ref_feature": None,
                       "ref_mv_feature": None, "ref_y": None, "ref_mv_y": None}
            else:
                r = p_net.forward_one_frame(x, dpb, q_in_ckpt, q_index, k % 4)
                dpb = r["dpb"]
            assert abs(float(r["bit"]) - float(g[f"est_bit_{k}"])) <= 1e-3 * float(g[f"est_bit_{k}"])
            np.testing.assert_allclose(dpb["ref_frame"].numpy(), g[f"est_x_hat_{k}"], atol=1e-6)


@pytest.mark.parametrize("warps,n", [(1, 0), (1, 1), (1, 33), (2, 100), (4, 1000)])
def test_interleaved_restatement_round_trips(warps, n):
    """The CPU restatement of the product's interleaved container is lossless (escapes included)
    and its size stays within the per-warp state overhead of the reference coder's size."""
    from oracle import rans_interleaved_oracle as rio
    tabs = eo.gaussian_tables("laplace")
    rng = np.random.default_rng(n + warps)
    calls = []
    for m in (n, n // 3 + 1):
        idx = rng.integers(0, 256, m).astype(np.int16)
        scale = 0.01 * (64.0 / 0.01) ** (idx / 255.0)
        sym = np.round(rng.laplace(0, scale)).clip(-30000, 30000).astype(np.int16)
        sym[rng.random(m) < 0.03] *= 41
        calls.append((sym, idx))
    stream = rio.encode(calls, tabs, warps)
    assert stream[0] == ((warps - 1) << 4 | 0xF)
    back = rio.decode(stream, [c[1] for c in calls], tabs)
    for (sym, _), out in zip(calls, back):
        assert np.array_equal(sym, out)
    enc = rans_oracle.RansEncoder(False, 1)
    enc.reset()
    for sym, idx in calls:
        enc.encode_with_indexes(sym, idx, *tabs)
    enc.flush()
    assert len(stream) <= enc.get_encoded_stream().size * 1.02 + 132 * warps + 8
