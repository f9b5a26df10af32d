"""test_video.py / dcvc_b200.harness: the reference-facing entry point.  CPU part: flag parsing,
rate selection, job layout, 4:2:0 helpers against the reference's scipy/numpy definitions, file
readers/writers, summary keys.  GPU part: the whole entry point on a synthetic PNG and YUV
sequence with stream files, against the oracle run on the same frames."""
import json
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)

import test_video as tv  # noqa: E402
from dcvc_b200 import arch, harness, synth  # noqa: E402


def _config(tmp_path, src_type, h, w, frames, gop, names=("seqA",)):
    cfg = {"root_path": str(tmp_path), "test_classes": {
        "SYN": {"test": 1, "base_path": "data", "src_type": src_type,
                "sequences": {n: {"width": w, "height": h, "frames": frames, "gop": gop}
                              for n in names}},
        "OFF": {"test": 0, "base_path": "none", "src_type": "png", "sequences": {}}}}
    path = tmp_path / "cfg.json"
    path.write_text(json.dumps(cfg))
    return cfg, str(path)


def test_rate_selection_and_jobs(tmp_path):
    cfg, path = _config(tmp_path, "png", 64, 96, 5, 4, ("a", "b"))
    args = tv.parse_args(["--test_config", path, "--output_path", "o.json", "--rate_num", "4"])
    assert tv.rate_indexes(args, [1, 2, 3, 4]) == (True, [0, 1, 2, 3])
    args = tv.parse_args(["--test_config", path, "--output_path", "o.json", "--rate_num", "6",
                          "--force_intra_period", "2", "--force_frame_num", "3",
                          "--write_stream", "1", "--stream_path", str(tmp_path / "bins")])
    q_in_ckpt, idx = tv.rate_indexes(args, [1, 2, 3, 4])
    assert q_in_ckpt is False and idx == [0, 13, 25, 38, 50, 63]
    jobs = tv.build_jobs(args, cfg, q_in_ckpt, idx)
    assert len(jobs) == 2 * 6
    j = jobs[7]
    assert j["video_path"] == "b" and j["rate_idx"] == 1 and j["i_frame_q_index"] == 13
    assert j["gop_size"] == 2 and j["frame_num"] == 3 and j["write_stream"]
    assert j["bin_folder"].endswith(os.path.join("bins", "b", "1"))
    assert j["src_path"].endswith(os.path.join("data", "b"))
    args = tv.parse_args(["--test_config", path, "--output_path", "o.json", "--rate_num", "2",
                          "--i_frame_q_indexes", "7", "9"])
    assert tv.rate_indexes(args, [1, 2, 3, 4]) == (False, [7, 9])


def test_yuv420_helpers_match_reference_definitions():
    scipy_ndimage = pytest.importorskip("scipy.ndimage")
    rng = np.random.default_rng(0)
    y = rng.integers(0, 256, (36, 52), dtype=np.uint8)
    uv = rng.integers(0, 256, (2, 18, 26), dtype=np.uint8)
    x, y_t, uv_t = harness.yuv420_u8_to_x(y, uv, torch.device("cpu"))
    ref_uv = scipy_ndimage.zoom(uv.astype(np.float32) / 255, (1, 2, 2), order=0)
    assert np.array_equal(x[0, 1:].numpy(), ref_uv)
    assert np.array_equal(x[0, 0].numpy(), y.astype(np.float32) / 255)
    yuv = rng.random((3, 36, 52)).astype(np.float32) * 1.2 - 0.1
    y2, uv2 = harness.x_to_yuv420(torch.from_numpy(yuv)[None])
    ref_u = np.mean(yuv[1].reshape(18, 2, 26, 2), axis=(-1, -3))
    assert np.allclose(uv2[0].numpy(), np.clip(ref_u, 0, 1), atol=1e-6)
    assert np.array_equal(y2.numpy(), np.clip(yuv[0], 0, 1))
    a = torch.from_numpy(yuv[0])
    assert harness.psnr(a, a) == 999.9
    mse = float(np.mean((yuv[0].astype(np.float64) - yuv[1].astype(np.float64)) ** 2))
    assert abs(harness.psnr(a, torch.from_numpy(yuv[1])) - 10 * np.log10(1 / mse)) < 1e-9


def test_readers_and_writers_round_trip(tmp_path):
    frames = (synth.synth_frames(2, 32, 48, seed=3) * 255).round().to(torch.uint8)
    folder = tmp_path / "png"
    w = harness.PNGWriter(str(folder), 48, 32)
    for k in range(2):
        w.write_rgb(frames[k:k + 1].float() / 255)
    r = harness.PNGReader(str(folder), 48, 32)
    for k in range(2):
        assert np.array_equal(r.read(), frames[k].numpy())
    assert r.read() is None
    yw = harness.YUV420Writer(str(tmp_path / "clip"), 48, 32)
    planes = []
    for k in range(2):
        y, uv = harness.x_to_yuv420(frames[k:k + 1].float() / 255)
        yw.write_420(y, uv)
        planes.append(((y * 255).round().to(torch.uint8).numpy(), (uv * 255).round().to(torch.uint8).numpy()))
    yw.close()
    yr = harness.YUV420Reader(str(tmp_path / "clip"), 48, 32)
    for k in range(2):
        y, uv = yr.read()
        assert np.array_equal(y, planes[k][0]) and np.array_equal(uv, planes[k][1])
    assert yr.read() is None
    with pytest.raises(ValueError):
        harness.PNGReader(str(tmp_path), 48, 32)


def test_summary_keys_and_arithmetic():
    s = harness.summarise([0, 1, 1, 0], [800.0, 100.0, 300.0, 1000.0], [40.0, 30.0, 32.0, 42.0],
                          [0.0] * 4, 100, 1.5)
    assert s["i_frame_num"] == 2 and s["p_frame_num"] == 2 and s["frame_pixel_num"] == 100
    assert s["ave_i_frame_bpp"] == 9.0 and s["ave_p_frame_bpp"] == 2.0
    assert s["ave_all_frame_bpp"] == 5.5 and s["ave_all_frame_psnr"] == 36.0
    assert s["ave_i_frame_psnr"] == 41.0 and s["ave_p_frame_psnr"] == 31.0
    want = {"frame_pixel_num", "i_frame_num", "p_frame_num", "test_time"} | {
        f"ave_{t}_frame_{m}" for t in ("i", "p", "all") for m in ("bpp", "psnr", "msssim")}
    assert set(s) == want
    s = harness.summarise([0], [80.0], [40.0], [0.0], 10, 0.1,
                          {f"{m}_{c}": [1.0] for m in ("psnr", "msssim") for c in "yuv"})
    assert s["ave_p_frame_bpp"] == 0 and s["ave_p_frame_psnr_u"] == 0
    assert s["ave_all_frame_psnr_y"] == 1.0 and "ave_i_frame_msssim_v" in s


def test_merge_units_equals_whole_sequence_summary():
    """GOP units coded on different ranks merge into the summary one process would write."""
    ft, bits = [0, 1, 1, 1, 0, 1, 1, 1], [800.0, 90, 80, 70, 810.0, 95, 85, 75]
    ps = [40.0, 38, 37, 36, 41.0, 39, 38, 37]
    planes = {f"{m}_{c}": [float(i) for i in range(8)] for m in ("psnr", "msssim") for c in "yuv"}
    whole = harness.summarise(ft, bits, ps, [0.0] * 8, 100, 2.0, planes)

    def part(a, b, t):
        return {"first_frame": a, "frame_types": ft[a:b], "bits": bits[a:b], "psnrs": ps[a:b],
                "msssims": [0.0] * (b - a), "planes": {k: v[a:b] for k, v in planes.items()},
                "pixels": 100, "test_time": t}
    merged = harness.merge_units([part(4, 8, 1.25), part(0, 4, 0.75)])
    assert merged == whole


def _checkpoints(tmp_path):
    paths = []
    for name, spec, seed in (("i", arch.intra_spec(), 0), ("p", arch.dmc_spec(), 100)):
        path = str(tmp_path / f"{name}.pth.tar")
        torch.save({"state_dict": arch.synth_state_dict(spec, seed)}, path)
        paths.append(path)
    return paths


@pytest.mark.gpu
def test_entry_point_png_with_streams_vs_oracle(tmp_path):
    from oracle import codec_oracle as co
    h, w, n = 64, 96, 3
    frames = (synth.synth_frames(n, h, w, seed=5) * 255).round().to(torch.uint8)
    pw = harness.PNGWriter(str(tmp_path / "data" / "seqA"), w, h)
    for k in range(n):
        pw.write_rgb(frames[k:k + 1].float() / 255)
    _, cfg_path = _config(tmp_path, "png", h, w, n, 4)
    ip, pp = _checkpoints(tmp_path)
    out = str(tmp_path / "res" / "out.json")
    tv.main(["--i_frame_model_path", ip, "--p_frame_model_path", pp, "--rate_num", "2",
             "--i_frame_q_indexes", "20", "40", "--test_config", cfg_path, "--cuda", "1",
             "--write_stream", "1", "--stream_path", str(tmp_path / "bins"), "--output_path", out])
    log = json.load(open(out))
    assert set(log) == {"SYN"} and set(log["SYN"]["seqA"]) == {"000", "001"}
    rec = log["SYN"]["seqA"]["000"]
    sizes = [os.path.getsize(tmp_path / "bins" / "seqA" / "0" / f"{k}.bin") for k in range(n)]
    assert abs(rec["ave_all_frame_bpp"] - sum(sizes) * 8 / (n * h * w)) < 1e-5
    assert rec["i_frame_num"] == 1 and rec["p_frame_num"] == 2
    # the oracle on the same 8-bit frames, same weights and rate
    sd_i = arch.synth_state_dict(arch.intra_spec(), 0)
    sd_p = arch.synth_state_dict(arch.dmc_spec(), 100)
    o_i, o_p = co.OracleIntra(sd_i, 1), co.OracleDMC(sd_p, 1)
    o_i.update()
    o_p.update()
    x = frames.float() / 255
    torch.set_num_threads(8)
    psnrs, bits = [], 0
    with torch.no_grad():
        r = o_i.compress(x[0:1], False, 20)
        bits += (len(r["bit_stream"]) + 13) * 8
        dpb = {"ref_frame": r["x_hat"].clamp(0, 1), "ref_feature": None, "ref_mv_feature": None,
               "ref_y": None, "ref_mv_y": None}
        psnrs.append(harness.psnr(dpb["ref_frame"], x[0:1]))
        for k in range(1, n):
            rp = o_p.compress(x[k:k + 1], dpb, False, 20, k % 4)
            bits += (len(rp["bit_stream"]) + 6) * 8
            dpb = rp["dpb"]
            dpb["ref_frame"] = dpb["ref_frame"].clamp(0, 1)
            psnrs.append(harness.psnr(dpb["ref_frame"], x[k:k + 1]))
    assert abs(rec["ave_all_frame_psnr"] - sum(psnrs) / n) < 0.02
    assert abs(rec["ave_all_frame_bpp"] - bits / (n * h * w)) <= 0.005 * bits / (n * h * w)


@pytest.mark.gpu
def test_entry_point_yuv420_vs_oracle(tmp_path):
    """--yuv420 1 through the entry point (YUV file -> pinned upload -> device 4:2:0 -> 4:4:4,
    coding with stream files, device 4:4:4 -> 4:2:0 + per-plane SSE) against the oracle run on the
    same planes with the reference's own numpy definitions of the colour handling and PSNR
    (functional.py:61-92, metrics.py:81-92, test_video.py:169-195)."""
    scipy_ndimage = pytest.importorskip("scipy.ndimage")
    from oracle import codec_oracle as co
    h, w, n = 64, 96, 3
    frames = synth.synth_frames(n, h, w, seed=6)
    os.makedirs(tmp_path / "data")
    yw = harness.YUV420Writer(str(tmp_path / "data" / "seqA"), w, h)
    for k in range(n):
        y, uv = harness.x_to_yuv420(synth.rgb_to_ycbcr444(frames[k:k + 1]))
        yw.write_420(y, uv)
    yw.close()
    _, cfg_path = _config(tmp_path, "yuv420", h, w, n, 4)
    ip, pp = _checkpoints(tmp_path)
    out = str(tmp_path / "out.json")
    tv.main(["--i_frame_model_path", ip, "--p_frame_model_path", pp, "--rate_num", "2",
             "--i_frame_q_indexes", "20", "40", "--test_config", cfg_path, "--yuv420", "1",
             "--cuda", "1", "--write_stream", "1", "--stream_path", str(tmp_path / "bins"),
             "--output_path", out])
    rec = json.load(open(out))["SYN"]["seqA"]["000"]
    want = (6 * rec["ave_all_frame_psnr_y"] + rec["ave_all_frame_psnr_u"] + rec["ave_all_frame_psnr_v"]) / 8
    assert abs(rec["ave_all_frame_psnr"] - want) < 1e-4
    # the oracle on the same file
    reader = harness.YUV420Reader(str(tmp_path / "data" / "seqA"), w, h)
    sd_i = arch.synth_state_dict(arch.intra_spec(), 0)
    sd_p = arch.synth_state_dict(arch.dmc_spec(), 100)
    o_i, o_p = co.OracleIntra(sd_i, 1), co.OracleDMC(sd_p, 1)
    o_i.update()
    o_p.update()
    torch.set_num_threads(8)

    def calc_psnr(a, b):
        return 10 * np.log10(1.0 / np.mean(np.square(a.astype(np.float64) - b.astype(np.float64))))

    ps = {c: [] for c in "yuv"}
    bits, dpb = 0, None
    with torch.no_grad():
        for k in range(n):
            y8, uv8 = reader.read()
            yf, uvf = y8.astype(np.float32) / 255, uv8.astype(np.float32) / 255
            yuv = np.concatenate((yf[None], scipy_ndimage.zoom(uvf, (1, 2, 2), order=0)), 0)
            x = torch.from_numpy(yuv)[None]
            if k == 0:
                r = o_i.compress(x, False, 20)
                bits += (len(r["bit_stream"]) + 13) * 8
                dpb = {"ref_frame": r["x_hat"].clamp(0, 1), "ref_feature": None,
                       "ref_mv_feature": None, "ref_y": None, "ref_mv_y": None}
            else:
                rp = o_p.compress(x, dpb, False, 20, k % 4)
                bits += (len(rp["bit_stream"]) + 6) * 8
                dpb = rp["dpb"]
                dpb["ref_frame"] = dpb["ref_frame"].clamp(0, 1)
            rec444 = dpb["ref_frame"][0].numpy()
            u = np.mean(np.reshape(rec444[1:2], (1, h // 2, 2, w // 2, 2)), axis=(-1, -3))
            v = np.mean(np.reshape(rec444[2:3], (1, h // 2, 2, w // 2, 2)), axis=(-1, -3))
            ps["y"].append(calc_psnr(yf, np.clip(rec444[0], 0, 1)))
            ps["u"].append(calc_psnr(uvf[0], np.clip(u[0], 0, 1)))
            ps["v"].append(calc_psnr(uvf[1], np.clip(v[0], 0, 1)))
    for c in "yuv":
        assert abs(rec[f"ave_all_frame_psnr_{c}"] - sum(ps[c]) / n) < 0.02, c
    assert abs(rec["ave_all_frame_bpp"] - bits / (n * h * w)) <= 0.005 * bits / (n * h * w)


@pytest.mark.gpu
def test_entry_point_graphs_equal_eager(tmp_path):
    """test_video.py --graphs 1 (P frames replayed from CUDA graphs) writes the same stream files
    and the same results as the eager launches; 14 frames so that graphs are recorded and replayed."""
    h, w, n = 64, 96, 14
    frames = (synth.synth_frames(n, h, w, seed=7) * 255).round().to(torch.uint8)
    pw = harness.PNGWriter(str(tmp_path / "data" / "seqA"), w, h)
    for k in range(n):
        pw.write_rgb(frames[k:k + 1].float() / 255)
    _, cfg_path = _config(tmp_path, "png", h, w, n, 32)
    ip, pp = _checkpoints(tmp_path)
    recs = []
    for g in ("0", "1"):
        out = str(tmp_path / f"res{g}" / "out.json")
        tv.main(["--i_frame_model_path", ip, "--p_frame_model_path", pp, "--rate_num", "1",
                 "--i_frame_q_indexes", "30", "--test_config", cfg_path, "--cuda", "1",
                 "--write_stream", "1", "--stream_path", str(tmp_path / f"bins{g}"),
                 "--graphs", g, "--output_path", out])
        recs.append(json.load(open(out))["SYN"]["seqA"]["000"])
    for k in range(n):
        a = open(tmp_path / "bins0" / "seqA" / "0" / f"{k}.bin", "rb").read()
        b = open(tmp_path / "bins1" / "seqA" / "0" / f"{k}.bin", "rb").read()
        assert a == b, f"frame {k}: stream differs under graphs"
    for key in ("ave_all_frame_bpp", "ave_all_frame_psnr", "frame_psnr", "frame_bpp"):
        if key in recs[0]:
            assert recs[0][key] == recs[1][key], key
