"""CPU: the DCVC-HEM oracle (oracle/hem_oracle.py) held to golden vectors produced by the
UNMODIFIED reference (tests/golden/make_golden_hem.py): streams byte-identical, reconstructions
bit-equal, estimate-mode bits equal -- the HEM oracle is pinned, not merely plausible.  Also the
host-side HEM pieces of the product: state_dict inventory, frame files, q rounding."""
import os

import numpy as np
import pytest
import torch

from dcvc_b200 import synth
from dcvc_b200.hem import arch as harch
from dcvc_b200.hem import stream_helper as sh
from oracle import hem_oracle as ho


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_hem_oracle_equals_reference_golden(golden_dir, name):
    g = np.load(os.path.join(golden_dir, f"hem_gop_{name}.npz"))
    h, w, n, seed = [int(v) for v in g["meta"]]
    gain, q_i, q_mv, q_y = [float(v) for v in g["gain_q"]]
    sd_i = harch.hem_synth_state_dict(harch.hem_intra_spec(), seed, gain)
    sd_p = harch.hem_synth_state_dict(harch.hem_dmc_spec(), seed + 100, gain)
    o_i, o_p = ho.OracleHemIntra(sd_i), ho.OracleHemDMC(sd_p)
    o_i.update()
    o_p.update()
    frames = synth.synth_frames(n, h, w, seed=seed)
    torch.set_num_threads(8)
    with torch.no_grad():
        qi, _ = ho.get_rounded_q(q_i)
        enc = o_i.compress(frames[0:1], qi)
        assert enc["bit_stream"] == g["stream_0"].tobytes()
        dec = o_i.decompress(enc["bit_stream"], h, w, qi)
        assert np.array_equal(dec["x_hat"].numpy(), g["x_hat_0"])
        est = o_i.forward(frames[0:1], q_i)
        assert abs(est["bit"] - float(g["est_bit_0"][0])) <= 1e-4 * float(g["est_bit_0"][0])
        dpb = {"ref_frame": dec["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
        qm, _ = ho.get_rounded_q(q_mv)
        qy, _ = ho.get_rounded_q(q_y)
        for k in (1, 2):
            est = o_p.forward_one_frame(frames[k:k + 1], dpb, q_mv, q_y)
            ref = g[f"est_bit_{k}"]
            for i, key in enumerate(("bit", "bit_y", "bit_z", "bit_mv_y", "bit_mv_z")):
                assert abs(float(est[key]) - ref[i]) <= 1e-4 * ref[i] + 1e-3, (k, key)
            encp = o_p.compress(frames[k:k + 1], dpb, qm, qy)
            assert encp["bit_stream"] == g[f"stream_{k}"].tobytes(), f"frame {k}"
            decp = o_p.decompress(dpb, encp["bit_stream"], h, w, qm, qy)
            assert np.array_equal(decp["dpb"]["ref_frame"].numpy(), g[f"x_hat_{k}"])
            dpb = decp["dpb"]


def test_hem_state_dict_inventory_and_init():
    from dcvc_b200 import hem
    i_net, p_net = hem.IntraNoAR(), hem.DMC()
    assert list(i_net.state_dict().keys()) and len(i_net.state_dict()) == 175
    assert len(p_net.state_dict()) == 406
    for net, spec in ((i_net, harch.hem_intra_spec()), (p_net, harch.hem_dmc_spec())):
        sd = net.state_dict()
        assert set(sd) == set(spec)
        assert all(tuple(sd[k].shape) == tuple(spec[k]) for k in spec)
    assert tuple(p_net.state_dict()["recon_generation_net.unet_1.conv3.conv.3.fc.0.weight"].shape) == (8, 128)


def test_hem_frame_files_and_rounded_q(tmp_path):
    assert sh.get_rounded_q(0.7312) == (0.73, 73)
    assert sh.get_rounded_q(1e-4) == (0.01, 1) and sh.get_rounded_q(1e4) == (655.0, 65500)
    assert sh.get_rounded_q(1.005)[1] == ho.get_rounded_q(1.005)[1]
    p = str(tmp_path / "f.bin")
    sh.encode_i(1080, 1920, 154, b"\x01\x02\x03", p)
    assert sh.filesize(p) == 8 + 2 + 4 + 3
    assert sh.decode_i(p) == (1080, 1920, 154, b"\x01\x02\x03")
    sh.encode_p(b"abcd", 101, 71, p)
    assert sh.filesize(p) == 4 + 4 + 4
    assert sh.decode_p(p) == (101, 71, b"abcd")
    with pytest.raises(ValueError):
        sh.filesize(str(tmp_path / "missing.bin"))
