"""Generates tests/golden/hem_gop_*.npz by running the UNMODIFIED DCVC-HEM reference
(/root/reference/DCVC-HEM, imported in place by oracle/ref_loader.py with its compiled coder from
oracle/_ref/hem) in the build container:  python tests/golden/make_golden_hem.py

Each file: one GOP (I + 2 P) at 64x64 / 64x128 through compress -> decompress with real streams,
for a (seed, weight gain, q_scales) combination -- gain 0.5 and 1.0 (HEM's init regime has many
bypass symbols), q_scales with and without more than two decimals (the stream carries
round(q * 100), stream_helper.py:41-45) -- plus the estimate-mode bits of the same frames.
DC and HEM both name their package `src`: run this in its own process."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)

from dcvc_b200 import synth  # noqa: E402
from dcvc_b200.hem import arch as harch  # noqa: E402
from oracle import ref_loader  # noqa: E402

CASES = [  # name, h, w, seed, gain, q_i, q_mv, q_y
    ("a", 64, 64, 0, 0.5, 1.08, 1.10, 0.96),
    ("b", 64, 128, 1, 1.0, 0.7312, 1.0149, 0.7134),
    ("c", 128, 64, 2, 0.5, 1.541, 0.919, 0.532),
]


def main():
    ref = ref_loader.import_reference("DCVC-HEM")
    from src.utils.stream_helper import get_rounded_q
    torch.manual_seed(0)
    torch.set_num_threads(8)
    for name, h, w, seed, gain, q_i, q_mv, q_y in CASES:
        sd_i = harch.hem_synth_state_dict(harch.hem_intra_spec(), seed, gain)
        sd_p = harch.hem_synth_state_dict(harch.hem_dmc_spec(), seed + 100, gain)
        i_net, p_net = ref.IntraNoAR(), ref.DMC()
        i_net.load_state_dict(sd_i)
        p_net.load_state_dict(sd_p)
        i_net, p_net = i_net.eval(), p_net.eval()
        i_net.update(force=True)
        p_net.update(force=True)
        frames = synth.synth_frames(3, h, w, seed=seed)
        out = {"meta": np.array([h, w, 3, seed], dtype=np.int64),
               "gain_q": np.array([gain, q_i, q_mv, q_y], dtype=np.float64)}
        with torch.no_grad():
            qi, _ = get_rounded_q(q_i)
            enc = i_net.compress(frames[0:1], qi)
            dec = i_net.decompress(enc["bit_stream"], h, w, qi)
            out["stream_0"] = np.frombuffer(enc["bit_stream"], dtype=np.uint8)
            out["x_hat_0"] = dec["x_hat"].numpy()
            est = i_net.forward(frames[0:1], q_i)             # estimate mode: q NOT rounded
            out["est_bit_0"] = np.array([float(est["bit"])])
            dpb = {"ref_frame": dec["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
            qm, _ = get_rounded_q(q_mv)
            qy, _ = get_rounded_q(q_y)
            for k in (1, 2):
                est = p_net.forward_one_frame(frames[k:k + 1], dpb, mv_y_q_scale=q_mv, y_q_scale=q_y)
                out[f"est_bit_{k}"] = np.array([float(est[n]) for n in
                                                ("bit", "bit_y", "bit_z", "bit_mv_y", "bit_mv_z")])
                enc = p_net.compress(frames[k:k + 1], dpb, qm, qy)
                dec = p_net.decompress(dpb, enc["bit_stream"], h, w, qm, qy)
                assert torch.equal(enc["dbp"]["ref_frame"].clamp(0, 1), dec["dpb"]["ref_frame"])
                out[f"stream_{k}"] = np.frombuffer(enc["bit_stream"], dtype=np.uint8)
                out[f"x_hat_{k}"] = dec["dpb"]["ref_frame"].numpy()
                dpb = dec["dpb"]
        path = os.path.join(ROOT, "tests", "golden", f"hem_gop_{name}.npz")
        np.savez_compressed(path, **out)
        print(name, [out[f"stream_{k}"].size for k in range(3)], os.path.getsize(path))


if __name__ == "__main__":
    main()
