"""-m gpu: CUDA-graph replay of the P-frame encode / decode (dcvc_b200.graphs.GraphedDMC) against
the eager launch sequence: byte-identical streams, bit-identical dpb entries, over two GOPs so that
every graph variant (first P frame after an intra frame, the three feature-adaptor indexes) is
warmed, recorded and replayed several times; also through the two-stream pipeline."""
import pytest
import torch

import dcvc_b200
from dcvc_b200 import arch, ops, synth
from dcvc_b200.graphs import GraphedDMC
from dcvc_b200.pipeline import EncDecPipeline

from gpu_util import dev

pytestmark = pytest.mark.gpu
KEYS = ("ref_frame", "ref_feature", "ref_mv_feature", "ref_y", "ref_mv_y")


def _models(mode="interleaved", parts=2):
    i_net = dcvc_b200.IntraNoAR(stream_part=parts, coder_mode=mode)
    p_net = dcvc_b200.DMC(stream_part=parts, coder_mode=mode)
    i_net.load_state_dict(arch.synth_state_dict(arch.intra_spec(), 5))
    p_net.load_state_dict(arch.synth_state_dict(arch.dmc_spec(), 105))
    i_net, p_net = i_net.to(dev()).eval(), p_net.to(dev()).eval()
    i_net.update(force=True)
    p_net.update(force=True)
    return i_net, p_net


def _sequence(i_net, p_enc, p_dec, frames, h, w, period):
    out = []
    dpb = None
    with torch.no_grad():
        for k in range(frames.shape[0]):
            x = frames[k:k + 1].to(dev())
            if k % period == 0:
                enc = i_net.compress(x, True, 1)
                dec = i_net.decompress(enc["bit_stream"], h, w, True, 1)
                dpb = {"ref_frame": dec["x_hat"], "ref_feature": None, "ref_mv_feature": None,
                       "ref_y": None, "ref_mv_y": None}
                out.append((enc["bit_stream"], {"ref_frame": dec["x_hat"].clone()}))
            else:
                enc = p_enc.compress(x, dpb, True, 1, k % 4)
                dec = p_dec.decompress(dpb, enc["bit_stream"], h, w, True, 1, k % 4)
                for key in KEYS:
                    assert torch.equal(enc["dpb"][key], dec["dpb"][key]), (k, key)
                dpb = dec["dpb"]
                out.append((enc["bit_stream"], {key: dpb[key].clone() for key in KEYS}))
    return out


@pytest.mark.parametrize("mode", [0, 1])
def test_graph_replay_equals_eager(mode):
    prev = ops.CONV_MODE
    ops.set_conv_mode(mode)
    try:
        h, w, n, period = 128, 192, 20, 10
        frames = synth.synth_frames(n, h, w, seed=7)
        i_net, p_net = _models()
        ref = _sequence(i_net, p_net, p_net, frames, h, w, period)
        g = GraphedDMC(p_net)
        got = _sequence(i_net, g, g, frames, h, w, period)
        assert g.replays >= 2 * (n - 2) - 2 * 8          # most P frames were replays
        for k in range(n):
            assert got[k][0] == ref[k][0], f"frame {k}: stream differs"
            for key, t in ref[k][1].items():
                assert torch.equal(got[k][1][key], t), f"frame {k}: {key} differs"
    finally:
        ops.set_conv_mode(prev)


def test_graphs_inside_the_pipeline():
    prev = ops.CONV_MODE
    ops.set_conv_mode(1)
    try:
        h, w, n, period = 64, 96, 14, 6
        frames = synth.synth_frames(n, h, w, seed=3)
        i_net, p_net = _models()
        i_dec, p_dec = _models()
        ref = _sequence(i_net, p_net, p_net, frames, h, w, period)
        pipe = EncDecPipeline(i_net, GraphedDMC(p_net), i_dec, GraphedDMC(p_dec), period, True, 1, dev())
        streams, decoded = [], []
        with torch.no_grad():
            for k in range(n):
                with torch.cuda.stream(pipe.enc_stream):
                    x = frames[k:k + 1].to(dev())
                s, d = pipe.push(x)
                streams.append(s)
                if d is not None:        # a graph's outputs are overwritten when it replays again:
                    with torch.cuda.stream(pipe.dec_stream):     # keep a copy, in stream order
                        decoded.append(d.clone())
            last = pipe.flush()
            with torch.cuda.stream(pipe.dec_stream):
                decoded.append(last.clone())
            pipe.join()
            torch.cuda.synchronize()
        for k in range(n):
            assert streams[k] == ref[k][0], f"frame {k}: stream differs"
            assert torch.equal(decoded[k], ref[k][1]["ref_frame"]), f"frame {k}: recon differs"
    finally:
        ops.set_conv_mode(prev)
