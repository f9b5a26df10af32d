"""TEST / MEASUREMENT INFRASTRUCTURE -- drives the UNMODIFIED reference models (imported by
oracle/ref_loader.py from /root/reference or from the staged copy in oracle/_ref/) through their
own public API: `IntraNoAR.compress/decompress`, `DMC.compress/decompress`
(reference DCVC-DC/src/models/image_model.py:198-252, video_model.py:425-520), set up the way the
reference's harness sets them up (DCVC-DC/test_video.py:276-316).

Used by bench.py for (a) the CPU reference arm (`--impl reference`, `cpu_baseline`) and (b) the
"unmodified reference on the same B200" figure (`gpu_reference`: PyTorch/cuDNN kernels + the
reference's host rANS, the `--cuda 1` path).  None of this repo's kernels are on that path.
"""
import os
import time

import torch

from . import ref_loader


def available():
    return ref_loader.reference_available("DCVC-DC")


def load_models(sd_i, sd_p, device, stream_part=1):
    """init_func of the reference harness (test_video.py:296-316): build, load, .to(device),
    .eval(), update(force=True)."""
    ref = ref_loader.import_reference("DCVC-DC")
    i_net = ref.IntraNoAR(ec_thread=False, stream_part=stream_part, inplace=True)
    p_net = ref.DMC(ec_thread=False, stream_part=stream_part, inplace=True)
    i_net.load_state_dict(sd_i)
    p_net.load_state_dict(sd_p)
    i_net, p_net = i_net.to(device).eval(), p_net.to(device).eval()
    i_net.update(force=True)
    p_net.update(force=True)
    return i_net, p_net


def _sync(device):
    if torch.device(device).type == "cuda":
        torch.cuda.synchronize(device)


def code_frames(i_net, p_net, frames, q_in_ckpt, q_index, device, dpb=None, first_idx=0,
                intra_period=32):
    """run_test's frame loop (test_video.py:138-165) on padded frames [n,3,H,W] (host tensor):
    per frame compress -> bytes -> decompress, dpb chained.  Returns ([seconds per frame],
    [bytes per frame], dpb, [kind per frame])."""
    secs, sizes, kinds = [], [], []
    h, w = frames.shape[2], frames.shape[3]
    with torch.no_grad():
        for k in range(frames.shape[0]):
            idx = first_idx + k
            x = frames[k:k + 1].to(device)
            _sync(device)
            t0 = time.perf_counter()
            if idx % intra_period == 0 or dpb is None:
                enc = i_net.compress(x, q_in_ckpt, q_index)
                dec = i_net.decompress(enc["bit_stream"], h, w, q_in_ckpt, q_index)
                dpb = {"ref_frame": dec["x_hat"], "ref_feature": None, "ref_mv_feature": None,
                       "ref_y": None, "ref_mv_y": None}
                kinds.append("I")
            else:
                enc = p_net.compress(x, dpb, q_in_ckpt, q_index, idx % 4)
                dpb = p_net.decompress(dpb, enc["bit_stream"], h, w, q_in_ckpt, q_index,
                                       idx % 4)["dpb"]
                kinds.append("P")
            _sync(device)
            secs.append(time.perf_counter() - t0)
            sizes.append(len(enc["bit_stream"]))
    return secs, sizes, dpb, kinds


def gpu_reference_fps(sd_i, sd_p, frames, q_in_ckpt, q_index, device, warm_p=2, timed_p=6):
    """The unmodified reference with the model on the GPU (test_video.py --cuda 1: cuDNN/ATen
    kernels, deterministic algorithms as test_video.py:277-281 sets them, host rANS through
    .cpu()): 1 I + warm_p untimed P frames, then timed_p P frames.  Returns a dict."""
    os.environ.setdefault("CUBLAS_WORKSPACE_CONFIG", ":4096:8")      # test_video.py:327
    prev_det = torch.are_deterministic_algorithms_enabled()
    prev_bench = torch.backends.cudnn.benchmark
    torch.backends.cudnn.benchmark = False
    torch.use_deterministic_algorithms(True)
    try:
        i_net, p_net = load_models(sd_i, sd_p, device)
        n = frames.shape[0]
        order = [k % n for k in range(1 + warm_p + timed_p)]
        seq = frames[order]
        secs, sizes, _, kinds = code_frames(i_net, p_net, seq, q_in_ckpt, q_index, device)
    finally:
        torch.use_deterministic_algorithms(prev_det)
        torch.backends.cudnn.benchmark = prev_bench
    t_p = secs[1 + warm_p:]
    del i_net, p_net
    torch.cuda.empty_cache()
    return {"value": len(t_p) / sum(t_p), "unit": "frames/s",
            "what": "UNMODIFIED reference (DCVC-DC IntraNoAR/DMC, PyTorch %s + cuDNN, deterministic "
                    "algorithms on as test_video.py:277-281, host rANS via .cpu()) with the model "
                    "on this GPU: compress+decompress per frame, same weights, same frames"
                    % torch.__version__,
            "frames_timed": len(t_p), "frame_kind": "P", "ms_per_frame": 1000.0 * sum(t_p) / len(t_p),
            "i_frame_ms": 1000.0 * secs[0], "bytes_per_p_frame": sum(sizes[1 + warm_p:]) / len(t_p)}


def cpu_reference_times(sd_i, sd_p, frames, q_in_ckpt, q_index, dpb=None, first_idx=0, threads=None):
    """The reference on the host cores (test_video.py --cuda 0), one process, torch on `threads`
    threads (default: all): codes `frames` (I first unless a dpb is given)."""
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    i_net, p_net = load_models(sd_i, sd_p, "cpu")
    if dpb is not None:
        dpb = {k: (v.detach().cpu().contiguous() if v is not None else None) for k, v in dpb.items()}
    secs, sizes, _, kinds = code_frames(i_net, p_net, frames, q_in_ckpt, q_index, "cpu", dpb=dpb,
                                        first_idx=first_idx)
    return secs, sizes, kinds, threads
