#!/usr/bin/env python
"""BASELINE.json configs[2]: DCVC-DC YUV420 1080p, 8 independent synthetic sequences x 96 frames,
intra period 32 -> 24 (sequence, GOP) work units dealt over 1/2/4/8 B200s
(`python bench.py --workload yuv420x8 --gpus N`; torchrun for N > 1).

The unit of parallel work is the GOP (reference DCVC-DC/test_video.py:140-159: every
GOP restarts from an intra frame): `dcvc_b200.sharding.assign_units(work_units(8, 96, 32), N)`
gives every rank its list, computed locally, no exchange.  STRONG scaling: the job (768 frames) is
fixed, `value` = 768 frames / (max over ranks of the device time).  The only collective is one
all-reduce of {bits, per-plane PSNR sums, frame counts} after the timed region.

Per frame, inside the timed region (e2e): 8-bit Y, U, V planes from pinned host memory -> device
(3.1 MB) -> `dcvc_frame_u8_to_view` (nearest chroma upsample, /255, replicate pad; reference
test_video.py:110-116,130-135) -> compress -> bytes to the host and back -> decompress ->
`dcvc_view_plane_sse` (2x2 chroma mean, clip, per-plane float64 SSE; test_video.py:169-195) ->
24 bytes to the host.  `value` is the same loop with the planes already resident in HBM.
"""
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = "DCVC-DC 1080p enc+dec frames/s"
UNIT = "frames/s"
N_SEQ, N_FRAMES, INTRA = 8, 96, 32
NF = 8                                       # distinct synthetic frames per sequence (ping-pong)
ORDER = list(range(NF)) + list(range(NF - 2, 0, -1))
Q_IN_CKPT, Q_INDEX = False, 32


def sequence_planes(seq, h, w):
    """8 distinct YUV420 8-bit frames of sequence `seq`: the RGB field through BT.709
    (functional.py:16-39), 2x2 chroma mean, 8-bit.  -> uint8 [NF, h*w*3/2] (pinned)."""
    from dcvc_b200 import harness, synth
    rgb = synth.synth_frames(NF, h, w, seed=100 + seq)
    out = torch.empty((NF, h * w * 3 // 2), dtype=torch.uint8).pin_memory()
    for k in range(NF):
        y, uv = harness.x_to_yuv420(synth.rgb_to_ycbcr444(rgb[k:k + 1]))
        planes = torch.cat(((y * 255).round().reshape(-1), (uv * 255).round().reshape(-1)))
        out[k] = planes.to(torch.uint8)
    return out


def run(args):
    import torch.distributed as dist
    import dcvc_b200
    from dcvc_b200 import _lib, arch, harness, ops, sharding
    from dcvc_b200.pipeline import EncDecPipeline
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dev = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(dev)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=dev)
    ops.set_conv_mode(args.conv_mode)
    H, W = args.height, args.width
    n_frames = N_FRAMES if args.steps >= N_FRAMES else max(INTRA, args.steps // INTRA * INTRA)

    def make_nets():
        i = dcvc_b200.IntraNoAR(stream_part=args.stream_part, coder_mode=args.coder)
        p = dcvc_b200.DMC(stream_part=args.stream_part, coder_mode=args.coder)
        i.load_state_dict(arch.synth_state_dict(arch.intra_spec(), 0))
        p.load_state_dict(arch.synth_state_dict(arch.dmc_spec(), 100))
        i, p = i.to(dev).eval(), p.to(dev).eval()
        i.update(force=True)
        p.update(force=True)
        return i, p

    i_net, p_net = make_nets()
    i_dec, p_dec = make_nets()
    # bench.py --graphs 1 (default): P frames replayed from CUDA graphs (bit-identical to the eager
    # launches; the first two P frames of every GOP continue from eager frames and run eagerly)
    use_graphs = bool(args.graphs)
    if use_graphs:
        from dcvc_b200.graphs import GraphedDMC
        p_net, p_dec = GraphedDMC(p_net), GraphedDMC(p_dec)
    units = sharding.work_units(N_SEQ, n_frames, INTRA)
    mine = sharding.assign_units(units, world)[rank]
    host = {s: sequence_planes(s, H, W) for s in sorted({u.sequence for u in mine})}
    my_frames = sum(u.n_frames for u in mine)
    fmt = harness.FMT_YUV420

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def one_pass(resident, warm_units):
        """Codes this rank's units (or `warm_units` untimed ones first); returns (ms, launches,
        bits, [sse float64 [3]] per frame in pinned memory)."""
        slots = 3
        feeder = harness.FrameFeeder(fmt, H, W, dev, slots=slots)
        res_planes = None
        if resident:      # the distinct source frames already in HBM: slot = frame id
            res_planes = {s: harness.FrameFeeder(fmt, H, W, dev, slots=NF) for s in host}
            for s, f in res_planes.items():
                for k in range(NF):
                    f.upload(k, host[s][k].numpy())
        todo = list(warm_units) + list(mine)
        n_total = sum(u.n_frames for u in todo)
        sse_host = torch.zeros((n_total, 3), dtype=torch.float64).pin_memory()
        pipe = EncDecPipeline(i_net, p_net, i_dec, p_dec, INTRA, Q_IN_CKPT, Q_INDEX, dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        count = 0
        prev = None                      # (feeder, slot, frame counter) of the frame in flight
        l0 = ms = None
        bits0 = 0

        def finish(decoded):
            nonlocal prev
            if decoded is not None and prev is not None:
                f, slot, idx = prev
                with torch.cuda.stream(pipe.dec_stream):
                    sse = f.distortion(slot, decoded)
                    sse_host[idx].copy_(sse, non_blocking=True)     # 24 bytes D2H per frame
            prev = None

        with torch.no_grad():
            for ui, u in enumerate(todo):
                if ui == len(warm_units):            # timed region starts here
                    finish(pipe.flush())
                    pipe.idx = 0                     # every unit opens with its intra frame
                    barrier()
                    l0 = _lib.launch_count()
                    bits0 = pipe.bits
                    e0.record(pipe.enc_stream)
                for k in range(u.n_frames):
                    src = ORDER[(u.first_frame + k) % len(ORDER)]
                    with torch.cuda.stream(pipe.enc_stream):
                        if resident:
                            f, slot = res_planes[u.sequence], src
                        else:
                            f, slot = feeder, count % slots
                            f.upload(slot, host[u.sequence][src].numpy())   # H2D from pinned
                        x = f.to_x(slot)
                    _, decoded = pipe.push(x)
                    keep = (f, slot, count)
                    finish(decoded)
                    prev = keep
                    count += 1
            finish(pipe.flush())
            pipe.join(pipe.enc_stream)
            e1.record(pipe.enc_stream)
            barrier()
        ms = e0.elapsed_time(e1)
        return ms, _lib.launch_count() - l0, pipe.bits - bits0, sse_host[n_total - my_frames:]

    # a short untimed warm-up unit (I + a few P frames) precedes the timed units of each pass
    n_warm = max(args.warmup, 18 if use_graphs else 3)   # graphs: every variant seen twice first
    warm = [sharding.Unit(mine[0].sequence if mine else 0, 0, 0, n_warm)] if mine else []
    if not mine:                                     # more ranks than units: nothing to code
        host[0] = sequence_planes(0, H, W)
        warm = [sharding.Unit(0, 0, 0, 3)]
    from bench import ClockSampler
    clocks = ClockSampler(local_rank)
    clocks.start()
    ms_r, launches, bits, _ = one_pass(True, warm)
    clock_info = clocks.stop()
    ms_e, _, bits_e, sse = one_pass(False, warm)
    torch.cuda.synchronize(dev)
    npx = (H * W, H * W // 4, H * W // 4)
    ps = [[harness.psnr_from_sse(float(sse[k, c]), npx[c]) for c in range(3)]
          for k in range(sse.shape[0])]
    sums = [sum(p[c] for p in ps) for c in range(3)]
    rd, ms_r = sharding.reduce_stats([bits, sums[0], sums[1], sums[2], len(ps), my_frames], ms_r, dev)
    _, ms_e = sharding.reduce_stats([0.0], ms_e, dev)
    total_frames = rd[5]
    if rank == 0:
        n = max(rd[4], 1)
        py, pu, pv = rd[1] / n, rd[2] / n, rd[3] / n
        plane_bytes = H * W * 3 // 2
        line = {"metric": METRIC, "value": total_frames / (ms_r / 1000.0), "unit": UNIT,
                "n_gpus": world, "steps": int(total_frames), "warmup": warm[0].n_frames,
                "ms_per_step": ms_r / max(total_frames, 1), "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None,
                "dtype": "tf32 + binary16 storage of conv-to-conv activations, fp32 accumulate"
                if args.conv_mode == 1 else "f32",
                "data": "synthetic",
                "config": {"workload": f"BASELINE configs[2]: DCVC-DC YUV420 {W}x{H}, {N_SEQ} "
                                       f"sequences x {n_frames} frames, intra period {INTRA}, "
                                       f"{len(units)} (sequence, GOP) units dealt over {world} "
                                       f"GPU(s) (dcvc_b200.sharding.assign_units), "
                                       f"compress+decompress per frame, real rANS bitstream "
                                       f"({args.coder}, {args.stream_part} streams), q_index {Q_INDEX}",
                           "units_per_rank": [len(x) for x in sharding.assign_units(units, world)],
                           "schedule": "pipelined (dcvc_b200.pipeline.EncDecPipeline)",
                           "cuda_graphs": bool(use_graphs),
                           "l2": "inputs larger than L2: every frame streams >= 6 GB of activations"},
                "clocks": clock_info, "gpu_launches": int(launches),
                "e2e": {"value": total_frames / (ms_e / 1000.0), "unit": UNIT,
                        "h2d_bytes_per_step": int(plane_bytes + bits / max(total_frames, 1) / 8),
                        "d2h_bytes_per_step": int(bits / max(total_frames, 1) / 8 + 24),
                        "ms_per_step": ms_e / max(total_frames, 1)},
                "rd": {"bpp": rd[0] / (total_frames * H * W), "psnr_y": py, "psnr_u": pu,
                       "psnr_v": pv, "psnr_yuv_6_1_1": (6 * py + pu + pv) / 8,
                       "frames": int(total_frames)},
                "roofline": None, "cpu_baseline": None}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()


if __name__ == "__main__":
    import bench
    a = bench.parse_args()
    a.workload = "yuv420x8"
    run(a)
