#!/usr/bin/env python
"""bench.py -- DCVC-DC 1080p encode+decode frames/s (BASELINE.json metric) on N B200s.

A "step" is one frame through the hot path: `compress` (real rANS bitstream) + `decompress` of a
1920x1080 RGB frame (replicate-padded to 1088 rows as the reference harness does,
DCVC-DC/test_video.py:130-135), frames chained through the decoded-picture buffer with an intra
frame every 32 (BASELINE.json configs[1]).  Random-init weights, synthetic frames (no datasets or
checkpoints offline).

  value : frames/s with the frames already resident in HBM when the timed region starts
          (the finished bitstream still goes to the host and comes back: that is the codec's
          output/input).
  e2e   : the same through the reference-facing module API from HOST buffers: every step copies
          the frame's 8-bit RGB planes from pinned host memory, converts and pads on the device,
          codes, decodes, and reads the frame's squared error back to the host.
  roofline     : the convolution group (the dominant kernels), algorithmic FLOPs / CUDA-event time.
  cpu_baseline : ONE P frame of this very 1080p sequence through the unmodified reference on the
                 host cores (oracle/_ref; the oracle port when that copy is absent), measured.
  parity       : symbol flips / x_hat RMS / bits of the benchmarked mode against mode 0, and
                 encoder == decoder, on the first frames of the sequence (outside the timing).
  arms         : the sequential schedule, the reference's own `encode_decode(output_path)` call
                 order (frame files on tmpfs) and the reference-compatible bitstream, timed beside
                 the pipelined headline.
  P frames are replayed from CUDA graphs (dcvc_b200.graphs; --graphs 0: eager launches).
  gpu_reference: the UNMODIFIED reference (PyTorch/cuDNN + host rANS) on this same GPU.

N > 1 (torchrun): every rank codes its own sequence (independent GOPs; weak scaling); the only
collective is one NCCL all-reduce of the rate/distortion statistics after the timed region.
`--impl reference` times the reference's own CPU implementation of the path instead (1 I + 1 P
frame at the full 1080p size).
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "DCVC-DC 1080p enc+dec frames/s"
UNIT = "frames/s"
HEIGHT, WIDTH, INTRA_PERIOD = 1080, 1920, 32
Q_IN_CKPT, Q_INDEX = False, 32


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=96)   # BASELINE configs[1]: 96 frames = 3 GOPs of 32
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--coder", default="interleaved", choices=["interleaved", "compat"])
    p.add_argument("--stream-part", type=int, default=8)
    p.add_argument("--conv-mode", type=int, default=int(os.environ.get("DCVC_CONV_MODE", "1")))
    p.add_argument("--height", type=int, default=HEIGHT)
    p.add_argument("--width", type=int, default=WIDTH)
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-gpu-reference", action="store_true",
                   help="skip timing the unmodified reference (PyTorch/cuDNN) on this GPU")
    p.add_argument("--no-parity", action="store_true")
    p.add_argument("--no-extra-arms", action="store_true",
                   help="skip the sequential / reference-compatible schedules timed beside the headline")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--pipeline", type=int, default=1,
                   help="1: encoder and decoder on two CUDA streams, decoder one frame behind "
                        "(dcvc_b200.pipeline); 0: the reference's strictly sequential frame loop")
    p.add_argument("--graphs", type=int, default=1,
                   help="1 (default): the P-frame encode / decode launches are replayed from CUDA "
                        "graphs (dcvc_b200.graphs; bit-identical to the eager launch sequence), "
                        "0: eager launches")
    p.add_argument("--profile-out", default=None, help="write the per-kernel table (json) here")
    p.add_argument("--workload", default="rgb1080p", choices=["rgb1080p", "entropy", "yuv420x8", "hem_sweep"],
                   help="rgb1080p: BASELINE configs[1] (the headline); entropy: configs[4] "
                        "microbench (bench_entropy.py); yuv420x8: configs[2], 8 YUV420 sequences "
                        "dealt by GOP over the ranks (bench_yuv.py); hem_sweep: configs[3], DCVC-HEM "
                        "4-point q_scale sweep (bench_hem.py)")
    return p.parse_args()


# ----------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.samples = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-i", str(self.index), "-lms", "200"], stdout=subprocess.PIPE,
                stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for s in self.samples:
            f = [t.strip() for t in s.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for name, val in zip(names, f[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------- CPU reference leg
def _synth_weights():
    from dcvc_b200 import arch
    return (arch.synth_state_dict(arch.intra_spec(), 0), arch.synth_state_dict(arch.dmc_spec(), 100))


NF = 8      # synthetic frames per sequence (played as a ping-pong)


def _padded_host_frames(h, w, seed):
    from dcvc_b200 import synth
    return pad_frame(synth.synth_frames(NF, h, w, seed=seed))


def cpu_reference(args, n_frames, dpb=None, first_idx=0, seed=0):
    """Times the reference's own CPU implementation of the path at the bench's OWN configuration
    (1080p frames, same weights, same rate point) on all host threads: `n_frames` frames of
    compress + decompress, I frame first unless a dpb is given.  kind "reference" = the
    UNMODIFIED reference Python + its C++ coder (oracle/_ref, built from /root/reference by
    oracle/build_ref.sh); kind "port" = the oracle restatement when that copy is absent.
    This is measurement infrastructure: the checker / baseline, never the product."""
    from oracle import ref_runner
    sd_i, sd_p = _synth_weights()
    frames = _padded_host_frames(args.height, args.width, seed)
    frames = frames[first_idx:first_idx + n_frames]
    cores = os.cpu_count() or 1
    if ref_runner.available():
        secs, sizes, kinds, cores = ref_runner.cpu_reference_times(
            sd_i, sd_p, frames, Q_IN_CKPT, Q_INDEX, dpb=dpb, first_idx=first_idx, threads=cores)
        kind = "reference"
        what = ("unmodified reference DCVC-DC (IntraNoAR/DMC compress+decompress, torch %s CPU "
                "fp32 + the reference's C++ rANS)" % torch.__version__)
    else:
        from oracle import codec_oracle as co
        torch.set_num_threads(cores)
        o_i, o_p = co.OracleIntra(sd_i, 1), co.OracleDMC(sd_p, 1)
        o_i.update()
        o_p.update()
        secs, sizes, kinds = [], [], []
        h, w = frames.shape[2], frames.shape[3]
        if dpb is not None:
            dpb = {k: (v.detach().cpu().contiguous() if v is not None else None)
                   for k, v in dpb.items()}
        with torch.no_grad():
            for k in range(frames.shape[0]):
                t0 = time.perf_counter()
                if dpb is None:
                    enc = o_i.compress(frames[k:k + 1], Q_IN_CKPT, Q_INDEX)
                    dec = o_i.decompress(enc["bit_stream"], h, w, Q_IN_CKPT, Q_INDEX)
                    dpb = {"ref_frame": dec["x_hat"], "ref_feature": None,
                           "ref_mv_feature": None, "ref_y": None, "ref_mv_y": None}
                    kinds.append("I")
                else:
                    fi = (first_idx + k) % 4
                    enc = o_p.compress(frames[k:k + 1], dpb, Q_IN_CKPT, Q_INDEX, fi)
                    dpb = o_p.decompress(dpb, enc["bit_stream"], h, w, Q_IN_CKPT, Q_INDEX,
                                         fi)["dpb"]
                    kinds.append("P")
                secs.append(time.perf_counter() - t0)
                sizes.append(len(enc["bit_stream"]))
        kind = "port"
        what = "oracle port (torch fp32 CPU restatement + C port of the reference rANS)"
    total = sum(secs)
    return {"value": len(secs) / total, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"{len(secs)} frame(s) [{' '.join(kinds)}] of the bench's own "
                      f"{args.width}x{args.height} sequence, compress+decompress, "
                      f"{', '.join('%.1f s' % t for t in secs)} per frame on {cores} threads; "
                      f"{what}",
            "seconds_per_frame": secs, "bytes_per_frame": sizes}


def workload_text(args):
    return (f"DCVC-DC {args.width}x{args.height} RGB, intra period {INTRA_PERIOD}, "
            f"compress+decompress per frame with real rANS bitstream "
            f"({args.coder}, {args.stream_part} streams), q_index {Q_INDEX}, one sequence per GPU")


def run_reference(args, rank):
    """--impl reference: the reference's own CPU implementation of the path on the box's host
    cores, at THIS arm's configuration (1080p, same weights / frames / rate point).  A step is one
    frame's compress + decompress; a 1080p frame costs tens of seconds of CPU, so the run is
    bounded to 1 I + 1 P frame (no warm-up: nothing to warm), whatever --steps asks; rank 0 only."""
    if rank != 0:
        return
    t0 = time.time()
    res = cpu_reference(args, 2)
    secs = res["seconds_per_frame"]
    value = res["value"]
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": len(secs), "warmup": 0, "ms_per_step": 1000.0 * sum(secs) / len(secs),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "impl": "reference",
            "config": {"workload": workload_text(args),
                       "note": "CPU arm: the reference's sequential frame loop, 1 process, all host "
                               "threads, coder = the reference's single-stream rANS; bounded to "
                               "1 I + 1 P frame at the full 1080p size (measured, not scaled)"},
            "cpu_baseline": {k: v for k, v in res.items()},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "wall_s": time.time() - t0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- GPU leg
def pad_frame(x):
    """replicate-pad to a multiple of 16 (test_video.py:130-135)"""
    import torch.nn.functional as F
    h, w = x.shape[2], x.shape[3]
    ph, pw = (h + 15) // 16 * 16 - h, (w + 15) // 16 * 16 - w
    return F.pad(x, (0, pw, 0, ph), mode="replicate") if (ph or pw) else x


class Sequence:
    """Frame loop of run_test (test_video.py:138-165) over a ping-pong of synthetic frames."""

    def __init__(self, i_net, p_net, h, w):
        self.i_net, self.p_net, self.h, self.w = i_net, p_net, h, w
        self.dpb = None
        self.idx = 0
        self.bits = 0

    def step(self, x_padded):
        ph, pw = x_padded.shape[2], x_padded.shape[3]
        if self.idx % INTRA_PERIOD == 0:
            enc = self.i_net.compress(x_padded, Q_IN_CKPT, Q_INDEX)
            stream = enc["bit_stream"]
            dec = self.i_net.decompress(stream, ph, pw, Q_IN_CKPT, Q_INDEX)
            self.dpb = {"ref_frame": dec["x_hat"], "ref_feature": None, "ref_mv_feature": None,
                        "ref_y": None, "ref_mv_y": None}
        else:
            fi = self.idx % 4
            enc = self.p_net.compress(x_padded, self.dpb, Q_IN_CKPT, Q_INDEX, fi)
            stream = enc["bit_stream"]
            self.dpb = self.p_net.decompress(self.dpb, stream, ph, pw, Q_IN_CKPT, Q_INDEX,
                                             fi)["dpb"]
        self.idx += 1
        self.bits += len(stream) * 8
        return self.dpb["ref_frame"], len(stream)


class ApiSequence:
    """The reference harness's own calls (test_video.py:141-143,155-158): `encode_decode` with an
    output_path, i.e. frame header + bytes written to a file and read back (here on tmpfs), the
    three device synchronisations DMC.encode_decode makes for its encoding/decoding times."""

    def __init__(self, i_net, p_net, h, w, tmpdir):
        self.i_net, self.p_net, self.h, self.w = i_net, p_net, h, w
        self.dpb, self.idx, self.bits = None, 0, 0
        self.path = os.path.join(tmpdir, f"dcvc_bench_{os.getpid()}.bin")

    def step(self, x_padded):
        if self.idx % INTRA_PERIOD == 0:
            r = self.i_net.encode_decode(x_padded, Q_IN_CKPT, Q_INDEX, self.path,
                                         pic_width=self.w, pic_height=self.h)
            self.dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_mv_feature": None,
                        "ref_y": None, "ref_mv_y": None}
        else:
            r = self.p_net.encode_decode(x_padded, self.dpb, Q_IN_CKPT, Q_INDEX, self.path,
                                         pic_width=self.w, pic_height=self.h,
                                         frame_idx=self.idx % 4)
            self.dpb = r["dpb"]
        self.idx += 1
        self.bits += r["bit"]
        return self.dpb["ref_frame"], r["bit"] // 8


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.workload == "entropy" and args.impl == "b200":
        import bench_entropy
        bench_entropy.run(bench_entropy.parse_args(
            ["--gpus", str(args.gpus), "--steps", str(min(args.steps, 20)),
             "--warmup", str(args.warmup)]))
        return
    if args.workload == "yuv420x8" and args.impl == "b200":
        import bench_yuv
        bench_yuv.run(args)
        return
    if args.workload == "hem_sweep" and args.impl == "b200":
        import bench_hem
        bench_hem.run(args)
        return
    if args.impl == "reference":
        run_reference(args, rank)
        return
    import torch.distributed as dist
    import dcvc_b200
    from dcvc_b200 import _lib, arch, ops, sharding, synth

    dev = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ops.set_conv_mode(args.conv_mode)
    H, W = args.height, args.width

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
    # pipelined mode: the decoder is a second model instance (own weights copy, coder context and
    # dpb), as it would be in a separate process; the two only share the byte stream
    i_dec, p_dec = make_nets() if args.pipeline else (None, None)
    use_graphs = bool(args.graphs)
    arm_graphs = use_graphs
    if use_graphs:
        from dcvc_b200.graphs import GraphedDMC
        p_net = GraphedDMC(p_net)
        p_dec = GraphedDMC(p_dec) if p_dec is not None else None
    # every graph variant (four phases of frame_idx % 4) is seen twice -- eager, then recorded --
    # and replayed once (the first launch uploads the graph) before the timed region: sixteen
    # untimed frames ahead of the warm-up
    prime = 16 if use_graphs else 0

    nf = NF
    host = synth.synth_frames(nf, H, W, seed=rank).pin_memory()      # [nf,3,H,W] fp32 in [0,1]
    order = list(range(nf)) + list(range(nf - 2, 0, -1))              # ping-pong: small motion
    resident = [pad_frame(host[k:k + 1].to(dev)) for k in range(nf)]
    total = args.warmup + args.steps

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(loop_body, pipeline=None, nets=None, steps=None, api=False):
        pipeline = args.pipeline if pipeline is None else pipeline
        ni, np_ = nets if nets is not None else (i_net, p_net)
        total = args.warmup + (args.steps if steps is None else steps)
        if pipeline:
            from dcvc_b200.pipeline import EncDecPipeline
            seq = EncDecPipeline(ni, np_, i_dec, p_dec, INTRA_PERIOD, Q_IN_CKPT, Q_INDEX, dev)
            tstream = seq.enc_stream
        elif api:
            tmp = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
            seq = ApiSequence(ni, np_, resident[0].shape[2], resident[0].shape[3], tmp)
            tstream = torch.cuda.current_stream(dev)
        else:
            seq = Sequence(ni, np_, H, W)
            tstream = torch.cuda.current_stream(dev)
        with torch.no_grad():
            for s in range(-prime, args.warmup):
                loop_body(seq, s)
            if pipeline:
                seq.flush()                      # the timed region holds K encodes and K decodes
            barrier()
            l0 = _lib.launch_count()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(tstream)
            for s in range(args.warmup, total):
                loop_body(seq, s)
            if pipeline:
                last = seq.flush()
                loop_body(None, last)            # end-of-run hook (pending metric reads)
                seq.join(tstream)                # the encoder stream waits for the decoder's
            e1.record(tstream)
            barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, _lib.launch_count() - l0, seq

    stats = {"sse": 0.0, "n": 0, "pending": [], "prev_x": None}

    def body_resident(seq, s):
        if seq is None:
            return
        if hasattr(seq, "push"):
            seq.push(resident[order[s % len(order)]])
        else:
            seq.step(resident[order[s % len(order)]])

    # e2e: the frame enters as raw 8-bit RGB planes from pinned host memory (what a reader hands
    # over: 6.2 MB instead of 25 MB of fp32), is converted and padded on the device
    # (dcvc_frame_u8_to_view), and the step's result -- the decoded frame's per-plane squared
    # error against the source (dcvc_view_plane_sse), 24 bytes -- returns to the host, read one
    # step late in the pipelined schedule so the read never stalls the streams
    from dcvc_b200 import harness
    host_u8 = (host * 255.0).round().to(torch.uint8).reshape(nf, -1).pin_memory()
    feeder = harness.FrameFeeder(harness.FMT_RGB444, H, W, dev, slots=3)
    sse_host = torch.zeros((total + 1, 3), dtype=torch.float64).pin_memory()

    def _metric(slot, x_hat, dstream, idx):
        with torch.cuda.stream(dstream):
            sse_host[idx].copy_(feeder.distortion(slot, x_hat), non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(dstream)
        stats["pending"].append((idx, ev))

    def _drain(keep):
        while len(stats["pending"]) > keep:
            idx, ev = stats["pending"].pop(0)
            ev.synchronize()                                 # that frame's SSE is on the host
            stats["sse"] += float(sse_host[idx].sum()) / (3 * H * W)
            stats["n"] += 1

    def body_e2e(seq, s):
        if seq is None:                           # end of a pipelined run: s = last decoded frame
            if s is not None and stats["prev_x"] is not None:
                slot, idx = stats["prev_x"]
                _metric(slot, s, stats["dstream"], idx)
            stats["prev_x"] = None
            _drain(0)
            return
        k = order[s % len(order)]
        slot = s % 3
        if not hasattr(seq, "push"):
            feeder.upload(slot, host_u8[k].numpy())                  # H2D from pinned memory
            x_hat, _ = seq.step(feeder.to_x(slot))
            _metric(slot, x_hat, torch.cuda.current_stream(dev), s)
            _drain(0)
            return
        with torch.cuda.stream(seq.enc_stream):
            feeder.upload(slot, host_u8[k].numpy())                  # H2D from pinned memory
            xp = feeder.to_x(slot)
        stats["dstream"] = seq.dec_stream
        _, decoded = seq.push(xp)                 # returns with the encoder stream synchronised
        if decoded is not None and stats["prev_x"] is not None:
            pslot, pidx = stats["prev_x"]
            _metric(pslot, decoded, seq.dec_stream, pidx)
        stats["prev_x"] = (slot, s)
        _drain(1)

    clocks = ClockSampler(local_rank)
    clocks.start()
    ms, launches, seq = timed(body_resident)
    clock_info = clocks.stop()
    frames = args.steps * world
    value = frames / (ms / 1000.0)
    bits_per_frame = seq.bits / max(seq.idx, 1)

    e2e = None
    if not args.no_e2e:
        ms_e, _, seq_e = timed(body_e2e)
        e2e = {"value": frames / (ms_e / 1000.0), "unit": UNIT,
               "h2d_bytes_per_step": int(3 * H * W + bits_per_frame / 8),
               "d2h_bytes_per_step": int(bits_per_frame / 8 + 24),
               "ms_per_step": ms_e / args.steps}

    # rate/distortion statistics: the only collective on this path
    rd, _ = sharding.reduce_stats([seq.bits, stats["sse"], stats["n"], seq.idx], ms, dev)

    # per-kernel profile of one P frame (CUDA events per launch, outside the timed region)
    roofline, table = None, None
    if rank == 0:
        # eager launches of the model itself (a graph replay has no per-launch hooks), a fresh
        # sequence: I, two P frames, then the profiled steady-state P frame (all dpb entries present)
        pseq = Sequence(i_net, getattr(p_net, "m", p_net), H, W)
        with torch.no_grad():
            for k in range(3):
                pseq.step(resident[order[k % len(order)]])
        roofline, table = profile_step(pseq, resident[order[3 % len(order)]], args)

    # the same workload in the schedules a reference user would compare with (1 GPU only; short)
    arms = None
    if world == 1 and not args.no_extra_arms:
        arms = {}
        k = min(args.steps, 48)
        p_arm = p_net
        if arm_graphs:      # own graphs: this schedule continues from the DECODER's dpb
            from dcvc_b200.graphs import GraphedDMC
            p_arm = GraphedDMC(p_net.m if use_graphs else p_net)
        gtxt = ("; P-frame launches replayed from CUDA graphs (dcvc_b200.graphs)"
                if arm_graphs else "")
        ms_s, _, seq_s = timed(body_resident, pipeline=0, nets=(i_net, p_arm), steps=k)
        arms["sequential"] = {
            "value": k / (ms_s / 1000.0), "unit": UNIT, "ms_per_step": ms_s / k, "steps": k,
            "what": "the drop-in call order of DMC.encode_decode / test_video.py: encode a frame, "
                    "then decode it, one model instance, one stream; same coder as the headline" + gtxt}
        ms_a, _, seq_a = timed(body_resident, pipeline=0, nets=(i_net, p_arm), steps=k, api=True)
        arms["encode_decode_api"] = {
            "value": k / (ms_a / 1000.0), "unit": UNIT, "ms_per_step": ms_a / k, "steps": k,
            "what": "IntraNoAR.encode_decode / DMC.encode_decode with an output_path, exactly as "
                    "test_video.py:141-143,155-158 calls them: frame file written and read back "
                    "(tmpfs), the call's own device synchronisations; the frame is reconstructed "
                    "once, by the decoder (the reference also continues from the decoder's dpb)" + gtxt}
        try:
            os.remove(seq_a.path)
        except OSError:
            pass
        ci = dcvc_b200.IntraNoAR(stream_part=16, coder_mode="compat")
        cp = dcvc_b200.DMC(stream_part=16, coder_mode="compat")
        ci.load_state_dict(arch.synth_state_dict(arch.intra_spec(), 0))
        cp.load_state_dict(arch.synth_state_dict(arch.dmc_spec(), 100))
        ci, cp = ci.to(dev).eval(), cp.to(dev).eval()
        ci.update(force=True)
        cp.update(force=True)
        kc = min(args.steps, 8)
        ms_c, _, seq_c = timed(body_resident, pipeline=0, nets=(ci, cp), steps=kc)
        arms["reference_compatible"] = {
            "value": kc / (ms_c / 1000.0), "unit": UNIT, "ms_per_step": ms_c / kc, "steps": kc,
            "bits_per_frame": seq_c.bits / max(seq_c.idx, 1),
            "what": "bitstream interchangeable with the reference: compat container "
                    "(--stream_part_i/p 16, py_rans.cpp:66-119; one device thread per part), "
                    "sequential encode -> decode"}
        del ci, cp, seq_c

    parity = None
    if rank == 0 and not args.no_parity:
        parity = parity_block(i_net, getattr(p_net, "m", p_net), resident, order, args)

    gpu_ref = None
    if rank == 0 and world == 1 and not args.no_gpu_reference:
        try:
            from oracle import ref_runner
            if ref_runner.available():
                sd_i, sd_p = _synth_weights()
                gpu_ref = ref_runner.gpu_reference_fps(sd_i, sd_p, _padded_host_frames(H, W, rank),
                                                       Q_IN_CKPT, Q_INDEX, dev)
            else:
                gpu_ref = {"unavailable": "oracle/_ref/DCVC-DC (staged reference Python) absent"}
        except Exception as e:      # measurement aid only: never takes the bench line down
            gpu_ref = {"unavailable": f"{type(e).__name__}: {e}"[:300]}

    if rank == 0:
        cpu = None
        if not args.no_cpu_baseline:
            # one P frame of this very sequence through the reference on the host cores, starting
            # from the decoded I frame (a P frame is ~97 % of a GOP-32; ~20-40 s of CPU)
            with torch.no_grad():
                enc = i_net.compress(resident[0], Q_IN_CKPT, Q_INDEX)
                dec = i_net.decompress(enc["bit_stream"], resident[0].shape[2],
                                       resident[0].shape[3], Q_IN_CKPT, Q_INDEX)
            dpb0 = {"ref_frame": dec["x_hat"].contiguous(), "ref_feature": None,
                    "ref_mv_feature": None, "ref_y": None, "ref_mv_y": None}
            cpu = cpu_reference(args, 1, dpb=dpb0, first_idx=1, seed=rank)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": ("tf32 + binary16 storage of conv-to-conv activations, fp32 accumulate"
                          if ops.F16 else "tf32, fp32 accumulate")
                if args.conv_mode == 1 else "f32", "data": "synthetic",
                "config": {"workload": workload_text(args),
                           "l2": "inputs larger than L2: every step streams a new frame through >= 6 GB of activations (full-resolution tensors are 130-400 MB each, the L2 is 126 MB); no explicit flush",
                           "conv_mode": args.conv_mode, "coder": args.coder,
                           "cuda_graphs": bool(use_graphs),
                           "graph_calls": ({"replayed": p_net.replays + p_dec.replays,
                                            "eager": p_net.eager + p_dec.eager}
                                           if use_graphs and p_dec is not None else None),
                           "schedule": ("PIPELINED (dcvc_b200.pipeline.EncDecPipeline, an API the "
                                        "reference does not have): encoder and decoder are two model "
                                        "instances on two CUDA streams, the decoder one frame behind "
                                        "the encoder; every frame is encoded and decoded inside the "
                                        "timed region; `arms.sequential` is the reference's call order")
                           if args.pipeline else "sequential encode -> decode per frame"},
                "clocks": clock_info, "gpu_launches": int(launches), "e2e": e2e,
                "roofline": roofline, "cpu_baseline": cpu, "parity": parity, "arms": arms,
                "gpu_reference": gpu_ref,
                "bits_per_frame": bits_per_frame,
                "psnr_db": (10 * math.log10(rd[2] / max(rd[1], 1e-12)) if rd[2] else None)}
        print(json.dumps(line), flush=True)
        if args.profile_out and table is not None:
            with open(args.profile_out, "w") as f:
                json.dump(table, f, indent=1)
    if world > 1:
        dist.destroy_process_group()


def parity_block(i_net, p_net, resident, order, args):
    """Parity figures printed beside the timing, computed on the first
    frames of the bench's own sequence, outside the timed region: 1 I + 2 P frames coded in the
    benchmarked conv mode and in mode 0 (fp32 FMA, which tests/test_gpu_1080p.py holds bit-equal
    in symbols to the CPU oracle at this size) FROM THE SAME dpb, so the figures are per-frame
    kernel error.  enc_dec_equal: the encoder's reconstruction (every dpb entry) equals the
    decoder's bit for bit in the benchmarked mode."""
    from dcvc_b200 import ops
    keys = ("ref_frame", "ref_feature", "ref_mv_feature", "ref_y", "ref_mv_y")
    prev = ops.CONV_MODE

    def code(kind, x, dpb, fi, mode):
        ops.set_conv_mode(mode)
        net = i_net if kind == "I" else p_net
        net.taps = {}
        ph, pw = x.shape[2], x.shape[3]
        if kind == "I":
            enc = net.compress(x, Q_IN_CKPT, Q_INDEX)
            dec = net.decompress(enc["bit_stream"], ph, pw, Q_IN_CKPT, Q_INDEX)
            equal = bool(torch.equal(enc["x_hat"], dec["x_hat"]))
            new = {"ref_frame": dec["x_hat"], "ref_feature": None, "ref_mv_feature": None,
                   "ref_y": None, "ref_mv_y": None}
        else:
            enc = net.compress(x, dpb, Q_IN_CKPT, Q_INDEX, fi)
            dec = net.decompress(dpb, enc["bit_stream"], ph, pw, Q_IN_CKPT, Q_INDEX, fi)
            equal = all(bool(torch.equal(enc["dpb"][k], dec["dpb"][k])) for k in keys)
            new = dec["dpb"]
        taps, net.taps = net.taps, None
        return len(enc["bit_stream"]), new, taps, equal

    def flips(ta, tb, prefixes):
        n = d = 0
        for pre in prefixes:
            for st in range(4):
                a, b = ta[f"{pre}_q_w_{st}"], tb[f"{pre}_q_w_{st}"]
                n += a.numel()
                d += int((a != b).sum())
        return d, n

    out = {"frames": "1 I + 2 P of this sequence, benchmarked mode vs mode 0 (fp32 FMA) from the same dpb",
           "enc_dec_equal": True, "symbol_flips": 0, "symbols": 0}
    bits_m, bits_0, sq, npx = 0, 0, 0.0, 0
    try:
        with torch.no_grad():
            dpb = None
            for k in range(3):
                x = resident[order[k]]
                kind = "I" if k == 0 else "P"
                b0, dpb0, t0, eq0 = code(kind, x, dpb, k % 4, 0)
                if args.conv_mode != 0:
                    b1, dpb1, t1, eq1 = code(kind, x, dpb, k % 4, args.conv_mode)
                else:
                    b1, dpb1, t1, eq1 = b0, dpb0, t0, eq0
                out["enc_dec_equal"] = out["enc_dec_equal"] and eq0 and eq1
                d, n = flips(t1, t0, ("i_y",) if kind == "I" else ("p_y", "p_mv_y"))
                out["symbol_flips"] += d
                out["symbols"] += n
                bits_m += b1
                bits_0 += b0
                diff = dpb1["ref_frame"] - dpb0["ref_frame"]
                sq += float(diff.pow(2).sum())
                npx += diff.numel()
                dpb = dpb0
    finally:
        ops.set_conv_mode(prev)
        i_net.taps = p_net.taps = None
    out["x_hat_rms_vs_mode0"] = (sq / max(npx, 1)) ** 0.5
    out["d_bits_pct"] = 100.0 * (bits_m - bits_0) / max(bits_0, 1)
    out["symbol_flip_frac"] = out["symbol_flips"] / max(out["symbols"], 1)
    out["oracle_link"] = ("mode 0 vs the CPU oracle at 1080p: symbols bit-equal, x_hat RMS 4e-8 "
                          "(tests/test_gpu_1080p.py, profiles/r02_parity_1080p.json)")
    return out


def profile_step(seq, x, args):
    """One more P frame with a CUDA-event pair around every launch of the library (events on the
    launching stream).  Returns the roofline object of the dominant kernel group and the table."""
    from dcvc_b200 import ops
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    with torch.no_grad():
        if seq.idx % INTRA_PERIOD == 0:
            seq.step(x)                          # the I frame that opens the GOP: not profiled
        ops.PROFILE = []
        seq.step(x)
    torch.cuda.synchronize()
    rows = {}
    for name, flops, nbytes, e0, e1 in ops.PROFILE:
        r = rows.setdefault(name, {"kernel": name, "launches": 0, "ms": 0.0, "flops": 0, "bytes": 0})
        r["launches"] += 1
        r["ms"] += e0.elapsed_time(e1)
        r["flops"] += flops
        r["bytes"] += nbytes
    ops.PROFILE = None
    table = sorted(rows.values(), key=lambda r: -r["ms"])
    total_ms = sum(r["ms"] for r in table)
    for r in table:
        r["share"] = r["ms"] / total_ms if total_ms else 0.0
        r["tflops"] = r["flops"] / (r["ms"] * 1e9) if r["ms"] else 0.0
        r["gbs"] = r["bytes"] / (r["ms"] * 1e6) if r["ms"] else 0.0
    # the dominant kernel is conv_tc_kernel (one __global__ function, every tensor-core
    # convolution shape of the frame): its launches are aggregated -- algorithmic flops / bytes
    # summed over the launches, divided by their summed duration
    conv = [r for r in table if "[tcgen05]" in r["kernel"] and r["kernel"].startswith("conv")]
    if not conv:
        conv = [r for r in table if r["kernel"].startswith("conv")]
    if not conv:
        return None, table
    top = {"kernel": "conv_tc_kernel (all %d convolution shapes of a P frame)" % len(conv),
           "launches": sum(r["launches"] for r in conv), "ms": sum(r["ms"] for r in conv),
           "flops": sum(r["flops"] for r in conv), "bytes": sum(r["bytes"] for r in conv)}
    top["share"] = top["ms"] / total_ms if total_ms else 0.0
    top["tflops"] = top["flops"] / (top["ms"] * 1e9) if top["ms"] else 0.0
    top["gbs"] = top["bytes"] / (top["ms"] * 1e6) if top["ms"] else 0.0
    n = max(top["launches"], 1)
    tf_peak = peaks.get("bf16_tflops_sustained")
    hbm_peak = peaks.get("hbm_gbs")
    which = "measured (MEASURED_PEAKS.json: bf16 sustained TFLOP/s, HBM copy GB/s)"
    if tf_peak is None or hbm_peak is None:
        tf_peak, hbm_peak, which = 1400.0, 6400.0, "fallback (B200_PROFILING.md)"
    # the roofline that bounds this kernel is the one with the larger minimum time
    t_tensor = top["flops"] / (tf_peak * 1e9)          # ms
    t_hbm = top["bytes"] / (hbm_peak * 1e6)            # ms
    # measured DRAM traffic per launch: mean over the conv_tc_kernel launches of the committed
    # ncu pass (dram__bytes_read.sum + dram__bytes_write.sum per launch, profiles/)
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "r02_conv_tc_traffic.json")) as f:
            traffic = json.load(f)["dram_bytes_per_launch"]
    except (OSError, ValueError, KeyError):
        pass
    heavy = max(conv, key=lambda r: r["ms"])
    common = {"kernel": top["kernel"], "traffic": traffic, "peak_source": which,
              "share_of_step": top["share"], "launches": top["launches"],
              "algorithmic_flops_per_launch": top["flops"] / n,
              "algorithmic_bytes_per_launch": top["bytes"] / n,
              "avg_launch_ms": top["ms"] / n,
              "tensor_tflops": top["tflops"], "tensor_frac_of_bf16_peak": top["tflops"] / tf_peak,
              "hbm_gbs": top["gbs"], "hbm_frac": top["gbs"] / hbm_peak,
              "heaviest_shape": {k: heavy[k] for k in ("kernel", "launches", "ms", "tflops", "gbs")}}
    if t_hbm >= t_tensor:
        roof = {"bound": "hbm", "achieved": top["gbs"], "peak": hbm_peak, "unit": "GB/s",
                "frac": top["gbs"] / hbm_peak}
    else:
        roof = {"bound": "tensor", "achieved": top["tflops"], "peak": tf_peak, "unit": "TFLOP/s",
                "frac": top["tflops"] / tf_peak}
    roof.update(common)
    return roof, table


if __name__ == "__main__":
    main()
