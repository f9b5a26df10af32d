"""Entry point with the flags, job layout and output JSON of the reference's
DCVC-DC/test_video.py (:26-56 flags, :318-465 job construction and result file), on the
B200-native models.

    python test_video.py --i_frame_model_path I.pth.tar --p_frame_model_path P.pth.tar \\
        --rate_num 4 --test_config cfg.json --cuda 1 --write_stream 1 --output_path out.json

One process per GPU instead of the reference's process pool: launched plainly it runs every
(sequence, rate) job on cuda:0 (or --cuda_device); launched under torchrun (RANK / WORLD_SIZE /
LOCAL_RANK in the environment) the (job, GOP) units are dealt to the ranks (every GOP restarts
from an intra frame, so GOPs are independent; `--shard job` deals whole jobs instead) and rank 0
merges the per-frame records and writes the one JSON.  `--worker` is accepted for command-line
compatibility and ignored; `--cuda 0` is refused (this implementation has no CPU path);
`--calc_ssim 1` is refused (MS-SSIM is outside the ported path).
"""
import argparse
import json
import os
import time

import numpy as np
import torch
import torch.distributed as dist


def str2bool(v):
    return str(v).lower() in ("yes", "y", "true", "t", "1")


def parse_args(argv=None):
    p = argparse.ArgumentParser(description="DCVC-DC test harness (dcvc_b200)")
    p.add_argument("--ec_thread", type=str2bool, nargs="?", const=True, default=False)
    p.add_argument("--stream_part_i", type=int, default=1)
    p.add_argument("--stream_part_p", type=int, default=1)
    p.add_argument("--i_frame_model_path", type=str)
    p.add_argument("--p_frame_model_path", type=str)
    p.add_argument("--rate_num", type=int, default=4)
    p.add_argument("--i_frame_q_indexes", type=int, nargs="+")
    p.add_argument("--p_frame_q_indexes", type=int, nargs="+")
    p.add_argument("--force_intra", type=str2bool, nargs="?", const=True, default=False)
    p.add_argument("--force_frame_num", type=int, default=-1)
    p.add_argument("--force_intra_period", type=int, default=-1)
    p.add_argument("--test_config", type=str, required=True)
    p.add_argument("--yuv420", type=str2bool, default=False, required=False)
    p.add_argument("--worker", "-w", type=int, default=1, help="ignored (one process per GPU)")
    p.add_argument("--cuda", type=str2bool, nargs="?", const=True, default=True)
    p.add_argument("--cuda_device", default=None)
    p.add_argument("--calc_ssim", type=str2bool, default=False, required=False)
    p.add_argument("--write_stream", type=str2bool, nargs="?", const=True, default=False)
    p.add_argument("--stream_path", type=str, default="out_bin")
    p.add_argument("--save_decoded_frame", type=str2bool, default=False)
    p.add_argument("--decoded_frame_path", type=str, default="decoded_frames")
    p.add_argument("--output_path", type=str, required=True)
    p.add_argument("--verbose", type=int, default=0)
    # extension: the product's second container (lossless, warp-interleaved)
    p.add_argument("--coder_mode", choices=("compat", "interleaved"), default="compat")
    # extension: what is dealt to the ranks under torchrun -- GOPs (default) or whole jobs
    p.add_argument("--shard", choices=("gop", "job"), default="gop")
    # extension: with --write_stream, replay the P frames' launches from CUDA graphs
    # (dcvc_b200.graphs; bit-identical to the eager launches)
    p.add_argument("--graphs", type=str2bool, nargs="?", const=True, default=False)
    return p.parse_args(argv)


def rate_indexes(args, anchors):
    """test_video.py:346-361: four rates -> the checkpoint's own anchors; otherwise rate_num
    points spread over the 64-entry fine table."""
    if args.i_frame_q_indexes is not None:
        assert len(args.i_frame_q_indexes) == args.rate_num
        return False, list(args.i_frame_q_indexes)
    if len(anchors) == args.rate_num:
        assert args.rate_num == 4
        return True, [0, 1, 2, 3]
    assert 2 <= args.rate_num <= 64
    return False, [int(i + 0.5) for i in np.linspace(0, 63, num=args.rate_num)]


def build_jobs(args, config, q_in_ckpt, q_indexes):
    jobs = []
    root = config["root_path"]
    for ds_name, ds in config["test_classes"].items():
        if ds["test"] == 0:
            continue
        for seq_name, seq in ds["sequences"].items():
            for rate_idx in range(args.rate_num):
                gop = seq["gop"]
                if args.force_intra:
                    gop = 1
                if args.force_intra_period > 0:
                    gop = args.force_intra_period
                bin_folder = os.path.join(args.stream_path, seq_name, str(rate_idx))
                jobs.append({
                    "rate_idx": rate_idx, "q_in_ckpt": q_in_ckpt,
                    "i_frame_q_index": q_indexes[rate_idx], "p_frame_q_index": q_indexes[rate_idx],
                    "force_intra": args.force_intra, "video_path": seq_name,
                    "dist_in_yuv420": args.yuv420, "src_type": ds["src_type"],
                    "src_height": seq["height"], "src_width": seq["width"], "gop_size": gop,
                    "frame_num": args.force_frame_num if args.force_frame_num > 0 else seq["frames"],
                    "src_path": os.path.join(root, ds["base_path"], seq_name),
                    "write_stream": args.write_stream, "bin_folder": bin_folder,
                    "save_decoded_frame": args.save_decoded_frame,
                    "recon_path": os.path.join(args.decoded_frame_path, seq_name, str(rate_idx)),
                    "ds_name": ds_name, "verbose": args.verbose})
    return jobs


def _round(obj, digits=6):
    if isinstance(obj, float):
        return round(obj, digits)
    if isinstance(obj, dict):
        return {k: _round(v, digits) for k, v in obj.items()}
    if isinstance(obj, list):
        return [_round(v, digits) for v in obj]
    return obj


def main(argv=None):
    begin = time.time()
    args = parse_args(argv)
    if not args.cuda:
        raise SystemExit("dcvc_b200 codes on the GPU only (--cuda 0 is not supported)")
    if args.calc_ssim:
        raise SystemExit("--calc_ssim is not supported (MS-SSIM is outside the ported path)")
    if args.cuda_device:
        os.environ["CUDA_VISIBLE_DEVICES"] = str(args.cuda_device)
    import dcvc_b200
    from dcvc_b200 import harness
    from dcvc_b200.stream_helper import get_state_dict

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    device = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    torch.manual_seed(0)
    np.random.seed(0)

    with open(args.test_config) as f:
        config = json.load(f)
    enc_anchors, dec_anchors = dcvc_b200.IntraNoAR.get_q_scales_from_ckpt(args.i_frame_model_path)
    q_in_ckpt, q_indexes = rate_indexes(args, enc_anchors)
    if rank == 0:
        print("q_scale_enc in intra ckpt: " + "".join(f"{float(q):.3f}, " for q in enc_anchors))
        print("q_scale_dec in intra ckpt: " + "".join(f"{float(q):.3f}, " for q in dec_anchors))
        print(f"testing {args.rate_num} rates, using q_indexes: " + "".join(f"{q}, " for q in q_indexes))

    i_net = dcvc_b200.IntraNoAR(ec_thread=args.ec_thread, stream_part=args.stream_part_i,
                                inplace=True, coder_mode=args.coder_mode)
    i_net.load_state_dict(get_state_dict(args.i_frame_model_path))
    i_net = i_net.to(device).eval()
    p_net = None
    if not args.force_intra:
        p_net = dcvc_b200.DMC(ec_thread=args.ec_thread, stream_part=args.stream_part_p,
                              inplace=True, coder_mode=args.coder_mode)
        p_net.load_state_dict(get_state_dict(args.p_frame_model_path))
        p_net = p_net.to(device).eval()
    i_net.update(force=True)
    if p_net is not None:
        p_net.update(force=True)
        if args.graphs:
            from dcvc_b200.graphs import GraphedDMC
            GraphedDMC(p_net)        # registers itself: DMC.encode_decode replays through it

    jobs = build_jobs(args, config, q_in_ckpt, q_indexes)
    # the unit of multi-GPU work is one GOP of one job (every GOP restarts from an intra frame,
    # reference test_video.py:140-159): all (job, GOP) units are dealt longest-first over the ranks
    # by the same deterministic map on every rank; decoded-frame dumps need whole sequences
    from dcvc_b200 import sharding
    units = []
    for ji, job in enumerate(jobs):
        per = job["frame_num"] if (args.save_decoded_frame or args.shard == "job") \
            else max(job["gop_size"], 1)
        units += [sharding.Unit(ji, u.gop, u.first_frame, u.n_frames)
                  for u in sharding.work_units(1, job["frame_num"], per)]
    parts = []
    for u in sharding.assign_units(units, world)[rank]:
        job = jobs[u.sequence]
        if job["write_stream"]:
            os.makedirs(job["bin_folder"], exist_ok=True)
        rec = harness.run_sequence(p_net, i_net, job, u.first_frame, u.n_frames, summary=False)
        rec["job"] = u.sequence
        parts.append(rec)
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, parts)      # per-frame records: a few floats per frame
        parts = [r for g in gathered for r in g]
        dist.destroy_process_group()
    if rank != 0:
        return
    results = []
    for ji, job in enumerate(jobs):
        res = harness.merge_units([r for r in parts if r["job"] == ji])
        res.update(ds_name=job["ds_name"], video_path=job["video_path"], rate_idx=job["rate_idx"])
        results.append(res)
    log = {}
    for ds_name, ds in config["test_classes"].items():
        if ds["test"] == 0:
            continue
        log[ds_name] = {}
        for seq in ds["sequences"]:
            log[ds_name][seq] = {}
            for rate in range(args.rate_num):
                for r in results:
                    if r["rate_idx"] == rate and r["ds_name"] == ds_name and r["video_path"] == seq:
                        log[ds_name][seq][f"{rate:03d}"] = r
    out_dir = os.path.dirname(args.output_path)
    if out_dir:
        os.makedirs(out_dir, exist_ok=True)
    with open(args.output_path, "w") as fp:
        json.dump(_round(log), fp, indent=2)
    frames = sum(j["frame_num"] for j in jobs)
    print("Test finished")
    print(f"Tested {frames} frames from {len(jobs) // max(args.rate_num, 1)} sequences")
    print(f"Total elapsed time: {(time.time() - begin) / 60:.1f} min")


if __name__ == "__main__":
    main()
