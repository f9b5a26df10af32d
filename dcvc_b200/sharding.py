"""GOP/sequence sharding across ranks.

Frames inside a GOP are a strict chain through the decoded-picture buffer and every GOP starts
with an intra frame that resets it (reference: DCVC-DC/test_video.py:140-159), so the unit of
parallel work is (sequence, GOP).  The reference maps worker i -> GPU i % n with a process pool
(test_video.py:282-294); here the map is static and deterministic so that every rank can compute
it locally with no exchange.  The data path has no collective: the only communication is one
all-reduce of the rate/distortion counters and a max over the ranks' device times.
"""
from collections import namedtuple

import torch
import torch.distributed as dist

Unit = namedtuple("Unit", "sequence gop first_frame n_frames")


def work_units(n_sequences, n_frames, intra_period):
    """All (sequence, GOP) units of a job; the last GOP of a sequence may be short."""
    if intra_period <= 0:
        raise ValueError("intra_period must be positive")
    units = []
    for s in range(n_sequences):
        for g, first in enumerate(range(0, n_frames, intra_period)):
            units.append(Unit(s, g, first, min(intra_period, n_frames - first)))
    return units


def assign_units(units, world_size):
    """Longest-first greedy assignment; returns world_size lists.  Ties break on (sequence, gop)
    and on the lowest rank so every rank derives the same map."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    load = [0] * world_size
    out = [[] for _ in range(world_size)]
    for u in sorted(units, key=lambda u: (-u.n_frames, u.sequence, u.gop)):
        r = min(range(world_size), key=lambda i: (load[i], i))
        out[r].append(u)
        load[r] += u.n_frames
    for r in range(world_size):
        out[r].sort(key=lambda u: (u.sequence, u.gop))
    return out


def reduce_stats(counters, device_ms, device=None):
    """counters: sequence of floats summed over ranks (bits, SSE, frames, ...); device_ms: this
    rank's timed-region duration, reduced with max.  Returns (list of sums, max ms).  With no
    process group (single process) it is the identity."""
    t = torch.tensor(list(counters), dtype=torch.float64, device=device)
    m = torch.tensor([device_ms], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        dist.all_reduce(m, op=dist.ReduceOp.MAX)
    return [float(v) for v in t.tolist()], float(m.item())
