"""DCVC-HEM frame files and q rounding (reference DCVC-HEM/src/utils/stream_helper.py:41-45,
102-143): big-endian headers; the continuous q_scale travels as round(q * 100) in a uint16.
    I frame: uint32 height, uint32 width, uint16 q_index, uint32 length, bytes
    P frame: uint16 mv_y_q_index, uint16 y_q_index, uint32 length, bytes
The bytes are one bare rANS stream (no part header)."""
import struct
from pathlib import Path

import numpy as np


def get_rounded_q(q_scale):
    """:41-45 -> (q_scale rounded to 1/100 inside [0.01, 655], its uint16 index)"""
    q_scale = np.clip(float(q_scale), 0.01, 655.)
    q_index = int(np.round(q_scale * 100))
    return q_index / 100, q_index


def filesize(filepath):
    if not Path(filepath).is_file():
        raise ValueError(f'Invalid file "{filepath}".')
    return Path(filepath).stat().st_size


def encode_i(height, width, q_index, bit_stream, output):
    with Path(output).open("wb") as f:
        f.write(struct.pack(">2I", height, width))
        f.write(struct.pack(">H", q_index))
        f.write(struct.pack(">I", len(bit_stream)))
        f.write(bit_stream)


def decode_i(inputpath):
    with Path(inputpath).open("rb") as f:
        height, width = struct.unpack(">2I", f.read(8))
        q_index, = struct.unpack(">H", f.read(2))
        n, = struct.unpack(">I", f.read(4))
        return height, width, q_index, f.read(n)


def encode_p(string, mv_y_q_index, y_q_index, output):
    with Path(output).open("wb") as f:
        f.write(struct.pack(">2H", mv_y_q_index, y_q_index))
        f.write(struct.pack(">I", len(string)))
        f.write(string)


def decode_p(inputpath):
    with Path(inputpath).open("rb") as f:
        mv_y_q_index, y_q_index = struct.unpack(">2H", f.read(4))
        n, = struct.unpack(">I", f.read(4))
        return mv_y_q_index, y_q_index, f.read(n)


def interpolate_log(min_val, max_val, num, decending=True):
    """DCVC-HEM/src/utils/common.py:23-31: `num` log-spaced q_scales between the checkpoint's
    extreme anchors (the rate sweep of DCVC-HEM/test_video.py:286-320)."""
    assert max_val > min_val > 0
    ends = (np.log(max_val), np.log(min_val)) if decending else (np.log(min_val), np.log(max_val))
    return np.exp(np.linspace(ends[0], ends[1], num))


def get_padding_size(height, width, p=64):
    """stream_helper.py:22-31; the HEM harness pads with ZEROS to a multiple of 64
    (DCVC-HEM/test_video.py:114-120)"""
    new_h = (height + p - 1) // p * p
    new_w = (width + p - 1) // p * p
    return 0, new_w - width, 0, new_h - height
