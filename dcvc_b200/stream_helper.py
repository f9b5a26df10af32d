"""Per-frame file format, padding/latent sizes and checkpoint loading -- the on-disk contract of
the reference's DCVC-DC/src/utils/stream_helper.py:22-139, kept bit-for-bit so frame files are
interchangeable with the reference in compat coder mode.

I frame : >I height, >I width, B flag, >I length, payload      (13-byte header)
P frame : B flag, B frame_idx, >I length, payload              (6-byte header)
flag    = (q_in_ckpt << 7) | (q_index << 1)                    (1-bit flag, 6-bit rate index)

There are also in-memory variants (pack_i / unpack_i / pack_p / unpack_p) so a caller can keep
GOP streams off the filesystem.
"""
import os
import struct

import torch

from .common_model import get_downsampled_shape, get_padding_size  # noqa: F401 (re-export)

_I_HEAD = struct.Struct(">IIBI")
_P_HEAD = struct.Struct(">BBI")


def get_state_dict(ckpt_path):
    """Unwrap {'state_dict'| 'net'} and strip a DataParallel 'module.' prefix (:40-47)."""
    ckpt = torch.load(ckpt_path, map_location=torch.device("cpu"))
    for key in ("state_dict", "net"):
        if key in ckpt:
            ckpt = ckpt[key]
    return {(k[len("module."):] if k.startswith("module.") else k): v for k, v in ckpt.items()}


def filesize(filepath):
    if not os.path.isfile(filepath):
        raise ValueError(f'Invalid file "{filepath}".')
    return os.stat(filepath).st_size


def _flag(q_in_ckpt, q_index):
    return ((1 if q_in_ckpt else 0) << 7) + (int(q_index) << 1)


def _unflag(flag):
    return (flag >> 7) > 0, (flag & 0x7F) >> 1


def pack_i(height, width, q_in_ckpt, q_index, bit_stream):
    return _I_HEAD.pack(height, width, _flag(q_in_ckpt, q_index), len(bit_stream)) + bit_stream


def unpack_i(blob):
    height, width, flag, n = _I_HEAD.unpack_from(blob, 0)
    q_in_ckpt, q_index = _unflag(flag)
    payload = bytes(blob[_I_HEAD.size:_I_HEAD.size + n])
    if len(payload) != n:
        raise ValueError("truncated I-frame payload")
    return height, width, q_in_ckpt, q_index, payload


def pack_p(string, q_in_ckpt, q_index, frame_idx):
    return _P_HEAD.pack(_flag(q_in_ckpt, q_index), frame_idx, len(string)) + string


def unpack_p(blob):
    flag, frame_idx, n = _P_HEAD.unpack_from(blob, 0)
    q_in_ckpt, q_index = _unflag(flag)
    payload = bytes(blob[_P_HEAD.size:_P_HEAD.size + n])
    if len(payload) != n:
        raise ValueError("truncated P-frame payload")
    return q_in_ckpt, q_index, frame_idx, payload


def encode_i(height, width, q_in_ckpt, q_index, bit_stream, output):
    with open(output, "wb") as f:
        f.write(pack_i(height, width, q_in_ckpt, q_index, bit_stream))


def decode_i(inputpath):
    with open(inputpath, "rb") as f:
        return unpack_i(f.read())


def encode_p(string, q_in_ckpt, q_index, frame_idx, output):
    with open(output, "wb") as f:
        f.write(pack_p(string, q_in_ckpt, q_index, frame_idx))


def decode_p(inputpath):
    with open(inputpath, "rb") as f:
        return unpack_p(f.read())
