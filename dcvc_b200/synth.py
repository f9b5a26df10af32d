"""Synthetic test/benchmark content (there are no datasets or checkpoints offline).

Frames: a seeded low-pass noise field translated by a few pixels per frame, so that SpyNet and
the warps see real, recoverable motion (SURVEY.md section 8(d) "Frames").  8-bit quantised so that
file and tensor paths agree.
"""
import torch
import torch.nn.functional as F


def synth_frames(num, height, width, seed=0, step=(2, 3), layers=2):
    """-> float32 tensor [num, 3, height, width] in [0,1], values k/255.  CPU, deterministic."""
    g = torch.Generator(device="cpu")
    g.manual_seed(1000 + seed)
    margin_y = step[0] * num + 8
    margin_x = step[1] * num + 8
    fields = []
    for layer in range(layers):
        big_h, big_w = height + 2 * margin_y, width + 2 * margin_x
        scale = 8 * (layer + 1)
        low = torch.rand(1, 3, big_h // scale + 2, big_w // scale + 2, generator=g)
        fields.append(F.interpolate(low, size=(big_h, big_w), mode="bicubic", align_corners=False))
    frames = []
    for k in range(num):
        acc = 0
        for layer, f in enumerate(fields):
            sgn = 1 if layer % 2 == 0 else -1
            oy = margin_y + sgn * step[0] * k
            ox = margin_x + sgn * step[1] * k
            acc = acc + f[:, :, oy:oy + height, ox:ox + width] / len(fields)
        frames.append(acc)
    x = torch.cat(frames, 0).clamp(0, 1)
    return torch.round(x * 255.0) / 255.0


def rgb_to_ycbcr444(rgb):
    """BT.709 full-range RGB -> YCbCr (reference: DCVC-DC/src/transforms/functional.py:16-39)."""
    kr, kg, kb = 0.2126, 0.7152, 0.0722
    r, g, b = rgb[:, 0:1], rgb[:, 1:2], rgb[:, 2:3]
    y = kr * r + kg * g + kb * b
    cb = 0.5 * (b - y) / (1 - kb) + 0.5
    cr = 0.5 * (r - y) / (1 - kr) + 0.5
    return torch.cat((y, cb, cr), 1).clamp(0, 1)
