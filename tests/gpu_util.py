"""Helpers for the -m gpu parity tests: move oracle (CPU, NCHW) tensors into the product's NHWC
Acts and back."""
import torch

from dcvc_b200.tensor import Act


def dev():
    return torch.device("cuda:0")


def to_act(t, cstride=None, coff=0):
    """CPU [1,C,H,W] -> Act on the GPU (optionally inside a wider buffer at channel coff)."""
    _, c, h, w = t.shape
    cs = cstride if cstride is not None else (c + 3) // 4 * 4
    buf = torch.full((h, w, cs), float("nan"), dtype=torch.float32, device=dev())
    buf[:, :, coff:coff + c] = t[0].permute(1, 2, 0).to(dev())
    return Act(buf, coff, c)


def from_act(a):
    """Act -> CPU [1,C,H,W] contiguous"""
    return a.nchw().detach().cpu().contiguous()


def maxdiff(a, b):
    return float((a - b).abs().max())


def psnr(a, b):
    mse = torch.mean((a - b) ** 2).item()
    return 10 * torch.log10(torch.tensor(1.0 / max(mse, 1e-20))).item()
