"""Device activations as NHWC channel-slice views (`Act`), the unit every kernel works on.
fp32 everywhere a kernel other than the tensor-core convolution reads them; private conv -> conv
intermediates may be binary16 (`Act.new(..., f16=True)`, see dcvc_conv.in_f16 / out_f16).

An `Act` is (buffer [h, w, cstride] fp32 on the GPU, first channel, channel count).  Several Acts
may alias disjoint channel ranges of one buffer, which is how torch.cat on the reference's hot
path disappears.  `Act.nchw()` exposes the logical [1, C, H, W] tensor (a strided torch view of
the same memory) for the reference-facing API; `Act.wrap()` recovers the Act from such a view
without a copy, so dpb tensors round-trip through the caller untouched.
"""
import torch

from . import _lib


def _ru4(c):
    return (c + 3) // 4 * 4


# Destinations for tagged allocations: {tag: Act}.  dcvc_b200.graphs sets it around a frame so
# that the tensors a frame returns (the next frame's dpb) are written straight into buffers that
# outlive the call, instead of being copied there afterwards.  None = allocate as usual.
PLACEMENT = None


def placed(tag, h, w, c, f16=False):
    """the preassigned Act for `tag`, or None; the layout must be the one asked for"""
    a = PLACEMENT.get(tag) if PLACEMENT else None
    if a is None:
        return None
    assert (a.h, a.w, a.c, a.f16) == (h, w, c, f16), (tag, (a.h, a.w, a.c, a.f16), (h, w, c, f16))
    return a


class Act:
    __slots__ = ("buf", "c0", "c")

    def __init__(self, buf, c0=0, c=None):
        assert buf.dim() == 3 and buf.dtype in (torch.float32, torch.float16) and buf.is_contiguous()
        self.buf = buf
        self.c0 = c0
        self.c = buf.shape[2] - c0 if c is None else c
        assert 0 <= self.c0 and self.c0 + self.c <= buf.shape[2]

    # ------------------------------------------------------------------ constructors
    @staticmethod
    def new(h, w, c, device, cstride=None, zero=False, f16=False):
        cs = ((c + 7) // 8 * 8 if f16 else _ru4(c)) if cstride is None else cstride
        buf = (torch.zeros if zero else torch.empty)(
            (h, w, cs), dtype=torch.float16 if f16 else torch.float32, device=device)
        return Act(buf, 0, c)

    @staticmethod
    def wrap(x, stream=None):
        """[1,C,H,W] torch tensor -> Act.  Zero-copy when x is a view produced by Act.nchw()
        (or any channels-last tensor); otherwise one NCHW->NHWC kernel."""
        assert x.dim() == 4 and x.shape[0] == 1 and x.dtype == torch.float32 and x.is_cuda
        _, c, h, w = x.shape
        cs = x.stride(3)
        if x.stride(1) == 1 and cs >= c and x.stride(2) == w * cs:
            c0 = x.storage_offset() % cs
            base = x.storage_offset() - c0
            if x.untyped_storage().nbytes() // 4 >= base + h * w * cs:
                buf = x.as_strided((h, w, cs), (w * cs, cs, 1), base)
                return Act(buf, c0, c)
        out = Act.new(h, w, c, x.device)
        xc = x.contiguous()
        _lib.check(_lib.lib().dcvc_nchw_to_nhwc(xc.data_ptr(), out.view(), _stream(stream)))
        return out

    # ------------------------------------------------------------------ accessors
    @property
    def h(self):
        return self.buf.shape[0]

    @property
    def w(self):
        return self.buf.shape[1]

    @property
    def cs(self):
        return self.buf.shape[2]

    @property
    def f16(self):
        return self.buf.dtype == torch.float16

    @property
    def device(self):
        return self.buf.device

    def slice(self, c0, c):
        return Act(self.buf, self.c0 + c0, c)

    def view(self):
        return _lib.View(self.buf.data_ptr(), self.h, self.w, self.c, self.cs, self.c0)

    def nchw(self):
        """Logical [1,C,H,W] view of the same memory (no copy)."""
        return self.buf[:, :, self.c0:self.c0 + self.c].permute(2, 0, 1).unsqueeze(0)

    def __repr__(self):
        return f"Act({self.h}x{self.w}x{self.c} @{self.c0}/{self.cs})"


NULL_VIEW = _lib.View(None, 0, 0, 0, 0, 0)


def _stream(stream=None):
    if stream is not None:
        return stream
    return torch.cuda.current_stream().cuda_stream
