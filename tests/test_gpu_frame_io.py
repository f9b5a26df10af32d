"""-m gpu: device-side frame ingest / egress (dcvc_frame_u8_to_view, dcvc_view_plane_sse) through
the C ABI against the reference's own definitions restated in numpy/scipy: astype(float32)/255,
scipy.ndimage.zoom(order=0), np.pad(edge) for F.pad(replicate), ycbcr444_to_420's 2x2 float32 mean +
clip (DCVC-DC/src/transforms/functional.py:61-92) and calc_psnr's float64 squared error
(src/utils/metrics.py:81-92).  u8 -> fp32 is bit-exact; SSE is float64 with a fixed reduction
order, compared at 1e-12 relative (luma / RGB) and 1e-6 (chroma: the order of the four float32
additions of the mean is numpy's business)."""
import numpy as np
import pytest
import torch

from dcvc_b200 import harness
from dcvc_b200.tensor import Act

from gpu_util import dev

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("fmt,h,w", [(harness.FMT_YUV420, 38, 52), (harness.FMT_RGB444, 37, 51),
                                     (harness.FMT_RGB444, 64, 96), (harness.FMT_YUV420, 1080, 1920)])
def test_u8_planes_to_padded_tensor(fmt, h, w):
    scipy_ndimage = pytest.importorskip("scipy.ndimage")
    rng = np.random.default_rng(h * w)
    if fmt == harness.FMT_YUV420:
        y = rng.integers(0, 256, (h, w), dtype=np.uint8)
        uv = rng.integers(0, 256, (2, h // 2, w // 2), dtype=np.uint8)
        raw = (y, uv)
        up = scipy_ndimage.zoom(uv.astype(np.float32) / 255, (1, 2, 2), order=0)
        ref = np.concatenate(((y.astype(np.float32) / 255)[None], up), 0)
    else:
        raw = rng.integers(0, 256, (3, h, w), dtype=np.uint8)
        ref = raw.astype(np.float32) / 255
    f = harness.FrameFeeder(fmt, h, w, dev())
    f.upload(0, harness.planes_of(raw, fmt))
    x = f.to_x(0)
    torch.cuda.synchronize()
    ph, pw = (h + 15) // 16 * 16, (w + 15) // 16 * 16
    assert tuple(x.shape) == (1, 3, ph, pw)
    want = np.pad(ref, ((0, 0), (0, ph - h), (0, pw - w)), mode="edge")
    assert np.array_equal(x[0].cpu().numpy(), want)


def _ref_420(yuv):
    """ycbcr444_to_420, functional.py:74-92"""
    c, h, w = yuv.shape
    u = np.mean(np.reshape(yuv[1:2], (1, h // 2, 2, w // 2, 2)), axis=(-1, -3))
    v = np.mean(np.reshape(yuv[2:3], (1, h // 2, 2, w // 2, 2)), axis=(-1, -3))
    return np.clip(yuv[0:1], 0., 1.), np.clip(np.concatenate((u, v), 0), 0., 1.)


def _sse(a, b):
    return float(np.sum(np.square(a.astype(np.float64) - b.astype(np.float64))))


@pytest.mark.parametrize("fmt,h,w", [(harness.FMT_YUV420, 38, 52), (harness.FMT_RGB444, 37, 51),
                                     (harness.FMT_YUV420, 1080, 1920)])
def test_plane_sse_and_decoded_planes(fmt, h, w):
    rng = np.random.default_rng(7 + h)
    ph, pw = (h + 15) // 16 * 16, (w + 15) // 16 * 16
    rec = (rng.random((3, ph, pw)).astype(np.float32) * 1.2 - 0.1)
    if fmt == harness.FMT_YUV420:
        y = rng.integers(0, 256, (h, w), dtype=np.uint8)
        uv = rng.integers(0, 256, (2, h // 2, w // 2), dtype=np.uint8)
        raw = (y, uv)
        y_rec, uv_rec = _ref_420(rec[:, :h, :w])
        want = [_sse(y.astype(np.float32) / 255, y_rec[0]),
                _sse(uv[0].astype(np.float32) / 255, uv_rec[0]),
                _sse(uv[1].astype(np.float32) / 255, uv_rec[1])]
        want_planes = np.concatenate((np.rint(y_rec * 255).reshape(-1),
                                      np.rint(uv_rec * 255).reshape(-1))).astype(np.uint8)
    else:
        raw = rng.integers(0, 256, (3, h, w), dtype=np.uint8)
        crop = np.clip(rec[:, :h, :w], 0., 1.)
        want = [_sse(raw[c].astype(np.float32) / 255, crop[c]) for c in range(3)]
        want_planes = np.rint(crop * 255).reshape(-1).astype(np.uint8)
    f = harness.FrameFeeder(fmt, h, w, dev())
    f.upload(1, harness.planes_of(raw, fmt))
    f.to_x(1)
    rec_act = Act(torch.from_numpy(np.ascontiguousarray(rec.transpose(1, 2, 0))).to(dev()).contiguous())
    rec_nchw = Act(torch.cat((rec_act.buf, torch.zeros(ph, pw, 1, device=dev())), 2).contiguous(),
                   0, 3).nchw()
    out = torch.zeros(f.nbytes, dtype=torch.uint8, device=dev())
    sse = f.distortion(1, rec_nchw, out).cpu().numpy()
    sse2 = f.distortion(1, rec_nchw).cpu().numpy()
    assert np.array_equal(sse, sse2)                       # fixed reduction order
    for c in range(3):
        tol = 1e-12 if (c == 0 or fmt == harness.FMT_RGB444) else 1e-6
        assert abs(sse[c] - want[c]) <= tol * want[c], (c, sse[c], want[c])
    got = out.cpu().numpy()
    n_luma = h * w if fmt == harness.FMT_YUV420 else 3 * h * w
    assert np.array_equal(got[:n_luma], want_planes[:n_luma])
    d = np.abs(got[n_luma:].astype(int) - want_planes[n_luma:].astype(int))
    assert d.size == 0 or (d.max() <= 1 and (d != 0).mean() < 1e-3)
    npx = f.plane_pixels()
    assert abs(harness.psnr_from_sse(sse[0], npx[0]) - 10 * np.log10(npx[0] / want[0])) < 1e-9
