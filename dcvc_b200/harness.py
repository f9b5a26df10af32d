"""Frame ingest/egress, colour-plane handling and the per-sequence coding loop behind
test_video.py (the reference's DCVC-DC/test_video.py:76-258 `run_test`, its readers/writers in
src/utils/video_reader.py / video_writer.py, the 4:2:0 helpers of src/transforms/functional.py:61-92
and the PSNR of src/utils/metrics.py:81-92).

Differences by design: frames are uploaded as raw 8-bit planes and converted on the device
(u8 -> fp32, nearest chroma upsample, replicate pad); the 4:4:4 -> 4:2:0 fold and the per-plane
PSNR also run on the device, so only three scalars per frame come back -- the reference makes a
device -> numpy -> scipy round trip per frame (test_video.py:110-116,169-195).
"""
import os
import time

import numpy as np
import torch
import torch.nn.functional as F

from .stream_helper import get_padding_size


# ----------------------------------------------------------------------------- readers / writers
class PNGReader:
    """Numbered PNG folder, im1.png.. or im00001.png.. (video_reader.py:45-80)."""

    def __init__(self, src_path, width, height):
        from PIL import Image
        self._open = Image.open
        names = set(os.listdir(src_path))
        if "im1.png" in names:
            self.digits = 1
        elif "im00001.png" in names:
            self.digits = 5
        else:
            raise ValueError("unknown image naming convention; please specify")
        self.src_path, self.width, self.height, self.index = src_path, width, height, 1

    def read(self):
        """-> uint8 [3,H,W] RGB, or None at the end."""
        path = os.path.join(self.src_path, f"im{str(self.index).zfill(self.digits)}.png")
        if not os.path.exists(path):
            return None
        rgb = np.asarray(self._open(path).convert("RGB"))
        assert rgb.shape[0] == self.height and rgb.shape[1] == self.width
        self.index += 1
        return np.ascontiguousarray(rgb.transpose(2, 0, 1))

    def skip(self, n):
        self.index += n

    def close(self):
        self.index = 1


class YUV420Reader:
    """Raw planar 8-bit 4:2:0 file (video_reader.py:121-161)."""

    def __init__(self, src_path, width, height):
        if not src_path.endswith(".yuv"):
            src_path += ".yuv"
        self.file = open(src_path, "rb")
        self.width, self.height = width, height
        self.y_size, self.uv_size = width * height, width * height // 2

    def read(self):
        """-> (uint8 [H,W], uint8 [2,H/2,W/2]) or None."""
        y = self.file.read(self.y_size)
        uv = self.file.read(self.uv_size)
        if len(y) < self.y_size or len(uv) < self.uv_size:
            return None
        return (np.frombuffer(y, dtype=np.uint8).copy().reshape(self.height, self.width),
                np.frombuffer(uv, dtype=np.uint8).copy().reshape(2, self.height // 2, self.width // 2))

    def skip(self, n):
        self.file.seek(n * (self.y_size + self.uv_size), os.SEEK_CUR)

    def close(self):
        self.file.close()


class PNGWriter:
    def __init__(self, dst_path, width, height):
        from PIL import Image
        self._from = Image.fromarray
        os.makedirs(dst_path, exist_ok=True)
        self.dst_path, self.index = dst_path, 1

    def write_planes(self, planes, fmt, h, w):
        """uint8 vector R|G|B from dcvc_view_plane_sse"""
        img = planes.reshape(3, h, w).transpose(1, 2, 0)
        self._from(np.ascontiguousarray(img)).save(
            os.path.join(self.dst_path, f"im{str(self.index).zfill(5)}.png"))
        self.index += 1

    def write_rgb(self, x_hat):
        """x_hat: device [1,3,H,W] in [0,1]."""
        img = (x_hat[0].clamp(0, 1) * 255).round().to(torch.uint8).permute(1, 2, 0).cpu().numpy()
        self._from(img).save(os.path.join(self.dst_path, f"im{str(self.index).zfill(5)}.png"))
        self.index += 1

    def close(self):
        pass


class YUV420Writer:
    def __init__(self, dst_path, width, height):
        if not dst_path.endswith(".yuv"):
            dst_path += ".yuv"
        os.makedirs(os.path.dirname(dst_path) or ".", exist_ok=True)
        self.file = open(dst_path, "wb")

    def write_planes(self, planes, fmt, h, w):
        """uint8 vector Y|U|V (4:2:0) from dcvc_view_plane_sse"""
        self.file.write(planes.tobytes())

    def write_420(self, y, uv):
        for plane in (y, uv):
            self.file.write((plane.clamp(0, 1) * 255).round().to(torch.uint8).cpu().numpy().tobytes())

    def close(self):
        self.file.close()


# ----------------------------------------------------------------------------- device-side planes
FMT_RGB444, FMT_YUV420 = 0, 1


def planes_of(raw, fmt):
    """reader output -> one contiguous uint8 vector in the layout dcvc_frame_u8_to_view reads"""
    if fmt == FMT_RGB444:
        return np.ascontiguousarray(raw).reshape(-1)
    y, uv = raw
    return np.concatenate((np.ascontiguousarray(y).reshape(-1), np.ascontiguousarray(uv).reshape(-1)))


class FrameFeeder:
    """Pinned, double-buffered ingest of raw 8-bit frames and their conversion on the device.

    upload(slot, planes) copies one frame's planes (RGB 4:4:4 or YUV 4:2:0, uint8) into a pinned
    staging buffer and enqueues the host->device copy on a private copy stream; to_x(slot) makes
    the caller's stream wait for that copy and runs `dcvc_frame_u8_to_view` (u8 -> fp32 / 255,
    nearest chroma upsample, replicate pad to a multiple of 16) into a fresh NHWC buffer, returned
    as the [1,3,H,W] tensor the models take.  Frame k+1 uploads while frame k is being coded.
    distortion(slot, recon) runs `dcvc_view_plane_sse` against the same device-resident source
    planes: three float64 sums (per plane) stay on the device until read."""

    def __init__(self, fmt, h, w, device, slots=2):
        from . import _lib
        self.fmt, self.h, self.w, self.device = fmt, h, w, torch.device(device)
        self.nbytes = int(_lib.lib().dcvc_frame_bytes(fmt, h, w))
        if self.nbytes == 0:
            raise ValueError(f"4:2:0 frames must be even in both dimensions, got {w}x{h}")
        self.ph, self.pw = (h + 15) // 16 * 16, (w + 15) // 16 * 16
        self.pinned = [torch.empty(self.nbytes, dtype=torch.uint8).pin_memory() for _ in range(slots)]
        self.dev = [torch.empty(self.nbytes, dtype=torch.uint8, device=self.device)
                    for _ in range(slots)]
        self.copy_stream = torch.cuda.Stream(self.device)
        self.copied = [torch.cuda.Event() for _ in range(slots)]
        self.consumed = [None] * slots
        self._ws = torch.empty(int(_lib.lib().dcvc_plane_sse_ws_bytes(h, w)) // 8 + 1,
                               dtype=torch.float64, device=self.device)

    def upload(self, slot, planes):
        if self.consumed[slot] is not None:      # the device copy may still be read by a kernel
            self.copy_stream.wait_event(self.consumed[slot])
        self.copied[slot].synchronize()          # the pinned buffer may still be in flight
        self.pinned[slot].numpy()[:] = planes
        with torch.cuda.stream(self.copy_stream):
            self.dev[slot].copy_(self.pinned[slot], non_blocking=True)
            self.copied[slot].record(self.copy_stream)

    def to_x(self, slot):
        from . import _lib
        from .tensor import Act, _stream
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self.copied[slot])
        a = Act.new(self.ph, self.pw, 3, self.device)
        _lib.check(_lib.lib().dcvc_frame_u8_to_view(self.dev[slot].data_ptr(), self.fmt, self.h,
                                                    self.w, a.view(), _stream()))
        return a.nchw()

    def distortion(self, slot, recon, out_planes=None):
        """recon: [1,3,H,W] decoded (padded) frame -> float64 [3] device tensor of per-plane SSE;
        out_planes: optional uint8 device vector receiving the decoded planes (writers)."""
        from . import _lib
        from .tensor import Act, _stream
        sse = torch.empty(3, dtype=torch.float64, device=self.device)
        _lib.check(_lib.lib().dcvc_view_plane_sse(
            Act.wrap(recon).view(), self.fmt, self.h, self.w, self.dev[slot].data_ptr(),
            sse.data_ptr(), out_planes.data_ptr() if out_planes is not None else None,
            self._ws.data_ptr(), _stream()))
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.consumed[slot] = ev
        return sse

    def plane_pixels(self):
        n = self.h * self.w
        return (n, n, n) if self.fmt == FMT_RGB444 else (n, n // 4, n // 4)


def psnr_from_sse(sse, pixels, data_range=1.0):
    """metrics.py:81-92: mse in float64; 999.9 for identical planes"""
    mse = sse / pixels
    return 10.0 * float(np.log10(data_range * data_range / mse)) if mse > 1e-10 else 999.9


def rgb_u8_to_x(rgb_u8, device):
    """uint8 [3,H,W] host -> fp32 [1,3,H,W] in [0,1] on the device."""
    t = torch.from_numpy(rgb_u8).to(device, non_blocking=True)
    return (t.to(torch.float32) / 255.0).unsqueeze(0)


def yuv420_u8_to_x(y_u8, uv_u8, device):
    """4:2:0 planes -> (x [1,3,H,W] 4:4:4 with nearest chroma upsample, y [H,W], uv [2,H/2,W/2]).
    scipy.ndimage.zoom(uv, (1,2,2), order=0) of functional.py:61-71 repeats every chroma sample
    2x2 (out index i reads in index round(i*(n-1)/(2n-1)) = i // 2)."""
    y = torch.from_numpy(np.ascontiguousarray(y_u8)).to(device, non_blocking=True).float() / 255.0
    uv = torch.from_numpy(np.ascontiguousarray(uv_u8)).to(device, non_blocking=True).float() / 255.0
    up = uv.repeat_interleave(2, dim=1).repeat_interleave(2, dim=2)
    return torch.cat((y.unsqueeze(0), up), 0).unsqueeze(0), y, uv


def x_to_yuv420(x_hat):
    """[1,3,H,W] -> (y [H,W], uv [2,H/2,W/2]): 2x2 chroma mean, clipped (functional.py:74-92)."""
    y = x_hat[0, 0].clamp(0, 1)
    uv = F.avg_pool2d(x_hat[:, 1:3], 2)[0].clamp(0, 1)
    return y, uv


def psnr(a, b, data_range=1.0):
    """metrics.py:81-92 on the device in float64; 999.9 for identical planes."""
    mse = torch.mean((a.double() - b.double()) ** 2)
    m = float(mse.item())
    return 10.0 * float(np.log10(data_range * data_range / m)) if m > 1e-10 else 999.9


# ----------------------------------------------------------------------------- summary
def summarise(frame_types, bits, psnrs, msssims, frame_pixel_num, test_time, planes=None):
    """Keys and arithmetic of generate_log_json (src/utils/common.py:44-157).  planes: optional
    dict name -> per-frame list for the _y/_u/_v variants of psnr and msssim."""
    n = len(frame_types)
    out = {"frame_pixel_num": frame_pixel_num}
    sel = {"i": [k for k in range(n) if frame_types[k] == 0],
           "p": [k for k in range(n) if frame_types[k] != 0]}
    out["i_frame_num"], out["p_frame_num"] = len(sel["i"]), len(sel["p"])
    series = {"psnr": psnrs, "msssim": msssims}
    if planes:
        series.update(planes)

    def put(tag, idx, denom_frames):
        out[f"ave_{tag}_frame_bpp"] = (sum(bits[k] for k in idx) / (denom_frames * frame_pixel_num)
                                       if denom_frames else 0)
        for name, vals in series.items():
            out[f"ave_{tag}_frame_{name}"] = (sum(vals[k] for k in idx) / denom_frames
                                              if denom_frames else 0)

    put("i", sel["i"], len(sel["i"]))
    out["test_time"] = test_time
    put("p", sel["p"], len(sel["p"]))
    put("all", list(range(n)), n)
    return out


# ----------------------------------------------------------------------------- coding loop
def merge_units(parts):
    """Per-frame records of the GOP units of ONE job (run_sequence(..., summary=False)), in any
    order -> the job's summary dict.  test_time is the sum of the units' times (what one process
    would have spent)."""
    parts = sorted(parts, key=lambda r: r["first_frame"])
    cat = lambda key: [v for r in parts for v in r[key]]          # noqa: E731
    planes = None
    if parts[0]["planes"] is not None:
        planes = {k: [v for r in parts for v in r["planes"][k]] for k in parts[0]["planes"]}
    return summarise(cat("frame_types"), cat("bits"), cat("psnrs"), cat("msssims"),
                     parts[0]["pixels"], sum(r["test_time"] for r in parts), planes)


def run_sequence(p_frame_net, i_frame_net, job, first_frame=0, frame_count=None, summary=True):
    """One (sequence, rate) job; `job` has the keys the reference's run_test reads
    (test_video.py:76-258).  Returns the summary dict.  first_frame / frame_count select a run of
    whole GOPs of the sequence (first_frame a multiple of gop_size): the unit of multi-GPU work,
    since every GOP restarts from an intra frame (test_video.py:140-159); with summary=False the
    per-frame records come back for merge_units()."""
    device = next(i_frame_net.parameters()).device
    yuv = bool(job["dist_in_yuv420"])
    if job["src_type"] == "png":
        reader = PNGReader(job["src_path"], job["src_width"], job["src_height"])
    elif job["src_type"] == "yuv420":
        reader = YUV420Reader(job["src_path"], job["src_width"], job["src_height"])
    else:
        raise ValueError(f"unknown src_type {job['src_type']!r}")
    if yuv and job["src_type"] != "yuv420":
        raise ValueError("--yuv420 needs a yuv420 source")
    writer = None
    if job.get("save_decoded_frame"):
        cls = YUV420Writer if yuv else PNGWriter
        writer = cls(job["recon_path"], job["src_width"], job["src_height"])
    write_stream = bool(job.get("write_stream"))
    q_in_ckpt, q_index = job["q_in_ckpt"], job["i_frame_q_index"]
    frame_types, bits, psnrs, msssims = [], [], [], []
    planes = {f"psnr_{c}": [] for c in "yuv"} if yuv else None
    if yuv:
        planes.update({f"msssim_{c}": [] for c in "yuv"})
    enc_t = dec_t = 0.0
    p_count = 0
    dpb = None
    fmt = FMT_YUV420 if job["src_type"] == "yuv420" else FMT_RGB444
    if job["src_type"] == "yuv420" and not yuv:
        raise ValueError("RGB distortion from a yuv420 source is not supported")
    h, w = job["src_height"], job["src_width"]
    pixels = h * w
    feeder = FrameFeeder(fmt, h, w, device)
    pl, pr, pt, pb = get_padding_size(h, w, 16)
    assert pl == 0 and pt == 0          # the reference pads right / bottom only
    if first_frame % max(job["gop_size"], 1) != 0:
        raise ValueError("a unit must start on an intra frame")
    n_frames = job["frame_num"] - first_frame if frame_count is None else frame_count
    reader.skip(first_frame)
    sse_host = torch.zeros((n_frames, 3), dtype=torch.float64).pin_memory()
    out_planes = torch.empty(feeder.nbytes, dtype=torch.uint8, device=device) if writer else None
    verbose = job.get("verbose", 0)

    def fetch(idx):
        raw = reader.read()
        if raw is None:
            raise ValueError(f"{job['src_path']}: only {idx} frames")
        feeder.upload(idx % 2, planes_of(raw, fmt))

    start = time.time()
    with torch.no_grad():
        fetch(0)
        for k in range(n_frames):
            frame_idx = first_frame + k
            t0 = time.time()
            slot = k % 2
            x_padded = feeder.to_x(slot)
            if k + 1 < n_frames:
                fetch(k + 1)                 # next frame: disk -> pinned -> device while this one codes
            bin_path = os.path.join(job["bin_folder"], f"{frame_idx}.bin") if write_stream else None
            if frame_idx % job["gop_size"] == 0 or p_frame_net is None:
                res = i_frame_net.encode_decode(x_padded, q_in_ckpt, q_index, bin_path,
                                                pic_height=h, pic_width=w)
                dpb = {"ref_frame": res["x_hat"], "ref_feature": None, "ref_mv_feature": None,
                       "ref_y": None, "ref_mv_y": None}
                recon = res["x_hat"]
                frame_types.append(0)
            else:
                res = p_frame_net.encode_decode(x_padded, dpb, q_in_ckpt, q_index, bin_path,
                                                pic_height=h, pic_width=w,
                                                frame_idx=frame_idx % 4)
                dpb = res["dpb"]
                recon = dpb["ref_frame"]
                frame_types.append(1)
                p_count += 1
                enc_t += res["encoding_time"]
                dec_t += res["decoding_time"]
            bits.append(float(res["bit"]))
            recon = recon.clamp_(0, 1)          # in place: this tensor IS dpb["ref_frame"]
            sse = feeder.distortion(slot, recon, out_planes)
            sse_host[k].copy_(sse, non_blocking=True)             # read after the loop
            if writer:
                writer.write_planes(out_planes.cpu().numpy(), fmt, h, w)
            if verbose >= 2:
                torch.cuda.synchronize(device)
                print(f"frame {frame_idx}, {time.time() - t0:.3f} seconds, bits: {bits[-1]:.3f}, "
                      f"SSE per plane: {sse_host[k].tolist()} ")
    torch.cuda.synchronize(device)
    npx = feeder.plane_pixels()
    for k in range(n_frames):
        e = [float(v) for v in sse_host[k]]
        if yuv:
            py, pu, pv = (psnr_from_sse(e[c], npx[c]) for c in range(3))
            psnrs.append((6 * py + pu + pv) / 8)
            for c, v in zip("yuv", (py, pu, pv)):
                planes[f"psnr_{c}"].append(v)
                planes[f"msssim_{c}"].append(0.0)
        else:
            # PSNR(x_hat, x) of test_video.py:65-68: one mean over the three planes
            psnrs.append(psnr_from_sse(sum(e), 3 * pixels))
        msssims.append(0.0)
    reader.close()
    if writer:
        writer.close()
    if job.get("verbose", 0) >= 1 and p_count:
        print(f"encoding/decoding {p_count} P frames, average encoding time "
              f"{enc_t / p_count * 1000:.0f} ms, average decoding time {dec_t / p_count * 1000:.0f} ms.")
    if not summary:
        return {"first_frame": first_frame, "frame_types": frame_types, "bits": bits,
                "psnrs": psnrs, "msssims": msssims, "planes": planes, "pixels": pixels,
                "test_time": time.time() - start}
    return summarise(frame_types, bits, psnrs, msssims, pixels, time.time() - start, planes)
