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

    def close(self):
        self.file.close()


class PNGWriter:
    def __init__(self, dst_path, width, height):
        from PIL import Image
        self._from = Image.fromarray
        os.makedirs(dst_path, exist_ok=True)
        self.dst_path, self.index = dst_path, 1

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

    def write_420(self, y, uv):
        for plane in (y, uv):
            self.file.write((plane.clamp(0, 1) * 255).round().to(torch.uint8).cpu().numpy().tobytes())

    def close(self):
        self.file.close()


# ----------------------------------------------------------------------------- device-side planes
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
def run_sequence(p_frame_net, i_frame_net, job):
    """One (sequence, rate) job; `job` has the keys the reference's run_test reads
    (test_video.py:76-258).  Returns the summary dict."""
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
    pixels = 0
    dpb = None
    start = time.time()
    with torch.no_grad():
        for frame_idx in range(job["frame_num"]):
            t0 = time.time()
            raw = reader.read()
            if raw is None:
                raise ValueError(f"{job['src_path']}: only {frame_idx} frames")
            if job["src_type"] == "yuv420":
                x, y_src, uv_src = yuv420_u8_to_x(raw[0], raw[1], device)
                if not yuv:
                    raise ValueError("RGB distortion from a yuv420 source is not supported")
            else:
                x = rgb_u8_to_x(raw, device)
            h, w = x.shape[2], x.shape[3]
            pixels = pixels or h * w
            assert pixels == h * w
            pl, pr, pt, pb = get_padding_size(h, w, 16)
            x_padded = F.pad(x, (pl, pr, pt, pb), mode="replicate")
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
            x_hat = recon[:, :, pt:pt + h, pl:pl + w]
            if yuv:
                y_rec, uv_rec = x_to_yuv420(x_hat)
                py, pu, pv = psnr(y_src, y_rec), psnr(uv_src[0], uv_rec[0]), psnr(uv_src[1], uv_rec[1])
                psnrs.append((6 * py + pu + pv) / 8)
                for c, v in zip("yuv", (py, pu, pv)):
                    planes[f"psnr_{c}"].append(v)
                    planes[f"msssim_{c}"].append(0.0)
                if writer:
                    writer.write_420(y_rec, uv_rec)
            else:
                mse = torch.mean((x_hat - x) ** 2)
                psnrs.append(float((20 * torch.log10(1 / torch.sqrt(mse))).item()))
                if writer:
                    writer.write_rgb(x_hat)
            msssims.append(0.0)
            if job.get("verbose", 0) >= 2:
                print(f"frame {frame_idx}, {time.time() - t0:.3f} seconds, bits: {bits[-1]:.3f}, "
                      f"PSNR: {psnrs[-1]:.4f}, MS-SSIM: {msssims[-1]:.4f} ")
    reader.close()
    if writer:
        writer.close()
    if job.get("verbose", 0) >= 1 and p_count:
        print(f"encoding/decoding {p_count} P frames, average encoding time "
              f"{enc_t / p_count * 1000:.0f} ms, average decoding time {dec_t / p_count * 1000:.0f} ms.")
    return summarise(frame_types, bits, psnrs, msssims, pixels, time.time() - start, planes)
