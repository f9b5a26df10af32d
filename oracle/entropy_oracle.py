"""TEST INFRASTRUCTURE -- CPU restatement of the reference's entropy model (torch fp32 CPU).

Restates, for the DCVC-DC path (paths relative to /root/reference/DCVC-DC):
  * CDF tables                 src/models/entropy_models.py:124-178 (factorised prior),
                               :228-267 (256-level Laplace / Gaussian scale table)
  * scale -> table index       src/models/entropy_models.py:269-273
  * quad-part spatial prior    src/models/common_model.py:92-321
  * bit estimates              src/models/common_model.py:39-61
Not used by the product.  Pinned against the reference by tests/test_oracle_golden.py.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

from . import rans_oracle

SCALE_LEVELS = 256
SCALE_MAX = 64.0
SCALE_MIN = {"laplace": 0.01, "gaussian": 0.11}

# step s codes channel-group g at mask position MASK_OF[s][g]  (common_model.py:168-220);
# mask m is 1 at (row&1, col&1) == (m>>1, m&1)               (common_model.py:102-129)
MASK_OF = ((0, 1, 2, 3), (3, 2, 1, 0), (2, 3, 0, 1), (1, 0, 3, 2))


def _cdf_fn(distribution):
    if distribution == "laplace":
        return torch.distributions.laplace.Laplace
    return torch.distributions.normal.Normal


def pmf_rows_to_cdf(pmf, tail_mass, pmf_length, max_length):
    """entropy_models.py:24-32"""
    cdf = torch.zeros((len(pmf_length), max_length + 2), dtype=torch.int32)
    for i in range(len(pmf_length)):
        prob = torch.cat((pmf[i][: int(pmf_length[i])], tail_mass[i]), dim=0)
        row = rans_oracle.pmf_to_quantized_cdf(prob.tolist(), 16)
        cdf[i, : row.size] = torch.from_numpy(row.astype(np.int32))
    return cdf


def scale_table(distribution):
    """entropy_models.py:224-226"""
    return torch.exp(torch.linspace(math.log(SCALE_MIN[distribution]), math.log(SCALE_MAX),
                                    SCALE_LEVELS))


def gaussian_tables(distribution):
    """GaussianEncoder.update, entropy_models.py:228-267 -> (cdf int32[256,L], sizes, offsets)"""
    table = scale_table(distribution)
    dist = _cdf_fn(distribution)
    center = torch.zeros_like(table) + 50
    d = dist(torch.zeros_like(table), table)
    for i in range(50, 1, -1):
        probs = d.cdf(torch.zeros_like(center) + i)
        center = torch.where(probs > 0.9999, torch.zeros_like(center) + i, center)
    center = center.int()
    length = 2 * center + 1
    max_length = int(length.max())
    samples = (torch.arange(max_length) - center[:, None]).float()
    scales = torch.zeros_like(samples) + table[:, None]
    d = dist(torch.zeros_like(scales), scales)
    upper = d.cdf(samples + 0.5)
    lower = d.cdf(samples - 0.5)
    pmf = upper - lower
    tail = 2 * lower[:, :1]
    cdf = pmf_rows_to_cdf(pmf, tail, length, max_length)
    return cdf.numpy(), (length + 2).numpy().astype(np.int32), (-center).numpy().astype(np.int32)


def factorized_cdf(sd, p, x):
    """BitEstimator.get_cdf, entropy_models.py:58-78,114-122.  x: [N,C,H,W]."""
    for k in ("f1", "f2", "f3"):
        x = x * F.softplus(sd[f"{p}.{k}.h"]) + sd[f"{p}.{k}.b"]
        x = x + torch.tanh(x) * torch.tanh(sd[f"{p}.{k}.a"])
    x = x * F.softplus(sd[f"{p}.f4.h"]) + sd[f"{p}.f4.b"]
    return torch.sigmoid(x)


def factorized_tables(sd, p):
    """BitEstimator.update, entropy_models.py:124-178"""
    channel = sd[p + ".f1.h"].numel()
    medians = torch.zeros(channel)
    minima = medians + 50
    for i in range(50, 1, -1):
        probs = factorized_cdf(sd, p, (torch.zeros_like(medians) - i)[None, :, None, None])
        probs = torch.squeeze(probs)
        minima = torch.where(probs < 0.0001, torch.zeros_like(medians) + i, minima)
    maxima = medians + 50
    for i in range(50, 1, -1):
        probs = factorized_cdf(sd, p, (torch.zeros_like(medians) + i)[None, :, None, None])
        probs = torch.squeeze(probs)
        maxima = torch.where(probs > 0.9999, torch.zeros_like(medians) + i, maxima)
    minima = minima.int()
    maxima = maxima.int()
    offset = -minima
    pmf_start = medians - minima
    pmf_length = maxima + minima + 1
    max_length = int(pmf_length.max())
    samples = torch.arange(max_length)
    samples = samples[None, :] + pmf_start[:, None, None]   # [C,1,L]
    lower = factorized_cdf(sd, p, samples - 0.5).squeeze(0)
    upper = factorized_cdf(sd, p, samples + 0.5).squeeze(0)
    pmf = (upper - lower)[:, 0, :]
    tail = lower[:, 0, :1] + (1.0 - upper[:, 0, -1:])
    cdf = pmf_rows_to_cdf(pmf, tail, pmf_length, max_length)
    return cdf.numpy(), (pmf_length + 2).numpy().astype(np.int32), offset.numpy().astype(np.int32)


def build_indexes(scales, distribution):
    """GaussianEncoder.build_indexes, entropy_models.py:269-273"""
    log_min = math.log(SCALE_MIN[distribution])
    step = (math.log(SCALE_MAX) - log_min) / (SCALE_LEVELS - 1)
    scales = torch.maximum(scales, torch.zeros_like(scales) + 1e-5)
    idx = (torch.log(scales) - log_min) / step
    return idx.clamp_(0, SCALE_LEVELS - 1).int()


def masks(h, w):
    """get_mask_four_parts, common_model.py:102-129"""
    r = torch.arange(h).view(h, 1) & 1
    c = torch.arange(w).view(1, w) & 1
    pos = r * 2 + c
    return [(pos == m).float().view(1, 1, h, w) for m in range(4)]


def probs_to_bits(probs):
    """common_model.py:39-43"""
    return torch.clamp_min(-1.0 * torch.log(probs + 1e-5) / math.log(2.0), 0)


def y_bits(y_q, sigma, distribution):
    """get_y_laplace_bits / get_y_gaussian_bits, common_model.py:45-57"""
    sigma = sigma.clamp(1e-5, 1e10)
    d = _cdf_fn(distribution)(torch.zeros_like(sigma), sigma)
    return probs_to_bits(d.cdf(y_q + 0.5) - d.cdf(y_q - 0.5))


def z_bits(sd, p, z_hat):
    """get_z_bits, common_model.py:59-61"""
    return probs_to_bits(factorized_cdf(sd, p, z_hat + 0.5) - factorized_cdf(sd, p, z_hat - 0.5))


def pad_for_y(y):
    """common_model.py:70-78 -> (padded, (pad_right, pad_bottom))"""
    h, w = y.shape[2:]
    pr = (w + 3) // 4 * 4 - w
    pb = (h + 3) // 4 * 4 - h
    return F.pad(y, (0, pr, 0, pb), mode="replicate"), (pr, pb)


def slice_to_y(params, pads):
    """common_model.py:85-86 (negative pad == crop)"""
    pr, pb = pads
    return F.pad(params, (0, -pr, 0, -pb))


def quad_forward(y, common_params, prior_fn):
    """forward_four_part_prior, common_model.py:142-252.

    prior_fn(step, cat(y_hat_so_far, common_params)) -> [1,2C,h,w] is the reference's
    y_spatial_prior(y_spatial_prior_adaptor_{step}(.)), step in 1..3.
    Returns dict(y_q, y_hat, scales_hat, y_q_w[4], scales_w[4]) with the reference's meanings
    (y_hat already multiplied back by quant_step).
    """
    quant_step, scales, means = common_params.chunk(3, 1)
    h, w = y.shape[2:]
    m = masks(h, w)
    quant_step = torch.clamp_min(quant_step, 0.5)
    y = y / quant_step
    yg = y.chunk(4, 1)
    sg = scales.chunk(4, 1)
    mg = means.chunk(4, 1)
    y_hat_so_far = None
    y_q_parts = [0, 0, 0, 0]
    s_parts = [0, 0, 0, 0]
    y_q_w, scales_w = [], []
    for step in range(4):
        if step > 0:
            out = prior_fn(step, torch.cat((y_hat_so_far, common_params), 1)).chunk(8, 1)
            sg, mg = out[:4], out[4:]
        hats = []
        qw, sw = 0, 0
        for g in range(4):
            mk = m[MASK_OF[step][g]]
            s_hat = sg[g] * mk
            m_hat = mg[g] * mk
            y_q = torch.round((yg[g] - m_hat) * mk)
            hats.append(y_q + m_hat)
            y_q_parts[g] = y_q_parts[g] + y_q
            s_parts[g] = s_parts[g] + s_hat
            qw = qw + y_q
            sw = sw + s_hat
        cur = torch.cat(hats, 1)
        y_hat_so_far = cur if y_hat_so_far is None else y_hat_so_far + cur
        y_q_w.append(qw)
        scales_w.append(sw)
    return {
        "y_q": torch.cat(y_q_parts, 1), "scales_hat": torch.cat(s_parts, 1),
        "y_hat": y_hat_so_far * quant_step, "y_q_w": y_q_w, "scales_w": scales_w,
    }


def quad_decompress(common_params, prior_fn, decode_fn):
    """decompress_four_part_prior, common_model.py:261-321.
    decode_fn(scales_r[1,C/4,h,w]) -> decoded symbols, same shape (float)."""
    quant_step, scales, means = common_params.chunk(3, 1)
    h, w = means.shape[2:]
    m = masks(h, w)
    quant_step = torch.clamp_min(quant_step, 0.5)
    sg = scales.chunk(4, 1)
    mg = means.chunk(4, 1)
    y_hat_so_far = None
    for step in range(4):
        if step > 0:
            out = prior_fn(step, torch.cat((y_hat_so_far, common_params), 1)).chunk(8, 1)
            sg, mg = out[:4], out[4:]
        scales_r = 0
        for g in range(4):
            scales_r = scales_r + sg[g] * m[MASK_OF[step][g]]
        y_q_r = decode_fn(scales_r)
        cur = torch.cat([(y_q_r + mg[g]) * m[MASK_OF[step][g]] for g in range(4)], 1)
        y_hat_so_far = cur if y_hat_so_far is None else y_hat_so_far + cur
    return y_hat_so_far * quant_step
