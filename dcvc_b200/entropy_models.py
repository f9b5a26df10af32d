"""Entropy-model host side: CDF tables, scale->index thresholds and the coder facade.

Mirrors the reference's DCVC-DC/src/models/entropy_models.py: `EntropyCoder` (:9-55),
`BitEstimator.update` (:124-178), `GaussianEncoder.update` (:203-267) and `build_indexes`
(:269-273).  Tables are built once per model on the host in fp32 exactly as the reference does
when its model lives on the CPU (the table content defines the bitstream), then uploaded; the
per-frame work (index lookup, quantisation, coding) runs in the CUDA kernels and never returns
to the host except for the finished byte stream.
"""
import ctypes
import os
import math

import numpy as np
import torch
import torch.nn.functional as F

from . import _lib
from .ops import _Prof

SCALE_MIN = {"laplace": 0.01, "gaussian": 0.11}
SCALE_MAX = 64.0
SCALE_LEVELS = 256
DIST_ID = {"laplace": 0, "gaussian": 1}


# DCVC_RANS_SIDE_STREAM=0: the encoder's coder kernels stay on the caller's stream
SIDE_STREAM = os.environ.get("DCVC_RANS_SIDE_STREAM", "1") != "0"


def pmf_to_quantized_cdf(pmf, precision=16):
    """MLCodec_CXX.pmf_to_quantized_cdf (ops.cpp:84-91) through the C ABI (host function)."""
    p = np.ascontiguousarray(np.asarray(pmf, dtype=np.float32))
    cdf = np.zeros(p.size + 1, dtype=np.uint32)
    _lib.check(_lib.lib().dcvc_pmf_to_quantized_cdf(p.ctypes.data, p.size, precision,
                                                    cdf.ctypes.data))
    return cdf


def _rows_to_cdf(pmf, tail_mass, pmf_length, max_length):
    cdf = np.zeros((len(pmf_length), max_length + 2), dtype=np.int32)
    for i in range(len(pmf_length)):
        n = int(pmf_length[i])
        prob = torch.cat((pmf[i][:n], tail_mass[i]), dim=0)
        row = pmf_to_quantized_cdf(prob.numpy())
        cdf[i, :row.size] = row.astype(np.int32)
    return cdf


def _distribution(name, scale):
    zero = torch.zeros_like(scale)
    if name == "laplace":
        return torch.distributions.laplace.Laplace(zero, scale)
    return torch.distributions.normal.Normal(zero, scale)


def scale_table(distribution):
    return torch.exp(torch.linspace(math.log(SCALE_MIN[distribution]), math.log(SCALE_MAX),
                                    SCALE_LEVELS))


def gaussian_tables(distribution):
    """GaussianEncoder.update: per scale level, support half-width = smallest i in [2,50] whose
    CDF exceeds 0.9999 (else 50), integer PMF, two-sided tail mass -> quantised CDF rows."""
    table = scale_table(distribution)
    dist = _distribution(distribution, table)
    half = torch.full_like(table, 50.0)
    for i in range(50, 1, -1):
        covered = dist.cdf(torch.full_like(table, float(i))) > 0.9999
        half = torch.where(covered, torch.full_like(table, float(i)), half)
    half = half.int()
    length = 2 * half + 1
    max_length = int(length.max())
    samples = (torch.arange(max_length) - half[:, None]).float()
    dist2 = _distribution(distribution, torch.zeros_like(samples) + table[:, None])
    upper = dist2.cdf(samples + 0.5)
    lower = dist2.cdf(samples - 0.5)
    cdf = _rows_to_cdf(upper - lower, 2 * lower[:, :1], length, max_length)
    return cdf, (length + 2).numpy().astype(np.int32), (-half).numpy().astype(np.int32)


def factorized_cdf(params, x):
    """BitEstimator.get_cdf: params = dict f1..f4 -> (h, b[, a]) tensors [1,C,1,1]; x [N,C,H,W]."""
    for k in ("f1", "f2", "f3"):
        h, b, a = params[k]
        x = x * F.softplus(h) + b
        x = x + torch.tanh(x) * torch.tanh(a)
    h, b = params["f4"]
    return torch.sigmoid(x * F.softplus(h) + b)


def factorized_tables(params):
    """BitEstimator.update on the host: per channel, support = [-(smallest i with cdf(-i) < 1e-4),
    smallest i with cdf(i) > 0.9999] searched downward from 50."""
    channel = params["f1"][0].numel()
    zeros = torch.zeros(channel)
    lo = zeros + 50
    hi = zeros + 50
    for i in range(50, 1, -1):
        p = factorized_cdf(params, (zeros - i)[None, :, None, None]).reshape(-1)
        lo = torch.where(p < 0.0001, zeros + i, lo)
    for i in range(50, 1, -1):
        p = factorized_cdf(params, (zeros + i)[None, :, None, None]).reshape(-1)
        hi = torch.where(p > 0.9999, zeros + i, hi)
    lo, hi = lo.int(), hi.int()
    length = hi + lo + 1
    max_length = int(length.max())
    samples = torch.arange(max_length)[None, :] + (zeros - lo)[:, None, None]   # [C,1,L]
    lower = factorized_cdf(params, samples - 0.5).squeeze(0)
    upper = factorized_cdf(params, samples + 0.5).squeeze(0)
    pmf = (upper - lower)[:, 0, :]
    tail = lower[:, 0, :1] + (1.0 - upper[:, 0, -1:])
    cdf = _rows_to_cdf(pmf, tail, length, max_length)
    return cdf, (length + 2).numpy().astype(np.int32), lo.neg().numpy().astype(np.int32)


def _index_of(scales, distribution):
    """GaussianEncoder.build_indexes evaluated with torch CPU ops (the definition)."""
    log_min = math.log(SCALE_MIN[distribution])
    step = (math.log(SCALE_MAX) - log_min) / (SCALE_LEVELS - 1)
    s = torch.maximum(scales, torch.zeros_like(scales) + 1e-5)
    return ((torch.log(s) - log_min) / step).clamp_(0, SCALE_LEVELS - 1).int()


def index_edges(distribution):
    """edges[k] = smallest fp32 scale whose table index is >= k (edges[0] = 0).  The device
    kernel's index = (#edges <= scale) - 1 is then bit-identical to build_indexes for every
    fp32 input, with no dependence on the device's logf rounding."""
    k = torch.arange(1, SCALE_LEVELS, dtype=torch.int32)
    lo = torch.full((SCALE_LEVELS - 1,), int(np.float32(1e-5).view(np.int32)), dtype=torch.int32)
    hi = torch.full((SCALE_LEVELS - 1,), int(np.float32(128.0).view(np.int32)), dtype=torch.int32)
    # invariant: index(lo) < k <= index(hi)   (positive floats order like their bit patterns)
    for _ in range(32):
        mid = lo + (hi - lo) // 2
        ge = _index_of(mid.view(torch.float32), distribution) >= k
        hi = torch.where(ge, mid, hi)
        lo = torch.where(ge, lo, mid)
    return torch.cat((torch.zeros(1), hi.view(torch.float32)))


def search_lut(cdf, sizes):
    """The interleaved / compat decoders' search table (dcvc_tables.lut): lut[r][b] = the largest
    symbol s of row r with cdf[r][s] <= 256 b, i.e. the symbol the first cumulative value of
    bucket b falls in; the symbol of any value of the bucket then lies in [lut[b], lut[b + 1]].
    uint8 [rows][256], or None when a row has more than 256 symbols (stride > 257)."""
    if cdf.shape[1] > 257:
        return None
    lut = np.zeros((cdf.shape[0], 256), dtype=np.uint8)
    starts = np.arange(256, dtype=np.int64) * 256
    for r in range(cdf.shape[0]):
        row = np.asarray(cdf[r, :max(int(sizes[r]) - 1, 1)], dtype=np.int64)
        lut[r] = np.clip(np.searchsorted(row, starts, side="right") - 1, 0, 255)
    return lut


class DeviceTables:
    """One table set {cdf, sizes, offsets} resident on the device."""

    def __init__(self, cdf, sizes, offsets, device):
        self.cdf_np, self.sizes_np, self.offsets_np = cdf, sizes, offsets
        self.cdf = torch.from_numpy(np.ascontiguousarray(cdf, dtype=np.int32)).to(device)
        self.sizes = torch.from_numpy(np.ascontiguousarray(sizes, dtype=np.int32)).to(device)
        self.offsets = torch.from_numpy(np.ascontiguousarray(offsets, dtype=np.int32)).to(device)
        # search table of the interleaved decoder (dcvc_tables.lut): per row and 256-wide bucket of
        # the 16-bit cumulative value, the symbol its first value falls in
        self.lut = None
        lut = search_lut(cdf, sizes)
        if lut is not None:
            self.lut = torch.from_numpy(lut).to(device)
        self.c = _lib.Tables(self.cdf.data_ptr(), self.sizes.data_ptr(), self.offsets.data_ptr(),
                             cdf.shape[0], cdf.shape[1],
                             self.lut.data_ptr() if self.lut is not None else None)

    def get_cdf_info(self):
        return self.cdf_np, self.sizes_np, self.offsets_np


class EntropyCoder:
    """Facade over the GPU coder; call sequence mirrors entropy_models.py:9-55
    (reset / encode* / flush / get_encoded_stream, set_stream / decode*).

    mode "compat": the reference's container, byte-identical, stream_part in 1..16.
    mode "interleaved": stream_part warps x 32 interleaved lanes (own container).
    mode "hem": DCVC-HEM's framing (the bare words of one compat stream, no flag byte;
                DCVC-HEM/src/cpp/rans/rans_interface.cpp:155-172), stream_part must be 1.
    """

    def __init__(self, ec_thread=False, stream_part=1, mode="compat", max_symbols=1 << 22,
                 device="cuda"):
        del ec_thread  # host threading does not exist here: the coder runs on the GPU
        self.mode = mode
        self.parts = int(stream_part)
        self.device = torch.device(device)
        self._h = None
        self._stream_bytes = b""
        self._pending = False
        self._side = None
        self._keep = []
        self.lazy_check = False      # pipeline.py: check a frame's decode errors one frame later
        self._create(int(max_symbols))

    def _create(self, max_symbols):
        """(Re)creates the coder context for frames of up to `max_symbols` symbols.  All device
        memory is a torch tensor owned here and lent to the library as its workspace."""
        L = _lib.lib()
        mode = {"compat": 0, "interleaved": 1, "hem": 2}[self.mode]
        nbytes = L.dcvc_rans_workspace_bytes(mode, self.parts, max_symbols)
        if nbytes == 0:
            raise _lib.DcvcError(
                f"EntropyCoder: unsupported configuration (mode {self.mode}, stream_part "
                f"{self.parts}: compat takes 1..16 parts, interleaved 1..256 warps)")
        with torch.cuda.device(self.device):
            ws = torch.empty(nbytes + 256, dtype=torch.uint8, device=self.device)
            off = (-ws.data_ptr()) % 256
            h = ctypes.c_void_p()
            _lib.check(L.dcvc_rans_create(ctypes.byref(h), mode, self.parts, max_symbols,
                                          ws.data_ptr() + off, nbytes,
                                          torch.cuda.current_stream().cuda_stream))
        self._destroy()
        self._h, self._ws = h, ws
        self.max_symbols = max_symbols
        self._cap = max_symbols * 4 + 1024 + 4 * self.parts
        self._out = torch.empty(self._cap, dtype=torch.uint8).pin_memory()

    def reserve(self, symbols):
        """Makes room for frames of `symbols` symbols (the reference coder has no limit): the
        context is rebuilt, larger, when a frame size beyond the current capacity shows up."""
        if symbols > self.max_symbols:
            if self._pending:
                self.get_encoded_stream()
            self.check()
            torch.cuda.synchronize(self.device)
            self._create(int(symbols))

    def _destroy(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                _lib.lib().dcvc_rans_destroy(h)
            except Exception:      # interpreter shutdown: modules may already be torn down
                pass
            self._h = None

    def __del__(self):
        self._destroy()

    def reset(self):
        if self._pending:        # an unread flush: finish it so the coder can start a new one
            self.get_encoded_stream()
        self._keep = []
        _lib.check(_lib.lib().dcvc_rans_enc_reset(self._h))

    def encode(self, symbols, indexes, n, plane, tables):
        """symbols int16 device tensor; indexes uint8 device tensor or None (index = pos//plane)."""
        self._keep.append((symbols, indexes))   # borrowed until flush
        _lib.check(_lib.lib().dcvc_rans_enc_push(
            self._h, symbols.data_ptr(), indexes.data_ptr() if indexes is not None else None,
            n, plane, ctypes.byref(tables.c)))

    def flush_enqueue(self):
        """flush() as pure stream work on the CURRENT stream, for CUDA-graph capture
        (dcvc_b200.graphs): no event, no side stream; mark_flushed() completes it per replay."""
        _lib.check(_lib.lib().dcvc_rans_enc_flush_enqueue(
            self._h, self._cap, torch.cuda.current_stream().cuda_stream))

    def mark_flushed_captured(self):
        """inside a stream capture: the coder's event as an event-record node of the graph"""
        _lib.check(_lib.lib().dcvc_rans_enc_flush_mark_captured(
            self._h, torch.cuda.current_stream().cuda_stream))

    def set_pending(self):
        """per graph replay: the recorded flush is in flight, get_encoded_stream() may wait on it"""
        _lib.check(_lib.lib().dcvc_rans_enc_flush_set_pending(self._h))
        self._pending = True

    def mark_flushed(self):
        _lib.check(_lib.lib().dcvc_rans_enc_flush_mark(
            self._h, torch.cuda.current_stream().cuda_stream))
        self._pending = True

    def flush(self):
        """Enqueues the coder kernels (RansEncoder.flush, py_rans.cpp:98-135); the host wait is
        in get_encoded_stream, so whatever the caller enqueues in between runs behind the coder
        on the GPU while the host already reads the bytes."""
        # The coder kernels occupy a handful of warps for ~1.2 ms: they run on the coder's own
        # stream, forked from the caller's here, so the caller's stream goes straight on with the
        # reconstruction.  The borrowed symbol tensors stay referenced until the host has waited
        # for the coder (get_encoded_stream), since the allocator only orders reuse on the
        # caller's stream.
        cur = torch.cuda.current_stream()
        from . import ops as _ops
        side = SIDE_STREAM and _ops.PROFILE is None     # (the per-kernel table times `cur`)
        if side:
            if self._side is None:
                self._side = torch.cuda.Stream(self.device)
            self._side.wait_stream(cur)
            run = self._side
        else:
            run = cur
        with _Prof("rans_enc_flush"):
            _lib.check(_lib.lib().dcvc_rans_enc_flush_begin(self._h, self._cap, run.cuda_stream))
        if not side:
            self._keep = []
        self._pending = True

    def get_encoded_stream(self):
        if self._pending:
            n = ctypes.c_size_t(0)
            self._pending = False
            _lib.check(_lib.lib().dcvc_rans_enc_flush_end(
                self._h, self._out.data_ptr(), self._cap, ctypes.byref(n)))
            self._stream_bytes = bytes(self._out[:n.value].numpy().tobytes())
            self._keep = []
        return self._stream_bytes

    def set_stream(self, stream):
        buf = np.frombuffer(stream, dtype=np.uint8)
        _lib.check(_lib.lib().dcvc_rans_dec_set_stream(
            self._h, buf.ctypes.data, buf.size, torch.cuda.current_stream().cuda_stream))

    def decode(self, indexes, n, plane, tables, out):
        with _Prof("rans_dec_pull"):
            _lib.check(_lib.lib().dcvc_rans_dec_pull(
                self._h, indexes.data_ptr() if indexes is not None else None, n, plane,
                ctypes.byref(tables.c), out.data_ptr(), torch.cuda.current_stream().cuda_stream))
        return out

    def finish_decode(self):
        """After the frame's last decode(): enqueue the copy of the decode kernels' error word.
        The matching check() (or the next set_stream) raises DcvcError for a truncated or
        corrupt frame; with lazy_check the host does not wait here."""
        _lib.check(_lib.lib().dcvc_rans_dec_finish(self._h,
                                                   torch.cuda.current_stream().cuda_stream))

    def check(self):
        _lib.check(_lib.lib().dcvc_rans_dec_check(self._h))

    def status(self):
        _lib.check(_lib.lib().dcvc_rans_status(self._h, torch.cuda.current_stream().cuda_stream))
