"""Encoder and decoder as a two-stage pipeline on two CUDA streams.

The reference's frame loop (DCVC-DC/test_video.py:138-165) codes a frame and decodes it before it
touches the next one.  Nothing forces that order: the encoder's own reconstruction (the `dpb`
`compress` returns) is bit-identical to what `decompress` produces from the stream -- both run the
same deterministic kernels (tests/test_gpu_codec.py) -- so the encoder can move on to frame t+1
while the decoder, a second model instance with its own coder context, weights copy and `dpb`,
works on the stream of frame t.  The two frame chains only meet in host memory (the byte stream),
exactly as an encoder and a decoder on two machines would.  On one GPU this fills the holes the
sequential order leaves: the rANS kernels (a handful of warps), the 1/16-resolution prior networks
(64 tiles on 148 SMs) and every kernel's tail run next to the other stream's full-width
convolutions.

`EncDecPipeline.push(x)` enqueues the decode of the previous frame on the decoder stream, then the
encode of `x` on the encoder stream, and returns when the encoder's stream bytes are on the host.
"""
import os

import torch

# DCVC_PIPE_ENCODE_FIRST=1: enqueue encode(t) before decode(t-1) inside push()
ENCODE_FIRST = int(os.environ.get("DCVC_PIPE_ENCODE_FIRST", "0"))


class EncDecPipeline:
    def __init__(self, i_enc, p_enc, i_dec, p_dec, intra_period, q_in_ckpt, q_index, device):
        self.i_enc, self.p_enc, self.i_dec, self.p_dec = i_enc, p_enc, i_dec, p_dec
        self.intra_period = int(intra_period)
        self.q_in_ckpt, self.q_index = q_in_ckpt, q_index
        self.device = torch.device(device)
        self.enc_stream = torch.cuda.Stream(self.device)
        self.dec_stream = torch.cuda.Stream(self.device)
        self.enc_dpb = None
        self.dec_dpb = None
        self.idx = 0              # frames pushed
        self.bits = 0
        self._pending = None      # (frame index, bytes, padded height, padded width)
        # decode errors (truncated / corrupt frame) are checked one frame late, at the next
        # set_stream, so that the host never waits for the decoder stream inside push()
        self._dec_models = (i_dec, p_dec)
        self._lazy(True)

    def _lazy(self, on):
        for m in self._dec_models:
            ec = getattr(m, "entropy_coder", None)
            if ec is not None:
                ec.lazy_check = on

    # ------------------------------------------------------------------ decoder stage
    def _decode(self, item):
        idx, stream, ph, pw = item
        with torch.cuda.stream(self.dec_stream):
            if idx % self.intra_period == 0:
                dec = self.i_dec.decompress(stream, ph, pw, self.q_in_ckpt, self.q_index)
                self.dec_dpb = {"ref_frame": dec["x_hat"], "ref_feature": None,
                                "ref_mv_feature": None, "ref_y": None, "ref_mv_y": None}
            else:
                self.dec_dpb = self.p_dec.decompress(self.dec_dpb, stream, ph, pw, self.q_in_ckpt,
                                                     self.q_index, idx % 4)["dpb"]
        return self.dec_dpb["ref_frame"]

    # ------------------------------------------------------------------ encoder stage
    def push(self, x_padded):
        """x_padded: [1,3,H,W] on the device, produced on (or synchronised with) `enc_stream`.
        Returns (stream bytes of this frame, decoded frame of the PREVIOUS push or None); the
        decoded frame is still being produced on `dec_stream` when this returns."""
        ph, pw = x_padded.shape[2], x_padded.shape[3]
        decoded, done = None, False
        if ENCODE_FIRST == 0 and self._pending is not None:
            decoded, done = self._decode(self._pending), True
        with torch.cuda.stream(self.enc_stream):
            if self.idx % self.intra_period == 0:
                enc = self.i_enc.compress(x_padded, self.q_in_ckpt, self.q_index)
                self.enc_dpb = {"ref_frame": enc["x_hat"], "ref_feature": None,
                                "ref_mv_feature": None, "ref_y": None, "ref_mv_y": None}
                stream = enc["bit_stream"]
            else:
                enc = self.p_enc.compress(x_padded, self.enc_dpb, self.q_in_ckpt, self.q_index,
                                          self.idx % 4, defer=True)
                self.enc_dpb = enc["dpb"]
                stream = None
        if not done and self._pending is not None:
            # the previous frame's decode is enqueued while the GPU works on this frame's encode;
            # only then does the host wait for the coder's event
            decoded = self._decode(self._pending)
        if stream is None:
            stream = enc["bit_stream"]()
        self._pending = (self.idx, stream, ph, pw)
        self.idx += 1
        self.bits += len(stream) * 8
        return stream, decoded

    def flush(self):
        """Decodes the last pushed frame; returns its decoded frame (or None)."""
        decoded = self._decode(self._pending) if self._pending is not None else None
        self._pending = None
        for m in self._dec_models:      # the last frames' deferred error checks
            m.entropy_coder.check()
        return decoded

    def join(self, stream=None):
        """Makes `stream` (default: the current stream) wait for both stages."""
        s = stream if stream is not None else torch.cuda.current_stream(self.device)
        s.wait_stream(self.enc_stream)
        s.wait_stream(self.dec_stream)
