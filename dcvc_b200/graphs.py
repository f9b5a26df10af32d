"""CUDA-graph replay of the P-frame encode and decode.

A 1080p P frame is ~785 kernel launches; issued one by one through ctypes they cost ~20 ms of
host time per frame, which bounds any schedule that cannot hide it (one stream, encode then
decode).  The shapes are static per (resolution, frame type), so the launches of `DMC.compress`
and `DMC.decompress` are recorded once into CUDA graphs and replayed:

    compress   = ONE graph: motion, contexts, latents; then two branches -- the entropy-coder
                 kernels + meta words -> host + an event-record node (cudaEventRecordExternal,
                 dcvc_rans_enc_flush_mark_captured), and the encoder's own reconstruction (the next
                 frame's dpb).  The host waits on the coder's event only, as in the eager path.
    decompress = host: parse + upload the stream (dcvc_rans_dec_set_stream)
                 | graph (everything else) | host: error-word copy (dcvc_rans_dec_finish)

The only host-visible points are the ones the rANS API already has.

Memory.  A frame's results (the five dpb tensors) are written by their producing kernels straight
into static buffers OUTSIDE the memory pool the graphs share (`tensor.PLACEMENT`), one set per
PHASE = frame_idx % 4, so no graph's intermediates can overwrite a dpb the caller still holds and
nothing is copied at the end of a graph.  A returned dpb stays valid until the same phase comes
round again (4 frames later) -- this differs from the reference's "new tensors each frame"
contract, which is why the graphs are opt-in (`GraphedDMC(model)`).  Inputs: a graph is recorded
reading the dpb it was given IN PLACE when those tensors are themselves static results of a
GraphedDMC (the steady state: phase p always continues from phase p - 1), so nothing is copied at
the start either; a call whose dpb lives elsewhere (the second P frame of a GOP continues from an
eager frame) runs eagerly, into the same static results.  Only the frame itself (25 MB) is copied
into a static input.  Results are bit-identical to the eager path (tests/test_gpu_graphs.py).

A key (shape, which dpb entries exist, phase, rate point) is run eagerly the first time it is
seen -- that fills every lazily built cache (packed weights, q vectors, kernel attributes) --
recorded the second time, replayed from then on.
"""
import torch

from . import _lib, tensor
from .tensor import Act

DPB_KEYS = ("ref_frame", "ref_feature", "ref_mv_feature", "ref_y", "ref_mv_y")


class _Static:
    """Static tensors with the exact memory layout of the ones given (dpb entries are NCHW views of
    NHWC buffers: the whole buffer is mirrored).  views[name]: the [1,C,H,W] tensor, acts[name]:
    the same memory as an Act."""

    def __init__(self, tensors):
        self.bufs, self.views, self.acts = {}, {}, {}
        for name, t in tensors.items():
            if t is None:
                self.views[name] = None
                continue
            a = _as_act(t)
            if a is not None:                       # channels-last view of an NHWC buffer
                buf = torch.empty_like(a.buf)
                self.bufs[name] = (buf, True)
                self.acts[name] = Act(buf, a.c0, a.c)
                self.views[name] = self.acts[name].nchw()
            else:
                buf = torch.empty_like(t)
                self.bufs[name] = (buf, False)
                self.views[name] = buf
            _STATIC_PTRS.add(buf.data_ptr())

    def load(self, tensors):
        for name, (buf, nhwc) in self.bufs.items():
            t = tensors[name]
            if nhwc:
                src = _as_act(t)
                if src is None or src.buf.shape != buf.shape:
                    self.views[name].copy_(t)       # same values, different layout
                elif src.buf.data_ptr() != buf.data_ptr():
                    buf.copy_(src.buf, non_blocking=True)
            elif t.data_ptr() != buf.data_ptr():
                buf.copy_(t, non_blocking=True)


# data pointers of every static result buffer of every GraphedDMC of this process: a dpb made of
# these can be read in place by a recorded graph
_STATIC_PTRS = set()


def _is_static(dpb):
    for n in DPB_KEYS:
        t = dpb[n]
        if t is None:
            continue
        a = _as_act(t)
        if (a.buf.data_ptr() if a is not None else t.data_ptr()) not in _STATIC_PTRS:
            return False
    return True


def _sig(dpb):
    """identity of a dpb's memory: what a graph recorded on it depends on"""
    return tuple(None if dpb[n] is None else (dpb[n].data_ptr(), tuple(dpb[n].shape),
                                              tuple(dpb[n].stride())) for n in DPB_KEYS)


def _as_act(t):
    """the NHWC buffer behind a [1,C,H,W] channels-last view, or None (see Act.wrap)"""
    if t.dim() != 4 or t.shape[0] != 1 or t.dtype != torch.float32:
        return None
    _, c, h, w = t.shape
    cs = t.stride(3)
    if t.stride(1) != 1 or cs < c or t.stride(2) != w * cs:
        return None
    c0 = t.storage_offset() % cs
    base = t.storage_offset() - c0
    if t.untyped_storage().nbytes() // 4 < base + h * w * cs:
        return None
    return Act(t.as_strided((h, w, cs), (w * cs, cs, 1), base), c0, c)


class GraphedDMC:
    """`compress` / `decompress` of a `dcvc_b200.DMC` replayed from CUDA graphs."""

    def __init__(self, model, first_p=False):
        """first_p: also record the first P frame of a GOP (dpb without features: once per intra
        period, so by default it simply runs eagerly)"""
        self.m = model
        model.__dict__["_graph_runner"] = self       # DMC.encode_decode routes through the graphs
        self.first_p = first_p
        self.pool = torch.cuda.graph_pool_handle()
        self.cap_stream = torch.cuda.Stream(model.device)
        self.side_stream = torch.cuda.Stream(model.device)
        self.enc, self.dec = {}, {}
        self.replays = self.eager = 0

    @property
    def entropy_coder(self):
        return self.m.entropy_coder

    def _key(self, kind, x_like, dpb, q_in_ckpt, q_index, frame_idx, extra=()):
        phase = -1 if dpb["ref_feature"] is None else frame_idx % 4
        pattern = tuple(dpb[n] is None for n in DPB_KEYS)
        return (kind, tuple(x_like), pattern, phase, bool(q_in_ckpt), int(q_index)) + tuple(extra)

    def _record(self, fn):
        """runs fn() under stream capture on the private stream; fn gets (begin, end) callables"""
        cur = torch.cuda.current_stream(self.m.device)
        torch.cuda.synchronize(self.m.device)
        self.cap_stream.wait_stream(cur)
        graphs = []

        def begin():
            g = torch.cuda.CUDAGraph()
            g.capture_begin(pool=self.pool)
            graphs.append(g)

        def end():
            graphs[-1].capture_end()

        l0 = _lib.launch_count()
        with torch.cuda.stream(self.cap_stream):
            out = fn(begin, end)
        cur.wait_stream(self.cap_stream)
        self._recorded = _lib.launch_count() - l0    # kernels one replay stands for
        _lib.lib().dcvc_launch_count_add(-self._recorded)            # recorded, not run
        return graphs, out

    def _placed(self, outs, fn):
        """fn() with the frame's results placed in `outs` (None: nowhere special)"""
        prev = tensor.PLACEMENT
        tensor.PLACEMENT = outs.acts if outs is not None else None
        try:
            return fn()
        finally:
            tensor.PLACEMENT = prev

    @staticmethod
    def _settle(outs, dpb):
        """the frame's results as static views: already there when placed, copied otherwise"""
        outs.load(dpb)
        return dict(outs.views)

    # ------------------------------------------------------------------ encoder
    def compress(self, x, dpb, q_in_ckpt, q_index, frame_idx, defer=False, reconstruct=True):
        m, ec = self.m, self.m.entropy_coder
        # the first P frame of a GOP, and any frame whose per-stage taps are being collected (a
        # replay has no host code between its kernels), run as plain launches
        if (dpb["ref_feature"] is None and not self.first_p) or m.taps is not None:
            return m.compress(x, dpb, q_in_ckpt, q_index, frame_idx, defer=defer,
                              reconstruct=reconstruct)
        key = self._key("enc", tuple(x.shape) + tuple(x.stride()), dpb, q_in_ckpt, q_index,
                        frame_idx, (bool(reconstruct),))
        ent = self.enc.get(key)

        def eager(outs):
            res = self._placed(outs, lambda: m.compress(x, dpb, q_in_ckpt, q_index, frame_idx,
                                                        defer=defer, reconstruct=reconstruct))
            self.eager += 1
            if reconstruct:
                res["dpb"] = self._settle(outs, res["dpb"])
            return res

        if ent is None:                              # first sight: eager (fills the lazy caches)
            res = m.compress(x, dpb, q_in_ckpt, q_index, frame_idx, defer=defer,
                             reconstruct=reconstruct)
            outs = _Static(res["dpb"]) if reconstruct else None
            self.enc[key] = ["warm", outs, 0]
            if reconstruct:
                res["dpb"] = self._settle(outs, res["dpb"])
            return res
        if ent[0] == "warm":
            outs = ent[1]
            in_place = _is_static(dpb)
            if not in_place and ent[2] < 2:          # wait for a steady-state dpb before recording
                ent[2] += 1
                return eager(outs)
            xs = _Static({"x": x})
            xs.load({"x": x})
            if in_place:
                sdpb = {n: dpb[n] for n in DPB_KEYS}
                static = None
            else:
                static = _Static(dict(dpb))
                static.load(dpb)
                sdpb = {n: static.views[n] for n in DPB_KEYS}

            def body(begin, end):
                def coder(mv_zs, zs, mv_coded, coded):
                    # the coder's handful of warps runs on a forked branch of the graph, beside the
                    # reconstruction; its event is an event-record node the host waits on
                    ec.reset()
                    m.encode_z(mv_zs[1], mv_zs[0], "bit_estimator_z_mv")
                    m.encode_z(zs[1], zs[0], "bit_estimator_z")
                    m.encode_y(mv_coded)
                    m.encode_y(coded)
                    self.side_stream.wait_stream(self.cap_stream)
                    with torch.cuda.stream(self.side_stream):
                        ec.flush_enqueue()
                        ec.mark_flushed_captured()
                begin()
                new_dpb = m._encode_side(xs.views["x"], sdpb, q_in_ckpt, q_index, frame_idx,
                                         True, coder=coder, reconstruct=reconstruct)[0]
                if reconstruct:
                    outs.load(new_dpb)               # no-op for placed results, else recorded copies
                self.cap_stream.wait_stream(self.side_stream)        # join the coder branch
                end()
                return dict(outs.views) if reconstruct else None

            zs_ = m._frame_symbols(x.shape[2], x.shape[3])
            ec.reserve(zs_)
            if ec._pending:                          # an unread flush of the previous frame
                ec.get_encoded_stream()
            graphs, new_dpb = self._placed(outs, lambda: self._record(body))
            ent = ["ready", {"x": xs, "static": static, "sig": _sig(dpb) if in_place else None,
                             "held": sdpb, "graphs": graphs, "dpb": new_dpb, "outs": outs,
                             "launches": self._recorded}]
            self.enc[key] = ent
        ent = ent[1]
        if ent["sig"] is not None and _sig(dpb) != ent["sig"]:
            return eager(ent["outs"])                # dpb from elsewhere (GOP start): this once eagerly
        if ec._pending:
            ec.get_encoded_stream()
        ent["x"].load({"x": x})
        if ent["static"] is not None:
            ent["static"].load(dpb)
        ent["graphs"][0].replay()
        ec.set_pending()
        _lib.lib().dcvc_launch_count_add(ent["launches"])
        self.replays += 1
        return {"dpb": ent["dpb"],
                "bit_stream": ec.get_encoded_stream if defer else ec.get_encoded_stream()}

    # ------------------------------------------------------------------ decoder
    def decompress(self, dpb, string, height, width, q_in_ckpt, q_index, frame_idx):
        m, ec = self.m, self.m.entropy_coder
        if (dpb["ref_feature"] is None and not self.first_p) or m.taps is not None:
            return m.decompress(dpb, string, height, width, q_in_ckpt, q_index, frame_idx)
        key = self._key("dec", (height, width), dpb, q_in_ckpt, q_index, frame_idx)
        ent = self.dec.get(key)

        def eager(outs):
            res = self._placed(outs, lambda: m.decompress(dpb, string, height, width, q_in_ckpt,
                                                          q_index, frame_idx))
            self.eager += 1
            return {"dpb": self._settle(outs, res["dpb"])}

        if ent is None:
            res = m.decompress(dpb, string, height, width, q_in_ckpt, q_index, frame_idx)
            outs = _Static(res["dpb"])
            self.dec[key] = ["warm", outs, 0]
            return {"dpb": self._settle(outs, res["dpb"])}
        if ent[0] == "warm":
            outs = ent[1]
            in_place = _is_static(dpb)
            if not in_place and ent[2] < 2:
                ent[2] += 1
                return eager(outs)
            if in_place:
                sdpb, static = {n: dpb[n] for n in DPB_KEYS}, None
            else:
                static = _Static(dict(dpb))
                static.load(dpb)
                sdpb = {n: static.views[n] for n in DPB_KEYS}

            def body(begin, end):
                begin()
                out = m._decompress_body(sdpb, height, width, q_in_ckpt, q_index, frame_idx, False)
                outs.load(out["dpb"])
                end()
                return {"dpb": dict(outs.views)}

            m._decode_begin(string, height, width)   # sizes the coder before the recording
            graphs, out = self._placed(outs, lambda: self._record(body))
            ent = ["ready", {"static": static, "sig": _sig(dpb) if in_place else None,
                             "held": sdpb, "graphs": graphs, "out": out, "outs": outs,
                             "launches": self._recorded}]
            self.dec[key] = ent
        ent = ent[1]
        if ent["sig"] is not None and _sig(dpb) != ent["sig"]:
            return eager(ent["outs"])
        m._decode_begin(string, height, width)
        if ent["static"] is not None:
            ent["static"].load(dpb)
        ent["graphs"][0].replay()
        _lib.lib().dcvc_launch_count_add(ent["launches"])
        ec.finish_decode()
        if not ec.lazy_check:
            ec.check()
        self.replays += 1
        return ent["out"]

    # everything else (encode_decode, forward_one_frame, parameters ...) is the model's own
    def __getattr__(self, name):
        return getattr(self.m, name)
