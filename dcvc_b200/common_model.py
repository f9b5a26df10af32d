"""`CompressionModel`: parameters, q handling, tables and the quad-part entropy model shared by
`IntraNoAR` and `DMC` -- the host-side mirror of the reference's
DCVC-DC/src/models/common_model.py:14-321.

The class is an nn.Module only for its parameter bookkeeping (state_dict keys identical to the
reference so the released checkpoints load unchanged; `.to()`, `.eval()`, `.parameters()` as the
harness uses them).  No torch op runs on the hot path: every step below is a kernel launch
through the C ABI.
"""
import math

import numpy as np
import torch
from torch import nn

from . import entropy_models as em
from . import layers, ops, tensor
from .tensor import Act


def _default_init(spec, anchor_num):
    """Random init of the reference's flavour (nn.Conv2d default scale; Bitparm N(0, 0.01);
    q parameters = 1, common_model/image_model/video_model constructors)."""
    sd = {}
    for name, shape in spec.items():
        if name.endswith(".weight"):
            fan_in = shape[1] * (shape[2] * shape[3] if len(shape) == 4 else 1)
            sd[name] = (torch.rand(shape) * 2 - 1) * (1.0 / math.sqrt(fan_in))
        elif name.endswith(".bias"):
            w = spec[name[:-5] + ".weight"]
            bound = 1.0 / math.sqrt(w[1] * w[2] * w[3])
            sd[name] = (torch.rand(shape) * 2 - 1) * bound
        elif "bit_estimator" in name:
            sd[name] = torch.randn(shape) * 0.01
        else:
            sd[name] = torch.ones(shape)
    return sd


def get_padding_size(height, width, p=64):
    """utils/stream_helper.py:22-31 (kept here too: the latent sizes depend on it)."""
    new_h = (height + p - 1) // p * p
    new_w = (width + p - 1) // p * p
    return 0, new_w - width, 0, new_h - height


def get_downsampled_shape(height, width, p):
    """utils/stream_helper.py:34-37"""
    new_h = (height + p - 1) // p * p
    new_w = (width + p - 1) // p * p
    return int(new_h / p + 0.5), int(new_w / p + 0.5)


class CompressionModel(nn.Module):
    def __init__(self, spec, y_distribution, z_names, ec_thread=False, stream_part=1,
                 anchor_num=4, coder_mode="compat"):
        super().__init__()
        self._spec = spec
        self.y_distribution = y_distribution
        self.z_names = tuple(z_names)            # bit-estimator prefixes
        self.ec_thread = ec_thread
        self.stream_part = stream_part
        self.anchor_num = anchor_num
        self.coder_mode = coder_mode
        for name, value in _default_init(spec, anchor_num).items():
            self._register(name, nn.Parameter(value))
        self.entropy_coder = None
        self._net = None
        self._fine = {}          # q_scale name -> float64 numpy table (load_state_dict)
        self._tables = {}
        self._edges = None
        self._scratch = {}
        self.taps = None         # tests: dict receiving named intermediates

    # ------------------------------------------------------------------ parameter tree
    def _register(self, name, param):
        mod = self
        parts = name.split(".")
        for p in parts[:-1]:
            child = mod._modules.get(p)
            if child is None:
                child = nn.Module()
                mod.add_module(p, child)
            mod = child
        mod.register_parameter(parts[-1], param)

    def _apply(self, fn, recurse=True):
        self._net = None
        self._tables = {}
        self._edges = None
        self._scratch = {}
        return super()._apply(fn, recurse)

    def load_state_dict(self, state_dict, strict=True):
        r = super().load_state_dict(state_dict, strict)
        self._net = None
        self._tables = {}
        self._scratch = {}
        self._build_fine_tables()
        return r

    def _q_scale_names(self):
        return [n for n in self._spec if "q_scale" in n]

    def _build_fine_tables(self):
        """64-point rate tables: exp(linspace(log a0, log a3, 64)) on fp32 anchors evaluated in
        float64 numpy (video_model.py:325-341, image_model.py:158-167)."""
        sd = self.state_dict()
        for n in self._q_scale_names():
            a = sd[n].reshape(-1).detach().cpu().numpy().astype(np.float32)
            self._fine[n] = np.exp(np.linspace(np.log(a[0]), np.log(a[3]), 64))

    @property
    def device(self):
        return next(self.parameters()).device

    @property
    def net(self):
        if self._net is None:
            self._net = layers.Net(self.state_dict(), self.device)
        return self._net

    # ------------------------------------------------------------------ q handling
    def get_curr_q(self, q_scale_name, q_basic_name, q_in_ckpt, q_index):
        """get_curr_q / get_q_for_inference (common_model.py:35-37): q_basic * q_scale[q_index];
        the fine table entry is a float64 scalar, the product stays fp32 (torch promotion)."""
        key = (q_scale_name, bool(q_in_ckpt), int(q_index))
        r = self._scratch.get(key)
        if r is None:
            basic = self.net.sd[q_basic_name]
            if q_in_ckpt:
                scale = self.net.sd[q_scale_name].reshape(-1)[q_index]
                r = (basic * scale).reshape(-1).contiguous()
            else:
                if q_scale_name not in self._fine:
                    self._build_fine_tables()
                r = (basic * self._fine[q_scale_name][q_index]).reshape(-1).contiguous()
            self._scratch[key] = r
        return r

    # ------------------------------------------------------------------ tables / coder
    def _bitparm(self, prefix, device=None):
        sd = self.state_dict()
        out = {}
        for k in ("f1", "f2", "f3", "f4"):
            names = ("h", "b", "a") if k != "f4" else ("h", "b")
            out[k] = tuple(sd[f"{prefix}.{k}.{n}"].detach().float().cpu() for n in names)
        return out

    def update(self, force=False):
        """CompressionModel.update (common_model.py:63-68): coder + CDF tables, once per model."""
        del force
        dev = self.device
        if dev.type != "cuda":
            raise RuntimeError("dcvc_b200 codes on the GPU only: move the model to a CUDA device "
                               "before update() (there is no CPU fallback)")
        self.entropy_coder = em.EntropyCoder(self.ec_thread, self.stream_part, self.coder_mode,
                                             device=dev)
        self._tables = {"y": em.DeviceTables(*em.gaussian_tables(self.y_distribution), dev)}
        for zn in self.z_names:
            self._tables[zn] = em.DeviceTables(*em.factorized_tables(self._bitparm(zn)), dev)
        self._ensure_edges()

    def _ensure_edges(self):
        if self._edges is None:
            self._edges = em.index_edges(self.y_distribution).to(self.device)
        return self._edges

    def _fparams(self, prefix):
        """[11][C] raw factorised-prior parameters for the bit-estimate kernel."""
        key = ("fparams", prefix)
        r = self._scratch.get(key)
        if r is None:
            sd = self.net.sd
            rows = []
            for k in ("f1", "f2", "f3"):
                rows += [sd[f"{prefix}.{k}.h"], sd[f"{prefix}.{k}.b"], sd[f"{prefix}.{k}.a"]]
            rows += [sd[f"{prefix}.f4.h"], sd[f"{prefix}.f4.b"]]
            r = torch.stack([t.reshape(-1) for t in rows], 0).contiguous()
            self._scratch[key] = r
        return r

    # ------------------------------------------------------------------ latent geometry
    @staticmethod
    def pad_for_y(y):
        """common_model.py:70-78: replicate-pad bottom/right to a multiple of 4."""
        ph = (y.h + 3) // 4 * 4
        pw = (y.w + 3) // 4 * 4
        if ph == y.h and pw == y.w:
            return y
        return ops.copy(y, h=ph, w=pw)

    # ------------------------------------------------------------------ hyper latent
    def quant_z(self, z, prefix, write, bits=None):
        """round(z) (+ symbols for the coder, + factorised bit estimate)."""
        n = z.h * z.w * z.c
        z_hat = Act.new(z.h, z.w, z.c, z.device)
        symbols = torch.empty(n, dtype=torch.int16, device=z.device) if write else None
        ws = None
        if bits is not None:
            ws = torch.empty((n + 255) // 256, dtype=torch.float32, device=z.device)
        ops.z_quant(z, z_hat, symbols, self._fparams(prefix) if bits is not None else None,
                    bits, ws)
        return z_hat, symbols

    def encode_z(self, symbols, z_hat, prefix):
        self.entropy_coder.encode(symbols, None, symbols.numel(), z_hat.h * z_hat.w,
                                  self._tables[prefix])

    def decode_z(self, channel, size, prefix):
        h, w = size
        n = channel * h * w
        symbols = torch.empty(n, dtype=torch.int16, device=self.device)
        self.entropy_coder.decode(None, n, h * w, self._tables[prefix], symbols)
        z_hat = Act.new(h, w, channel, self.device)
        ops.z_dequant(symbols, z_hat)
        return z_hat

    # ------------------------------------------------------------------ quad-part prior
    def new_prior_buffer(self, h, w, c):
        """[y_hat_so_far (C, zero) | common_params (3C)]: the spatial-prior conv input
        (torch.cat((y_hat_so_far, common_params)), common_model.py:183) as one buffer."""
        catb = Act.new(h, w, 4 * c, self.device)
        ops.fill(catb.slice(0, c), 0.0)
        return catb

    def forward_four_part_prior(self, y, catb, prior_fn, write=False, bits=None, tag=None):
        """forward_four_part_prior / compress_four_part_prior (common_model.py:142-259).
        Returns (y_hat Act, [(symbols, indexes)] * 4 when write)."""
        c = y.c
        n = (c // 4) * y.h * y.w
        dev = y.device
        y_hat_so_far = catb.slice(0, c)
        params = catb.slice(c, 3 * c)
        y_out = tensor.placed(tag, y.h, y.w, c) or Act.new(y.h, y.w, c, dev)
        edges = self._ensure_edges()
        dist = em.DIST_ID[self.y_distribution]
        ws = torch.empty((n + 255) // 256, dtype=torch.float32, device=dev) if bits is not None \
            else None
        coded = []
        for step in range(4):
            prior = prior_fn(step, catb) if step > 0 else None
            sym = idx = None
            if write:
                sym = torch.empty(n, dtype=torch.int16, device=dev)
                idx = torch.empty(n, dtype=torch.uint8, device=dev)
                coded.append((sym, idx))
            ops.quad_step_enc(y, params, prior, y_hat_so_far, y_out, edges, step, dist, sym, idx,
                              bits, ws)
        return y_out, coded

    def encode_y(self, coded):
        for sym, idx in coded:
            self.entropy_coder.encode(sym, idx, sym.numel(), 1, self._tables["y"])

    def decompress_four_part_prior(self, catb, c, prior_fn, tag=None):
        """decompress_four_part_prior (common_model.py:261-321): per step, indexes -> rANS pull ->
        scatter, all on the device."""
        h, w = catb.h, catb.w
        n = (c // 4) * h * w
        dev = catb.device
        y_hat_so_far = catb.slice(0, c)
        params = catb.slice(c, 3 * c)
        y_out = tensor.placed(tag, h, w, c) or Act.new(h, w, c, dev)
        edges = self._ensure_edges()
        idx = torch.empty(n, dtype=torch.uint8, device=dev)
        sym = torch.empty(n, dtype=torch.int16, device=dev)
        for step in range(4):
            prior = prior_fn(step, catb) if step > 0 else None
            ops.quad_step_indexes(params, prior, y_hat_so_far, y_out, edges, step, idx)
            self.entropy_coder.decode(idx, n, 1, self._tables["y"], sym)
            ops.quad_step_dec(params, prior, y_hat_so_far, y_out, edges, step, sym)
        return y_out
