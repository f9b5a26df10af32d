"""`IntraNoAR` -- the DCVC-DC intra model behind the reference's module API
(DCVC-DC/src/models/image_model.py:61-252), executed as fused sm_100a kernel launches.

Same constructor signature, state_dict keys, method signatures and return dicts as the reference;
`x` and `x_hat` are [1,3,H,W] fp32 CUDA tensors (x_hat is a strided view of an NHWC buffer).
"""
import torch

from . import arch, layers, ops
from .common_model import CompressionModel, get_downsampled_shape
from .stream_helper import encode_i, decode_i, filesize
from .tensor import Act


class IntraNoAR(CompressionModel):
    def __init__(self, N=256, anchor_num=4, ec_thread=False, stream_part=1, inplace=False,
                 coder_mode="compat"):
        del inplace  # activations are fused into the convolutions
        super().__init__(arch.intra_spec(N, anchor_num), "gaussian", ("bit_estimator_z",),
                         ec_thread, stream_part, anchor_num, coder_mode)
        self.N = N

    # ------------------------------------------------------------------ q
    @staticmethod
    def get_q_scales_from_ckpt(ckpt_path):
        """image_model.py:151-156"""
        from .stream_helper import get_state_dict
        ckpt = get_state_dict(ckpt_path)
        return ckpt["q_scale_enc"].reshape(-1), ckpt["q_scale_dec"].reshape(-1)

    def get_q_for_inference(self, q_in_ckpt, q_index):
        """image_model.py:107-112"""
        return (self.get_curr_q("q_scale_enc", "q_basic_enc", q_in_ckpt, q_index),
                self.get_curr_q("q_scale_dec", "q_basic_dec", q_in_ckpt, q_index))

    # ------------------------------------------------------------------ sub-networks
    def _analysis(self, x, q_enc):
        """IntraEncoder.forward (image_model.py:32-35)"""
        net = self.net
        t = layers.rbws(net, "enc.enc_1.0", x)
        t = layers.dcb2(net, "enc.enc_1.1", t, scale=q_enc)
        t = layers.dcb2(net, "enc.enc_2.1", layers.rbws(net, "enc.enc_2.0", t))
        t = layers.dcb2(net, "enc.enc_2.3", layers.rbws(net, "enc.enc_2.2", t))
        return ops.conv(t, net.cw("enc.enc_2.4"), stride=2)

    def _hyper_enc(self, y_pad):
        """hyper_enc (image_model.py:68-73)"""
        net = self.net
        t = layers.dcb2(net, "hyper_enc.0", y_pad)
        t = ops.conv(t, net.cw("hyper_enc.1"), stride=2, act=0.01)
        return ops.conv(t, net.cw("hyper_enc.3"), stride=2)

    def _params(self, z_hat, catb, c):
        """hyper_dec + y_prior_fusion + slice_to_y (image_model.py:74-92,123-125), written
        straight into the params slice of the spatial-prior buffer."""
        net = self.net
        t = layers.rbu(net, "hyper_dec.0", z_hat)
        t = layers.rbu(net, "hyper_dec.1", t)
        t = layers.dcb2(net, "hyper_dec.2", t)
        t = layers.dcb2(net, "y_prior_fusion.0", t)
        dst = catb.slice(c, 3 * c)
        if t.h == catb.h and t.w == catb.w:
            layers.dcb2(net, "y_prior_fusion.1", t, out=dst)
        else:
            ops.copy(layers.dcb2(net, "y_prior_fusion.1", t), out=dst)

    def _prior_fn(self):
        net = self.net

        def fn(step, catb):
            t = ops.conv(catb, net.cw(f"y_spatial_prior_adaptor_{step}"))
            for i in range(3):
                t = layers.dcb2(net, f"y_spatial_prior.{i}", t)
            return t
        return fn

    def _synthesis(self, y_hat, q_dec, clamp):
        """IntraDecoder + refine (image_model.py:55-58,95-98)"""
        net = self.net
        t = layers.rbu(net, "dec.dec_1.1", layers.dcb2(net, "dec.dec_1.0", y_hat))
        t = layers.rbu(net, "dec.dec_1.3", layers.dcb2(net, "dec.dec_1.2", t))
        t = layers.rbu(net, "dec.dec_1.5", layers.dcb2(net, "dec.dec_1.4", t), scale=q_dec)
        t = layers.rbu(net, "dec.dec_2.1", layers.dcb2(net, "dec.dec_2.0", t))
        t = layers.unet(net, "refine.0", t, layers.dcb2)
        # 3 -> 4 output channels (a zero filter): whole 16-byte pixels take the vectorised epilogue
        return ops.conv(t, net.cw("refine.1", cout_pad=4), clamp01=clamp).slice(0, 3)

    def _encode_side(self, x, q_in_ckpt, q_index, write, bits_y=None, bits_z=None):
        q_enc, q_dec = self.get_q_for_inference(q_in_ckpt, q_index)
        xa = Act.wrap(x)
        y = self._analysis(xa, q_enc)
        z = self._hyper_enc(self.pad_for_y(y))
        z_hat, z_sym = self.quant_z(z, "bit_estimator_z", write, bits_z)
        catb = self.new_prior_buffer(y.h, y.w, y.c)
        self._params(z_hat, catb, y.c)
        y_hat, coded = self.forward_four_part_prior(y, catb, self._prior_fn(), write, bits_y)
        if self.taps is not None:
            for k, v in (("y", y), ("z_hat", z_hat), ("params", catb.slice(y.c, 3 * y.c)),
                         ("y_hat", y_hat)):
                self.taps["i_" + k] = v.nchw().detach().float().clone()
            for st, (sym, idx) in enumerate(coded):
                shp = (1, y.c // 4, y.h, y.w)
                self.taps[f"i_y_q_w_{st}"] = sym.view(shp).clone()
                self.taps[f"i_y_idx_w_{st}"] = idx.view(shp).clone()
        return y_hat, coded, z_hat, z_sym, q_dec

    # ------------------------------------------------------------------ reference API
    def forward(self, x, q_in_ckpt=False, q_index=None):
        """image_model.py:114-149 (estimate mode; x_hat is not clamped)."""
        dev = x.device
        bits = torch.zeros(2, dtype=torch.float32, device=dev)
        y_hat, _, _, _, q_dec = self._encode_side(x, q_in_ckpt, q_index, False,
                                                  bits[0:1], bits[1:2])
        x_hat = self._synthesis(y_hat, q_dec, clamp=False)
        pixel_num = x.shape[2] * x.shape[3]
        bpp_y = bits[0:1] / pixel_num
        bpp_z = bits[1:2] / pixel_num
        return {"x_hat": x_hat.nchw(), "bit": torch.sum(bpp_y + bpp_z) * pixel_num,
                "bpp": bpp_y + bpp_z, "bpp_y": bpp_y, "bpp_z": bpp_z}

    def compress(self, x, q_in_ckpt, q_index, reconstruct=True):
        """image_model.py:198-229; stream order z, y parts 0..3.  reconstruct=False (used by
        encode_decode, which returns the decoder's x_hat as the reference does, :187-196) skips the
        encoder-side synthesis and returns x_hat None."""
        zs = get_downsampled_shape(x.shape[2], x.shape[3], 64)
        self.entropy_coder.reserve(self.N * (x.shape[2] // 16 * (x.shape[3] // 16) + zs[0] * zs[1]))
        y_hat, coded, z_hat, z_sym, q_dec = self._encode_side(x, q_in_ckpt, q_index, True)
        self.entropy_coder.reset()
        self.encode_z(z_sym, z_hat, "bit_estimator_z")
        self.encode_y(coded)
        self.entropy_coder.flush()
        if not reconstruct:
            return {"bit_stream": self.entropy_coder.get_encoded_stream(), "x_hat": None}
        x_hat = self._synthesis(y_hat, q_dec, clamp=True)
        return {"bit_stream": self.entropy_coder.get_encoded_stream(), "x_hat": x_hat.nchw()}

    def decompress(self, bit_stream, height, width, q_in_ckpt, q_index):
        """image_model.py:231-252"""
        _, q_dec = self.get_q_for_inference(q_in_ckpt, q_index)
        ec = self.entropy_coder
        z_size = get_downsampled_shape(height, width, 64)
        yh, yw = get_downsampled_shape(height, width, 16)
        ec.reserve(self.N * (yh * yw + z_size[0] * z_size[1]))
        ec.set_stream(bit_stream)
        z_hat = self.decode_z(self.N, z_size, "bit_estimator_z")
        catb = self.new_prior_buffer(yh, yw, self.N)
        self._params(z_hat, catb, self.N)
        y_hat = self.decompress_four_part_prior(catb, self.N, self._prior_fn())
        ec.finish_decode()
        x_hat = self._synthesis(y_hat, q_dec, clamp=True)
        if not ec.lazy_check:       # a truncated / corrupt / wrong-table frame raises here
            ec.check()
        return {"x_hat": x_hat.nchw()}

    def encode_decode(self, x, q_in_ckpt, q_index, output_path=None, pic_width=None,
                      pic_height=None):
        """image_model.py:169-196"""
        if output_path is None:
            encoded = self.forward(x, q_in_ckpt, q_index)
            return {"bit": encoded["bit"].item(), "x_hat": encoded["x_hat"]}
        assert pic_height is not None
        assert pic_width is not None
        compressed = self.compress(x, q_in_ckpt, q_index, reconstruct=False)
        encode_i(pic_height, pic_width, q_in_ckpt, q_index, compressed["bit_stream"], output_path)
        bit = filesize(output_path) * 8
        height, width, q_in_ckpt, q_index, bit_stream = decode_i(output_path)
        decompressed = self.decompress(bit_stream, height, width, q_in_ckpt, q_index)
        return {"bit": bit, "x_hat": decompressed["x_hat"]}
