/* dcvc_b200.h -- C ABI of libdcvc_b200.so (sm_100a CUDA kernels for the DCVC-DC per-frame
 * coding path).
 *
 * This is the inner boundary of the drop-in: the reference reaches native code only through two
 * pybind11 modules (MLCodec_rans / MLCodec_CXX, reference DCVC-DC/src/cpp/py_rans/py_rans.cpp:227-243,
 * DCVC-DC/src/cpp/ops/ops.cpp:84-91) and otherwise through stock torch ops (cuDNN/ATen).  Here
 * every one of those native calls and stock ops on the hot path is an `extern "C"` entry point
 * taking plain pointers and sizes.  The Python host (dcvc_b200/*.py) binds them with ctypes;
 * INTEGRATION.md shows the stubs a reference maintainer would add.
 *
 * Conventions
 *   - All device tensors are fp32 NHWC "views": element (y, x, ch) of a view lives at
 *     data[(y * w + x) * cstride + coff + ch].  cstride >= coff + c lets several producers
 *     write disjoint channel ranges of one buffer, which replaces torch.cat.
 *   - Every function returns 0 on success, a negative code on error (message through
 *     dcvc_last_error()); nothing throws, nothing allocates device memory behind the caller
 *     (workspaces are caller-provided and sized by *_bytes queries).
 *   - All kernels are enqueued on the `stream` argument (a cudaStream_t passed as void*);
 *     calls are asynchronous unless stated otherwise.
 *   - There is no CPU fallback: without a CUDA device every compute entry returns
 *     DCVC_ERR_CUDA.
 */
#ifndef DCVC_B200_H
#define DCVC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DCVC_OK 0
#define DCVC_ERR_ARG (-1)
#define DCVC_ERR_CUDA (-2)
#define DCVC_ERR_UNSUPPORTED (-3)
#define DCVC_ERR_STREAM (-4) /* malformed bitstream / container */

typedef struct dcvc_view {
  float *data;
  int32_t h, w, c;   /* logical extent */
  int32_t cstride;   /* floats between consecutive pixels */
  int32_t coff;      /* first channel of the view inside the pixel */
} dcvc_view;

const char *dcvc_last_error(void);
int dcvc_version(void);
/* number of kernels this library has launched since load (bench.py's gpu_launches) */
int64_t dcvc_launch_count(void);
/* a replayed CUDA graph recorded from n of this library's launches adds n (dcvc_b200/graphs.py) */
void dcvc_launch_count_add(int64_t n);

/* ------------------------------------------------------------------------------------------
 * Convolution with fused prologue/epilogue.  Replaces nn.Conv2d (+ LeakyReLU/ReLU, residual
 * adds, `* q`, nn.PixelShuffle, the ConvFFN2 gate) -- reference DCVC-DC/src/models/layers.py:18-222,
 * video_net.py:58-251.
 *
 *   acc[n]  = bias[n] + sum_{tap,c} lrelu(in[.., c], in_slope) * weight[(tap*cin + c)*cout_w + n]
 *   gate:   v[i] = acc[2i] * lrelu(acc[2i+1], gate_slope)          (cout_w/2 values)
 *   v       = lrelu(v, out_slope)                                   (slope 1 = identity, 0 = ReLU)
 *   v      += res1[dst] + res2[dst]                                 (when present)
 *   v      *= scale[dst channel]                                    (when present)
 *   shuffle: raw channel n goes to pixel (2*oy + ((n>>1)&1), 2*ox + (n&1)), channel n>>2
 *   clamp01: v = min(max(v, 0), 1)                                   (x_hat.clamp_(0, 1))
 *   out[dst] = v ;  out2[dst] = lrelu(v, out2_slope)                (when present)
 *
 * weight layout: [ksize*ksize][cin][cout_w] fp32 ("KN"); for the tensor-core path additionally
 * weight_tc: [cout_pad][ksize*ksize][cin_pad] ("NK", K-major), see dcvc_conv_tc_pack_bytes().
 * `mode`: 0 = fp32 FMA on CUDA cores (exact mode), 1 = tcgen05 kind::tf32 with fp32 TMEM
 * accumulators where the shape is eligible (falls back to 0 otherwise; the decision depends
 * only on the shape, never on data, so encoder and decoder always agree).
 */
typedef struct dcvc_conv {
  dcvc_view in, out, res1, res2, out2;
  const float *weight;
  const float *weight_tc; /* may be NULL -> mode 0 */
  const float *bias;      /* may be NULL */
  const float *scale;     /* may be NULL */
  int32_t ksize, stride, pad;
  int32_t cout_w;         /* raw conv output channels */
  int32_t gate;
  int32_t shuffle;        /* 0 none; 1 depth-to-space; 2 the same, and weight_tc is packed with
                             its cout rows ordered (subpixel, channel) instead of (channel,
                             subpixel) -- `weight` and `bias` always keep the reference order */
  float in_slope, out_slope, gate_slope, out2_slope;
  int32_t mode;
  int32_t clamp01;
  /* half-precision storage of private conv -> conv intermediates (tensor-core path only; the
   * call fails with DCVC_ERR_UNSUPPORTED where that path is not eligible, it never falls back):
   * in_f16:  in.data points at IEEE binary16 elements (cstride/coff count elements) and the MMA
   *          runs kind::f16 on weight_tc16; both operands then carry the 11-bit significand
   *          kind::tf32 would have read from fp32 storage, accumulation stays fp32 in TMEM.
   * out_f16: out.data points at binary16 elements, written round-to-nearest-even after the
   *          fused epilogue; no gate; depth-to-space only with planar-packed weights and no
   *          residual; out2, when present, then holds binary16 elements too (out2 always has
   *          out's element type) and needs a residual-free plain convolution.
   * res1_f16: res1.data points at binary16 elements (same epilogue restriction). */
  const void *weight_tc16; /* [cout_pad][ksize*ksize][cin_pad16] binary16, dcvc_conv_tc_geometry16 */
  int32_t in_f16;
  int32_t out_f16;
  int32_t res1_f16;
  int32_t res2_f16;       /* res2.data points at binary16 elements (TMA-store epilogue only) */
} dcvc_conv;

int dcvc_conv2d(const dcvc_conv *d, void *stream);
/* returns 1 if the tensor-core kernel would take this descriptor */
int dcvc_conv_tc_eligible(const dcvc_conv *d);
/* geometry of the NK packed weight for the tensor-core path: rows n_pad = cout_w rounded up to
 * 16, each row [ksize*ksize][cin_pad] with cin_pad = cin rounded up to the K chunk kc (32, 16 or
 * 8 floats = one 128/64/32-byte swizzle row).  Padding entries must be zero. */
int dcvc_conv_tc_geometry(int cout_w, int cin, int *n_pad, int *cin_pad, int *kc);
/* the same for weight_tc16: kc in {64, 32, 16} binary16 elements */
int dcvc_conv_tc_geometry16(int cout_w, int cin, int *n_pad, int *cin_pad, int *kc);

/* Fused ConvFFN of a DepthConvBlock (reference DCVC-DC/src/models/layers.py:166-179, called from
 * DepthConvBlock.forward :199-209):
 *     out = (in + lrelu(W2 . lrelu(W1 . in + b1, slope1) + b2, slope2)) * scale
 * as one two-GEMM tcgen05 kernel; the 4C-wide hidden activation stays in shared memory.
 * in: binary16 view, c in {32, 48, 64, 128}, >= 512 pixels (dcvc_ffn_eligible); out: fp32 or
 * binary16 view (out_f16) of the same shape.  w1: binary16 [ceil(c/64)][4c][64] (atom a holds input
 * channels 64a..64a+63 zero-padded to 64, row n = output channel of the 1x1 up-projection
 * conv.0; one atom for c <= 64); w2: binary16 [4c/64][c][64] (chunk j holds input
 * channels 64j..64j+63 of the down-projection conv.2, row n = its output channel); b1 [4c],
 * b2 [c], scale [c] or NULL: fp32.  Replaces the two dcvc_conv2d calls of the unfused block. */
typedef struct dcvc_ffn_desc {
  dcvc_view in, out;
  const void *w1, *w2;
  const float *b1, *b2, *scale;
  float slope1, slope2;
  int32_t out_f16;
} dcvc_ffn_desc;
int dcvc_ffn(const dcvc_ffn_desc *d, void *stream);
int dcvc_ffn_eligible(int channels, int64_t pixels);

/* depthwise 3x3, pad 1, stride 1 (DepthConv.depth_conv, layers.py:148); weight [9][c], bias [c] */
int dcvc_dwconv3x3(const dcvc_view *in, const dcvc_view *out, const float *weight,
                   const float *bias, void *stream);
/* the same with both views holding binary16 elements (fp32 weights, fp32 arithmetic, result
 * rounded to nearest even): the private conv1 -> depth_conv -> conv2 intermediates of DepthConv */
int dcvc_dwconv3x3_f16(const dcvc_view *in, const dcvc_view *out, const float *weight,
                       const float *bias, void *stream);

/* ------------------------------------------------------------------------------------------
 * Motion / resampling group (reference DCVC-DC/src/models/video_net.py:8-55,104-126,
 * video_model.py:26-63).
 */
/* NCHW (contiguous, c planes of h*w) <-> NHWC view */
int dcvc_nchw_to_nhwc(const float *src, const dcvc_view *dst, void *stream);
int dcvc_nhwc_to_nchw(const dcvc_view *src, float *dst, int clamp01, void *stream);
/* dst(y,x,:) = mul * lrelu(src(min(y,h-1), min(x,w-1), :), slope): channel-slice copy,
 * replicate pad (pad_for_y, common_model.py:70-78) and crop (slice_to_y, :85-86) */
int dcvc_copy(const dcvc_view *src, const dcvc_view *dst, float mul, float slope, void *stream);
/* dcvc_copy (equal sizes only) with a binary16 destination: dst->data points at binary16
 * elements, dst->cstride/coff count elements; feeds dcvc_conv.in_f16 */
int dcvc_copy_f16(const dcvc_view *src, const dcvc_view *dst, float mul, float slope, void *stream);
/* the same with a binary16 source view (LeakyReLU copies of binary16 ResBlock inputs) */
int dcvc_copy_f16_f16(const dcvc_view *src, const dcvc_view *dst, float mul, float slope,
                      void *stream);
/* fewer than 8 fp32 channels into ONE 16-byte binary16 group of dst (zero-padded): the image next
 * to a binary16 context in ContextualEncoder's first concatenation (video_model.py:186) */
int dcvc_copy_f16_narrow(const dcvc_view *src, const dcvc_view *dst, void *stream);
int dcvc_fill(const dcvc_view *dst, float value, void *stream);
/* 2x2/2 pooling; mode 0 = average (F.avg_pool2d), 1 = max (nn.MaxPool2d) */
int dcvc_pool2(const dcvc_view *src, const dcvc_view *dst, int mode, void *stream);
/* the same between two binary16 views (the UNet's skip buffers, video_net.py:129-214) */
int dcvc_pool2_f16(const dcvc_view *src, const dcvc_view *dst, int mode, void *stream);
/* F.interpolate(mode='bilinear', align_corners=False) to dst's size, times `mul` */
int dcvc_resize_bilinear(const dcvc_view *src, const dcvc_view *dst, float mul, void *stream);
/* flow_warp: grid_sample(bilinear, border, align_corners=True) at (x + flow[0], y + flow[1]) */
int dcvc_flow_warp(const dcvc_view *feat, const dcvc_view *flow, const dcvc_view *out,
                   void *stream);
/* OffsetDiversity from the 96-channel offset/mask tensor at half resolution onward:
 * bilinear x2, 40*tanh + flow, sigmoid mask, 32 three-channel warps, grouped 1x1 fusion
 * (video_model.py:46-63).  fusion_w [48][6], fusion_b [48]. */
int dcvc_offset_diversity(const dcvc_view *x, const dcvc_view *offs_half, const dcvc_view *flow,
                          const float *fusion_w, const float *fusion_b, const dcvc_view *out,
                          void *stream);
/* dcvc_flow_warp / dcvc_offset_diversity writing a binary16 `out` view (fp32 inputs, fp32
 * arithmetic, result rounded to nearest even): the motion-compensated contexts that only
 * tensor-core convolutions read (MultiScaleContextFusion's concatenations). */
int dcvc_flow_warp_f16(const dcvc_view *feat, const dcvc_view *flow, const dcvc_view *out,
                       void *stream);
int dcvc_offset_diversity_f16(const dcvc_view *x, const dcvc_view *offs_half,
                              const dcvc_view *flow, const float *fusion_w, const float *fusion_b,
                              const dcvc_view *out, void *stream);
/* the same with a binary16 source view too (FeatureExtractor outputs, video_model.py:66-86, that
 * only these two gathers and convolutions read); offsets, flow and arithmetic stay fp32 */
int dcvc_flow_warp_f16_f16(const dcvc_view *feat, const dcvc_view *flow, const dcvc_view *out,
                           void *stream);
int dcvc_offset_diversity_f16_f16(const dcvc_view *x, const dcvc_view *offs_half,
                                  const dcvc_view *flow, const float *fusion_w,
                                  const float *fusion_b, const dcvc_view *out, void *stream);

/* ------------------------------------------------------------------------------------------
 * Frame ingest / egress on the device.  Raw 8-bit planes in device memory
 * (uploaded by the caller, e.g. from pinned host buffers) <-> the fp32 4:4:4 tensor the models
 * read and write.  Replaces np_image_to_tensor / astype(float32)/255 of the readers
 * (DCVC-DC/src/utils/video_reader.py:45-161), ycbcr420_to_444(order=0) and ycbcr444_to_420
 * (src/transforms/functional.py:61-92), the replicate pad of test_video.py:130-135 and the
 * squared-error sums of calc_psnr (src/utils/metrics.py:81-92, test_video.py:169-195).
 *   DCVC_FMT_RGB444: planes R, G, B of h*w bytes each.
 *   DCVC_FMT_YUV420: planes Y (h*w), U, V ((h/2)*(w/2) each), h and w even.
 */
#define DCVC_FMT_RGB444 0
#define DCVC_FMT_YUV420 1
size_t dcvc_frame_bytes(int fmt, int h, int w);
/* dst (3 channels, at least h x w; the excess is the replicate pad): plane value / 255, chroma
 * repeated 2x2 for 4:2:0 (scipy.ndimage.zoom order 0) */
int dcvc_frame_u8_to_view(const uint8_t *planes, int fmt, int h, int w, const dcvc_view *dst,
                          void *stream);
/* sse3[c] = sum over plane c of (clip(rec) - src/255)^2 in float64, rec cropped to h x w; for
 * 4:2:0 the chroma planes of rec are the clipped 2x2 means.  rec_planes_out (may be NULL):
 * the same planes rounded to 8 bits (the decoded-frame writers, src/utils/video_writer.py).
 * ws: >= dcvc_plane_sse_ws_bytes(h, w) bytes; the reduction order is fixed. */
size_t dcvc_plane_sse_ws_bytes(int h, int w);
int dcvc_view_plane_sse(const dcvc_view *rec, int fmt, int h, int w, const uint8_t *src_planes,
                        double *sse3, uint8_t *rec_planes_out, void *ws, void *stream);

/* ------------------------------------------------------------------------------------------
 * Quad-part entropy model (reference DCVC-DC/src/models/common_model.py:39-61,92-321,
 * entropy_models.py:269-273).
 *
 * Latents are NHWC views with C channels; C4 = C/4.  `params` holds [quant_step | scales |
 * means] (3C channels).  Step s codes channel group g at the 2x2 position MASK_OF[s][g].
 * For s >= 1 `prior` is the y_spatial_prior output [scales(C) | means(C)]; for s == 0 pass
 * prior.data == NULL and scales/means come from `params`.
 *
 * edges: 256 fp32 thresholds, index(scale) = (number of edges <= max(scale,1e-5)) - 1, built by
 * the host so that the result is bit-identical to GaussianEncoder.build_indexes.
 *
 * symbols/indexes (may be NULL in estimate mode): int16 / uint8 arrays of C4*h*w entries in the
 * reference's [1,C4,h,w] order ((j*h + y)*w + x).
 */
typedef struct dcvc_quad {
  dcvc_view y;        /* encoder input (C); unused by the decode step */
  dcvc_view params;   /* 3C */
  dcvc_view prior;    /* 2C or data == NULL */
  dcvc_view y_hat;    /* C, un-scaled y_hat_so_far slice (conv input of the next step) */
  dcvc_view y_out;    /* C, y_hat * quant_step (final latent), written at coded positions */
  const float *edges; /* device, 256 */
  int32_t step;
  int32_t distribution; /* 0 laplace, 1 gaussian (bit estimate only) */
} dcvc_quad;

/* encoder: quantise the step's positions, write y_hat/y_out, emit symbols + CDF indexes,
 * and (bits != NULL) add the estimated bits of the step to *bits via per-block partials
 * reduced in a fixed order. bits_ws: >= dcvc_quad_bits_ws_bytes(C4*h*w) */
size_t dcvc_quad_bits_ws_bytes(int64_t n);
int dcvc_quad_step_enc(const dcvc_quad *q, int16_t *symbols, uint8_t *indexes, float *bits,
                       void *bits_ws, void *stream);
/* dcvc_quad_step_enc for callers that cannot allocate (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same inputs, outputs and results as dcvc_quad_step_enc, but the per-block
 * bit partials live in a static device array inside the library (64 KiB, held by every GPU that
 * loads the library until exit) instead of a caller workspace; the call allocates nothing.
 * Limit: C4*h*w <= dcvc_quad_step_enc_static_max() (= 1 << 22 positions per step); larger steps
 * return DCVC_ERR_ARG.  All calls share the array: calls that overlap on the GPU (two streams)
 * race on it -- serialise them on one stream.  Nothing has to be cleared between calls (each call
 * overwrites the partials it later reads); *bits is accumulated into, as in dcvc_quad_step_enc. */
int64_t dcvc_quad_step_enc_static_max(void);
int dcvc_quad_step_enc_static(const dcvc_quad *q, int16_t *symbols, uint8_t *indexes, float *bits,
                              void *stream);
/* decoder, phase 1: CDF indexes of the step's positions (scales_r -> build_indexes) */
int dcvc_quad_step_indexes(const dcvc_quad *q, uint8_t *indexes, void *stream);
/* decoder, phase 2: scatter decoded symbols: y_hat = sym + mean at the step's positions */
int dcvc_quad_step_dec(const dcvc_quad *q, const int16_t *symbols, void *stream);

/* ------------------------------------------------------------------------------------------
 * DCVC-HEM: two-step checkerboard ("dual") spatial prior (reference
 * DCVC-HEM/src/models/common_model.py:84-188: get_mask, process_with_mask, forward_dual_prior,
 * decompress_dual_prior).  Channels split in halves; mask_0 = 1 where (y + x) is even.  Step 0
 * codes half 0 at mask_0 and half 1 at mask_1 positions with (scales, means) from the prior
 * fusion; step 1 codes the complement with y_spatial_prior's output.
 *   fused : prior-fusion output [q_step | scales | means] (3C; read by step 0)
 *   cat   : the spatial-prior conv input [y_hat_0_0 | y_hat_1_1 | means | scales | max(q_step,.5)]
 *           (4C), WRITTEN by step 0 (it replaces torch.cat, :125), read by step 1
 *   prior : y_spatial_prior output [scales_0 | means_0 | scales_1 | means_1] (2C; step 1)
 *   y_out : y_hat * q_step (* post_scale[c] when given: the `* curr_q` that follows, e.g.
 *           video_model.py:288), written at the step's positions
 *   symbols / indexes: [1, C/2, h, w] order; entry (jj, y, x) is the one channel of the pair
 *           {jj, jj + C/2} that the step codes at (y, x) -- y_q_w_0 / y_q_w_1 of :148-153.
 */
typedef struct dcvc_dual {
  dcvc_view y;        /* encoder input (C), already divided by curr_q; unused by the decoder */
  dcvc_view fused;    /* 3C */
  dcvc_view prior;    /* 2C (step 1) */
  dcvc_view cat;      /* 4C */
  dcvc_view y_out;    /* C */
  const float *edges; /* device, 256 (see dcvc_quad) */
  const float *post_scale; /* device [C] or NULL */
  int32_t step;       /* 0 or 1 */
  int32_t distribution; /* 0 laplace, 1 gaussian (bit estimate: sigma floor 1e-5 / 0.11) */
} dcvc_dual;
int dcvc_dual_step_enc(const dcvc_dual *q, int16_t *symbols, uint8_t *indexes, float *bits,
                       void *bits_ws, void *stream);   /* bits_ws: dcvc_quad_bits_ws_bytes(C/2*h*w) */
int dcvc_dual_step_indexes(const dcvc_dual *q, uint8_t *indexes, void *stream);
int dcvc_dual_step_dec(const dcvc_dual *q, const int16_t *symbols, void *stream);
/* dst = src * vec[c] (divide = 0) or src / vec[c] (divide = 1, IEEE division): y / curr_q and
 * y_hat * curr_q of DCVC-HEM (video_model.py:263-271,288; image_model.py:61-63,76), x * gate of
 * SELayer (video_net.py:162) */
int dcvc_channel_scale(const dcvc_view *src, const dcvc_view *dst, const float *vec, int divide,
                       void *stream);
/* the same on binary16 views (whole 8-channel groups; fp32 arithmetic, round to nearest even):
 * the SE branch of ConvBlockResidual in tensor-core mode */
int dcvc_channel_scale_f16(const dcvc_view *src, const dcvc_view *dst, const float *vec, int divide,
                           void *stream);
/* SELayer gate (DCVC-HEM/src/models/video_net.py:149-163): gate = sigmoid(W2 relu(W1 mean_hw(x)));
 * w1 [hidden][c], w2 [c][hidden] (nn.Linear weights, no bias); the mean is accumulated in float64
 * in a fixed order.  x: 16-byte aligned view, c <= 256.  ws: >= dcvc_se_ws_bytes(h, w, c). */
size_t dcvc_se_ws_bytes(int h, int w, int c);
int dcvc_se_gate(const dcvc_view *x, const float *w1, const float *w2, int hidden, float *gate,
                 void *ws, void *stream);
/* the same gate of a binary16 view (whole 4-channel groups) */
int dcvc_se_gate_f16(const dcvc_view *x, const float *w1, const float *w2, int hidden, float *gate,
                     void *ws, void *stream);

/* hyper latent: z_hat = round(z) (NHWC in place allowed), symbols in [1,C,h,w] order, and
 * (bits != NULL) factorised-prior bit estimate (BitEstimator, entropy_models.py:58-122;
 * fparams = [h1 b1 a1 h2 b2 a2 h3 b3 a3 h4 b4][C] raw parameters) */
int dcvc_z_quant(const dcvc_view *z, const dcvc_view *z_hat, int16_t *symbols,
                 const float *fparams, float *bits, void *bits_ws, void *stream);
/* decoded z symbols [1,C,h,w] int16 -> NHWC fp32 view */
int dcvc_z_dequant(const int16_t *symbols, const dcvc_view *z_hat, void *stream);

/* ------------------------------------------------------------------------------------------
 * Tables (host side, run once per model).  Replaces MLCodec_CXX.pmf_to_quantized_cdf
 * (DCVC-DC/src/cpp/ops/ops.cpp:24-82).  cdf has n+1 entries.  Host pointers. */
int dcvc_pmf_to_quantized_cdf(const float *pmf, int n, int precision, uint32_t *cdf);

/* ------------------------------------------------------------------------------------------
 * rANS coder on the GPU.  Replaces MLCodec_rans.RansEncoder / RansDecoder
 * (DCVC-DC/src/cpp/py_rans/py_rans.cpp:11-225, rans/rans.cpp:76-331).
 *
 * A table set is {cdf int32 [T][L], sizes int32 [T], offsets int32 [T]} in device memory.
 * Encoder protocol (mirrors reset / encode_with_indexes* / flush / get_encoded_stream):
 *   dcvc_rans_enc_reset, then dcvc_rans_enc_push per call (device symbol/index arrays are
 *   borrowed until flush), then dcvc_rans_enc_flush which runs the coder kernels, assembles the
 *   container on the device, and copies it to `out_host` (synchronises the stream).
 * mode DCVC_RANS_COMPAT: byte-identical to the reference container with `parts` in 1..16
 *   (one device thread per part: this mode exists for interoperability, not speed).
 * mode DCVC_RANS_INTERLEAVED: `parts` warps (1..256) x 32 lanes of 32-bit rANS sharing one word
 *   stream per warp (own container; see DESIGN.md "Interleaved stream").
 * Decoder protocol (mirrors set_stream / decode_stream*): dcvc_rans_dec_set_stream copies the
 *   container to the device and initialises the lanes; each dcvc_rans_dec_pull decodes n symbols
 *   for device-resident indexes into a device int16 array, continuing from the previous call.
 */
#define DCVC_RANS_COMPAT 0
#define DCVC_RANS_INTERLEAVED 1
/* DCVC-HEM's framing (DCVC-HEM/src/cpp/rans/rans_interface.cpp:85-244: BufferedRansEncoder.flush
 * returns the bare rANS words, RansDecoder.set_stream takes them): the compat coder with one part
 * and no flag byte.  Same arithmetic, same tables, symbols clamped to +-30000 like DCVC-DC
 * (HEM itself passes int32 symbols unclamped; such values do not occur with trained weights). */
#define DCVC_RANS_COMPAT_BARE 2
#define DCVC_RANS_MAX_CALLS 32

typedef struct dcvc_rans dcvc_rans; /* opaque coder context (encoder + decoder state) */

typedef struct dcvc_tables {
  const int32_t *cdf;     /* device [T][stride] */
  const int32_t *sizes;   /* device [T] */
  const int32_t *offsets; /* device [T] */
  int32_t count, stride;
  /* optional search table for the interleaved decoder, device [T][256] uint8, 16-byte aligned,
   * lut[t][b] = largest s with cdf[t][s] <= 256 b (any stride <= 257); NULL = none.  The
   * reference searches linearly (rans.cpp:289-292); same result. */
  const uint8_t *lut;
} dcvc_tables;

/* `parts`: 1..16 stream parts in compat mode (the reference container keeps parts-1 in a nibble,
 * py_rans.cpp:99), 1..256 warps in interleaved mode.  Device memory comes from the caller:
 * `workspace` is >= dcvc_rans_workspace_bytes(mode, parts, max_symbols_per_frame) bytes of
 * 256-byte aligned device memory that the context uses until dcvc_rans_destroy (0 = bad
 * arguments).  The context itself only owns pinned host staging buffers, two events and a copy
 * stream.  `stream` orders the workspace initialisation. */
size_t dcvc_rans_workspace_bytes(int mode, int parts, int64_t max_symbols_per_frame);
int dcvc_rans_create(dcvc_rans **out, int mode, int parts, int64_t max_symbols_per_frame,
                     void *workspace, size_t workspace_bytes, void *stream);
void dcvc_rans_destroy(dcvc_rans *r);
int dcvc_rans_enc_reset(dcvc_rans *r);
/* index_is_u8: indexes are uint8 table rows (y parts); otherwise NULL indexes means
 * "index = position / plane" (the factorised z prior: channel id), plane = h*w */
int dcvc_rans_enc_push(dcvc_rans *r, const int16_t *symbols, const uint8_t *indexes, int64_t n,
                       int64_t plane, const dcvc_tables *t);
int dcvc_rans_enc_flush(dcvc_rans *r, uint8_t *out_host, size_t out_capacity, size_t *out_len,
                        void *stream);
/* The same flush in two halves, for callers that keep enqueuing work on `stream` after the coder
 * (compress enqueues the frame's reconstruction between the two: py_rans.cpp's encoder thread,
 * :98-135, plays this role in the reference).  begin: kernels + meta copy + event on `stream`, no
 * host wait.  end: waits for that event only and copies the container to `out_host` on a private
 * stream.  One flush may be pending per coder. */
int dcvc_rans_enc_flush_begin(dcvc_rans *r, size_t out_capacity, void *stream);
/* flush_begin = flush_enqueue + flush_mark.  For callers that record a frame's launches into a
 * CUDA graph: `enqueue` is pure stream work (capturable); `mark` records the event flush_end waits
 * on and is called outside the capture, after the graph launch, on every replay. */
int dcvc_rans_enc_flush_enqueue(dcvc_rans *r, size_t out_capacity, void *stream);
int dcvc_rans_enc_flush_mark(dcvc_rans *r, void *stream);
/* `mark` inside a stream capture: the coder's event becomes an event-record node of the graph
 * (cudaEventRecordExternal) that the host can wait on after each launch while later nodes of the
 * same graph still run; per replay the host then calls dcvc_rans_enc_flush_set_pending. */
int dcvc_rans_enc_flush_mark_captured(dcvc_rans *r, void *stream);
int dcvc_rans_enc_flush_set_pending(dcvc_rans *r);
int dcvc_rans_enc_flush_end(dcvc_rans *r, uint8_t *out_host, size_t out_capacity, size_t *out_len);
int dcvc_rans_dec_set_stream(dcvc_rans *r, const uint8_t *host, size_t len, void *stream);
int dcvc_rans_dec_pull(dcvc_rans *r, const uint8_t *indexes, int64_t n, int64_t plane,
                       const dcvc_tables *t, int16_t *symbols_out, void *stream);
/* Deferred decoder error check (RansDecoder has none: the reference decodes garbage silently,
 * rans.cpp:272-331).  dcvc_rans_dec_finish, called after the frame's last dcvc_rans_dec_pull,
 * enqueues a 4-byte copy of the decode kernels' error word plus an event on `stream`, without a
 * host wait; dcvc_rans_dec_check waits for THAT event only and returns DCVC_ERR_STREAM when a pull
 * ran out of words or met a malformed escape (truncated / corrupt / wrong-table frame).  A check
 * still pending when the next dcvc_rans_dec_set_stream or dcvc_rans_dec_finish arrives is done
 * there, so an error is never lost. */
int dcvc_rans_dec_finish(dcvc_rans *r, void *stream);
int dcvc_rans_dec_check(dcvc_rans *r);
/* device error flag of the decoder/encoder kernels (0 ok); synchronises the stream */
int dcvc_rans_status(dcvc_rans *r, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* DCVC_B200_H */
