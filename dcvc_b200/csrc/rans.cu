// rans.cu -- GPU-resident rANS coder (kernel group 4) and the host-side CDF quantiser.
//
// Replaces MLCodec_rans.RansEncoder/RansDecoder (reference DCVC-DC/src/cpp/py_rans/py_rans.cpp:11-225,
// rans/rans.cpp:76-331) and MLCodec_CXX.pmf_to_quantized_cdf (ops/ops.cpp:24-82).
//
// Symbols and CDF indexes never leave the device: the entropy kernels write them to device
// arrays, the coder kernels read them there, and only the finished container (encode) or the
// received container (decode) crosses PCIe -- versus 10 + 20 blocking .cpu()/.to(device) round
// trips per P frame in the reference (entropy_models.py:37-55).
//
// Containers
//   compat      : byte 0 = ((parts-1)<<4) | (1 if part sizes are 2-byte else 0); parts-1 LE sizes;
//                 part streams (py_rans.cpp:74-119).  Byte-identical to the reference.
//   interleaved : byte 0 = ((warps-1)<<4) | 0xF; warps-1 LE uint32 sizes; per warp: 32 LE uint32
//                 lane states, then LE uint16 words in decode order.  More than 16 warps (up to
//                 256) do not fit the nibble: byte 0 = 0x0E, bytes 1-2 = warps-1 (LE uint16),
//                 then the same sizes and payloads.  Symbol k of a warp's chunk
//                 belongs to lane k%32, round k/32; within a round, operation slot s (0 = the
//                 symbol, >=1 = escape nibbles) is processed for all lanes that have it, and the
//                 lanes that renormalise in that slot take consecutive words in lane order.
#include <string.h>

#include <new>
#include <vector>

#include "common.cuh"
#include "rans_core.h"

namespace {

constexpr int MAX_PARTS = 256;         // interleaved warps
constexpr int MAX_COMPAT_PARTS = 16;   // the reference container keeps parts-1 in one nibble
constexpr size_t DEC_TAB_SMEM_MAX = 200 * 1024;  // decoder's shared-memory table budget
constexpr int SLACK_WORDS = 4096;  // per part, 32-bit words beyond one word per symbol

struct CallRec {
  const int16_t *sym;
  const uint8_t *idx;
  int64_t n, plane;
  rans::Tables t;
};


struct EncJob {
  CallRec calls[DCVC_RANS_MAX_CALLS];
  int ncalls, parts;
  int64_t each_sum;  // sum over the calls of n / parts (symbols of every part but the last)
  int64_t total;     // sum over the calls of n
  int64_t op_off[DCVC_RANS_MAX_CALLS + 1];  // first prepared operation of each call
};

// region bounds of part i inside the encode buffer, in 32-bit words: one word per symbol of the
// part plus SLACK_WORDS (every part but the last holds each_sum symbols)
__host__ __device__ __forceinline__ int64_t reg_off(const EncJob &job, int i) {
  return i < job.parts ? (int64_t)i * (job.each_sum + SLACK_WORDS)
                       : job.total + (int64_t)job.parts * SLACK_WORDS;
}

// meta layout (uint32): [0] error flags, [1] container length, [2+i] part length in bytes,
// [2+MAX_PARTS+i] part start in bytes inside the encode buffer
constexpr int META_WORDS = 2 + 2 * MAX_PARTS;
constexpr uint32_t ERR_ENC_OVERFLOW = 1, ERR_DEC_UNDERFLOW = 2, ERR_OUT_CAPACITY = 4, ERR_ESCAPE = 8;

__device__ __forceinline__ void part_range(int64_t n, int parts, int i, int64_t &off, int64_t &cur) {
  const int64_t each = n / parts;
  off = (int64_t)i * each;
  cur = i < parts - 1 ? each : n - each * (parts - 1);
}

__device__ __forceinline__ int call_index(const CallRec &c, int64_t pos) {
  return c.idx ? (int)c.idx[pos] : (int)(pos / c.plane);
}

// ------------------------------------------------------------------ compat kernels
// Pass 1 (whole grid): per symbol {start | freq << 16, escape ? (raw << 1) | 1 : 0, m} with
// m = floor((2^64 - 1) / freq), so that the sequential pass replaces its 64-bit division by a
// multiply-high and a short correction and never waits on a table load.
__global__ void compat_prepare_kernel(const EncJob job, uint4 *ops) {
  const int64_t total = job.op_off[job.ncalls];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    int c = 0;
    while (c + 1 < job.ncalls && i >= job.op_off[c + 1]) ++c;
    const CallRec &cr = job.calls[c];
    const int64_t pos = i - job.op_off[c];
    const int index = call_index(cr, pos);
    const int32_t *cdf = cr.t.cdf + (int64_t)index * cr.t.stride;
    const int32_t max_value = cr.t.sizes[index] - 2;
    int32_t value = (int32_t)cr.sym[pos] - cr.t.offsets[index];
    uint32_t esc = 0, raw = 0;
    if (value < 0) { raw = (uint32_t)(-2 * value - 1); value = max_value; esc = 1; }
    else if (value >= max_value) { raw = (uint32_t)(2 * (value - max_value)); value = max_value; esc = 1; }
    const uint32_t start = (uint32_t)cdf[value], freq = (uint32_t)(cdf[value + 1] - cdf[value]);
    const uint64_t m = freq ? 0xffffffffffffffffull / freq : 0;
    ops[i] = make_uint4(start | (freq << 16), esc ? ((raw << 1) | 1u) : 0u, (uint32_t)m, (uint32_t)(m >> 32));
  }
}

// Pass 2: one warp per stream part.  The warp fetches 32 prepared operations at a time
// (coalesced, the next batch in flight while lane 0 works); lane 0 runs the dependent rANS chain
// of rans.cpp:141-168 (LIFO: last symbol first, within an escape the digits before the count
// before the escape interval).
__global__ void __launch_bounds__(32) compat_encode_kernel(const EncJob job, const uint4 *ops_all,
                                                           uint32_t *buf, uint32_t *meta) {
  // One sequential 64-bit state per part (rans.cpp:76-168).  The warp runs the chain redundantly
  // in all 32 lanes (same state, same instructions): lane l holds prepared operation l of the
  // current group of 32 in registers and every lane fetches operation j by shuffle, off the
  // state's dependent chain, so there is no shared-memory round trip and no warp barrier per
  // group; lane 0 alone writes the emitted words.
  const int i = blockIdx.x, lane = threadIdx.x;
  rans::Enc64 e;
  rans::enc64_init(e, buf + reg_off(job, i + 1), buf + reg_off(job, i));
  auto emit = [&]() {
    if (e.ptr <= e.limit) e.overflow = 1;
    else { --e.ptr; if (lane == 0) *e.ptr = (uint32_t)e.x; }
    e.x >>= 32;
  };
  for (int c = job.ncalls - 1; c >= 0; --c) {
    const CallRec &cr = job.calls[c];
    int64_t off, cur;
    part_range(cr.n, job.parts, i, off, cur);
    const uint4 *ops = ops_all + job.op_off[c] + off;
    // groups of 32 from the top: the group ending at hi covers k in [hi - 32, hi), lane = k - (hi - 32)
    int64_t hi = cur;
    uint4 nxt = make_uint4(0, 0, 0, 0);
    if (hi - 32 + lane >= 0 && hi > 0) nxt = __ldg(ops + hi - 32 + lane);
    while (hi > 0) {
      const uint4 mine = nxt;
      const int64_t lo = hi - 32 > 0 ? hi - 32 : 0;
      if (lo > 0 && lo - 32 + lane >= 0) nxt = __ldg(ops + lo - 32 + lane);
      const int jlo = (int)(lo - (hi - 32));               // first valid lane of this group
#pragma unroll 4
      for (int j = 31; j >= jlo; --j) {
        const uint32_t ox = __shfl_sync(0xffffffffu, mine.x, j), oy = __shfl_sync(0xffffffffu, mine.y, j);
        const uint32_t oz = __shfl_sync(0xffffffffu, mine.z, j), ow = __shfl_sync(0xffffffffu, mine.w, j);
        const uint32_t freq = ox >> 16, start = ox & 0xffffu;
        if (oy & 1u) {                                     // escape: digits (high first), count
          const uint32_t raw = oy >> 1;
          const int nb = raw ? (35 - __clz(raw)) >> 2 : 0;
          for (int d = nb - 1; d >= -1; --d) {
            const uint32_t nib = d < 0 ? (uint32_t)nb : (raw >> (4 * d)) & 15u;
            if (e.x >= (1ull << 59)) emit();
            e.x = (e.x << rans::BYPASS_BITS) | nib;
          }
        }
        if (e.x >= ((uint64_t)freq << 47)) emit();
        const uint64_t m = (uint64_t)oz | ((uint64_t)ow << 32);
        uint64_t q = __umul64hi(e.x, m);                   // <= x / freq, short by at most 2
        uint64_t rem = e.x - q * freq;
        while (rem >= freq) { rem -= freq; ++q; }
        e.x = (q << rans::PRECISION) + rem + start;
      }
      hi = lo;
    }
  }
  if (lane != 0) return;
  rans::enc64_flush(e);
  if (e.overflow) atomicOr(meta, ERR_ENC_OVERFLOW);
  meta[2 + i] = (uint32_t)((buf + reg_off(job, i + 1) - e.ptr) * 4);
  meta[2 + MAX_PARTS + i] = (uint32_t)((e.ptr - buf) * 4);
}

struct DecPart {
  uint64_t x;            // compat state
  uint32_t lane_x[32];   // interleaved states
  uint32_t rp, end;      // read position / end, in words (compat: 32-bit, interleaved: 16-bit)
  uint32_t base;         // byte offset of the part inside the device stream buffer
  uint32_t pad;
};

__global__ void compat_dec_init_kernel(const uint8_t *stream, DecPart *parts, uint32_t *meta) {
  if (threadIdx.x != 0) return;
  DecPart &p = parts[blockIdx.x];
  const uint32_t *w = reinterpret_cast<const uint32_t *>(stream + p.base);
  rans::Dec64 d;
  rans::dec64_init(d, w, w + p.end);
  p.x = d.x;
  p.rp = (uint32_t)(d.ptr - w);
  if (d.underflow) atomicOr(meta, ERR_DEC_UNDERFLOW);
}

// One block per stream part: all threads stage the call's CDF rows in shared memory (the symbol
// search of rans.cpp:289-292 is the dependent chain), then warp 0 decodes: the warp fetches 32
// table indexes at a time, lane 0 runs the sequential rANS state, the warp stores 32 symbols.
constexpr int CDEC_THREADS = 128;
template <bool TAB_SMEM>
__global__ void __launch_bounds__(CDEC_THREADS) compat_decode_kernel(const CallRec cr, int nparts,
                                                                     const uint8_t *stream,
                                                                     DecPart *parts, int16_t *out,
                                                                     uint32_t *meta) {
  extern __shared__ int32_t tab[];
  __shared__ int s_idx[32];
  __shared__ int16_t s_out[32];
  rans::Tables t = cr.t;
  if (TAB_SMEM) {
    const int ncdf = cr.t.count * cr.t.stride;
    for (int k = threadIdx.x; k < ncdf; k += CDEC_THREADS) tab[k] = cr.t.cdf[k];
    for (int k = threadIdx.x; k < cr.t.count; k += CDEC_THREADS) {
      tab[ncdf + k] = cr.t.sizes[k];
      tab[ncdf + cr.t.count + k] = cr.t.offsets[k];
    }
    __syncthreads();
    t.cdf = tab; t.sizes = tab + ncdf; t.offsets = tab + ncdf + cr.t.count;
  }
  if (threadIdx.x >= 32) return;
  const int i = blockIdx.x, lane = threadIdx.x;
  DecPart &p = parts[i];
  const uint32_t *w = reinterpret_cast<const uint32_t *>(stream + p.base);
  rans::Dec64 d;
  d.x = p.x; d.ptr = w + p.rp; d.end = w + p.end; d.underflow = 0;
  int64_t off, cur;
  part_range(cr.n, nparts, i, off, cur);
  int nxt = (lane < cur) ? call_index(cr, off + lane) : -1;
  for (int64_t k0 = 0; k0 < cur; k0 += 32) {
    s_idx[lane] = nxt;
    const int64_t kn = k0 + 32 + lane;
    nxt = kn < cur ? call_index(cr, off + kn) : -1;
    __syncwarp();
    if (lane == 0) {
      const int nb = cur - k0 < 32 ? (int)(cur - k0) : 32;
      for (int j = 0; j < nb; ++j) s_out[j] = (int16_t)rans::dec64_symbol(d, s_idx[j], t);
    }
    __syncwarp();
    if (k0 + lane < cur) out[off + k0 + lane] = s_out[lane];
    __syncwarp();
  }
  if (lane != 0) return;
  p.x = d.x;
  p.rp = (uint32_t)(d.ptr - w);
  if (d.underflow) atomicOr(meta, ERR_DEC_UNDERFLOW);
}

__device__ __forceinline__ uint32_t lds32(uint32_t a) {
  uint32_t v;
  asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ uint32_t lds8(uint32_t a) {
  uint32_t v;
  asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}

constexpr int LUT_PITCH = 260;   // bytes per LUT row in shared memory: 256 buckets + the row's last symbol

// words of shared memory in front of the LUT: cdf rows + one {size, offset} word per row, padded
// to 16 bytes
__host__ __device__ __forceinline__ size_t dec_tab_words(int count, int stride) {
  return ((size_t)count * stride + count + 3) & ~(size_t)3;
}


// The decoders' shared-memory image of a call's table set: [count][stride] cdf | [count]
// size << 16 | (offset & 0xffff) | (LUT) [count][LUT_PITCH] bytes: the 256 buckets of
// dcvc_tables.lut and, at byte 256, the row's last symbol.  Returns shared-space addresses.
template <bool LUT, int THREADS>
__device__ __forceinline__ void stage_tables(const CallRec &cr, int32_t *tab, uint32_t &s_cdf,
                                             uint32_t &s_row, uint32_t &s_lut) {
  const int ncdf = cr.t.count * cr.t.stride;
  if ((ncdf & 3) == 0 && (reinterpret_cast<uintptr_t>(cr.t.cdf) & 15) == 0) {
    const int4 *src = reinterpret_cast<const int4 *>(cr.t.cdf);
    int4 *dst = reinterpret_cast<int4 *>(tab);
#pragma unroll 4
    for (int i = threadIdx.x; i < (ncdf >> 2); i += THREADS) dst[i] = __ldg(src + i);
  } else {
    for (int i = threadIdx.x; i < ncdf; i += THREADS) tab[i] = cr.t.cdf[i];
  }
  for (int i = threadIdx.x; i < cr.t.count; i += THREADS)
    tab[ncdf + i] = (cr.t.sizes[i] << 16) | (cr.t.offsets[i] & 0xffff);
  const size_t lw = dec_tab_words(cr.t.count, cr.t.stride);
  if (LUT) {
    const uint32_t *src = reinterpret_cast<const uint32_t *>(cr.t.lut);
    uint32_t *dst = reinterpret_cast<uint32_t *>(tab + lw);
#pragma unroll 8
    for (int i = threadIdx.x; i < cr.t.count * 64; i += THREADS)
      dst[(i >> 6) * (LUT_PITCH / 4) + (i & 63)] = __ldg(src + i);
    for (int i = threadIdx.x; i < cr.t.count; i += THREADS)
      dst[i * (LUT_PITCH / 4) + 64] = (uint32_t)(cr.t.sizes[i] - 2);
  }
  __syncthreads();
  s_cdf = (uint32_t)__cvta_generic_to_shared(tab);
  asm volatile("" : "+r"(s_cdf));      // keep the base in a register (else re-derived every round)
  s_row = s_cdf + 4u * (uint32_t)ncdf;
  s_lut = s_cdf + 4u * (uint32_t)lw;
}

// Compat decoder with the tables in shared memory.  The reference's stream is ONE sequential
// rANS state per part (rans.cpp:272-331), so the warp runs that chain redundantly in all 32 lanes
// (same state, same instructions): a table index comes from the lane that prefetched it by
// shuffle, lane j keeps symbol j of a group of 32 for the coalesced store, stream words come from
// a shared-memory ring filled a window ahead (see ilv_decode_kernel), and the symbol search starts
// from the host-built LUT.  ~120 dependent cycles per symbol instead of ~900 for the generic
// single-lane loop (compat_decode_kernel, kept for table sets that do not fit).
template <bool LUT>
__global__ void __launch_bounds__(CDEC_THREADS) compat_decode_fast_kernel(const CallRec cr, int nparts,
                                                                          const uint8_t *stream,
                                                                          DecPart *parts, int16_t *out,
                                                                          uint32_t *meta) {
  extern __shared__ __align__(16) int32_t tab[];
  __shared__ uint32_t ring[128];
  uint32_t s_cdf, s_row, s_lut;
  stage_tables<LUT, CDEC_THREADS>(cr, tab, s_cdf, s_row, s_lut);
  if (threadIdx.x >= 32) return;
  const int i = blockIdx.x, lane = threadIdx.x;
  DecPart &p = parts[i];
  const uint32_t *w = reinterpret_cast<const uint32_t *>(stream + p.base);
  uint64_t x = p.x;
  uint32_t rp = p.rp;
  const uint32_t end = p.end;
  uint32_t err = 0;
  auto fetch = [&](uint32_t pos) -> uint32_t { return pos < end ? w[pos] : 0u; };
  uint32_t cw = rp >> 5;
  ring[(cw * 32 + lane) & 127] = fetch(cw * 32 + lane);
  ring[((cw + 1) * 32 + lane) & 127] = fetch((cw + 1) * 32 + lane);
  uint32_t pend = fetch((cw + 2) * 32 + lane);
  __syncwarp();
  auto word = [&]() -> uint32_t {                          // dec64_word, warp-uniform
    if (rp >= end) { err |= ERR_DEC_UNDERFLOW; return 0u; }
    const uint32_t v = ring[rp & 127];
    ++rp;
    if ((rp >> 5) != cw) {
      ++cw;
      __syncwarp();
      ring[((cw + 1) * 32 + lane) & 127] = pend;
      pend = fetch((cw + 2) * 32 + lane);
      __syncwarp();
    }
    return v;
  };
  auto bits4 = [&]() -> uint32_t {                         // dec64_bits(d, BYPASS_BITS)
    const uint32_t val = (uint32_t)x & 15u;
    x >>= 4;
    if (x < rans::L64) x = (x << 32) | word();
    return val;
  };
  int64_t off, cur;
  part_range(cr.n, nparts, i, off, cur);
  const uint32_t n = (uint32_t)cur;
  const uint8_t *ip = cr.idx ? cr.idx + off : nullptr;
  int16_t *o = out + off + lane;
  const uint32_t stride = (uint32_t)cr.t.stride;
  auto load_idx = [&](uint32_t k) -> int {
    if (k >= n) return 0;
    return ip ? (int)ip[k] : (int)((uint64_t)(off + k) / (uint64_t)cr.plane);
  };
  int nxt = load_idx(lane);
  for (uint32_t k0 = 0; k0 < n; k0 += 32) {
    const int mine_idx = nxt;
    nxt = load_idx(k0 + 32 + lane);
    const int nb = n - k0 < 32u ? (int)(n - k0) : 32;
    int mine = 0;
#pragma unroll 4
    for (int j = 0; j < nb; ++j) {
      const uint32_t index = (uint32_t)__shfl_sync(0xffffffffu, mine_idx, j);
      const uint32_t row = lds32(s_row + 4u * index);
      const uint32_t c = s_cdf + 4u * index * stride;
      const int maxv = (int)(row >> 16) - 2;
      const uint32_t cum = (uint32_t)x & 0xffffu;
      int lo = 0, hi = maxv;
      if (LUT) {
        const uint32_t l = s_lut + (uint32_t)LUT_PITCH * index + (cum >> 8);
        lo = (int)lds8(l);
        hi = (int)lds8(l + 1);
      }
      while (lo < hi) {                                    // largest s with cdf[s] <= cum
        const int mid = (lo + hi + 1) >> 1;
        if (lds32(c + 4u * mid) <= cum) lo = mid; else hi = mid - 1;
      }
      const uint32_t start = lds32(c + 4u * lo), next = lds32(c + 4u * lo + 4u);
      x = (uint64_t)(next - start) * (x >> 16) + (cum - start);
      if (x < rans::L64) x = (x << 32) | word();
      int value = lo;
      if (lo == maxv) {                                    // bypass escape, rans.cpp:296-318
        uint32_t val = bits4();
        int nbv = (int)val;
        while (val == 15u && !(err & ERR_DEC_UNDERFLOW)) {
          val = bits4();
          nbv += (int)val;
        }
        uint32_t raw = 0;
        for (int q = 0; q < nbv && q < 8; ++q) raw |= bits4() << (4 * q);
        value = (int)(raw >> 1);
        if (raw & 1u) value = -value - 1; else value += maxv;
      }
      if (lane == j) mine = value + (int)(int16_t)(row & 0xffffu);
    }
    if (k0 + lane < n) *o = (int16_t)mine;
    o += 32;
  }
  if (lane != 0) return;
  p.x = x;
  p.rp = rp;
  if (err) atomicOr(meta, err);
}

// ------------------------------------------------------------------ interleaved kernels
__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// Pass 1 (whole grid): every symbol's coding interval, looked up in parallel so that the
// sequential pass below never waits on a table load.  ops[i] = {start | freq << 16,
// (floor((2^32-1)/freq) & ~1) | escape flag}; the reciprocal turns the per-symbol x / freq of the
// state update into a multiply plus a short correction.
__global__ void ilv_prepare_kernel(const EncJob job, uint2 *ops) {
  const int64_t total = job.op_off[job.ncalls];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    int c = 0;
    while (c + 1 < job.ncalls && i >= job.op_off[c + 1]) ++c;
    const CallRec &cr = job.calls[c];
    const int64_t pos = i - job.op_off[c];
    const int index = call_index(cr, pos);
    const int32_t *cdf = cr.t.cdf + (int64_t)index * cr.t.stride;
    const int32_t max_value = cr.t.sizes[index] - 2;
    int32_t value = (int32_t)cr.sym[pos] - cr.t.offsets[index];
    // out of the table's support: the escape interval, and instead of the reciprocal the raw
    // value whose nibbles follow it (rans.cpp:88-101): word y = (raw << 1) | 1, raw < 2^17
    uint32_t esc = 0, raw = 0;
    if (value < 0) { raw = (uint32_t)(-2 * value - 1); value = max_value; esc = 1; }
    else if (value >= max_value) { raw = (uint32_t)(2 * (value - max_value)); value = max_value; esc = 1; }
    const uint32_t start = (uint32_t)cdf[value], freq = (uint32_t)(cdf[value + 1] - cdf[value]);
    ops[i] = make_uint2(start | (freq << 16), esc ? ((raw << 1) | 1u) : ((0xffffffffu / freq) & ~1u));
  }
}

constexpr int ENC_BATCH = 8;  // rounds whose operations are fetched before the dependent chain

// Pass 2 (one warp per stream): the dependent rANS chain, fed from the prepared operations.
__global__ void __launch_bounds__(32) ilv_encode_kernel(const EncJob job, const uint2 *ops_all,
                                                        uint32_t *buf, uint32_t *meta) {
  const int w = blockIdx.x, lane = threadIdx.x;
  uint16_t *base = reinterpret_cast<uint16_t *>(buf + reg_off(job, w));
  const int64_t cap = (reg_off(job, w + 1) - reg_off(job, w)) * 2;
  int64_t wp = cap;
  uint32_t x = rans::L32;
  bool overflow = false;
  const uint32_t lt = lanemask_lt();
  for (int c = job.ncalls - 1; c >= 0; --c) {
    const CallRec &cr = job.calls[c];
    int64_t off, cur;
    part_range(cr.n, job.parts, w, off, cur);
    const uint2 *ops = ops_all + job.op_off[c] + off;
    const int64_t rounds = (cur + 31) >> 5;
    for (int64_t r_hi = rounds; r_hi > 0; r_hi -= ENC_BATCH) {
      uint2 o[ENC_BATCH];
#pragma unroll
      for (int j = 0; j < ENC_BATCH; ++j) {
        const int64_t k = (r_hi - 1 - j) * 32 + lane;
        o[j] = (k >= 0 && k < cur) ? __ldg(ops + k) : make_uint2(0u, 0u);   // freq 0: inactive
      }
#pragma unroll
      for (int j = 0; j < ENC_BATCH; ++j) {
        if (r_hi - 1 - j < 0) break;                     // warp-uniform
        const uint32_t freq = o[j].x >> 16, start = o[j].x & 0xffffu;
        const bool active = freq != 0;
        if (__ballot_sync(0xffffffffu, active && (o[j].y & 1u)) == 0) {
          // common case: one interval per lane
          const bool emit = active && x >= (freq << 16);
          const uint32_t m = __ballot_sync(0xffffffffu, emit);
          if (m) {
            const int cnt = __popc(m);
            if (wp - cnt < 64) {
              overflow = true;
            } else {
              wp -= cnt;
              if (emit) base[wp + __popc(m & lt)] = (uint16_t)x;
            }
            if (emit) x >>= 16;
          }
          if (active) {
            uint32_t q = __umulhi(x, o[j].y & ~1u);      // <= x / freq, short by at most 3
            uint32_t rem = x - q * freq;
            while (rem >= freq) { rem -= freq; ++q; }
            x = (q << 16) + rem + start;
          }
        } else {
          // some lane escapes its table's support: its operations are the escape interval
          // (slot 0), the digit count (slot 1; an int16 symbol has at most 5 hex digits, so the
          // count is one nibble) and the digits of raw, low first (rans.cpp:76-139); walk the
          // slots backwards, all lanes in step
          const bool esc = active && (o[j].y & 1u);
          const uint32_t raw = o[j].y >> 1;
          const int nb = esc ? (raw ? (35 - __clz(raw)) >> 2 : 0) : 0;
          const int nops = active ? (esc ? 2 + nb : 1) : 0;
          const int maxops = __reduce_max_sync(0xffffffffu, nops);
          for (int s = maxops - 1; s >= 0; --s) {
            const bool has = s < nops;
            const bool bypass = s > 0;
            const uint32_t x_max = bypass ? (1u << 28) : (freq << 16);
            const bool emit = has && x >= x_max;
            const uint32_t m = __ballot_sync(0xffffffffu, emit);
            const int cnt = __popc(m);
            if (cnt) {
              if (wp - cnt < 64) {
                overflow = true;
              } else {
                wp -= cnt;
                if (emit) base[wp + __popc(m & lt)] = (uint16_t)x;
              }
              if (emit) x >>= 16;
            }
            if (has) {
              if (bypass) {
                const uint32_t nib = s == 1 ? (uint32_t)nb : (raw >> (4 * (s - 2))) & 15u;
                x = (x << rans::BYPASS_BITS) | nib;
              } else if (esc) {
                x = ((x / freq) << 16) + (x % freq) + start;
              } else {
                uint32_t q = __umulhi(x, o[j].y & ~1u);
                uint32_t rem = x - q * freq;
                while (rem >= freq) { rem -= freq; ++q; }
                x = (q << 16) + rem + start;
              }
            }
          }
        }
      }
    }
  }
  wp -= 64;
  base[wp + 2 * lane] = (uint16_t)x;
  base[wp + 2 * lane + 1] = (uint16_t)(x >> 16);
  if (overflow && lane == 0) atomicOr(meta, ERR_ENC_OVERFLOW);
  if (lane == 0) {
    meta[2 + w] = (uint32_t)((cap - wp) * 2);
    meta[2 + MAX_PARTS + w] = (uint32_t)(reg_off(job, w) * 4 + wp * 2);
  }
}

__global__ void ilv_dec_init_kernel(const uint8_t *stream, DecPart *parts, uint32_t *meta) {
  DecPart &p = parts[blockIdx.x];
  const uint16_t *w = reinterpret_cast<const uint16_t *>(stream + p.base);
  const int lane = threadIdx.x;
  if (p.end < 64) {
    if (lane == 0) atomicOr(meta, ERR_DEC_UNDERFLOW);
    p.lane_x[lane] = rans::L32;
  } else {
    p.lane_x[lane] = (uint32_t)w[2 * lane] | ((uint32_t)w[2 * lane + 1] << 16);
  }
  if (lane == 0) p.rp = 64;
}

constexpr int DEC_THREADS = 128;   // warp 0 decodes; all four warps stage the tables
constexpr int DEC_BATCH = 8;       // rounds whose table indexes are fetched ahead of the chain

// One warp per stream.  The CDF rows of this call's table set are staged in shared memory first
// (TAB_SMEM; the symbol search per round is the dependent chain, and a shared-memory probe costs a
// tenth of an L2 one); table sets larger than the budget are searched in place.
//
// A stream is decoded by ONE warp, one round of 32 symbols after the other, and every round
// depends on the previous one through the lanes' states: the kernel is bound by the number of
// dependent instructions per round (ncu: 6.5 cycles per instruction, one warp per scheduler,
// "wait" = fixed-latency dependency is the top stall), so the common round is kept short, free of
// memory waits and with two branches only (search miss, renormalise-or-escape):
//  * WORD RING: the stream's renormalisation words pass through a 128-word ring in shared memory,
//    filled one 32-word window at a time; the window loaded at one rotation sits in a register
//    until the next rotation stores it, so no instruction ever waits for that global load.  A
//    round consumes at most 32 words: the ring always holds the current and the next window.
//  * SEARCH LUT (LUT; dcvc_tables.lut, built once per table set by the host): per CDF row,
//    lut[b] = largest s with cdf[s] <= 256 b; the symbol for cum lies in [lut[cum >> 8],
//    lut[(cum >> 8) + 1]]: no probe at all for the peaked rows instead of seven.
//  * table rows are addressed with 32-bit shared-space addresses, {size, offset} of a row share
//    one word, table indexes are fetched one batch (8 rounds) ahead, whole batches run without
//    per-round bounds checks; the last rounds of a part go through the checked round.
template <bool TAB_SMEM, bool LUT>
__global__ void __launch_bounds__(DEC_THREADS) ilv_decode_kernel(const CallRec cr, int nparts,
                                                                 const uint8_t *stream,
                                                                 DecPart *parts, int16_t *out,
                                                                 uint32_t *meta) {
  extern __shared__ __align__(16) int32_t tab[];   // [count][stride] cdf | [count] size<<16|offset | lut
  __shared__ uint16_t ring[128];
  uint32_t s_cdf = 0, s_row = 0, s_lut = 0;
  if (TAB_SMEM) stage_tables<LUT, DEC_THREADS>(cr, tab, s_cdf, s_row, s_lut);
  if (threadIdx.x >= 32) return;
  const int wi = blockIdx.x, lane = threadIdx.x;
  DecPart &p = parts[wi];
  const uint16_t *w = reinterpret_cast<const uint16_t *>(stream + p.base);
  uint32_t x = p.lane_x[lane];
  uint32_t rp = p.rp;
  const uint32_t end = p.end;
  const uint32_t lt = lanemask_lt();
  uint32_t err = 0;
  auto fetch = [&](uint32_t pos) -> uint16_t { return pos < end ? w[pos] : (uint16_t)0; };
  uint32_t cw = rp >> 5;                                  // window of the read position
  ring[(cw * 32 + lane) & 127] = fetch(cw * 32 + lane);
  ring[((cw + 1) * 32 + lane) & 127] = fetch((cw + 1) * 32 + lane);
  uint16_t pend = fetch((cw + 2) * 32 + lane);            // window cw + 2, stored at the next rotation
  __syncwarp();
  // renormalisation of the lanes in `need` (ballot m != 0, warp-uniform): 16 bits each, lane order
  auto refill = [&](bool need, uint32_t m) {
    const uint32_t pos = rp + __popc(m & lt);
    if (need) {
      if (pos >= end) err |= ERR_DEC_UNDERFLOW;
      x = (x << 16) | ring[pos & 127];
    }
    rp += __popc(m);
    if ((rp >> 5) != cw) {
      ++cw;
      __syncwarp();
      ring[((cw + 1) * 32 + lane) & 127] = pend;
      pend = fetch((cw + 2) * 32 + lane);
      __syncwarp();
    }
  };
  // bypass escape of the lanes with phase == 1 (rans.cpp:296-318): count nibbles, then raw nibbles
  auto escape = [&](int &phase, int &value, int maxv) {
    int nb = 0, jn = 0;
    uint32_t raw = 0;
    do {
      uint32_t val = 0;
      if (phase) {
        val = x & 15u;
        x >>= 4;
      }
      const bool need2 = phase != 0 && x < rans::L32;
      const uint32_t m2 = __ballot_sync(0xffffffffu, need2);
      if (m2) refill(need2, m2);
      bool done = false;
      if (phase == 1) {
        nb += (int)val;
        if (val != 15u) {
          if (nb == 0) done = true; else phase = 2;
          if (nb > 8) { err |= ERR_ESCAPE; done = true; }
        }
      } else if (phase == 2) {
        raw |= val << (4 * jn);
        if (++jn == nb) done = true;
      }
      if (done) {
        value = (int)(raw >> 1);
        if (raw & 1u) value = -value - 1; else value += maxv;
        phase = 0;
      }
      if (err & ERR_DEC_UNDERFLOW) phase = 0;  // never spin on a truncated stream
    } while (__any_sync(0xffffffffu, phase != 0));
  };
  int64_t off, cur;
  part_range(cr.n, nparts, wi, off, cur);
  const uint32_t n = (uint32_t)cur;                      // < 2^31: dcvc_rans_dec_pull checks
  const uint8_t *ip = cr.idx ? cr.idx + off : nullptr;
  int16_t *o = out + off + lane;
  const uint32_t stride = (uint32_t)cr.t.stride;
  auto load_idx = [&](uint32_t k) -> int {
    if (k >= n) return -1;
    return ip ? (int)ip[k] : (int)((uint64_t)(off + k) / (uint64_t)cr.plane);
  };
  const uint32_t rounds = (n + 31) >> 5;
  uint32_t r0 = 0;
  if (TAB_SMEM) {
    const uint32_t full = n >> 5;                        // rounds in which every lane has a symbol
    int nxt[DEC_BATCH];
#pragma unroll
    for (int j = 0; j < DEC_BATCH; ++j) nxt[j] = load_idx((uint32_t)j * 32 + lane);
    for (; r0 + DEC_BATCH <= full; r0 += DEC_BATCH) {
      uint32_t rowm[DEC_BATCH], cbase[DEC_BATCH], lbase[DEC_BATCH];
#pragma unroll
      for (int j = 0; j < DEC_BATCH; ++j) {
        const uint32_t index = (uint32_t)nxt[j];
        rowm[j] = lds32(s_row + 4u * index);
        cbase[j] = s_cdf + 4u * index * stride;
        lbase[j] = s_lut + (uint32_t)LUT_PITCH * index;
      }
#pragma unroll
      for (int j = 0; j < DEC_BATCH; ++j) nxt[j] = load_idx((r0 + DEC_BATCH + j) * 32 + lane);
#pragma unroll
      for (int j = 0; j < DEC_BATCH; ++j) {
        const int maxv = (int)(rowm[j] >> 16) - 2;
        const uint32_t cum = x & 0xffffu;
        int lo = 0, hi = maxv;
        if (LUT) {
          const uint32_t l = lbase[j] + (cum >> 8);
          lo = (int)lds8(l);
          hi = (int)lds8(l + 1);
        }
        while (lo < hi) {                                // largest s with cdf[s] <= cum
          const int mid = (lo + hi + 1) >> 1;
          if (lds32(cbase[j] + 4u * mid) <= cum) lo = mid; else hi = mid - 1;
        }
        const uint32_t start = lds32(cbase[j] + 4u * lo), next = lds32(cbase[j] + 4u * lo + 4u);
        x = (next - start) * (x >> 16) + (cum - start);
        int value = lo;
        int phase = lo == maxv ? 1 : 0;
        const bool need = x < rans::L32;
        if (__any_sync(0xffffffffu, need || phase)) {
          const uint32_t m = __ballot_sync(0xffffffffu, need);
          if (m) refill(need, m);
          if (__any_sync(0xffffffffu, phase != 0)) escape(phase, value, maxv);
        }
        *o = (int16_t)(value + (int)(int16_t)(rowm[j] & 0xffffu));
        o += 32;
      }
    }
  }
  // checked rounds: the tail of a part, or every round when the tables stay in global memory
  for (; r0 < rounds; ++r0) {
    const int index = load_idx(r0 * 32 + lane);
    const bool active = index >= 0;
    int phase = 0, maxv = 0, offset = 0, value = 0;
    if (active) {
      const uint32_t cum = x & 0xffffu;
      uint32_t start, next;
      int sidx;
      if (TAB_SMEM) {
        const uint32_t row = lds32(s_row + 4u * (uint32_t)index);
        const uint32_t c = s_cdf + 4u * (uint32_t)index * stride;
        maxv = (int)(row >> 16) - 2;
        offset = (int)(int16_t)(row & 0xffffu);
        int lo = 0, hi = maxv;
        while (lo < hi) {
          const int mid = (lo + hi + 1) >> 1;
          if (lds32(c + 4u * mid) <= cum) lo = mid; else hi = mid - 1;
        }
        sidx = lo;
        start = lds32(c + 4u * sidx);
        next = lds32(c + 4u * sidx + 4u);
      } else {
        const int32_t *cdf = cr.t.cdf + (int64_t)index * cr.t.stride;
        const int size = cr.t.sizes[index];
        maxv = size - 2;
        offset = cr.t.offsets[index];
        sidx = rans::find_symbol(cdf, size, cum);
        start = (uint32_t)cdf[sidx];
        next = (uint32_t)cdf[sidx + 1];
      }
      x = (next - start) * (x >> 16) + cum - start;
      value = sidx;
      if (sidx == maxv) phase = 1;
    }
    const bool need = active && x < rans::L32;
    const uint32_t m = __ballot_sync(0xffffffffu, need);
    if (m) refill(need, m);
    if (__any_sync(0xffffffffu, phase != 0)) escape(phase, value, maxv);
    if (active) *o = (int16_t)(value + offset);
    o += 32;
  }
  p.lane_x[lane] = x;
  __syncwarp();
  if (lane == 0) p.rp = rp;
  if (err) atomicOr(meta, err);
}

// ------------------------------------------------------------------ container assembly
__global__ void assemble_kernel(int mode, int parts, const uint8_t *buf, uint32_t *meta,
                                uint8_t *out, uint32_t out_cap) {
  __shared__ uint32_t dst_off[MAX_PARTS + 1];
  __shared__ uint32_t s_per, s_overhead;
  __shared__ int s_fits;
  const bool ext = mode != DCVC_RANS_COMPAT && parts > 16;
  if (threadIdx.x == 0) {
    uint32_t maximum = 0, total = 0;
    for (int i = 0; i < parts; ++i) {
      const uint32_t len = meta[2 + i];
      if (i < parts - 1 && len > maximum) maximum = len;
      dst_off[i] = total;
      total += len;
    }
    dst_off[parts] = total;
    s_per = (mode == DCVC_RANS_COMPAT && maximum <= 65535u) ? 2 : 4;
    s_overhead = (ext ? 3 : 1) + (uint32_t)(parts - 1) * s_per;
    s_fits = total + s_overhead <= out_cap;
    if (blockIdx.x == 0) {
      meta[1] = total + s_overhead;
      if (!s_fits) atomicOr(meta, ERR_OUT_CAPACITY);
    }
  }
  __syncthreads();
  if (!s_fits) return;
  const uint32_t per = s_per, overhead = s_overhead;
  if (blockIdx.x == 0) {
    const uint32_t hdr = ext ? 3 : 1;
    if (threadIdx.x == 0) {
      if (ext) {
        out[0] = 0x0E;
        out[1] = (uint8_t)((parts - 1) & 0xff);
        out[2] = (uint8_t)((parts - 1) >> 8);
      } else {
        out[0] = (uint8_t)(((parts - 1) << 4) | (mode == DCVC_RANS_COMPAT ? (per == 2 ? 1 : 0) : 0xF));
      }
    }
    for (int i = threadIdx.x; i < parts - 1; i += blockDim.x) {
      const uint32_t len = dst_off[i + 1] - dst_off[i];
      for (uint32_t b = 0; b < per; ++b) out[hdr + i * per + b] = (uint8_t)(len >> (8 * b));
    }
  }
  // payload copy: block b takes slice b / parts of part b % parts (all blocks busy for few parts);
  // with more parts than blocks every block walks whole parts
  const int nslices = (int)gridDim.x / parts > 0 ? (int)gridDim.x / parts : 1;
  for (int job = blockIdx.x; job < parts * nslices; job += gridDim.x) {
    const int i = job % parts, sl = job / parts;
    const uint32_t len = dst_off[i + 1] - dst_off[i];
    const uint32_t b0 = (uint32_t)((uint64_t)len * sl / nslices), b1 = (uint32_t)((uint64_t)len * (sl + 1) / nslices);
    const uint8_t *src = buf + meta[2 + MAX_PARTS + i];
    uint8_t *dst = out + overhead + dst_off[i];
    for (uint32_t b = b0 + threadIdx.x; b < b1; b += blockDim.x) dst[b] = src[b];
  }
}

}  // namespace

// ------------------------------------------------------------------ host context
struct dcvc_rans {
  int mode = 0, parts = 1;
  bool bare = false;     // DCVC_RANS_COMPAT_BARE: compat, one part, no flag byte on either side
  int64_t max_symbols = 0;
  EncJob job;
  // device memory: slices of the caller's workspace (dcvc_rans_workspace_bytes)
  uint32_t *d_buf = nullptr;
  uint2 *d_ops = nullptr;          // interleaved: prepared operations, one per symbol
  size_t buf_words = 0;
  uint8_t *d_out = nullptr;
  size_t out_cap = 0;
  uint32_t *d_meta = nullptr;
  uint8_t *d_stream = nullptr;
  size_t stream_cap = 0;
  DecPart *d_parts = nullptr;
  // pinned host staging (host memory, owned by the context)
  uint32_t *h_meta = nullptr, *h_dec_flag = nullptr;
  uint8_t *h_stage = nullptr;
  DecPart *h_parts = nullptr;
  int dec_parts = 0;
  cudaEvent_t staged = nullptr;
  bool stage_pending = false;
  cudaEvent_t flushed = nullptr;   // encoder: after the meta words of flush_begin reached the host
  cudaStream_t side = nullptr;     // encoder: payload copy of flush_end, not behind later work
  bool flush_pending = false;
  cudaEvent_t dec_done = nullptr;  // decoder: after the error word of dec_finish reached the host
  bool dec_check_pending = false;
};

namespace {
struct WsLayout {
  size_t buf_words, out_cap, stream_cap;
  size_t o_buf, o_ops, o_out, o_meta, o_stream, o_parts, total;
};
inline size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }
WsLayout ws_layout(int mode, int parts, int64_t max_symbols) {
  WsLayout w;
  w.buf_words = (size_t)max_symbols + (size_t)SLACK_WORDS * parts;
  w.out_cap = w.buf_words * 4 + 256 + 4 * (size_t)parts;
  w.stream_cap = w.out_cap + 16 * (size_t)parts;
  size_t o = 0;
  w.o_buf = o; o = al256(o + w.buf_words * 4);
  // prepared operations, one per symbol: 8 bytes (interleaved) or 16 bytes (compat)
  w.o_ops = o; o = al256(o + (size_t)max_symbols * (mode == DCVC_RANS_INTERLEAVED ? sizeof(uint2) : sizeof(uint4)));
  w.o_out = o; o = al256(o + w.out_cap);
  w.o_meta = o; o = al256(o + META_WORDS * 4);
  w.o_stream = o; o = al256(o + w.stream_cap);
  w.o_parts = o; o = al256(o + sizeof(DecPart) * MAX_PARTS);
  w.total = o;
  return w;
}
bool create_args_ok(int mode, int parts, int64_t max_symbols) {
  if (mode == DCVC_RANS_COMPAT_BARE) return parts == 1 && max_symbols >= 1;
  if (mode != DCVC_RANS_COMPAT && mode != DCVC_RANS_INTERLEAVED) return false;
  if (parts < 1 || parts > (mode == DCVC_RANS_COMPAT ? MAX_COMPAT_PARTS : MAX_PARTS)) return false;
  return max_symbols >= 1;
}
}  // namespace

extern "C" void dcvc_rans_destroy(dcvc_rans *r) {
  if (!r) return;
  cudaFreeHost(r->h_meta); cudaFreeHost(r->h_dec_flag); cudaFreeHost(r->h_stage); cudaFreeHost(r->h_parts);
  if (r->staged) cudaEventDestroy(r->staged);
  if (r->flushed) cudaEventDestroy(r->flushed);
  if (r->dec_done) cudaEventDestroy(r->dec_done);
  if (r->side) cudaStreamDestroy(r->side);
  delete r;
}

#define RANS_TRY(expr)                                                                     \
  do {                                                                                     \
    cudaError_t e_ = (expr);                                                               \
    if (e_ != cudaSuccess) {                                                               \
      dcvc_fail(DCVC_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #expr,                  \
                cudaGetErrorString(e_));                                                   \
      dcvc_rans_destroy(r);                                                                \
      return DCVC_ERR_CUDA;                                                                \
    }                                                                                      \
  } while (0)

extern "C" size_t dcvc_rans_workspace_bytes(int mode, int parts, int64_t max_symbols) {
  if (!create_args_ok(mode, parts, max_symbols)) return 0;
  return ws_layout(mode == DCVC_RANS_COMPAT_BARE ? DCVC_RANS_COMPAT : mode, parts, max_symbols).total;
}

extern "C" int dcvc_rans_create(dcvc_rans **out, int mode, int parts, int64_t max_symbols,
                                void *workspace, size_t workspace_bytes, void *stream) {
  if (!out || !create_args_ok(mode, parts, max_symbols))
    return dcvc_fail(DCVC_ERR_ARG, "rans_create: bad arguments (mode %d parts %d; compat takes 1..%d parts, "
                     "interleaved 1..%d warps)", mode, parts, MAX_COMPAT_PARTS, MAX_PARTS);
  const bool bare = mode == DCVC_RANS_COMPAT_BARE;
  if (bare) mode = DCVC_RANS_COMPAT;
  const WsLayout w = ws_layout(mode, parts, max_symbols);
  if (!workspace || workspace_bytes < w.total || (reinterpret_cast<uintptr_t>(workspace) & 255))
    return dcvc_fail(DCVC_ERR_ARG, "rans_create: workspace of %zu bytes (256-byte aligned device memory) needed, got %zu",
                     w.total, workspace_bytes);
  dcvc_rans *r = new (std::nothrow) dcvc_rans();
  if (!r) return dcvc_fail(DCVC_ERR_ARG, "rans_create: out of host memory");
  r->mode = mode; r->parts = parts; r->max_symbols = max_symbols; r->bare = bare;
  r->job.ncalls = 0; r->job.parts = parts;
  uint8_t *base = static_cast<uint8_t *>(workspace);
  r->buf_words = w.buf_words; r->out_cap = w.out_cap; r->stream_cap = w.stream_cap;
  r->d_buf = reinterpret_cast<uint32_t *>(base + w.o_buf);
  r->d_ops = reinterpret_cast<uint2 *>(base + w.o_ops);
  r->d_out = base + w.o_out;
  r->d_meta = reinterpret_cast<uint32_t *>(base + w.o_meta);
  r->d_stream = base + w.o_stream;
  r->d_parts = reinterpret_cast<DecPart *>(base + w.o_parts);
  RANS_TRY(cudaMemsetAsync(r->d_meta, 0, META_WORDS * 4, (cudaStream_t)stream));
  RANS_TRY(cudaMallocHost(&r->h_meta, META_WORDS * 4));
  RANS_TRY(cudaMallocHost(&r->h_dec_flag, 16));
  RANS_TRY(cudaMallocHost(&r->h_stage, r->stream_cap));
  RANS_TRY(cudaMallocHost(&r->h_parts, sizeof(DecPart) * MAX_PARTS));
  RANS_TRY(cudaEventCreateWithFlags(&r->staged, cudaEventDisableTiming));
  RANS_TRY(cudaEventCreateWithFlags(&r->flushed, cudaEventDisableTiming));
  RANS_TRY(cudaEventCreateWithFlags(&r->dec_done, cudaEventDisableTiming));
  RANS_TRY(cudaStreamCreateWithFlags(&r->side, cudaStreamNonBlocking));
  *out = r;
  return DCVC_OK;
}

extern "C" int dcvc_rans_enc_reset(dcvc_rans *r) {
  if (!r) return dcvc_fail(DCVC_ERR_ARG, "rans: null context");
  r->job.ncalls = 0;
  return DCVC_OK;
}

extern "C" int dcvc_rans_enc_push(dcvc_rans *r, const int16_t *symbols, const uint8_t *indexes,
                                  int64_t n, int64_t plane, const dcvc_tables *t) {
  if (!r || (!symbols && n > 0) || !t || !t->cdf || !t->sizes || !t->offsets || n < 0)
    return dcvc_fail(DCVC_ERR_ARG, "rans_enc_push: bad arguments");
  if (n == 0) return DCVC_OK;  // an empty call codes nothing (the reference pushes no symbols)
  if (!indexes && plane < 1) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_push: plane needed without indexes");
  if (r->job.ncalls >= DCVC_RANS_MAX_CALLS) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_push: too many calls");
  int64_t total = n;
  for (int c = 0; c < r->job.ncalls; ++c) total += r->job.calls[c].n;
  if (total > r->max_symbols) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_push: more than max_symbols");
  CallRec &cr = r->job.calls[r->job.ncalls++];
  cr.sym = symbols; cr.idx = indexes; cr.n = n; cr.plane = plane < 1 ? 1 : plane;
  cr.t.cdf = t->cdf; cr.t.sizes = t->sizes; cr.t.offsets = t->offsets; cr.t.stride = t->stride;
  cr.t.count = t->count;
  return DCVC_OK;
}

static int rans_errors(uint32_t flags) {
  if (!flags) return DCVC_OK;
  if (flags & ERR_OUT_CAPACITY) return dcvc_fail(DCVC_ERR_ARG, "rans: output capacity too small");
  if (flags & ERR_ENC_OVERFLOW) return dcvc_fail(DCVC_ERR_ARG, "rans: encode buffer overflow");
  return dcvc_fail(DCVC_ERR_STREAM, "rans: truncated or corrupt stream (flags %u)", flags);
}

// flush in two halves: `begin` enqueues the coding kernels and the copy of the meta words on
// `stream` and records an event; `end` waits for THAT EVENT ONLY, then fetches the payload on a
// side stream.  Work the caller enqueues on `stream` between the two (the encoder's own
// reconstruction of the frame) keeps running while the host already holds the bytes.
extern "C" int dcvc_rans_enc_flush_begin(dcvc_rans *r, size_t out_capacity, void *stream) {
  int rc = dcvc_rans_enc_flush_enqueue(r, out_capacity, stream);
  if (rc) return rc;
  return dcvc_rans_enc_flush_mark(r, stream);
}

// flush_begin in two steps for callers that record the coder's work into a CUDA graph: `enqueue`
// is pure stream work (memset, kernels, the copy of the meta words to pinned host memory) and may
// be captured; `mark` records the event flush_end waits on and must run outside the capture, after
// the graph launch, on every replay.
extern "C" int dcvc_rans_enc_flush_mark(dcvc_rans *r, void *stream) {
  if (!r) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_flush_mark: bad arguments");
  if (r->flush_pending) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_flush_mark: a flush is already pending");
  DCVC_CUDA(cudaEventRecord(r->flushed, (cudaStream_t)stream));
  r->flush_pending = true;
  return DCVC_OK;
}

// `mark` for a stream that is being captured into a CUDA graph: the event becomes an event-record
// NODE of the graph (cudaEventRecordExternal), so that the host can wait on it after every launch
// of the graph while the graph's later nodes still run; on replays the host side only has to
// re-arm the pending flag (dcvc_rans_enc_flush_set_pending).
extern "C" int dcvc_rans_enc_flush_mark_captured(dcvc_rans *r, void *stream) {
  if (!r) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_flush_mark_captured: bad arguments");
  DCVC_CUDA(cudaEventRecordWithFlags(r->flushed, (cudaStream_t)stream, cudaEventRecordExternal));
  return DCVC_OK;
}

extern "C" int dcvc_rans_enc_flush_set_pending(dcvc_rans *r) {
  if (!r) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_flush_set_pending: bad arguments");
  if (r->flush_pending) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_flush_set_pending: a flush is already pending");
  r->flush_pending = true;
  return DCVC_OK;
}

extern "C" int dcvc_rans_enc_flush_enqueue(dcvc_rans *r, size_t out_capacity, void *stream) {
  if (!r) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_flush_enqueue: bad arguments");
  if (r->flush_pending) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_flush_enqueue: a flush is already pending");
  cudaStream_t s = (cudaStream_t)stream;
  r->job.each_sum = 0;
  r->job.total = 0;
  for (int c = 0; c < r->job.ncalls; ++c) {
    r->job.each_sum += r->job.calls[c].n / r->parts;
    r->job.total += r->job.calls[c].n;
  }
  if ((size_t)reg_off(r->job, r->parts) > r->buf_words)
    return dcvc_fail(DCVC_ERR_ARG, "rans_enc_flush: buffer too small");
  DCVC_CUDA(cudaMemsetAsync(r->d_meta, 0, META_WORDS * 4, s));
  int64_t total = 0;
  for (int c = 0; c < r->job.ncalls; ++c) { r->job.op_off[c] = total; total += r->job.calls[c].n; }
  r->job.op_off[r->job.ncalls] = total;
  const int pblocks = (int)((total + 255) / 256 < 2048 ? (total + 255) / 256 : 2048);
  if (r->mode == DCVC_RANS_COMPAT) {
    if (total > 0) {
      compat_prepare_kernel<<<pblocks, 256, 0, s>>>(r->job, reinterpret_cast<uint4 *>(r->d_ops));
      DCVC_LAUNCHED();
    }
    compat_encode_kernel<<<r->parts, 32, 0, s>>>(r->job, reinterpret_cast<const uint4 *>(r->d_ops),
                                                 r->d_buf, r->d_meta);
  } else {
    if (total > 0) {
      ilv_prepare_kernel<<<pblocks, 256, 0, s>>>(r->job, r->d_ops);
      DCVC_LAUNCHED();
    }
    ilv_encode_kernel<<<r->parts, 32, 0, s>>>(r->job, r->d_ops, r->d_buf, r->d_meta);
  }
  DCVC_LAUNCHED();
  const uint32_t cap = (uint32_t)(out_capacity < r->out_cap ? out_capacity : r->out_cap);
  assemble_kernel<<<64, 256, 0, s>>>(r->mode, r->parts, reinterpret_cast<const uint8_t *>(r->d_buf),
                                     r->d_meta, r->d_out, cap);
  DCVC_LAUNCHED();
  DCVC_CUDA(cudaMemcpyAsync(r->h_meta, r->d_meta, 8, cudaMemcpyDeviceToHost, s));
  r->job.ncalls = 0;
  return DCVC_OK;
}

extern "C" int dcvc_rans_enc_flush_end(dcvc_rans *r, uint8_t *out_host, size_t out_capacity,
                                       size_t *out_len) {
  if (!r || !out_host || !out_len) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_flush_end: bad arguments");
  if (!r->flush_pending) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_flush_end: no flush pending");
  r->flush_pending = false;
  DCVC_CUDA(cudaEventSynchronize(r->flushed));
  int rc = rans_errors(r->h_meta[0]);
  if (rc) return rc;
  const size_t skip = r->bare ? 1 : 0;     // the bare framing drops the (constant) flag byte
  const size_t len = r->h_meta[1] - skip;
  if (len > out_capacity) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_flush_end: output buffer too small");
  DCVC_CUDA(cudaMemcpyAsync(out_host, r->d_out + skip, len, cudaMemcpyDeviceToHost, r->side));
  DCVC_CUDA(cudaStreamSynchronize(r->side));
  *out_len = len;
  return DCVC_OK;
}

extern "C" int dcvc_rans_enc_flush(dcvc_rans *r, uint8_t *out_host, size_t out_capacity,
                                   size_t *out_len, void *stream) {
  if (!r || !out_host || !out_len) return dcvc_fail(DCVC_ERR_ARG, "rans_enc_flush: bad arguments");
  int rc = dcvc_rans_enc_flush_begin(r, out_capacity, stream);
  if (rc) return rc;
  return dcvc_rans_enc_flush_end(r, out_host, out_capacity, out_len);
}

extern "C" int dcvc_rans_dec_set_stream(dcvc_rans *r, const uint8_t *host, size_t len, void *stream) {
  if (!r || !host || len < 1) return dcvc_fail(DCVC_ERR_STREAM, "rans_dec_set_stream: empty stream");
  cudaStream_t s = (cudaStream_t)stream;
  if (r->dec_check_pending) {      // the previous frame's error word (dcvc_rans_dec_finish)
    int rc = dcvc_rans_dec_check(r);
    if (rc) return rc;
  }
  const uint8_t flag = r->bare ? 0x01 : host[0];     // bare: one part, nothing precedes the words
  int n = (flag >> 4) + 1;
  const int low = flag & 0x0f;
  size_t hdr = r->bare ? 0 : 1;
  if (!r->bare && flag == 0x0E) {  // extended interleaved header: warp count in two bytes
    if (len < 3) return dcvc_fail(DCVC_ERR_STREAM, "rans_dec_set_stream: truncated header");
    n = ((int)host[1] | ((int)host[2] << 8)) + 1;
    hdr = 3;
    if (n <= 16 || n > MAX_PARTS) return dcvc_fail(DCVC_ERR_STREAM, "rans_dec_set_stream: bad warp count %d", n);
  }
  const int mode = (!r->bare && (low == 0xF || flag == 0x0E)) ? DCVC_RANS_INTERLEAVED : DCVC_RANS_COMPAT;
  if (mode != r->mode) return dcvc_fail(DCVC_ERR_STREAM, "rans_dec_set_stream: container mode %d, coder mode %d", mode, r->mode);
  if (mode == DCVC_RANS_COMPAT && low > 1) return dcvc_fail(DCVC_ERR_STREAM, "rans_dec_set_stream: bad flag byte");
  const int per = (mode == DCVC_RANS_COMPAT && low == 1) ? 2 : 4;
  size_t off = hdr, total = 0;
  size_t sizes[MAX_PARTS];
  for (int i = 0; i < n - 1; ++i) {
    if (off + per > len) return dcvc_fail(DCVC_ERR_STREAM, "rans_dec_set_stream: truncated header");
    size_t v = 0;
    for (int b = 0; b < per; ++b) v |= (size_t)host[off + b] << (8 * b);
    sizes[i] = v; total += v; off += per;
  }
  if (off + total > len) return dcvc_fail(DCVC_ERR_STREAM, "rans_dec_set_stream: part sizes exceed stream");
  sizes[n - 1] = len - off - total;
  if (len + 16 * (size_t)n > r->stream_cap) return dcvc_fail(DCVC_ERR_ARG, "rans_dec_set_stream: stream larger than coder capacity");
  if (r->stage_pending) { DCVC_CUDA(cudaEventSynchronize(r->staged)); r->stage_pending = false; }
  size_t dst = 0;
  const size_t unit = mode == DCVC_RANS_COMPAT ? 4 : 2;
  for (int i = 0; i < n; ++i) {
    memcpy(r->h_stage + dst, host + off, sizes[i]);
    memset(&r->h_parts[i], 0, sizeof(DecPart));
    r->h_parts[i].base = (uint32_t)dst;
    r->h_parts[i].end = (uint32_t)(sizes[i] / unit);
    off += sizes[i];
    dst = (dst + sizes[i] + 15) & ~(size_t)15;
  }
  r->dec_parts = n;
  DCVC_CUDA(cudaMemcpyAsync(r->d_stream, r->h_stage, dst ? dst : 16, cudaMemcpyHostToDevice, s));
  DCVC_CUDA(cudaMemcpyAsync(r->d_parts, r->h_parts, sizeof(DecPart) * n, cudaMemcpyHostToDevice, s));
  DCVC_CUDA(cudaEventRecord(r->staged, s));
  r->stage_pending = true;
  DCVC_CUDA(cudaMemsetAsync(r->d_meta, 0, 4, s));
  if (mode == DCVC_RANS_COMPAT)
    compat_dec_init_kernel<<<n, 32, 0, s>>>(r->d_stream, r->d_parts, r->d_meta);
  else
    ilv_dec_init_kernel<<<n, 32, 0, s>>>(r->d_stream, r->d_parts, r->d_meta);
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" int dcvc_rans_dec_pull(dcvc_rans *r, const uint8_t *indexes, int64_t n, int64_t plane,
                                  const dcvc_tables *t, int16_t *symbols_out, void *stream) {
  if (!r || !t || !t->cdf || (!symbols_out && n > 0) || n < 0 || r->dec_parts < 1)
    return dcvc_fail(DCVC_ERR_ARG, "rans_dec_pull: bad arguments or no stream set");
  if (n == 0) return DCVC_OK;
  if (!indexes && plane < 1) return dcvc_fail(DCVC_ERR_ARG, "rans_dec_pull: plane needed without indexes");
  CallRec cr;
  cr.sym = nullptr; cr.idx = indexes; cr.n = n; cr.plane = plane < 1 ? 1 : plane;
  cr.t.cdf = t->cdf; cr.t.sizes = t->sizes; cr.t.offsets = t->offsets; cr.t.stride = t->stride;
  cr.t.count = t->count; cr.t.lut = t->lut;
  if (n / r->dec_parts + r->dec_parts >= ((int64_t)1 << 31))
    return dcvc_fail(DCVC_ERR_ARG, "rans_dec_pull: more than 2^31 symbols per stream part");
  cudaStream_t s = (cudaStream_t)stream;
  // decoder's shared-memory image: cdf rows, one {size, offset} word per row [, the search LUT]
  const size_t base_bytes = dec_tab_words(cr.t.count, cr.t.stride) * 4;
  const size_t lut_bytes = (size_t)cr.t.count * LUT_PITCH;
  const bool smem_ok = cr.t.count > 0 && base_bytes <= DEC_TAB_SMEM_MAX && cr.t.stride < 32768;
  // the LUT costs 256 bytes per row to stage: only for parts that decode enough symbols
  const bool lut_ok = t->lut && n / r->dec_parts >= 4096 && cr.t.stride <= 257 &&
                      (reinterpret_cast<uintptr_t>(t->lut) & 3) == 0 &&
                      base_bytes + lut_bytes <= DEC_TAB_SMEM_MAX;
  if (r->mode == DCVC_RANS_COMPAT) {
    if (smem_ok) {
      int rc = dcvc_once_per_device(DCVC_ONCE_RANS_DEC2, [] {
        cudaError_t e = cudaFuncSetAttribute(compat_decode_fast_kernel<true>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DEC_TAB_SMEM_MAX);
        if (e != cudaSuccess) return e;
        return cudaFuncSetAttribute(compat_decode_fast_kernel<false>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DEC_TAB_SMEM_MAX);
      });
      if (rc) return rc;
      if (lut_ok)
        compat_decode_fast_kernel<true><<<r->dec_parts, CDEC_THREADS, base_bytes + lut_bytes, s>>>(
            cr, r->dec_parts, r->d_stream, r->d_parts, symbols_out, r->d_meta);
      else
        compat_decode_fast_kernel<false><<<r->dec_parts, CDEC_THREADS, base_bytes, s>>>(
            cr, r->dec_parts, r->d_stream, r->d_parts, symbols_out, r->d_meta);
    } else {
      compat_decode_kernel<false><<<r->dec_parts, CDEC_THREADS, 0, s>>>(
          cr, r->dec_parts, r->d_stream, r->d_parts, symbols_out, r->d_meta);
    }
  } else {
    if (smem_ok) {
      int rc = dcvc_once_per_device(DCVC_ONCE_RANS_DEC, [] {
        cudaError_t e = cudaFuncSetAttribute(ilv_decode_kernel<true, true>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DEC_TAB_SMEM_MAX);
        if (e != cudaSuccess) return e;
        return cudaFuncSetAttribute(ilv_decode_kernel<true, false>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DEC_TAB_SMEM_MAX);
      });
      if (rc) return rc;
      if (lut_ok)
        ilv_decode_kernel<true, true><<<r->dec_parts, DEC_THREADS, base_bytes + lut_bytes, s>>>(
            cr, r->dec_parts, r->d_stream, r->d_parts, symbols_out, r->d_meta);
      else
        ilv_decode_kernel<true, false><<<r->dec_parts, DEC_THREADS, base_bytes, s>>>(
            cr, r->dec_parts, r->d_stream, r->d_parts, symbols_out, r->d_meta);
    } else {
      ilv_decode_kernel<false, false><<<r->dec_parts, DEC_THREADS, 0, s>>>(
          cr, r->dec_parts, r->d_stream, r->d_parts, symbols_out, r->d_meta);
    }
  }
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" int dcvc_rans_dec_finish(dcvc_rans *r, void *stream) {
  if (!r) return dcvc_fail(DCVC_ERR_ARG, "rans: null context");
  cudaStream_t s = (cudaStream_t)stream;
  if (r->dec_check_pending) {
    int rc = dcvc_rans_dec_check(r);
    if (rc) return rc;
  }
  DCVC_CUDA(cudaMemcpyAsync(r->h_dec_flag, r->d_meta, 4, cudaMemcpyDeviceToHost, s));
  DCVC_CUDA(cudaEventRecord(r->dec_done, s));
  r->dec_check_pending = true;
  return DCVC_OK;
}

extern "C" int dcvc_rans_dec_check(dcvc_rans *r) {
  if (!r) return dcvc_fail(DCVC_ERR_ARG, "rans: null context");
  if (!r->dec_check_pending) return DCVC_OK;
  r->dec_check_pending = false;
  DCVC_CUDA(cudaEventSynchronize(r->dec_done));
  return rans_errors(r->h_dec_flag[0]);
}

extern "C" int dcvc_rans_status(dcvc_rans *r, void *stream) {
  if (!r) return dcvc_fail(DCVC_ERR_ARG, "rans: null context");
  cudaStream_t s = (cudaStream_t)stream;
  DCVC_CUDA(cudaMemcpyAsync(r->h_meta, r->d_meta, 4, cudaMemcpyDeviceToHost, s));
  DCVC_CUDA(cudaStreamSynchronize(s));
  return rans_errors(r->h_meta[0]);
}

// ------------------------------------------------------------------ CDF quantiser (host)
// Semantics of ops.cpp:24-82: scale each probability by 2^precision with round-half-away,
// renormalise the integer frequencies so they sum to 2^precision (truncating division), take the
// running sum, pin the last entry to 2^precision, and give every empty bin one count taken from
// the bin with the smallest frequency above one, shifting the entries in between.
extern "C" int dcvc_pmf_to_quantized_cdf(const float *pmf, int n, int precision, uint32_t *cdf) {
  if (!pmf || !cdf || n < 1 || precision < 1 || precision > 16)
    return dcvc_fail(DCVC_ERR_ARG, "pmf_to_quantized_cdf: bad arguments");
  const uint32_t top = 1u << precision;
  std::vector<uint32_t> f((size_t)n + 1, 0u);
  for (int i = 0; i < n; ++i) {
    const float scaled = roundf(pmf[i] * (float)top);
    f[i + 1] = (uint32_t)((double)scaled + 0.5);
  }
  uint32_t sum = 0;
  for (uint32_t v : f) sum += v;
  if (sum == 0) return dcvc_fail(DCVC_ERR_ARG, "pmf_to_quantized_cdf: all-zero pmf");
  uint32_t run = 0;
  for (int i = 0; i <= n; ++i) {
    run += (uint32_t)(((uint64_t)top * f[i]) / sum);
    cdf[i] = run;
  }
  cdf[n] = top;
  for (int i = 0; i < n; ++i) {
    if (cdf[i] != cdf[i + 1]) continue;
    int donor = -1;
    uint32_t donor_freq = 0xffffffffu;
    for (int j = 0; j < n; ++j) {
      const uint32_t fr = cdf[j + 1] - cdf[j];
      if (fr > 1 && fr < donor_freq) { donor_freq = fr; donor = j; }
    }
    if (donor < 0) return dcvc_fail(DCVC_ERR_ARG, "pmf_to_quantized_cdf: no bin to steal from");
    if (donor < i) {
      for (int j = donor + 1; j <= i; ++j) cdf[j] -= 1;
    } else {
      for (int j = i + 1; j <= donor; ++j) cdf[j] += 1;
    }
  }
  return DCVC_OK;
}
