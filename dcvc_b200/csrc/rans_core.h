// rans_core.h -- coder arithmetic shared by the device kernels in rans.cu.
//
// Two coders live here:
//  (1) "compat": the reference's 64-bit rANS with 32-bit words, 16-bit precision, 4-bit bypass
//      escapes and LIFO symbol order, byte-identical to DCVC-DC/src/cpp/rans/rans.cpp:37-168,
//      265-331 (rANS64 primitives: ryg_rans @ c9d162d, restated from their definition: state in
//      [2^31, 2^63), word stream written backwards, read forwards).  One sequential state per
//      stream part, so one device thread per part.
//  (2) "interleaved": 32 lanes of 32-bit rANS (state in [2^16, 2^32), 16-bit words, same 16-bit
//      precision, same escape scheme) sharing one word stream per warp; see rans.cu.
//
// Everything is DCVC_HD so that tests/ can compile this header for the host (g++) and check the
// very same arithmetic against the oracle without a GPU.  The product only ever runs it on the
// device.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define DCVC_HD __host__ __device__ __forceinline__
#else
#define DCVC_HD static inline
#endif

namespace rans {

constexpr int PRECISION = 16;
constexpr int BYPASS_BITS = 4;
constexpr int MAX_BYPASS = 15;
constexpr uint64_t L64 = 1ull << 31;
constexpr uint32_t L32 = 1u << 16;
constexpr int MAX_OPS = 24;  // main + prefix + nibbles for any int16 symbol (<= 1+2+5)

struct Tables {
  const int32_t *cdf;
  const int32_t *sizes;
  const int32_t *offsets;
  int32_t stride;
  int32_t count;   // rows (table sets of `stride` entries)
  const uint8_t *lut = nullptr;  // optional [count][256] search LUT (dcvc_tables.lut)
};

// One coding operation: bits==16 -> (start,freq) interval; bits==4 -> raw nibble `start`.
struct Op {
  uint16_t start, freq;
  uint8_t bypass;
};

// Expand one symbol into its operations in *forward* (decode) order
// (rans.cpp:76-139).  Returns the number of ops (>= 1), or 0 when index < 0 (skipped).
DCVC_HD int expand_symbol(int32_t symbol, int32_t index, const Tables &t, Op *ops) {
  if (index < 0) return 0;
  const int32_t *cdf = t.cdf + (int64_t)index * t.stride;
  const int32_t max_value = t.sizes[index] - 2;
  int32_t value = symbol - t.offsets[index];
  uint32_t raw = 0;
  if (value < 0) {
    raw = (uint32_t)(-2 * value - 1);
    value = max_value;
  } else if (value >= max_value) {
    raw = (uint32_t)(2 * (value - max_value));
    value = max_value;
  }
  int n = 0;
  ops[n].start = (uint16_t)cdf[value];
  ops[n].freq = (uint16_t)(cdf[value + 1] - cdf[value]);
  ops[n].bypass = 0;
  ++n;
  if (value == max_value) {
    int32_t nb = 0;
    while ((raw >> (nb * BYPASS_BITS)) != 0) ++nb;
    int32_t val = nb;
    while (val >= MAX_BYPASS) {
      ops[n].start = MAX_BYPASS; ops[n].freq = 0; ops[n].bypass = 1; ++n;
      val -= MAX_BYPASS;
    }
    ops[n].start = (uint16_t)val; ops[n].freq = 0; ops[n].bypass = 1; ++n;
    for (int32_t j = 0; j < nb; ++j) {
      ops[n].start = (uint16_t)((raw >> (j * BYPASS_BITS)) & MAX_BYPASS);
      ops[n].freq = 0; ops[n].bypass = 1; ++n;
    }
  }
  return n;
}

// first k with cdf[k] > cum, minus 1 (the reference's linear find_if, rans.cpp:289-292; the
// cdf is increasing so a binary search returns the same k)
DCVC_HD int find_symbol(const int32_t *cdf, int size, uint32_t cum) {
  int lo = 0, hi = size;  // first position with cdf[pos] > cum
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if ((uint32_t)cdf[mid] > cum) hi = mid; else lo = mid + 1;
  }
  return lo - 1;
}

// ------------------------------------------------------------------ compat (64-bit) coder
struct Enc64 {
  uint64_t x;
  uint32_t *ptr;    // next word is written at --ptr
  uint32_t *limit;  // lowest legal address
  int overflow;
};

DCVC_HD void enc64_init(Enc64 &e, uint32_t *end, uint32_t *limit) {
  e.x = L64; e.ptr = end; e.limit = limit; e.overflow = 0;
}
DCVC_HD void enc64_emit(Enc64 &e) {
  if (e.ptr <= e.limit) { e.overflow = 1; } else { *--e.ptr = (uint32_t)e.x; }
  e.x >>= 32;
}
DCVC_HD void enc64_put(Enc64 &e, const Op &op) {
  if (!op.bypass) {
    const uint64_t x_max = ((L64 >> PRECISION) << 32) * (uint64_t)op.freq;
    if (e.x >= x_max) enc64_emit(e);
    e.x = ((e.x / op.freq) << PRECISION) + (e.x % op.freq) + op.start;
  } else {
    const uint64_t x_max = ((L64 >> 16) << 32) * (uint64_t)(1u << (16 - BYPASS_BITS));
    if (e.x >= x_max) enc64_emit(e);
    e.x = (e.x << BYPASS_BITS) | op.start;
  }
}
DCVC_HD void enc64_flush(Enc64 &e) {
  if (e.ptr - 2 < e.limit) { e.overflow = 1; return; }
  e.ptr -= 2;
  e.ptr[0] = (uint32_t)e.x;
  e.ptr[1] = (uint32_t)(e.x >> 32);
}

struct Dec64 {
  uint64_t x;
  const uint32_t *ptr;
  const uint32_t *end;
  int underflow;
};
DCVC_HD uint32_t dec64_word(Dec64 &d) {
  if (d.ptr >= d.end) { d.underflow = 1; return 0; }
  return *d.ptr++;
}
DCVC_HD void dec64_init(Dec64 &d, const uint32_t *p, const uint32_t *end) {
  d.ptr = p; d.end = end; d.underflow = 0;
  const uint64_t lo = dec64_word(d);
  const uint64_t hi = dec64_word(d);
  d.x = lo | (hi << 32);
}
DCVC_HD uint32_t dec64_bits(Dec64 &d, uint32_t nbits) {
  const uint32_t val = (uint32_t)(d.x & ((1u << nbits) - 1));
  d.x >>= nbits;
  if (d.x < L64) d.x = (d.x << 32) | dec64_word(d);
  return val;
}
// rans.cpp:272-331 for one symbol
DCVC_HD int32_t dec64_symbol(Dec64 &d, int32_t index, const Tables &t) {
  if (index < 0) return 0;
  const int32_t *cdf = t.cdf + (int64_t)index * t.stride;
  const int32_t size = t.sizes[index];
  const int32_t max_value = size - 2;
  const uint32_t cum = (uint32_t)(d.x & 0xffffu);
  const int s = find_symbol(cdf, size, cum);
  const uint32_t start = (uint32_t)cdf[s], freq = (uint32_t)(cdf[s + 1] - cdf[s]);
  d.x = (uint64_t)freq * (d.x >> PRECISION) + cum - start;
  if (d.x < L64) d.x = (d.x << 32) | dec64_word(d);
  int32_t value = s;
  if (value == max_value) {
    int32_t val = (int32_t)dec64_bits(d, BYPASS_BITS);
    int32_t nb = val;
    while (val == MAX_BYPASS && !d.underflow) {
      val = (int32_t)dec64_bits(d, BYPASS_BITS);
      nb += val;
    }
    int32_t raw = 0;
    for (int32_t j = 0; j < nb && j < 8; ++j) {
      val = (int32_t)dec64_bits(d, BYPASS_BITS);
      raw |= val << (j * BYPASS_BITS);
    }
    value = raw >> 1;
    if (raw & 1) value = -value - 1; else value += max_value;
  }
  return value + t.offsets[index];
}

// ------------------------------------------------------------------ 32-bit lane arithmetic
// encode side: returns true when a 16-bit word (the low half of x before the call) must be
// emitted first; the caller emits it and passes x >> 16.
DCVC_HD bool enc32_needs_emit(uint32_t x, const Op &op) {
  const uint32_t x_max = op.bypass ? (1u << 28) : ((uint32_t)op.freq << 16);
  return x >= x_max;
}
DCVC_HD uint32_t enc32_update(uint32_t x, const Op &op) {
  if (op.bypass) return (x << BYPASS_BITS) | op.start;
  return ((x / op.freq) << PRECISION) + (x % op.freq) + op.start;
}

}  // namespace rans
