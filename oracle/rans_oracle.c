/* rans_oracle.c -- TEST INFRASTRUCTURE (CPU oracle; never linked into the product).
 *
 * Plain-C restatement of the reference's native entropy-coding path:
 *   - pmf_to_quantized_cdf        DCVC-DC/src/cpp/ops/ops.cpp:24-82
 *   - symbol -> {start,range,bypass} expansion with escape nibbles
 *                                 DCVC-DC/src/cpp/rans/rans.cpp:76-139
 *   - LIFO rANS64 flush           DCVC-DC/src/cpp/rans/rans.cpp:141-168 (+ :37-55)
 *   - sequential decode           DCVC-DC/src/cpp/rans/rans.cpp:265-331 (+ :57-73)
 *   - N-part container            DCVC-DC/src/cpp/py_rans/py_rans.cpp:22-119,133-225
 * rANS64 primitives: ryg_rans @ c9d162d (un-vendored third party; algorithm restated in
 * oracle/third_party/rans64.h, see that header).
 *
 * Pinned against the reference's own compiled C++ (oracle/_ref/MLCodec_rans*.so) by
 * tests/test_oracle_rans.py whenever oracle/_ref exists, and against committed golden
 * byte streams (tests/golden/) everywhere.
 *
 * Deliberate deviations from reference *bugs* (not from its format):
 *   - py_rans.cpp:61-62 sizes the index vector with eachSymbolSize but copies
 *     currSymbolSize elements (heap overflow for the last part when n % parts != 0);
 *     here every part gets correctly sized arrays.  Output bytes are identical.
 *   - rans.cpp:146 allocates nsyms words, which under-runs for < 2 symbols; here the
 *     buffer always has room for the final 2-word flush.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "third_party/rans64.h"

#define PRECISION 16
#define BYPASS_PRECISION 4
#define MAX_BYPASS_VAL ((1 << BYPASS_PRECISION) - 1)
#define MAX_PARTS 16

/* ------------------------------------------------------------------ CDF quantiser */
/* ops.cpp:24-82.  pmf[n] float -> cdf[n+1] uint32, total 1<<precision. Returns 0 ok. */
int oracle_pmf_to_quantized_cdf(const float *pmf, int n, int precision, uint32_t *cdf) {
  int i, j;
  cdf[0] = 0;
  for (i = 0; i < n; ++i) {
    /* static_cast<uint32_t>(std::round(p * (1 << precision)) + 0.5): float product,
       float round, then promoted to double by "+ 0.5", truncated. */
    float scaled = pmf[i] * (float)(1 << precision);
    cdf[i + 1] = (uint32_t)((double)roundf(scaled) + 0.5);
  }
  /* std::accumulate(..., 0): int accumulator, wraps like unsigned for our magnitudes */
  uint32_t total = 0;
  for (i = 0; i <= n; ++i) total += cdf[i];
  if (total == 0) return -1;
  for (i = 0; i <= n; ++i)
    cdf[i] = (uint32_t)((((uint64_t)1 << precision) * (uint64_t)cdf[i]) / total);
  for (i = 1; i <= n; ++i) cdf[i] += cdf[i - 1];
  cdf[n] = 1u << precision;
  for (i = 0; i < n; ++i) {
    if (cdf[i] == cdf[i + 1]) {
      uint32_t best_freq = ~0u;
      int best_steal = -1;
      for (j = 0; j < n; ++j) {
        uint32_t freq = cdf[j + 1] - cdf[j];
        if (freq > 1 && freq < best_freq) {
          best_freq = freq;
          best_steal = j;
        }
      }
      if (best_steal < 0) return -2;
      if (best_steal < i) {
        for (j = best_steal + 1; j <= i; ++j) cdf[j]--;
      } else {
        for (j = i + 1; j <= best_steal; ++j) cdf[j]++;
      }
    }
  }
  return 0;
}

/* ------------------------------------------------------------------ encoder */
typedef struct {
  uint16_t start;
  uint16_t range;
  uint8_t bypass;
} sym_t;

typedef struct {
  sym_t *syms;
  size_t n, cap;
  uint8_t *stream;
  size_t stream_len;
} part_enc_t;

typedef struct {
  int parts;
  part_enc_t p[MAX_PARTS];
  uint8_t *out;
  size_t out_len;
} oracle_enc_t;

static void push_sym(part_enc_t *pe, uint16_t start, uint16_t range, uint8_t bypass) {
  if (pe->n == pe->cap) {
    pe->cap = pe->cap ? pe->cap * 2 : 4096;
    pe->syms = (sym_t *)realloc(pe->syms, pe->cap * sizeof(sym_t));
  }
  pe->syms[pe->n].start = start;
  pe->syms[pe->n].range = range;
  pe->syms[pe->n].bypass = bypass;
  pe->n++;
}

oracle_enc_t *oracle_enc_new(int parts) {
  if (parts < 1 || parts > MAX_PARTS) return NULL;
  oracle_enc_t *e = (oracle_enc_t *)calloc(1, sizeof(oracle_enc_t));
  e->parts = parts;
  return e;
}

void oracle_enc_reset(oracle_enc_t *e) {
  for (int i = 0; i < e->parts; ++i) e->p[i].n = 0;
}

void oracle_enc_free(oracle_enc_t *e) {
  if (!e) return;
  for (int i = 0; i < e->parts; ++i) {
    free(e->p[i].syms);
    free(e->p[i].stream);
  }
  free(e->out);
  free(e);
}

/* rans.cpp:76-139 for one part */
static void expand_symbols(part_enc_t *pe, const int16_t *symbols, const int16_t *indexes,
                           int n, const int32_t *cdfs, int cdf_stride,
                           const int32_t *cdf_sizes, const int32_t *offsets) {
  for (int i = 0; i < n; ++i) {
    const int32_t cdf_idx = indexes[i];
    if (cdf_idx < 0) continue;
    const int32_t *cdf = cdfs + (size_t)cdf_idx * cdf_stride;
    const int32_t max_value = cdf_sizes[cdf_idx] - 2;
    int32_t value = symbols[i] - offsets[cdf_idx];
    uint32_t raw_val = 0;
    if (value < 0) {
      raw_val = (uint32_t)(-2 * value - 1);
      value = max_value;
    } else if (value >= max_value) {
      raw_val = (uint32_t)(2 * (value - max_value));
      value = max_value;
    }
    push_sym(pe, (uint16_t)cdf[value], (uint16_t)(cdf[value + 1] - cdf[value]), 0);
    if (value == max_value) {
      int32_t n_bypass = 0;
      while ((raw_val >> (n_bypass * BYPASS_PRECISION)) != 0) ++n_bypass;
      int32_t val = n_bypass;
      while (val >= MAX_BYPASS_VAL) {
        push_sym(pe, MAX_BYPASS_VAL, MAX_BYPASS_VAL + 1, 1);
        val -= MAX_BYPASS_VAL;
      }
      push_sym(pe, (uint16_t)val, (uint16_t)(val + 1), 1);
      for (int32_t j = 0; j < n_bypass; ++j) {
        const int32_t v1 = (raw_val >> (j * BYPASS_PRECISION)) & MAX_BYPASS_VAL;
        push_sym(pe, (uint16_t)v1, (uint16_t)(v1 + 1), 1);
      }
    }
  }
}

/* py_rans.cpp:22-66: one call's symbols are cut into `parts` contiguous chunks */
void oracle_enc_push(oracle_enc_t *e, const int16_t *symbols, const int16_t *indexes, int n,
                     const int32_t *cdfs, int cdf_stride, const int32_t *cdf_sizes,
                     const int32_t *offsets) {
  const int each = n / e->parts;
  const int last = n - each * (e->parts - 1);
  for (int i = 0; i < e->parts; ++i) {
    const int cur = i < e->parts - 1 ? each : last;
    const int off = i * each;
    expand_symbols(&e->p[i], symbols + off, indexes + off, cur, cdfs, cdf_stride, cdf_sizes,
                   offsets);
  }
}

/* rans.cpp:37-55 */
static void enc_put_bits(Rans64State *r, uint32_t **pptr, uint32_t val, uint32_t nbits) {
  uint64_t x = *r;
  const uint32_t freq = 1u << (16 - nbits);
  const uint64_t x_max = ((RANS64_L >> 16) << 32) * (uint64_t)freq;
  if (x >= x_max) {
    *pptr -= 1;
    **pptr = (uint32_t)x;
    x >>= 32;
  }
  *r = (x << nbits) | val;
}

/* rans.cpp:141-168 for one part */
static void flush_part(part_enc_t *pe) {
  Rans64State rans;
  Rans64EncInit(&rans);
  const size_t words = pe->n + 2;
  uint32_t *buf = (uint32_t *)malloc(words * sizeof(uint32_t));
  uint32_t *ptr = buf + words;
  for (size_t k = pe->n; k-- > 0;) {
    const sym_t s = pe->syms[k];
    if (!s.bypass)
      Rans64EncPut(&rans, &ptr, s.start, s.range, PRECISION);
    else
      enc_put_bits(&rans, &ptr, s.start, BYPASS_PRECISION);
  }
  pe->n = 0;
  Rans64EncFlush(&rans, &ptr);
  const size_t nbytes = (size_t)((buf + words) - ptr) * sizeof(uint32_t);
  free(pe->stream);
  pe->stream = (uint8_t *)malloc(nbytes ? nbytes : 1);
  memcpy(pe->stream, ptr, nbytes);
  pe->stream_len = nbytes;
  free(buf);
}

/* flush every part, then assemble the container (py_rans.cpp:74-119).
 * Returns the container length; bytes via oracle_enc_stream(). */
long oracle_enc_flush(oracle_enc_t *e) {
  size_t maximum = 0, total = 0;
  for (int i = 0; i < e->parts; ++i) {
    flush_part(&e->p[i]);
    if (i < e->parts - 1 && e->p[i].stream_len > maximum) maximum = e->p[i].stream_len;
    total += e->p[i].stream_len;
  }
  const int per = maximum > 65535 ? 4 : 2;
  size_t overhead = 1;
  if (e->parts > 1) overhead += (size_t)(e->parts - 1) * per;
  free(e->out);
  e->out = (uint8_t *)malloc(total + overhead);
  e->out_len = total + overhead;
  e->out[0] = (uint8_t)(((e->parts - 1) << 4) + (per == 2 ? 1 : 0));
  for (int i = 0; i < e->parts - 1; ++i) {
    if (per == 2) {
      uint16_t s = (uint16_t)e->p[i].stream_len;
      memcpy(e->out + 1 + 2 * i, &s, 2);
    } else {
      uint32_t s = (uint32_t)e->p[i].stream_len;
      memcpy(e->out + 1 + 4 * i, &s, 4);
    }
  }
  size_t off = overhead;
  for (int i = 0; i < e->parts; ++i) {
    memcpy(e->out + off, e->p[i].stream, e->p[i].stream_len);
    off += e->p[i].stream_len;
  }
  return (long)e->out_len;
}

const uint8_t *oracle_enc_stream(const oracle_enc_t *e) { return e->out; }

/* ------------------------------------------------------------------ decoder */
typedef struct {
  Rans64State rans;
  uint32_t *ptr;
  uint32_t *buf; /* owned, word-aligned copy (+ zero padding) */
} part_dec_t;

typedef struct {
  int parts;
  part_dec_t p[MAX_PARTS];
} oracle_dec_t;

oracle_dec_t *oracle_dec_new(void) { return (oracle_dec_t *)calloc(1, sizeof(oracle_dec_t)); }

void oracle_dec_free(oracle_dec_t *d) {
  if (!d) return;
  for (int i = 0; i < MAX_PARTS; ++i) free(d->p[i].buf);
  free(d);
}

/* py_rans.cpp:133-166 + rans.cpp:265-270.  Returns number of parts, <0 on error. */
int oracle_dec_set_stream(oracle_dec_t *d, const uint8_t *enc, long len) {
  if (len < 1) return -1;
  const uint8_t flag = enc[0];
  const int n = (flag >> 4) + 1;
  const int per = (flag & 0x0f) == 1 ? 2 : 4;
  uint32_t sizes[MAX_PARTS];
  long off = 1, total = 0;
  for (int i = 0; i < n - 1; ++i) {
    if (off + per > len) return -2;
    if (per == 2) {
      uint16_t s;
      memcpy(&s, enc + off, 2);
      sizes[i] = s;
    } else {
      uint32_t s;
      memcpy(&s, enc + off, 4);
      sizes[i] = s;
    }
    off += per;
    total += sizes[i];
  }
  if (len - off - total < 0) return -3;
  sizes[n - 1] = (uint32_t)(len - off - total);
  d->parts = n;
  for (int i = 0; i < n; ++i) {
    free(d->p[i].buf);
    d->p[i].buf = (uint32_t *)calloc(sizes[i] / 4 + 4, sizeof(uint32_t));
    memcpy(d->p[i].buf, enc + off, sizes[i]);
    d->p[i].ptr = d->p[i].buf;
    Rans64DecInit(&d->p[i].rans, &d->p[i].ptr);
    off += sizes[i];
  }
  return n;
}

/* rans.cpp:57-73 */
static uint32_t dec_get_bits(Rans64State *r, uint32_t **pptr, uint32_t n_bits) {
  uint64_t x = *r;
  const uint32_t val = (uint32_t)(x & ((1u << n_bits) - 1));
  x = x >> n_bits;
  if (x < RANS64_L) {
    x = (x << 32) | **pptr;
    *pptr += 1;
  }
  *r = x;
  return val;
}

/* rans.cpp:272-331 for one part */
static void decode_part(part_dec_t *pd, const int16_t *indexes, int n, const int32_t *cdfs,
                        int cdf_stride, const int32_t *cdf_sizes, const int32_t *offsets,
                        int16_t *out) {
  for (int i = 0; i < n; ++i) {
    const int32_t cdf_idx = indexes[i];
    if (cdf_idx < 0) { /* reference reads offsets[cdf_idx] (UB); emit 0 */
      out[i] = 0;
      continue;
    }
    const int32_t offset = offsets[cdf_idx];
    const int32_t *cdf = cdfs + (size_t)cdf_idx * cdf_stride;
    const int32_t size = cdf_sizes[cdf_idx];
    const int32_t max_value = size - 2;
    const uint32_t cum = Rans64DecGet(&pd->rans, PRECISION);
    int k = 0;
    while (k < size && !((uint32_t)cdf[k] > cum)) ++k; /* linear find_if */
    const uint32_t s = (uint32_t)(k - 1);
    Rans64DecAdvance(&pd->rans, &pd->ptr, (uint32_t)cdf[s], (uint32_t)(cdf[s + 1] - cdf[s]),
                     PRECISION);
    int32_t value = (int32_t)s;
    if (value == max_value) {
      int32_t val = (int32_t)dec_get_bits(&pd->rans, &pd->ptr, BYPASS_PRECISION);
      int32_t n_bypass = val;
      while (val == MAX_BYPASS_VAL) {
        val = (int32_t)dec_get_bits(&pd->rans, &pd->ptr, BYPASS_PRECISION);
        n_bypass += val;
      }
      int32_t raw_val = 0;
      for (int j = 0; j < n_bypass; ++j) {
        val = (int32_t)dec_get_bits(&pd->rans, &pd->ptr, BYPASS_PRECISION);
        raw_val |= val << (j * BYPASS_PRECISION);
      }
      value = raw_val >> 1;
      if (raw_val & 1)
        value = -value - 1;
      else
        value += max_value;
    }
    out[i] = (int16_t)(value + offset);
  }
}

/* py_rans.cpp:168-225 (parts decoded one after the other here; the reference uses
 * std::async per part, results are identical) */
void oracle_dec_decode(oracle_dec_t *d, const int16_t *indexes, int n, const int32_t *cdfs,
                       int cdf_stride, const int32_t *cdf_sizes, const int32_t *offsets,
                       int16_t *out) {
  const int each = n / d->parts;
  const int last = n - each * (d->parts - 1);
  for (int i = 0; i < d->parts; ++i) {
    const int cur = i < d->parts - 1 ? each : last;
    decode_part(&d->p[i], indexes + (size_t)i * each, cur, cdfs, cdf_stride, cdf_sizes, offsets,
                out + (size_t)i * each);
  }
}
