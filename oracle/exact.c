/*
 * exact.c -- plain-C CPU restatement of the paged-KV attention path.  TEST INFRASTRUCTURE ONLY
 * (see oracle/__init__.py): linked by nothing in the product; called from tests/, smoke() and
 * the cpu_baseline leg of bench.py through oracle/exact.py (ctypes).
 *
 * What each function restates (paths relative to /root/reference):
 *   orc_f32_to_e4m3 / orc_kv_write   python/sglang/srt/mem_cache/memory_pool.py:436-458
 *       (cache_k.div_(scale) in the tensor's 16-bit dtype, .to(float8_e4m3fn), buf[loc] = x).
 *       The cast itself lives in a dependency that is not under /root/reference: PyTorch
 *       (torch 2.11 in this image), c10/util/Float8_e4m3fn.h `fp8e4m3fn_from_fp32_value`
 *       -- round to nearest even, |x| >= 480 (after rounding: > 464) or NaN -> 0x7F | sign,
 *       values below 2^-6 land on the subnormal grid k * 2^-9.  Pinned exhaustively against
 *       torch's own cast over all 65536 bf16 bit patterns (tests/test_oracle_exact.py).
 *   orc_gather_rows                  layers/attention/torch_native_backend.py:92-94,155-157
 *   orc_flatten_kv_indices           layers/attention/utils.py:5-41
 *   orc_decode                       layers/attention/torch_native_backend.py:112-180
 *   orc_extend                       layers/attention/torch_native_backend.py:27-110
 *   orc_merge_state                  sgl-kernel/csrc/attention/merge_attn_states.cu:32-106
 * Attention here accumulates in double: it is the "exact" answer used to show on which side
 * (CUDA path: fp32 accumulate; oracle #1: bf16 SDPA) a residual error lies.
 *
 * Element types: 0 = bf16, 1 = f16, 2 = e4m3 (uint8), 3 = f32  (same codes as include/b200attn.h).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_BF16 0
#define ORC_F16 1
#define ORC_E4M3 2
#define ORC_F32 3

static float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

float orc_bf16_to_f32(uint16_t b) { return u2f((uint32_t)b << 16); }

uint16_t orc_f32_to_bf16(float f) { /* round to nearest even; NaN stays NaN */
  uint32_t u = f2u(f);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x0040u);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

float orc_f16_to_f32(uint16_t h) {
  const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  uint32_t e = (h >> 10) & 0x1fu, m = h & 0x3ffu;
  if (e == 0) {
    if (m == 0) return u2f(sign);
    float v = (float)m * 5.9604644775390625e-08f; /* 2^-24 */
    return sign ? -v : v;
  }
  if (e == 31) return u2f(sign | 0x7f800000u | (m << 13));
  return u2f(sign | ((e + 112u) << 23) | (m << 13));
}

uint16_t orc_f32_to_f16(float f) { /* RNE with subnormals */
  const uint32_t u = f2u(f), sign = (u >> 16) & 0x8000u, a = u & 0x7fffffffu;
  if (a > 0x7f800000u) return (uint16_t)(sign | 0x7e00u);
  if (a >= 0x47800000u) return (uint16_t)(sign | 0x7c00u); /* >= 65536 -> inf (65520 rounds up below) */
  if (a < 0x38800000u) {                                    /* < 2^-14: subnormal grid 2^-24 */
    const float v = u2f(a) * 16777216.0f;                   /* exact scaling */
    const float r = nearbyintf(v);                          /* default rounding mode = RNE */
    return (uint16_t)(sign | (uint32_t)r);
  }
  uint32_t r = a + 0xfffu + ((a >> 13) & 1u);
  r = (r - (112u << 23)) >> 13;
  return (uint16_t)(sign | r); /* carries into the exponent (and to inf at 65520) correctly */
}

/* torch.float8_e4m3fn cast of an fp32 value */
uint8_t orc_f32_to_e4m3(float f) {
  uint32_t u = f2u(f);
  const uint32_t sign = (u >> 24) & 0x80u;
  u &= 0x7fffffffu;
  uint32_t r;
  if (u >= 0x43f00000u) { /* >= 480, inf, NaN */
    r = 0x7fu;
  } else if (u < 0x3c800000u) { /* < 2^-6: add-the-magic-number rounding onto the 2^-9 grid */
    const float magic = u2f(141u << 23);
    r = f2u(u2f(u) + magic) - (141u << 23);
  } else {
    const uint32_t odd = (u >> 20) & 1u;
    u += ((uint32_t)(7 - 127) << 23) + 0x7ffffu;
    u += odd;
    r = u >> 20;
  }
  return (uint8_t)(r | sign);
}

float orc_e4m3_to_f32(uint8_t b) {
  const uint32_t e = (b >> 3) & 0xfu, m = b & 7u;
  float v;
  if (e == 15 && m == 7) return u2f(0x7fc00000u | ((uint32_t)(b & 0x80u) << 24));
  if (e == 0) v = (float)m * 0.001953125f; /* 2^-9 */
  else v = ldexpf((float)(8u + m), (int)e - 10);
  return (b & 0x80u) ? -v : v;
}

static double load_elem(const void* p, int64_t i, int dtype) {
  switch (dtype) {
    case ORC_BF16: return (double)orc_bf16_to_f32(((const uint16_t*)p)[i]);
    case ORC_F16: return (double)orc_f16_to_f32(((const uint16_t*)p)[i]);
    case ORC_E4M3: return (double)orc_e4m3_to_f32(((const uint8_t*)p)[i]);
    default: return (double)((const float*)p)[i];
  }
}

static int64_t load_idx(const void* p, int is_i64, int64_t i) {
  return is_i64 ? ((const int64_t*)p)[i] : (int64_t)((const int32_t*)p)[i];
}

/* ------------------------------------------------------------------ KV write
 * pool[loc[t], h, :] = cast(x[t, h, :]); sequential over t, so the LAST duplicate loc wins,
 * as index_put does on CPU.  in_dtype bf16/f16; pool_dtype == in_dtype (bit copy) or e4m3.
 * scale == 0 means "no scale" (None).  scale_mode 0: x / s, 1: x * (1/s); both rounded once to
 * the input's 16-bit dtype before the fp8 cast (the in-place div_ of memory_pool.py:437-440).
 * If writeback != 0 the rounded quotient is stored back into x. */
void orc_kv_write(uint16_t* x, int64_t tok_stride, int64_t head_stride, const int64_t* loc, int T, int H,
                  int D, void* pool, int64_t pool_tok_stride, int in_dtype, int pool_dtype,
                  const float* scale_per_head, int n_scales, int scale_mode, int writeback) {
  for (int t = 0; t < T; ++t) {
    for (int h = 0; h < H; ++h) {
      uint16_t* src = x + (int64_t)t * tok_stride + (int64_t)h * head_stride;
      const int64_t dst = loc[t] * pool_tok_stride + (int64_t)h * D;
      if (pool_dtype == in_dtype) {
        memcpy((uint16_t*)pool + dst, src, (size_t)D * 2);
        continue;
      }
      const float s = scale_per_head ? scale_per_head[n_scales > 1 ? h : 0] : 0.0f;
      for (int d = 0; d < D; ++d) {
        uint16_t b = src[d];
        float f = in_dtype == ORC_BF16 ? orc_bf16_to_f32(b) : orc_f16_to_f32(b);
        if (s != 0.0f) {
          const float q = scale_mode == 1 ? f * (1.0f / s) : f / s;
          b = in_dtype == ORC_BF16 ? orc_f32_to_bf16(q) : orc_f32_to_f16(q);
          f = in_dtype == ORC_BF16 ? orc_bf16_to_f32(b) : orc_f16_to_f32(b);
          if (writeback) src[d] = b;
        }
        ((uint8_t*)pool)[dst + d] = orc_f32_to_e4m3(f);
      }
    }
  }
}

/* dst[j, :] = pool[indices[j], :]  (rows of row_bytes) */
void orc_gather_rows(const uint8_t* pool, int64_t tok_stride_bytes, const int32_t* indices, int64_t n,
                     int64_t row_bytes, uint8_t* dst) {
  for (int64_t j = 0; j < n; ++j) memcpy(dst + j * row_bytes, pool + (int64_t)indices[j] * tok_stride_bytes, (size_t)row_bytes);
}

/* kv_indices[kv_indptr[i] + j] = table[rpi[i], start[i] + j],  j < lens[i] */
void orc_flatten_kv_indices(const int32_t* table, int64_t table_stride, const void* rpi, const void* lens,
                            const void* start, int is_i64, const int32_t* kv_indptr, int32_t* kv_indices, int bs) {
  for (int i = 0; i < bs; ++i) {
    const int64_t row = load_idx(rpi, is_i64, i), n = load_idx(lens, is_i64, i);
    const int64_t s = start ? load_idx(start, is_i64, i) : 0;
    for (int64_t j = 0; j < n; ++j) kv_indices[kv_indptr[i] + j] = table[row * table_stride + s + j];
  }
}

/* one query row against keys [0, n_vis) of request row `trow` of the table */
static void attend_row(const void* q, int64_t q_off, int q_dtype, const void* kp, const void* vp,
                       int64_t pool_tok_stride, int kv_dtype, int kv_head, int D, const int32_t* trow,
                       int64_t n_vis, double scale, double logit_cap, double v_scale, double* logits,
                       float* out, float* lse_out) {
  double mx = -INFINITY;
  for (int64_t j = 0; j < n_vis; ++j) {
    const int64_t base = (int64_t)trow[j] * pool_tok_stride + (int64_t)kv_head * D;
    double acc = 0.0;
    for (int d = 0; d < D; ++d) acc += load_elem(q, q_off + d, q_dtype) * load_elem(kp, base + d, kv_dtype);
    acc *= scale;
    if (logit_cap > 0.0) acc = logit_cap * tanh(acc / logit_cap);
    logits[j] = acc;
    if (acc > mx) mx = acc;
  }
  double sum = 0.0;
  for (int64_t j = 0; j < n_vis; ++j) { logits[j] = exp(logits[j] - mx); sum += logits[j]; }
  for (int d = 0; d < D; ++d) {
    double acc = 0.0;
    for (int64_t j = 0; j < n_vis; ++j)
      acc += logits[j] * load_elem(vp, (int64_t)trow[j] * pool_tok_stride + (int64_t)kv_head * D + d, kv_dtype);
    out[d] = n_vis > 0 ? (float)(acc / sum * v_scale) : 0.0f;
  }
  if (lse_out) *lse_out = n_vis > 0 ? (float)(mx + log(sum)) : -INFINITY;
}

/* out fp32 [bs, Hq, D]; lse fp32 [bs, Hq] or NULL.  scale = sm_scale * k_scale. */
int orc_decode(const void* q, int64_t q_tok_stride, int q_dtype, const void* k_pool, const void* v_pool,
               int64_t pool_tok_stride, int kv_dtype, const int32_t* table, int64_t table_stride,
               const void* rpi, const void* seq_lens, int is_i64, int bs, int Hq, int Hkv, int D,
               double scale, double logit_cap, double v_scale, float* out, float* lse) {
  const int G = Hq / Hkv;
  int64_t max_len = 1;
  for (int i = 0; i < bs; ++i) { const int64_t L = load_idx(seq_lens, is_i64, i); if (L > max_len) max_len = L; }
  double* logits = (double*)malloc(sizeof(double) * (size_t)max_len);
  if (!logits) return -1;
  for (int i = 0; i < bs; ++i) {
    const int64_t L = load_idx(seq_lens, is_i64, i);
    const int32_t* trow = table + load_idx(rpi, is_i64, i) * table_stride;
    for (int h = 0; h < Hq; ++h)
      attend_row(q, (int64_t)i * q_tok_stride + (int64_t)h * D, q_dtype, k_pool, v_pool, pool_tok_stride, kv_dtype,
                 h / G, D, trow, L, scale, logit_cap, v_scale, logits, out + ((int64_t)i * Hq + h) * D,
                 lse ? lse + (int64_t)i * Hq + h : NULL);
  }
  free(logits);
  return 0;
}

/* out fp32 [T, Hq, D].  Query t of request i (row start_loc[i] + t) sees keys j <= prefix_i + t
 * when causal, all seq_len_i keys otherwise. */
int orc_extend(const void* q, int64_t q_tok_stride, int q_dtype, const void* k_pool, const void* v_pool,
               int64_t pool_tok_stride, int kv_dtype, const int32_t* table, int64_t table_stride,
               const void* rpi, const void* seq_lens, int is_i64, const int32_t* extend_seq_lens,
               const int32_t* extend_prefix_lens, const int32_t* extend_start_loc, int bs, int Hq, int Hkv,
               int D, double scale, double logit_cap, double v_scale, int causal, float* out, float* lse) {
  const int G = Hq / Hkv;
  int64_t max_len = 1;
  for (int i = 0; i < bs; ++i) { const int64_t L = load_idx(seq_lens, is_i64, i); if (L > max_len) max_len = L; }
  double* logits = (double*)malloc(sizeof(double) * (size_t)max_len);
  if (!logits) return -1;
  for (int i = 0; i < bs; ++i) {
    const int64_t L = load_idx(seq_lens, is_i64, i);
    const int32_t* trow = table + load_idx(rpi, is_i64, i) * table_stride;
    for (int t = 0; t < extend_seq_lens[i]; ++t) {
      const int64_t row = (int64_t)extend_start_loc[i] + t;
      const int64_t n_vis = causal ? (int64_t)extend_prefix_lens[i] + t + 1 : L;
      for (int h = 0; h < Hq; ++h)
        attend_row(q, row * q_tok_stride + (int64_t)h * D, q_dtype, k_pool, v_pool, pool_tok_stride, kv_dtype,
                   h / G, D, trow, n_vis, scale, logit_cap, v_scale, logits, out + (row * Hq + h) * D,
                   lse ? lse + row * Hq + h : NULL);
    }
  }
  free(logits);
  return 0;
}

/* 2-way LSE merge; a, b [rows, D] of dtype; out fp32 [rows, D]; out_lse [rows] or NULL */
void orc_merge_state(const void* a, const float* lse_a, const void* b, const float* lse_b, int64_t rows, int D,
                     int dtype, float* out, float* out_lse) {
  for (int64_t r = 0; r < rows; ++r) {
    double la = lse_a[r], lb = lse_b[r];
    if (isinf(la) && la > 0) la = -INFINITY;
    if (isinf(lb) && lb > 0) lb = -INFINITY;
    const double m = la > lb ? la : lb;
    const double ea = isinf(m) ? 0.0 : exp(la - m), eb = isinf(m) ? 0.0 : exp(lb - m);
    const double s = ea + eb;
    for (int d = 0; d < D; ++d)
      out[r * D + d] = s > 0 ? (float)((load_elem(a, r * D + d, dtype) * ea + load_elem(b, r * D + d, dtype) * eb) / s) : 0.0f;
    if (out_lse) out_lse[r] = s > 0 ? (float)(log(s) + m) : -INFINITY;
  }
}

/* ---------------------------------------------------------------- HiCache KV transfer
 * Restates sgl-kernel/csrc/kvcacheio/transfer.cu:9-110: for every item and layer, copy item_bytes bytes from
 * src(layer, src_idx[i]) to dst(layer, dst_idx[i]); layouts 0 = layer first (:25-35), 1 = page first (:37-47),
 * 2 = per-layer pointer table (:49-59).  One buffer per call (K, V or the single MLA buffer). */
static const uint8_t* orc_item_addr(const uint8_t* base, int layout, int64_t dim, int64_t layer, int64_t index,
                                    int64_t item_bytes) {
  if (layout == 0) return base + layer * dim + index * item_bytes;
  if (layout == 1) return base + index * dim + layer * item_bytes;
  return (const uint8_t*)(uintptr_t)(((const uint64_t*)base)[layer]) + index * item_bytes;
}

void orc_transfer_kv(const uint8_t* src, uint8_t* dst, int src_layout, int dst_layout, int64_t src_dim, int64_t dst_dim,
                     const int64_t* src_idx, const int64_t* dst_idx, int64_t n_items, int64_t item_bytes,
                     int64_t start_layer, int64_t num_layers) {
  for (int64_t i = 0; i < n_items; ++i)
    for (int64_t l = start_layer; l < start_layer + num_layers; ++l)
      memcpy((uint8_t*)orc_item_addr(dst, dst_layout, dst_dim, l, dst_idx[i], item_bytes),
             orc_item_addr(src, src_layout, src_dim, l, src_idx[i], item_bytes), (size_t)item_bytes);
}
