// Paged extend / prefill attention, register-MMA kernel (B200ATTN_EXTEND_MMA).
//
// Reference semantics: python/sglang/srt/layers/attention/torch_native_backend.py:27-110 -- for
// request i the E_i new queries sit at absolute positions P_i .. P_i+E_i-1 and query t sees keys
// j <= P_i + t (causal) of the L_i = P_i + E_i tokens addressed by req_to_token[rpi[i], :L_i]; ALL
// keys (cached prefix and new tokens) are read from the paged pool, i.e. the KV write happened first
// (torch_native_backend.py:196-199).  GPU design it supersedes: triton_ops/extend_attention.py:42-303.
//
// This is the general kernel: any ragged batch, D in {64,128}, bf16/f16 or FP8-E4M3 KV, rows with
// E_i = 1 (the decode rows of a MIXED batch).  The dense tcgen05/TMEM kernel for D = 128 bf16 lives in
// extend_tc.cu; b200attn_extend (extend.cu) picks between the two.
//
// One CTA = 64 queries of one q head (4 warps x 16 rows); KV walked in 64-token tiles gathered row by
// row through the page table with 16-byte cp.async into a double-buffered XOR-swizzled smem ring;
// S = Q K^T and O += P V on mma.sync.m16n8k16 with fp32 online softmax in the log2 domain.
#include "common.cuh"
#include "extend.cuh"

namespace b200attn {

constexpr int kExtThreads = 128;
constexpr int kExtBM = 64;
constexpr int kExtBN = 64;

template <int D, int KVB>
struct ExtCfg {
  static constexpr int HEADB = D * KVB;        // bytes of one head row
  static constexpr int CPH = HEADB / 16;       // 16-byte chunks per row
  static constexpr int TILEB = kExtBN * HEADB; // K (or V) tile
  static constexpr int STAGEB = 2 * TILEB;
  static constexpr int SWZ = CPH >= 8 ? 7 : (CPH - 1);
  static constexpr int CHUNKS = kExtBN * CPH;
  static constexpr int ITERS = CHUNKS / kExtThreads;
  static constexpr int NSTAGE = 2;
  static constexpr int SMEM = NSTAGE * STAGEB + 3 * kExtBN * 4;
  static_assert(CHUNKS % kExtThreads == 0, "tile must divide over the CTA");
};

template <int D, int KVB, bool kQBf16, bool kMasked>
__global__ void __launch_bounds__(kExtThreads) extend_mma_kernel(const ExtendParams p) {
  using Cfg = ExtCfg<D, KVB>;
  constexpr bool kFp8 = (KVB == 1);
  constexpr bool kMmaBf16 = kQBf16 && !kFp8;
  constexpr int KSTEPS = D / 16, NBLK = D / 8;
  constexpr int BN = kExtBN;

  extern __shared__ __align__(128) uint8_t smem[];
  int32_t* idx_ring = reinterpret_cast<int32_t*>(smem + Cfg::NSTAGE * Cfg::STAGEB);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int qd = lane & 3, row = lane >> 2;
  const int req = blockIdx.z, head = blockIdx.y;
  const int E = p.extend_seq_lens[req];
  const int qt = static_cast<int>(gridDim.x) - 1 - static_cast<int>(blockIdx.x);  // heavy (late) tiles first
  const int q0 = qt * kExtBM;
  if (q0 >= E) return;
  const int P = p.extend_prefix_lens[req];
  const int L = static_cast<int>(load_index(p.seq_lens, p.idx_is_i64, req));
  const int start = p.extend_start_loc[req];
  const int kv_head = head / p.G;
  const int32_t* trow = p.req_to_token + load_index(p.req_pool_indices, p.idx_is_i64, req) * p.table_stride;
  const int q_hi = min(q0 + kExtBM, E);                       // one past the last query of this tile
  const int kv_end = p.causal ? min(L, P + q_hi) : L;         // keys any row of the tile can see
  const int ntiles = (kv_end + BN - 1) / BN;
  // masked variant: first prefix key inside the sliding window (0 without a window) and its tile
  int kv_begin = 0;
  if constexpr (kMasked) {
    if (p.window >= 0) kv_begin = P - min(P, p.window + 1);
    else if (p.window == B200ATTN_WINDOW_NEW_TOKENS_ONLY) kv_begin = P;  // prefix part computed elsewhere
  }
  const int tile0 = kv_begin / BN;

  const uint8_t* kpool = reinterpret_cast<const uint8_t*>(p.k_pool) + static_cast<int64_t>(kv_head) * Cfg::HEADB;
  const uint8_t* vpool = reinterpret_cast<const uint8_t*>(p.v_pool) + static_cast<int64_t>(kv_head) * Cfg::HEADB;
  const int64_t pool_row_bytes = p.pool_tok_stride * KVB;

  auto stage_indices = [&](int t) {
    if (t < ntiles && tid < BN) {
      const int tok = t * BN + tid;
      idx_ring[(t % 3) * BN + tid] = (tok < kv_end && tok >= kv_begin) ? __ldg(trow + tok) : -1;
    }
  };
  auto issue_tile = [&](int t) {
    if (t < ntiles) {
      const int32_t* idx = idx_ring + (t % 3) * BN;
      const uint32_t k_s = smem_u32(smem + (t % Cfg::NSTAGE) * Cfg::STAGEB), v_s = k_s + Cfg::TILEB;
#pragma unroll
      for (int it = 0; it < Cfg::ITERS; ++it) {
        const int c = it * kExtThreads + tid;
        const int tok = c / Cfg::CPH, ch = c % Cfg::CPH;
        const int32_t slot = idx[tok];
        const int valid = slot >= 0 ? 16 : 0;
        const int64_t src = static_cast<int64_t>(slot >= 0 ? slot : 0) * pool_row_bytes + ch * 16;
        const uint32_t dst = tok * Cfg::HEADB + ((ch ^ (tok & Cfg::SWZ)) << 4);
        cp_async16(k_s + dst, kpool + src, valid);
        cp_async16(v_s + dst, vpool + src, valid);
      }
    }
    cp_async_commit();
  };

  stage_indices(tile0);
  stage_indices(tile0 + 1);
  __syncthreads();
  issue_tile(tile0);

  // ---- Q fragments: rows = 16 queries of this warp (zero for rows past E)
  const int t0 = q0 + warp * 16 + row, t1 = t0 + 8;  // query index inside the request
  uint32_t qa[KSTEPS][4];
  {
    const uint16_t* qbase = reinterpret_cast<const uint16_t*>(p.q) + static_cast<int64_t>(head) * D;
    const uint16_t* qrow0 = qbase + static_cast<int64_t>(start + t0) * p.q_tok_stride;
    const uint16_t* qrow1 = qbase + static_cast<int64_t>(start + t1) * p.q_tok_stride;
    const bool v0 = t0 < E, v1 = t1 < E;
#pragma unroll
    for (int ks = 0; ks < KSTEPS; ++ks) {
      uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
      if constexpr (!kFp8) {
        if (v0) {
          a0 = *reinterpret_cast<const uint32_t*>(qrow0 + ks * 16 + 2 * qd);
          a2 = *reinterpret_cast<const uint32_t*>(qrow0 + ks * 16 + 8 + 2 * qd);
        }
        if (v1) {
          a1 = *reinterpret_cast<const uint32_t*>(qrow1 + ks * 16 + 2 * qd);
          a3 = *reinterpret_cast<const uint32_t*>(qrow1 + ks * 16 + 8 + 2 * qd);
        }
      } else {
        // fp8 K: a b16 ldmatrix register holds d = 16*ks + 4*qd + {0..3}; permute Q's contraction index alike
        if (v0) { const uint2 t = *reinterpret_cast<const uint2*>(qrow0 + ks * 16 + 4 * qd); a0 = t.x; a2 = t.y; }
        if (v1) { const uint2 t = *reinterpret_cast<const uint2*>(qrow1 + ks * 16 + 4 * qd); a1 = t.x; a3 = t.y; }
        if constexpr (kQBf16) {
          auto cv = [](uint32_t x) {
            const float lo = __uint_as_float(x << 16), hi = __uint_as_float(x & 0xffff0000u);
            return pack2<false>(fminf(fmaxf(lo, -65504.f), 65504.f), fminf(fmaxf(hi, -65504.f), 65504.f));
          };
          a0 = cv(a0); a1 = cv(a1); a2 = cv(a2); a3 = cv(a3);
        }
      }
      qa[ks][0] = a0; qa[ks][1] = a1; qa[ks][2] = a2; qa[ks][3] = a3;
    }
  }
  // keys visible to the two rows this thread owns
  const int vis0 = t0 < E ? (p.causal ? min(P + t0 + 1, L) : L) : 0;
  const int vis1 = t1 < E ? (p.causal ? min(P + t1 + 1, L) : L) : 0;
  const int warp_vis = p.causal ? min(P + min(q0 + warp * 16 + 16, E), L) : L;  // max over the warp's rows

  float o[NBLK][4];
#pragma unroll
  for (int nb = 0; nb < NBLK; ++nb) { o[nb][0] = o[nb][1] = o[nb][2] = o[nb][3] = 0.f; }
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

  // masked variant: per-row predicate pieces
  //   prefix key j (kv_begin <= j < P) needs j >= lo (window rule t <= (j - kv_begin) + window)
  int lo0 = kv_begin, lo1 = kv_begin;
  const uint8_t* mrow0 = nullptr;
  const uint8_t* mrow1 = nullptr;
  if constexpr (kMasked) {
    if (p.window > 0) {
      lo0 = kv_begin + max(0, t0 - p.window);
      lo1 = kv_begin + max(0, t1 - p.window);
    }
    if (p.custom_mask != nullptr) {
      const uint8_t* mb = p.custom_mask + p.mask_indptr[req];
      mrow0 = mb + static_cast<int64_t>(min(t0, E - 1)) * L;
      mrow1 = mb + static_cast<int64_t>(min(t1, E - 1)) * L;
    }
  }

  for (int t = tile0; t < ntiles; ++t) {
    issue_tile(t + 1);
    stage_indices(t + 2);
    cp_async_wait<1>();
    __syncthreads();
    const uint32_t k_s = smem_u32(smem + (t % Cfg::NSTAGE) * Cfg::STAGEB), v_s = k_s + Cfg::TILEB;
    const int kv0 = t * BN;
    if (kv0 < warp_vis && q0 + warp * 16 < E) {
      // ---- S = Q K^T : 16 x 64
      float s[8][4];
#pragma unroll
      for (int nb = 0; nb < 8; ++nb) { s[nb][0] = s[nb][1] = s[nb][2] = s[nb][3] = 0.f; }
      const int mj = lane >> 3, mr = lane & 7;
#pragma unroll
      for (int g = 0; g < 4; ++g) {  // 16-token groups
        if constexpr (!kFp8) {
#pragma unroll
          for (int ks = 0; ks < KSTEPS; ++ks) {
            const int tok = g * 16 + (mj >> 1) * 8 + mr;
            const int ch = 2 * ks + (mj & 1);
            uint32_t r[4];
            ldmatrix_x4(r, k_s + tok * Cfg::HEADB + ((ch ^ (tok & Cfg::SWZ)) << 4));
            mma_16816<kMmaBf16>(s[2 * g], qa[ks], r[0], r[1]);
            mma_16816<kMmaBf16>(s[2 * g + 1], qa[ks], r[2], r[3]);
          }
        } else {
#pragma unroll
          for (int kp = 0; kp < KSTEPS / 2; ++kp) {
            const int tok = g * 16 + (mj & 1) * 8 + mr;
            const int ch = 2 * kp + (mj >> 1);
            uint32_t r[4];
            ldmatrix_x4(r, k_s + tok * Cfg::HEADB + ((ch ^ (tok & Cfg::SWZ)) << 4));
#pragma unroll
            for (int half = 0; half < 2; ++half) {
              const uint32_t r0 = r[2 * half], r1 = r[2 * half + 1];
              const int ks = 2 * kp + half;
              mma_16816<false>(s[2 * g], qa[ks], e4m3x2_to_f16x2(r0 & 0xffff), e4m3x2_to_f16x2(r0 >> 16));
              mma_16816<false>(s[2 * g + 1], qa[ks], e4m3x2_to_f16x2(r1 & 0xffff), e4m3x2_to_f16x2(r1 >> 16));
            }
          }
        }
      }
      // ---- scale, cap, mask
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int nb = 0; nb < 8; ++nb) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float x;
          if (p.logit_cap > 0.f) x = p.logit_cap * fast_tanh(s[nb][e] * p.scale_raw / p.logit_cap) * kLog2e;
          else x = s[nb][e] * p.scale_log2;
          const int j = kv0 + nb * 8 + 2 * qd + (e & 1);
          bool ok = j < ((e & 2) ? vis1 : vis0);
          if constexpr (kMasked) {
            if (j < P) ok = ok && j >= ((e & 2) ? lo1 : lo0);
            if (mrow0 != nullptr && ok && (j >= P || !p.skip_prefix_mask)) ok = __ldg(((e & 2) ? mrow1 : mrow0) + j) != 0;
          }
          x = ok ? x : -INFINITY;
          s[nb][e] = x;
          if (e & 2) mx1 = fmaxf(mx1, x); else mx0 = fmaxf(mx0, x);
        }
      }
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
      const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
      const float base0 = (mn0 == -INFINITY) ? 0.f : mn0, base1 = (mn1 == -INFINITY) ? 0.f : mn1;
      const float al0 = fast_exp2(m0 - base0), al1 = fast_exp2(m1 - base1);  // m == -inf -> 0
      m0 = mn0; m1 = mn1;
      float rs0 = 0.f, rs1 = 0.f;
      uint32_t pa[4][4];
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const float p00 = fast_exp2(s[2 * g][0] - base0), p01 = fast_exp2(s[2 * g][1] - base0);
        const float p02 = fast_exp2(s[2 * g][2] - base1), p03 = fast_exp2(s[2 * g][3] - base1);
        const float p10 = fast_exp2(s[2 * g + 1][0] - base0), p11 = fast_exp2(s[2 * g + 1][1] - base0);
        const float p12 = fast_exp2(s[2 * g + 1][2] - base1), p13 = fast_exp2(s[2 * g + 1][3] - base1);
        rs0 += p00 + p01 + p10 + p11;
        rs1 += p02 + p03 + p12 + p13;
        pa[g][0] = pack2<kMmaBf16>(p00, p01);
        pa[g][1] = pack2<kMmaBf16>(p02, p03);
        pa[g][2] = pack2<kMmaBf16>(p10, p11);
        pa[g][3] = pack2<kMmaBf16>(p12, p13);
      }
      l0 = l0 * al0 + rs0;
      l1 = l1 * al1 + rs1;
#pragma unroll
      for (int nb = 0; nb < NBLK; ++nb) { o[nb][0] *= al0; o[nb][1] *= al0; o[nb][2] *= al1; o[nb][3] *= al1; }

      // ---- O += P V
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        if constexpr (!kFp8) {
#pragma unroll
          for (int jp = 0; jp < NBLK / 2; ++jp) {
            const int tok = g * 16 + (mj & 1) * 8 + mr;
            const int ch = 2 * jp + (mj >> 1);
            uint32_t r[4];
            ldmatrix_x4_trans(r, v_s + tok * Cfg::HEADB + ((ch ^ (tok & Cfg::SWZ)) << 4));
            mma_16816<kMmaBf16>(o[2 * jp], pa[g], r[0], r[1]);
            mma_16816<kMmaBf16>(o[2 * jp + 1], pa[g], r[2], r[3]);
          }
        } else {
#pragma unroll
          for (int cp = 0; cp < D / 32; ++cp) {
            const int tok = g * 16 + (mj & 1) * 8 + mr;
            const int ch = 2 * cp + (mj >> 1);
            uint32_t r[4];
            ldmatrix_x4_trans(r, v_s + tok * Cfg::HEADB + ((ch ^ (tok & Cfg::SWZ)) << 4));
#pragma unroll
            for (int half = 0; half < 2; ++half) {
              const uint32_t lo = r[2 * half], hi = r[2 * half + 1];
              const uint32_t e_lo = __byte_perm(lo, 0, 0x0020), o_lo = __byte_perm(lo, 0, 0x0031);
              const uint32_t e_hi = __byte_perm(hi, 0, 0x0020), o_hi = __byte_perm(hi, 0, 0x0031);
              const int c = 2 * cp + half;
              mma_16816<false>(o[2 * c], pa[g], e4m3x2_to_f16x2(e_lo), e4m3x2_to_f16x2(e_hi));
              mma_16816<false>(o[2 * c + 1], pa[g], e4m3x2_to_f16x2(o_lo), e4m3x2_to_f16x2(o_hi));
            }
          }
        }
      }
    }
    __syncthreads();  // the stage is refilled by the next iteration's issue_tile
  }
  cp_async_wait<0>();

  // ---- normalise and store
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const int t = rr ? t1 : t0;
    if (t >= E) continue;
    const float l = rr ? l1 : l0, m = rr ? m1 : m0;
    const float inv = l > 0.f ? p.v_scale / l : 0.f;
    uint16_t* dst = reinterpret_cast<uint16_t*>(p.out) + static_cast<int64_t>(start + t) * p.out_tok_stride +
                    static_cast<int64_t>(head) * D;
    if constexpr (!kFp8) {
#pragma unroll
      for (int nb = 0; nb < NBLK; ++nb)
        *reinterpret_cast<uint32_t*>(dst + nb * 8 + 2 * qd) = pack2<kQBf16>(o[nb][2 * rr] * inv, o[nb][2 * rr + 1] * inv);
    } else {
#pragma unroll
      for (int c = 0; c < D / 16; ++c) {
        const uint32_t w0 = pack2<kQBf16>(o[2 * c][2 * rr] * inv, o[2 * c + 1][2 * rr] * inv);
        const uint32_t w1 = pack2<kQBf16>(o[2 * c][2 * rr + 1] * inv, o[2 * c + 1][2 * rr + 1] * inv);
        *reinterpret_cast<uint2*>(dst + c * 16 + 4 * qd) = make_uint2(w0, w1);
      }
    }
    if (p.lse_out != nullptr && qd == 0)
      p.lse_out[static_cast<int64_t>(start + t) * p.Hq + head] = l > 0.f ? (m + log2f(l)) * kLn2 : -INFINITY;
  }
}

template <int D, int KVB, bool kQBf16, bool kMasked = false>
static int launch_mma(const ExtendParams& p, cudaStream_t st) {
  using Cfg = ExtCfg<D, KVB>;
  auto kern = extend_mma_kernel<D, KVB, kQBf16, kMasked>;
  static bool configured = false;
  if (!configured) {
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    configured = true;
  }
  dim3 grid((p.max_extend_len + kExtBM - 1) / kExtBM, p.Hq, p.bs);
  kern<<<grid, kExtThreads, Cfg::SMEM, st>>>(p);
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}

int extend_mma_launch(const ExtendParams& p, int D, int kvb, bool q_bf16, cudaStream_t st) {
  if (D == 128) {
    if (kvb == 2) return q_bf16 ? launch_mma<128, 2, true>(p, st) : launch_mma<128, 2, false>(p, st);
    return q_bf16 ? launch_mma<128, 1, true>(p, st) : launch_mma<128, 1, false>(p, st);
  }
  if (kvb == 2) return q_bf16 ? launch_mma<64, 2, true>(p, st) : launch_mma<64, 2, false>(p, st);
  return q_bf16 ? launch_mma<64, 1, true>(p, st) : launch_mma<64, 1, false>(p, st);
}

int extend_mma_masked_launch(const ExtendParams& p, int D, int kvb, bool q_bf16, cudaStream_t st) {
  if (D == 128) {
    if (kvb == 2) return q_bf16 ? launch_mma<128, 2, true, true>(p, st) : launch_mma<128, 2, false, true>(p, st);
    return q_bf16 ? launch_mma<128, 1, true, true>(p, st) : launch_mma<128, 1, false, true>(p, st);
  }
  if (kvb == 2) return q_bf16 ? launch_mma<64, 2, true, true>(p, st) : launch_mma<64, 2, false, true>(p, st);
  return q_bf16 ? launch_mma<64, 1, true, true>(p, st) : launch_mma<64, 1, false, true>(p, st);
}

}  // namespace b200attn
