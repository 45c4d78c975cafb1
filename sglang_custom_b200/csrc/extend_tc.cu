// Paged extend / prefill attention on the 5th-generation tensor cores (B200ATTN_EXTEND_TCGEN05).
//
// Reference semantics: python/sglang/srt/layers/attention/torch_native_backend.py:27-110 (query t of request i,
// at absolute position P_i + t, sees keys j <= P_i + t of the L_i tokens addressed by req_to_token; every key --
// cached radix prefix and new tokens alike -- is read from the paged pool, torch_native_backend.py:196-199).
// GPU design it supersedes: triton_ops/extend_attention.py:42-303 (tl.dot -> mma.sync, num_stages = 1, prefix K/V
// re-gathered by every 128-row q tile).
//
// Design (D = 128, bf16 / f16 Q and KV; the bound is the bf16 tensor pipe, co-limited by exp2 on the SFUs)
//   * one CTA owns TWO 128-row Q tiles that share every K/V tile: two q heads of one GQA group over the same
//     128 tokens (G even) or two consecutive 128-token tiles of one head (G odd).  K/V smem traffic per FLOP is
//     halved and the softmax of one tile overlaps the MMAs of the other.
//   * tensor memory (512 columns): S0 | S1 | O0 | O1, 128 fp32 columns each.  S = Q K^T is an SS-mode
//     tcgen05.mma (both operands from 128B-swizzled shared memory); P (bf16) is written back over the first 64
//     columns of S and O += P V runs in TS mode (A from tensor memory, V as an MN-major B operand), so P never
//     touches shared memory.  P is handed over in two 64-key halves, so the first four PV k-steps start while
//     the softmax warps still work on the second half.
//   * exp2 is the co-limit: 2 x 128 x 128 exponentials per K/V tile are 2048 MUFU cycles per SM at 16/clk -- as
//     long as the 2048 tensor cycles of the four MMAs.  kEmu of every four element pairs are therefore computed
//     on the FMA pipe instead (ptx::exp2_poly2: Cody-Waite split + degree-3 minimax polynomial on packed f32x2,
//     7.5e-5 relative error, invisible after the bf16 rounding of P); masked (diagonal / last) tiles use MUFU only
//     so that masked keys are exact zeros.
//   * K/V tiles of 128 tokens are gathered row by row through req_to_token by three producer warps with 16-byte
//     cp.async (LDGSTS) straight into the UMMA canonical 128B-swizzle layout; completion is counted on an
//     mbarrier with cp.async.mbarrier.arrive.noinc, so the producers never wait for their own copies -- no
//     flattened kv_indices copy and no register staging.  Two K and two V buffers of 32 KiB form the ring.
//     (The first cut used TMA gather4, four rows x 128 B per instruction: measured 63 cycles per instruction,
//     i.e. 8 000 cycles per 128-key K+V tile against a 2 048-cycle MMA budget -- profiles/README.md -- so the
//     page_size = 1 gather moved to LDGSTS; tma_gather4 stays in tc_ptx.cuh for the paged-tile path.)
//   * warp roles (three warpgroups, registers re-balanced with setmaxnreg): warps 0-3 / 4-7 = softmax + output
//     of Q tile 0 / 1 (one query row per thread, fp32 online softmax in the log2 domain on packed f32x2 FMAs, the
//     O rescale is lazy: only when the row maximum grows by more than 2^8), warp 8 = MMA issuer (the whole warp
//     stays converged so descriptors live in uniform registers; one elected lane issues), warps 9-11 = producers.
//     All hand-offs are mbarriers; tcgen05.commit tracks MMA completion.  Every wait is bounded (ptx::mbar_wait
//     traps instead of hanging).
//   * grid order: head pair fastest, then q tile (heavy first), then request -- the CTAs in flight at any time
//     stream the K/V of one request, which therefore stays in L2 across its 16 head pairs x q tiles.
#include "extend.cuh"
#include "tc_ptx.cuh"
#include "tma.cuh"

namespace b200attn {
namespace tc {

constexpr int kThreads = 384;       // 3 warpgroups
constexpr int kProducerWarps = 3;
// setmaxnreg targets.  The CTA is launched with 168 registers per thread (64 512 in its pool), so the 128 threads of the
// third warpgroup must release at least what the 256 softmax threads claim: 128 x (168 - 88) = 256 x (208 - 168).
constexpr int kRegsSoftmax = 208;
constexpr int kRegsOther = 88;
constexpr int kTile = 128;           // BM = BN = D = 128
constexpr int kTileBytes = 32768;    // 128 rows x 256 B
constexpr int kHalfBytes = 16384;    // 128 rows x 128 B (one 64-element swizzle atom column)
constexpr int kOffQ = 0;             // Q0, Q1
constexpr int kOffK = 2 * kTileBytes;
constexpr int kOffV = 4 * kTileBytes;
constexpr int kOffBar = 6 * kTileBytes;
constexpr int kSmemBytes = kOffBar + 32 * 8 + 16 + 1024;  // + barriers + tmem slot + alignment slack
constexpr float kRescaleThreshold = 8.0f;                 // log2 units

// P_FULL + 2 * t + h: half h (64 keys) of P_t is in tensor memory
enum Bar { Q_FULL = 0, K_FULL = 1, K_EMPTY = 3, V_FULL = 5, V_EMPTY = 7, S_FULL = 9, O_FULL = 11, P_FULL = 13, kNumBars = 17 };

// dev-only phase timing (kernel variant kProf): every warp accumulates clock64 deltas, CTA (0,0,0) stores them
template <bool kOn>
struct Prof {
  unsigned long long acc[8];
  long long t0;
  __device__ __forceinline__ void begin() {
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0;
    t0 = clock64();
  }
  __device__ __forceinline__ void mark(int slot) {
    const long long now = clock64();
    acc[slot] += static_cast<unsigned long long>(now - t0);
    t0 = now;
  }
  __device__ __forceinline__ void flush(unsigned long long* dst, int warp, int lane) {
    if (dst != nullptr && lane == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0)
      for (int i = 0; i < 8; ++i) dst[warp * 8 + i] = acc[i];
  }
};
template <>
struct Prof<false> {
  __device__ __forceinline__ void begin() {}
  __device__ __forceinline__ void mark(int) {}
  __device__ __forceinline__ void flush(unsigned long long*, int, int) {}
};

// One K/V tile of one query row: S (fp32, tensor memory) -> online-softmax update -> P (bf16, tensor memory).
//   kMask: some of the 128 keys are invisible to some row (diagonal / last tile) -> mask, MUFU only
//   kEmu : of every 4 element pairs, kEmu take the FMA-pipe exp2
//   kWin : sliding-window variant -- keys [0, nlo) of the tile are invisible too, and a row may see nothing at all
//          in its first tiles: the running maximum then stays -inf until the first visible key (no rescale needed,
//          l and O are still exact zeros)
template <bool kBf16, bool kMask, int kEmu, bool kProf, bool kWin = false>
__device__ __forceinline__ void softmax_tile(const uint32_t tS, const uint32_t tO, const uint32_t bar_s,
                                             const uint32_t bar_p, const int j, const int nvis,
                                             const float scale_log2, float& m_used, float& l, Prof<kProf>& prof,
                                             const int nlo = 0) {
  ptx::mbar_wait(bar_s, j & 1);
  ptx::tc_fence_after();
  prof.mark(0);
  uint32_t s[128];
#pragma unroll
  for (int c = 0; c < 4; ++c) ptx::tmem_ld32(tS + c * 32, s + 32 * c);
  ptx::tmem_wait_ld();
  prof.mark(1);

  if constexpr (kMask) {
#pragma unroll
    for (int i = 0; i < 128; ++i) s[i] = (i < nvis && (!kWin || i >= nlo)) ? s[i] : 0xff800000u;  // -inf
  }
  float mx8[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) mx8[i] = fmaxf(__uint_as_float(s[i]), __uint_as_float(s[8 + i]));
#pragma unroll
  for (int i = 16; i < 128; i += 16)
#pragma unroll
    for (int k = 0; k < 8; ++k)
      mx8[k] = fmaxf(fmaxf(mx8[k], __uint_as_float(s[i + k])), __uint_as_float(s[i + 8 + k]));
  float mx = fmaxf(fmaxf(fmaxf(mx8[0], mx8[1]), fmaxf(mx8[2], mx8[3])),
                   fmaxf(fmaxf(mx8[4], mx8[5]), fmaxf(mx8[6], mx8[7])));
  mx *= scale_log2;  // scale > 0: the maximum commutes with the scaling
  prof.mark(2);

  if (j == 0) {
    m_used = (!kWin && mx == -INFINITY) ? 0.f : mx;
  } else {
    const bool grow = mx > m_used + kRescaleThreshold;
    if (__any_sync(0xffffffffu, grow)) {
      // S_t(j) is complete, so PV_t(j-1) -- issued before it -- is complete too: O_t is stable until this
      // thread group hands P_t(j) to the MMA warp
      const float m_new = grow ? mx : m_used;
      const float f = (kWin && !grow) ? 1.f : fast_exp2(m_used - m_new);  // kWin: m_used may still be -inf
      l *= f;
      m_used = m_new;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t o[32];
        ptx::tmem_ld32(tO + c * 32, o);
        ptx::tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * f);
        ptx::tmem_st32(tO + c * 32, o);
      }
      ptx::tmem_wait_st();
    }
  }
  prof.mark(3);

  const float m_sub = (kWin && kMask && m_used == -INFINITY) ? 0.f : m_used;  // nothing seen yet: every P is exp2(-inf) = 0
  const float2 sc2 = make_float2(scale_log2, scale_log2), nm2 = make_float2(-m_sub, -m_sub);
  float2 sum2[4] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    uint32_t pk[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const int e = h * 64 + 2 * i;
      float2 x = ptx::ffma2(make_float2(__uint_as_float(s[e]), __uint_as_float(s[e + 1])), sc2, nm2);
      if (!kMask && (i & 3) < kEmu) {
        x = ptx::exp2_poly2(x);
      } else {
        x.x = fast_exp2(x.x);
        x.y = fast_exp2(x.y);
      }
      sum2[i & 3] = ptx::fadd2(sum2[i & 3], x);
      pk[i] = pack2<kBf16>(x.x, x.y);
    }
    ptx::tmem_st32(tS + h * 32, pk);
    ptx::tmem_wait_st();
    ptx::tc_fence_before();
    ptx::mbar_arrive(bar_p + 8u * h);
  }
  const float2 st = ptx::fadd2(ptx::fadd2(sum2[0], sum2[1]), ptx::fadd2(sum2[2], sum2[3]));
  l += st.x + st.y;
  prof.mark(4);
}

// kBf16: MMA operand format (bf16 or f16).  kFp8: the pool holds e4m3 bytes; the producers expand them to f16 on
// the way into shared memory (F2FP.F16.E4M3.UNPACK_B, exact; it issues on the ALU pipe and overlaps the MUFU
// exponentials -- scripts/ubench/pipes.cu), so the MMA runs in f16 and kBf16 is false.  kQCvt: q/out are bf16 while
// the MMA runs in f16 (bf16 model over an FP8 KV cache): Q is converted on its way into shared memory.
// kWin: sliding-window layers (p.window >= 0, semantics in extend.cuh): the K/V walk starts at the 128-key tile that
// holds the first prefix key inside the window, and the first tiles carry a per-row lower bound.
struct ExtendTmaps {
  CUtensorMap k, v;
};

// kTma: paged-tile loads (page_size % 16 == 0, 16-bit pools): one producer warp reads one table entry per 16-key page
// and fetches the page's K (V) rows of this kv head as two bulk-tensor boxes of 16 rows x 128 B -- the 128-byte
// swizzle of the copy IS the UMMA canonical layout the row gathers build by hand.
template <bool kBf16, int kEmu, bool kProf, bool kFp8 = false, bool kQCvt = false, bool kWin = false, bool kTma = false>
__global__ void __launch_bounds__(kThreads, 1)
extend_tc_kernel(const ExtendParams p, const int pair_heads, const __grid_constant__ ExtendTmaps tm) {
  static_assert(!(kFp8 && kBf16) && !(kQCvt && kBf16), "FP8 KV runs the MMA in f16");
  static_assert(!(kTma && kFp8), "FP8 pools are expanded in registers: no bulk-tensor path");
  constexpr bool kOutBf16 = kBf16 || kQCvt;  // dtype of q and out in global memory
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;  // 128B swizzle atoms repeat every 1024 B
  const uint32_t bars = smem + kOffBar;
  auto bar = [&](int i) { return bars + 8u * static_cast<uint32_t>(i); };
  const uint32_t tmem_slot = bars + 32 * 8;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  Prof<kProf> prof;
  // KV splits (p.kv_splits > 1, b200attn_extend_split): the request dimension of the grid carries (request, split); a
  // CTA walks only its share of the item's K/V tiles and writes a normalised partial + LSE (merged by
  // extend_split_merge_kernel).  Few long rows -- chunked prefill over a long prefix, head-partitioned ranks --
  // otherwise leave most SMs idle: the grid has one CTA per (request, head pair, 128 q rows).
  const int S = p.kv_splits;
  const int zr = p.grid_order ? blockIdx.z : blockIdx.y;
  const int req = zr / S, split = zr - req * S;
  const int E = p.extend_seq_lens[req];
  // heavy (late) tiles first
  const int zt = p.grid_order ? static_cast<int>(gridDim.y) - 1 - static_cast<int>(blockIdx.y)
                              : static_cast<int>(gridDim.z) - 1 - static_cast<int>(blockIdx.z);
  int head[2], q0[2];
  if (pair_heads) {
    head[0] = 2 * blockIdx.x; head[1] = head[0] + 1;
    q0[0] = q0[1] = zt * kTile;
  } else {
    head[0] = head[1] = blockIdx.x;
    q0[0] = 2 * zt * kTile; q0[1] = q0[0] + kTile;
  }
  if (q0[0] >= E) return;  // CTA-uniform, before any barrier or allocation
  const bool valid1 = q0[1] < E;
  const int P = p.extend_prefix_lens[req];
  const int L = static_cast<int>(load_index(p.seq_lens, p.idx_is_i64, req));
  const int start = p.extend_start_loc[req];
  const int kv_head = head[0] / p.G;
  int kv_end[2], nt[2];
  kv_end[0] = p.causal ? min(L, P + min(q0[0] + kTile, E)) : L;
  kv_end[1] = valid1 ? (p.causal ? min(L, P + min(q0[1] + kTile, E)) : L) : 0;
  // sliding window: first prefix key inside the window, and the tile-aligned key the walk starts at (tile j of this
  // CTA covers keys [kv_base + 128 j, kv_base + 128 (j + 1)))
  int kv_begin = 0, kv_base = 0;
  if constexpr (kWin) {
    kv_begin = P - min(P, p.window + 1);
    kv_base = kv_begin & ~(kTile - 1);
  }
  nt[0] = (kv_end[0] - kv_base + kTile - 1) / kTile;
  nt[1] = valid1 ? (kv_end[1] - kv_base + kTile - 1) / kTile : 0;
  int n = max(nt[0], nt[1]);
  int kv_len = max(kv_end[0], kv_end[1]);
  int64_t out_off = 0, lse_off = 0;  // partial buffers of this split
  if (S > 1) {
    const int j_lo = split * n / S, j_hi = (split + 1) * n / S;  // this split's tiles of the item (CTA-uniform)
    const int hi_key = kv_base + j_hi * kTile;
    kv_base += j_lo * kTile;
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      nt[t] = max(0, min(nt[t], j_hi) - j_lo);
      kv_end[t] = min(kv_end[t], hi_key);
    }
    n = max(nt[0], nt[1]);
    kv_len = max(kv_end[0], kv_end[1]);
    out_off = static_cast<int64_t>(split) * p.split_out_stride;
    lse_off = static_cast<int64_t>(split) * p.split_lse_stride;
    if (n == 0) {  // fewer tiles than splits: this split is empty for every row (CTA-uniform, before any barrier)
      if (tid < 2 * kTile) {
        const int t = tid >> 7, row = tid & (kTile - 1);
        if ((t == 0 || valid1) && q0[t] + row < E)
          p.lse_out[lse_off + static_cast<int64_t>(start + q0[t] + row) * p.Hq + head[t]] = -INFINITY;
      }
      return;
    }
  }

  // ---------------------------------------------------------------- setup
  if (warp == 8) ptx::tmem_alloc<512>(tmem_slot);
  if (tid == 0) {
    ptx::mbar_init(bar(Q_FULL), 256);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(bar(K_FULL + i), kTma ? 1 : kProducerWarps * 32);
      ptx::mbar_init(bar(K_EMPTY + i), 1);
      ptx::mbar_init(bar(V_FULL + i), kTma ? 1 : kProducerWarps * 32);
      ptx::mbar_init(bar(V_EMPTY + i), 1);
      ptx::mbar_init(bar(S_FULL + i), 1);
      ptx::mbar_init(bar(O_FULL + i), 1);
      ptx::mbar_init(bar(P_FULL + 2 * i), 128);
      ptx::mbar_init(bar(P_FULL + 2 * i + 1), 128);
    }
    ptx::fence_mbar_init();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];\n" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp < 8) {
    ptx::setmaxnreg_inc<kRegsSoftmax>();
    // ================================================================ softmax + output of Q tile t
    const int t = warp >> 2;
    const int wq = warp & 3;          // tensor-memory lane quarter of this warp
    const int row = wq * 32 + lane;   // query row of this thread inside the tile
    const int tix = tid - 128 * t;
    const bool tile_valid = (t == 0) || valid1;

    // ---- Q tile -> shared memory, K-major 128B-swizzle layout (two 64-element column halves)
    {
      const uint16_t* qbase = reinterpret_cast<const uint16_t*>(p.q) + static_cast<int64_t>(head[t]) * kTile;
      const uint32_t sq = smem + kOffQ + t * kTileBytes;
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        const int id = tix + 128 * c;
        const int r = id >> 4, ch = id & 15;
        const bool ok = tile_valid && (q0[t] + r < E);
        const uint16_t* src = ok ? qbase + static_cast<int64_t>(start + q0[t] + r) * p.q_tok_stride + ch * 8
                                 : reinterpret_cast<const uint16_t*>(p.q);
        const uint32_t dst = sq + (ch >> 3) * kHalfBytes + r * 128 + (((ch & 7) ^ (r & 7)) << 4);
        if constexpr (kQCvt) {
          uint4 w = ok ? __ldg(reinterpret_cast<const uint4*>(src)) : make_uint4(0u, 0u, 0u, 0u);
          w.x = ptx::bf16x2_to_f16x2(w.x); w.y = ptx::bf16x2_to_f16x2(w.y);
          w.z = ptx::bf16x2_to_f16x2(w.z); w.w = ptx::bf16x2_to_f16x2(w.w);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(dst), "r"(w.x), "r"(w.y), "r"(w.z), "r"(w.w) : "memory");
        } else {
          cp_async16(dst, src, ok ? 16 : 0);
        }
      }
      if constexpr (!kQCvt) {
        cp_async_commit();
        cp_async_wait<0>();
      }
      ptx::fence_proxy_async();  // generic-proxy writes -> visible to the tensor core (async proxy)
      ptx::mbar_arrive(bar(Q_FULL));
    }

    if (tile_valid) {
      const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(wq * 32) << 16);
      const uint32_t tS = lane_base + t * 128;        // S_t, and P_t over its first 64 columns
      const uint32_t tO = lane_base + 256 + t * 128;  // O_t
      const bool row_valid = q0[t] + row < E;
      const int lim = row_valid ? (p.causal ? min(P + q0[t] + row + 1, L) : L) : L;  // keys visible to this row
      const int lim_row0 = p.causal ? min(P + q0[t] + 1, L) : L;                     // the tile's smallest limit
      float m_used = 0.f, l = 0.f;
      const int my_n = nt[t];
      if (my_n == 0) {  // (splits only) this Q tile has no key in this split's range: empty partial, no MMA will arrive
        if (row_valid) p.lse_out[lse_off + static_cast<int64_t>(start + q0[t] + row) * p.Hq + head[t]] = -INFINITY;
      } else {
      const int n_unmasked = min(my_n, (lim_row0 - kv_base) / kTile);  // tiles every row of this Q tile sees completely
      const uint32_t bar_s = bar(S_FULL + t), bar_p = bar(P_FULL + 2 * t);
      // window: prefix keys below lo are invisible to this row (new-token keys, >= P, have no lower bound); the
      // first n_lead tiles hold such a key for some row of the Q tile
      int lo = 0, n_lead = 0;
      if constexpr (kWin) {
        const int tq = q0[t] + row;
        lo = (p.window > 0 ? min(kv_begin + max(0, tq - p.window), P) : kv_begin) - kv_base;
        const int lo_max = (p.window > 0 ? min(kv_begin + max(0, q0[t] + kTile - 1 - p.window), P) : kv_begin) - kv_base;
        n_lead = min(my_n, (lo_max + kTile - 1) / kTile);
      }
      const int lim_rel = lim - kv_base;

      prof.begin();
      int j = 0;
      if constexpr (kWin) {
#pragma unroll 1
        for (; j < n_lead; ++j)
          softmax_tile<kBf16, true, kEmu, kProf, true>(tS, tO, bar_s, bar_p, j, lim_rel - j * kTile, p.scale_log2, m_used,
                                                       l, prof, lo - j * kTile);
      }
#pragma unroll 1
      for (; j < n_unmasked; ++j)
        softmax_tile<kBf16, false, kEmu, kProf, kWin>(tS, tO, bar_s, bar_p, j, kTile, p.scale_log2, m_used, l, prof);
#pragma unroll 1
      for (; j < my_n; ++j)
        softmax_tile<kBf16, true, kEmu, kProf, kWin>(tS, tO, bar_s, bar_p, j, lim_rel - j * kTile, p.scale_log2, m_used, l,
                                                     prof, lo - j * kTile);

      // ---- epilogue: O_t / l -> global
      ptx::mbar_wait(bar(O_FULL + t), 0);
      ptx::tc_fence_after();
      prof.mark(5);
      const float inv = l > 0.f ? p.v_scale / l : 0.f;
      uint16_t* dst = reinterpret_cast<uint16_t*>(p.out) + out_off +
                      static_cast<int64_t>(start + q0[t] + row) * p.out_tok_stride + static_cast<int64_t>(head[t]) * kTile;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t o[32];
        ptx::tmem_ld32(tO + c * 32, o);
        ptx::tmem_wait_ld();
        if (row_valid) {
#pragma unroll
          for (int v4 = 0; v4 < 4; ++v4) {
            uint4 w;
            w.x = pack2<kOutBf16>(__uint_as_float(o[8 * v4 + 0]) * inv, __uint_as_float(o[8 * v4 + 1]) * inv);
            w.y = pack2<kOutBf16>(__uint_as_float(o[8 * v4 + 2]) * inv, __uint_as_float(o[8 * v4 + 3]) * inv);
            w.z = pack2<kOutBf16>(__uint_as_float(o[8 * v4 + 4]) * inv, __uint_as_float(o[8 * v4 + 5]) * inv);
            w.w = pack2<kOutBf16>(__uint_as_float(o[8 * v4 + 6]) * inv, __uint_as_float(o[8 * v4 + 7]) * inv);
            *reinterpret_cast<uint4*>(dst + c * 32 + v4 * 8) = w;
          }
        }
      }
      if (p.lse_out != nullptr && row_valid)
        p.lse_out[lse_off + static_cast<int64_t>(start + q0[t] + row) * p.Hq + head[t]] =
            l > 0.f ? (m_used + log2f(l)) * kLn2 : -INFINITY;
      prof.mark(6);
      prof.flush(p.prof, warp, lane);
      ptx::tc_fence_before();
      }
    }
  } else if (warp == 8) {
    ptx::setmaxnreg_dec<kRegsOther>();
    // ================================================================ MMA issuer
    // The whole warp walks the loop (mbarrier waits included) so that every operand stays warp-uniform; one
    // elected lane -- the same one throughout, tcgen05.commit tracks the MMAs of the issuing thread -- issues.
    const bool leader = ptx::elect_one();
    constexpr int kFmt = kBf16 ? 1 : 0;
    constexpr uint32_t idesc_qk = ptx::instr_desc_f16(kFmt, 128, 128, 0);  // A = Q, B = K: both K-major
    constexpr uint32_t idesc_pv = ptx::instr_desc_f16(kFmt, 128, 128, 1);  // A = P (tmem), B = V: MN-major
    // K-major SW128: 8-row groups 1024 B apart; a 16-element k step is 32 B inside the 128 B atom row,
    // the second 64-element half of D lives in the next 16 KiB sub-tile.
    // MN-major SW128 (V: rows = keys = the contraction index, 128 B of D per row): 8-key groups 1024 B apart
    // (SBO), the two 64-element halves of D 16 KiB apart (LBO); a 16-key k step is 2048 B.
    const uint64_t qd0 = ptx::smem_desc_sw128(smem + kOffQ, 16, 1024);
    const uint64_t kd0 = ptx::smem_desc_sw128(smem + kOffK, 16, 1024);
    const uint64_t vd0 = ptx::smem_desc_sw128(smem + kOffV, kHalfBytes, 1024);
    constexpr uint64_t kTileStep = kTileBytes >> 4;  // descriptor address units of 16 B

    auto issue_qk = [&](int t, int kb) {
      const uint64_t qd = qd0 + static_cast<uint64_t>(t) * kTileStep;
      const uint64_t kd = kd0 + static_cast<uint64_t>(kb) * kTileStep;
      const uint32_t d = tmem_base + t * 128;
      if (leader) {
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint64_t off = static_cast<uint64_t>(((kk >> 2) * kHalfBytes + (kk & 3) * 32) >> 4);
          ptx::umma_ss(d, qd + off, kd + off, idesc_qk, kk > 0 ? 1u : 0u);
        }
        ptx::umma_commit(bar(S_FULL + t));
      }
      __syncwarp();
    };
    // four k-steps (64 keys) of O_t += P_t V
    auto issue_pv_half = [&](int t, int vb, int h, bool acc) {
      const uint64_t vd = vd0 + static_cast<uint64_t>(vb) * kTileStep + static_cast<uint64_t>(h) * ((4 * 2048) >> 4);
      const uint32_t d = tmem_base + 256 + t * 128;
      const uint32_t a = tmem_base + t * 128 + h * 32;
      if (leader) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          ptx::umma_ts(d, a + kk * 8, vd + static_cast<uint64_t>((kk * 2048) >> 4), idesc_pv, (acc || kk > 0) ? 1u : 0u);
      }
      __syncwarp();
    };
    auto commit = [&](int b) {
      if (leader) ptx::umma_commit(bar(b));
      __syncwarp();
    };

    prof.begin();
    ptx::mbar_wait(bar(Q_FULL), 0);
    ptx::mbar_wait(bar(K_FULL + 0), 0);
    ptx::fence_proxy_async();  // K/V arrive through cp.async (generic proxy); the tensor core reads through the async proxy
    ptx::tc_fence_after();
    prof.mark(0);
    for (int t = 0; t < 2; ++t)
      if (nt[t] > 0) issue_qk(t, 0);
    commit(K_EMPTY + 0);
    prof.mark(7);

#pragma unroll 1
    for (int j = 0; j < n; ++j) {
      const int vb = j & 1, kb = (j + 1) & 1;
      bool v_ready = false, k_ready = false;
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        if (j < nt[t]) {
          ptx::mbar_wait(bar(P_FULL + 2 * t), j & 1);
          prof.mark(1 + t);
          if (!v_ready) {
            ptx::mbar_wait(bar(V_FULL + vb), (j >> 1) & 1);
            ptx::fence_proxy_async();
            v_ready = true;
            prof.mark(3);
          }
          ptx::tc_fence_after();
          issue_pv_half(t, vb, 0, j > 0);
          prof.mark(7);
          ptx::mbar_wait(bar(P_FULL + 2 * t + 1), j & 1);
          prof.mark(5);
          ptx::tc_fence_after();
          issue_pv_half(t, vb, 1, true);
          if (j == nt[t] - 1) commit(O_FULL + t);
          prof.mark(7);
        }
        if (j + 1 < nt[t]) {
          if (!k_ready) {
            ptx::mbar_wait(bar(K_FULL + kb), ((j + 1) >> 1) & 1);
            ptx::fence_proxy_async();
            k_ready = true;
            prof.mark(4);
          }
          ptx::tc_fence_after();
          issue_qk(t, kb);
          prof.mark(7);
        }
      }
      commit(V_EMPTY + vb);
      if (j + 1 < n) commit(K_EMPTY + kb);
      prof.mark(7);
    }
    prof.flush(p.prof, warp, lane);
  } else {
    ptx::setmaxnreg_dec<kRegsOther>();
    // ================================================================ producers: gather K/V rows through the page table
    // A warp instruction copies two pool rows of this kv head (2 x 256 B, 16 lanes x 16 B each).  The 128 rows of a
    // tile are 16 groups of 8 (one swizzle period); producer warp pw owns groups m = pw, pw + 3, ... and walks a
    // group as four row pairs q, so the swizzled destination of (group k', pair q) is base_q + k' * 3072: four
    // loop-invariant registers instead of one per copy.
    const int pw = warp - 9;
    const int32_t* trow = p.req_to_token + load_index(p.req_pool_indices, p.idx_is_i64, req) * p.table_stride;
    constexpr int kGroups = 6;                       // groups per producer warp (the last one only for pw == 0)
    // Page-table slots: lane i of load c holds the slot of row 8 * (pw + 3 * (4 c + i / 8)) + i % 8 -- the rows of
    // this warp only, two coalesced-by-eight loads per tile, fetched one tile ahead.
    int sreg[2];
    auto load_slots = [&](int j) {
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int m = pw + 3 * (4 * c + (lane >> 3));
        const int tok = kv_base + j * kTile + 8 * m + (lane & 7);
        sreg[c] = (m < 16 && tok < kv_len) ? __ldg(trow + tok) : 0;  // past the end: the dummy slot (masked later)
      }
    };
    if constexpr (kTma) {
      if (pw == 0) {
        const int64_t row_bytes = p.pool_tok_stride * 2;
        const uint8_t* kbase = reinterpret_cast<const uint8_t*>(p.k_pool) + static_cast<int64_t>(kv_head) * (kTile * 2);
        const uint8_t* vbase = reinterpret_cast<const uint8_t*>(p.v_pool) + static_cast<int64_t>(kv_head) * (kTile * 2);
        const int seg0 = 2 * kv_head;  // a 256-byte head row = two 128-byte segments of the pool row
        // lane i < 8 owns page i of a tile: the slot of its first row, fetched one tile ahead
        auto page_slot = [&](int j) {
          const int tok = kv_base + j * kTile + 16 * lane;
          return (lane < 8 && tok < kv_len) ? __ldg(trow + tok) : 0;
        };
        // Pages that lie completely below kv_len arrive as boxes.  The rest of the tile (the ragged last page and
        // everything behind it -- only the last tile of a CTA has any) is written through the generic proxy: real rows
        // copied, rows past kv_len zeroed, so that neither another request's rows nor stale shared memory (0 * NaN)
        // can reach O.  One arrival covers both: the stores are fenced before expect_tx.
        auto load_tile = [&](const CUtensorMap* map, const uint8_t* base, uint32_t tile, uint32_t full, int slot, int j) {
          const int tok0 = kv_base + j * kTile;
          const int nvalid = min(kTile, kv_len - tok0);
          const int n_full = nvalid >> 4;
          if (n_full < 8) {
            for (int i = n_full * 256 + lane; i < kTile * 16; i += 32) {
              const int r = i >> 4, ch = i & 15;
              uint4 w = make_uint4(0u, 0u, 0u, 0u);
              if (r < nvalid) w = ld_nc_16(base + static_cast<int64_t>(__ldg(trow + tok0 + r)) * row_bytes + ch * 16);
              const uint32_t dst = tile + (ch >> 3) * kHalfBytes + r * 128 + (((ch & 7) ^ (r & 7)) << 4);
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(dst), "r"(w.x), "r"(w.y), "r"(w.z), "r"(w.w) : "memory");
            }
            ptx::fence_proxy_async();
          }
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive_expect_tx(full, static_cast<uint32_t>(n_full) * 4096u);
          __syncwarp();
          if (lane < n_full) {
            ptx::tma_load_3d(tile + lane * 2048, map, 0, slot, seg0, full);
            ptx::tma_load_3d(tile + kHalfBytes + lane * 2048, map, 0, slot, seg0 + 1, full);
          }
        };
        int next = page_slot(0);
        prof.begin();
#pragma unroll 1
        for (int j = 0; j < n; ++j) {
          const int b = j & 1, u = j >> 1;
          const int slot = next;
          if (j + 1 < n) next = page_slot(j + 1);
          ptx::mbar_wait(bar(K_EMPTY + b), (u & 1) ^ 1);
          prof.mark(1);
          load_tile(&tm.k, kbase, smem + kOffK + b * kTileBytes, bar(K_FULL + b), slot, j);
          prof.mark(2);
          ptx::mbar_wait(bar(V_EMPTY + b), (u & 1) ^ 1);
          prof.mark(3);
          load_tile(&tm.v, vbase, smem + kOffV + b * kTileBytes, bar(V_FULL + b), slot, j);
          prof.mark(4);
        }
      }
    } else if constexpr (!kFp8) {
      const int64_t row_bytes = p.pool_tok_stride * 2;
      const uint8_t* kbase = reinterpret_cast<const uint8_t*>(p.k_pool) + static_cast<int64_t>(kv_head) * (kTile * 2) + (lane & 15) * 16;
      const uint8_t* vbase = reinterpret_cast<const uint8_t*>(p.v_pool) + static_cast<int64_t>(kv_head) * (kTile * 2) + (lane & 15) * 16;
      const int sub = lane >> 4;                       // which row of the pair
      const int ch = lane & 15;                        // 16-byte chunk of the 256-byte row
      uint32_t dst_q[4];                               // destination of (group pw, pair q) relative to the tile base
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int r = 8 * pw + 2 * q + sub;
        dst_q[q] = (ch >> 3) * kHalfBytes + r * 128 + (((ch & 7) ^ (r & 7)) << 4);
      }
      // the copy of (group k', pair q) takes its slot from lane 8 * (k' % 4) + 2 q + sub of register k' / 4
      int slot[4 * kGroups];
      auto spread_slots = [&]() {
#pragma unroll
        for (int kq = 0; kq < 4 * kGroups; ++kq) {
          const int kp = kq >> 2, q = kq & 3;
          slot[kq] = __shfl_sync(0xffffffffu, sreg[kp >> 2], 8 * (kp & 3) + 2 * q + sub);
        }
      };
      auto gather = [&](const uint8_t* base, uint32_t tile) {
#pragma unroll
        for (int kq = 0; kq < 4 * kGroups; ++kq) {
          const int kp = kq >> 2, q = kq & 3;
          if (pw + 3 * kp < 16)  // warp-uniform
            cp_async16(tile + dst_q[q] + kp * 3072, base + static_cast<int64_t>(slot[kq]) * row_bytes);
        }
      };
      load_slots(0);
      prof.begin();
#pragma unroll 1
      for (int j = 0; j < n; ++j) {
        const int b = j & 1, u = j >> 1;
        spread_slots();
        if (j + 1 < n) load_slots(j + 1);  // issued before this tile's copies so that it never queues behind them
        prof.mark(0);
        // K_j
        ptx::mbar_wait(bar(K_EMPTY + b), (u & 1) ^ 1);
        prof.mark(1);
        gather(kbase, smem + kOffK + b * kTileBytes);
        ptx::cp_async_mbar_arrive_noinc(bar(K_FULL + b));
        prof.mark(2);
        // V_j
        ptx::mbar_wait(bar(V_EMPTY + b), (u & 1) ^ 1);
        prof.mark(3);
        gather(vbase, smem + kOffV + b * kTileBytes);
        ptx::cp_async_mbar_arrive_noinc(bar(V_FULL + b));
        prof.mark(4);
      }
    } else {
      // FP8-E4M3 pool: a row of one kv head is 128 bytes; a warp instruction loads four rows (8 lanes x 16 B each),
      // every lane expands its 16 e4m3 values to 16 f16 (two 16-byte chunks of the 256-byte shared-memory row) and
      // stores them into the same UMMA canonical 128B-swizzle layout the 16-bit path fills with LDGSTS.  The loads of
      // a half tile (three 8-row groups = six warp loads) are all issued before the first conversion.  Register
      // stores go through the generic proxy: fence.proxy.async, then a plain mbarrier arrive.
      const int64_t row_bytes = p.pool_tok_stride;
      const uint8_t* kbase = reinterpret_cast<const uint8_t*>(p.k_pool) + static_cast<int64_t>(kv_head) * kTile + (lane & 7) * 16;
      const uint8_t* vbase = reinterpret_cast<const uint8_t*>(p.v_pool) + static_cast<int64_t>(kv_head) * kTile + (lane & 7) * 16;
      const int sub = lane >> 3;                       // which row of the quad
      const int c0 = 2 * (lane & 7);                   // first of the two 16-byte f16 chunks this lane produces
      uint32_t dst_q[2];                               // destination of (group pw, quad q), chunk c0; chunk c0+1 is ^ 16
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int r = 8 * pw + 4 * q + sub;
        dst_q[q] = (c0 >> 3) * kHalfBytes + r * 128 + (((c0 & 7) ^ (r & 7)) << 4);
      }
      int scur[2];  // this tile's slots (sreg holds the next tile's); spread by shuffle at the point of use
      // the twelve warp loads of a tile share are issued before the wait for the empty buffer: registers do not
      // depend on shared memory, so the L2 / HBM latency overlaps the MMAs still reading the previous tile
      auto issue_loads = [&](const uint8_t* base, uint4 (&w)[12]) {
#pragma unroll
        for (int kq = 0; kq < 12; ++kq) {
          const int kp = kq >> 1;
          if (pw + 3 * kp < 16) {  // warp-uniform
            const int sl = __shfl_sync(0xffffffffu, scur[kp >> 2], 8 * (kp & 3) + 4 * (kq & 1) + sub);
            w[kq] = ld_nc_16(base + static_cast<int64_t>(sl) * row_bytes);
          }
        }
      };
      auto expand_store = [&](uint32_t tile, const uint4 (&w)[12]) {
#pragma unroll
        for (int kq = 0; kq < 12; ++kq) {
          const int kp = kq >> 1, q = kq & 1;
          if (pw + 3 * kp < 16) {
            const uint32_t d = tile + dst_q[q] + kp * 3072;
            const uint32_t in[4] = {w[kq].x, w[kq].y, w[kq].z, w[kq].w};
            uint32_t o[8];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              o[2 * e] = e4m3x2_to_f16x2(static_cast<uint16_t>(in[e] & 0xffffu));
              o[2 * e + 1] = e4m3x2_to_f16x2(static_cast<uint16_t>(in[e] >> 16));
            }
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(d), "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]) : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(d ^ 16u), "r"(o[4]), "r"(o[5]), "r"(o[6]), "r"(o[7]) : "memory");
          }
        }
      };
      load_slots(0);
      prof.begin();
      uint4 w[12];
      scur[0] = sreg[0]; scur[1] = sreg[1];
      if (n > 0) issue_loads(kbase, w);
#pragma unroll 1
      for (int j = 0; j < n; ++j) {
        const int b = j & 1, u = j >> 1;
        if (j + 1 < n) load_slots(j + 1);
        ptx::mbar_wait(bar(K_EMPTY + b), (u & 1) ^ 1);
        expand_store(smem + kOffK + b * kTileBytes, w);
        issue_loads(vbase, w);  // V_j on its way while K_j is published
        ptx::fence_proxy_async();
        ptx::mbar_arrive(bar(K_FULL + b));
        ptx::mbar_wait(bar(V_EMPTY + b), (u & 1) ^ 1);
        expand_store(smem + kOffV + b * kTileBytes, w);
        scur[0] = sreg[0]; scur[1] = sreg[1];
        if (j + 1 < n) issue_loads(kbase, w);  // K_{j+1}
        ptx::fence_proxy_async();
        ptx::mbar_arrive(bar(V_FULL + b));
      }
    }
    prof.flush(p.prof, warp, lane);
  }

  // ---------------------------------------------------------------- teardown
  __syncthreads();
  if (warp == 8) {
    __syncwarp();
    ptx::tc_fence_after();
    ptx::tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace tc

bool extend_tc_supported(const ExtendParams& p, int D, int kvb) {
  return D == 128 && (kvb == 2 || kvb == 1) && p.logit_cap == 0.f && p.Hq >= 1 && (p.pool_tok_stride * kvb) % 16 == 0;
}

namespace {
using KernelFn = void (*)(const ExtendParams, const int, const tc::ExtendTmaps);

// B200ATTN_EXTEND_TMA=0 keeps page_size >= 16 pools on the row-gather producers (for A/B timing; default 1)
bool extend_tma_enabled() {
  static const bool on = [] { const char* e = getenv("B200ATTN_EXTEND_TMA"); return e ? atoi(e) != 0 : true; }();
  return on;
}

template <bool kBf16>
KernelFn pick_kernel(int emu, bool prof) {
  if (prof) return tc::extend_tc_kernel<kBf16, 2, true>;
  switch (emu) {
    case 0: return tc::extend_tc_kernel<kBf16, 0, false>;
    case 1: return tc::extend_tc_kernel<kBf16, 1, false>;
    case 3: return tc::extend_tc_kernel<kBf16, 3, false>;
    default: return tc::extend_tc_kernel<kBf16, 2, false>;
  }
}
}  // namespace

int extend_tc_launch(const ExtendParams& p, bool q_bf16, int kvb, cudaStream_t st) {
  const int emu = (p.exp_emu >= 0 && p.exp_emu <= 3) ? p.exp_emu : 2;
  KernelFn kern;
  static const tc::ExtendTmaps no_maps{};
  tc::ExtendTmaps maps = no_maps;
  const bool tma = kvb == 2 && p.page_size >= 16 && p.page_size % 16 == 0 && p.prof == nullptr && extend_tma_enabled();
  if (tma) {
    const int64_t row_bytes = p.pool_tok_stride * 2;
    int rc = encode_pool_tmap(&maps.k, p.k_pool, 2, row_bytes, p.Hkv * 2, 16, 1);
    if (rc != B200ATTN_OK) return rc;
    rc = encode_pool_tmap(&maps.v, p.v_pool, 2, row_bytes, p.Hkv * 2, 16, 1);
    if (rc != B200ATTN_OK) return rc;
    if (p.window >= 0)
      kern = q_bf16 ? tc::extend_tc_kernel<true, 0, false, false, false, true, true>
                    : tc::extend_tc_kernel<false, 0, false, false, false, true, true>;
    else
      kern = q_bf16 ? tc::extend_tc_kernel<true, 0, false, false, false, false, true>
                    : tc::extend_tc_kernel<false, 0, false, false, false, false, true>;
  } else if (p.window >= 0) {
    if (kvb == 1)
      kern = q_bf16 ? tc::extend_tc_kernel<false, 0, false, true, true, true>
                    : tc::extend_tc_kernel<false, 0, false, true, false, true>;
    else
      kern = q_bf16 ? tc::extend_tc_kernel<true, 0, false, false, false, true>
                    : tc::extend_tc_kernel<false, 0, false, false, false, true>;
  } else if (kvb == 1)  // FP8-E4M3 pool: f16 MMA, e4m3 expanded by the producers; bf16 q converted on load
    kern = q_bf16 ? tc::extend_tc_kernel<false, 0, false, true, true> : tc::extend_tc_kernel<false, 0, false, true, false>;
  else
    kern = q_bf16 ? pick_kernel<true>(emu, p.prof != nullptr) : pick_kernel<false>(emu, p.prof != nullptr);
  // opt in to the large dynamic shared memory once per kernel instance (cheap host call, no device sync)
  B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::kSmemBytes));
  const int pair_heads = (p.G % 2 == 0) ? 1 : 0;
  const int q_tiles = (p.max_extend_len + tc::kTile - 1) / tc::kTile;
  const int nz = pair_heads ? q_tiles : (q_tiles + 1) / 2;
  const int zs = p.bs * p.kv_splits;  // (request, KV split)
  dim3 grid(pair_heads ? p.Hq / 2 : p.Hq, p.grid_order ? nz : zs, p.grid_order ? zs : nz);
  B2_REQUIRE(grid.y <= 65535 && grid.z <= 65535, B200ATTN_ERR_UNSUPPORTED, "extend: batch %d or %d q tiles exceed the grid",
             p.bs, q_tiles);
  kern<<<grid, tc::kThreads, tc::kSmemBytes, st>>>(p, pair_heads, maps);
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}

}  // namespace b200attn
