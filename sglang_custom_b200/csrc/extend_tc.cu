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
//     touches shared memory.
//   * K/V tiles of 128 tokens are gathered row by row through req_to_token by three producer warps with 16-byte
//     cp.async (LDGSTS) straight into the UMMA canonical 128B-swizzle layout; completion is counted on an
//     mbarrier with cp.async.mbarrier.arrive.noinc, so the producers never wait for their own copies -- no
//     flattened kv_indices copy and no register staging.  Two K and two V buffers of 32 KiB form the ring.
//     (The first cut used TMA gather4, four rows x 128 B per instruction: measured 63 cycles per instruction,
//     i.e. 8 000 cycles per 128-key K+V tile against a 2 048-cycle MMA budget -- profiles/README.md -- so the
//     page_size = 1 gather moved to LDGSTS; tma_gather4 stays in tc_ptx.cuh for the paged-tile path.)
//   * warp roles (three warpgroups, registers re-balanced with setmaxnreg): warps 0-3 / 4-7 = softmax + output
//     of Q tile 0 / 1 (one query row per thread, fp32 online softmax in the log2 domain on packed f32x2 FMAs, the
//     O rescale is lazy: only when the row maximum grows by more than 2^8), warp 8 = MMA issuer (one thread),
//     warps 9-11 = producers.  All hand-offs are mbarriers; tcgen05.commit tracks MMA completion.  Every wait
//     is bounded (ptx::mbar_wait traps instead of hanging).
#include "extend.cuh"
#include "tc_ptx.cuh"

namespace b200attn {
namespace tc {

constexpr int kThreads = 384;       // 3 warpgroups
constexpr int kProducerWarps = 3;
constexpr int kRegsSoftmax = 208;   // setmaxnreg targets: 2 x 128 x 208 + 128 x 80 <= 64 Ki registers
constexpr int kRegsOther = 80;
constexpr int kTile = 128;           // BM = BN = D = 128
constexpr int kTileBytes = 32768;    // 128 rows x 256 B
constexpr int kHalfBytes = 16384;    // 128 rows x 128 B (one 64-element swizzle atom column)
constexpr int kOffQ = 0;             // Q0, Q1
constexpr int kOffK = 2 * kTileBytes;
constexpr int kOffV = 4 * kTileBytes;
constexpr int kOffBar = 6 * kTileBytes;
constexpr int kSmemBytes = kOffBar + 16 * 8 + 16 + 1024;  // + barriers + tmem slot + alignment slack
constexpr float kRescaleThreshold = 8.0f;                 // log2 units

enum Bar { Q_FULL = 0, K_FULL = 1, K_EMPTY = 3, V_FULL = 5, V_EMPTY = 7, S_FULL = 9, P_FULL = 11, O_FULL = 13 };

// dev-only phase timing: CTA (0,0,0) accumulates clock64 deltas into p.prof[slot]
#define TC_PROF_BEGIN() long long prof_t0 = prof_on ? clock64() : 0
#define TC_PROF_MARK(slot)                                                          \
  do {                                                                              \
    if (prof_on) {                                                                  \
      const long long now__ = clock64();                                            \
      prof_acc[slot] += static_cast<unsigned long long>(now__ - prof_t0);           \
      prof_t0 = now__;                                                              \
    }                                                                               \
  } while (0)

template <bool kBf16>
__global__ void __launch_bounds__(kThreads, 1)
extend_tc_kernel(const ExtendParams p, const int pair_heads) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;  // 128B swizzle atoms repeat every 1024 B
  const uint32_t bars = smem + kOffBar;
  auto bar = [&](int i) { return bars + 8u * static_cast<uint32_t>(i); };
  const uint32_t tmem_slot = bars + 16 * 8;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool prof_on = p.prof != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
  unsigned long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const int req = blockIdx.y;
  const int E = p.extend_seq_lens[req];
  const int zt = static_cast<int>(gridDim.z) - 1 - static_cast<int>(blockIdx.z);  // heavy (late) tiles first
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
  nt[0] = (kv_end[0] + kTile - 1) / kTile;
  nt[1] = (kv_end[1] + kTile - 1) / kTile;
  const int n = max(nt[0], nt[1]);
  const int kv_len = max(kv_end[0], kv_end[1]);

  // ---------------------------------------------------------------- setup
  if (warp == 8) ptx::tmem_alloc<512>(tmem_slot);
  if (tid == 0) {
    ptx::mbar_init(bar(Q_FULL), 256);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(bar(K_FULL + i), kProducerWarps * 32);
      ptx::mbar_init(bar(K_EMPTY + i), 1);
      ptx::mbar_init(bar(V_FULL + i), kProducerWarps * 32);
      ptx::mbar_init(bar(V_EMPTY + i), 1);
      ptx::mbar_init(bar(S_FULL + i), 1);
      ptx::mbar_init(bar(P_FULL + i), 128);
      ptx::mbar_init(bar(O_FULL + i), 1);
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
        cp_async16(dst, src, ok ? 16 : 0);
      }
      cp_async_commit();
      cp_async_wait<0>();
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

      TC_PROF_BEGIN();
      for (int j = 0; j < my_n; ++j) {
        ptx::mbar_wait(bar(S_FULL + t), j & 1);
        ptx::tc_fence_after();
        TC_PROF_MARK(0);
        uint32_t s[4][32];
#pragma unroll
        for (int c = 0; c < 4; ++c) ptx::tmem_ld32(tS + c * 32, s[c]);
        ptx::tmem_wait_ld();
        TC_PROF_MARK(1);

        const int kv0 = j * kTile;
        float mx8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) mx8[i] = -INFINITY;
        if (kv0 + kTile > lim_row0) {  // diagonal or last tile: mask (warp-uniform branch)
          const int nvis = lim - kv0;
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              float x = __uint_as_float(s[c][i]);
              x = (c * 32 + i < nvis) ? x : -INFINITY;
              s[c][i] = __float_as_uint(x);
              mx8[i & 7] = fmaxf(mx8[i & 7], x);
            }
        } else {
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) mx8[i & 7] = fmaxf(mx8[i & 7], __uint_as_float(s[c][i]));
        }
        float mx = fmaxf(fmaxf(fmaxf(mx8[0], mx8[1]), fmaxf(mx8[2], mx8[3])),
                         fmaxf(fmaxf(mx8[4], mx8[5]), fmaxf(mx8[6], mx8[7])));
        mx *= p.scale_log2;  // scale > 0: the maximum commutes with the scaling
        TC_PROF_MARK(2);

        if (j == 0) {
          m_used = (mx == -INFINITY) ? 0.f : mx;
        } else {
          const bool grow = mx > m_used + kRescaleThreshold;
          if (__any_sync(0xffffffffu, grow)) {
            // S_t(j) is complete, so PV_t(j-1) -- issued before it -- is complete too: O_t is stable until this
            // thread group hands P_t(j) to the MMA warp
            const float m_new = grow ? mx : m_used;
            const float f = fast_exp2(m_used - m_new);
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

        TC_PROF_MARK(3);
        const float2 sc2 = make_float2(p.scale_log2, p.scale_log2), nm2 = make_float2(-m_used, -m_used);
        float2 sum2[4] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t pk[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int c = h * 2 + (i >> 4), e = (i & 15) * 2;
            float2 x = ptx::ffma2(make_float2(__uint_as_float(s[c][e]), __uint_as_float(s[c][e + 1])), sc2, nm2);
            if (!(p.debug & 2)) { x.x = fast_exp2(x.x); x.y = fast_exp2(x.y); }
            sum2[i & 3] = ptx::fadd2(sum2[i & 3], x);
            pk[i] = pack2<kBf16>(x.x, x.y);
          }
          ptx::tmem_st32(tS + h * 32, pk);
        }
        const float2 st = ptx::fadd2(ptx::fadd2(sum2[0], sum2[1]), ptx::fadd2(sum2[2], sum2[3]));
        const float sum = st.x + st.y;
        ptx::tmem_wait_st();
        l += sum;
        ptx::tc_fence_before();
        ptx::mbar_arrive(bar(P_FULL + t));
        TC_PROF_MARK(4);
      }

      // ---- epilogue: O_t / l -> global
      ptx::mbar_wait(bar(O_FULL + t), 0);
      ptx::tc_fence_after();
      TC_PROF_MARK(5);
      const float inv = l > 0.f ? p.v_scale / l : 0.f;
      uint16_t* dst = reinterpret_cast<uint16_t*>(p.out) +
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
            w.x = pack2<kBf16>(__uint_as_float(o[8 * v4 + 0]) * inv, __uint_as_float(o[8 * v4 + 1]) * inv);
            w.y = pack2<kBf16>(__uint_as_float(o[8 * v4 + 2]) * inv, __uint_as_float(o[8 * v4 + 3]) * inv);
            w.z = pack2<kBf16>(__uint_as_float(o[8 * v4 + 4]) * inv, __uint_as_float(o[8 * v4 + 5]) * inv);
            w.w = pack2<kBf16>(__uint_as_float(o[8 * v4 + 6]) * inv, __uint_as_float(o[8 * v4 + 7]) * inv);
            *reinterpret_cast<uint4*>(dst + c * 32 + v4 * 8) = w;
          }
        }
      }
      if (p.lse_out != nullptr && row_valid)
        p.lse_out[static_cast<int64_t>(start + q0[t] + row) * p.Hq + head[t]] =
            l > 0.f ? (m_used + log2f(l)) * kLn2 : -INFINITY;
      TC_PROF_MARK(6);
      if (prof_on && lane == 0)
        for (int i = 0; i < 8; ++i) p.prof[warp * 8 + i] = prof_acc[i];
      ptx::tc_fence_before();
    }
  } else if (warp == 8) {
    ptx::setmaxnreg_dec<kRegsOther>();
    // ================================================================ MMA issuer (one thread)
    if (lane == 0) {
      constexpr int kFmt = kBf16 ? 1 : 0;
      constexpr uint32_t idesc_qk = ptx::instr_desc_f16(kFmt, 128, 128, 0);  // A = Q, B = K: both K-major
      constexpr uint32_t idesc_pv = ptx::instr_desc_f16(kFmt, 128, 128, 1);  // A = P (tmem), B = V: MN-major
      const uint32_t tS[2] = {tmem_base, tmem_base + 128};
      const uint32_t tO[2] = {tmem_base + 256, tmem_base + 384};
      // K-major SW128: 8-row groups 1024 B apart; a 16-element k step is 32 B inside the 128 B atom row,
      // the second 64-element half of D lives in the next 16 KiB sub-tile
      auto desc_k_major = [&](uint32_t tile) { return ptx::smem_desc_sw128(tile, 16, 1024); };
      // MN-major SW128 (V: rows = keys = the contraction index, 128 B of D per row): 8-key groups 1024 B apart
      // (SBO), the two 64-element halves of D 16 KiB apart (LBO); a 16-key k step is 2048 B
      auto desc_mn_major = [&](uint32_t tile) { return ptx::smem_desc_sw128(tile, kHalfBytes, 1024); };
      const uint64_t qd[2] = {desc_k_major(smem + kOffQ), desc_k_major(smem + kOffQ + kTileBytes)};
      const uint64_t kd[2] = {desc_k_major(smem + kOffK), desc_k_major(smem + kOffK + kTileBytes)};
      const uint64_t vd[2] = {desc_mn_major(smem + kOffV), desc_mn_major(smem + kOffV + kTileBytes)};

      auto issue_qk = [&](int t, int kb) {
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint64_t off = static_cast<uint64_t>(((kk >> 2) * kHalfBytes + (kk & 3) * 32) >> 4);
          ptx::umma_ss(tS[t], qd[t] + off, kd[kb] + off, idesc_qk, kk > 0 ? 1u : 0u);
        }
        ptx::umma_commit(bar(S_FULL + t));
      };
      auto issue_pv = [&](int t, int vb, bool acc) {
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint64_t off = static_cast<uint64_t>((kk * 2048) >> 4);
          ptx::umma_ts(tO[t], tS[t] + kk * 8, vd[vb] + off, idesc_pv, (acc || kk > 0) ? 1u : 0u);
        }
      };

      TC_PROF_BEGIN();
      ptx::mbar_wait(bar(Q_FULL), 0);
      ptx::mbar_wait(bar(K_FULL + 0), 0);
      ptx::fence_proxy_async();  // K/V arrive through cp.async (generic proxy); the tensor core reads through the async proxy
      ptx::tc_fence_after();
      TC_PROF_MARK(0);
      for (int t = 0; t < 2; ++t)
        if (nt[t] > 0) issue_qk(t, 0);
      ptx::umma_commit(bar(K_EMPTY + 0));

      for (int j = 0; j < n; ++j) {
        const int vb = j & 1, kb = (j + 1) & 1;
        bool v_ready = false, k_ready = false;
        for (int t = 0; t < 2; ++t) {
          if (j < nt[t]) {
            TC_PROF_MARK(7);
            ptx::mbar_wait(bar(P_FULL + t), j & 1);
            TC_PROF_MARK(1 + t);
            if (!v_ready) { ptx::mbar_wait(bar(V_FULL + vb), (j >> 1) & 1); v_ready = true; }
            TC_PROF_MARK(3);
            ptx::fence_proxy_async();
            ptx::tc_fence_after();
            issue_pv(t, vb, j > 0);
            if (j == nt[t] - 1) ptx::umma_commit(bar(O_FULL + t));
          }
          if (j + 1 < nt[t]) {
            TC_PROF_MARK(7);
            if (!k_ready) { ptx::mbar_wait(bar(K_FULL + kb), ((j + 1) >> 1) & 1); k_ready = true; }
            TC_PROF_MARK(4);
            ptx::fence_proxy_async();
            ptx::tc_fence_after();
            issue_qk(t, kb);
          }
        }
        ptx::umma_commit(bar(V_EMPTY + vb));
        if (j + 1 < n) ptx::umma_commit(bar(K_EMPTY + kb));
      }
      TC_PROF_MARK(7);
      if (prof_on)
        for (int i = 0; i < 8; ++i) p.prof[warp * 8 + i] = prof_acc[i];
    }
  } else {
    ptx::setmaxnreg_dec<kRegsOther>();
    // ================================================================ producers: gather K/V rows through the page table
    // A warp instruction copies two pool rows of this kv head (2 x 256 B, 16 lanes x 16 B each); row pair i of a
    // tile goes to producer warp i % 3.  Each lane reads the slots of its own rows straight from the page table
    // (a broadcast load per half warp), once per tile, and uses them for both the K and the V gather.
    const int pw = warp - 9;
    const int32_t* trow = p.req_to_token + load_index(p.req_pool_indices, p.idx_is_i64, req) * p.table_stride;
    const int64_t row_bytes = p.pool_tok_stride * 2;
    const uint8_t* kbase = reinterpret_cast<const uint8_t*>(p.k_pool) + static_cast<int64_t>(kv_head) * (kTile * 2) + (lane & 15) * 16;
    const uint8_t* vbase = reinterpret_cast<const uint8_t*>(p.v_pool) + static_cast<int64_t>(kv_head) * (kTile * 2) + (lane & 15) * 16;
    const int sub = lane >> 4;                       // which row of the pair
    const int ch = lane & 15;                        // 16-byte chunk of the 256-byte row
    constexpr int kPairsPerWarp = (64 + kProducerWarps - 1) / kProducerWarps;  // 22
    // slots of the rows this lane copies: row r_k = 2 * (pw + 3 k) + sub  (16 lanes share an address: broadcast).
    // The slots of tile j + 1 are fetched before the gathers of tile j are issued, so the index load never queues
    // behind this warp's own copies.
    int slot[kPairsPerWarp], slot_next[kPairsPerWarp];
    auto load_slots = [&](int j, int (&dst)[kPairsPerWarp]) {
#pragma unroll
      for (int k = 0; k < kPairsPerWarp; ++k) {
        const int r = 2 * (pw + kProducerWarps * k) + sub;
        const int tok = j * kTile + r;
        dst[k] = (r < kTile && tok < kv_len) ? __ldg(trow + tok) : 0;  // past the end: the dummy slot (masked later)
      }
    };
    load_slots(0, slot_next);
    TC_PROF_BEGIN();
    for (int j = 0; j < n; ++j) {
      const int b = j & 1, u = j >> 1;
#pragma unroll
      for (int k = 0; k < kPairsPerWarp; ++k) slot[k] = slot_next[k];
      if (j + 1 < n) load_slots(j + 1, slot_next);
      TC_PROF_MARK(0);
      auto gather = [&](const uint8_t* base, uint32_t tile) {
#pragma unroll
        for (int k = 0; k < kPairsPerWarp; ++k) {
          const int r = 2 * (pw + kProducerWarps * k) + sub;
          if (2 * (pw + kProducerWarps * k) < kTile) {  // warp-uniform
            const uint32_t dst = tile + (ch >> 3) * kHalfBytes + r * 128 + (((ch & 7) ^ (r & 7)) << 4);
            cp_async16(dst, base + static_cast<int64_t>(slot[k]) * row_bytes);
          }
        }
      };
      // K_j
      ptx::mbar_wait(bar(K_EMPTY + b), (u & 1) ^ 1);
      TC_PROF_MARK(1);
      if (!(p.debug & 1)) gather(kbase, smem + kOffK + b * kTileBytes);
      TC_PROF_MARK(2);
      ptx::cp_async_mbar_arrive_noinc(bar(K_FULL + b));
      TC_PROF_MARK(5);
      // V_j
      ptx::mbar_wait(bar(V_EMPTY + b), (u & 1) ^ 1);
      TC_PROF_MARK(3);
      if (!(p.debug & 1)) gather(vbase, smem + kOffV + b * kTileBytes);
      TC_PROF_MARK(4);
      ptx::cp_async_mbar_arrive_noinc(bar(V_FULL + b));
      TC_PROF_MARK(6);
    }
    if (prof_on && lane == 0)
      for (int i = 0; i < 8; ++i) p.prof[warp * 8 + i] = prof_acc[i];
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
  return D == 128 && kvb == 2 && p.logit_cap == 0.f && p.Hq >= 1 && (p.pool_tok_stride * 2) % 16 == 0;
}

int extend_tc_launch(const ExtendParams& p, bool q_bf16, cudaStream_t st) {
  auto kern = q_bf16 ? tc::extend_tc_kernel<true> : tc::extend_tc_kernel<false>;
  static bool configured[2] = {false, false};
  if (!configured[q_bf16]) {
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::kSmemBytes));
    configured[q_bf16] = true;
  }
  const int pair_heads = (p.G % 2 == 0) ? 1 : 0;
  const int q_tiles = (p.max_extend_len + tc::kTile - 1) / tc::kTile;
  dim3 grid(pair_heads ? p.Hq / 2 : p.Hq, p.bs, pair_heads ? q_tiles : (q_tiles + 1) / 2);
  B2_REQUIRE(grid.y <= 65535 && grid.z <= 65535, B200ATTN_ERR_UNSUPPORTED, "extend: batch %d or %d q tiles exceed the grid",
             p.bs, q_tiles);
  kern<<<grid, tc::kThreads, tc::kSmemBytes, st>>>(p, pair_heads);
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}

}  // namespace b200attn
