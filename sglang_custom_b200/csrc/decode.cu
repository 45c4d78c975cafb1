// Paged split-KV decode attention for sm_100a.
//
// Reference semantics: python/sglang/srt/layers/attention/torch_native_backend.py:112-180
// (per request: gather K/V rows through req_to_token, softmax(q K^T * scale) V, no causal mask);
// GPU designs it supersedes: triton_ops/decode_attention.py:45-552 (stage1 + stage2) and the
// sgl-kernel CPU flash-decoding sgl-kernel/csrc/cpu/decode.cpp:863-1373.
//
// Design (HBM-bound: the job is to keep the row gathers streaming and the per-tile math off their critical path)
//   * persistent grid: n_ctas CTAs (one per SM); the device-side plan (decode_plan.cu) hands every CTA an equal
//     share of the flattened 16-token units of the batch, so ragged batches stay balanced without a host sync.
//   * a CTA owns HB kv heads (HB = 8 for Llama-3-8B: the K row of one token for all 8 heads is one contiguous
//     2 KiB segment of the pool) and walks its share in tiles of NT tokens; rows are gathered straight from the
//     pool through req_to_token with 16-byte cp.async (LDGSTS) into a 3..6-stage XOR-swizzled shared-memory
//     ring -- no flattened kv_indices copy is ever built.
//   * warp-specialised: warps 8-11 are producers (gather only), warps 0-7 consumers; stages are handed over with
//     full/empty mbarriers, registers re-balanced with setmaxnreg.  (The first cut was monolithic: every warp
//     issued its share of the gather, then computed, one __syncthreads per tile -- 6.09 TB/s at C2 against
//     6.92 TB/s for this design; ncu showed 31 % of warp time at that barrier and a ~2000-cycle serial chain per
//     64 KiB tile against a 2560-cycle HBM budget.)
//   * consumer warp (hb, ts) computes 16 tokens of head hb per tile: the GQA group is packed as the M dimension
//     of mma.sync.m16n8k16 (S = Q K^T, O += P V), fp32 online softmax with quad shuffles; FP8-E4M3 K/V are
//     dequantised in registers (cvt.rn.f16x2.e4m3x2) on their way into the MMA.
//   * a request handled entirely by one CTA is normalised and written directly; otherwise the CTA writes a
//     normalised fp32 partial + LSE and the merge kernel (below) combines the pieces.
#include <stdlib.h>

#include "common.cuh"
#include "decode_plan.cuh"
#include "tc_ptx.cuh"
#include "tma.cuh"

namespace b200attn {

struct DecodeParams {
  const void* q;
  void* out;
  const void* k_pool;
  const void* v_pool;
  const int32_t* req_to_token;
  const void* req_pool_indices;
  const void* seq_lens;
  const int32_t* plan;
  float* ws_o;    // [2*n_ctas*HG][Hq... see slot layout
  float* ws_lse;
  float* lse_out;
  int64_t q_tok_stride, out_tok_stride, pool_tok_stride, table_stride;
  int idx_is_i64;
  int bs, Hq, Hkv, G;
  // multi-query rows (b200attn_decode_multiq): the G = nq * Gr MMA rows of a kv head are nq query tokens x Gr q heads
  // that share the request's keys.  Row r -> token r / Gr, q head kv_head * Gr + r % Gr.  Plain decode: nq = 1, Gr = G.
  int Gr, Hq_real;
  int merge_hpc;  // q heads served by one CTA of the merge kernel (set by launch_merge)
  int64_t q_sub_stride, out_sub_stride;  // stride between the nq tokens of a request (elements)
  int64_t lse_req_stride;                // lse_out: [request][token][Hq_real]
  int n_ctas;
  float scale_log2;  // sm_scale * k_scale * log2(e)   (logit_cap == 0)
  float scale_raw;   // sm_scale * k_scale
  float logit_cap;
  float v_scale;
  // paged-tile loads (page_size % 16 == 0): segments of a (page, K or V) box covered by one bulk-tensor copy, and the
  // dimension order of the tensor map (0: [128 B][rows][segments], 1: [128 B][segments][rows] with one segment per copy)
  int tma_box_segs, tma_order;
  int page_size;  // allocator page size (b200attn_decode_paged); 1 = token-granular slots
};

struct DecodeTmaps {
  CUtensorMap k, v;
};

constexpr int kTileWarps = 8;                                   // consumer warps that share one tile (stage)
constexpr int kProducerThreads = 128;                           // 4 producer warps
constexpr int kMaxSeg = 32;

// CW consumer warps = TG = CW / 8 tile groups of 8 warps; group g consumes the tiles t with t % TG == g, so that with
// CW = 16 two stages of the ring are being consumed at any time (four consumer warps per SM sub-partition).
// kBig (token-major FP8 consumer only): all CW = 16 consumer warps share one tile -- a stage of 64 KiB instead of two
// tile groups over 32 KiB stages.  An FP8 row is half as long as a bf16 row, so a 16-token tile is half the bytes and the
// per-tile fixed cost of the gather side (slot ring, producer barrier, mbarrier hand-overs) weighs twice as much: ncu
// showed the 16-warp kernel's consumers waiting for data with DRAM at 74 % whatever the page table looked like.
template <int D, int HB, int KVB /*bytes per kv elem*/, int NSTAGE, int CW = 8, bool kTokM = false, bool kBig = false>
struct DecodeCfg {
  static constexpr int TW = (kTokM && kBig) ? CW : kTileWarps;  // consumer warps that share one tile
  static constexpr int TG = CW / TW;                 // tile groups
  static constexpr int THREADS = CW * 32 + kProducerThreads;
  static constexpr int TS = TW / HB;                 // token sub-tiles (warps) per head and tile
  static constexpr int NT = 16 * TS;                 // tokens per tile
  static constexpr int HEADB = D * KVB;              // bytes of one head row
  static constexpr int ROWB = HB * HEADB;            // bytes gathered per token (K or V)
  static constexpr int CPH = HEADB / 16;             // 16-byte chunks per head row
  static constexpr int CPR = ROWB / 16;              // chunks per token row
  static constexpr int TILEB = NT * ROWB;            // bytes of a K (or V) tile
  static constexpr int STAGEB = 2 * TILEB;
  static constexpr int CHUNKS = NT * CPR;            // 16-byte chunks per K tile
  static constexpr int P_ITERS = CHUNKS / kProducerThreads;
  static_assert(CHUNKS % kProducerThreads == 0, "tile must divide over the producer warps");
  static constexpr int SWZ = CPH >= 8 ? 7 : (CPH - 1);
  static constexpr int RED_ROWS = 16;
  // cross-warp merge buffer of the TS * TG warps of a head (dedicated: the producers refill stages asynchronously).
  // Token-major consumers (kTokM) park their whole register state lane by lane: (D/16*4 + 4) floats per lane.
  static constexpr int RED_SLOT_F = kTokM ? (D / 16 * 4 + 4) * 32 : RED_ROWS * (D + 2);
  // slots of the merge buffer; a merge round with more writers than slots runs in passes (kBig: the 64 KiB stages leave
  // room for four slots)
  static constexpr int RED_SLOTS = (kTokM && kBig) ? 4 : CW / 2;
  static constexpr int REDB = (TS * TG > 1) ? RED_SLOTS * RED_SLOT_F * 4 : 0;
  static constexpr int IDXB = 2 * NT * 4;            // double-buffered token-slot ring
  static constexpr int BARB = 2 * NSTAGE * 8;        // full[NSTAGE], empty[NSTAGE]
  static constexpr int OFF_RED = NSTAGE * STAGEB;
  static constexpr int OFF_IDX = OFF_RED + REDB;
  static constexpr int OFF_BAR = OFF_IDX + IDXB;
  static constexpr int SMEM = OFF_BAR + BARB + 128 + 1024;  // + slack: the stage ring starts on a 1024-byte boundary
  static constexpr int SEGS = HEADB / 128;                  // 128-byte swizzle segments per head row (paged-tile loads)
};

struct Segment {
  int batch;      // request position in the batch
  int tok_begin;  // first kv position of this piece
  int tok_end;    // one past the last kv position
  int slot;       // -1: write final output; else workspace slot
  int64_t table_off;  // req_pool_index * table_stride
};

// ---------------------------------------------------------------------------------------
// The decode kernel.
//   * warps 0-7  = consumers: wait for a stage on its `full` mbarrier, compute, release it on `empty` -- no CTA-wide
//     barrier and no gather issue on their critical path;
//   * warps 8-11 = producers: gather the K/V rows of a tile through req_to_token with 16-byte cp.async as soon
//     as the stage is released; completion is counted on the `full` mbarrier (cp.async.mbarrier.arrive.noinc), so
//     producers never wait for their own copies.  Token slots are prefetched one tile ahead into a register and
//     published to the other producer warps through a double-buffered shared-memory ring.
//   * registers are re-balanced with setmaxnreg (consumers 208, producers 88).
__device__ __forceinline__ void named_barrier(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(nthreads) : "memory");
}

template <int D, int HB, int KVB, bool kQBf16, bool kG8, int NSTAGE, int CW, bool kTokM, bool kTma, bool kBig = false>
__global__ void __launch_bounds__(CW * 32 + kProducerThreads, 1)
decode_kernel(const DecodeParams p, const __grid_constant__ DecodeTmaps tm) {
  using Cfg = DecodeCfg<D, HB, KVB, NSTAGE, CW, kTokM, kBig>;
  constexpr int TW = Cfg::TW;
  constexpr bool kFp8 = (KVB == 1);
  constexpr int TG = Cfg::TG;
  static_assert(NSTAGE % TG == 0, "a stage must always serve the same tile group (see decode_stages)");
  static_assert(!kTokM || (kFp8 && kG8), "the token-major consumer is built for FP8 K/V with a GQA group <= 8");
  constexpr bool kMmaBf16 = kQBf16 && !kFp8;  // fp8 K/V are dequantised to f16 (exact)
  constexpr int TS = Cfg::TS, NT = Cfg::NT;
  constexpr int KSTEPS = D / 16;
  constexpr int NBLK = D / 8;

  extern __shared__ __align__(128) uint8_t smem_raw[];
  // 1024-byte aligned: the 128-byte swizzle of the paged-tile (TMA) loads is keyed on shared-memory address bits 7-9
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* stage_base = smem;
  float* red_buf = reinterpret_cast<float*>(smem + Cfg::OFF_RED);
  int32_t* idx_ring = reinterpret_cast<int32_t*>(smem + Cfg::OFF_IDX);
  const uint32_t bars = smem_u32(smem + Cfg::OFF_BAR);
  auto full_bar = [&](int s) { return bars + 8u * static_cast<uint32_t>(s); };
  auto empty_bar = [&](int s) { return bars + 8u * static_cast<uint32_t>(NSTAGE + s); };
  __shared__ Segment segs[kMaxSeg];
  __shared__ int n_segs_sh;
  __shared__ int next_req_sh, next_unit_sh;

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int cta = blockIdx.x;
  const int G = p.G;

  const int32_t* plan = p.plan;
  const int total_units = plan[kPlanTotalUnits];
  const int W = plan[kPlanUnitsPerCta];
  const int bs = p.bs;
  const int32_t* cum = plan + kPlanHeader;                 // [bs + 1]
  const int32_t* first_req = plan + kPlanHeader + bs + 1;  // [n_ctas]
  const int u_begin = cta * W;
  if (u_begin >= total_units) return;  // CTA-uniform, before any barrier
  if constexpr (kTma) {
    // a sliding-window plan shifts the first token of a request off its page boundary: the backend keeps window layers
    // on the row-gather instantiation, anything else is a caller bug that must not pass silently
    if (plan[kPlanWindow] >= 0) __trap();
  }
  const int u_end = min(u_begin + W, total_units);

  if (tid == 0) {
    next_req_sh = first_req[cta];
    next_unit_sh = u_begin;
    for (int s = 0; s < NSTAGE; ++s) {
      ptx::mbar_init(full_bar(s), kTma ? 1 : kProducerThreads);
      ptx::mbar_init(empty_bar(s), TW);
    }
    ptx::fence_mbar_init();
  }
  __syncthreads();
  const bool is_producer = warp >= CW;
  int stage = 0;        // ring position of the next tile (both roles walk the same tile sequence)
  uint32_t phase = 0;
  auto advance = [&]() { if (++stage == NSTAGE) { stage = 0; phase ^= 1u; } };
  auto seg_tiles = [&](int s) { return (segs[s].tok_end - segs[s].tok_begin + NT - 1) / NT; };

  // The two roles run textually separate loops (same sequence of CTA barriers) so that the register budget set by
  // setmaxnreg holds for everything a role executes.
  if (is_producer) {
    // (16 consumer warps: the kernel as a whole fits the 96 registers 640 threads leave -- no re-balancing.  A
    //  setmaxnreg.inc that asks for more than the producers released blocks forever.)
    if constexpr (CW == 8) ptx::setmaxnreg_dec<88>();
    while (true) {
      __syncthreads();  // thread 0 (a consumer) has built the next batch of segments
      const int n_segs = n_segs_sh;
      if (n_segs == 0) break;
      // ============================================================ producers
      if constexpr (kTma) {
        // paged-tile loads: the first producer warp alone; lane i owns page i of a tile (NT / 16 pages): one table read
        // (the slot of the page's first row, fetched one tile ahead) and one bulk-tensor copy per box of K and of V,
        // completion counted in bytes on the stage's `full` mbarrier.
        if (warp == CW) {
          constexpr int PG = NT / 16;
          constexpr uint32_t PAGEB = 16 * Cfg::ROWB;  // bytes of a page of K (or V) for this CTA's head block
          const int seg0 = blockIdx.y * HB * Cfg::SEGS;
          int i_seg = 0, i_tile = 0;
          int32_t pend = 0;
          int pend_pages = 0;
          auto fetch = [&]() {
            pend_pages = 0;
            if (i_seg >= n_segs) return;
            const Segment& sg = segs[i_seg];
            const int tok0 = sg.tok_begin + i_tile * NT;
            pend_pages = min(PG, (sg.tok_end - tok0 + 15) >> 4);
            pend = (lane < pend_pages) ? __ldg(p.req_to_token + sg.table_off + tok0 + lane * 16) : 0;
            if (++i_tile >= seg_tiles(i_seg)) { i_tile = 0; ++i_seg; }
          };
          fetch();
          for (int sg = 0; sg < n_segs; ++sg) {
            const int ntiles = seg_tiles(sg);
            for (int t = 0; t < ntiles; ++t) {
              const int32_t slot = pend;
              const int npg = pend_pages;
              fetch();
              ptx::mbar_wait(empty_bar(stage), phase ^ 1u);
              const uint32_t k_s = smem_u32(stage_base + stage * Cfg::STAGEB), v_s = k_s + Cfg::TILEB;
              if (lane == 0) ptx::mbar_arrive_expect_tx(full_bar(stage), static_cast<uint32_t>(npg) * 2u * PAGEB);
              __syncwarp();
              if (lane < npg) {
                const uint32_t dk = k_s + lane * PAGEB, dv = v_s + lane * PAGEB;
                for (int sgm = 0; sgm < HB * Cfg::SEGS; sgm += p.tma_box_segs) {
                  const int c1 = p.tma_order ? seg0 + sgm : slot, c2 = p.tma_order ? slot : seg0 + sgm;
                  ptx::tma_load_3d(dk + sgm * 2048, &tm.k, 0, c1, c2, full_bar(stage));
                  ptx::tma_load_3d(dv + sgm * 2048, &tm.v, 0, c1, c2, full_bar(stage));
                }
              }
              advance();
            }
          }
        } else {
          // idle producer warps still walk the ring so that `stage` / `phase` stay in step for the next segment batch
          for (int sg = 0; sg < n_segs; ++sg) {
            const int ntiles = seg_tiles(sg);
            for (int t = 0; t < ntiles; ++t) advance();
          }
        }
      } else {
      const int ptid = tid - CW * 32;
      const uint8_t* kpool = reinterpret_cast<const uint8_t*>(p.k_pool) + static_cast<int64_t>(blockIdx.y) * Cfg::ROWB;
      const uint8_t* vpool = reinterpret_cast<const uint8_t*>(p.v_pool) + static_cast<int64_t>(blockIdx.y) * Cfg::ROWB;
      const int64_t pool_row_bytes = p.pool_tok_stride * KVB;
      static_assert(Cfg::CPR <= kProducerThreads && kProducerThreads % Cfg::CPR == 0, "a token row must divide the producer threads");
      constexpr int kTpi = kProducerThreads / Cfg::CPR;  // tokens covered per iteration
      const int tok_lo = ptid / Cfg::CPR;
      const int cr = ptid % Cfg::CPR;
      const int ch = cr % Cfg::CPH;
      const uint32_t dst_col = (cr / Cfg::CPH) * Cfg::HEADB;
      const uint8_t* kcol = kpool + cr * 16;
      const uint8_t* vcol = vpool + cr * 16;
      int i_seg = 0, i_tile = 0;  // next tile whose token slots get fetched
      constexpr int SPT = (NT + kProducerThreads - 1) / kProducerThreads;  // token slots fetched per producer thread and tile
      int32_t pend[SPT];
#pragma unroll
      for (int i = 0; i < SPT; ++i) pend[i] = -1;
      auto fetch = [&]() {
        if (i_seg >= n_segs) return;
        const Segment& s = segs[i_seg];
#pragma unroll
        for (int i = 0; i < SPT; ++i) {
          const int tt = ptid + i * kProducerThreads;
          if (tt < NT) {
            const int tok = s.tok_begin + i_tile * NT + tt;
            pend[i] = (tok < s.tok_end) ? __ldg(p.req_to_token + s.table_off + tok) : -1;
          }
        }
        if (++i_tile >= seg_tiles(i_seg)) { i_tile = 0; ++i_seg; }
      };
      auto publish = [&](int buf) {
#pragma unroll
        for (int i = 0; i < SPT; ++i) {
          const int tt = ptid + i * kProducerThreads;
          if (tt < NT) idx_ring[buf * NT + tt] = pend[i];
        }
      };
      fetch();
      publish(0);
      named_barrier(2, kProducerThreads);
      fetch();
      int cur = 0;
      for (int sg = 0; sg < n_segs; ++sg) {
        const int ntiles = seg_tiles(sg);
        for (int t = 0; t < ntiles; ++t) {
          ptx::mbar_wait(empty_bar(stage), phase ^ 1u);
          const int32_t* idx = idx_ring + cur * NT;
          const uint32_t k_s = smem_u32(stage_base + stage * Cfg::STAGEB), v_s = k_s + Cfg::TILEB;
          // chunk c = it * 128 + ptid of the tile: the 16-byte column (cr) of a thread never changes, its token
          // advances by TPI per iteration -- a rolled loop with a handful of live registers
#pragma unroll 2
          for (int it = 0; it < Cfg::P_ITERS; ++it) {
            const int tok = it * kTpi + tok_lo;
            const int32_t slot = idx[tok];
            const int valid = slot >= 0 ? 16 : 0;
            const int64_t src = static_cast<int64_t>(slot >= 0 ? slot : 0) * pool_row_bytes;
            const uint32_t dst = tok * Cfg::ROWB + dst_col + ((ch ^ (tok & Cfg::SWZ)) << 4);
            cp_async16(k_s + dst, kcol + src, valid);
            cp_async16(v_s + dst, vcol + src, valid);
          }
          ptx::cp_async_mbar_arrive_noinc(full_bar(stage));
          publish(cur ^ 1);  // slots of the next tile (fetched a tile ago)
          fetch();
          named_barrier(2, kProducerThreads);
          cur ^= 1;
          advance();
        }
      }
      }
      __syncthreads();  // the segment list may be rebuilt
    }
    asm volatile("cp.async.wait_all;\n" ::: "memory");
  } else {
    if constexpr (CW == 8) ptx::setmaxnreg_inc<208>();
    int tile_no = 0;  // running tile number of this CTA (the tile groups alternate)
    while (true) {
      // ---------------- build the next batch of segments (thread 0), at most kMaxSeg
      if (tid == 0) {
        int i = next_req_sh, u = next_unit_sh, n = 0;
        while (u < u_end && n < kMaxSeg) {
          while (cum[i + 1] <= u) ++i;
          const int c0 = cum[i], c1 = cum[i + 1];
          const int seg_end = min(c1, u_end);
          const int Lfull = static_cast<int>(load_index(p.seq_lens, p.idx_is_i64, i));
          const int win = p.plan[kPlanWindow];  // sliding-window plan: only the last min(L, win + 1) tokens
          const int L = (win >= 0 && Lfull > win + 1) ? win + 1 : Lfull;
          Segment s;
          s.batch = i;
          s.tok_begin = (u - c0) * kUnit;
          s.tok_end = min((seg_end - c0) * kUnit, L);
          const bool is_first = (u == c0), is_last = (seg_end == c1);
          s.slot = (is_first && is_last) ? -1 : (is_first ? 2 * cta + 1 : 2 * cta);
          s.table_off = load_index(p.req_pool_indices, p.idx_is_i64, i) * p.table_stride + (Lfull - L);
          segs[n++] = s;
          u = seg_end;
        }
        n_segs_sh = n;
        next_req_sh = i;
        next_unit_sh = u;
      }
      __syncthreads();
      const int n_segs = n_segs_sh;
      if (n_segs == 0) break;
      // ============================================================ consumers
      const int tg = warp / TW;            // tile group: consumes the tiles with tile_no % TG == tg
      const int hb = (warp % TW) % HB;     // head within the CTA's head block
      const int ts = (warp % TW) / HB;     // token sub-tile inside a tile
      const int qd = lane & 3;    // quad lane
      const int row = lane >> 2;  // MMA row 0..7
      const int kv_head = blockIdx.y * HB + hb;
      auto consumer_barrier = [&]() { named_barrier(1, CW * 32); };
      // byte offset of 16-byte chunk ch of this warp's head, token tok of the tile, inside a K (or V) tile
      auto kv_off = [&](int tok, int ch) -> uint32_t {
        if constexpr (kTma)  // [page of the tile][segment][16 rows][128 B], swizzle keyed on the row (TMA SWIZZLE_128B)
          return (tok >> 4) * (16 * Cfg::ROWB) + (hb * Cfg::SEGS + (ch >> 3)) * 2048 + (tok & 15) * 128 +
                 (((ch & 7) ^ (tok & 7)) << 4);
        else
          return tok * Cfg::ROWB + hb * Cfg::HEADB + ((ch ^ (tok & Cfg::SWZ)) << 4);
      };
      // A paged-tile load brings whole pages: the rows of the last page past the end of the piece hold other tokens'
      // (or never written) data.  K garbage is masked by select; V garbage is zeroed here so that 0 * NaN cannot
      // reach O.  Rare (once per piece), the rows belong to this warp alone.
      auto zero_v_tail = [&](uint32_t v_s, int nvalid) {
        if constexpr (kTma) {
          for (int i = lane; i < (16 - nvalid) * Cfg::CPH; i += 32) {
            const int tok = ts * 16 + nvalid + i / Cfg::CPH;
            asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};\n" ::"r"(v_s + kv_off(tok, i % Cfg::CPH)), "r"(0u) : "memory");
          }
          ptx::fence_proxy_async();  // the stage is refilled through the async proxy
          __syncwarp();
        }
      };
      for (int sg = 0; sg < n_segs; ++sg) {
        const Segment seg = segs[sg];
        const int ntiles = (seg.tok_end - seg.tok_begin + NT - 1) / NT;

        if constexpr (!kTokM) {
        // ---- Q fragments of this request (rows = q heads of the group, zero padded to 16)
        uint32_t qa[KSTEPS][4];
        {
          const uint16_t* qb = reinterpret_cast<const uint16_t*>(p.q) + seg.batch * p.q_tok_stride;
          const int Gr = p.Gr;
          const uint16_t* qrow0 = qb + (row / Gr) * p.q_sub_stride + static_cast<int64_t>(kv_head * Gr + row % Gr) * D;
          const uint16_t* qrow1 = qb + ((row + 8) / Gr) * p.q_sub_stride +
                                  static_cast<int64_t>(kv_head * Gr + (row + 8) % Gr) * D;
          const bool v0 = row < G, v1 = !kG8 && (row + 8 < G);
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
              // fp8 K: a thread's ldmatrix register holds d = 16*ks + 4*qd + {0,1,2,3}; the
              // contraction index is permuted identically on the Q side.
              if (v0) {
                const uint2 t = *reinterpret_cast<const uint2*>(qrow0 + ks * 16 + 4 * qd);
                a0 = t.x; a2 = t.y;
              }
              if (v1) {
                const uint2 t = *reinterpret_cast<const uint2*>(qrow1 + ks * 16 + 4 * qd);
                a1 = t.x; a3 = t.y;
              }
              if constexpr (kQBf16) {  // bf16 -> f16 (saturating), the MMA runs in f16 for fp8 KV
                auto cv = [](uint32_t x) {
                  const float lo = __uint_as_float(x << 16), hi = __uint_as_float(x & 0xffff0000u);
                  const float l2 = fminf(fmaxf(lo, -65504.f), 65504.f), h2 = fminf(fmaxf(hi, -65504.f), 65504.f);
                  return pack2<false>(l2, h2);
                };
                a0 = cv(a0); a1 = cv(a1); a2 = cv(a2); a3 = cv(a3);
              }
            }
            qa[ks][0] = a0; qa[ks][1] = a1; qa[ks][2] = a2; qa[ks][3] = a3;
          }
        }

        float o[NBLK][4];
  #pragma unroll
        for (int nb = 0; nb < NBLK; ++nb) { o[nb][0] = o[nb][1] = o[nb][2] = o[nb][3] = 0.f; }
        float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

        for (int t = 0; t < ntiles; ++t, ++tile_no) {
          if (TG > 1 && tile_no % TG != tg) { advance(); continue; }  // the other tile group's stage
          ptx::mbar_wait(full_bar(stage), phase);
          const int tok0 = seg.tok_begin + t * NT + ts * 16;  // first kv position of this warp's 16 tokens
          const int nvalid = seg.tok_end - tok0;               // <= 0: nothing to do
          const uint32_t k_s = smem_u32(stage_base + stage * Cfg::STAGEB);
          const uint32_t v_s = k_s + Cfg::TILEB;
          if (nvalid > 0) {
          if (nvalid < 16) zero_v_tail(v_s, nvalid);
          // ---- S = Q K^T for 16 tokens (2 n-blocks of 8)
          float s[2][4];
  #pragma unroll
          for (int nb = 0; nb < 2; ++nb) { s[nb][0] = s[nb][1] = s[nb][2] = s[nb][3] = 0.f; }
          if constexpr (!kFp8) {
  #pragma unroll
            for (int ks = 0; ks < KSTEPS; ++ks) {
              const int mj = lane >> 3, mr = lane & 7;
              const int tok = ts * 16 + (mj >> 1) * 8 + mr;
              const int ch = 2 * ks + (mj & 1);
              uint32_t r[4];
              ldmatrix_x4(r, k_s + kv_off(tok, ch));
              mma_16816<kMmaBf16>(s[0], qa[ks], r[0], r[1]);
              mma_16816<kMmaBf16>(s[1], qa[ks], r[2], r[3]);
            }
          } else {
            // fp8: one 16-byte chunk = 16 d; matrices (tok 0-7, ch), (tok 8-15, ch), (tok 0-7, ch+1), (tok 8-15, ch+1)
  #pragma unroll
            for (int kp = 0; kp < KSTEPS / 2; ++kp) {
              const int mj = lane >> 3, mr = lane & 7;
              const int tok = ts * 16 + (mj & 1) * 8 + mr;
              const int ch = 2 * kp + (mj >> 1);
              uint32_t r[4];
              ldmatrix_x4(r, k_s + kv_off(tok, ch));
  #pragma unroll
              for (int half = 0; half < 2; ++half) {
                const uint32_t r0 = r[2 * half], r1 = r[2 * half + 1];
                const int ks = 2 * kp + half;
                mma_16816<false>(s[0], qa[ks], e4m3x2_to_f16x2(r0 & 0xffff), e4m3x2_to_f16x2(r0 >> 16));
                mma_16816<false>(s[1], qa[ks], e4m3x2_to_f16x2(r1 & 0xffff), e4m3x2_to_f16x2(r1 >> 16));
              }
            }
          }

          // ---- scale, cap, mask, online softmax (log2 domain)
  #pragma unroll
          for (int nb = 0; nb < 2; ++nb) {
  #pragma unroll
            for (int e = 0; e < 4; ++e) {
              float x;
              if (p.logit_cap > 0.f) {
                x = p.logit_cap * fast_tanh(s[nb][e] * p.scale_raw / p.logit_cap) * kLog2e;
              } else {
                x = s[nb][e] * p.scale_log2;
              }
              const int col = nb * 8 + 2 * qd + (e & 1);
              s[nb][e] = (col < nvalid) ? x : -INFINITY;
            }
          }
          float mx0 = fmaxf(fmaxf(s[0][0], s[0][1]), fmaxf(s[1][0], s[1][1]));
          mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
          mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
          const float mn0 = fmaxf(m0, mx0);  // finite: at least one valid column
          const float al0 = fast_exp2(m0 - mn0);
          m0 = mn0;
          float p00 = fast_exp2(s[0][0] - mn0), p01 = fast_exp2(s[0][1] - mn0);
          float p10 = fast_exp2(s[1][0] - mn0), p11 = fast_exp2(s[1][1] - mn0);
          l0 = l0 * al0 + (p00 + p01 + p10 + p11);
          uint32_t pa[4];
          pa[0] = pack2<kMmaBf16>(p00, p01);
          pa[2] = pack2<kMmaBf16>(p10, p11);
          float al1 = 1.f;
          if constexpr (!kG8) {
            float mx1 = fmaxf(fmaxf(s[0][2], s[0][3]), fmaxf(s[1][2], s[1][3]));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
            const float mn1 = fmaxf(m1, mx1);
            al1 = fast_exp2(m1 - mn1);
            m1 = mn1;
            const float q00 = fast_exp2(s[0][2] - mn1), q01 = fast_exp2(s[0][3] - mn1);
            const float q10 = fast_exp2(s[1][2] - mn1), q11 = fast_exp2(s[1][3] - mn1);
            l1 = l1 * al1 + (q00 + q01 + q10 + q11);
            pa[1] = pack2<kMmaBf16>(q00, q01);
            pa[3] = pack2<kMmaBf16>(q10, q11);
          } else {
            pa[1] = 0u; pa[3] = 0u;
          }
  #pragma unroll
          for (int nb = 0; nb < NBLK; ++nb) {
            o[nb][0] *= al0; o[nb][1] *= al0;
            if constexpr (!kG8) { o[nb][2] *= al1; o[nb][3] *= al1; }
          }

          // ---- O += P V
          if constexpr (!kFp8) {
  #pragma unroll
            for (int jp = 0; jp < NBLK / 2; ++jp) {
              const int mj = lane >> 3, mr = lane & 7;
              const int tok = ts * 16 + (mj & 1) * 8 + mr;
              const int ch = 2 * jp + (mj >> 1);
              uint32_t r[4];
              ldmatrix_x4_trans(r, v_s + kv_off(tok, ch));
              mma_16816<kMmaBf16>(o[2 * jp], pa, r[0], r[1]);
              mma_16816<kMmaBf16>(o[2 * jp + 1], pa, r[2], r[3]);
            }
          } else {
            // fp8 V: a transposed b16 ldmatrix hands a thread (tok 2qd, d pair), (tok 2qd+1, d pair);
            // regroup bytes so each half-word is one d for two consecutive tokens.
            // chunk c (16 d) -> n-block "even d" o[4c'+..]; accumulator column order is fixed up at the store.
  #pragma unroll
            for (int cp = 0; cp < D / 32; ++cp) {
              const int mj = lane >> 3, mr = lane & 7;
              const int tok = ts * 16 + (mj & 1) * 8 + mr;
              const int ch = 2 * cp + (mj >> 1);
              uint32_t r[4];
              ldmatrix_x4_trans(r, v_s + kv_off(tok, ch));
  #pragma unroll
              for (int half = 0; half < 2; ++half) {
                const uint32_t lo = r[2 * half], hi = r[2 * half + 1];  // tokens 0-7, tokens 8-15
                const uint32_t e_lo = __byte_perm(lo, 0, 0x0020), o_lo = __byte_perm(lo, 0, 0x0031);
                const uint32_t e_hi = __byte_perm(hi, 0, 0x0020), o_hi = __byte_perm(hi, 0, 0x0031);
                const int c = 2 * cp + half;  // 16-d chunk
                mma_16816<false>(o[2 * c], pa, e4m3x2_to_f16x2(e_lo), e4m3x2_to_f16x2(e_hi));
                mma_16816<false>(o[2 * c + 1], pa, e4m3x2_to_f16x2(o_lo), e4m3x2_to_f16x2(o_hi));
              }
            }
          }
          }
          // every ldmatrix of this tile has returned (the MMAs above consumed them): hand the stage back
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(empty_bar(stage));
          advance();
        }  // tiles

        // ---- finish the piece: reduce l over the quad, merge the TS warps of a head, store
        l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
        l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
        if constexpr (!kG8) {
          l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
          l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
        }
        if constexpr (TS * TG > 1) {
          // tree merge: in each round the upper half of the token sub-tiles hands its state to the lower half
          constexpr int RSTRIDE = D + 2;
          float* red = red_buf;  // dedicated: the producers refill stages asynchronously
          const int tsa = tg * TS + ts;  // this warp's place among the TS * TG partial states of the head
  #pragma unroll
          for (int half = TS * TG / 2; half >= 1; half >>= 1) {
            consumer_barrier();
            if (tsa >= half && tsa < 2 * half) {
              float* dst = red + ((tsa - half) * HB + hb) * Cfg::RED_ROWS * RSTRIDE;
  #pragma unroll
              for (int nb = 0; nb < NBLK; ++nb) {
                *reinterpret_cast<float2*>(dst + row * RSTRIDE + nb * 8 + 2 * qd) = make_float2(o[nb][0], o[nb][1]);
                if constexpr (!kG8)
                  *reinterpret_cast<float2*>(dst + (row + 8) * RSTRIDE + nb * 8 + 2 * qd) = make_float2(o[nb][2], o[nb][3]);
              }
              if (qd == 0) {
                *reinterpret_cast<float2*>(dst + row * RSTRIDE + D) = make_float2(m0, l0);
                if constexpr (!kG8) *reinterpret_cast<float2*>(dst + (row + 8) * RSTRIDE + D) = make_float2(m1, l1);
              }
            }
            consumer_barrier();
            if (tsa < half) {
              const float* src = red + (tsa * HB + hb) * Cfg::RED_ROWS * RSTRIDE;
              {
                const float2 ml = *reinterpret_cast<const float2*>(src + row * RSTRIDE + D);
                const float mn = fmaxf(m0, ml.x);
                const float a = (mn == -INFINITY) ? 0.f : fast_exp2(m0 - mn);
                const float b = (mn == -INFINITY) ? 0.f : fast_exp2(ml.x - mn);
                m0 = mn; l0 = l0 * a + ml.y * b;
  #pragma unroll
                for (int nb = 0; nb < NBLK; ++nb) {
                  const float2 x = *reinterpret_cast<const float2*>(src + row * RSTRIDE + nb * 8 + 2 * qd);
                  o[nb][0] = o[nb][0] * a + x.x * b;
                  o[nb][1] = o[nb][1] * a + x.y * b;
                }
              }
              if constexpr (!kG8) {
                const float2 ml = *reinterpret_cast<const float2*>(src + (row + 8) * RSTRIDE + D);
                const float mn = fmaxf(m1, ml.x);
                const float a = (mn == -INFINITY) ? 0.f : fast_exp2(m1 - mn);
                const float b = (mn == -INFINITY) ? 0.f : fast_exp2(ml.x - mn);
                m1 = mn; l1 = l1 * a + ml.y * b;
  #pragma unroll
                for (int nb = 0; nb < NBLK; ++nb) {
                  const float2 x = *reinterpret_cast<const float2*>(src + (row + 8) * RSTRIDE + nb * 8 + 2 * qd);
                  o[nb][2] = o[nb][2] * a + x.x * b;
                  o[nb][3] = o[nb][3] * a + x.y * b;
                }
              }
            }
          }
        }

        if (ts == 0 && tg == 0) {
  #pragma unroll
          for (int rr = 0; rr < (kG8 ? 1 : 2); ++rr) {
            const int r = row + rr * 8;
            if (r >= G) continue;
            const float m = rr ? m1 : m0, l = rr ? l1 : l0;
            const float inv = (l > 0.f) ? p.v_scale / l : 0.f;
            const float lse = (l > 0.f) ? (m + log2f(l)) * kLn2 : -INFINITY;
            const int head = kv_head * G + r;  // workspace index (virtual head)
            const int sub = r / p.Gr, head_real = kv_head * p.Gr + r % p.Gr;
            if (seg.slot < 0) {
              uint16_t* dst = reinterpret_cast<uint16_t*>(p.out) + seg.batch * p.out_tok_stride +
                              sub * p.out_sub_stride + static_cast<int64_t>(head_real) * D;
              if constexpr (!kFp8) {
  #pragma unroll
                for (int nb = 0; nb < NBLK; ++nb)
                  *reinterpret_cast<uint32_t*>(dst + nb * 8 + 2 * qd) =
                      pack2<kQBf16>(o[nb][2 * rr] * inv, o[nb][2 * rr + 1] * inv);
              } else {
  #pragma unroll
                for (int c = 0; c < D / 16; ++c) {  // thread holds d = 16c + 4qd + {0,2} (even blk), {1,3} (odd blk)
                  const uint32_t w0 = pack2<kQBf16>(o[2 * c][2 * rr] * inv, o[2 * c + 1][2 * rr] * inv);
                  const uint32_t w1 = pack2<kQBf16>(o[2 * c][2 * rr + 1] * inv, o[2 * c + 1][2 * rr + 1] * inv);
                  *reinterpret_cast<uint2*>(dst + c * 16 + 4 * qd) = make_uint2(w0, w1);
                }
              }
              if (p.lse_out != nullptr && qd == 0)
                p.lse_out[static_cast<int64_t>(seg.batch) * p.lse_req_stride + sub * p.Hq_real + head_real] = lse;
            } else {
              float* dst = p.ws_o + (static_cast<int64_t>(seg.slot) * p.Hq + head) * D;
              if constexpr (!kFp8) {
  #pragma unroll
                for (int nb = 0; nb < NBLK; ++nb)
                  *reinterpret_cast<float2*>(dst + nb * 8 + 2 * qd) =
                      make_float2(o[nb][2 * rr] * inv, o[nb][2 * rr + 1] * inv);
              } else {
  #pragma unroll
                for (int c = 0; c < D / 16; ++c)
                  *reinterpret_cast<float4*>(dst + c * 16 + 4 * qd) =
                      make_float4(o[2 * c][2 * rr] * inv, o[2 * c + 1][2 * rr] * inv,
                                  o[2 * c][2 * rr + 1] * inv, o[2 * c + 1][2 * rr + 1] * inv);
              }
              if (qd == 0) p.ws_lse[static_cast<int64_t>(seg.slot) * p.Hq + head] = lse;
            }
          }
        }
        } else {
          // ================================================================ token-major consumer (FP8 K/V, G <= 8)
          // The 16 tokens of the warp are the M dimension of mma.sync.m16n8k16 and the (up to 8) q heads of the GQA
          // group its N dimension: S^T = K Q^T and O^T = V^T P^T.  Against the head-major form above (heads padded
          // to 16 MMA rows) this halves the MMA count and the accumulator registers, which is what lets 16 consumer
          // warps fit -- the FP8 stream needs the extra warps because every byte pays an e4m3 -> f16 conversion
          // (F2FP) in front of the MMA and the chain LDSM -> F2FP -> HMMA -> softmax -> HMMA of one warp is long.
          // Thread (r = lane / 4, qd): S^T rows r, r + 8 (tokens), columns 2qd, 2qd + 1 (heads).
          constexpr int NB = D / 16;  // accumulator blocks: even cp -> "even d" rows, odd -> "odd d" rows (see PV)
          uint32_t qb[KSTEPS][2];
          {
            const uint16_t* qbase = reinterpret_cast<const uint16_t*>(p.q) + seg.batch * p.q_tok_stride;
            const int Gr = p.Gr;
            const uint16_t* qrow = qbase + (row / Gr) * p.q_sub_stride + static_cast<int64_t>(kv_head * Gr + row % Gr) * D;
            const bool vq = row < G;
  #pragma unroll
            for (int ks = 0; ks < KSTEPS; ++ks) {
              // a K register of ldmatrix holds d = 16*ks' + 4*qd + {0,1,2,3} (bytes); the contraction index is
              // permuted identically here: b0 = d 4qd + {0,1}, b1 = d 4qd + {2,3}
              uint2 t = make_uint2(0u, 0u);
              if (vq) t = *reinterpret_cast<const uint2*>(qrow + ks * 16 + 4 * qd);
              if constexpr (kQBf16) { t.x = ptx::bf16x2_to_f16x2(t.x); t.y = ptx::bf16x2_to_f16x2(t.y); }
              qb[ks][0] = t.x; qb[ks][1] = t.y;
            }
          }
          float o[NB][4];
  #pragma unroll
          for (int nb = 0; nb < NB; ++nb) { o[nb][0] = o[nb][1] = o[nb][2] = o[nb][3] = 0.f; }
          float mh0 = -INFINITY, mh1 = -INFINITY;  // running maxima of heads 2qd, 2qd + 1 (log2 domain, scaled)
          float lh0 = 0.f, lh1 = 0.f;              // this thread's share of the row sums (reduced at the end)
          const int mj = lane >> 3, mr = lane & 7;

          for (int t = 0; t < ntiles; ++t, ++tile_no) {
            if (TG > 1 && tile_no % TG != tg) { advance(); continue; }  // the other tile group's stage
            ptx::mbar_wait(full_bar(stage), phase);
            const int tok0 = seg.tok_begin + t * NT + ts * 16;
            const int nvalid = seg.tok_end - tok0;
            const uint32_t k_s = smem_u32(stage_base + stage * Cfg::STAGEB);
            const uint32_t v_s = k_s + Cfg::TILEB;
            if (nvalid > 0) {
              if (nvalid < 16) zero_v_tail(v_s, nvalid);
              // ---- S^T = K Q^T: A = K (16 tokens x 16 d per step, dequantised), B = Q^T from registers
              float sa[4] = {0.f, 0.f, 0.f, 0.f}, sb[4] = {0.f, 0.f, 0.f, 0.f};  // two chains (even / odd k-steps)
  #pragma unroll
              for (int kp = 0; kp < KSTEPS / 2; ++kp) {
                const int tok = ts * 16 + (mj & 1) * 8 + mr;
                const int ch = 2 * kp + (mj >> 1);
                uint32_t r[4];  // (tok 0-7, ch), (tok 8-15, ch), (tok 0-7, ch + 1), (tok 8-15, ch + 1)
                ldmatrix_x4(r, k_s + kv_off(tok, ch));
                uint32_t a[4];
                a[0] = e4m3x2_to_f16x2(r[0] & 0xffff); a[2] = e4m3x2_to_f16x2(r[0] >> 16);
                a[1] = e4m3x2_to_f16x2(r[1] & 0xffff); a[3] = e4m3x2_to_f16x2(r[1] >> 16);
                mma_16816<false>(sa, a, qb[2 * kp][0], qb[2 * kp][1]);
                a[0] = e4m3x2_to_f16x2(r[2] & 0xffff); a[2] = e4m3x2_to_f16x2(r[2] >> 16);
                a[1] = e4m3x2_to_f16x2(r[3] & 0xffff); a[3] = e4m3x2_to_f16x2(r[3] >> 16);
                mma_16816<false>(sb, a, qb[2 * kp + 1][0], qb[2 * kp + 1][1]);
              }
              float x0 = sa[0] + sb[0], x1 = sa[1] + sb[1], x2 = sa[2] + sb[2], x3 = sa[3] + sb[3];
              if (p.logit_cap > 0.f) {
                const float c = p.scale_raw / p.logit_cap, d = p.logit_cap * kLog2e;
                x0 = d * fast_tanh(x0 * c); x1 = d * fast_tanh(x1 * c);
                x2 = d * fast_tanh(x2 * c); x3 = d * fast_tanh(x3 * c);
              } else {
                x0 *= p.scale_log2; x1 *= p.scale_log2; x2 *= p.scale_log2; x3 *= p.scale_log2;
              }
              if (nvalid < 16) {  // ragged end of the piece (warp-uniform)
                if (row >= nvalid) { x0 = -INFINITY; x1 = -INFINITY; }
                if (row + 8 >= nvalid) { x2 = -INFINITY; x3 = -INFINITY; }
              }
              // ---- online softmax per head: the maximum runs over the token rows (lane bits 2..4)
              float mx0 = fmaxf(x0, x2), mx1 = fmaxf(x1, x3);
  #pragma unroll
              for (int sh = 4; sh <= 16; sh <<= 1) {
                mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, sh));
                mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, sh));
              }
              const float mn0 = fmaxf(mh0, mx0), mn1 = fmaxf(mh1, mx1);  // finite: token tok0 is valid
              const bool grew = (mn0 > mh0) || (mn1 > mh1);
              if (__any_sync(0xffffffffu, grew)) {  // lazy rescale: a running maximum rarely moves after the first tiles
                const float al0 = fast_exp2(mh0 - mn0), al1 = fast_exp2(mh1 - mn1);
                lh0 *= al0; lh1 *= al1;
  #pragma unroll
                for (int nb = 0; nb < NB; ++nb) { o[nb][0] *= al0; o[nb][1] *= al1; o[nb][2] *= al0; o[nb][3] *= al1; }
                mh0 = mn0; mh1 = mn1;
              }
              const float p0 = fast_exp2(x0 - mh0), p1 = fast_exp2(x1 - mh1);
              const float p2 = fast_exp2(x2 - mh0), p3 = fast_exp2(x3 - mh1);
              lh0 += p0 + p2; lh1 += p1 + p3;
              // P^T as the B operand: transpose the two 8x8 (token x head) blocks across the warp
              uint32_t b0 = pack2<false>(p0, p1), b1 = pack2<false>(p2, p3);
              asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;\n" : "=r"(b0) : "r"(b0));
              asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;\n" : "=r"(b1) : "r"(b1));
              // ---- O^T += V^T P^T: a transposed b16 ldmatrix of fp8 rows hands a thread (tok 2qd: d0 d1), (tok 2qd+1:
              // d0 d1) with d0 = 16*ch + 2*r; one byte permute groups them per d, and the "even d" / "odd d" halves
              // feed two MMAs whose rows are d = 32cp + 2r (+16 for rows 8-15) resp. + 1.
  #pragma unroll
              for (int cp = 0; cp < D / 32; ++cp) {
                const int tok = ts * 16 + (mj >> 1) * 8 + mr;
                const int ch = 2 * cp + (mj & 1);
                uint32_t r[4];  // (tok 0-7, ch), (tok 0-7, ch + 1), (tok 8-15, ch), (tok 8-15, ch + 1)
                ldmatrix_x4_trans(r, v_s + kv_off(tok, ch));
                uint32_t ae[4], ao[4];
  #pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const uint32_t g = __byte_perm(r[i], 0, 0x3120);  // (t0 d0, t1 d0, t0 d1, t1 d1)
                  ae[i] = e4m3x2_to_f16x2(g & 0xffff);
                  ao[i] = e4m3x2_to_f16x2(g >> 16);
                }
                mma_16816<false>(o[2 * cp], ae, b0, b1);
                mma_16816<false>(o[2 * cp + 1], ao, b0, b1);
              }
            }
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(empty_bar(stage));
            advance();
          }  // tiles

          // ---- finish the piece: merge the TS * TG warps of a head (whole register state, lane by lane), store
          if constexpr (TS * TG > 1) {
            const int tsa = tg * TS + ts;
            constexpr int SLOT = Cfg::RED_SLOT_F;
            constexpr int RS = Cfg::RED_SLOTS;
  #pragma unroll
            for (int half = TS * TG / 2; half >= 1; half >>= 1) {
              // the half * HB warps with tsa in [half, 2 half) hand their state to the warps below them, RS at a time
  #pragma unroll
              for (int pass = 0; pass * RS < half * HB; ++pass) {
                consumer_barrier();
                const int wslot = (tsa - half) * HB + hb;
                if (tsa >= half && tsa < 2 * half && wslot / RS == pass) {
                  float* dst = red_buf + (wslot % RS) * SLOT + lane;
  #pragma unroll
                  for (int nb = 0; nb < NB; ++nb) {
  #pragma unroll
                    for (int e = 0; e < 4; ++e) dst[(nb * 4 + e) * 32] = o[nb][e];
                  }
                  dst[(NB * 4 + 0) * 32] = mh0; dst[(NB * 4 + 1) * 32] = mh1;
                  dst[(NB * 4 + 2) * 32] = lh0; dst[(NB * 4 + 3) * 32] = lh1;
                }
                consumer_barrier();
                const int rslot = tsa * HB + hb;
                if (tsa < half && rslot / RS == pass) {
                  const float* src = red_buf + (rslot % RS) * SLOT + lane;
                  const float om0 = src[(NB * 4 + 0) * 32], om1 = src[(NB * 4 + 1) * 32];
                  const float mn0 = fmaxf(mh0, om0), mn1 = fmaxf(mh1, om1);
                  const float a0 = (mn0 == -INFINITY) ? 0.f : fast_exp2(mh0 - mn0);
                  const float c0 = (mn0 == -INFINITY) ? 0.f : fast_exp2(om0 - mn0);
                  const float a1 = (mn1 == -INFINITY) ? 0.f : fast_exp2(mh1 - mn1);
                  const float c1 = (mn1 == -INFINITY) ? 0.f : fast_exp2(om1 - mn1);
                  mh0 = mn0; mh1 = mn1;
                  lh0 = lh0 * a0 + src[(NB * 4 + 2) * 32] * c0;
                  lh1 = lh1 * a1 + src[(NB * 4 + 3) * 32] * c1;
  #pragma unroll
                  for (int nb = 0; nb < NB; ++nb) {
                    o[nb][0] = o[nb][0] * a0 + src[(nb * 4 + 0) * 32] * c0;
                    o[nb][1] = o[nb][1] * a1 + src[(nb * 4 + 1) * 32] * c1;
                    o[nb][2] = o[nb][2] * a0 + src[(nb * 4 + 2) * 32] * c0;
                    o[nb][3] = o[nb][3] * a1 + src[(nb * 4 + 3) * 32] * c1;
                  }
                }
              }
            }
          }
          if (ts == 0 && tg == 0) {
  #pragma unroll
            for (int sh = 4; sh <= 16; sh <<= 1) {
              lh0 += __shfl_xor_sync(0xffffffffu, lh0, sh);
              lh1 += __shfl_xor_sync(0xffffffffu, lh1, sh);
            }
  #pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              const int r = 2 * qd + hh;  // virtual head (row of the GQA group)
              if (r >= G) continue;
              const float m = hh ? mh1 : mh0, l = hh ? lh1 : lh0;
              const float inv = (l > 0.f) ? p.v_scale / l : 0.f;
              const float lse = (l > 0.f) ? (m + log2f(l)) * kLn2 : -INFINITY;
              const int head = kv_head * G + r;
              const int sub = r / p.Gr, head_real = kv_head * p.Gr + r % p.Gr;
              // this thread holds d = 32cp + 2*row + {0 (block 2cp), 1 (block 2cp+1)} and the same + 16
              if (seg.slot < 0) {
                uint16_t* dst = reinterpret_cast<uint16_t*>(p.out) + seg.batch * p.out_tok_stride +
                                sub * p.out_sub_stride + static_cast<int64_t>(head_real) * D + 2 * row;
  #pragma unroll
                for (int cp = 0; cp < D / 32; ++cp) {
                  *reinterpret_cast<uint32_t*>(dst + 32 * cp) =
                      pack2<kQBf16>(o[2 * cp][hh] * inv, o[2 * cp + 1][hh] * inv);
                  *reinterpret_cast<uint32_t*>(dst + 32 * cp + 16) =
                      pack2<kQBf16>(o[2 * cp][2 + hh] * inv, o[2 * cp + 1][2 + hh] * inv);
                }
                if (p.lse_out != nullptr && row == 0)
                  p.lse_out[static_cast<int64_t>(seg.batch) * p.lse_req_stride + sub * p.Hq_real + head_real] = lse;
              } else {
                float* dst = p.ws_o + (static_cast<int64_t>(seg.slot) * p.Hq + head) * D + 2 * row;
  #pragma unroll
                for (int cp = 0; cp < D / 32; ++cp) {
                  *reinterpret_cast<float2*>(dst + 32 * cp) = make_float2(o[2 * cp][hh] * inv, o[2 * cp + 1][hh] * inv);
                  *reinterpret_cast<float2*>(dst + 32 * cp + 16) =
                      make_float2(o[2 * cp][2 + hh] * inv, o[2 * cp + 1][2 + hh] * inv);
                }
                if (row == 0) p.ws_lse[static_cast<int64_t>(seg.slot) * p.Hq + head] = lse;
              }
            }
          }
        }
      }  // segments
      __syncthreads();  // every role is done with the segment list
    }
  }
}

// ---------------------------------------------------------------------------------------
// LSE merge of the pieces of every request that was cut across CTAs
// (reference: triton_ops/decode_attention.py:492-552, sgl-kernel/csrc/cpu/decode.cpp:812-860).
// One warp per (request, q head).
constexpr int kMergeWarps = 8;
constexpr int kMergeCoopMin = 8;  // up to this many partials a single warp merges a head on its own

// One CTA (8 warps) per (request, group of p.merge_hpc q heads): the warps split the partials of a request between
// them -- at low batch a single request is cut across all persistent CTAs (up to n_ctas partials), and walking them
// with one warp made this kernel longer than stage 1 (72 us at bs 1 against 9 us; profiles/README.md) -- and combine
// through shared memory, one head after the other.  merge_hpc is chosen on the host so that large batches launch
// about a thousand CTAs (one CTA per (request, head) cost 10 us of pure CTA turnover at bs 256 x 32 heads): requests
// that one stage-1 CTA finished -- the common case there -- leave after two loads for the whole head group.
template <int D, bool kBf16>
__global__ void __launch_bounds__(kMergeWarps * 32) decode_merge_kernel(const DecodeParams p) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int groups = p.Hq / p.merge_hpc;
  const int i = blockIdx.x / groups, head0 = (blockIdx.x % groups) * p.merge_hpc;
  const int32_t* plan = p.plan;
  const int W = plan[kPlanUnitsPerCta];
  const int32_t* cum = plan + kPlanHeader;
  const int c0 = cum[i], c1 = cum[i + 1];
  constexpr int EPL = D / 32;  // elements per lane
  const bool empty = c1 == c0;
  const int cta0 = c0 / W, cta1 = empty ? cta0 : (c1 - 1) / W;
  const int n = cta1 - cta0 + 1;
  if (!empty && n == 1) return;  // written directly by the decode kernel
  __shared__ float s_red[kMergeWarps];
  __shared__ float s_acc[kMergeWarps][D];
  auto slot_of = [&](int k) { return k == 0 ? 2 * cta0 + 1 : 2 * (cta0 + k); };
  if (empty || n <= kMergeCoopMin) {
    // few partials (the usual 2 of a request that straddles one share boundary) or none: one warp per head, no
    // shared memory, no CTA barrier -- the heads of the group proceed in parallel
    for (int head = head0 + warp; head < head0 + p.merge_hpc; head += kMergeWarps) {
      const int kvh = head / p.G, r = head % p.G;
      const int sub = r / p.Gr, head_real = kvh * p.Gr + r % p.Gr;
      uint16_t* dst = reinterpret_cast<uint16_t*>(p.out) + i * p.out_tok_stride + sub * p.out_sub_stride +
                      static_cast<int64_t>(head_real) * D;
      float* lse_dst = p.lse_out ? p.lse_out + static_cast<int64_t>(i) * p.lse_req_stride + sub * p.Hq_real + head_real : nullptr;
      if (empty) {  // empty request: defined as zeros
#pragma unroll
        for (int e = 0; e < EPL; ++e) dst[lane * EPL + e] = 0;
        if (lse_dst != nullptr && lane == 0) *lse_dst = -INFINITY;
        continue;
      }
      float mx = -INFINITY;
      for (int k = lane; k < n; k += 32) mx = fmaxf(mx, p.ws_lse[static_cast<int64_t>(slot_of(k)) * p.Hq + head]);
#pragma unroll
      for (int s = 16; s >= 1; s >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, s));
      float acc[EPL];
#pragma unroll
      for (int e = 0; e < EPL; ++e) acc[e] = 0.f;
      float wsum = 0.f;
#pragma unroll 4
      for (int k = 0; k < n; ++k) {
        const int64_t slot = slot_of(k);
        const float lse = p.ws_lse[slot * p.Hq + head];
        const float w = (lse == -INFINITY) ? 0.f : __expf(lse - mx);
        const float* src = p.ws_o + (slot * p.Hq + head) * D + lane * EPL;
        if constexpr (EPL == 4) {
          const float4 x = *reinterpret_cast<const float4*>(src);
          acc[0] += w * x.x; acc[1] += w * x.y; acc[2] += w * x.z; acc[3] += w * x.w;
        } else {
          const float2 x = *reinterpret_cast<const float2*>(src);
          acc[0] += w * x.x; acc[1] += w * x.y;
        }
        wsum += w;
      }
      const float inv = wsum > 0.f ? 1.f / wsum : 0.f;
#pragma unroll
      for (int e = 0; e < EPL; e += 2)
        *reinterpret_cast<uint32_t*>(dst + lane * EPL + e) = pack2<kBf16>(acc[e] * inv, acc[e + 1] * inv);
      if (lse_dst != nullptr && lane == 0) *lse_dst = wsum > 0.f ? mx + __logf(wsum) : -INFINITY;
    }
    return;
  }
  // many partials: the eight warps split them, one head after the other
  for (int head = head0; head < head0 + p.merge_hpc; ++head) {
    const int kvh = head / p.G, r = head % p.G;  // virtual head -> (token, real q head), see DecodeParams
    const int sub = r / p.Gr, head_real = kvh * p.Gr + r % p.Gr;
    uint16_t* dst = reinterpret_cast<uint16_t*>(p.out) + i * p.out_tok_stride + sub * p.out_sub_stride +
                    static_cast<int64_t>(head_real) * D;
    float* lse_dst = p.lse_out ? p.lse_out + static_cast<int64_t>(i) * p.lse_req_stride + sub * p.Hq_real + head_real : nullptr;
    float mx = -INFINITY;
    for (int k = threadIdx.x; k < n; k += kMergeWarps * 32)
      mx = fmaxf(mx, p.ws_lse[static_cast<int64_t>(slot_of(k)) * p.Hq + head]);
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, s));
    if (lane == 0) s_red[warp] = mx;
    __syncthreads();
#pragma unroll
    for (int w = 0; w < kMergeWarps; ++w) mx = fmaxf(mx, s_red[w]);
    __syncthreads();  // s_red is reused for the weight sums
    float acc[EPL];
#pragma unroll
    for (int e = 0; e < EPL; ++e) acc[e] = 0.f;
    float wsum = 0.f;
#pragma unroll 4
    for (int k = warp; k < n; k += kMergeWarps) {
      const int64_t slot = slot_of(k);
      const float lse = p.ws_lse[slot * p.Hq + head];
      const float w = (lse == -INFINITY) ? 0.f : __expf(lse - mx);
      const float* src = p.ws_o + (slot * p.Hq + head) * D + lane * EPL;
      if constexpr (EPL == 4) {
        const float4 x = *reinterpret_cast<const float4*>(src);
        acc[0] += w * x.x; acc[1] += w * x.y; acc[2] += w * x.z; acc[3] += w * x.w;
      } else {
        const float2 x = *reinterpret_cast<const float2*>(src);
        acc[0] += w * x.x; acc[1] += w * x.y;
      }
      wsum += w;
    }
#pragma unroll
    for (int e = 0; e < EPL; ++e) s_acc[warp][lane * EPL + e] = acc[e];
    if (lane == 0) s_red[warp] = wsum;
    __syncthreads();
    if (warp == 0) {
      wsum = 0.f;
#pragma unroll
      for (int e = 0; e < EPL; ++e) acc[e] = 0.f;
#pragma unroll
      for (int w = 0; w < kMergeWarps; ++w) {
        wsum += s_red[w];
#pragma unroll
        for (int e = 0; e < EPL; ++e) acc[e] += s_acc[w][lane * EPL + e];
      }
      const float inv = wsum > 0.f ? 1.f / wsum : 0.f;
#pragma unroll
      for (int e = 0; e < EPL; e += 2)
        *reinterpret_cast<uint32_t*>(dst + lane * EPL + e) = pack2<kBf16>(acc[e] * inv, acc[e + 1] * inv);
      if (lse_dst != nullptr && lane == 0) *lse_dst = wsum > 0.f ? mx + __logf(wsum) : -INFINITY;
    }
    __syncthreads();  // s_red / s_acc are rewritten for the next head
  }
}

// stage count: as many stages as fit next to the merge buffer (at least 2, at most 6; 3 are enough for 64 KiB stages)
template <int D, int HB, int KVB, int CW, bool kTokM, bool kBig = false>
constexpr int decode_stages() {
  constexpr int TW = (kTokM && kBig) ? CW : kTileWarps;
  constexpr int kStatic = 2048;
  int n = (2 * (16 * (TW / HB)) * HB * D * KVB >= 65536) ? 3 : 6;
  while (n > 2) {
    const int stageb = 2 * (16 * (TW / HB)) * HB * D * KVB;
    const int tstg = (TW / HB) * (CW / TW);
    const int slot_f = kTokM ? (D / 16 * 4 + 4) * 32 : 16 * (D + 2);
    const int redb = tstg > 1 ? ((kTokM && kBig) ? 4 : CW / 2) * slot_f * 4 : 0;
    const int total = n * stageb + redb + 2 * 16 * (TW / HB) * 4 + 2 * n * 8 + 128 + 1024;
    if (total + kStatic <= 227 * 1024) break;
    --n;
  }
  // Two tile groups (CW = 16): the stage count must be EVEN so that a stage always serves the same group.  With an odd
  // ring a group meets a stage only on every second phase and waits on the same parity each time; a `full` barrier
  // that is still one fill behind then reads as complete (parity aliasing).  The row gathers complete in issue order,
  // which hid this; bulk-tensor copies complete out of order -- tile t + 3 can land before tile t -- and large FP8
  // batches on the paged-tile path read stale stages (found with bench.py's full-size config-5 leg at N = 2).
  if (CW / TW > 1) n -= n % (CW / TW);
  return n;
}

// B200ATTN_DECODE_TMA=0 keeps page_size >= 16 pools on the row-gather producers (for A/B timing; default 1)
static bool decode_tma_enabled() {
  static const bool on = [] { const char* e = getenv("B200ATTN_DECODE_TMA"); return e ? atoi(e) != 0 : true; }();
  return on;
}

// Tensor maps of the K and V pools for the paged-tile producers.  Encoding is host-only work (about a microsecond);
// the last few (pool, geometry) pairs are cached per thread -- a decode step walks the same layers every time.
struct TmapKey {
  const void* k; const void* v; int64_t row_bytes; int kvb, row_segs, box_segs;
  bool operator==(const TmapKey& o) const {
    return k == o.k && v == o.v && row_bytes == o.row_bytes && kvb == o.kvb && row_segs == o.row_segs && box_segs == o.box_segs;
  }
};
static int decode_tmaps(const DecodeParams& p, int kvb, int D, int HB, const DecodeTmaps** out, int* box_segs_out) {
  constexpr int kSlots = 256;
  struct Entry { TmapKey key; DecodeTmaps maps; bool used; };
  static thread_local Entry cache[kSlots];
  const int row_segs = p.Hkv * D * kvb / 128;
  const int box_segs = HB * D * kvb / 128;
  const TmapKey key{p.k_pool, p.v_pool, p.pool_tok_stride * kvb, kvb, row_segs, box_segs};
  const size_t h = (reinterpret_cast<uintptr_t>(p.k_pool) >> 8) * 0x9E3779B97F4A7C15ull;
  Entry& e = cache[(h >> 40) % kSlots];
  if (!e.used || !(e.key == key)) {
    int rc = encode_pool_tmap(&e.maps.k, p.k_pool, kvb, key.row_bytes, row_segs, 16, box_segs);
    if (rc != B200ATTN_OK) return rc;
    rc = encode_pool_tmap(&e.maps.v, p.v_pool, kvb, key.row_bytes, row_segs, 16, box_segs);
    if (rc != B200ATTN_OK) return rc;
    e.key = key;
    e.used = true;
  }
  *out = &e.maps;
  *box_segs_out = box_segs;
  return B200ATTN_OK;
}

template <int D, int HB, int KVB, bool kQBf16, bool kG8, int CW, bool kTokM, bool kTma, bool kBig = false>
static int launch_decode_cw(const DecodeParams& p_in, int head_groups, cudaStream_t st) {
  constexpr int kStatic = 2048;
  constexpr int NSTAGE = decode_stages<D, HB, KVB, CW, kTokM, kBig>();
  using Cfg = DecodeCfg<D, HB, KVB, NSTAGE, CW, kTokM, kBig>;
  static_assert(Cfg::SMEM + kStatic <= 227 * 1024, "decode staging exceeds the shared memory of one CTA");
  DecodeParams p = p_in;
  static const DecodeTmaps no_maps{};
  const DecodeTmaps* maps = &no_maps;
  if constexpr (kTma) {
    const int rc = decode_tmaps(p, KVB, D, HB, &maps, &p.tma_box_segs);
    if (rc != B200ATTN_OK) return rc;
    p.tma_order = 0;
  }
  auto kern = decode_kernel<D, HB, KVB, kQBf16, kG8, NSTAGE, CW, kTokM, kTma, kBig>;
  static bool configured = false;  // per instantiation
  if (!configured) {
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    configured = true;
  }
  dim3 grid(p.n_ctas, head_groups, 1);
  kern<<<grid, Cfg::THREADS, Cfg::SMEM, st>>>(p, *maps);
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}

// B200ATTN_DECODE_FP8_TOKM=0 keeps FP8 pools on the head-major 8-warp consumer (for A/B timing; default 1)
static bool fp8_tokm_enabled() {
  static const bool on = [] { const char* e = getenv("B200ATTN_DECODE_FP8_TOKM"); return e ? atoi(e) != 0 : true; }();
  return on;
}

// B200ATTN_DECODE_FP8_BIG: 0 = two tile groups over 32 KiB stages (the form before round 2's last change), 1 = 64 KiB
// stages for head blocks >= 2, 2 = for every head block (default; one head per CTA: 256-token tiles, +2 % at C4 / TP 8,
// neutral at low batch)
static int fp8_big_level() {
  static const int lv = [] { const char* e = getenv("B200ATTN_DECODE_FP8_BIG"); return e ? atoi(e) : 2; }();
  return lv;
}

template <int D, int HB, int KVB, bool kQBf16, bool kG8, bool kTma>
static int launch_decode_t(const DecodeParams& p, int head_groups, cudaStream_t st) {
  if constexpr (KVB == 1 && kG8) {
    if (fp8_tokm_enabled()) {
      // 64 KiB stages shared by all 16 warps (32 / 64 / 128 / 256 tokens per tile for 8 / 4 / 2 / 1 heads per CTA)
      if constexpr (D == 128) {
        if (fp8_big_level() >= (HB >= 2 ? 1 : 2))
          return launch_decode_cw<D, HB, KVB, kQBf16, kG8, 16, true, kTma, true>(p, head_groups, st);
      }
      return launch_decode_cw<D, HB, KVB, kQBf16, kG8, 16, true, kTma>(p, head_groups, st);
    }
  }
  return launch_decode_cw<D, HB, KVB, kQBf16, kG8, 8, false, kTma>(p, head_groups, st);
}

template <int D, int HB, int KVB, bool kQBf16, bool kG8>
static int launch_decode(const DecodeParams& p, int head_groups, cudaStream_t st) {
  // paged-tile loads: D = 128 (head rows of 128 or 256 bytes = whole swizzle segments), pages of a multiple of 16
  if constexpr (D == 128) {
    if (p.page_size >= 16 && p.page_size % 16 == 0 && decode_tma_enabled())
      return launch_decode_t<D, HB, KVB, kQBf16, kG8, true>(p, head_groups, st);
  }
  return launch_decode_t<D, HB, KVB, kQBf16, kG8, false>(p, head_groups, st);
}

template <int D, int KVB, bool kQBf16, bool kG8>
static int dispatch_hb(const DecodeParams& p, int Hkv, cudaStream_t st) {
  if (Hkv % 8 == 0) return launch_decode<D, 8, KVB, kQBf16, kG8>(p, Hkv / 8, st);
  if (Hkv % 4 == 0) return launch_decode<D, 4, KVB, kQBf16, kG8>(p, Hkv / 4, st);
  if (Hkv % 2 == 0) return launch_decode<D, 2, KVB, kQBf16, kG8>(p, Hkv / 2, st);
  return launch_decode<D, 1, KVB, kQBf16, kG8>(p, Hkv, st);
}

template <int D, int KVB>
static int dispatch_q(const DecodeParams& p, int Hkv, bool q_bf16, cudaStream_t st) {
  const bool g8 = p.G <= 8;
  if (q_bf16) return g8 ? dispatch_hb<D, KVB, true, true>(p, Hkv, st) : dispatch_hb<D, KVB, true, false>(p, Hkv, st);
  return g8 ? dispatch_hb<D, KVB, false, true>(p, Hkv, st) : dispatch_hb<D, KVB, false, false>(p, Hkv, st);
}

}  // namespace b200attn

using namespace b200attn;

extern "C" B2_EXPORT size_t b200attn_decode_workspace_bytes(int n_ctas, int Hq, int Dv) {
  if (n_ctas <= 0 || Hq <= 0 || Dv <= 0) return 0;
  return static_cast<size_t>(2) * n_ctas * Hq * (static_cast<size_t>(Dv) + 1) * sizeof(float);
}

namespace b200attn {

// argument checks shared by the three decode entry points; fills DecodeParams
static int decode_setup(DecodeParams& p, const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride,
                        const void* k_pool, const void* v_pool, int64_t pool_tok_stride, const int32_t* req_to_token,
                        int64_t table_stride, const void* req_pool_indices, const void* seq_lens, int idx_is_i64,
                        int bs, int Hq, int Hkv, int D, int q_dtype, int kv_dtype, float sm_scale, float logit_cap,
                        float k_scale, float v_scale, const int32_t* plan, int n_ctas, void* workspace,
                        size_t workspace_bytes, float* lse_out, bool need_inputs) {
  B2_REQUIRE(bs >= 0 && Hq > 0 && Hkv > 0, B200ATTN_ERR_INVALID, "decode: bad sizes bs=%d Hq=%d Hkv=%d", bs, Hq, Hkv);
  B2_REQUIRE(out && plan && workspace, B200ATTN_ERR_INVALID, "decode: null pointer");
  if (need_inputs)
    B2_REQUIRE(q && k_pool && v_pool && req_to_token && req_pool_indices && seq_lens, B200ATTN_ERR_INVALID,
               "decode: null pointer");
  B2_REQUIRE(Hq % Hkv == 0, B200ATTN_ERR_INVALID, "decode: Hq=%d not a multiple of Hkv=%d", Hq, Hkv);
  const int G = Hq / Hkv;
  B2_REQUIRE(G <= 16, B200ATTN_ERR_UNSUPPORTED, "decode: GQA group %d > 16 is not built", G);
  B2_REQUIRE(D == 64 || D == 128, B200ATTN_ERR_UNSUPPORTED, "decode: head dim %d (64/128 only)", D);
  B2_REQUIRE(q_dtype == B200ATTN_BF16 || q_dtype == B200ATTN_F16, B200ATTN_ERR_UNSUPPORTED, "decode: q dtype %d", q_dtype);
  B2_REQUIRE(kv_dtype == q_dtype || kv_dtype == B200ATTN_FP8_E4M3, B200ATTN_ERR_UNSUPPORTED,
             "decode: kv dtype %d with q dtype %d", kv_dtype, q_dtype);
  const int kvb = elem_size(kv_dtype);
  B2_REQUIRE(aligned16(q) && aligned16(out) && aligned16(k_pool) && aligned16(v_pool) && q_tok_stride % 8 == 0 &&
                 out_tok_stride % 8 == 0 && (pool_tok_stride * kvb) % 16 == 0,
             B200ATTN_ERR_INVALID, "decode: tensors must be 16-byte aligned");
  B2_REQUIRE(pool_tok_stride >= static_cast<int64_t>(Hkv) * D, B200ATTN_ERR_INVALID, "decode: pool_tok_stride < Hkv*D");
  B2_REQUIRE(n_ctas > 0, B200ATTN_ERR_INVALID, "decode: n_ctas=%d", n_ctas);
  B2_REQUIRE(workspace_bytes >= b200attn_decode_workspace_bytes(n_ctas, Hq, D), B200ATTN_ERR_WORKSPACE,
             "decode: workspace %zu < %zu bytes", workspace_bytes, b200attn_decode_workspace_bytes(n_ctas, Hq, D));
  B2_REQUIRE(k_scale > 0.f && v_scale > 0.f, B200ATTN_ERR_INVALID, "decode: k_scale/v_scale must be > 0");

  p.q = q; p.out = out; p.k_pool = k_pool; p.v_pool = v_pool;
  p.req_to_token = req_to_token; p.req_pool_indices = req_pool_indices; p.seq_lens = seq_lens;
  p.plan = plan;
  p.ws_o = reinterpret_cast<float*>(workspace);
  p.ws_lse = p.ws_o + static_cast<size_t>(2) * n_ctas * Hq * D;
  p.lse_out = lse_out;
  p.q_tok_stride = q_tok_stride; p.out_tok_stride = out_tok_stride;
  p.pool_tok_stride = pool_tok_stride; p.table_stride = table_stride;
  p.idx_is_i64 = idx_is_i64;
  p.bs = bs; p.Hq = Hq; p.Hkv = Hkv; p.G = G;
  p.merge_hpc = 1;
  p.Gr = G; p.Hq_real = Hq; p.q_sub_stride = 0; p.out_sub_stride = 0; p.lse_req_stride = Hq;
  p.n_ctas = n_ctas;
  p.scale_raw = sm_scale * k_scale;
  p.scale_log2 = sm_scale * k_scale * kLog2e;
  p.logit_cap = logit_cap;
  p.v_scale = v_scale;
  p.tma_box_segs = 1; p.tma_order = 0; p.page_size = 1;
  return B200ATTN_OK;
}

static int launch_partial(const DecodeParams& p, int D, int kvb, bool q_bf16, cudaStream_t st) {
  if (D == 128) return (kvb == 2) ? dispatch_q<128, 2>(p, p.Hkv, q_bf16, st) : dispatch_q<128, 1>(p, p.Hkv, q_bf16, st);
  return (kvb == 2) ? dispatch_q<64, 2>(p, p.Hkv, q_bf16, st) : dispatch_q<64, 1>(p, p.Hkv, q_bf16, st);
}

static int launch_merge(const DecodeParams& p, int D, bool q_bf16, cudaStream_t st) {
  // heads per CTA: the largest power of two dividing Hq that keeps about a thousand CTAs in the grid
  int hpc = 1;
  while (hpc * 2 <= p.Hq && p.Hq % (hpc * 2) == 0 && static_cast<int64_t>(p.bs) * p.Hq / (hpc * 2) >= 1024) hpc *= 2;
  DecodeParams pm = p;
  pm.merge_hpc = hpc;
  const int blocks = p.bs * (p.Hq / hpc);
  constexpr int kT = kMergeWarps * 32;
  if (D == 128) {
    if (q_bf16) decode_merge_kernel<128, true><<<blocks, kT, 0, st>>>(pm);
    else decode_merge_kernel<128, false><<<blocks, kT, 0, st>>>(pm);
  } else {
    if (q_bf16) decode_merge_kernel<64, true><<<blocks, kT, 0, st>>>(pm);
    else decode_merge_kernel<64, false><<<blocks, kT, 0, st>>>(pm);
  }
  B2_CUDA(cudaGetLastError());
  return B200ATTN_OK;
}

}  // namespace b200attn

#define B2_DECODE_ARGS                                                                                            \
  q, q_tok_stride, out, out_tok_stride, k_pool, v_pool, pool_tok_stride, req_to_token, table_stride,              \
      req_pool_indices, seq_lens, idx_is_i64, bs, Hq, Hkv, D, q_dtype, kv_dtype, sm_scale, logit_cap, k_scale,    \
      v_scale, plan, n_ctas, workspace, workspace_bytes, lse_out

#define B2_DECODE_PARAMS                                                                                          \
  const void *q, int64_t q_tok_stride, void *out, int64_t out_tok_stride, const void *k_pool, const void *v_pool, \
      int64_t pool_tok_stride, const int32_t *req_to_token, int64_t table_stride, const void *req_pool_indices,  \
      const void *seq_lens, int idx_is_i64, int bs, int Hq, int Hkv, int D, int q_dtype, int kv_dtype,           \
      float sm_scale, float logit_cap, float k_scale, float v_scale, const int32_t *plan, int n_ctas,            \
      void *workspace, size_t workspace_bytes, float *lse_out

static int decode_entry(B2_DECODE_PARAMS, int page_size, bool merge, b200attn_stream_t stream) {
  if (bs == 0) return B200ATTN_OK;
  B2_REQUIRE(page_size >= 1, B200ATTN_ERR_INVALID, "decode: page_size=%d", page_size);
  DecodeParams p;
  const int rc = decode_setup(p, B2_DECODE_ARGS, true);
  if (rc != B200ATTN_OK) return rc;
  p.page_size = page_size;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool q_bf16 = q_dtype == B200ATTN_BF16;
  const int rc2 = launch_partial(p, D, elem_size(kv_dtype), q_bf16, st);
  if (rc2 != B200ATTN_OK || !merge) return rc2;
  return launch_merge(p, D, q_bf16, st);
}

extern "C" B2_EXPORT int b200attn_decode(B2_DECODE_PARAMS, b200attn_stream_t stream) {
  return decode_entry(B2_DECODE_ARGS, 1, true, stream);
}

// stage 1 only: the persistent split-KV kernel (final outputs of unsplit requests + fp32 partials)
extern "C" B2_EXPORT int b200attn_decode_partial(B2_DECODE_PARAMS, b200attn_stream_t stream) {
  return decode_entry(B2_DECODE_ARGS, 1, false, stream);
}

// b200attn_decode (stage = 0) / b200attn_decode_partial (stage = 1) for a pool whose allocator hands out pages of
// page_size consecutive slots: with page_size % 16 == 0 and D = 128 the producers read one table entry per 16-token
// unit and fetch the unit's K and V rows as bulk-tensor (TMA) boxes.
extern "C" B2_EXPORT int b200attn_decode_paged(B2_DECODE_PARAMS, int page_size, int stage, b200attn_stream_t stream) {
  B2_REQUIRE(stage == 0 || stage == 1, B200ATTN_ERR_INVALID, "decode_paged: stage=%d", stage);
  return decode_entry(B2_DECODE_ARGS, page_size, stage == 0, stream);
}

// stage 2 only: LSE merge of the partials of requests that were cut across CTAs
extern "C" B2_EXPORT int b200attn_decode_merge(void* out, int64_t out_tok_stride, int bs, int Hq, int D, int q_dtype,
                                               const int32_t* plan, int n_ctas, void* workspace,
                                               size_t workspace_bytes, float* lse_out, b200attn_stream_t stream) {
  if (bs == 0) return B200ATTN_OK;
  DecodeParams p;
  const int rc = decode_setup(p, out, 8, out, out_tok_stride, out, out, static_cast<int64_t>(Hq) * D, nullptr, 0,
                              nullptr, nullptr, 0, bs, Hq, Hq, D, q_dtype, q_dtype, 1.f, 0.f, 1.f, 1.f, plan, n_ctas,
                              workspace, workspace_bytes, lse_out, false);
  if (rc != B200ATTN_OK) return rc;
  return launch_merge(p, D, q_dtype == B200ATTN_BF16, static_cast<cudaStream_t>(stream));
}

// nq query tokens per request over the SAME seq_lens[i] keys, no mask among them: the cached-prefix part of a
// speculative-decoding verify batch (TARGET_VERIFY, triton_backend.py:227-264), where every draft token sees the whole
// prefix.  The nq * G rows of a kv head ride the MMA row dimension the GQA group already uses, so K/V are streamed from
// HBM once for all of them (nq * Hq / Hkv <= 16).  The caller merges the draft-token part with b200attn_merge_state.
static int decode_multiq_entry(int page_size,
    const void* q, int64_t q_req_stride, int64_t q_tok_stride, void* out, int64_t out_req_stride, int64_t out_tok_stride,
    const void* k_pool, const void* v_pool, int64_t pool_tok_stride, const int32_t* req_to_token, int64_t table_stride,
    const void* req_pool_indices, const void* seq_lens, int idx_is_i64, int bs, int nq, int Hq, int Hkv, int D,
    int q_dtype, int kv_dtype, float sm_scale, float logit_cap, float k_scale, float v_scale, const int32_t* plan,
    int n_ctas, void* workspace, size_t workspace_bytes, float* lse_out, int64_t lse_req_stride,
    b200attn_stream_t stream) {
  if (bs == 0 || nq == 0) return B200ATTN_OK;
  B2_REQUIRE(page_size >= 1, B200ATTN_ERR_INVALID, "decode_multiq: page_size=%d", page_size);
  B2_REQUIRE(nq > 0 && Hq > 0 && Hkv > 0 && Hq % Hkv == 0 && nq * (Hq / Hkv) <= 16, B200ATTN_ERR_UNSUPPORTED,
             "decode_multiq: nq * Hq / Hkv = %d * %d / %d must be <= 16", nq, Hq, Hkv);
  B2_REQUIRE(q_tok_stride % 8 == 0 && out_tok_stride % 8 == 0, B200ATTN_ERR_INVALID,
             "decode_multiq: token strides must be multiples of 8 elements");
  DecodeParams p;
  const int rc = decode_setup(p, q, q_req_stride, out, out_req_stride, k_pool, v_pool, pool_tok_stride, req_to_token,
                              table_stride, req_pool_indices, seq_lens, idx_is_i64, bs, Hq * nq, Hkv, D, q_dtype,
                              kv_dtype, sm_scale, logit_cap, k_scale, v_scale, plan, n_ctas, workspace,
                              workspace_bytes, lse_out, true);
  if (rc != B200ATTN_OK) return rc;
  p.Gr = Hq / Hkv; p.Hq_real = Hq;
  p.q_sub_stride = q_tok_stride; p.out_sub_stride = out_tok_stride;
  p.lse_req_stride = lse_req_stride;
  p.page_size = page_size;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool q_bf16 = q_dtype == B200ATTN_BF16;
  const int rc2 = launch_partial(p, D, elem_size(kv_dtype), q_bf16, st);
  if (rc2 != B200ATTN_OK) return rc2;
  return launch_merge(p, D, q_bf16, st);
}

#define B2_MULTIQ_PARAMS                                                                                         \
  const void *q, int64_t q_req_stride, int64_t q_tok_stride, void *out, int64_t out_req_stride,                 \
      int64_t out_tok_stride, const void *k_pool, const void *v_pool, int64_t pool_tok_stride,                  \
      const int32_t *req_to_token, int64_t table_stride, const void *req_pool_indices, const void *seq_lens,    \
      int idx_is_i64, int bs, int nq, int Hq, int Hkv, int D, int q_dtype, int kv_dtype, float sm_scale,        \
      float logit_cap, float k_scale, float v_scale, const int32_t *plan, int n_ctas, void *workspace,          \
      size_t workspace_bytes, float *lse_out, int64_t lse_req_stride
#define B2_MULTIQ_ARGS                                                                                           \
  q, q_req_stride, q_tok_stride, out, out_req_stride, out_tok_stride, k_pool, v_pool, pool_tok_stride,          \
      req_to_token, table_stride, req_pool_indices, seq_lens, idx_is_i64, bs, nq, Hq, Hkv, D, q_dtype, kv_dtype, \
      sm_scale, logit_cap, k_scale, v_scale, plan, n_ctas, workspace, workspace_bytes, lse_out, lse_req_stride

extern "C" B2_EXPORT int b200attn_decode_multiq(B2_MULTIQ_PARAMS, b200attn_stream_t stream) {
  return decode_multiq_entry(1, B2_MULTIQ_ARGS, stream);
}

// b200attn_decode_multiq for a paged pool (see b200attn_decode_paged)
extern "C" B2_EXPORT int b200attn_decode_multiq_paged(B2_MULTIQ_PARAMS, int page_size, b200attn_stream_t stream) {
  return decode_multiq_entry(page_size, B2_MULTIQ_ARGS, stream);
}

// ---------------------------------------------------------------------------------------------------------------
// b200attn_decode_static_ws: b200attn_decode / b200attn_decode_paged for callers that have no way to get device
// memory for the split-KV partials.  No counterpart in the reference.
//
// Who that is: a host that binds this library from another runtime (cgo, JNI, a ctypes stub inside a server whose
// own allocator has already claimed the GPU's memory for the KV pool) holds only the tensors it was handed -- q, out,
// the pools, the tables.  It cannot call cudaMalloc itself: it may not link the CUDA runtime at all, the memory may
// simply be gone (SGLang sizes the pool to fill what is free), and cudaMalloc / cudaFree stall every stream of the
// device while they run, which a serving loop cannot afford per step.  Such a caller has nowhere to put the fp32
// partials the split-KV kernel hands to the merge kernel.
//
// So the partials live in the array below, which is part of the library image: every GPU on which the library's
// kernels get loaded carries these B200ATTN_DECODE_STATIC_WS_BYTES until the process exits, whether or not this
// entry point is ever called.  Nothing in the array has to start at zero (stage 1 writes every slot the merge
// reads, exactly as with a caller-provided torch.empty workspace), so nothing is cleared between calls.
//
// `g_decode_static_ws` in HOST code is only a handle for the runtime's symbol table, not an address a kernel could
// use; the device address is looked up once per device with cudaGetSymbolAddress (static_ws_for_current_device).
static __device__ __align__(256) unsigned char g_decode_static_ws[B200ATTN_DECODE_STATIC_WS_BYTES];

namespace b200attn {

// Device address of g_decode_static_ws on the CURRENT device; resolved on the first call per device and remembered.
// The lookup makes the runtime load this module onto the device if it has not yet, which is why the first call on a
// device belongs before any stream capture.
static int static_ws_for_current_device(void** out) {
  constexpr int kMaxDevices = 64;
  static void* ptrs[kMaxDevices] = {};  // written once per slot; racing first calls store the same value
  int dev = -1;
  B2_CUDA(cudaGetDevice(&dev));
  B2_REQUIRE(dev >= 0 && dev < kMaxDevices, B200ATTN_ERR_UNSUPPORTED, "decode_static_ws: device ordinal %d >= %d", dev,
             kMaxDevices);
  void* ptr = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
  if (ptr == nullptr) {
    B2_CUDA(cudaGetSymbolAddress(&ptr, g_decode_static_ws));
    B2_REQUIRE(ptr != nullptr && (reinterpret_cast<uintptr_t>(ptr) & 255u) == 0, B200ATTN_ERR_CUDA,
               "decode_static_ws: cudaGetSymbolAddress returned %p for the static workspace", ptr);
    __atomic_store_n(&ptrs[dev], ptr, __ATOMIC_RELEASE);
  }
  *out = ptr;
  return B200ATTN_OK;
}

}  // namespace b200attn

extern "C" B2_EXPORT size_t b200attn_decode_static_ws_capacity(void) { return B200ATTN_DECODE_STATIC_WS_BYTES; }

extern "C" B2_EXPORT int b200attn_decode_static_ws(
    const void* q, int64_t q_tok_stride, void* out, int64_t out_tok_stride, const void* k_pool, const void* v_pool,
    int64_t pool_tok_stride, const int32_t* req_to_token, int64_t table_stride, const void* req_pool_indices,
    const void* seq_lens, int idx_is_i64, int bs, int Hq, int Hkv, int D, int q_dtype, int kv_dtype, float sm_scale,
    float logit_cap, float k_scale, float v_scale, const int32_t* plan, int n_ctas, float* lse_out, int page_size,
    b200attn_stream_t stream) {
  if (bs == 0) return B200ATTN_OK;
  const size_t need = b200attn_decode_workspace_bytes(n_ctas, Hq, D);
  B2_REQUIRE(need > 0, B200ATTN_ERR_INVALID, "decode_static_ws: bad sizes n_ctas=%d Hq=%d D=%d", n_ctas, Hq, D);
  B2_REQUIRE(need <= static_cast<size_t>(B200ATTN_DECODE_STATIC_WS_BYTES), B200ATTN_ERR_WORKSPACE,
             "decode_static_ws: n_ctas=%d x Hq=%d x D=%d needs %zu bytes of partials, the built-in workspace holds %zu; "
             "use b200attn_decode with a caller-provided workspace (or fewer CTAs)",
             n_ctas, Hq, D, need, static_cast<size_t>(B200ATTN_DECODE_STATIC_WS_BYTES));
  void* workspace = nullptr;
  const int rc = static_ws_for_current_device(&workspace);
  if (rc != B200ATTN_OK) return rc;
  const size_t workspace_bytes = B200ATTN_DECODE_STATIC_WS_BYTES;
  return decode_entry(B2_DECODE_ARGS, page_size, true, stream);
}

#ifdef B200ATTN_DEBUG_WAIT
// development build only: timed-out mbarrier waits recorded by the decode kernels (see tc_ptx.cuh); returns the count
extern "C" B2_EXPORT int b200attn_debug_wait_dump(unsigned int* host_out, int clear) {
  unsigned int buf[4 * 64 + 1];
  if (cudaMemcpyFromSymbol(buf, b200attn::ptx::g_wait_dbg, sizeof(buf)) != cudaSuccess) return -1;
  for (int i = 0; i < 4 * 64 + 1; ++i) host_out[i] = buf[i];
  if (clear) {
    unsigned int zero[4 * 64 + 1] = {0};
    cudaMemcpyToSymbol(b200attn::ptx::g_wait_dbg, zero, sizeof(zero));
  }
  return static_cast<int>(buf[0]);
}
#endif
