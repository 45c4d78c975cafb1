"""GPU parity of paged extend / prefill attention (prefix + new tokens, causal) through B200AttnBackend ->
the C ABI, against oracle #1 (torch_native) and the fp64 restatement; both kernels (register-MMA and
tcgen05/TMEM) are exercised explicitly through the `impl` switch of b200attn_extend."""
from __future__ import annotations

import pytest
import torch

import golden_util as gu
import helpers as H
from sglang_custom_b200 import _lib, ops, synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", [n for n in gu.ATTN_FIXTURES if n.startswith("extend")])
def test_extend_golden_fixtures(name, device):
    fx = gu.load(name)
    r = H.Runner(fx, device)
    mode = r.itf.ForwardMode.MIXED if name == "extend_mixed" else None
    out = r.forward(mode)
    kp, vp = r.pools()
    assert H.pools_equal(kp, fx.ref_k_pool_after) and H.pools_equal(vp, fx.ref_v_pool_after)
    H.assert_close(out, fx.ref_out, what=name)


CASES = [
    # bs, Hq, Hkv, D, seq_lens, extend_lens, kwargs
    (2, 32, 8, 128, [512, 384], [256, 384], {}),                       # C3-shaped heads: prefix + new
    (1, 16, 1, 128, [123], [123], {}),                                 # test_extend.py:183
    (4, 16, 4, 128, [1230, 700, 64, 129], [615, 1, 64, 128], {}),       # test_extend.py:185, ragged + a decode row
    (3, 8, 2, 64, [300, 200, 100], [100, 200, 37], {}),                # D = 64
    (2, 8, 1, 128, [600, 333], [200, 333], {"fp8": True}),             # FP8 KV (register-MMA kernel)
    (2, 8, 1, 128, [640, 320], [128, 64], {"fp8": True, "page_size": 16, "kv_scale": 0.1135}),
    (2, 12, 4, 128, [777, 130], [700, 130], {}),                       # 12/4 heads (test_triton_attention_kernels.py:174)
    (3, 4, 4, 128, [50, 257, 128], [50, 129, 1], {}),                  # MHA
    (2, 8, 2, 128, [90, 31], [40, 31], {"dtype": torch.float16}),
    (2, 32, 8, 128, [260, 384], [130, 256], {"idx_dtype": torch.int32}),
]


@pytest.mark.parametrize("bs,Hq,Hkv,D,lens,ext,kw", CASES)
def test_extend_vs_oracles(bs, Hq, Hkv, D, lens, ext, kw, device):
    case = synth.make_case(bs, Hq, Hkv, D, lens, extend_lens=ext, seed=1234 + bs + Hq, **kw)
    kp, vp = H.pool_after_write(case)
    r = H.Runner(case, device)
    out = r.forward()
    dk, dv = r.pools()
    assert H.pools_equal(dk, kp) and H.pools_equal(dv, vp), "KV write must be bit-exact"
    H.assert_close(out, H.oracle_extend(case, kp, vp), what="vs torch_native")
    H.assert_close(out, H.exact_extend(case, kp, vp), H.EXACT_MAX_ABS_TOL, H.EXACT_MEAN_ABS_TOL, what="vs fp64 restatement", rel=True)


def _direct(case, device, kp, vp, impl, causal=True, logit_cap=0.0, want_lse=False):
    c = H.to_device(case, device)
    out = torch.empty_like(c.q)
    lse = torch.empty((case.T, case.Hq), dtype=torch.float32, device=device) if want_lse else None
    s = case.kv_scale if case.fp8 else 1.0
    ops.extend(c.q, out, kp.to(device), vp.to(device), c.req_to_token, c.req_pool_indices, c.seq_lens,
               c.extend_seq_lens, c.extend_prefix_lens, c.extend_start_loc, case.max_extend_len, case.sm_scale,
               logit_cap=logit_cap, k_scale=s, v_scale=s, causal=causal,
               kv_dtype=torch.float8_e4m3fn if case.fp8 else None, impl=impl, lse_out=lse)
    torch.cuda.synchronize()
    return (out, lse) if want_lse else out


@pytest.mark.parametrize("impl", [_lib.EXTEND_MMA, _lib.EXTEND_TCGEN05])
@pytest.mark.parametrize("lens,ext", [([512, 300], [256, 300]), ([129], [129]), ([2048 + 64], [64]), ([1000, 1, 260], [999, 1, 130])])
def test_both_extend_kernels_agree_with_fp64(impl, lens, ext, device):
    case = synth.make_case(len(lens), 8, 2, 128, lens, extend_lens=ext, seed=sum(lens))
    kp, vp = H.pool_after_write(case)
    out, lse = _direct(case, device, kp, vp, impl, want_lse=True)
    want, want_lse = H.exact_extend(case, kp, vp, return_lse=True)
    H.assert_close(out, want, H.EXACT_MAX_ABS_TOL, H.EXACT_MEAN_ABS_TOL, what=f"impl={impl}", rel=True)
    assert torch.allclose(lse.cpu(), want_lse, atol=2e-3, rtol=1e-4)


@pytest.mark.parametrize("impl", [_lib.EXTEND_MMA, _lib.EXTEND_TCGEN05])
@pytest.mark.parametrize("kw", [dict(kv_scale=1.0), dict(kv_scale=0.1135), dict(kv_scale=0.1135, page_size=16),
                                dict(kv_scale=0.5, dtype=torch.float16)])
def test_fp8_kv_extend_both_kernels(impl, kw, device):
    """FP8-E4M3 KV cache under extend: the tcgen05 kernel expands e4m3 to f16 in its producer warps (exact) and
    runs the MMA in f16; the register-MMA kernel expands in registers.  Both against the fp64 restatement on the
    same pool bytes and against oracle #1 on the dequantised pool."""
    lens, ext = [640, 333, 1100, 130], [256, 333, 129, 2]
    case = synth.make_case(len(lens), 16, 2, 128, lens, extend_lens=ext, fp8=True, seed=4242, **kw)
    kp, vp = H.pool_after_write(case)
    out, lse = _direct(case, device, kp, vp, impl, want_lse=True)
    want, want_lse = H.exact_extend(case, kp, vp, return_lse=True)
    H.assert_close(out, want, H.EXACT_MAX_ABS_TOL, H.EXACT_MEAN_ABS_TOL, what=f"fp8 impl={impl}", rel=True)
    assert torch.allclose(lse.cpu(), want_lse, atol=2e-3, rtol=1e-4)
    H.assert_close(out, H.oracle_extend(case, kp, vp), what="fp8 vs torch_native on the dequantised pool")


@pytest.mark.parametrize("impl", [_lib.EXTEND_MMA, _lib.EXTEND_TCGEN05])
def test_extend_non_causal_encoder_only(impl, device):
    """ENCODER_ONLY layers attend bidirectionally (torch_native_backend.py:206-208)."""
    case = synth.make_case(2, 8, 2, 128, [200, 131], extend_lens=[200, 131], seed=9)
    kp, vp = H.pool_after_write(case)
    out = _direct(case, device, kp, vp, impl, causal=False)
    H.assert_close(out, H.oracle_extend(case, kp, vp, causal=False), what="non-causal vs torch_native")
    H.assert_close(out, H.exact_extend(case, kp, vp, causal=False), H.EXACT_MAX_ABS_TOL, H.EXACT_MEAN_ABS_TOL, rel=True)


def test_extend_logit_cap(device):
    case = synth.make_case(2, 8, 2, 128, [300, 77], extend_lens=[150, 77], seed=10)
    kp, vp = H.pool_after_write(case)
    out = _direct(case, device, kp, vp, _lib.EXTEND_AUTO, logit_cap=30.0)
    H.assert_close(out, H.exact_extend(case, kp, vp, logit_cap=30.0), H.EXACT_MAX_ABS_TOL, H.EXACT_MEAN_ABS_TOL, rel=True)


def test_mixed_batch_equals_decode_for_single_token_rows(device):
    """MIXED batches carry decode requests as E_i = 1 rows (schedule_batch.py:1310-1338): same attention as decode."""
    lens = [700, 33, 1500, 1]
    dec = synth.make_case(4, 16, 4, 128, lens, seed=31)
    ext = synth.make_case(4, 16, 4, 128, lens, extend_lens=[1, 1, 1, 1], seed=31)
    assert torch.equal(dec.q, ext.q) and torch.equal(dec.out_cache_loc, ext.out_cache_loc)
    o_dec = H.Runner(dec, device).forward()
    r = H.Runner(ext, device)
    o_ext = r.forward(r.itf.ForwardMode.MIXED)
    H.assert_close(o_ext, o_dec, 4e-3, 3e-4, what="mixed vs decode")


@pytest.mark.parametrize("kw", [dict(), dict(fp8=True, kv_scale=0.1135), dict(page_size=16)])
def test_mixed_batch_prefill_rows_plus_decode_tail(kw, device):
    """A real MIXED forward (schedule_batch.py:1310-1338): prefill requests first, the running decode requests
    appended with extend_len 1.  The backend serves the tail with the decode kernel and the head with the extend
    kernel; both routes must equal the oracle's single forward_extend."""
    lens = [384, 200, 1500, 33, 700, 1, 129]
    ext = [256, 200, 1, 1, 1, 1, 1]
    case = synth.make_case(len(lens), 16, 4, 128, lens, extend_lens=ext, seed=77, **kw)
    kp, vp = H.pool_after_write(case)
    want = H.oracle_extend(case, kp, vp)
    r = H.Runner(case, device)
    got = r.forward(r.itf.ForwardMode.MIXED)
    assert r.backend.forward_metadata.n_decode_tail == 5 and r.backend.forward_metadata.max_extend_len == 256
    H.assert_close(got, want, what="mixed (decode tail routed)")
    r2 = H.Runner(case, device)
    r2.backend.route_decode_tail = False
    got2 = r2.forward(r2.itf.ForwardMode.MIXED)
    assert r2.backend.forward_metadata.n_decode_tail == 0
    H.assert_close(got2, want, what="mixed (all rows through extend)")
    H.assert_close(got, got2, 8e-3, 6e-4, what="mixed: the two routes")
    # a batch of only 1-token rows under EXTEND mode: everything goes to the decode kernel
    case1 = synth.make_case(3, 16, 4, 128, [90, 1, 400], extend_lens=[1, 1, 1], seed=78, **kw)
    kp, vp = H.pool_after_write(case1)
    r3 = H.Runner(case1, device)
    got3 = r3.forward()
    assert r3.backend.forward_metadata.n_decode_tail == 3
    H.assert_close(got3, H.oracle_extend(case1, kp, vp), what="all-tail extend")


def test_c5_shaped_mixed_forward_scaled_down(device):
    """BASELINE config 5 at 1/8 scale, per-rank TP 8 shape (8 Q / 1 KV head, FP8 KV, page_size 16): one MIXED
    forward of chunked-prefill rows (prefix a multiple of the page size, 128-token chunk) followed by running decode
    rows with ragged lengths; the generator is the one scripts/time_mixed.py runs at full size."""
    g = torch.Generator().manual_seed(5)
    n_dec, n_ext = 31, 2
    ext_prefix = (16 * torch.randint(0, 1536 // 8, (n_ext,), generator=g)).tolist()
    dec_lens = torch.randint(1024 // 8, 32768 // 8 + 1, (n_dec,), generator=g).tolist()
    E = 128
    lens = [p + E for p in ext_prefix] + dec_lens
    ext = [E] * n_ext + [1] * n_dec
    case = synth.make_case(len(lens), 8, 1, 128, lens, extend_lens=ext, fp8=True, kv_scale=0.1135, page_size=16, seed=1239)
    kp, vp = H.pool_after_write(case)
    r = H.Runner(case, device)
    got = r.forward(r.itf.ForwardMode.MIXED)
    dk, dv = r.pools()
    assert H.pools_equal(dk, kp) and H.pools_equal(dv, vp), "FP8 KV write of the mixed forward"
    assert r.backend.forward_metadata.n_decode_tail == n_dec
    H.assert_close(got, H.oracle_extend(case, kp, vp), what="C5-shaped mixed forward")
    # the two-stream routes (decode tail on a side stream, either launch order) are bit-identical to the serial one
    for mode in (True, 2):
        r.backend.overlap_mixed = mode
        again = r.forward(r.itf.ForwardMode.MIXED)
        assert torch.equal(again, got), f"overlap_mixed={mode}"


def test_extend_chunked_prefill_equals_one_shot(device):
    """Chunked prefill (scheduler.py:452-457): chunk 2 sees chunk 1 as cached prefix; same result as one extend."""
    L = 700
    one = synth.make_case(1, 8, 2, 128, [L], extend_lens=[L], seed=41)
    r = H.Runner(one, device)
    o_one = r.forward()
    # chunk 1: first 300 tokens; chunk 2: remaining 400 with prefix 300 -- reuse the same table/pool/q
    c1 = synth.make_case(1, 8, 2, 128, [L], extend_lens=[L], seed=41)
    outs = []
    r2 = H.Runner(c1, device)
    import types
    for (p0, p1) in ((0, 300), (300, L)):
        ch = types.SimpleNamespace(**vars(c1))
        ch.T = p1 - p0
        ch.q, ch.k_new, ch.v_new = c1.q[p0:p1], c1.k_new[p0:p1], c1.v_new[p0:p1]
        ch.seq_lens = torch.tensor([p1]); ch.seq_lens_sum = p1
        ch.extend_seq_lens = torch.tensor([p1 - p0], dtype=torch.int32)
        ch.extend_prefix_lens = torch.tensor([p0], dtype=torch.int32)
        ch.extend_start_loc = torch.tensor([0], dtype=torch.int32)
        ch.out_cache_loc = c1.out_cache_loc[p0:p1]
        ch.max_extend_len = p1 - p0
        r2.case = ch
        outs.append(r2.forward())
    H.assert_close(torch.cat(outs), o_one, 4e-3, 3e-4, what="chunked vs one-shot")


def test_extend_full_size_c3_sampled_oracle(device):
    """BASELINE config 3 at full size (bs 16, prefix 2048 + new 2048, 32/8, D 128): a sample of query rows of two
    requests is checked against the fp64 restatement computed on the same bytes."""
    bs = 16
    case = synth.make_case_device(bs, 32, 8, 128, [4096] * bs, device, extend_lens=[2048] * bs, seed=1237)
    r = H.Runner(case, device)
    out = r.forward()
    dk, dv = r.pools()
    for req in (3, 15):
        sub = synth.extract_subcase(case, [req], dk, dv)
        # keep the oracle cheap: only 64 query rows spread over the chunk (positions matter for the causal mask)
        rows = torch.arange(0, 2048, 32)
        import types
        want_rows = []
        # per-row oracle: a query at position P + t is a decode over the first P + t + 1 keys
        for t in rows.tolist():
            one = types.SimpleNamespace(**vars(sub))
            one.q = sub.q[t:t + 1]
            one.seq_lens = torch.tensor([2048 + t + 1])
            one.req_pool_indices = torch.tensor([0])
            want_rows.append(H.exact_decode(one, sub.k_pool, sub.v_pool))
        want = torch.cat(want_rows)
        got = out[sub.token_rows[rows].to(device)].cpu()
        H.assert_close(got, want, H.EXACT_MAX_ABS_TOL, H.EXACT_MEAN_ABS_TOL, what=f"C3 request {req}", rel=True)


PAGED_EXTEND_CASES = [
    # bs, Hq, Hkv, seq_lens, extend_lens, page_size, dtype, window
    (2, 32, 8, [512, 400], [256, 400], 16, torch.bfloat16, -1),      # ragged last page (400 % 16 = 0, 512): full pages
    (3, 8, 2, [1237, 130, 2049], [700, 130, 1000], 16, torch.bfloat16, -1),   # ragged ends, multi-tile prefix
    (2, 6, 2, [333, 777], [200, 500], 32, torch.float16, -1),        # page 32, G = 3 (two q tiles of one head per CTA)
    (2, 8, 1, [1500, 900], [300, 257], 64, torch.bfloat16, -1),      # page 64
    (2, 8, 2, [2000, 1111], [512, 300], 16, torch.bfloat16, 255),    # sliding window on the TMA instantiation
]


@pytest.mark.parametrize("bs,Hq,Hkv,lens,ext,page,dtype,window", PAGED_EXTEND_CASES)
def test_extend_paged_tile_loads_match_row_gathers(bs, Hq, Hkv, lens, ext, page, dtype, window, device):
    """page_size % 16 == 0, 16-bit pool: the tcgen05 kernel fed by TMA boxes (b200attn_extend_paged) is bit-identical
    to the same kernel fed by row gathers, and matches the oracle; pool rows no request position maps to are NaN."""
    case = synth.make_case(bs, Hq, Hkv, 128, lens, extend_lens=ext, seed=977 + page + Hq, page_size=page, dtype=dtype)
    kp, vp = H.pool_after_write(case)
    c = H.to_device(case, device)
    dk, dv = kp.to(device).clone(), vp.to(device).clone()
    used = torch.zeros(dk.shape[0], dtype=torch.bool, device=device)
    for i in range(bs):
        used[c.req_to_token[c.req_pool_indices[i], :lens[i]].long()] = True
    used[:page] = True
    dk[~used] = float("nan")
    dv[~used] = float("nan")
    outs = []
    for ps in (page, 1):
        out = torch.empty_like(c.q)
        lse = torch.empty((case.T, Hq), dtype=torch.float32, device=device)
        ops.extend(c.q, out, dk, dv, c.req_to_token, c.req_pool_indices, c.seq_lens, c.extend_seq_lens,
                   c.extend_prefix_lens, c.extend_start_loc, case.max_extend_len, case.sm_scale,
                   impl=_lib.EXTEND_TCGEN05, lse_out=lse, sliding_window_size=window, page_size=ps)
        torch.cuda.synchronize()
        assert torch.isfinite(out.float()).all() and torch.isfinite(lse).all(), f"page_size={ps}: NaN leaked"
        outs.append((out, lse))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1]), "TMA boxes vs row gathers"
    if window < 0:
        H.assert_close(outs[0][0], H.oracle_extend(case, kp, vp), what="paged extend vs torch_native")


SPLIT_CASES = [
    # bs, Hq, Hkv, seq_lens, extend_lens, kwargs, kv_splits
    (1, 8, 1, [4096 + 300], [300], {}, 4),                                   # chunked prefill over a long prefix, one KV head
    (2, 32, 8, [3000, 700], [700, 129], {}, 2),                               # ragged; the second row has fewer tiles than rows
    (3, 16, 4, [400, 130, 200], [300, 130, 1], {}, 4),                        # more splits than tiles for some items (empty splits)
    (2, 28, 4, [2500, 1025], [1500, 1025], {}, 3),                            # G = 7: two consecutive q tiles per CTA
    (2, 8, 1, [2048 + 512, 640], [512, 128], {"fp8": True, "kv_scale": 0.1135}, 3),
    (2, 16, 2, [16 * 100 + 256, 16 * 3 + 256], [256, 256], {"page_size": 16}, 2),  # TMA page loads
    (1, 8, 2, [1500], [1500], {"dtype": torch.float16}, 8),
]


@pytest.mark.parametrize("bs,Hq,Hkv,lens,ext,kw,S", SPLIT_CASES)
def test_extend_kv_splits_match_unsplit_and_oracle(bs, Hq, Hkv, lens, ext, kw, S, device):
    """b200attn_extend_split: every item's K/V range cut into S shares (separate CTAs), partials merged by LSE.  The
    partials are rounded to the q dtype, so the result is compared with the unsplit tcgen05 kernel and with the oracle
    inside the parity bound (not bit for bit); the LSE must agree too."""
    case = synth.make_case(bs, Hq, Hkv, 128, lens, extend_lens=ext, seed=977 + S + Hq, **kw)
    kp, vp = H.pool_after_write(case)
    want = H.oracle_extend(case, kp, vp)
    base, lse0 = _direct(case, device, kp, vp, _lib.EXTEND_TCGEN05, want_lse=True)
    c = H.to_device(case, device)
    s = case.kv_scale if case.fp8 else 1.0
    ws = torch.full((ops.extend_split_workspace_bytes(case.T, Hq, 128, S),), 0xFF, dtype=torch.uint8, device=device)  # NaN-filled
    out = torch.empty_like(c.q)
    lse = torch.empty((case.T, Hq), dtype=torch.float32, device=device)
    ops.extend(c.q, out, kp.to(device), vp.to(device), c.req_to_token, c.req_pool_indices, c.seq_lens,
               c.extend_seq_lens, c.extend_prefix_lens, c.extend_start_loc, case.max_extend_len, case.sm_scale,
               k_scale=s, v_scale=s, kv_dtype=torch.float8_e4m3fn if case.fp8 else None, lse_out=lse,
               page_size=kw.get("page_size", 1), kv_splits=S, workspace=ws)
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all(), "an empty split's (unwritten) partial leaked into the merge"
    H.assert_close(out, want, what=f"kv_splits={S} vs torch_native")
    H.assert_close(out, base.cpu(), 4e-3, 4e-4, what=f"kv_splits={S} vs the unsplit kernel")
    assert torch.allclose(lse, lse0, atol=2e-3, rtol=1e-4)


def test_extend_split_rejects_bad_arguments(device):
    case = synth.make_case(1, 8, 1, 128, [512], extend_lens=[256], seed=5)
    kp, vp = H.pool_after_write(case)
    c = H.to_device(case, device)
    out = torch.empty_like(c.q)
    args = (c.q, out, kp.to(device), vp.to(device), c.req_to_token, c.req_pool_indices, c.seq_lens, c.extend_seq_lens,
            c.extend_prefix_lens, c.extend_start_loc, case.max_extend_len, case.sm_scale)
    with pytest.raises(ops.B200AttnError):  # no workspace
        ops.extend(*args, kv_splits=2)
    small = torch.empty(16, dtype=torch.uint8, device=device)
    with pytest.raises(ops.B200AttnError):  # workspace too small
        ops.extend(*args, kv_splits=2, workspace=small)
    ws = torch.empty(ops.extend_split_workspace_bytes(case.T, 8, 128, 2), dtype=torch.uint8, device=device)
    with pytest.raises(ops.B200AttnError):  # plain batches only
        ops.extend(*args, kv_splits=2, workspace=ws, sliding_window_size=64)


def test_backend_picks_kv_splits_for_a_long_prefix_chunk(device):
    """Chunked prefill of one request over a long prefix with one KV head per rank: 4 x 2 items for 148 SMs.  The backend
    must choose a split factor from the host-side length lists (no sync) and still match the oracle."""
    case = synth.make_case(1, 8, 1, 128, [6000 + 256], extend_lens=[256], seed=31)
    kp, vp = H.pool_after_write(case)
    r = H.Runner(case, device)
    out = r.forward()
    assert r.backend.forward_metadata.kv_splits > 1
    H.assert_close(out, H.oracle_extend(case, kp, vp), what="backend with KV splits")
    r.backend.extend_kv_splits = 1
    out1 = r.forward()
    assert r.backend.forward_metadata.kv_splits == 1
    H.assert_close(out, out1, 4e-3, 4e-4, what="split vs unsplit through the backend")


def test_backend_longest_request_first_is_bit_identical(device):
    """The backend launches the prefill rows longest request first (permuted per-request arrays; q / out rows are addressed
    through extend_start_loc): same bits as batch order, including a MIXED batch with a decode tail behind the rows."""
    for lens, ext, mode in (([300, 2600, 129, 1400], [200, 600, 129, 700], None),
                            ([260, 1900, 700, 45, 333], [130, 900, 1, 1, 1], "MIXED")):
        case = synth.make_case(len(lens), 16, 4, 128, lens, extend_lens=ext, seed=61)
        kp, vp = H.pool_after_write(case)
        r = H.Runner(case, device)
        r.backend.extend_kv_splits = 1
        fm = getattr(r.itf.ForwardMode, mode) if mode else None
        r.backend.sort_extend_requests = True
        a = r.forward(fm).clone()
        assert r.backend.forward_metadata.ext_order is not None
        r.backend.sort_extend_requests = False
        b = r.forward(fm)
        assert r.backend.forward_metadata.ext_order is None
        assert torch.equal(a, b)
        want = H.oracle_extend(case, kp, vp)
        H.assert_close(a, want, what="sorted launch vs torch_native")
