"""GPU parity of paged split-KV decode (plan + decode + LSE merge), driven through B200AttnBackend ->
the C ABI, against oracle #1 (the reference's torch_native attention) and the fp64 restatement."""
from __future__ import annotations

import pytest
import torch

import golden_util as gu
import helpers as H
from oracle import exact
from sglang_custom_b200 import ops, synth

pytestmark = pytest.mark.gpu


def _run_backend(case, device, **kw):
    r = H.Runner(case, device)
    out = r.forward(**kw)
    return r, out


@pytest.mark.parametrize("name", [n for n in gu.ATTN_FIXTURES if n.startswith("decode")])
def test_decode_golden_fixtures(name, device):
    """Inputs and outputs recorded from the unmodified reference (tests/golden/make_golden.py)."""
    fx = gu.load(name)
    r, out = _run_backend(fx, device)
    kp, vp = r.pools()
    assert H.pools_equal(kp, fx.ref_k_pool_after) and H.pools_equal(vp, fx.ref_v_pool_after)
    H.assert_close(out, fx.ref_out, what=name)


CASES = [
    # bs, Hq, Hkv, D, lens, kwargs
    (8, 32, 8, 64, [512] * 8, {}),                                   # BASELINE config 1 (C1)
    (4, 32, 8, 128, [4096, 1000, 17, 2049], {}),                      # C2-shaped heads, ragged
    (3, 32, 32, 128, [300, 1, 77], {}),                               # MHA (G = 1)
    (2, 40, 8, 128, [1024, 555], {}),                                 # G = 5 (test_decode.py:151)
    (2, 16, 1, 128, [257, 1023], {}),                                 # G = 16
    (5, 8, 1, 128, [8192, 100, 3000, 16, 15], {"fp8": True}),         # C4 at TP 8: 8/1 heads, FP8 KV
    (4, 16, 2, 128, [700, 701, 33, 1], {"fp8": True, "kv_scale": 0.1135}),   # calibrated scale (kv_cache_scales_llama3_8b.json)
    (3, 8, 1, 128, [5000, 48, 333], {"fp8": True, "page_size": 16}),  # C5-style pages
    (3, 32, 8, 128, [640, 64, 1000], {"page_size": 16}),
    (6, 32, 8, 64, [100, 200, 1, 2, 3, 64], {"fp8": True}),
    (2, 32, 8, 128, [100, 300], {"table": "arange"}),                 # the reference tests' easy layout
    (2, 32, 8, 128, [100, 300], {"idx_dtype": torch.int32}),          # CUDA-graph index dtype
    (2, 8, 4, 128, [90, 31], {"dtype": torch.float16}),
    (3, 12, 4, 128, [500, 100, 128], {}),                             # 12/4 heads (test_triton_attention_kernels.py:174)
    (2, 6, 2, 64, [50, 60], {}),                                      # Hkv = 2
]


@pytest.mark.parametrize("bs,Hq,Hkv,D,lens,kw", CASES)
def test_decode_vs_oracles(bs, Hq, Hkv, D, lens, kw, device):
    case = synth.make_case(bs, Hq, Hkv, D, lens, seed=1234 + bs + Hq, **kw)
    kp, vp = H.pool_after_write(case)
    r, out = _run_backend(case, device)
    dk, dv = r.pools()
    assert H.pools_equal(dk, kp) and H.pools_equal(dv, vp), "KV write must be bit-exact"
    H.assert_close(out, H.oracle_decode(case, kp, vp), what="vs torch_native")
    H.assert_close(out, H.exact_decode(case, kp, vp), H.EXACT_MAX_ABS_TOL, H.EXACT_MEAN_ABS_TOL, what="vs fp64 restatement", rel=True)


def test_decode_lse_and_logit_cap(device):
    case = synth.make_case(3, 16, 4, 128, [300, 5000, 31], seed=77)
    c = H.to_device(case, device)
    kp, vp = H.pool_after_write(case)
    dk, dv = kp.to(device), vp.to(device)
    info_ctas = 148
    plan = torch.zeros(ops.decode_plan_bytes(3, info_ctas) // 4, dtype=torch.int32, device=device)
    ws = torch.empty(ops.decode_workspace_bytes(info_ctas, 16, 128) // 4, dtype=torch.float32, device=device)
    for cap in (0.0, 30.0):
        ops.decode_plan(c.seq_lens, info_ctas, plan)
        out = torch.empty_like(c.q)
        lse = torch.empty((3, 16), dtype=torch.float32, device=device)
        ops.decode(c.q, out, dk, dv, c.req_to_token, c.req_pool_indices, c.seq_lens, plan, info_ctas, ws,
                   case.sm_scale, logit_cap=cap, lse_out=lse)
        want, want_lse = H.exact_decode(case, kp, vp, logit_cap=cap, return_lse=True)
        H.assert_close(out, want, H.EXACT_MAX_ABS_TOL, H.EXACT_MEAN_ABS_TOL, what=f"cap={cap}", rel=True)
        assert torch.allclose(lse.cpu(), want_lse, atol=2e-3, rtol=1e-4), f"lse cap={cap}"


@pytest.mark.parametrize("n_ctas", [1, 7, 148, 296, 1000])
def test_decode_is_split_invariant(n_ctas, device):
    """However the flattened KV range is cut across persistent CTAs, the merged result is the same attention."""
    case = synth.make_case(5, 8, 2, 128, [1, 2000, 333, 16, 4097], seed=5)
    c = H.to_device(case, device)
    kp, vp = H.pool_after_write(case)
    dk, dv = kp.to(device), vp.to(device)
    plan = torch.zeros(ops.decode_plan_bytes(5, n_ctas) // 4, dtype=torch.int32, device=device)
    ws = torch.empty(ops.decode_workspace_bytes(n_ctas, 8, 128) // 4, dtype=torch.float32, device=device)
    ops.decode_plan(c.seq_lens, n_ctas, plan, min_units_per_cta=1)
    out = torch.empty_like(c.q)
    ops.decode(c.q, out, dk, dv, c.req_to_token, c.req_pool_indices, c.seq_lens, plan, n_ctas, ws, case.sm_scale)
    H.assert_close(out, H.exact_decode(case, kp, vp), H.EXACT_MAX_ABS_TOL, H.EXACT_MEAN_ABS_TOL, what=f"n_ctas={n_ctas}", rel=True)


def test_decode_plan_contents(device):
    lens = torch.tensor([1, 16, 17, 0, 4096, 33], dtype=torch.int64, device=device)
    n_ctas = 10
    plan = torch.zeros(ops.decode_plan_bytes(6, n_ctas) // 4, dtype=torch.int32, device=device)
    ops.decode_plan(lens, n_ctas, plan, min_units_per_cta=1)
    p = plan.cpu()
    units = [1, 1, 2, 0, 256, 3]
    total = sum(units)
    W = -(-total // n_ctas)
    assert p[0] == total and p[1] == W and p[2] == 6 and p[3] == n_ctas
    cum = [0]
    for u in units:
        cum.append(cum[-1] + u)
    assert p[8:8 + 7].tolist() == cum
    first = p[8 + 7:8 + 7 + n_ctas].tolist()
    for cta in range(n_ctas):
        u = cta * W
        if u < total:
            i = first[cta]
            assert cum[i] <= u and any(cum[j + 1] > u for j in range(i, 6))


def test_decode_padded_cuda_graph_rows_and_empty_batch(device):
    """Padded rows arrive as seq_len = 1, req_pool_idx = 0, loc = 0 (cuda_graph_runner.py:647-655)."""
    case = synth.make_case(4, 8, 2, 128, [50, 1, 1, 1], seed=3, max_reqs=6)
    if int(case.req_pool_indices[0]) == 0:  # keep the live request off table row 0 (the padded rows' row)
        unused = [r for r in range(6) if r not in case.req_pool_indices.tolist()][0]
        case.req_to_token[unused] = case.req_to_token[0]
        case.req_pool_indices[0] = unused
    case.req_pool_indices[1:] = 0
    case.out_cache_loc[1:] = 0
    case.req_to_token[0] = 0
    r, out = _run_backend(case, device)
    # the live row is right; padded rows attend over the dummy slot only (whichever padded writer won) -> finite
    want = H.oracle_decode(case)
    H.assert_close(out[:1], want[:1], what="live row")
    assert torch.isfinite(out.float()).all()
    # bs == 0
    c0 = synth.make_case(0, 8, 2, 128, [], seed=1, max_reqs=2)
    r0 = H.Runner(c0, device)
    assert r0.forward().shape[0] == 0


def test_decode_under_cuda_graph_capture_and_replay(device):
    """The CUDA-graph hooks (cuda_graph_runner.py:251-256,550-558,686-695): fixed int32 buffers, plan refreshed
    in place before every replay, no allocation or host sync inside the captured region."""
    from sglang_custom_b200 import interfaces as itf
    case_a = synth.make_case(4, 32, 8, 128, [300, 64, 2000, 17], seed=21, max_reqs=8, idx_dtype=torch.int32, pool_slack=2000)
    r = H.Runner(case_a, device)
    be = r.backend
    be.init_cuda_graph_state(8, 8)
    bs = 6  # captured batch size; live rows 4, padded 2
    rpi = torch.zeros(8, dtype=torch.int32, device=device)
    sl = torch.full((8,), be.get_cuda_graph_seq_len_fill_value(), dtype=torch.int32, device=device)
    loc = torch.zeros(8, dtype=torch.int64, device=device)
    q = torch.zeros((8, 32 * 128), dtype=torch.bfloat16, device=device)
    k = torch.zeros((8, 8 * 128), dtype=torch.bfloat16, device=device)
    v = torch.zeros((8, 8 * 128), dtype=torch.bfloat16, device=device)

    def fill(case):
        n = case.bs
        rpi.zero_(); sl.fill_(1); loc.zero_()
        rpi[:n] = case.req_pool_indices.to(device); sl[:n] = case.seq_lens.to(device)
        loc[:n] = case.out_cache_loc.to(device)
        q[:n] = case.q.reshape(n, -1).to(device); k[:n] = case.k_new.reshape(n, -1).to(device)
        v[:n] = case.v_new.reshape(n, -1).to(device)

    fb = itf.ForwardBatch(forward_mode=itf.ForwardMode.DECODE, batch_size=bs, req_pool_indices=rpi[:bs],
                          seq_lens=sl[:bs], out_cache_loc=loc[:bs], seq_lens_sum=0,
                          req_to_token_pool=r.mr.req_to_token_pool, token_to_kv_pool=r.mr.token_to_kv_pool,
                          attn_backend=be)
    fill(case_a)
    be.init_forward_metadata_capture_cuda_graph(bs, bs, rpi, sl, None, itf.ForwardMode.DECODE, None)
    stream = torch.cuda.Stream(device)
    stream.wait_stream(torch.cuda.current_stream(device))
    with torch.cuda.stream(stream):
        for _ in range(2):
            r.layer.forward(q[:bs], k[:bs], v[:bs], fb)
    torch.cuda.current_stream(device).wait_stream(stream)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=stream):
        out = r.layer.forward(q[:bs], k[:bs], v[:bs], fb)
    # replay 1: same batch
    be.init_forward_metadata_replay_cuda_graph(bs, rpi, sl, 0, None, itf.ForwardMode.DECODE, None, None)
    graph.replay()
    torch.cuda.synchronize()
    H.assert_close(out[:4].view(4, 32, 128), H.oracle_decode(case_a), what="replay A")
    # replay 2: different lengths and tables, 3 live rows
    case_b = synth.make_case(3, 32, 8, 128, [2000, 5, 128], seed=22, max_reqs=8, idx_dtype=torch.int32)
    Lb = case_b.req_to_token.shape[1]
    assert Lb <= r.mr.req_to_token_pool.req_to_token.shape[1] and case_b.P <= r.mr.token_to_kv_pool.k_buffer[0].shape[0]
    r.mr.req_to_token_pool.req_to_token.zero_()
    r.mr.req_to_token_pool.req_to_token[:, :Lb].copy_(case_b.req_to_token.to(device))
    r.mr.token_to_kv_pool.k_buffer[0][:case_b.P].copy_(case_b.k_pool.to(device))
    r.mr.token_to_kv_pool.v_buffer[0][:case_b.P].copy_(case_b.v_pool.to(device))
    fill(case_b)
    be.init_forward_metadata_replay_cuda_graph(bs, rpi, sl, 0, None, itf.ForwardMode.DECODE, None, None)
    graph.replay()
    torch.cuda.synchronize()
    H.assert_close(out[:3].view(3, 32, 128), H.oracle_decode(case_b), what="replay B")
    assert torch.isfinite(out.float()).all()


def test_decode_full_size_c2_sampled_oracle_and_relocation_invariance(device):
    """BASELINE config 2 at full size (bs 256, ctx 4096, 32/8, D 128, bf16, P = 1M + 1 rows).
    (a) a sample of requests is checked against both CPU oracles on the same bytes;
    (b) size-independent property: moving every KV row to a different physical slot (and rewriting the
        page table accordingly) leaves the output bit-identical."""
    bs, L = 256, 4096
    case = synth.make_case_device(bs, 32, 8, 128, [L] * bs, device, seed=1236)
    r = H.Runner(case, device)
    out = r.forward()
    dk, dv = r.pools()
    sample = [0, 17, 101, 255]
    sub = synth.extract_subcase(case, sample, dk, dv)
    got = out[sample].cpu()
    H.assert_close(got, H.oracle_decode(sub, sub.k_pool, sub.v_pool), what="C2 sample vs torch_native")
    H.assert_close(got, H.exact_decode(sub, sub.k_pool, sub.v_pool), H.EXACT_MAX_ABS_TOL, H.EXACT_MEAN_ABS_TOL,
                   what="C2 sample vs fp64", rel=True)
    # (b) relocate: new_pool[perm[s]] = pool[s]; table entries s -> perm[s]
    P = dk.shape[0]
    perm = torch.randperm(P - 1, device=device) + 1
    perm = torch.cat([torch.zeros(1, dtype=perm.dtype, device=device), perm])
    nk, nv = torch.empty_like(dk), torch.empty_like(dv)
    nk[perm] = dk
    nv[perm] = dv
    table2 = perm[r.mr.req_to_token_pool.req_to_token.long()].to(torch.int32)
    c = case
    plan, ws, be = r.backend.plan, r.backend.workspace, r.backend
    out2 = torch.empty_like(out)
    ops.decode_plan(c.seq_lens, be.n_ctas, plan, be.min_units_per_cta)
    ops.decode(c.q, out2, nk, nv, table2, c.req_pool_indices, c.seq_lens, plan, be.n_ctas, ws, c.sm_scale)
    torch.cuda.synchronize()
    assert torch.equal(out2, out)


PAGED_CASES = [
    # bs, Hq, Hkv, lens, fp8, page_size
    (3, 32, 8, [640, 64, 1000], False, 16),        # bf16, 8 KV heads: one 32 KiB box per page and K/V
    (4, 8, 1, [5000, 48, 333, 17], True, 16),      # C5 per-rank heads: FP8, 128-byte head rows
    (3, 64, 8, [1500, 31, 2049], True, 16),        # C5 at TP 1 (token-major consumer, 16 warps)
    (3, 16, 2, [100, 1, 4097], False, 32),         # page 32: two units per page
    (2, 16, 4, [777, 4000], True, 64),             # page 64 (test_flashattn_backend.py page_size 64)
    (2, 16, 1, [300, 129], False, 16),             # G = 16 (head-major consumer, both MMA halves)
]


@pytest.mark.parametrize("bs,Hq,Hkv,lens,fp8,page", PAGED_CASES)
def test_decode_paged_tile_loads_match_row_gathers(bs, Hq, Hkv, lens, fp8, page, device):
    """page_size % 16 == 0: the TMA box loads (b200attn_decode_paged) and the row gathers read the same bytes in the
    same order, so the two outputs are bit-identical -- and both match the oracle.  The rows of every request's last
    page past its end are poisoned with NaN patterns: whole-page loads must not let them reach the output."""
    case = synth.make_case(bs, Hq, Hkv, 128, lens, seed=4321 + page + Hq, fp8=fp8, page_size=page,
                           kv_scale=0.1135 if fp8 else 1.0)
    kp, vp = H.pool_after_write(case)
    want = H.oracle_decode(case, kp, vp)
    c = H.to_device(case, device)
    dk, dv = kp.to(device).clone(), vp.to(device).clone()
    # poison: slots of allocated pages that no request position maps to
    used = torch.zeros(dk.shape[0], dtype=torch.bool, device=device)
    for i in range(bs):
        used[c.req_to_token[c.req_pool_indices[i], :lens[i]].long()] = True
    used[:page] = True
    if fp8:
        dk.view(torch.uint8)[~used] = 0x7F
        dv.view(torch.uint8)[~used] = 0x7F
    else:
        dk[~used] = float("nan")
        dv[~used] = float("nan")
    n_ctas = 148
    plan = torch.zeros(ops.decode_plan_bytes(bs, n_ctas) // 4, dtype=torch.int32, device=device)
    ws = torch.empty(ops.decode_workspace_bytes(n_ctas, Hq, 128) // 4, dtype=torch.float32, device=device)
    outs = []
    for ps, min_units in ((page, 4), (1, 4), (page, 1)):
        ops.decode_plan(c.seq_lens, n_ctas, plan, min_units_per_cta=min_units)
        out = torch.empty_like(c.q)
        lse = torch.empty((bs, Hq), dtype=torch.float32, device=device)
        ops.decode(c.q, out, dk, dv, c.req_to_token, c.req_pool_indices, c.seq_lens, plan, n_ctas, ws, case.sm_scale,
                   k_scale=case.kv_scale, v_scale=case.kv_scale, kv_dtype=torch.float8_e4m3fn if fp8 else None,
                   lse_out=lse, page_size=ps)
        torch.cuda.synchronize()
        assert torch.isfinite(out.float()).all() and torch.isfinite(lse).all(), f"page_size={ps}: NaN leaked from a page tail"
        H.assert_close(out, want, what=f"page_size={ps} vs torch_native")
        outs.append((out, lse))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1]), "TMA boxes vs row gathers"


def test_decode_paged_rejects_bad_arguments(device):
    case = synth.make_case(1, 8, 1, 128, [64], seed=3, page_size=16)
    c = H.to_device(case, device)
    plan = torch.zeros(ops.decode_plan_bytes(1, 4) // 4, dtype=torch.int32, device=device)
    ws = torch.empty(ops.decode_workspace_bytes(4, 8, 128) // 4, dtype=torch.float32, device=device)
    ops.decode_plan(c.seq_lens, 4, plan)
    with pytest.raises(ops.B200AttnError):
        from sglang_custom_b200 import _lib
        _lib.check(_lib.load().b200attn_decode_paged(
            c.q.data_ptr(), c.q.stride(0), c.q.data_ptr(), c.q.stride(0), c.k_pool.data_ptr(), c.v_pool.data_ptr(),
            c.k_pool.stride(0), c.req_to_token.data_ptr(), c.req_to_token.stride(0), c.req_pool_indices.data_ptr(),
            c.seq_lens.data_ptr(), 1, 1, 8, 1, 128, 0, 0, 0.1, 0.0, 1.0, 1.0, plan.data_ptr(), 4, ws.data_ptr(),
            ws.numel() * 4, None, 16, 7, None), "b200attn_decode_paged")


def test_decode_paged_large_fp8_batch_two_tile_groups(device):
    """Regression (round 2): the 16-warp FP8 consumer runs two tile groups over the stage ring.  With an odd ring a
    group met a stage on every second phase only -- the same parity every time -- and bulk-tensor copies complete out
    of order, so a `full` barrier one fill behind read as complete: stale stages at best, a faulted launch at worst
    (bs 64 x 8000 tokens, 64/8 heads, page 16; small cases never showed it).  Boxes and row gathers must agree bit for
    bit at a size where the ring wraps hundreds of times per CTA."""
    bs, Hq, Hkv, L = 64, 64, 8, 8000
    case = synth.make_case_device(bs, Hq, Hkv, 128, [L] * bs, device, page_size=16, fp8=True, kv_scale=0.1135, seed=77)
    n_ctas = 148
    plan = torch.zeros(ops.decode_plan_bytes(bs, n_ctas) // 4, dtype=torch.int32, device=device)
    ws = torch.empty(ops.decode_workspace_bytes(n_ctas, Hq, 128) // 4, dtype=torch.float32, device=device)
    ops.decode_plan(case.seq_lens, n_ctas, plan)
    outs = []
    for ps in (1, 16, 16):
        out = torch.empty_like(case.q)
        ops.decode(case.q, out, case.k_pool, case.v_pool, case.req_to_token, case.req_pool_indices, case.seq_lens, plan,
                   n_ctas, ws, case.sm_scale, k_scale=case.kv_scale, v_scale=case.kv_scale,
                   kv_dtype=torch.float8_e4m3fn, page_size=ps)
        torch.cuda.synchronize()
        outs.append(out)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[1], outs[2])
