"""GPU parity of the speculative-decoding draft modes through B200AttnBackend / B200MultiStepDraftBackend -> C ABI:
DRAFT_EXTEND (eagle_utils.py:107-164, triton_backend.py:265-281,476-499,609-625) and the multi-step draft decode
(triton_backend.py:735-873, eagle_utils.py:821-899), against the torch_native oracle on explicit key lists."""
from __future__ import annotations

import types

import pytest
import torch

import golden_util as gu
import helpers as H
from sglang_custom_b200 import interfaces as itf
from sglang_custom_b200 import synth
from sglang_custom_b200.draft_backend import B200MultiStepDraftBackend, build_branch_table, draft_token_start

pytestmark = pytest.mark.gpu


def _draft_extend_batch(r, accept):
    c, dev = r.case, r.device
    return itf.ForwardBatch(
        forward_mode=itf.ForwardMode.DRAFT_EXTEND, batch_size=c.bs, req_pool_indices=c.req_pool_indices.to(dev),
        seq_lens=c.seq_lens.to(dev), out_cache_loc=c.out_cache_loc.to(dev), seq_lens_sum=c.seq_lens_sum,
        req_to_token_pool=r.mr.req_to_token_pool, token_to_kv_pool=r.mr.token_to_kv_pool, attn_backend=r.backend,
        seq_lens_cpu=c.seq_lens.cpu(),
        spec_info=types.SimpleNamespace(accept_length=torch.tensor(accept, dtype=torch.int32, device=dev)))


@pytest.mark.parametrize("kw", [dict(), dict(fp8=True, kv_scale=0.1135), dict(page_size=16)])
def test_draft_extend_is_causal_extend_over_the_accepted_tokens(kw, device):
    """seq_lens include the accept_length accepted tokens; query t of request i sees keys <= seq_lens - accept + t."""
    lens, accept = [300, 77, 1030, 6], [4, 1, 6, 6]
    case = synth.make_case(4, 16, 4, 128, lens, extend_lens=accept, seed=701, **kw)
    kp, vp = H.pool_after_write(case)
    r = H.Runner(case, device)
    r.mr.server_args.speculative_num_steps = 5
    r.backend.speculative_num_steps = 5
    out = r.forward(fb=_draft_extend_batch(r, accept))
    dk, dv = r.pools()
    assert H.pools_equal(dk, kp) and H.pools_equal(dv, vp)
    H.assert_close(out, H.oracle_extend(case, kp, vp), what="draft extend vs torch_native")
    # without speculative_num_steps the grid bound is max(accept_length) (the reference's own sync)
    r.backend.speculative_num_steps = None
    out2 = r.forward(fb=_draft_extend_batch(r, accept))
    assert torch.equal(out, out2)


@pytest.mark.parametrize("name", [n for n in gu.DRAFT_EXTEND_FIXTURES if "f32" not in n])
def test_draft_extend_matches_the_reference_causal_result_not_the_triton_one(name, device):
    """Outputs of the unmodified reference kernel for the same DRAFT_EXTEND batch, both ways
    (tests/golden/make_golden_draft_extend.py): the backend reproduces the FlashInfer-style causal result and is
    measurably away from the Triton backend's (which counts the accepted tokens twice) -- DESIGN.md 3.10."""
    fx = gu.load(name)
    r = H.Runner(fx, device)
    out = r.forward(fb=_draft_extend_batch(r, fx.accept_length.tolist()))
    dk, dv = r.pools()
    assert H.pools_equal(dk, fx.ref_k_pool_after) and H.pools_equal(dv, fx.ref_v_pool_after)
    H.assert_close(out, fx.ref_out_causal, what=f"{name}: vs reference causal (FlashInfer semantics)")
    gap = (out.float().cpu() - fx.ref_out_triton.float()).abs().max()
    assert float(gap) > 5 * H.MAX_ABS_TOL, f"{name}: unexpectedly equal to the Triton backend's double-counted result"


def test_draft_extend_cuda_graph_hooks(device):
    """Captured with speculative_num_steps + 1 rows per request, replayed with other accept lengths
    (triton_backend.py:476-499, 609-625): the lens buffers are refilled in place."""
    steps = 3
    lens_a, acc_a = [64, 200, 33], [4, 4, 4]
    lens_b, acc_b = [60, 190, 30], [2, 4, 1]   # fits inside case A's pool and table
    case_a = synth.make_case(3, 8, 2, 128, lens_a, extend_lens=acc_a, seed=711)
    case_b = synth.make_case(3, 8, 2, 128, lens_b, extend_lens=acc_b, seed=712)
    assert case_b.P <= case_a.P and case_b.req_to_token.shape[1] <= case_a.req_to_token.shape[1]
    r = H.Runner(case_a, device)
    r.mr.server_args.speculative_num_steps = steps
    be = r.backend
    be.speculative_num_steps = steps
    be.init_cuda_graph_state(3, 3 * (steps + 1))
    dev = device
    rpi = case_a.req_pool_indices.to(dev).clone()
    sl = case_a.seq_lens.to(dev).clone()
    acc = torch.tensor(acc_a, dtype=torch.int32, device=dev)
    spec = types.SimpleNamespace(accept_length=acc)
    T = 3 * (steps + 1)
    q = torch.zeros((T, 8 * 128), dtype=case_a.dtype, device=dev)
    out_static = torch.zeros_like(q)
    fb = _draft_extend_batch(r, acc_a)
    fb.req_pool_indices, fb.seq_lens, fb.spec_info = rpi, sl, spec
    mode = itf.ForwardMode.DRAFT_EXTEND

    def write_pool(case):
        kp, vp = H.pool_after_write(case)
        pool = r.mr.token_to_kv_pool
        pool.k_buffer[0][:kp.shape[0]].copy_(kp.to(dev))
        pool.v_buffer[0][:vp.shape[0]].copy_(vp.to(dev))
        r.mr.req_to_token_pool.req_to_token[:, :case.req_to_token.shape[1]].copy_(case.req_to_token.to(dev))
        return kp, vp

    kp, vp = write_pool(case_a)
    q[:case_a.T] = case_a.q.reshape(case_a.T, -1).to(dev)
    be.init_forward_metadata_capture_cuda_graph(3, T, rpi, sl, None, mode, spec)
    side = torch.cuda.Stream(dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        r.layer.forward(q, None, None, fb, save_kv_cache=False)
    torch.cuda.current_stream(dev).wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        out_static.copy_(r.layer.forward(q, None, None, fb, save_kv_cache=False))
    for case, accs in ((case_a, acc_a), (case_b, acc_b)):
        kp, vp = write_pool(case)
        rpi.copy_(case.req_pool_indices.to(dev))
        sl.copy_(case.seq_lens.to(dev))
        acc.copy_(torch.tensor(accs, dtype=torch.int32, device=dev))
        q.zero_()
        q[:case.T] = case.q.reshape(case.T, -1).to(dev)
        be.init_forward_metadata_replay_cuda_graph(3, rpi, sl, -1, None, mode, spec, None)
        g.replay()
        torch.cuda.synchronize(dev)
        H.assert_close(out_static[:case.T].view(case.T, 8, 128), H.oracle_extend(case, kp, vp), what=f"replay {accs}")


def _multistep_case(bs, Hq, Hkv, lens, topk, steps, page, fp8, seed):
    """A decode-shaped case whose requests have room for topk * steps draft slots behind their prefix."""
    total = [int(draft_token_start(torch.tensor([L]), topk, steps, page)[0, -1]) + steps for L in lens]
    case = synth.make_case(bs, Hq, Hkv, 128, total, seed=seed, page_size=page, fp8=fp8, kv_scale=0.1135 if fp8 else 1.0)
    return case


@pytest.mark.parametrize("topk,steps,page,fp8", [(1, 3, 1, False), (4, 3, 1, False), (2, 4, 16, True), (1, 5, 16, False)])
def test_multistep_draft_backend_matches_oracle_on_explicit_key_lists(topk, steps, page, fp8, device):
    lens = [100, 33, 257]
    bs, Hq, Hkv = len(lens), 16, 4
    case = _multistep_case(bs, Hq, Hkv, lens, topk, steps, page, fp8, seed=900 + topk + steps)
    kp, vp = case.k_pool, case.v_pool   # nothing is written: the draft tokens' K/V are already in the pool
    r = H.Runner(case, device)
    r.mr.server_args.page_size = page
    msb = B200MultiStepDraftBackend(r.mr, topk, steps)
    dev = device
    sl = torch.tensor(lens, dtype=torch.int64, device=dev)
    rpi = case.req_pool_indices.to(dev)
    g = torch.Generator().manual_seed(5)
    n = bs * topk
    q = torch.randn((n, Hq, 128), generator=g).to(case.dtype)
    fb = itf.ForwardBatch(
        forward_mode=itf.ForwardMode.DECODE, batch_size=bs, req_pool_indices=rpi, seq_lens=sl,
        out_cache_loc=torch.zeros(n, dtype=torch.int64, device=dev), seq_lens_sum=int(sl.sum()),
        req_to_token_pool=r.mr.req_to_token_pool, token_to_kv_pool=r.mr.token_to_kv_pool, attn_backend=None,
        seq_lens_cpu=sl.cpu(), spec_info=types.SimpleNamespace())
    for graph in (False, True):
        if graph:
            msb.init_cuda_graph_state(bs, n)
            msb.init_forward_metadata_capture_cuda_graph(fb)
            msb.init_forward_metadata_replay_cuda_graph(fb, bs)
        else:
            msb.init_forward_metadata(fb)
        table = build_branch_table(case.req_to_token, case.req_pool_indices, torch.tensor(lens), topk, steps, page)
        for i in range(steps):
            fb.attn_backend = msb.attn_backends[i]
            out = r.layer.forward(q.reshape(n, -1).to(dev), None, None, fb, save_kv_cache=False)
            torch.cuda.synchronize(dev)
            # oracle: a decode batch of bs * topk rows over the branch table
            sub = types.SimpleNamespace(**vars(case))
            sub.bs, sub.T = n, n
            sub.q = q
            sub.req_to_token = table
            sub.req_pool_indices = torch.arange(n)
            sub.seq_lens = (torch.tensor(lens).repeat_interleave(topk) + i + 1)
            sub.decode = True
            want = H.oracle_decode(sub, kp, vp)
            H.assert_close(out.view(n, Hq, 128), want, what=f"step {i} topk {topk} graph {graph}")
