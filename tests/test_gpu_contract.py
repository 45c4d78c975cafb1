"""GPU tests of the backend's contract edges (round-2 review items): save_kv_cache=False, padded rows behind a MIXED
tail, in-place LSE merge, duplicate KV-write slots."""
from __future__ import annotations

import types

import pytest
import torch

import helpers as H
from sglang_custom_b200 import ops, synth
from sglang_custom_b200._lib import B200AttnError

pytestmark = pytest.mark.gpu


def test_save_kv_cache_false_with_kv_given_needs_the_fused_write_marker(device):
    """The backend reads every key from the pool (like torch_native_backend.py:196-199); the Triton backend attends
    over the passed k/v instead (extend_attention_fwd K_Extend/V_Extend).  forward_*(k, v, save_kv_cache=False) is
    therefore only legal right after rotary.forward_fused_kv wrote this layer's K/V of this batch; otherwise raise."""
    case = synth.make_case(2, 8, 2, 128, [300, 200], extend_lens=[100, 200], seed=11)
    r = H.Runner(case, device)
    fb = r.batch()
    r.backend.init_forward_metadata(fb)
    q = case.q.reshape(case.T, -1).to(device)
    k = case.k_new.reshape(case.T, -1).to(device)
    v = case.v_new.reshape(case.T, -1).to(device)
    with pytest.raises(B200AttnError, match="save_kv_cache=False"):
        r.layer.forward(q, k, v, fb, save_kv_cache=False)
    # k=None: "the pool already holds them" (radix_attention.py:91-99) -- allowed
    r.layer.forward(q, None, None, fb, save_kv_cache=False)
    # after the marker (what rotary.forward_fused_kv sets) the call is legal, for this layer and batch only
    r.backend.note_kv_written(r.layer, fb)
    r.layer.forward(q, k, v, fb, save_kv_cache=False)
    fb2 = r.batch()
    r.backend.init_forward_metadata(fb2)
    with pytest.raises(B200AttnError):
        r.layer.forward(q, k, v, fb2, save_kv_cache=False)
    # decode path: same rule
    dcase = synth.make_case(3, 8, 2, 128, [50, 70, 90], seed=12)
    rd = H.Runner(dcase, device)
    fbd = rd.batch()
    rd.backend.init_forward_metadata(fbd)
    qd = dcase.q.reshape(3, -1).to(device)
    kd = dcase.k_new.reshape(3, -1).to(device)
    with pytest.raises(B200AttnError):
        rd.layer.forward(qd, kd, kd, fbd, save_kv_cache=False)


def test_mixed_tail_rows_are_found_with_padded_q_rows_behind_them(device):
    """MIXED batch whose q carries padded rows behind the real tokens (DP attention / num_token_non_padded): the
    decode tail starts at sum(extend lens of the prefill rows), not at q.shape[0] - n_tail."""
    case = synth.make_case(4, 8, 2, 128, [260, 90, 33, 400], extend_lens=[130, 1, 1, 1], seed=13)
    kp, vp = H.pool_after_write(case)
    want = H.oracle_extend(case, kp, vp)
    r = H.Runner(case, device)
    fb = r.batch(r.itf.ForwardMode.MIXED)
    r.backend.init_forward_metadata(fb)
    assert r.backend.forward_metadata.n_decode_tail == 3
    assert r.backend.forward_metadata.n_extend_tokens == 130
    pad = 5
    T = case.T

    def padded(x):
        x = x.reshape(T, -1).to(device)
        return torch.cat([x, torch.zeros((pad, x.shape[1]), dtype=x.dtype, device=device)])
    q = padded(case.q)
    # KV write through the pool stand-in first (out_cache_loc has T entries), then attention with k=None
    pool = r.mr.token_to_kv_pool
    pool.set_kv_buffer(r.layer, case.out_cache_loc.to(device), case.k_new.to(device), case.v_new.to(device))
    out = r.layer.forward(q, None, None, fb, save_kv_cache=False)
    torch.cuda.synchronize()
    H.assert_close(out[:T].view(T, case.Hq, case.D), want, what="mixed batch with padded q rows")


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_merge_state_in_place_equals_out_of_place(dtype, device):
    """ops.merge_state(a, ..., out=a) -- what the TARGET_VERIFY route does -- is bit-identical to a fresh output."""
    g = torch.Generator().manual_seed(3)
    a = torch.randn((777, 8, 128), generator=g).to(dtype).to(device)
    b = torch.randn((777, 8, 128), generator=g).to(dtype).to(device)
    la = (torch.randn((777, 8), generator=g) * 4).to(device)
    lb = (torch.randn((777, 8), generator=g) * 4).to(device)
    o, l = ops.merge_state(a, la, b, lb)
    a2 = a.clone()
    o2, l2 = ops.merge_state(a2, la, b, lb, out=a2)
    assert o2.data_ptr() == a2.data_ptr()
    assert torch.equal(o, o2) and torch.equal(l, l2)
    b3 = b.clone()
    o3, _ = ops.merge_state(a, la, b3, lb, out=b3)
    assert torch.equal(o, o3)


def test_check_loc_finds_duplicates_outside_the_dummy_page(device):
    """Duplicates in out_cache_loc are last-writer-wins in the reference (memory_pool.py:457-458: buf[loc] = x); the
    CUDA write leaves the order unspecified, so any duplicate outside the dummy page [0, page_size) is reported."""
    loc = torch.tensor([5, 9, 0, 0, 0, 17], dtype=torch.int64, device=device)
    assert ops.check_loc(loc, 1, 100) == (0, 0)          # padded rows all hit the dummy slot: legal
    loc = torch.tensor([5, 9, 0, 9, 0, 17, 9], dtype=torch.int64, device=device)
    assert ops.check_loc(loc, 1, 100) == (3, 0)          # slot 9 three times = 3 colliding pairs
    loc = torch.tensor([5, 9, 3, 3, 100, -1], dtype=torch.int64, device=device)
    assert ops.check_loc(loc, 16, 100) == (0, 2)         # 3 is inside a 16-row dummy page; 100 and -1 out of range
    assert ops.check_loc(torch.zeros(0, dtype=torch.int64, device=device), 1, 10) == (0, 0)
    big = torch.randperm(20000, device=device)[:8192].to(torch.int64) + 1
    assert ops.check_loc(big, 1, 20001) == (0, 0)
    big[4000] = big[77]
    assert ops.check_loc(big, 1, 20001) == (1, 0)


def test_backend_rejects_write_races_when_check_loc_is_on(device):
    case = synth.make_case(3, 8, 2, 128, [50, 70, 90], seed=14)
    r = H.Runner(case, device)
    r.backend.check_loc = True
    r.forward()  # a well-formed decode batch passes
    fb = r.batch()
    fb.out_cache_loc = fb.out_cache_loc.clone()
    fb.out_cache_loc[2] = fb.out_cache_loc[0]  # two requests writing the same live slot
    with pytest.raises(B200AttnError, match="duplicate"):
        r.forward(fb=fb)


def test_cuda_graph_padding_only_duplicates_the_dummy_slot(device):
    """The only duplicates SGLang produces are padded CUDA-graph rows: seq_len = fill value 1, req_pool_index 0,
    out_cache_loc 0 (cuda_graph_runner.py:686-695) -- all inside the dummy page."""
    case = synth.make_case(4, 8, 2, 128, [50, 70, 90, 30], seed=15, max_reqs=8)
    r = H.Runner(case, device)
    be = r.backend
    fill = be.get_cuda_graph_seq_len_fill_value()
    bs_pad = 8
    loc = torch.zeros(bs_pad, dtype=torch.int64, device=device)
    loc[:4] = case.out_cache_loc.to(device)
    sl = torch.full((bs_pad,), fill, dtype=torch.int64, device=device)
    sl[:4] = case.seq_lens.to(device)
    assert ops.check_loc(loc, case.page_size, case.P) == (0, 0)
    assert int((loc[4:] < case.page_size).all())


def test_idle_batch_with_a_spec_info_launches_nothing(device):
    """EAGLE with DP attention hands idle ranks an IDLE / bs == 0 batch that still carries an (idle) EagleDraftInput or
    EagleVerifyInput.  The reference reads spec_info.kv_indptr and launches nothing; so must this backend, in the eager
    path and in both CUDA-graph hooks, instead of asking for draft-step rows that an idle input does not have."""
    case = synth.make_case(2, 8, 2, 128, [40, 50], seed=21)
    r = H.Runner(case, device)
    itf = r.itf
    empty = torch.zeros(0, dtype=torch.int64, device=device)
    for mode, bs in ((itf.ForwardMode.IDLE, 0), (itf.ForwardMode.DECODE, 0), (itf.ForwardMode.IDLE, 2)):
        fb = r.batch(itf.ForwardMode.DECODE)
        fb.forward_mode, fb.batch_size = mode, bs
        if bs == 0:
            fb.req_pool_indices, fb.seq_lens = empty, empty
        fb.spec_info = types.SimpleNamespace()  # no b200_draft_step: an idle input
        r.backend.init_forward_metadata(fb)
        md = r.backend.forward_metadata
        assert md.is_decode and md.bs == 0
        r.backend.init_forward_metadata_capture_cuda_graph(bs, bs, fb.req_pool_indices, fb.seq_lens, None, mode,
                                                           fb.spec_info)
        assert r.backend.forward_metadata.bs == 0
        r.backend.init_forward_metadata_replay_cuda_graph(bs, fb.req_pool_indices, fb.seq_lens, 0, None, mode,
                                                          fb.spec_info, None)
    # a non-empty decode batch with a spec_info but no draft-step rows is still an error
    fb = r.batch(itf.ForwardMode.DECODE)
    fb.spec_info = types.SimpleNamespace()
    with pytest.raises(B200AttnError, match="draft-step rows"):
        r.backend.init_forward_metadata(fb)


def test_fused_write_marker_is_tied_to_the_batch_object_not_its_id(device):
    """The marker of rotary.forward_fused_kv must not be satisfiable by a LATER batch that CPython gave the id of a
    collected one: it is stamped on the batch object itself."""
    case = synth.make_case(2, 8, 2, 128, [300, 200], extend_lens=[100, 200], seed=13)
    r = H.Runner(case, device)
    q = case.q.reshape(case.T, -1).to(device)
    k = case.k_new.reshape(case.T, -1).to(device)
    v = case.v_new.reshape(case.T, -1).to(device)
    fb = r.batch()
    r.backend.init_forward_metadata(fb)
    r.backend.note_kv_written(r.layer, fb)
    r.layer.forward(q, k, v, fb, save_kv_cache=False)
    # a stand-in for "same id, different object": copy the stamp-free attributes into a new batch
    fb2 = r.batch()
    assert getattr(fb2, "b200_kv_written", None) is None
    r.backend.init_forward_metadata(fb2)
    with pytest.raises(B200AttnError, match="save_kv_cache=False"):
        r.layer.forward(q, k, v, fb2, save_kv_cache=False)
    # a stale stamp (an earlier write, another layer's sequence number) does not pass either
    r.backend.note_kv_written(r.layer, fb2)
    r.backend.note_kv_written(r.layer, fb)
    with pytest.raises(B200AttnError, match="save_kv_cache=False"):
        r.layer.forward(q, k, v, fb2, save_kv_cache=False)
    r.layer.forward(q, k, v, fb, save_kv_cache=False)
