"""Host logic of the speculative-decoding draft modes (CPU): the branch table of B200MultiStepDraftBackend against the
reference's generate_draft_decode_kv_indices (golden fixture made by tests/golden/make_golden_draft.py from the
unmodified kernel, speculative/eagle_utils.py:821-899), and the DRAFT_EXTEND length algebra."""
from __future__ import annotations

from pathlib import Path

import pytest
import torch

from sglang_custom_b200.draft_backend import build_branch_table, draft_token_start

GOLDEN = Path(__file__).resolve().parent / "golden" / "draft_kv_indices.pt"


@pytest.mark.parametrize("idx", range(5))
def test_branch_table_equals_reference_kv_indices(idx):
    fx = torch.load(GOLDEN)[idx]
    sl, topk, steps, page = fx["seq_lens"], fx["topk"], fx["num_steps"], fx["page_size"]
    table = build_branch_table(fx["req_to_token"], fx["req_pool_indices"], sl, topk, steps, page)
    bs = sl.numel()
    assert table.shape == (bs * topk, fx["req_to_token"].shape[1])
    for i in range(steps):
        indptr, indices = fx["ref_kv_indptr"][i], fx["ref_kv_indices"][i]
        for b in range(bs):
            for t in range(topk):
                r = b * topk + t
                n = int(sl[b]) + i + 1
                lo, hi = int(indptr[r]), int(indptr[r + 1])
                assert hi - lo == n, (i, b, t, lo, hi, n)
                assert torch.equal(indices[lo:hi], table[r, :n]), (i, b, t)


def test_branch_table_into_fixed_buffer_and_topk1():
    fx = torch.load(GOLDEN)[1]
    sl, topk, steps, page = fx["seq_lens"], fx["topk"], fx["num_steps"], fx["page_size"]
    want = build_branch_table(fx["req_to_token"], fx["req_pool_indices"], sl, topk, steps, page)
    buf = torch.full((sl.numel() * topk + 5, fx["req_to_token"].shape[1]), -7, dtype=torch.int32)
    got = build_branch_table(fx["req_to_token"], fx["req_pool_indices"], sl, topk, steps, page, out=buf)
    assert got.data_ptr() == buf.data_ptr() and torch.equal(got, want) and (buf[sl.numel() * topk:] == -7).all()
    # topk == 1: the draft tokens already follow the prefix, the table is the request's own row
    fx1 = torch.load(GOLDEN)[0]
    t1 = build_branch_table(fx1["req_to_token"], fx1["req_pool_indices"], fx1["seq_lens"], 1, fx1["num_steps"], 1)
    assert torch.equal(t1, fx1["req_to_token"][fx1["req_pool_indices"]])


def test_draft_token_start_formulas():
    sl = torch.tensor([48, 17, 100])
    assert draft_token_start(sl, 2, 3, 1).tolist() == [[48, 51], [17, 20], [100, 103]]
    # page 16: branch t starts ceil((L % 16 + steps) / 16) pages further (eagle_utils.py:875-883)
    assert draft_token_start(sl, 2, 3, 16).tolist() == [[48, 64], [17, 33], [100, 116]]
    assert draft_token_start(torch.tensor([30]), 3, 5, 16).tolist() == [[30, 62, 94]]


@pytest.mark.parametrize("L,topk,steps,page", [(29, 2, 4, 1), (31, 3, 3, 1), (32, 2, 2, 1), (26, 2, 5, 16), (30, 4, 2, 1)])
def test_branch_table_drops_slots_past_the_row_instead_of_aliasing_them(L, topk, steps, page):
    """seq_len + num_steps (or a branch's start + j) beyond the table width: out-of-range draft slots are masked, as in
    the reference's kernel (eagle_utils.py:889-905 loads with `mask=`), never folded onto the last column."""
    width = 32
    r2t = (torch.arange(3 * width, dtype=torch.int32).reshape(3, width) + 1000)
    rpi = torch.tensor([2, 0])
    sl = torch.tensor([L, 5])
    start = draft_token_start(sl, topk, steps, page)
    want = r2t[rpi].repeat_interleave(topk, 0).clone()
    for b in range(2):
        for t in range(topk):
            for j in range(steps):
                s, d = int(start[b, t]) + j, int(sl[b]) + j
                if s < width and d < width:
                    want[b * topk + t, d] = r2t[rpi[b], s]
    for _ in range(3):  # same answer every time: no dependence on which duplicate scatter_ keeps
        assert torch.equal(build_branch_table(r2t, rpi, sl, topk, steps, page), want)
    buf = torch.zeros((2 * topk, width), dtype=torch.int32)
    assert torch.equal(build_branch_table(r2t, rpi, sl, topk, steps, page, out=buf), want)
