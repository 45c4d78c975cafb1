"""Pin oracle/triton_masked.py (sliding window + speculative-decoding custom mask) against the reference.

The reference's CPU backends have neither predicate; its Triton kernels do (extend_attention.py:158-175,236-247,
triton_backend.py:927-955).  tests/golden/masked_*.pt are outputs of those unmodified kernels, executed by Triton's
CPU interpreter in the build container (tests/golden/make_golden_masked.py).  Here:
  * the restatement reproduces every fixture (fp32: 2e-5; fp16: 2e-3 -- the reference rounds P to fp16 before P.V),
  * window_kv_indices is bit-exact against update_sliding_window_buffer's output stored in the fixtures,
  * where /root/reference is mounted, a fresh seeded case is run through the live interpreter in a subprocess,
  * properties the reference's own test states: a mask equal to the causal mask == no mask
    (test/srt/test_triton_attention_kernels.py:121-172), window >= L == no window.
"""
from __future__ import annotations

import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

import golden_util as gu
import helpers as H
from oracle import ref_loader
from oracle import torch_native as tn
from oracle import triton_masked as tm
from sglang_custom_b200 import synth

ROOT = Path(__file__).resolve().parents[1]


def _tol(dtype):
    return (2e-5, 2e-6) if dtype == torch.float32 else (2e-3, 2e-4)


def _oracle_extend(fx):
    return tm.extend(fx.q, fx.ref_k_pool_after, fx.ref_v_pool_after, fx.req_to_token, fx.req_pool_indices, fx.seq_lens,
                     fx.extend_prefix_lens, fx.extend_seq_lens, fx.sm_scale, True, fx.window,
                     getattr(fx, "custom_mask", None), getattr(fx, "mask_indptr", None), fx.skip_prefix_custom_mask)


@pytest.mark.parametrize("name", gu.MASKED_EXTEND_FIXTURES)
def test_masked_extend_restatement_matches_triton_golden(name):
    fx = gu.load(name)
    out = _oracle_extend(fx)
    mx, mean = _tol(fx.dtype)
    H.assert_close(out, fx.ref_out, max_abs=mx, mean_abs=mean, what=name)


@pytest.mark.parametrize("name", gu.MASKED_DECODE_FIXTURES)
def test_window_decode_restatement_matches_triton_golden(name):
    fx = gu.load(name)
    out = tm.decode(fx.q, fx.ref_k_pool_after, fx.ref_v_pool_after, fx.req_to_token, fx.req_pool_indices, fx.seq_lens,
                    fx.sm_scale, fx.window)
    mx, mean = _tol(fx.dtype)
    H.assert_close(out, fx.ref_out, max_abs=mx, mean_abs=mean, what=name)
    indptr, idx = tm.window_kv_indices(fx.req_to_token, fx.req_pool_indices, fx.seq_lens, fx.window)
    assert torch.equal(indptr, fx.ref_window_kv_indptr)
    assert torch.equal(idx, fx.ref_window_kv_indices[: idx.numel()])


def test_masks_in_fixtures_are_causal_subsets():
    """The domain the CUDA kernel promises reference-identical results on (include/b200attn.h)."""
    for name in gu.MASKED_EXTEND_FIXTURES:
        fx = gu.load(name)
        if not hasattr(fx, "custom_mask"):
            continue
        for i in range(fx.bs):
            P, E = int(fx.extend_prefix_lens[i]), int(fx.extend_seq_lens[i])
            m = fx.custom_mask[int(fx.mask_indptr[i]):int(fx.mask_indptr[i + 1])].reshape(E, P + E)
            assert not bool(torch.triu(m[:, P:], diagonal=1).any()), name
            assert bool(m[:, P:].diagonal().all()), name


@pytest.mark.parametrize("seed", [11, 12])
def test_causal_mask_equals_no_mask_and_torch_native(seed):
    case = synth.make_case(3, 8, 2, 64, [50, 21, 9], extend_lens=[13, 21, 4], seed=seed, dtype=torch.float32)
    kp, vp = H.pool_after_write(case)
    blocks = []
    for i in range(case.bs):
        P, E = int(case.extend_prefix_lens[i]), int(case.extend_seq_lens[i])
        blocks.append(torch.cat([torch.ones(E, P, dtype=torch.bool), torch.tril(torch.ones(E, E)).bool()], 1))
    mask, indptr = tm.flatten_masks(blocks)
    args = (case.q, kp, vp, case.req_to_token, case.req_pool_indices, case.seq_lens, case.extend_prefix_lens,
            case.extend_seq_lens, case.sm_scale)
    plain = tm.extend(*args)
    masked = tm.extend(*args, custom_mask=mask, mask_indptr=indptr)
    assert torch.equal(plain, masked)
    native = tn.extend(*args, causal=True)
    H.assert_close(plain, native, max_abs=2e-5, mean_abs=2e-6, what="restatement vs torch_native")
    wide = tm.extend(*args, sliding_window_size=10_000)
    assert torch.equal(plain, wide)


def test_window_decode_equals_truncated_decode():
    case = synth.make_case(3, 8, 2, 64, [70, 5, 33], seed=21, dtype=torch.float32)
    kp, vp = H.pool_after_write(case)
    W = 15
    got = tm.decode(case.q, kp, vp, case.req_to_token, case.req_pool_indices, case.seq_lens, case.sm_scale, W)
    # torch_native decode over a page table whose rows were shifted so that only the window remains
    table = torch.zeros_like(case.req_to_token)
    lens = case.seq_lens.clone()
    for i in range(case.bs):
        first, n = tm.window_range(int(case.seq_lens[i]), W)
        r = int(case.req_pool_indices[i])
        table[r, :n] = case.req_to_token[r, first:first + n]
        lens[i] = n
    want = tn.decode(case.q, kp, vp, table, case.req_pool_indices, lens, case.sm_scale)
    H.assert_close(got, want, max_abs=2e-5, mean_abs=2e-6, what="window decode")


_LIVE = r"""
import sys
sys.path.insert(0, {root!r}); sys.path.insert(0, {root!r} + "/tests")
from oracle import ref_triton
import torch
import helpers as H
from oracle import triton_masked as tm
from sglang_custom_b200 import synth
g = torch.Generator().manual_seed(5)
case = synth.make_case(2, 4, 2, 64, [61, 30], extend_lens=[9, 9], seed=77, dtype=torch.float32)
kp, vp = H.pool_after_write(case)
blocks = [tm.tree_mask(int(case.extend_prefix_lens[i]), 9, g) for i in range(2)]
mask, indptr = tm.flatten_masks(blocks)
for window, m, ip in ((-1, mask, indptr), (20, None, None)):
    ref = ref_triton.ref_extend(case.q, case.k_new, case.v_new, kp, vp, case.req_to_token, case.req_pool_indices,
                                case.extend_prefix_lens, case.extend_seq_lens, case.sm_scale, True, window, m, ip)
    got = tm.extend(case.q, kp, vp, case.req_to_token, case.req_pool_indices, case.seq_lens, case.extend_prefix_lens,
                    case.extend_seq_lens, case.sm_scale, True, window, m, ip)
    assert (ref - got).abs().max() < 2e-5, (window, float((ref - got).abs().max()))
dcase = synth.make_case(3, 4, 2, 64, [80, 3, 40], seed=78, dtype=torch.float32)
kp, vp = H.pool_after_write(dcase)
ref = ref_triton.ref_decode(dcase.q, kp, vp, dcase.req_to_token, dcase.req_pool_indices, dcase.seq_lens, dcase.sm_scale, 17)
got = tm.decode(dcase.q, kp, vp, dcase.req_to_token, dcase.req_pool_indices, dcase.seq_lens, dcase.sm_scale, 17)
assert (ref - got).abs().max() < 2e-5
print("LIVE_OK")
"""


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference is only mounted in the build container")
def test_live_triton_interpreter_equals_restatement():
    env = dict(os.environ, TRITON_INTERPRET="1")
    r = subprocess.run([sys.executable, "-c", _LIVE.format(root=str(ROOT))], capture_output=True, text=True, env=env,
                       timeout=600)
    assert "LIVE_OK" in r.stdout, r.stderr[-2000:]


def _triton_draft_extend(fx):
    """What the reference's TRITON backend computes for a DRAFT_EXTEND batch (triton_backend.py:265-281): all
    seq_lens[i] cached slots -- the accepted tokens included -- as an unmasked prefix, then the accepted tokens once
    more under the causal mask.  Restated in fp32 so the committed reference output can be recognised."""
    out = torch.zeros_like(fx.q)
    G = fx.q.shape[1] // fx.ref_k_pool_after.shape[1]
    start = 0
    for i in range(fx.seq_lens.numel()):
        L, E = int(fx.seq_lens[i]), int(fx.accept_length[i])
        idx = fx.req_to_token[int(fx.req_pool_indices[i]), :L].long()
        K = torch.cat([fx.ref_k_pool_after[idx], fx.k_new[start:start + E]]).float()
        V = torch.cat([fx.ref_v_pool_after[idx], fx.v_new[start:start + E]]).float()
        vis = torch.cat([torch.ones(E, L, dtype=torch.bool), torch.ones(E, E, dtype=torch.bool).tril()], dim=1)
        for h in range(fx.q.shape[1]):
            s = (fx.q[start:start + E, h].float() @ K[:, h // G].T * fx.sm_scale).masked_fill(~vis, float("-inf"))
            out[start:start + E, h] = (torch.softmax(s, -1) @ V[:, h // G]).to(fx.q.dtype)
        start += E
    return out


@pytest.mark.parametrize("name", gu.DRAFT_EXTEND_FIXTURES)
def test_draft_extend_deviation_is_pinned_against_both_reference_backends(name):
    """DRAFT_EXTEND: the reference's two backends disagree (tests/golden/make_golden_draft_extend.py).  The oracle the
    CUDA path is held to (torch_native extend with prefix = seq_lens - accept_length) is the FlashInfer-style causal
    result; the Triton-style result counts the accepted tokens twice and is a different number.  Both are outputs of
    the unmodified reference kernel; this test keeps the difference on record."""
    fx = gu.load(name)
    mx, mean = _tol(fx.dtype)
    prefix = fx.seq_lens - fx.accept_length
    causal = tn.extend(fx.q, fx.ref_k_pool_after, fx.ref_v_pool_after, fx.req_to_token, fx.req_pool_indices,
                       fx.seq_lens, prefix, fx.accept_length, fx.sm_scale)
    H.assert_close(causal, fx.ref_out_causal, max_abs=mx, mean_abs=mean, what=f"{name}: causal")
    H.assert_close(_triton_draft_extend(fx), fx.ref_out_triton, max_abs=mx, mean_abs=mean, what=f"{name}: triton")
    gap = (fx.ref_out_causal.float() - fx.ref_out_triton.float()).abs().max()
    assert float(gap) > 10 * H.MAX_ABS_TOL / 2, "the two reference results must be distinguishable at the parity bound"
