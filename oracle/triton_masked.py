"""CPU restatement of the reference's sliding-window and custom-mask attention semantics.

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Only tests/ may import this; the product never does.

The reference's two CPU attention paths (torch_native, sgl-kernel CPU) know neither a sliding window nor a
speculative-decoding mask; those predicates exist only in its Triton backend.  This file restates them as one
explicit visibility matrix per request followed by an fp32 softmax -- no tiling, no online rescaling -- and is
pinned against the unmodified Triton kernels executed by Triton's CPU interpreter (oracle/ref_triton.py; fixtures
tests/golden/masked_*.pt made by tests/golden/make_golden_masked.py).

Reference lines restated:
  window range   update_sliding_window_buffer, layers/attention/triton_backend.py:927-955
                 window_kv_lens = min(lens, W + 1); the LAST window_kv_lens positions are kept.  `lens` is seq_lens in
                 decode (:170-201) and extend_prefix_lens in extend (:299-309).
  window mask    _fwd_kernel stage 1, triton_ops/extend_attention.py:168-174: q_id <= kv_id + W with q_id counted from
                 the request's first NEW token and kv_id from the first key of the window range; only when W > 0.
                 Stage 2 (new tokens, :196-262) has no window term.
  custom mask    :158-167 (prefix, unless SKIP_PREFIX_CUSTOM_MASK, which the backend always leaves True: :322) and
                 :236-247 (new tokens; replaces the causal test, restricted to the keys stage 2 visits).
  decode         forward_decode just swaps kv_indices for window_kv_indices (triton_backend.py:711-716).

Two behaviours of the reference kernel that are NOT reproduced (documented in DESIGN.md):
  * a query whose stage-1 keys are all masked in the first key block turns NaN there (exp(-inf - -inf),
    extend_attention.py:178-180).  Here such a row attends to whatever it can see; a row that sees nothing gives 0.
  * with a custom mask that is not a subset of the causal mask, which "future" keys are honoured depends on BLOCK_M
    (stage 2 stops at the end of the query block when IS_CAUSAL, :191-195).  Here mask AND causal is used; EAGLE tree
    masks are causal subsets, where both agree.
"""
from __future__ import annotations

import math
from typing import Optional

import torch


def window_range(length: int, sliding_window_size: int) -> tuple:
    """(first kept position, number kept) of `length` positions (triton_backend.py:936-944)."""
    if sliding_window_size is None or sliding_window_size < 0:
        return 0, length
    n = min(length, sliding_window_size + 1)
    return length - n, n


def window_kv_indices(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, lens: torch.Tensor,
                      sliding_window_size: int):
    """update_sliding_window_buffer restated: (window_kv_indptr int32 [bs+1], window_kv_indices int32)."""
    bs = lens.numel()
    indptr = torch.zeros(bs + 1, dtype=torch.int32)
    pieces = []
    for i in range(bs):
        start, n = window_range(int(lens[i]), sliding_window_size)
        indptr[i + 1] = indptr[i] + n
        pieces.append(req_to_token[int(req_pool_indices[i]), start:start + n].to(torch.int32))
    return indptr, (torch.cat(pieces) if pieces else torch.zeros(0, dtype=torch.int32))


def visibility(P: int, E: int, causal: bool, sliding_window_size: int, mask_block: Optional[torch.Tensor],
               skip_prefix_custom_mask: bool) -> torch.Tensor:
    """bool [E, P+E]: key j (absolute position in the request) visible to new-token query t."""
    L = P + E
    t = torch.arange(E)[:, None]
    j = torch.arange(L)[None, :]
    is_prefix = j < P
    vis = torch.ones((E, L), dtype=torch.bool)
    if causal:
        vis &= j <= P + t
    W = -1 if sliding_window_size is None else sliding_window_size
    if W >= 0:
        start, _ = window_range(P, W)
        vis &= ~is_prefix | (j >= start)
        if W > 0:
            vis &= ~is_prefix | (t <= (j - start) + W)
    if mask_block is not None:
        m = mask_block.reshape(E, L).to(torch.bool)
        use = ~is_prefix if skip_prefix_custom_mask else torch.ones_like(is_prefix)
        vis &= ~use | m
    return vis


def _cap(s: torch.Tensor, logit_cap: float) -> torch.Tensor:
    return logit_cap * torch.tanh(s / logit_cap) if logit_cap and logit_cap > 0 else s


def extend(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, req_to_token: torch.Tensor,
           req_pool_indices: torch.Tensor, seq_lens: torch.Tensor, extend_prefix_lens: torch.Tensor,
           extend_seq_lens: torch.Tensor, scaling: float, causal: bool = True, sliding_window_size: int = -1,
           custom_mask: Optional[torch.Tensor] = None, mask_indptr: Optional[torch.Tensor] = None,
           skip_prefix_custom_mask: bool = True, logit_cap: float = 0.0) -> torch.Tensor:
    """q [T,Hq,D]; pools [P,Hkv,D] already holding the new tokens (KV write first, triton_backend.py:647-650).
    Returns [T,Hq,Dv] in q.dtype; softmax and both products in fp32."""
    T, Hq, D = q.shape
    Hkv = k_cache.shape[1]
    G = Hq // Hkv
    out = torch.zeros((T, Hq, v_cache.shape[2]), dtype=q.dtype)
    start = 0
    for i in range(seq_lens.numel()):
        P, E = int(extend_prefix_lens[i]), int(extend_seq_lens[i])
        L = P + E
        assert L == int(seq_lens[i])
        if E == 0:
            continue
        idx = req_to_token[int(req_pool_indices[i]), :L].long()
        K = k_cache[idx].float()  # [L,Hkv,D]
        V = v_cache[idx].float()
        blk = None
        if custom_mask is not None:
            m0 = int(mask_indptr[i])
            blk = custom_mask[m0:m0 + E * L]
        vis = visibility(P, E, causal, sliding_window_size, blk, skip_prefix_custom_mask)
        qi = q[start:start + E].float()  # [E,Hq,D]
        for h in range(Hq):
            s = _cap(qi[:, h] @ K[:, h // G].T * scaling, logit_cap)
            s = s.masked_fill(~vis, float("-inf"))
            any_vis = vis.any(dim=1, keepdim=True)
            p = torch.softmax(torch.where(any_vis, s, torch.zeros_like(s)), dim=-1)
            p = torch.where(any_vis, p, torch.zeros_like(p))
            out[start:start + E, h] = (p @ V[:, h // G]).to(q.dtype)
        start += E
    return out


def decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, req_to_token: torch.Tensor,
           req_pool_indices: torch.Tensor, seq_lens: torch.Tensor, scaling: float, sliding_window_size: int = -1,
           logit_cap: float = 0.0) -> torch.Tensor:
    """q [bs,Hq,D]; request i attends to its last min(L_i, W+1) tokens."""
    bs, Hq, D = q.shape
    Hkv = k_cache.shape[1]
    G = Hq // Hkv
    out = torch.zeros((bs, Hq, v_cache.shape[2]), dtype=q.dtype)
    for i in range(bs):
        first, n = window_range(int(seq_lens[i]), sliding_window_size)
        if n == 0:
            continue
        idx = req_to_token[int(req_pool_indices[i]), first:first + n].long()
        K, V = k_cache[idx].float(), v_cache[idx].float()
        for h in range(Hq):
            s = _cap(q[i, h].float() @ K[:, h // G].T * scaling, logit_cap)
            out[i, h] = (torch.softmax(s, dim=-1) @ V[:, h // G]).to(q.dtype)
    return out


def tree_mask(P: int, E: int, gen: torch.Generator) -> torch.Tensor:
    """A random EAGLE-style verify mask [E, P+E] (bool): every draft token sees the whole prefix, itself and the
    ancestors of a random tree rooted at draft token 0 (speculative/eagle_utils.py builds the same shape:
    seq_mask_len = num_draft_tokens * (seq_len + num_draft_tokens), triton_backend.py:254-256)."""
    m = torch.zeros((E, P + E), dtype=torch.bool)
    m[:, :P] = True
    parent = [-1] + [int(torch.randint(0, t, (1,), generator=gen)) for t in range(1, E)]
    for t in range(E):
        a = t
        while a >= 0:
            m[t, P + a] = True
            a = parent[a]
    return m


def flatten_masks(blocks) -> tuple:
    """-> (custom_mask bool [sum E_i*L_i], mask_indptr int64 [bs+1]) as TritonAttnBackend lays them out (:253-259)."""
    indptr = torch.zeros(len(blocks) + 1, dtype=torch.int64)
    for i, b in enumerate(blocks):
        indptr[i + 1] = indptr[i] + b.numel()
    flat = torch.cat([b.reshape(-1) for b in blocks]) if blocks else torch.zeros(0, dtype=torch.bool)
    return flat, indptr


__all__ = ["window_range", "window_kv_indices", "visibility", "extend", "decode", "tree_mask", "flatten_masks", "math"]
