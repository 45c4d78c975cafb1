"""Oracle #1: restatement of the reference's torch_native attention path (CPU, torch SDPA).

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Pinned against the unmodified reference by
tests/test_oracle_pin.py (bit-exact, same torch build).

Reference lines restated here (relative to /root/reference/python/sglang/srt):
  set_kv_buffer   mem_cache/memory_pool.py:420-458
  decode          layers/attention/torch_native_backend.py:112-180 (+ wrapper :226-267)
  extend          layers/attention/torch_native_backend.py:27-110 (+ wrapper :182-224)
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F


def set_kv_buffer(k_buffer: torch.Tensor, v_buffer: torch.Tensor, loc: torch.Tensor, cache_k: torch.Tensor,
                  cache_v: torch.Tensor, pool_dtype: torch.dtype, k_scale=None, v_scale=None) -> None:
    """memory_pool.py:436-458.  `k_buffer` may be the uint8 store view of an fp8 pool.

    Like the reference, the scale division is IN PLACE on the caller's tensors
    (memory_pool.py:437-440), the cast is torch's `.to(pool_dtype)` and the scatter is
    index_put (`buf[loc] = x`, last writer wins on duplicate loc)."""
    if cache_k.dtype != pool_dtype:
        if k_scale is not None:
            cache_k.div_(k_scale)
        if v_scale is not None:
            cache_v.div_(v_scale)
        cache_k = cache_k.to(pool_dtype)
        cache_v = cache_v.to(pool_dtype)
    if k_buffer.dtype != pool_dtype:  # store_dtype != dtype: fp8 kept as uint8
        cache_k = cache_k.view(k_buffer.dtype)
        cache_v = cache_v.view(v_buffer.dtype)
    k_buffer[loc] = cache_k
    v_buffer[loc] = cache_v


def _sdpa(q_hld: torch.Tensor, k_hld: torch.Tensor, v_hld: torch.Tensor, scaling, enable_gqa: bool,
          causal: bool) -> torch.Tensor:
    """[H, L, D] operands -> [H, Lq, Dv]; one batch element, exactly the call the reference makes."""
    return F.scaled_dot_product_attention(q_hld.unsqueeze(0), k_hld.unsqueeze(0), v_hld.unsqueeze(0),
                                          enable_gqa=enable_gqa, scale=scaling, is_causal=causal).squeeze(0)


def decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, req_to_token: torch.Tensor,
           req_pool_indices: torch.Tensor, seq_lens: torch.Tensor, scaling: Optional[float],
           out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """torch_native_backend.py:112-180.  q [bs, Hq, D]; caches [P, Hkv, D]; returns [bs, Hq, Dv].

    One query token per request; all seq_lens[i] cached tokens are visible (causal=False)."""
    bs, Hq, _ = q.shape
    gqa = Hq != k_cache.shape[1]
    if out is None:
        out = torch.empty((bs, Hq, v_cache.shape[2]), dtype=q.dtype)
    q_heads_first = q.movedim(0, 1)  # [Hq, bs, D]
    for i in range(bs):
        n_kv = int(seq_lens[i])
        slots = req_to_token[int(req_pool_indices[i]), :n_kv]
        keys = k_cache[slots].movedim(0, 1)  # the page gather: [Hkv, n_kv, D]
        vals = v_cache[slots].movedim(0, 1)
        o = _sdpa(q_heads_first[:, i:i + 1, :], keys, vals, scaling, gqa, False)
        out[i:i + 1] = o.movedim(1, 0)
    return out


def extend(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, req_to_token: torch.Tensor,
           req_pool_indices: torch.Tensor, seq_lens: torch.Tensor, extend_prefix_lens: torch.Tensor,
           extend_seq_lens: torch.Tensor, scaling: Optional[float], causal: bool = True,
           out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """torch_native_backend.py:27-110.  q [T, Hq, D] (T = sum E_i), returns [T, Hq, Dv].

    The reference pads each request's queries to seq_len rows (the first prefix_len rows are
    uninitialised memory whose outputs are thrown away, :80-88) so that SDPA's `is_causal`
    top-left mask lands on the right keys; the pad rows are zeros here, which changes nothing
    for the kept rows (SDPA rows are independent)."""
    T, Hq, D = q.shape
    gqa = Hq != k_cache.shape[1]
    if out is None:
        out = torch.empty((T, Hq, v_cache.shape[2]), dtype=q.dtype)
    q_heads_first = q.movedim(0, 1)  # [Hq, T, D]
    row = 0
    for i in range(seq_lens.shape[0]):
        n_new, n_prefix, n_kv = int(extend_seq_lens[i]), int(extend_prefix_lens[i]), int(seq_lens[i])
        padded = torch.zeros((Hq, n_kv, D), dtype=q.dtype)
        padded[:, n_prefix:, :] = q_heads_first[:, row:row + n_new, :]
        slots = req_to_token[int(req_pool_indices[i]), :n_kv]
        keys = k_cache[slots].movedim(0, 1)
        vals = v_cache[slots].movedim(0, 1)
        o = _sdpa(padded, keys, vals, scaling, gqa, causal).movedim(1, 0)  # [n_kv, Hq, Dv]
        out[row:row + n_new] = o[n_prefix:]
        row += n_new
    return out


def dequantised_pool(pool_u8: torch.Tensor, scale: float, dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """FP8 oracle input: pool.view(e4m3).to(bf16) * scale."""
    x = pool_u8.view(torch.float8_e4m3fn).to(dtype)
    return x if scale == 1.0 else x * scale


def merge_state(a: torch.Tensor, lse_a: torch.Tensor, b: torch.Tensor, lse_b: torch.Tensor):
    """Pure-torch statement of the 2-way merge (sgl-kernel/tests/test_merge_state_v2.py:100-133,
    merge_attn_states.cu:32-106): +inf LSE counts as -inf; fp32 math; output in a's dtype."""
    la = torch.where(lse_a == float("inf"), torch.full_like(lse_a, float("-inf")), lse_a).float()
    lb = torch.where(lse_b == float("inf"), torch.full_like(lse_b, float("-inf")), lse_b).float()
    m = torch.maximum(la, lb)
    ea, eb = torch.exp(la - m), torch.exp(lb - m)
    s = ea + eb
    out = a.float() * (ea / s).unsqueeze(-1) + b.float() * (eb / s).unsqueeze(-1)
    return out.to(a.dtype), torch.log(s) + m
