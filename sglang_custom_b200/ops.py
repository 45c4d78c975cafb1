"""Tensor-level wrappers over the C ABI: torch supplies device memory and streams only.

Every function validates dtypes/devices, passes raw pointers + strides to
`libb200attn.so`, and raises `B200AttnError` on a non-zero return code.
"""
from __future__ import annotations

from typing import Optional, Union

import torch

from . import _lib
from ._lib import B200AttnError, check

_DT = {torch.bfloat16: _lib.BF16, torch.float16: _lib.F16, torch.float32: _lib.F32,
       torch.float8_e4m3fn: _lib.FP8_E4M3}


def _dt(t: torch.dtype) -> int:
    if t not in _DT:
        raise B200AttnError(f"unsupported dtype {t}")
    return _DT[t]


def _ptr(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()


def _cuda(*ts: Optional[torch.Tensor]) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise B200AttnError("b200attn operates on CUDA tensors only (no CPU fallback exists)")


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _is_i64(t: torch.Tensor) -> int:
    if t.dtype == torch.int64:
        return 1
    if t.dtype == torch.int32:
        return 0
    raise B200AttnError(f"index tensor must be int32 or int64, got {t.dtype}")


def _pool_dtype(pool: torch.Tensor, logical: Optional[torch.dtype]) -> int:
    """MHATokenToKVPool stores fp8 as uint8 (`store_dtype`, memory_pool.py:146-150)."""
    if pool.dtype == torch.uint8:
        if logical not in (None, torch.float8_e4m3fn):
            raise B200AttnError(f"uint8 pool with logical dtype {logical} (only float8_e4m3fn is built)")
        return _lib.FP8_E4M3
    return _dt(pool.dtype)


def _last_dim_contig(t: torch.Tensor, name: str) -> None:
    if t.stride(-1) != 1:
        raise B200AttnError(f"{name}: last dimension must be contiguous")


ScaleT = Union[None, float, torch.Tensor]


def _scale_args(s: ScaleT):
    """-> (host float, device tensor or None, n). 0.0 host scale means 'no scaling'."""
    if s is None:
        return 0.0, None, 1
    if isinstance(s, torch.Tensor):
        if s.is_cuda:
            t = s.detach().to(torch.float32).reshape(-1).contiguous()
            return 0.0, t, t.numel()
        return float(s), None, 1
    return float(s), None, 1


def kv_write(k: torch.Tensor, v: torch.Tensor, loc: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor,
             k_scale: ScaleT = None, v_scale: ScaleT = None, pool_dtype: Optional[torch.dtype] = None,
             scale_mode: int = _lib.SCALE_DIV, writeback_scaled: bool = False) -> None:
    """pool[loc] = cast(k), cast(v): MHATokenToKVPool.set_kv_buffer (memory_pool.py:420-458)."""
    _cuda(k, v, loc, k_pool, v_pool)
    if k.dim() != 3 or v.shape != k.shape:
        raise B200AttnError(f"kv_write: k/v must be [T,Hkv,D] of equal shape, got {tuple(k.shape)} {tuple(v.shape)}")
    T, H, D = k.shape
    if loc.dtype != torch.int64 or loc.numel() != T or not loc.is_contiguous():
        raise B200AttnError("kv_write: loc must be a contiguous int64 [T] tensor")
    _last_dim_contig(k, "k"); _last_dim_contig(v, "v")
    if k_pool.dim() != 3 or k_pool.shape[1:] != (H, D) or v_pool.shape != k_pool.shape:
        raise B200AttnError("kv_write: pool must be [P,Hkv,D]")
    if k_pool.stride(2) != 1 or k_pool.stride(1) != D:
        raise B200AttnError("kv_write: pool head/dim strides must be (D, 1)")
    ks_h, ks_d, n1 = _scale_args(k_scale)
    vs_h, vs_d, n2 = _scale_args(v_scale)
    keep = (ks_d, vs_d)  # keep temporaries alive until the launch is enqueued
    check(_lib.load().b200attn_kv_write(
        k.data_ptr(), v.data_ptr(), k.stride(0), k.stride(1), v.stride(0), v.stride(1), loc.data_ptr(),
        T, H, D, k_pool.data_ptr(), v_pool.data_ptr(), k_pool.stride(0), _dt(k.dtype),
        _pool_dtype(k_pool, pool_dtype), ks_h, vs_h, _ptr(ks_d), _ptr(vs_d), max(n1, n2),
        int(scale_mode), int(bool(writeback_scaled)), _stream(k)), "b200attn_kv_write")
    del keep


def set_kv_buffer(k_buffer: torch.Tensor, v_buffer: torch.Tensor, loc: torch.Tensor, cache_k: torch.Tensor,
                  cache_v: torch.Tensor, pool_dtype: torch.dtype, k_scale: ScaleT = None, v_scale: ScaleT = None,
                  scale_mode: int = _lib.SCALE_DIV, writeback_scaled: bool = False) -> None:
    """MHATokenToKVPool.set_kv_buffer (memory_pool.py:420-458) on raw layer buffers.

    `k_buffer` is the stored buffer (uint8 for an fp8 pool); `pool_dtype` the logical dtype.
    As in the reference, scales only apply when a cast happens (cache_k.dtype != pool dtype).
    scale_mode DIV is torch's CPU `div_` (the oracle); RCP is what torch CUDA does for a python
    float.  writeback_scaled=True reproduces the reference's in-place `cache_k.div_(k_scale)`."""
    if loc.dtype != torch.int64:
        loc = loc.to(torch.int64)
    if cache_k.dtype == pool_dtype:
        k_scale = v_scale = None
    kv_write(cache_k, cache_v, loc.contiguous(), k_buffer, v_buffer, k_scale, v_scale, pool_dtype, scale_mode,
             writeback_scaled)


def rope_kv_write(positions: torch.Tensor, q: torch.Tensor, k: torch.Tensor, cos_sin_cache: torch.Tensor,
                  is_neox: bool, v: Optional[torch.Tensor] = None, loc: Optional[torch.Tensor] = None,
                  k_pool: Optional[torch.Tensor] = None, v_pool: Optional[torch.Tensor] = None,
                  k_scale: ScaleT = None, v_scale: ScaleT = None, pool_dtype: Optional[torch.dtype] = None,
                  scale_mode: int = _lib.SCALE_DIV, math_mode: int = _lib.ROPE_FP32) -> None:
    """In place: q, k = rotary_emb(positions, q, k) (layers/rotary_embedding.py:138-166 / :218-244) and, when `loc`
    is given, set_kv_buffer(loc, k, v) (memory_pool.py:420-458) in the same launch.
    q [T,Hq,D], k/v [T,Hkv,D] (views with head stride D); cos_sin_cache [max_pos, rotary_dim] fp32 or q.dtype."""
    _cuda(positions, q, k, v, cos_sin_cache, loc, k_pool, v_pool)
    if q.dim() != 3 or k.dim() != 3 or q.shape[0] != k.shape[0] or q.shape[2] != k.shape[2] or k.dtype != q.dtype:
        raise B200AttnError(f"rope_kv_write: q [T,Hq,D] / k [T,Hkv,D] expected, got {tuple(q.shape)} {tuple(k.shape)}")
    T, Hq, D = q.shape
    Hkv = k.shape[1]
    for name, t in (("q", q), ("k", k)) + ((("v", v),) if v is not None else ()):
        if t.stride(2) != 1 or t.stride(1) != D:
            raise B200AttnError(f"rope_kv_write: {name} must have strides (*, D, 1)")
    if positions.numel() != T or not positions.is_contiguous():
        raise B200AttnError("rope_kv_write: positions must be a contiguous [T] tensor")
    if cos_sin_cache.dim() != 2 or not cos_sin_cache.is_contiguous():
        raise B200AttnError("rope_kv_write: cos_sin_cache must be contiguous [max_position, rotary_dim]")
    ks_h = vs_h = 0.0
    ks_d = vs_d = None
    n = 1
    pool_dt = _dt(q.dtype)
    if loc is not None:
        if v is None or k_pool is None or v_pool is None or v.shape != k.shape:
            raise B200AttnError("rope_kv_write: v, k_pool and v_pool are required with loc")
        if loc.dtype != torch.int64 or loc.numel() != T or not loc.is_contiguous():
            raise B200AttnError("rope_kv_write: loc must be a contiguous int64 [T] tensor")
        if k_pool.dim() != 3 or k_pool.shape[1:] != (Hkv, D) or v_pool.shape != k_pool.shape or \
                k_pool.stride(2) != 1 or k_pool.stride(1) != D:
            raise B200AttnError("rope_kv_write: pool must be [P,Hkv,D] with strides (*, D, 1)")
        pool_dt = _pool_dtype(k_pool, pool_dtype)
        if pool_dt != _dt(q.dtype):
            ks_h, ks_d, n1 = _scale_args(k_scale)
            vs_h, vs_d, n2 = _scale_args(v_scale)
            n = max(n1, n2)
    check(_lib.load().b200attn_rope_kv_write(
        q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0), _ptr(v), v.stride(0) if v is not None else 0,
        positions.data_ptr(), _is_i64(positions), cos_sin_cache.data_ptr(), _dt(cos_sin_cache.dtype),
        cos_sin_cache.shape[1], int(bool(is_neox)), T, Hq, Hkv, D, _dt(q.dtype), _ptr(loc), _ptr(k_pool),
        _ptr(v_pool), k_pool.stride(0) if k_pool is not None else 0, pool_dt, ks_h, vs_h, _ptr(ks_d), _ptr(vs_d), n,
        int(scale_mode), int(math_mode), _stream(q)), "b200attn_rope_kv_write")


def decode_plan_bytes(max_bs: int, n_ctas: int) -> int:
    return _lib.load().b200attn_decode_plan_bytes(int(max_bs), int(n_ctas))


def decode_workspace_bytes(n_ctas: int, Hq: int, Dv: int) -> int:
    return _lib.load().b200attn_decode_workspace_bytes(int(n_ctas), int(Hq), int(Dv))


def decode_plan(seq_lens: torch.Tensor, n_ctas: int, plan: torch.Tensor, min_units_per_cta: int = 4,
                sliding_window_size: int = -1) -> None:
    """sliding_window_size > -1: plan over the last min(L, window + 1) tokens of every request
    (update_sliding_window_buffer, triton_backend.py:927-955); decode() then reads the window from the plan."""
    _cuda(seq_lens, plan)
    if plan.dtype != torch.int32 or not plan.is_contiguous():
        raise B200AttnError("decode_plan: plan must be a contiguous int32 tensor")
    if sliding_window_size is not None and sliding_window_size > -1:
        check(_lib.load().b200attn_decode_plan_window(
            seq_lens.data_ptr(), _is_i64(seq_lens), seq_lens.numel(), int(n_ctas), int(min_units_per_cta),
            int(sliding_window_size), plan.data_ptr(), plan.numel() * 4, _stream(plan)), "b200attn_decode_plan_window")
        return
    check(_lib.load().b200attn_decode_plan(
        seq_lens.data_ptr(), _is_i64(seq_lens), seq_lens.numel(), int(n_ctas), int(min_units_per_cta),
        plan.data_ptr(), plan.numel() * 4, _stream(plan)), "b200attn_decode_plan")


def decode(q: torch.Tensor, out: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor,
           req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, seq_lens: torch.Tensor,
           plan: torch.Tensor, n_ctas: int, workspace: torch.Tensor, sm_scale: float, logit_cap: float = 0.0,
           k_scale: float = 1.0, v_scale: float = 1.0, kv_dtype: Optional[torch.dtype] = None,
           lse_out: Optional[torch.Tensor] = None, stage: str = "both", page_size: int = 1) -> torch.Tensor:
    """q,out [bs,Hq,D]; pools [P,Hkv,D]; see b200attn_decode in include/b200attn.h.
    stage: "both" (b200attn_decode), "partial" (stage 1 only) or "merge" (stage 2 only).
    page_size > 1: the allocator's page size (b200attn_decode_paged; multiples of 16 take the TMA box loads)."""
    _cuda(q, out, k_pool, v_pool, req_to_token, req_pool_indices, seq_lens, plan, workspace, lse_out)
    bs, Hq, D = q.shape
    Hkv = k_pool.shape[1]
    if out.shape != q.shape or out.dtype != q.dtype:
        raise B200AttnError("decode: out must match q")
    if q.stride(2) != 1 or q.stride(1) != D or out.stride(2) != 1 or out.stride(1) != D:
        raise B200AttnError("decode: q/out head stride must be D")
    if req_to_token.dtype != torch.int32 or req_to_token.stride(1) != 1:
        raise B200AttnError("decode: req_to_token must be int32 with contiguous rows")
    if req_pool_indices.dtype != seq_lens.dtype:
        raise B200AttnError("decode: req_pool_indices and seq_lens must share a dtype")
    if k_pool.stride(1) != D or k_pool.stride(2) != 1 or v_pool.stride() != k_pool.stride():
        raise B200AttnError("decode: pool must be [P,Hkv,D] with strides (*, D, 1)")
    if lse_out is not None and (lse_out.dtype != torch.float32 or not lse_out.is_contiguous()):
        raise B200AttnError("decode: lse_out must be contiguous fp32 [bs,Hq]")
    if stage == "merge":
        check(_lib.load().b200attn_decode_merge(
            out.data_ptr(), out.stride(0), bs, Hq, D, _dt(q.dtype), plan.data_ptr(), int(n_ctas),
            workspace.data_ptr(), workspace.numel() * workspace.element_size(), _ptr(lse_out), _stream(q)),
            "b200attn_decode_merge")
        return out
    if page_size > 1:
        check(_lib.load().b200attn_decode_paged(
            q.data_ptr(), q.stride(0), out.data_ptr(), out.stride(0), k_pool.data_ptr(), v_pool.data_ptr(),
            k_pool.stride(0), req_to_token.data_ptr(), req_to_token.stride(0), req_pool_indices.data_ptr(),
            seq_lens.data_ptr(), _is_i64(seq_lens), bs, Hq, Hkv, D, _dt(q.dtype), _pool_dtype(k_pool, kv_dtype),
            float(sm_scale), float(logit_cap or 0.0), float(k_scale), float(v_scale), plan.data_ptr(), int(n_ctas),
            workspace.data_ptr(), workspace.numel() * workspace.element_size(), _ptr(lse_out), int(page_size),
            0 if stage == "both" else 1, _stream(q)), "b200attn_decode_paged")
        return out
    fn = _lib.load().b200attn_decode if stage == "both" else _lib.load().b200attn_decode_partial
    check(fn(
        q.data_ptr(), q.stride(0), out.data_ptr(), out.stride(0), k_pool.data_ptr(), v_pool.data_ptr(),
        k_pool.stride(0), req_to_token.data_ptr(), req_to_token.stride(0), req_pool_indices.data_ptr(),
        seq_lens.data_ptr(), _is_i64(seq_lens), bs, Hq, Hkv, D, _dt(q.dtype), _pool_dtype(k_pool, kv_dtype),
        float(sm_scale), float(logit_cap or 0.0), float(k_scale), float(v_scale), plan.data_ptr(), int(n_ctas),
        workspace.data_ptr(), workspace.numel() * workspace.element_size(), _ptr(lse_out), _stream(q)),
        "b200attn_decode")
    return out


def decode_static_ws_capacity() -> int:
    """Bytes of split-KV partials the library's built-in workspace can hold (b200attn_decode_static_ws_capacity)."""
    return _lib.load().b200attn_decode_static_ws_capacity()


def decode_static_ws(q: torch.Tensor, out: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor,
                     req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, seq_lens: torch.Tensor,
                     plan: torch.Tensor, n_ctas: int, sm_scale: float, logit_cap: float = 0.0,
                     k_scale: float = 1.0, v_scale: float = 1.0, kv_dtype: Optional[torch.dtype] = None,
                     lse_out: Optional[torch.Tensor] = None, page_size: int = 1) -> torch.Tensor:
    """`decode(..., stage="both")` without the `workspace` argument: b200attn_decode_static_ws.

    For embedders that cannot hand the library a scratch buffer -- a binding with no CUDA allocator of its own, or a
    server that has already given all free HBM to the KV pool and must not cudaMalloc per step.  The partials go into
    a 10 MiB array inside the library (`decode_static_ws_capacity()`); a problem whose
    `decode_workspace_bytes(n_ctas, Hq, D)` is larger raises B200AttnError [WORKSPACE].  Neither this wrapper nor the
    C call allocates.  Calls share that one array per GPU: keep them on one stream (or order them with events), and
    make the first call on each GPU during start-up, before capturing any stream there."""
    _cuda(q, out, k_pool, v_pool, req_to_token, req_pool_indices, seq_lens, plan, lse_out)
    bs, Hq, D = q.shape
    Hkv = k_pool.shape[1]
    if out.shape != q.shape or out.dtype != q.dtype:
        raise B200AttnError("decode_static_ws: out must match q")
    if q.stride(2) != 1 or q.stride(1) != D or out.stride(2) != 1 or out.stride(1) != D:
        raise B200AttnError("decode_static_ws: q/out head stride must be D")
    if req_to_token.dtype != torch.int32 or req_to_token.stride(1) != 1:
        raise B200AttnError("decode_static_ws: req_to_token must be int32 with contiguous rows")
    if req_pool_indices.dtype != seq_lens.dtype:
        raise B200AttnError("decode_static_ws: req_pool_indices and seq_lens must share a dtype")
    if k_pool.stride(1) != D or k_pool.stride(2) != 1 or v_pool.stride() != k_pool.stride():
        raise B200AttnError("decode_static_ws: pool must be [P,Hkv,D] with strides (*, D, 1)")
    if lse_out is not None and (lse_out.dtype != torch.float32 or not lse_out.is_contiguous()):
        raise B200AttnError("decode_static_ws: lse_out must be contiguous fp32 [bs,Hq]")
    with torch.cuda.device(q.device):  # the symbol lookup and the launch both act on the current device
        check(_lib.load().b200attn_decode_static_ws(
            q.data_ptr(), q.stride(0), out.data_ptr(), out.stride(0), k_pool.data_ptr(), v_pool.data_ptr(),
            k_pool.stride(0), req_to_token.data_ptr(), req_to_token.stride(0), req_pool_indices.data_ptr(),
            seq_lens.data_ptr(), _is_i64(seq_lens), bs, Hq, Hkv, D, _dt(q.dtype), _pool_dtype(k_pool, kv_dtype),
            float(sm_scale), float(logit_cap or 0.0), float(k_scale), float(v_scale), plan.data_ptr(), int(n_ctas),
            _ptr(lse_out), int(page_size), _stream(q)), "b200attn_decode_static_ws")
    return out


def decode_multiq(q: torch.Tensor, out: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor,
                  req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, seq_lens: torch.Tensor,
                  plan: torch.Tensor, n_ctas: int, workspace: torch.Tensor, sm_scale: float, logit_cap: float = 0.0,
                  k_scale: float = 1.0, v_scale: float = 1.0, kv_dtype: Optional[torch.dtype] = None,
                  lse_out: Optional[torch.Tensor] = None, page_size: int = 1) -> torch.Tensor:
    """q,out [bs,nq,Hq,D] (any request / token strides, head stride D); every one of the nq tokens of request i attends
    to keys [0, seq_lens[i]) -- see b200attn_decode_multiq.  lse_out: fp32 [bs,nq,Hq] (request stride free)."""
    _cuda(q, out, k_pool, v_pool, req_to_token, req_pool_indices, seq_lens, plan, workspace, lse_out)
    bs, nq, Hq, D = q.shape
    Hkv = k_pool.shape[1]
    if out.shape != q.shape or out.dtype != q.dtype:
        raise B200AttnError("decode_multiq: out must match q")
    if q.stride(3) != 1 or q.stride(2) != D or out.stride(3) != 1 or out.stride(2) != D:
        raise B200AttnError("decode_multiq: q/out head stride must be D")
    if req_to_token.dtype != torch.int32 or req_to_token.stride(1) != 1:
        raise B200AttnError("decode_multiq: req_to_token must be int32 with contiguous rows")
    if req_pool_indices.dtype != seq_lens.dtype:
        raise B200AttnError("decode_multiq: req_pool_indices and seq_lens must share a dtype")
    if lse_out is not None and (lse_out.dtype != torch.float32 or lse_out.shape != (bs, nq, Hq)
                                or lse_out.stride(2) != 1 or lse_out.stride(1) != Hq):
        raise B200AttnError("decode_multiq: lse_out must be fp32 [bs,nq,Hq] with contiguous (token, head)")
    args = (
        q.data_ptr(), q.stride(0), q.stride(1), out.data_ptr(), out.stride(0), out.stride(1), k_pool.data_ptr(),
        v_pool.data_ptr(), k_pool.stride(0), req_to_token.data_ptr(), req_to_token.stride(0),
        req_pool_indices.data_ptr(), seq_lens.data_ptr(), _is_i64(seq_lens), bs, nq, Hq, Hkv, D, _dt(q.dtype),
        _pool_dtype(k_pool, kv_dtype), float(sm_scale), float(logit_cap or 0.0), float(k_scale), float(v_scale),
        plan.data_ptr(), int(n_ctas), workspace.data_ptr(), workspace.numel() * workspace.element_size(),
        _ptr(lse_out), lse_out.stride(0) if lse_out is not None else 0)
    if page_size > 1:
        check(_lib.load().b200attn_decode_multiq_paged(*args, int(page_size), _stream(q)), "b200attn_decode_multiq_paged")
    else:
        check(_lib.load().b200attn_decode_multiq(*args, _stream(q)), "b200attn_decode_multiq")
    return out


def merge_state(a: torch.Tensor, lse_a: torch.Tensor, b: torch.Tensor, lse_b: torch.Tensor,
                out: Optional[torch.Tensor] = None, out_lse: Optional[torch.Tensor] = None):
    """merge_state / merge_state_v2 (layers/attention/merge_state.py:26-46)."""
    _cuda(a, lse_a, b, lse_b, out, out_lse)
    if a.dim() != 3 or a.shape != b.shape or a.dtype != b.dtype:
        raise B200AttnError("merge_state: a,b must be [T,H,D] of equal shape/dtype")
    T, H, D = a.shape
    a, b = a.contiguous(), b.contiguous()
    lse_a = lse_a.to(torch.float32).contiguous(); lse_b = lse_b.to(torch.float32).contiguous()
    if out is None:
        out = torch.empty_like(a)
    if out_lse is None:
        out_lse = torch.empty((T, H), dtype=torch.float32, device=a.device)
    if not out.is_contiguous() or not out_lse.is_contiguous():
        raise B200AttnError("merge_state: outputs must be contiguous")
    check(_lib.load().b200attn_merge_state(
        a.data_ptr(), lse_a.data_ptr(), b.data_ptr(), lse_b.data_ptr(), out.data_ptr(), out_lse.data_ptr(),
        T, H, D, _dt(a.dtype), _stream(a)), "b200attn_merge_state")
    return out, out_lse


def extend_split_workspace_bytes(T: int, Hq: int, D: int, kv_splits: int) -> int:
    return int(_lib.load().b200attn_extend_split_workspace_bytes(int(T), int(Hq), int(D), int(kv_splits)))


def extend(q: torch.Tensor, out: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor,
           req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, seq_lens: torch.Tensor,
           extend_seq_lens: torch.Tensor, extend_prefix_lens: torch.Tensor, extend_start_loc: torch.Tensor,
           max_extend_len: int, sm_scale: float, logit_cap: float = 0.0, k_scale: float = 1.0,
           v_scale: float = 1.0, causal: bool = True, kv_dtype: Optional[torch.dtype] = None,
           impl: int = _lib.EXTEND_AUTO, lse_out: Optional[torch.Tensor] = None, sliding_window_size: int = -1,
           custom_mask: Optional[torch.Tensor] = None, mask_indptr: Optional[torch.Tensor] = None,
           skip_prefix_custom_mask: bool = True, page_size: int = 1, kv_splits: int = 1,
           workspace: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q,out [T,Hq,D]; see b200attn_extend / b200attn_extend_masked / b200attn_extend_paged in include/b200attn.h.
    page_size > 1: the allocator's page size (16-bit pools with page_size % 16 == 0 take TMA box loads).
    sliding_window_size / custom_mask / mask_indptr / skip_prefix_custom_mask carry the meaning they have in
    extend_attention_fwd (triton_ops/extend_attention.py:306-324)."""
    _cuda(q, out, k_pool, v_pool, req_to_token, req_pool_indices, seq_lens, extend_seq_lens,
          extend_prefix_lens, extend_start_loc)
    T, Hq, D = q.shape
    Hkv = k_pool.shape[1]
    bs = seq_lens.numel()
    if out.shape != q.shape or out.dtype != q.dtype:
        raise B200AttnError("extend: out must match q")
    if q.stride(2) != 1 or q.stride(1) != D or out.stride(2) != 1 or out.stride(1) != D:
        raise B200AttnError("extend: q/out head stride must be D")
    for name, t in (("extend_seq_lens", extend_seq_lens), ("extend_prefix_lens", extend_prefix_lens),
                    ("extend_start_loc", extend_start_loc)):
        if t.dtype != torch.int32 or not t.is_contiguous() or t.numel() != bs:
            raise B200AttnError(f"extend: {name} must be contiguous int32 [bs]")
    if req_to_token.dtype != torch.int32 or req_to_token.stride(1) != 1:
        raise B200AttnError("extend: req_to_token must be int32 with contiguous rows")
    if req_pool_indices.dtype != seq_lens.dtype:
        raise B200AttnError("extend: req_pool_indices and seq_lens must share a dtype")
    masked = custom_mask is not None or (sliding_window_size is not None and sliding_window_size > -1)
    if kv_splits > 1:
        # K/V range of every item cut into kv_splits shares (b200attn_extend_split); workspace: uint8 / any dtype tensor of
        # at least extend_split_workspace_bytes(T, Hq, D, kv_splits) bytes
        if masked or (logit_cap or 0.0) != 0.0:
            raise B200AttnError("extend: kv_splits > 1 takes plain batches only (no window, mask or logit cap)")
        need = extend_split_workspace_bytes(T, Hq, D, kv_splits)
        if workspace is None or not workspace.is_cuda or workspace.numel() * workspace.element_size() < need:
            raise B200AttnError(f"extend: kv_splits={kv_splits} needs a CUDA workspace of {need} bytes")
        check(_lib.load().b200attn_extend_split(
            q.data_ptr(), q.stride(0), out.data_ptr(), out.stride(0), k_pool.data_ptr(), v_pool.data_ptr(),
            k_pool.stride(0), req_to_token.data_ptr(), req_to_token.stride(0), req_pool_indices.data_ptr(),
            seq_lens.data_ptr(), _is_i64(seq_lens), extend_seq_lens.data_ptr(), extend_prefix_lens.data_ptr(),
            extend_start_loc.data_ptr(), bs, int(max_extend_len), Hq, Hkv, D, _dt(q.dtype),
            _pool_dtype(k_pool, kv_dtype), float(sm_scale), float(k_scale), float(v_scale), int(bool(causal)),
            _ptr(lse_out), int(page_size), int(T), int(kv_splits), workspace.data_ptr(),
            workspace.numel() * workspace.element_size(), _stream(q)), "b200attn_extend_split")
        return out
    if masked or page_size > 1:
        if custom_mask is not None:
            _cuda(custom_mask, mask_indptr)
            if custom_mask.dtype not in (torch.bool, torch.uint8) or not custom_mask.is_contiguous():
                raise B200AttnError("extend: custom_mask must be a contiguous bool/uint8 tensor")
            if mask_indptr is None or mask_indptr.dtype != torch.int64 or mask_indptr.numel() < bs + 1 \
                    or not mask_indptr.is_contiguous():
                raise B200AttnError("extend: mask_indptr must be contiguous int64 [bs+1]")
        args = (
            q.data_ptr(), q.stride(0), out.data_ptr(), out.stride(0), k_pool.data_ptr(), v_pool.data_ptr(),
            k_pool.stride(0), req_to_token.data_ptr(), req_to_token.stride(0), req_pool_indices.data_ptr(),
            seq_lens.data_ptr(), _is_i64(seq_lens), extend_seq_lens.data_ptr(), extend_prefix_lens.data_ptr(),
            extend_start_loc.data_ptr(), bs, int(max_extend_len), Hq, Hkv, D, _dt(q.dtype),
            _pool_dtype(k_pool, kv_dtype), float(sm_scale), float(logit_cap or 0.0), float(k_scale), float(v_scale),
            int(bool(causal)), int(impl), int(sliding_window_size if sliding_window_size is not None else -1),
            _ptr(custom_mask), _ptr(mask_indptr), int(bool(skip_prefix_custom_mask)), _ptr(lse_out))
        if page_size > 1:
            check(_lib.load().b200attn_extend_paged(*args, int(page_size), _stream(q)), "b200attn_extend_paged")
        else:
            check(_lib.load().b200attn_extend_masked(*args, _stream(q)), "b200attn_extend_masked")
        return out
    check(_lib.load().b200attn_extend(
        q.data_ptr(), q.stride(0), out.data_ptr(), out.stride(0), k_pool.data_ptr(), v_pool.data_ptr(),
        k_pool.stride(0), req_to_token.data_ptr(), req_to_token.stride(0), req_pool_indices.data_ptr(),
        seq_lens.data_ptr(), _is_i64(seq_lens), extend_seq_lens.data_ptr(), extend_prefix_lens.data_ptr(),
        extend_start_loc.data_ptr(), bs, int(max_extend_len), Hq, Hkv, D, _dt(q.dtype),
        _pool_dtype(k_pool, kv_dtype), float(sm_scale), float(logit_cap or 0.0), float(k_scale), float(v_scale),
        int(bool(causal)), int(impl), _ptr(lse_out), _stream(q)), "b200attn_extend")
    return out


def check_loc(loc: torch.Tensor, dummy_rows: int, pool_rows: int) -> tuple:
    """(duplicate pairs outside the dummy page, entries out of range) of a KV-write slot list -- a debug check that
    SYNCHRONISES; see b200attn_check_loc.  The reference is last-writer-wins for duplicates (memory_pool.py:457-458)."""
    _cuda(loc)
    if loc.dtype != torch.int64 or not loc.is_contiguous():
        raise B200AttnError("check_loc: loc must be a contiguous int64 tensor")
    flags = torch.empty(2, dtype=torch.int32, device=loc.device)
    check(_lib.load().b200attn_check_loc(loc.data_ptr(), loc.numel(), int(dummy_rows), int(pool_rows),
                                         flags.data_ptr(), _stream(loc)), "b200attn_check_loc")
    d, o = flags.tolist()
    return int(d), int(o)


def flatten_kv_indices(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, lens: torch.Tensor,
                       kv_indptr: torch.Tensor, kv_start: Optional[torch.Tensor], kv_indices: torch.Tensor) -> None:
    """create_flashinfer_kv_indices_triton (layers/attention/utils.py:5-41)."""
    _cuda(req_to_token, req_pool_indices, lens, kv_indptr, kv_start, kv_indices)
    if req_to_token.dtype != torch.int32 or kv_indptr.dtype != torch.int32 or kv_indices.dtype != torch.int32:
        raise B200AttnError("flatten_kv_indices: table, kv_indptr and kv_indices must be int32")
    if lens.dtype != req_pool_indices.dtype or (kv_start is not None and kv_start.dtype != lens.dtype):
        raise B200AttnError("flatten_kv_indices: req_pool_indices, lens and kv_start must share a dtype")
    check(_lib.load().b200attn_flatten_kv_indices(
        req_to_token.data_ptr(), req_to_token.stride(0), req_pool_indices.data_ptr(), lens.data_ptr(),
        _ptr(kv_start), _is_i64(lens), kv_indptr.data_ptr(), kv_indices.data_ptr(), lens.numel(),
        _stream(req_to_token)), "b200attn_flatten_kv_indices")


def move_kv(data_ptrs: torch.Tensor, data_strides: torch.Tensor, max_stride_bytes: int, tgt_loc: torch.Tensor,
            src_loc: torch.Tensor) -> None:
    """buf[tgt_loc] = buf[src_loc] in every buffer listed in data_ptrs (uint64 addresses) / data_strides (int64 row bytes),
    one launch -- MHATokenToKVPool.move_kv_cache (memory_pool.py:460-468).  Sources are read as they were before the
    call (the reference's in-place rule, :1170-1171)."""
    _cuda(data_ptrs, data_strides, tgt_loc, src_loc)
    if data_ptrs.dtype not in (torch.uint64, torch.int64) or data_strides.dtype != torch.int64:
        raise B200AttnError("move_kv: data_ptrs must be uint64 / int64 and data_strides int64")
    if not (data_ptrs.is_contiguous() and data_strides.is_contiguous() and tgt_loc.is_contiguous()
            and src_loc.is_contiguous()):
        raise B200AttnError("move_kv: arguments must be contiguous")
    if data_ptrs.numel() != data_strides.numel() or tgt_loc.numel() != src_loc.numel() or tgt_loc.dtype != src_loc.dtype:
        raise B200AttnError("move_kv: data_ptrs / data_strides and tgt_loc / src_loc must pair up")
    check(_lib.load().b200attn_move_kv(
        data_ptrs.data_ptr(), data_strides.data_ptr(), data_ptrs.numel(), int(max_stride_bytes), tgt_loc.data_ptr(),
        src_loc.data_ptr(), _is_i64(tgt_loc), tgt_loc.numel(), _stream(tgt_loc)), "b200attn_move_kv")


def gather_kv(pool: torch.Tensor, indices: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out[j] = pool[indices[j]] -- the page gather `k_cache[per_req_tokens]`, bit exact."""
    _cuda(pool, indices, out)
    if indices.dtype != torch.int32 or not indices.is_contiguous():
        raise B200AttnError("gather_kv: indices must be contiguous int32")
    row = pool[0]
    if not row.is_contiguous():
        raise B200AttnError("gather_kv: pool rows must be contiguous")
    if out is None:
        out = torch.empty((indices.numel(),) + tuple(pool.shape[1:]), dtype=pool.dtype, device=pool.device)
    row_bytes = row.numel() * pool.element_size()
    check(_lib.load().b200attn_gather_kv(
        pool.data_ptr(), pool.stride(0) * pool.element_size(), indices.data_ptr(), indices.numel(), row_bytes,
        out.data_ptr(), _stream(pool)), "b200attn_gather_kv")
    return out
