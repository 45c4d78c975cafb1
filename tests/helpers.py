"""Shared helpers of the parity tests: run a synthetic case through the oracle and through the CUDA path."""
from __future__ import annotations

import types
from typing import Optional

import torch

from oracle import exact
from oracle import torch_native as tn

# north_star tolerances (BASELINE.json): bf16 attention outputs within max-abs 2e-2 and mean-abs 2e-3 of
# the reference's torch_native CPU attention; FP8-KV outputs vs the oracle on the dequantised pool: the
# same bound is used (the quantisation error sits on both sides; P.V is never done in fp8).
MAX_ABS_TOL = 2e-2
MEAN_ABS_TOL = 2e-3
# against the fp64-accumulating restatement the CUDA path (fp32 accumulate, bf16 P) must be much closer
EXACT_MAX_ABS_TOL = 8e-3
EXACT_MEAN_ABS_TOL = 6e-4


def pool_after_write(case) -> tuple:
    """Oracle KV write (memory_pool.py:420-458 restated): returns (k_pool, v_pool) after the forward's write.
    For an fp8 case the pool is uint8 and the write divides by kv_scale (when != 1) before the cast."""
    kp, vp = case.k_pool.clone(), case.v_pool.clone()
    if case.T == 0:
        return kp, vp
    if case.fp8:
        s = None if case.kv_scale == 1.0 else case.kv_scale
        tn.set_kv_buffer(kp, vp, case.out_cache_loc, case.k_new.clone(), case.v_new.clone(), torch.float8_e4m3fn, s, s)
    else:
        tn.set_kv_buffer(kp, vp, case.out_cache_loc, case.k_new.clone(), case.v_new.clone(), case.dtype)
    return kp, vp


def oracle_pools_for_attention(case, kp, vp):
    """What oracle #1 attends over: the pool itself, or its dequantised copy for fp8."""
    if case.fp8:
        return tn.dequantised_pool(kp, case.kv_scale, case.dtype), tn.dequantised_pool(vp, case.kv_scale, case.dtype)
    return kp, vp


def oracle_decode(case, kp=None, vp=None) -> torch.Tensor:
    if kp is None:
        kp, vp = pool_after_write(case)
    k, v = oracle_pools_for_attention(case, kp, vp)
    return tn.decode(case.q, k, v, case.req_to_token, case.req_pool_indices, case.seq_lens, case.sm_scale)


def oracle_extend(case, kp=None, vp=None, causal: bool = True) -> torch.Tensor:
    if kp is None:
        kp, vp = pool_after_write(case)
    k, v = oracle_pools_for_attention(case, kp, vp)
    return tn.extend(case.q, k, v, case.req_to_token, case.req_pool_indices, case.seq_lens,
                     case.extend_prefix_lens, case.extend_seq_lens, case.sm_scale, causal=causal)


def exact_decode(case, kp, vp, logit_cap: float = 0.0, return_lse: bool = False):
    s = case.kv_scale if case.fp8 else 1.0
    return exact.decode(case.q, kp, vp, case.req_to_token, case.req_pool_indices, case.seq_lens, case.sm_scale,
                        logit_cap=logit_cap, k_scale=s, v_scale=s, return_lse=return_lse)


def exact_extend(case, kp, vp, causal: bool = True, logit_cap: float = 0.0, return_lse: bool = False):
    s = case.kv_scale if case.fp8 else 1.0
    return exact.extend(case.q, kp, vp, case.req_to_token, case.req_pool_indices, case.seq_lens,
                        case.extend_seq_lens, case.extend_prefix_lens, case.extend_start_loc, case.sm_scale,
                        logit_cap=logit_cap, k_scale=s, v_scale=s, causal=causal, return_lse=return_lse)


def assert_close(got: torch.Tensor, want: torch.Tensor, max_abs: float = MAX_ABS_TOL, mean_abs: float = MEAN_ABS_TOL,
                 what: str = "", rel: bool = False) -> None:
    """rel=True scales the max-abs bound by max(1, |want|): the output is rounded to bf16 (half an ulp of a value in
    [2, 4) is already 7.8e-3), which matters only for the tight comparison against the fp64 restatement."""
    g, w = got.detach().float().cpu(), want.detach().float().cpu()
    assert g.shape == w.shape, f"{what}: shape {tuple(g.shape)} vs {tuple(w.shape)}"
    assert torch.isfinite(g).all(), f"{what}: non-finite values in the CUDA output"
    d = (g - w).abs()
    if rel:
        d = d / w.abs().clamp(min=1.0)
    mx, mn = float(d.max()) if d.numel() else 0.0, float(d.mean()) if d.numel() else 0.0
    assert mx <= max_abs and mn <= mean_abs, f"{what}: max-abs {mx:.3e} (tol {max_abs}) mean-abs {mn:.3e} (tol {mean_abs})"


def canon_e4m3(u8: torch.Tensor) -> torch.Tensor:
    """e4m3fn has one NaN per sign (0x7F / 0xFF).  Which sign a NaN input produces depends on how the
    host's FPU propagated it through `div_` (x86 hands back the negative default NaN in some paths), so
    bit comparisons of fp8 pools treat the two NaN encodings as one.  Every other byte is compared as is."""
    u8 = u8.cpu()
    return torch.where((u8 & 0x7F) == 0x7F, torch.full_like(u8, 0x7F), u8)


def canon_16(x: torch.Tensor) -> torch.Tensor:
    """bf16/f16 bit pattern with every NaN mapped to one value (sign/payload of NaN is not part of parity)."""
    x = x.cpu()
    bits = x.view(torch.int16).clone()
    bits[torch.isnan(x)] = 0x7FC0 if x.dtype == torch.bfloat16 else 0x7E00
    return bits


def pools_equal(a: torch.Tensor, b: torch.Tensor) -> bool:
    a, b = a.cpu(), b.cpu()
    if a.dtype == torch.uint8:
        return torch.equal(canon_e4m3(a), canon_e4m3(b))
    return torch.equal(a.view(torch.int16), b.view(torch.int16))


# ------------------------------------------------------------------ CUDA side
from sglang_custom_b200.harness import Runner, to_device  # noqa: E402,F401  (oracle-free; bench.py uses it directly)
