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
                 what: str = "") -> None:
    g, w = got.detach().float().cpu(), want.detach().float().cpu()
    assert g.shape == w.shape, f"{what}: shape {tuple(g.shape)} vs {tuple(w.shape)}"
    assert torch.isfinite(g).all(), f"{what}: non-finite values in the CUDA output"
    d = (g - w).abs()
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
def to_device(case, device) -> types.SimpleNamespace:
    d = types.SimpleNamespace(**vars(case))
    for k, v in vars(case).items():
        if isinstance(v, torch.Tensor):
            setattr(d, k, v.to(device))
    return d


class Runner:
    """Drives the product path exactly as SGLang would: ModelRunner stand-in -> B200AttnBackend ->
    RadixAttention.forward -> backend.forward_decode / forward_extend -> C ABI."""

    def __init__(self, case, device, max_reqs: Optional[int] = None):
        from sglang_custom_b200 import interfaces as itf
        from sglang_custom_b200.backend import B200AttnBackend
        self.itf = itf
        self.case = case
        self.device = device
        kv_dtype = torch.float8_e4m3fn if case.fp8 else case.dtype
        cfg = itf.ModelConfig(context_len=case.req_to_token.shape[1], num_attention_heads=case.Hq,
                              num_key_value_heads=case.Hkv, head_dim=case.D, num_hidden_layers=1)
        self.mr = itf.ModelRunner(cfg, max_reqs=case.req_to_token.shape[0], pool_size=case.P - case.page_size,
                                  device=str(device), kv_dtype=kv_dtype, page_size=case.page_size, dtype=case.dtype)
        self.mr.req_to_token_pool.req_to_token.copy_(case.req_to_token.to(device))
        pool = self.mr.token_to_kv_pool
        pool.k_buffer[0].copy_(case.k_pool.to(device))
        pool.v_buffer[0].copy_(case.v_pool.to(device))
        self.backend = B200AttnBackend(self.mr)
        s = case.kv_scale if case.fp8 and case.kv_scale != 1.0 else None
        self.layer = itf.RadixAttention(num_heads=case.Hq, head_dim=case.D, scaling=case.sm_scale,
                                        num_kv_heads=case.Hkv, layer_id=0, k_scale_float=s, v_scale_float=s)

    def batch(self, mode=None):
        itf, c, dev = self.itf, self.case, self.device
        if mode is None:
            mode = itf.ForwardMode.DECODE if c.decode else itf.ForwardMode.EXTEND
        fb = itf.ForwardBatch(
            forward_mode=mode, batch_size=c.bs, req_pool_indices=c.req_pool_indices.to(dev),
            seq_lens=c.seq_lens.to(dev), out_cache_loc=c.out_cache_loc.to(dev), seq_lens_sum=c.seq_lens_sum,
            req_to_token_pool=self.mr.req_to_token_pool, token_to_kv_pool=self.mr.token_to_kv_pool,
            attn_backend=self.backend, seq_lens_cpu=c.seq_lens.cpu())
        if not mode.is_decode():
            fb.extend_seq_lens = c.extend_seq_lens.to(dev)
            fb.extend_prefix_lens = c.extend_prefix_lens.to(dev)
            fb.extend_start_loc = c.extend_start_loc.to(dev)
            fb.extend_seq_lens_cpu = c.extend_seq_lens.tolist()
            fb.extend_prefix_lens_cpu = c.extend_prefix_lens.tolist()
            fb.extend_num_tokens = c.T
        return fb

    def forward(self, mode=None, save_kv_cache: bool = True, strided_qkv: bool = True) -> torch.Tensor:
        """One attention forward; q/k/v arrive as strided views of one qkv buffer, as in LlamaAttention
        (models/llama.py:182-193) and as the reference's CPU tests make them on purpose (test_decode.py:105-110)."""
        c, dev = self.case, self.device
        fb = self.batch(mode)
        self.backend.init_forward_metadata(fb)
        if strided_qkv and c.T > 0:
            qkv = torch.cat([c.q.reshape(c.T, -1), c.k_new.reshape(c.T, -1), c.v_new.reshape(c.T, -1)], dim=1).to(dev)
            q, k, v = qkv.split([c.Hq * c.D, c.Hkv * c.D, c.Hkv * c.D], dim=1)
        else:
            q = c.q.reshape(c.T, -1).to(dev)
            k = c.k_new.reshape(c.T, -1).to(dev)
            v = c.v_new.reshape(c.T, -1).to(dev)
        out = self.layer.forward(q, k, v, fb, save_kv_cache=save_kv_cache)
        torch.cuda.synchronize(dev)
        return out.view(c.T, c.Hq, c.D)

    def pools(self):
        pool = self.mr.token_to_kv_pool
        return pool.k_buffer[0], pool.v_buffer[0]
