"""Drive the backend the way SGLang does, without an SGLang install: a ModelRunner stand-in, the
B200AttnBackend, one RadixAttention layer and hand-built ForwardBatch objects (the pattern of the reference's
python/sglang/test/attention/test_flashattn_backend.py:15-70).  Used by tests/, smoke() and bench.py; imports
nothing from oracle/."""
from __future__ import annotations

import types
from typing import Optional

import torch


def to_device(case, device) -> types.SimpleNamespace:
    d = types.SimpleNamespace(**vars(case))
    for k, v in vars(case).items():
        if isinstance(v, torch.Tensor):
            setattr(d, k, v.to(device))
    return d


class Runner:
    """Drives the product path exactly as SGLang would: ModelRunner stand-in -> B200AttnBackend ->
    RadixAttention.forward -> backend.forward_decode / forward_extend -> C ABI."""

    def __init__(self, case, device, max_reqs: Optional[int] = None, sliding_window_size: Optional[int] = None,
                 num_draft_tokens: Optional[int] = None, tp_size: int = 1, tp_rank: int = 0):
        """tp_size > 1: `case` is rank tp_rank's slice (synth.make_case_device(head_shard=...) / tp.shard_case) of a
        model with case.Hq * tp_size query heads; the ModelRunner stand-in is built with the GLOBAL head counts and
        tp_size, and the backend derives its per-rank heads from them as TritonAttnBackend does
        (triton_backend.py:69-76)."""
        from sglang_custom_b200 import interfaces as itf
        from sglang_custom_b200.backend import B200AttnBackend
        self.itf = itf
        self.case = case
        self.device = device
        kv_dtype = torch.float8_e4m3fn if case.fp8 else case.dtype
        hkv_global = getattr(case, "Hkv_global", case.Hkv * tp_size)
        cfg = itf.ModelConfig(context_len=case.req_to_token.shape[1], num_attention_heads=case.Hq * tp_size,
                              num_key_value_heads=hkv_global, head_dim=case.D, num_hidden_layers=1)
        self.mr = itf.ModelRunner(cfg, max_reqs=case.req_to_token.shape[0], pool_size=case.P - case.page_size,
                                  device=str(device), kv_dtype=kv_dtype, page_size=case.page_size, dtype=case.dtype,
                                  tp_size=tp_size, tp_rank=tp_rank)
        assert cfg.get_num_kv_heads(tp_size) == case.Hkv, "case is not this rank's KV-head slice"
        self.mr.req_to_token_pool.req_to_token.copy_(case.req_to_token.to(device))
        pool = self.mr.token_to_kv_pool
        pool.k_buffer[0].copy_(case.k_pool.to(device))
        pool.v_buffer[0].copy_(case.v_pool.to(device))
        # sliding-window models (model_runner.sliding_window_size) / speculative decoding (server_args)
        self.mr.sliding_window_size = sliding_window_size
        self.mr.server_args.speculative_num_draft_tokens = num_draft_tokens
        self.backend = B200AttnBackend(self.mr)
        s = case.kv_scale if case.fp8 and case.kv_scale != 1.0 else None
        self.layer = itf.RadixAttention(num_heads=case.Hq, head_dim=case.D, scaling=case.sm_scale,
                                        num_kv_heads=case.Hkv, layer_id=0, k_scale_float=s, v_scale_float=s,
                                        sliding_window_size=-1 if sliding_window_size is None else sliding_window_size)

    def verify_batch(self, custom_mask: torch.Tensor):
        """TARGET_VERIFY batch as EagleVerifyInput.prepare_for_verify leaves it (speculative/eagle_utils.py:255-276):
        seq_lens = tokens already cached, the draft tokens' slots sit behind them in req_to_token, the mask comes in
        spec_info.custom_mask.  The case must have extend_lens == num_draft_tokens for every request."""
        itf, c, dev = self.itf, self.case, self.device
        prefix = c.extend_prefix_lens.to(c.seq_lens.dtype)
        return itf.ForwardBatch(
            forward_mode=itf.ForwardMode.TARGET_VERIFY, batch_size=c.bs, req_pool_indices=c.req_pool_indices.to(dev),
            seq_lens=prefix.to(dev), out_cache_loc=c.out_cache_loc.to(dev), seq_lens_sum=int(prefix.sum()),
            req_to_token_pool=self.mr.req_to_token_pool, token_to_kv_pool=self.mr.token_to_kv_pool,
            attn_backend=self.backend, seq_lens_cpu=prefix.cpu(),
            spec_info=types.SimpleNamespace(custom_mask=custom_mask.to(dev), draft_token_num=c.max_extend_len))

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

    def forward(self, mode=None, save_kv_cache: bool = True, strided_qkv: bool = True, fb=None) -> torch.Tensor:
        """One attention forward; q/k/v arrive as strided views of one qkv buffer, as in LlamaAttention
        (models/llama.py:182-193) and as the reference's CPU tests make them on purpose (test_decode.py:105-110)."""
        c, dev = self.case, self.device
        fb = self.batch(mode) if fb is None else fb
        self.backend.init_forward_metadata(fb)
        if strided_qkv and c.T > 0:
            qkv = torch.cat([c.q.reshape(c.T, c.Hq * c.D), c.k_new.reshape(c.T, c.Hkv * c.D),
                             c.v_new.reshape(c.T, c.Hkv * c.D)], dim=1).to(dev)
            q, k, v = qkv.split([c.Hq * c.D, c.Hkv * c.D, c.Hkv * c.D], dim=1)
        else:
            q = c.q.reshape(c.T, c.Hq * c.D).to(dev)
            k = c.k_new.reshape(c.T, c.Hkv * c.D).to(dev)
            v = c.v_new.reshape(c.T, c.Hkv * c.D).to(dev)
        if not save_kv_cache:
            k = v = None  # "the pool already holds the new tokens" (radix_attention.py:91-99 accepts k=None)
        out = self.layer.forward(q, k, v, fb, save_kv_cache=save_kv_cache)
        torch.cuda.synchronize(dev)
        return out.view(c.T, c.Hq, c.D)

    def pools(self):
        pool = self.mr.token_to_kv_pool
        return pool.k_buffer[0], pool.v_buffer[0]
