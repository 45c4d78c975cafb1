"""Rotary embedding whose forward is the fused b200attn RoPE (+ KV-write) kernel.

Host-side mirror of the reference's `RotaryEmbedding` (python/sglang/srt/layers/rotary_embedding.py:69-250):
same constructor arguments, the same fp32 cos|sin table (`_compute_inv_freq` :104-118, `_compute_cos_sin_cache`
:120-136 -- table construction is one-off init work done with torch ops, as in the reference) and the same
`forward(positions, query, key, offsets=None)` contract: query/key are rotated IN PLACE and returned, like the
reference's CUDA path (`forward_cuda` :218-244 -> sgl-kernel `apply_rope_with_cos_sin_cache_inplace`).

`_compute_inv_freq`, `_compute_cos_sin_cache` and the Llama-3 frequency-scaling formula follow the reference line by
line on purpose: the cos|sin table has to be bit-identical to the reference's, so there is no second way to write them.

`forward_fused_kv` is the SURVEY 8f rank-2 row: RoPE and `set_kv_buffer` (memory_pool.py:420-458) in one launch; the
model then calls attention with `save_kv_cache=False` (the hook RadixAttention.forward already has,
radix_attention.py:82-108).  There is no torch fallback: without libb200attn.so these raise.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple, Union

import torch

from . import _lib, ops


class RotaryEmbedding:
    def __init__(self, head_size: int, rotary_dim: int, max_position_embeddings: int, base: Union[int, float],
                 is_neox_style: bool, dtype: torch.dtype, device="cuda", cache_dtype: torch.dtype = torch.float32,
                 math_mode: int = _lib.ROPE_FP32) -> None:
        self.head_size = head_size
        self.rotary_dim = rotary_dim
        self.max_position_embeddings = max_position_embeddings
        self.base = base
        self.is_neox_style = is_neox_style
        self.dtype = dtype
        self.math_mode = math_mode
        # the reference keeps the table in fp32 on CUDA and casts it to `dtype` elsewhere (:98-102)
        self.cos_sin_cache = self._compute_cos_sin_cache().to(cache_dtype).to(device).contiguous()

    def _compute_inv_freq(self, base: Union[int, float]) -> torch.Tensor:
        return 1.0 / (base ** (torch.arange(0, self.rotary_dim, 2, dtype=torch.float) / self.rotary_dim))

    def _compute_cos_sin_cache(self) -> torch.Tensor:
        inv_freq = self._compute_inv_freq(self.base)
        t = torch.arange(self.max_position_embeddings, dtype=torch.float)
        freqs = torch.einsum("i,j -> ij", t, inv_freq)
        return torch.cat((freqs.cos(), freqs.sin()), dim=-1)

    def _views(self, positions, query, key):
        positions = positions.flatten()
        T = positions.shape[0]
        return positions.contiguous(), query.view(T, -1, self.head_size), key.view(T, -1, self.head_size)

    def forward(self, positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                offsets: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        if offsets is not None:
            positions = positions + offsets
        pos, q3, k3 = self._views(positions, query, key)
        ops.rope_kv_write(pos, q3, k3, self.cos_sin_cache, self.is_neox_style, math_mode=self.math_mode)
        return query, key

    __call__ = forward

    def forward_fused_kv(self, positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                         layer, forward_batch, scale_mode: int = _lib.SCALE_DIV) -> Tuple[torch.Tensor, torch.Tensor]:
        """q, k = rotary_emb(positions, q, k); token_to_kv_pool.set_kv_buffer(layer, out_cache_loc, k, v, scales):
        one launch.  Follow with `attn(q, k, v, forward_batch, save_kv_cache=False)`."""
        pos, q3, k3 = self._views(positions, query, key)
        v3 = value.view(pos.shape[0], -1, self.head_size)
        pool = forward_batch.token_to_kv_pool
        lid = layer.layer_id - getattr(pool, "start_layer", 0)
        ks = layer.k_scale if getattr(layer, "k_scale", None) is not None else getattr(layer, "k_scale_float", None)
        vs = layer.v_scale if getattr(layer, "v_scale", None) is not None else getattr(layer, "v_scale_float", None)
        ops.rope_kv_write(pos, q3, k3, self.cos_sin_cache, self.is_neox_style, v=v3, loc=forward_batch.out_cache_loc,
                          k_pool=pool.k_buffer[lid], v_pool=pool.v_buffer[lid], k_scale=ks, v_scale=vs,
                          pool_dtype=pool.dtype, scale_mode=scale_mode, math_mode=self.math_mode)
        # tell the backend that this layer's new K/V of this batch are in the pool: forward_*(save_kv_cache=False)
        # with k/v given raises otherwise (it reads every key from the pool, never from k/v)
        note = getattr(getattr(forward_batch, "attn_backend", None), "note_kv_written", None)
        if note is not None:
            note(layer, forward_batch)
        return query, key


class Llama3RotaryEmbedding(RotaryEmbedding):
    """Llama-3.1 frequency scaling (rotary_embedding.py:783-827)."""

    def __init__(self, head_size, rotary_dim, max_position_embeddings, base, is_neox_style, dtype,
                 scaling_factor: float, low_freq_factor: float, high_freq_factor: float, orig_max_position: int,
                 **kw) -> None:
        self.scaling_factor = scaling_factor
        self.low_freq_factor = low_freq_factor
        self.high_freq_factor = high_freq_factor
        self.orig_max_position = orig_max_position
        super().__init__(head_size, rotary_dim, max_position_embeddings, base, is_neox_style, dtype, **kw)

    def _compute_inv_freq(self, base):
        inv_freqs = super()._compute_inv_freq(base)
        low_freq_wavelen = self.orig_max_position / self.low_freq_factor
        high_freq_wavelen = self.orig_max_position / self.high_freq_factor
        wave_len = 2 * math.pi / inv_freqs
        if self.low_freq_factor != self.high_freq_factor:
            smooth = (self.orig_max_position / wave_len - self.low_freq_factor) / (
                self.high_freq_factor - self.low_freq_factor)
        else:
            smooth = 0
        return torch.where(
            wave_len < high_freq_wavelen, inv_freqs,
            torch.where(wave_len > low_freq_wavelen, inv_freqs / self.scaling_factor,
                        (1 - smooth) * inv_freqs / self.scaling_factor + smooth * inv_freqs))
