"""Oracle of the rotary embedding that precedes the KV write.

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Restates, for the CPU, the reference's
python/sglang/srt/layers/rotary_embedding.py:
  cos|sin table          RotaryEmbedding._compute_inv_freq / _compute_cos_sin_cache   :104-136
  Llama-3 scaling        Llama3RotaryEmbedding._compute_inv_freq                      :806-827
  forward_native         :138-166 with _apply_rotary_emb :38-66 (op-by-op in the tensor dtype)
Pinned bit-exact against the unmodified reference class by tests/test_oracle_pin.py (build container) and
through the committed fixtures tests/golden/rope_*.pt (made by tests/golden/make_golden.py).

`rope_fp32` is the second semantics: the products in fp32 with one final rounding, which is what the reference's
CUDA op computes from its fp32 table (forward_cuda :218-244 -> flashinfer pos_enc, un-vendored: that one is
"parity unpinned" bit-wise and is checked against forward_native within the reference's own test tolerance,
sgl-kernel/tests/test_rotary_embedding.py).
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch


def inv_freq(rotary_dim: int, base: float, llama3: Optional[dict] = None) -> torch.Tensor:
    f = 1.0 / (base ** (torch.arange(0, rotary_dim, 2, dtype=torch.float) / rotary_dim))
    if llama3 is None:
        return f
    orig, lo, hi, factor = (llama3["original_max_position_embeddings"], llama3["low_freq_factor"],
                            llama3["high_freq_factor"], llama3["factor"])
    low_wavelen, high_wavelen = orig / lo, orig / hi
    wave_len = 2 * math.pi / f
    smooth = (orig / wave_len - lo) / (hi - lo) if lo != hi else 0
    return torch.where(wave_len < high_wavelen, f,
                       torch.where(wave_len > low_wavelen, f / factor, (1 - smooth) * f / factor + smooth * f))


def cos_sin_cache(rotary_dim: int, max_position: int, base: float, llama3: Optional[dict] = None) -> torch.Tensor:
    """fp32 [max_position, rotary_dim] = cos | sin."""
    t = torch.arange(max_position, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq(rotary_dim, base, llama3))
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1)


def _apply(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, is_neox: bool) -> torch.Tensor:
    cos = cos.unsqueeze(-2).to(x.dtype)
    sin = sin.unsqueeze(-2).to(x.dtype)
    if is_neox:
        x1, x2 = torch.chunk(x, 2, dim=-1)
    else:
        x1, x2 = x[..., ::2], x[..., 1::2]
    o1 = x1 * cos - x2 * sin
    o2 = x2 * cos + x1 * sin
    if is_neox:
        return torch.cat((o1, o2), dim=-1)
    return torch.stack((o1, o2), dim=-1).flatten(-2)


def rope_native(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, cache: torch.Tensor, head_size: int,
                is_neox: bool) -> Tuple[torch.Tensor, torch.Tensor]:
    """forward_native (:138-166): cache is whatever dtype the module holds (x dtype on CPU)."""
    positions = positions.flatten()
    T = positions.shape[0]
    rotary_dim = cache.shape[1]
    cos, sin = cache.index_select(0, positions).chunk(2, dim=-1)
    outs = []
    for x in (query, key):
        shape = x.shape
        x = x.view(T, -1, head_size)
        rot = _apply(x[..., :rotary_dim], cos, sin, is_neox)
        outs.append(torch.cat((rot, x[..., rotary_dim:]), dim=-1).reshape(shape))
    return outs[0], outs[1]


def rope_fp32(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, cache: torch.Tensor, head_size: int,
              is_neox: bool) -> Tuple[torch.Tensor, torch.Tensor]:
    """Same formula, fp32 products (each fp32 op rounded once, no fusion), a single rounding to the tensor dtype."""
    positions = positions.flatten()
    T = positions.shape[0]
    rotary_dim = cache.shape[1]
    cos, sin = cache.float().index_select(0, positions).chunk(2, dim=-1)
    outs = []
    for x in (query, key):
        shape, dt = x.shape, x.dtype
        x = x.view(T, -1, head_size)
        rot = _apply(x[..., :rotary_dim].float(), cos, sin, is_neox).to(dt)
        outs.append(torch.cat((rot, x[..., rotary_dim:]), dim=-1).reshape(shape))
    return outs[0], outs[1]
