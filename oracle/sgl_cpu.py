"""Oracle #2 / CPU baseline: the reference's own sgl-kernel CPU attention, compiled by
oracle/build_ref.py into oracle/_ref/.  TEST INFRASTRUCTURE -- see oracle/__init__.py.

Call contract (sgl-kernel/csrc/cpu/decode.cpp:1363-1450, extend.cpp:566-667, and the way the
reference's own backend drives them, python/sglang/srt/layers/attention/intel_amx_backend.py:36-45,
test/srt/cpu/test_decode.py:62-163, test/srt/cpu/test_extend.py:78-186):
  * decode ALSO performs the KV write of key/value at `loc` before attending (decode.cpp:1468-1486)
  * attn_logits is a zero fp32 workspace [bs, Hq, num_kv_splits, Dv + 1], num_kv_splits = 8
  * loc, req_pool_indices, seq_lens are int64; req_to_token int32 or int64
  * extend reads the new tokens from k_extend/v_extend and the prefix from the pool
"""
from __future__ import annotations

import importlib.machinery
import importlib.util
from pathlib import Path
from typing import Optional

import torch

_REF_DIR = Path(__file__).resolve().parent / "_ref"
NUM_KV_SPLITS = 8  # intel_amx_backend.py:36-45
_mod = None
_variant = None


def _cpu_flags() -> set:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def pick_variant() -> Optional[Path]:
    flags = _cpu_flags()
    want_avx512 = {"avx512f", "avx512bw", "avx512vl", "avx512dq", "avx512_bf16", "amx_bf16", "amx_tile"} <= flags
    order = ["sgl_cpu_attn_avx512", "sgl_cpu_attn_generic"] if want_avx512 else ["sgl_cpu_attn_generic"]
    for name in order:
        p = _REF_DIR / f"{name}.so"
        if p.is_file():
            return p
    return None


def available() -> bool:
    return pick_variant() is not None


def variant() -> Optional[str]:
    load()
    return _variant


def load():
    global _mod, _variant
    if _mod is None:
        p = pick_variant()
        if p is None:
            raise FileNotFoundError("oracle/_ref/sgl_cpu_attn_*.so not built (python -m oracle.build_ref)")
        name = p.stem
        loader = importlib.machinery.ExtensionFileLoader(name, str(p))
        spec = importlib.util.spec_from_loader(name, loader)
        mod = importlib.util.module_from_spec(spec)
        loader.exec_module(mod)
        _mod, _variant = mod, name
    return _mod


def decode(q: torch.Tensor, k_buffer: torch.Tensor, v_buffer: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
           loc: torch.Tensor, req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, seq_lens: torch.Tensor,
           sm_scale: float, logit_cap: float = 0.0, out: Optional[torch.Tensor] = None,
           attn_logits: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q [bs,Hq,D]; buffers [P,Hkv,D] (written at loc with key/value first); returns [bs,Hq,Dv]."""
    bs, Hq, _ = q.shape
    Dv = v_buffer.shape[2]
    if out is None:
        out = torch.empty((bs, Hq, Dv), dtype=q.dtype)
    if attn_logits is None:
        attn_logits = torch.zeros((bs, Hq, NUM_KV_SPLITS, Dv + 1), dtype=torch.float32)
    load().decode_attention_cpu(q, k_buffer, v_buffer, out, key, value, loc.to(torch.int64), attn_logits,
                                req_to_token, req_pool_indices.to(torch.int64), seq_lens.to(torch.int64),
                                float(sm_scale), float(logit_cap))
    return out


def extend(q_extend: torch.Tensor, k_extend: torch.Tensor, v_extend: torch.Tensor, k_buffer: torch.Tensor,
           v_buffer: torch.Tensor, req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
           seq_lens: torch.Tensor, extend_seq_lens: torch.Tensor, extend_start_loc: torch.Tensor,
           max_len_extend: int, sm_scale: float, logit_cap: float = 0.0,
           out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q_extend [T,Hq,D], k/v_extend [T,Hkv,D]; the prefix (seq_len - extend_len tokens) comes from the pool."""
    if out is None:
        out = torch.empty((q_extend.shape[0], q_extend.shape[1], v_buffer.shape[2]), dtype=q_extend.dtype)
    idx_dt = req_to_token.dtype  # extend.cpp:661-663: extend lens share the table's index dtype
    load().extend_attention_cpu(q_extend, k_extend, v_extend, out, k_buffer, v_buffer, req_to_token,
                                req_pool_indices.to(torch.int64), seq_lens.to(torch.int64),
                                extend_seq_lens.to(idx_dt), extend_start_loc.to(idx_dt), int(max_len_extend),
                                float(sm_scale), float(logit_cap))
    return out
