"""Run the UNMODIFIED reference Triton attention kernels on the CPU through Triton's interpreter.

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Only works where /root/reference exists (the build
container).  It pins oracle/triton_masked.py (sliding window, speculative-decoding custom mask), for which the
reference's two CPU paths (torch_native, sgl-kernel CPU) have no argument: those semantics live only in
  python/sglang/srt/layers/attention/triton_ops/extend_attention.py:42-438   (_fwd_kernel, extend_attention_fwd)
  python/sglang/srt/layers/attention/triton_ops/decode_attention.py:677-728  (decode_attention_fwd)
  python/sglang/srt/layers/attention/triton_backend.py:927-955               (update_sliding_window_buffer)
  python/sglang/srt/layers/attention/utils.py:5-41                           (create_flashinfer_kv_indices_triton)
TRITON_INTERPRET=1 makes `@triton.jit` execute the kernel source with numpy, so the files above run here without a
GPU.  The variable is read when the decorator runs: this module must be imported in a process that has not imported
triton yet (tests/golden/make_golden_masked.py and the live pin test run it in a subprocess).

Interpreter limits found by probing: bf16 tensors are mishandled (numpy has no bf16; results are garbage), fp16 and
fp32 are exact -- so the fixtures are fp16 / fp32.
"""
from __future__ import annotations

import os
import sys
import types

if "triton" in sys.modules and os.environ.get("TRITON_INTERPRET") != "1":
    raise RuntimeError("oracle.ref_triton must be imported before triton, in a process with TRITON_INTERPRET=1")
os.environ["TRITON_INTERPRET"] = "1"

import torch  # noqa: E402

from . import ref_loader  # noqa: E402
from .ref_loader import _SRT, _load, _stub  # noqa: E402

_cache = {}


def available() -> bool:
    return os.path.isfile(os.path.join(_SRT, "layers", "attention", "triton_ops", "extend_attention.py"))


def load():
    """-> namespace(ext=extend_attention module, dec=decode_attention module, tb=triton_backend module, utils=...)"""
    if "ns" in _cache:
        return _cache["ns"]
    base = ref_loader.load()  # registers the stub packages and AttentionType
    att = os.path.join(_SRT, "layers", "attention")
    _stub("sglang.srt.layers.attention.triton_ops")
    _stub("sglang.srt.layers.attention.triton_ops.prefill_attention", context_attention_fwd=None)
    _stub("sglang.srt.layers.dp_attention", get_attention_tp_size=lambda: 1)
    _stub("sglang.srt.utils", is_cuda=lambda: False, is_hip=lambda: False, get_bool_env_var=lambda *a, **k: False,
          get_device_core_count=lambda *a, **k: 148,
          next_power_of_2=lambda n: 1 << (max(int(n), 1) - 1).bit_length())
    utils = _load("sglang.srt.layers.attention.utils", os.path.join(att, "utils.py"))
    ext = _load("sglang.srt.layers.attention.triton_ops.extend_attention",
                os.path.join(att, "triton_ops", "extend_attention.py"))
    dec = _load("sglang.srt.layers.attention.triton_ops.decode_attention",
                os.path.join(att, "triton_ops", "decode_attention.py"))
    tb = _load("sglang.srt.layers.attention.triton_backend", os.path.join(att, "triton_backend.py"))
    ns = types.SimpleNamespace(ext=ext, dec=dec, tb=tb, utils=utils, base=base)
    _cache["ns"] = ns
    return ns


def _flatten(ns, req_to_token, req_pool_indices, lens):
    """kv_indptr / kv_indices as TritonAttnBackend.init_forward_metadata builds them (triton_backend.py:171-186)."""
    bs = lens.numel()
    kv_indptr = torch.zeros(bs + 1, dtype=torch.int32)
    kv_indptr[1:] = torch.cumsum(lens, 0)
    kv_indices = torch.zeros(max(int(kv_indptr[-1]), 1), dtype=torch.int32)
    ns.utils.create_flashinfer_kv_indices_triton[(bs,)](req_to_token, req_pool_indices, lens, kv_indptr, None,
                                                        kv_indices, req_to_token.stride(0))
    return kv_indptr, kv_indices


def ref_extend(q, k_new, v_new, k_pool, v_pool, req_to_token, req_pool_indices, prefix_lens, extend_lens, scaling,
               causal=True, sliding_window_size=-1, custom_mask=None, mask_indptr=None, skip_prefix_custom_mask=True,
               logit_cap=0.0):
    """TritonAttnBackend.forward_extend's call (triton_backend.py:658-684) with the metadata of
    init_forward_metadata's EXTEND branch (:283-322): prefix keys through kv_indices (or window_kv_indices), new
    keys from the contiguous k_new / v_new."""
    ns = load()
    bs = prefix_lens.numel()
    prefix_lens = prefix_lens.to(torch.int32)
    extend_lens = extend_lens.to(torch.int32)
    rpi = req_pool_indices.to(torch.int32)
    qo_indptr = torch.zeros(bs + 1, dtype=torch.int32)
    qo_indptr[1:] = torch.cumsum(extend_lens, 0)
    if sliding_window_size is not None and sliding_window_size > -1:
        window_kv_indptr = torch.zeros(bs + 1, dtype=torch.int32)
        kv_indptr, kv_indices, _ = ns.tb.update_sliding_window_buffer(
            window_kv_indptr, req_to_token, sliding_window_size, prefix_lens, rpi, bs, "cpu")
        if kv_indices.numel() == 0:
            kv_indices = torch.zeros(1, dtype=torch.int32)
    else:
        sliding_window_size = -1
        kv_indptr, kv_indices = _flatten(ns, req_to_token, rpi, prefix_lens)
    out = torch.zeros_like(q)
    ns.ext.extend_attention_fwd(q, k_new.contiguous(), v_new.contiguous(), out, k_pool, v_pool, qo_indptr, kv_indptr,
                                kv_indices, custom_mask, causal, mask_indptr, int(extend_lens.max()), scaling,
                                logit_cap, skip_prefix_custom_mask, sliding_window_size)
    return out


def ref_decode(q, k_pool, v_pool, req_to_token, req_pool_indices, seq_lens, scaling, sliding_window_size=-1,
               logit_cap=0.0, max_kv_splits=8):
    """TritonAttnBackend.forward_decode's call (triton_backend.py:711-731) with the metadata of
    init_forward_metadata's DECODE branch (:170-226), sliding-window layers included."""
    ns = load()
    bs, Hq, D = q.shape
    seq_lens = seq_lens.to(torch.int32)
    rpi = req_pool_indices.to(torch.int32)
    if sliding_window_size is not None and sliding_window_size > -1:
        window_kv_indptr = torch.zeros(bs + 1, dtype=torch.int32)
        kv_indptr, kv_indices, lens = ns.tb.update_sliding_window_buffer(
            window_kv_indptr, req_to_token, sliding_window_size, seq_lens, rpi, bs, "cpu")
    else:
        kv_indptr, kv_indices = _flatten(ns, req_to_token, rpi, seq_lens)
        lens = seq_lens
    # any split count gives the same result up to rounding; a fixed even split keeps the run deterministic
    num_kv_splits = torch.clamp((lens.to(torch.int32) + 63) // 64, 1, max_kv_splits).to(torch.int32)
    attn_logits = torch.zeros((bs, Hq, max_kv_splits, v_pool.shape[-1]), dtype=torch.float32)
    attn_lse = torch.zeros((bs, Hq, max_kv_splits), dtype=torch.float32)
    out = torch.zeros_like(q)
    ns.dec.decode_attention_fwd(q, k_pool, v_pool, out, kv_indptr, kv_indices, attn_logits, attn_lse, num_kv_splits,
                                max_kv_splits, scaling, logit_cap)
    return out
