"""Load the UNMODIFIED reference attention files from /root/reference under stub modules.

TEST INFRASTRUCTURE -- see oracle/__init__.py.  Only works where /root/reference exists (the
build container); the GPU box never calls this.  It is how oracle/torch_native.py is pinned and
how tests/golden/ fixtures are generated (tests/golden/make_golden.py).

`import sglang` as a package fails here (IPython, sgl_kernel and a torch-private symbol are
missing -- SURVEY.md section 8c), but the files on the attention path only need torch:
  python/sglang/srt/layers/attention/base_attn_backend.py
  python/sglang/srt/layers/attention/torch_native_backend.py
  python/sglang/srt/mem_cache/memory_pool.py            (MHATokenToKVPool.set_kv_buffer)
so their other imports are replaced by empty stand-ins registered in sys.modules.
"""
from __future__ import annotations

import enum
import importlib.util
import os
import sys
import types

REF_ROOT = os.environ.get("B200ATTN_REFERENCE", "/root/reference")
_SRT = os.path.join(REF_ROOT, "python", "sglang", "srt")


def available() -> bool:
    return os.path.isfile(os.path.join(_SRT, "layers", "attention", "torch_native_backend.py"))


def _stub(name: str, **attrs) -> types.ModuleType:
    m = sys.modules.get(name)
    if m is None:
        m = types.ModuleType(name)
        m.__path__ = []  # behave as a package so dotted children resolve
        sys.modules[name] = m
    for k, v in attrs.items():
        setattr(m, k, v)
    return m


def _load(name: str, path: str) -> types.ModuleType:
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


_cache = {}


def load():
    """-> namespace(backend=TorchNativeAttnBackend instance, tn=module, MHATokenToKVPool=class)"""
    if "ns" in _cache:
        return _cache["ns"]
    if not available():
        raise FileNotFoundError(f"{_SRT} not found: the reference is only mounted in the build container")
    if "sglang" in sys.modules and not getattr(sys.modules["sglang"], "__b200attn_stub__", False):
        raise RuntimeError("a real `sglang` package is already imported; refusing to shadow it")

    for n in ("sglang", "sglang.srt", "sglang.srt.layers", "sglang.srt.layers.attention",
              "sglang.srt.model_executor", "sglang.srt.mem_cache", "sgl_kernel"):
        _stub(n).__b200attn_stub__ = True

    class AttentionType(enum.Enum):  # values as in layers/radix_attention.py:25-34
        DECODER = "decoder"
        ENCODER_ONLY = "encoder_only"

    _stub("sglang.srt.layers.radix_attention", AttentionType=AttentionType, RadixAttention=object)
    _stub("sglang.srt.model_executor.forward_batch_info", ForwardBatch=object, ForwardMode=object)
    _stub("sglang.srt.model_executor.cuda_graph_runner", get_is_capture_mode=lambda: False)
    class _NoMemorySaver:  # stand-in for torch_memory_saver_adapter.TorchMemorySaverAdapter (an optional dependency):
        @staticmethod      # create(enable) -> object whose region(tag) is a no-op context manager
        def create(enable=False):
            return _NoMemorySaver()

        def region(self, *a, **k):
            import contextlib
            return contextlib.nullcontext()

    _stub("sglang.srt.torch_memory_saver_adapter", TorchMemorySaverAdapter=_NoMemorySaver)
    _stub("sglang.srt.constants", GPU_MEMORY_TYPE_KV_CACHE="kv_cache")
    _stub("sglang.srt.utils", get_bool_env_var=lambda *a, **k: False, is_cuda=lambda: False,
          is_npu=lambda: False, next_power_of_2=lambda n: 1 << (max(int(n), 1) - 1).bit_length())
    _stub("sgl_kernel.kvcacheio", transfer_kv_per_layer=None, transfer_kv_per_layer_mla=None)

    _load("sglang.srt.layers.attention.base_attn_backend",
          os.path.join(_SRT, "layers", "attention", "base_attn_backend.py"))
    tn = _load("sglang.srt.layers.attention.torch_native_backend",
               os.path.join(_SRT, "layers", "attention", "torch_native_backend.py"))
    mp = _load("sglang.srt.mem_cache.memory_pool", os.path.join(_SRT, "mem_cache", "memory_pool.py"))

    ns = types.SimpleNamespace(
        tn=tn, backend=tn.TorchNativeAttnBackend(types.SimpleNamespace(device="cpu")),
        MHATokenToKVPool=mp.MHATokenToKVPool, AttentionType=AttentionType)
    _cache["ns"] = ns
    return ns


def ref_set_kv_buffer(k_buffer, v_buffer, loc, cache_k, cache_v, pool_dtype, k_scale=None, v_scale=None):
    """Run the reference's own MHATokenToKVPool.set_kv_buffer (memory_pool.py:420-458) on a
    one-layer stand-in for the pool object (only the attributes that method reads)."""
    ns = load()
    pool = types.SimpleNamespace(dtype=pool_dtype, store_dtype=k_buffer.dtype, k_buffer=[k_buffer],
                                 v_buffer=[v_buffer], start_layer=0, alt_stream=None, device_module=None)
    layer = types.SimpleNamespace(layer_id=0)
    ns.MHATokenToKVPool.set_kv_buffer(pool, layer, loc, cache_k, cache_v, k_scale, v_scale)


def ref_decode(q, k_cache, v_cache, req_to_token, req_pool_indices, seq_lens, scaling):
    """The reference's _run_sdpa_forward_decode (torch_native_backend.py:112-180)."""
    import torch
    ns = load()
    out = torch.empty((q.shape[0], q.shape[1], v_cache.shape[2]), dtype=q.dtype)
    ns.backend._run_sdpa_forward_decode(q, out, k_cache, v_cache, req_to_token, req_pool_indices, seq_lens,
                                        scaling=scaling, enable_gqa=q.shape[1] != k_cache.shape[1],
                                        causal=False)
    return out


def ref_extend(q, k_cache, v_cache, req_to_token, req_pool_indices, seq_lens, extend_prefix_lens,
               extend_seq_lens, scaling, causal=True):
    """The reference's _run_sdpa_forward_extend (torch_native_backend.py:27-110)."""
    import torch
    ns = load()
    out = torch.empty((q.shape[0], q.shape[1], v_cache.shape[2]), dtype=q.dtype)
    ns.backend._run_sdpa_forward_extend(q, out, k_cache, v_cache, req_to_token, req_pool_indices, seq_lens,
                                        extend_prefix_lens, extend_seq_lens, scaling=scaling,
                                        enable_gqa=q.shape[1] != k_cache.shape[1], causal=causal)
    return out


def ref_rotary_embedding(head_size: int, rotary_dim: int, max_position: int, base, is_neox: bool, dtype,
                         rope_scaling=None):
    """The reference's own rotary embedding module (layers/rotary_embedding.py, `get_rope` :1385-) built for the CPU:
    its `forward_native(positions, query, key)` is what oracle/rope.py restates."""
    import torch.nn as nn
    load()
    name = "sglang.srt.layers.rotary_embedding"
    if name not in sys.modules or not hasattr(sys.modules[name], "get_rope"):
        class CustomOp(nn.Module):  # sglang/srt/custom_op.py: dispatches forward() to a backend method
            def forward(self, *a, **k):
                return self.forward_native(*a, **k)
        _stub("sglang.srt.custom_op", CustomOp=CustomOp)
        _stub("sglang.srt.utils", cpu_has_amx_support=lambda: True, is_cpu=lambda: True, is_hip=lambda: False,
              is_cuda=lambda: False, is_npu=lambda: False, get_bool_env_var=lambda *a, **k: False)
        _load(name, os.path.join(_SRT, "layers", "rotary_embedding.py"))
    mod = sys.modules[name]
    if rope_scaling is None:
        return mod.RotaryEmbedding(head_size, rotary_dim, max_position, base, is_neox, dtype)
    return mod.get_rope(head_size, rotary_dim, max_position, base, is_neox, rope_scaling=dict(rope_scaling), dtype=dtype)
