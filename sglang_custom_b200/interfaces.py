"""Stand-ins for the SGLang objects the attention backend touches, so the backend runs (and is
tested) without an SGLang install.  Attribute names and meanings mirror the reference exactly; in
a real deployment these are SGLang's own objects and this module is not imported.

Mirrored reference types (python/sglang/srt/...):
  ForwardMode        model_executor/forward_batch_info.py:60-130
  ForwardBatch       model_executor/forward_batch_info.py:160-300   (only the fields a backend reads)
  AttentionType      layers/radix_attention.py:25-34
  RadixAttention     layers/radix_attention.py:37-108               (descriptor + forward dispatch)
  ReqToTokenPool     mem_cache/memory_pool.py:51-110
  MHATokenToKVPool   mem_cache/memory_pool.py:184-458
  ModelRunner        model_executor/model_runner.py (attributes read by TritonAttnBackend.__init__,
                     layers/attention/triton_backend.py:41-122)
"""
from __future__ import annotations

import enum
from dataclasses import dataclass, field
from typing import List, Optional

import torch


class ForwardMode(enum.IntEnum):
    EXTEND = enum.auto()
    DECODE = enum.auto()
    MIXED = enum.auto()
    IDLE = enum.auto()
    SPLIT_PREFILL = enum.auto()
    TARGET_VERIFY = enum.auto()
    DRAFT_EXTEND = enum.auto()
    DUMMY_FIRST = enum.auto()

    def is_extend(self) -> bool:
        return self in (ForwardMode.EXTEND, ForwardMode.MIXED, ForwardMode.DRAFT_EXTEND, ForwardMode.TARGET_VERIFY)

    is_prefill = is_extend

    def is_decode(self) -> bool:
        return self == ForwardMode.DECODE

    def is_mixed(self) -> bool:
        return self == ForwardMode.MIXED

    def is_idle(self) -> bool:
        return self == ForwardMode.IDLE

    def is_decode_or_idle(self) -> bool:
        return self in (ForwardMode.DECODE, ForwardMode.IDLE)

    def is_target_verify(self) -> bool:
        return self == ForwardMode.TARGET_VERIFY

    def is_draft_extend(self) -> bool:
        return self == ForwardMode.DRAFT_EXTEND

    def is_cuda_graph(self) -> bool:
        return self in (ForwardMode.DECODE, ForwardMode.TARGET_VERIFY, ForwardMode.IDLE)


class AttentionType(enum.Enum):
    DECODER = "decoder"
    ENCODER_ONLY = "encoder_only"


class ReqToTokenPool:
    """The page table: req_to_token[r, j] = pool slot of token j of the request in req-slot r."""

    def __init__(self, size: int, max_context_len: int, device: str):
        self.size = size
        self.max_context_len = max_context_len
        self.device = device
        self.req_to_token = torch.zeros((size, max_context_len), dtype=torch.int32, device=device)
        self.free_slots = list(range(size))

    def write(self, indices, values):
        self.req_to_token[indices] = values

    def alloc(self, need_size: int) -> Optional[List[int]]:
        if need_size > len(self.free_slots):
            return None
        out, self.free_slots = self.free_slots[:need_size], self.free_slots[need_size:]
        return out

    def free(self, idx):
        self.free_slots.extend([idx] if isinstance(idx, int) else list(idx))


class MHATokenToKVPool:
    """[size + page_size, head_num, head_dim] K and V buffers per layer; fp8 stored as uint8.

    set_kv_buffer keeps the reference signature (memory_pool.py:420-428) and runs the
    b200attn KV-write kernel; there is no torch fallback."""

    def __init__(self, size: int, page_size: int, dtype: torch.dtype, head_num: int, head_dim: int,
                 layer_num: int, device: str, start_layer: int = 0):
        self.size, self.page_size, self.dtype = size, page_size, dtype
        self.head_num, self.head_dim, self.layer_num = head_num, head_dim, layer_num
        self.device, self.start_layer = device, start_layer
        self.store_dtype = torch.uint8 if dtype in (torch.float8_e4m3fn, torch.float8_e5m2) else dtype
        self.layer_transfer_counter = None
        shape = (size + page_size, head_num, head_dim)
        self.k_buffer = [torch.zeros(shape, dtype=self.store_dtype, device=device) for _ in range(layer_num)]
        self.v_buffer = [torch.zeros(shape, dtype=self.store_dtype, device=device) for _ in range(layer_num)]
        # memory_pool.py:256-268: base address and row size (bytes) of every K then every V buffer, on the device
        bufs = self.k_buffer + self.v_buffer
        self.token_stride = head_num * head_dim
        self.data_ptrs = torch.tensor([x.data_ptr() for x in bufs], dtype=torch.uint64, device=device)
        self._row_bytes = [x[0].numel() * x.element_size() for x in bufs]
        self.data_strides = torch.tensor(self._row_bytes, dtype=torch.int64, device=device)

    def move_kv_cache(self, tgt_loc: torch.Tensor, src_loc: torch.Tensor) -> None:
        """memory_pool.py:460-468: all layers, K and V, in one launch (b200attn_move_kv); no torch fallback."""
        from . import ops
        ops.move_kv(self.data_ptrs, self.data_strides, max(self._row_bytes), tgt_loc, src_loc)

    def _view(self, buf: torch.Tensor) -> torch.Tensor:
        return buf.view(self.dtype) if self.store_dtype != self.dtype else buf

    def get_key_buffer(self, layer_id: int) -> torch.Tensor:
        if self.layer_transfer_counter is not None:
            self.layer_transfer_counter.wait_until(layer_id - self.start_layer)
        return self._view(self.k_buffer[layer_id - self.start_layer])

    def get_value_buffer(self, layer_id: int) -> torch.Tensor:
        if self.layer_transfer_counter is not None:
            self.layer_transfer_counter.wait_until(layer_id - self.start_layer)
        return self._view(self.v_buffer[layer_id - self.start_layer])

    def get_kv_buffer(self, layer_id: int):
        return self.get_key_buffer(layer_id), self.get_value_buffer(layer_id)

    def set_kv_buffer(self, layer, loc: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor,
                      k_scale=None, v_scale=None, layer_id_override: Optional[int] = None) -> None:
        from . import ops
        layer_id = layer.layer_id if layer_id_override is None else layer_id_override
        ops.set_kv_buffer(self.k_buffer[layer_id - self.start_layer], self.v_buffer[layer_id - self.start_layer],
                          loc, cache_k, cache_v, self.dtype, k_scale, v_scale)


@dataclass
class RadixAttention:
    """Per-layer attention descriptor (layers/radix_attention.py:40-80)."""
    num_heads: int
    head_dim: int
    scaling: float
    num_kv_heads: int
    layer_id: int
    logit_cap: float = 0.0
    v_head_dim: int = -1
    sliding_window_size: int = -1
    is_cross_attention: bool = False
    attn_type: AttentionType = AttentionType.DECODER
    k_scale: Optional[torch.Tensor] = None
    v_scale: Optional[torch.Tensor] = None
    k_scale_float: Optional[float] = None
    v_scale_float: Optional[float] = None

    def __post_init__(self):
        self.tp_q_head_num = self.num_heads
        self.tp_k_head_num = self.num_kv_heads
        self.tp_v_head_num = self.num_kv_heads
        self.qk_head_dim = self.head_dim
        self.v_head_dim = self.v_head_dim if self.v_head_dim != -1 else self.head_dim

    def forward(self, q, k, v, forward_batch, save_kv_cache: bool = True, **kwargs):
        if k is not None:  # radix_attention.py:91-99
            k = k.view(-1, self.tp_k_head_num, self.qk_head_dim)
            v = v.view(-1, self.tp_v_head_num, self.v_head_dim)
        return forward_batch.attn_backend.forward(q, k, v, self, forward_batch, save_kv_cache, **kwargs)


@dataclass
class ForwardBatch:
    forward_mode: ForwardMode
    batch_size: int
    req_pool_indices: torch.Tensor
    seq_lens: torch.Tensor
    out_cache_loc: torch.Tensor
    seq_lens_sum: int
    req_to_token_pool: ReqToTokenPool = None
    token_to_kv_pool: MHATokenToKVPool = None
    attn_backend: object = None
    seq_lens_cpu: Optional[torch.Tensor] = None
    extend_seq_lens: Optional[torch.Tensor] = None
    extend_prefix_lens: Optional[torch.Tensor] = None
    extend_start_loc: Optional[torch.Tensor] = None
    extend_seq_lens_cpu: Optional[List[int]] = None
    extend_prefix_lens_cpu: Optional[List[int]] = None
    extend_num_tokens: Optional[int] = None
    encoder_lens: Optional[torch.Tensor] = None
    spec_info: object = None
    positions: Optional[torch.Tensor] = None


@dataclass
class ModelConfig:
    context_len: int
    num_attention_heads: int
    num_key_value_heads: int
    head_dim: int
    num_hidden_layers: int = 1
    is_encoder_decoder: bool = False

    def get_num_kv_heads(self, tp_size: int) -> int:
        return max(1, self.num_key_value_heads // tp_size)  # configs/model_config.py:342-349


@dataclass
class ServerArgs:
    kv_cache_dtype: str = "auto"
    page_size: int = 1
    disable_cuda_graph: bool = False
    speculative_num_draft_tokens: Optional[int] = None
    speculative_num_steps: Optional[int] = None
    triton_attention_num_kv_splits: int = 8
    attention_backend: str = "b200"


class ModelRunner:
    """The slice of ModelRunner an attention backend reads (triton_backend.py:41-122)."""

    def __init__(self, model_config: ModelConfig, max_reqs: int, pool_size: int, device: str = "cuda",
                 kv_dtype: torch.dtype = torch.bfloat16, page_size: int = 1, tp_size: int = 1, tp_rank: int = 0,
                 gpu_id: int = 0, dtype: torch.dtype = torch.bfloat16):
        self.model_config = model_config
        self.device = device
        self.gpu_id = gpu_id
        self.dtype = dtype
        self.kv_cache_dtype = kv_dtype
        self.page_size = page_size
        self.tp_size, self.tp_rank = tp_size, tp_rank
        self.sliding_window_size = None
        self.server_args = ServerArgs(page_size=page_size)
        self.req_to_token_pool = ReqToTokenPool(max_reqs, model_config.context_len, device)
        self.token_to_kv_pool = MHATokenToKVPool(
            pool_size, page_size, kv_dtype, model_config.get_num_kv_heads(tp_size), model_config.head_dim,
            model_config.num_hidden_layers, device)
        self.attn_backend = None
