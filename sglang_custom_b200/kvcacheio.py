"""HiCache KV page transfers on the b200attn transfer kernel.

Same function names, argument order and meaning as the reference's `sgl_kernel.kvcacheio`
(sgl-kernel/python/sgl_kernel/kvcacheio.py:6-237 -> sgl-kernel/csrc/kvcacheio/transfer.cu:177-430), so
`mem_cache/memory_pool_host.py` and `memory_pool.py:350-389` can import this module in its place:

  transfer_kv_per_layer(src_k, dst_k, src_v, dst_v, src_indices, dst_indices, io_backend, page_size, item_size, ...)
  transfer_kv_per_layer_pf_lf(...)      page-first source (host) -> layer-first destination (device), one layer
  transfer_kv_all_layer(...)            per-layer pointer tables on both sides
  transfer_kv_all_layer_lf_pf(...)      pointer table (device) -> page-first destination (host)
  transfer_kv_per_layer_mla / _mla_pf_lf / transfer_kv_all_layer_mla / _mla_lf_pf    single buffer (MLA pools)
  transfer_kv_direct(...)               the reference's "direct" backend: per-page tensor.copy_ (cudaMemcpy), no kernel

As in the reference, `item_size` of the `io_backend="kernel"` per-layer entry points is in ELEMENTS (it is multiplied
by the element size, kvcacheio.py:25,157) while every other entry point takes BYTES.  Host-side tensors must be pinned
(the kernel dereferences them through the mapped address, as the reference does).  `num_warps_per_block` is accepted
for signature compatibility and ignored: the kernel flattens the job into 16-byte chunks instead of giving a warp
an item.
"""
from __future__ import annotations

from typing import List

import torch

from . import _lib
from ._lib import B200AttnError, check

LF, PF, TBL = _lib.LAYOUT_LAYER_FIRST, _lib.LAYOUT_PAGE_FIRST, _lib.LAYOUT_LAYER_TABLE


def _idx(t: torch.Tensor, name: str) -> torch.Tensor:
    # reference checks, transfer.cu:137-141
    if not t.is_cuda:
        raise B200AttnError(f"{name} must be a CUDA tensor")
    if t.dtype != torch.int64:
        raise B200AttnError(f"{name} must be of type long")
    return t.contiguous()


def _stream(idx: torch.Tensor) -> int:
    return torch.cuda.current_stream(idx.device).cuda_stream


def _launch(src_k, dst_k, src_v, dst_v, src_layout, dst_layout, src_dim, dst_dim, src_indices, dst_indices,
            item_bytes, num_layers, block_quota) -> None:
    src_indices, dst_indices = _idx(src_indices, "Source indices"), _idx(dst_indices, "Destination indices")
    if src_indices.numel() != dst_indices.numel():
        raise B200AttnError("Source and destination indices must have the same length")
    ptr = lambda t: 0 if t is None else t.data_ptr()
    check(_lib.load().b200attn_transfer_kv(
        ptr(src_k), ptr(dst_k), ptr(src_v), ptr(dst_v), src_layout, dst_layout, int(src_dim), int(dst_dim),
        src_indices.data_ptr(), dst_indices.data_ptr(), src_indices.numel(), int(item_bytes), 0, int(num_layers),
        int(block_quota), _stream(src_indices)), "b200attn_transfer_kv")


def _table(t: torch.Tensor, num_layers: int, what: str) -> torch.Tensor:
    if not t.is_cuda or t.dtype not in (torch.uint64, torch.int64) or not t.is_contiguous():
        raise B200AttnError(f"{what} must be a contiguous CUDA uint64 tensor of per-layer base pointers")
    if t.shape[0] != num_layers:
        raise B200AttnError(f"Number of layers in {what} does not match num_layers")
    return t


def transfer_kv_per_layer(src_k, dst_k, src_v, dst_v, src_indices, dst_indices, io_backend: str, page_size: int,
                          item_size: int, block_quota: int = 2, num_warps_per_block: int = 32):
    if io_backend == "kernel":
        _launch(src_k, dst_k, src_v, dst_v, LF, LF, 0, 0, src_indices, dst_indices,
                item_size * src_k.element_size(), 1, block_quota)
    elif io_backend == "direct":
        transfer_kv_direct([src_k, src_v], [dst_k, dst_v], src_indices, dst_indices, page_size)
    else:
        raise ValueError("Unsupported io backend")


def transfer_kv_per_layer_pf_lf(src_k, dst_k, src_v, dst_v, src_indices, dst_indices, item_size: int,
                                src_layout_dim: int, block_quota: int = 2, num_warps_per_block: int = 32):
    _launch(src_k, dst_k, src_v, dst_v, PF, LF, src_layout_dim, 0, src_indices, dst_indices, item_size, 1, block_quota)


def transfer_kv_all_layer(src_k_layers, dst_k_layers, src_v_layers, dst_v_layers, src_indices, dst_indices,
                          io_backend: str, item_size: int, num_layers: int, block_quota: int = 2,
                          num_warps_per_block: int = 32):
    if io_backend == "kernel":
        for t, w in ((src_k_layers, "source k tensor"), (dst_k_layers, "destination k tensor"),
                     (src_v_layers, "source v tensor"), (dst_v_layers, "destination v tensor")):
            _table(t, num_layers, w)
        _launch(src_k_layers, dst_k_layers, src_v_layers, dst_v_layers, TBL, TBL, 0, 0, src_indices, dst_indices,
                item_size, num_layers, block_quota)
    elif io_backend == "direct":
        raise NotImplementedError("Deprecated interface")
    else:
        raise ValueError("Unsupported io backend")


def transfer_kv_all_layer_lf_pf(src_k_layers, dst_k, src_v_layers, dst_v, src_indices, dst_indices, item_size: int,
                                dst_layout_dim: int, num_layers: int, block_quota: int = 2,
                                num_warps_per_block: int = 32):
    _table(src_k_layers, num_layers, "source k tensor")
    _table(src_v_layers, num_layers, "source v tensor")
    _launch(src_k_layers, dst_k, src_v_layers, dst_v, TBL, PF, 0, dst_layout_dim, src_indices, dst_indices, item_size,
            num_layers, block_quota)


def transfer_kv_direct(src_layers: List[torch.Tensor], dst_layers: List[torch.Tensor], src_indices: torch.Tensor,
                       dst_indices: torch.Tensor, page_size: int):
    """The reference's copy-engine path (transfer.cu:432-465): one `copy_` per (page, layer); indices are read on the
    host.  No kernel of ours is involved -- this is cudaMemcpyAsync plumbing, kept for interface completeness."""
    if len(src_layers) != len(dst_layers):
        raise B200AttnError("Source and destination layers must have the same number of layers")
    if src_indices.numel() != dst_indices.numel():
        raise B200AttnError("Source and destination indices must have the same length")
    if page_size <= 0:
        raise B200AttnError("Page size must be positive")
    if src_indices.numel() % page_size != 0:
        raise B200AttnError("Source indices size must be divisible by page size")
    s_cpu, d_cpu = src_indices.cpu().tolist(), dst_indices.cpu().tolist()
    for i in range(0, len(s_cpu), page_size):
        s, d = s_cpu[i], d_cpu[i]
        for src, dst in zip(src_layers, dst_layers):
            dst[d:d + page_size].copy_(src[s:s + page_size], non_blocking=True)


def transfer_kv_per_layer_mla(src, dst, src_indices, dst_indices, io_backend: str, page_size: int, item_size: int,
                              block_quota: int = 2, num_warps_per_block: int = 32):
    if io_backend == "kernel":
        _launch(src, dst, None, None, LF, LF, 0, 0, src_indices, dst_indices, item_size * src.element_size(), 1,
                block_quota)
    elif io_backend == "direct":
        transfer_kv_direct([src], [dst], src_indices, dst_indices, page_size)
    else:
        raise ValueError("Unsupported io backend")


def transfer_kv_per_layer_mla_pf_lf(src, dst, src_indices, dst_indices, item_size: int, src_layout_dim: int,
                                    block_quota: int = 2, num_warps_per_block: int = 32):
    _launch(src, dst, None, None, PF, LF, src_layout_dim, 0, src_indices, dst_indices, item_size, 1, block_quota)


def transfer_kv_all_layer_mla(src_layers, dst_layers, src_indices, dst_indices, io_backend: str, item_size: int,
                              num_layers: int, block_quota: int = 2, num_warps_per_block: int = 32):
    if io_backend == "kernel":
        _table(src_layers, num_layers, "source tensor")
        _table(dst_layers, num_layers, "destination tensor")
        _launch(src_layers, dst_layers, None, None, TBL, TBL, 0, 0, src_indices, dst_indices, item_size, num_layers,
                block_quota)
    elif io_backend == "direct":
        raise NotImplementedError("Deprecated interface")
    else:
        raise ValueError("Unsupported io backend")


def transfer_kv_all_layer_mla_lf_pf(src_layers, dst, src_indices, dst_indices, item_size: int, dst_layout_dim: int,
                                    num_layers: int, block_quota: int = 2, num_warps_per_block: int = 32):
    _table(src_layers, num_layers, "source tensor")
    _launch(src_layers, dst, None, None, TBL, PF, 0, dst_layout_dim, src_indices, dst_indices, item_size, num_layers,
            block_quota)
