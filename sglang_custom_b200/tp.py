"""Tensor-parallel harness pieces: how the attention path shards over GPUs.

Attention is independent per KV head, so SGLang tensor parallelism gives rank r the KV heads
[r*Hkv/tp, (r+1)*Hkv/tp) and the matching Q heads (python/sglang/srt/models/llama.py:118-131,
configs/model_config.py:342-349); every rank owns its own MHATokenToKVPool slice and an identical
replica of req_to_token.  There is NO collective inside the path -- the only collectives here are the
harness's own: a barrier around the timed region, a MAX-reduce of the per-rank device time, and an
all-gather of the outputs for the concatenated parity check against the global oracle.
Works with backend "nccl" (GPU box) and "gloo" (CPU tests, world_size 2).
"""
from __future__ import annotations

import os
from typing import List, Optional

import torch
import torch.distributed as dist

from .synth import slice_heads


def env_rank_world() -> tuple:
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def init(backend: Optional[str] = None) -> tuple:
    """torchrun-style rendezvous from RANK/WORLD_SIZE/MASTER_*; returns (rank, world, local_rank)."""
    rank, world, local = env_rank_world()
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29517")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, world, local


def shard_case(case, rank: int, world: int):
    """Rank r's slice of a global case.  When world > Hkv the KV heads are replicated
    (model_config.py:346-349): ranks [g*rep, (g+1)*rep) share KV head g and split its Q heads."""
    if world == 1:
        return case
    if case.Hkv % world == 0:
        return slice_heads(case, rank, world)
    assert world % case.Hkv == 0 and case.Hq % world == 0, "tp must divide or be a multiple of the KV head count"
    import types
    rep = world // case.Hkv
    kv = rank // rep
    hq = case.Hq // world
    out = types.SimpleNamespace(**vars(case))
    out.Hq, out.Hkv = hq, 1
    out.q = case.q[:, rank * hq:(rank + 1) * hq].contiguous()
    for name in ("k_new", "v_new", "k_pool", "v_pool"):
        setattr(out, name, getattr(case, name)[:, kv:kv + 1].contiguous())
    return out


def gather_heads(out_local: torch.Tensor, world: int) -> torch.Tensor:
    """[T, Hq/world, D] per rank -> [T, Hq, D] on every rank (harness-only collective)."""
    if world == 1:
        return out_local
    parts: List[torch.Tensor] = [torch.empty_like(out_local) for _ in range(world)]
    dist.all_gather(parts, out_local.contiguous())
    return torch.cat(parts, dim=1)


def barrier(world: int) -> None:
    if world > 1:
        dist.barrier()


def max_over_ranks(value: float, world: int, device="cpu") -> float:
    """Device-timed milliseconds -> the slowest rank's (multi-GPU numbers are max over ranks)."""
    if world == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value: float, world: int, device="cpu") -> float:
    if world == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())
