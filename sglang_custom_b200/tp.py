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


def _gather_device():
    return torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")


def gather_cat(t: torch.Tensor, world: int, dim: int = 1) -> torch.Tensor:
    """Concatenate equally shaped per-rank tensors along `dim` in rank order; the result is a CPU tensor on every
    rank (harness-only collective, used to rebuild a GLOBAL sample for the parity check)."""
    if world == 1:
        return t.cpu()
    x = t.contiguous().to(_gather_device())
    if x.dtype == torch.float8_e4m3fn:
        x = x.view(torch.uint8)
    parts: List[torch.Tensor] = [torch.empty_like(x) for _ in range(world)]
    dist.all_gather(parts, x)
    return torch.cat(parts, dim=dim).cpu()


def gather_subcase(sub, world: int, kv_replicas: int = 1):
    """Per-rank CPU sub-cases (synth.extract_subcase of the same requests on every rank's head slice) -> the GLOBAL
    sub-case: q / k_new / v_new / pools concatenated over heads in rank order (= SGLang's head order,
    models/llama.py:118-131).  With replicated KV heads (world > Hkv) every `kv_replicas`-th rank contributes its KV."""
    if world == 1:
        return sub
    import types
    out = types.SimpleNamespace(**vars(sub))
    out.q = gather_cat(sub.q, world)
    for name in ("k_new", "v_new", "k_pool", "v_pool"):
        full = gather_cat(getattr(sub, name), world)
        if kv_replicas > 1:  # [.., world * 1, D]: keep one copy per KV head
            full = full[:, ::kv_replicas].contiguous()
        setattr(out, name, full)
    out.Hq = out.q.shape[1]
    out.Hkv = out.k_pool.shape[1]
    return out


def all_true(flag: bool, world: int, device="cpu") -> bool:
    if world == 1:
        return bool(flag)
    t = torch.tensor([1.0 if flag else 0.0], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item() > 0.5)


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
