"""Seeded synthetic workloads for the attention path.

Everything is generated on the CPU in bf16 (the identical bytes are then given to the CUDA
path and to the CPU oracle).  No oracle code is imported here.

A case is a plain namespace of CPU tensors:
  q [T,Hq,D], k_new/v_new [T,Hkv,D]                 inputs of one forward (T = bs for decode)
  k_pool/v_pool [P,Hkv,D]                           pool state BEFORE the forward's KV write (bf16, or uint8
                                                    e4m3 when fp8); rows [0,page_size) are the dummy page
  req_to_token [max_reqs, ctx] int32                page table (already holds the new tokens' slots)
  req_pool_indices [bs], seq_lens [bs] (int64)      seq_lens INCLUDE the new tokens
  out_cache_loc [T] int64                           slots of the new tokens
  extend_seq_lens / extend_prefix_lens / extend_start_loc [bs] int32 (extend cases)
"""
from __future__ import annotations

import types
from typing import Optional, Sequence

import torch

BASELINE_SEED = 1234  # seed = 1234 + config id, as test/srt/cpu/test_decode.py:62 uses 1234


def _rand_bf16(gen: torch.Generator, shape, chunk_rows: int = 1 << 16) -> torch.Tensor:
    """N(0,1) in bf16 without materialising an fp32 copy of a multi-GB pool."""
    out = torch.empty(shape, dtype=torch.bfloat16)
    flat = out.view(shape[0], -1)
    for r0 in range(0, shape[0], chunk_rows):
        r1 = min(shape[0], r0 + chunk_rows)
        flat[r0:r1] = torch.randn((r1 - r0, flat.shape[1]), generator=gen, dtype=torch.float32).to(torch.bfloat16)
    return out


def quantise_pool(x_bf16: torch.Tensor, scale: float) -> torch.Tensor:
    """(x / s).to(float8_e4m3fn) viewed as uint8 -- how an fp8 pool is filled."""
    y = x_bf16 if scale == 1.0 else (x_bf16.float() / scale).to(torch.bfloat16)
    return y.to(torch.float8_e4m3fn).view(torch.uint8)


def make_case(bs: int, Hq: int, Hkv: int, D: int, seq_lens: Sequence[int], extend_lens: Optional[Sequence[int]] = None,
              page_size: int = 1, fp8: bool = False, kv_scale: float = 1.0, seed: int = BASELINE_SEED,
              table: str = "permuted", idx_dtype: torch.dtype = torch.int64, pool_slack: int = 0,
              max_reqs: Optional[int] = None, dtype: torch.dtype = torch.bfloat16) -> types.SimpleNamespace:
    """seq_lens[i] = L_i (total kv length incl. new tokens); extend_lens[i] = E_i (None -> decode, E_i = 1).

    table: "permuted" scatters token slots randomly over the pool (worst-case gather; pages of
    `page_size` consecutive slots stay together, as the allocator guarantees), "arange" lays the
    requests out back to back (the reference tests' layout, test_decode.py:95-99)."""
    gen = torch.Generator().manual_seed(seed)
    seq = torch.as_tensor(list(seq_lens), dtype=torch.int64)
    assert seq.numel() == bs
    decode = extend_lens is None
    ext = torch.ones(bs, dtype=torch.int64) if decode else torch.as_tensor(list(extend_lens), dtype=torch.int64)
    assert bool((ext <= seq).all()) and bool((ext >= 0).all())
    T = int(ext.sum())
    max_reqs = max_reqs or bs
    ctx = int(((int(seq.max()) if bs else 1) + page_size - 1) // page_size * page_size)

    pages_per_req = (seq + page_size - 1) // page_size
    n_pages = int(pages_per_req.sum())
    total_pages = n_pages + (pool_slack + page_size - 1) // page_size
    P = (total_pages + 1) * page_size  # + the dummy page [0, page_size)

    if table == "permuted":
        page_ids = torch.randperm(total_pages, generator=gen)[:n_pages] + 1
    else:
        page_ids = torch.arange(n_pages) + 1
    req_rows = torch.randperm(max_reqs, generator=gen)[:bs]
    req_to_token = torch.zeros((max_reqs, ctx), dtype=torch.int32)
    off = 0
    within = torch.arange(page_size)
    for i in range(bs):
        npg = int(pages_per_req[i])
        slots = (page_ids[off:off + npg, None] * page_size + within[None, :]).reshape(-1)
        req_to_token[req_rows[i], :npg * page_size] = slots.to(torch.int32)
        off += npg

    k_pool = _rand_bf16(gen, (P, Hkv, D))
    v_pool = _rand_bf16(gen, (P, Hkv, D))
    k_pool[:page_size] = 0
    v_pool[:page_size] = 0
    q = _rand_bf16(gen, (max(T, 1), Hq, D))[:T]
    k_new = _rand_bf16(gen, (max(T, 1), Hkv, D))[:T]
    v_new = _rand_bf16(gen, (max(T, 1), Hkv, D))[:T]
    if dtype != torch.bfloat16:
        k_pool, v_pool, q, k_new, v_new = (t.to(dtype) for t in (k_pool, v_pool, q, k_new, v_new))

    prefix = seq - ext
    start = torch.cumsum(ext, 0) - ext
    loc = torch.cat([req_to_token[req_rows[i], int(prefix[i]):int(seq[i])].to(torch.int64) for i in range(bs)]) \
        if bs else torch.zeros(0, dtype=torch.int64)

    case = types.SimpleNamespace(
        bs=bs, Hq=Hq, Hkv=Hkv, D=D, T=T, P=P, page_size=page_size, fp8=fp8, kv_scale=float(kv_scale),
        decode=decode, sm_scale=float(D) ** -0.5, dtype=dtype,
        q=q, k_new=k_new, v_new=v_new, k_pool=k_pool, v_pool=v_pool, req_to_token=req_to_token,
        req_pool_indices=req_rows.to(idx_dtype), seq_lens=seq.to(idx_dtype), out_cache_loc=loc,
        extend_seq_lens=ext.to(torch.int32), extend_prefix_lens=prefix.to(torch.int32),
        extend_start_loc=start.to(torch.int32), max_extend_len=int(ext.max()) if bs else 0,
        seq_lens_sum=int(seq.sum()))
    if fp8:
        case.k_pool = quantise_pool(k_pool, kv_scale)
        case.v_pool = quantise_pool(v_pool, kv_scale)
    return case


def slice_heads(case: types.SimpleNamespace, rank: int, tp: int) -> types.SimpleNamespace:
    """KV-head partition of a global case, exactly SGLang tensor parallelism (models/llama.py:118-131):
    rank r owns KV heads [r*Hkv/tp, (r+1)*Hkv/tp) and the matching Q heads; the page table is replicated."""
    assert case.Hkv % tp == 0
    hk = case.Hkv // tp
    hq = case.Hq // tp
    out = types.SimpleNamespace(**vars(case))
    out.Hq, out.Hkv = hq, hk
    out.q = case.q[:, rank * hq:(rank + 1) * hq].contiguous()
    for name in ("k_new", "v_new", "k_pool", "v_pool"):
        setattr(out, name, getattr(case, name)[:, rank * hk:(rank + 1) * hk].contiguous())
    return out


# ------------------------------------------------------------------ BASELINE.json configs
def config_shapes(cfg: int) -> dict:
    """Shapes of BASELINE.json configs[cfg-1] (C1..C5)."""
    return {
        1: dict(bs=8, Hq=32, Hkv=8, D=64, ctx=512, fp8=False, page_size=1),
        2: dict(bs=256, Hq=32, Hkv=8, D=128, ctx=4096, fp8=False, page_size=1),
        3: dict(bs=16, Hq=32, Hkv=8, D=128, prefix=2048, extend=2048, fp8=False, page_size=1),
        4: dict(bs=512, Hq=64, Hkv=8, D=128, ctx=8192, fp8=True, page_size=1),
        5: dict(bs=256, Hq=64, Hkv=8, D=128, ctx_max=32768, fp8=True, page_size=16, n_extend=8, extend=1024),
    }[cfg]


def algorithmic_decode_bytes(seq_lens: Sequence[int], Hq: int, Hkv: int, D: int, kv_bytes: int, q_bytes: int = 2) -> int:
    """KV read + page-table read (once per layer) + Q in + O out + new K/V read and pool write."""
    tot = int(sum(int(x) for x in seq_lens))
    bs = len(seq_lens)
    return (tot * Hkv * 2 * D * kv_bytes + tot * 4 + 2 * bs * Hq * D * q_bytes
            + bs * Hkv * 2 * D * (q_bytes + kv_bytes))


def algorithmic_extend_flops(extend_lens: Sequence[int], prefix_lens: Sequence[int], Hq: int, D: int,
                             causal: bool = True) -> int:
    """4 * Hq * D * sum_i (E_i * P_i + E_i (E_i + 1) / 2); masked-out products are not counted."""
    tot = 0
    for e, pfx in zip(extend_lens, prefix_lens):
        e, pfx = int(e), int(pfx)
        tot += e * pfx + (e * (e + 1) // 2 if causal else e * e)
    return 4 * Hq * D * tot


# ------------------------------------------------------------------ full-size cases, built on the device
def make_case_device(bs: int, Hq: int, Hkv: int, D: int, seq_lens: Sequence[int], device,
                     extend_lens: Optional[Sequence[int]] = None, page_size: int = 1, fp8: bool = False,
                     kv_scale: float = 1.0, seed: int = BASELINE_SEED, table: str = "permuted",
                     idx_dtype: torch.dtype = torch.int64, dtype: torch.dtype = torch.bfloat16,
                     pool_rows: Optional[int] = None, head_shard: Optional[tuple] = None) -> types.SimpleNamespace:
    """Same layout as make_case, generated with the device's RNG so BASELINE-size pools (GiBs) never cross PCIe.
    The pool state is AFTER-write-agnostic: rows of the new tokens hold random data that the forward overwrites.

    head_shard = (rank, world): build only rank's slice of the GLOBAL case with Hq / Hkv heads, partitioned the way
    SGLang tensor parallelism does (models/llama.py:118-131; KV heads replicated when world > Hkv,
    model_config.py:342-349).  Every head is drawn from its own generator (seeded by the global head index), so
    the concatenation of the ranks' slices is one and the same global case whatever `world` is -- including
    world = 1 -- and no rank ever materialises more than its slice.  The page table is identical on every rank."""
    dev = torch.device(device)
    q_ids = kv_ids = None
    if head_shard is not None:
        rank, world = head_shard
        assert Hq % world == 0 and (Hkv % world == 0 or world % Hkv == 0), "tp must divide / be a multiple of Hkv"
        hq_l, hkv_l = Hq // world, max(1, Hkv // world)
        q_ids = list(range(rank * hq_l, (rank + 1) * hq_l))
        kv0 = rank * hkv_l if Hkv % world == 0 else rank // (world // Hkv)
        kv_ids = list(range(kv0, kv0 + hkv_l))
        hkv_global = Hkv
        Hq, Hkv = hq_l, hkv_l
    gen = torch.Generator(device=dev).manual_seed(seed)
    cpu_gen = torch.Generator().manual_seed(seed)
    seq = torch.as_tensor(list(seq_lens), dtype=torch.int64)
    decode = extend_lens is None
    ext = torch.ones(bs, dtype=torch.int64) if decode else torch.as_tensor(list(extend_lens), dtype=torch.int64)
    T = int(ext.sum())
    ctx = int(((int(seq.max()) if bs else 1) + page_size - 1) // page_size * page_size)
    pages_per_req = (seq + page_size - 1) // page_size
    n_pages = int(pages_per_req.sum())
    total_pages = max(n_pages, ((pool_rows or 0) + page_size - 1) // page_size - 1)
    P = (total_pages + 1) * page_size
    if table == "permuted":
        page_ids = torch.randperm(total_pages, generator=cpu_gen)[:n_pages] + 1
    else:
        page_ids = torch.arange(n_pages) + 1
    req_rows = torch.randperm(bs, generator=cpu_gen)
    req_to_token = torch.zeros((bs, ctx), dtype=torch.int32)
    within = torch.arange(page_size)
    off = 0
    for i in range(bs):
        npg = int(pages_per_req[i])
        slots = (page_ids[off:off + npg, None] * page_size + within[None, :]).reshape(-1)
        req_to_token[req_rows[i], :npg * page_size] = slots.to(torch.int32)
        off += npg

    def rnd(shape, ids=None, salt=0):
        out = torch.empty(shape, dtype=dtype, device=dev)
        step = 1 << 18
        if ids is None:
            flat = out.view(shape[0], -1)
            for r0 in range(0, shape[0], step):
                r1 = min(shape[0], r0 + step)
                flat[r0:r1] = torch.randn((r1 - r0, flat.shape[1]), generator=gen, device=dev, dtype=torch.float32).to(dtype)
            return out
        for j, h in enumerate(ids):  # one generator per GLOBAL head: the data do not depend on the partition
            gh = torch.Generator(device=dev).manual_seed(seed * 1_000_003 + salt * 10_007 + h)
            for r0 in range(0, shape[0], step * 4):
                r1 = min(shape[0], r0 + step * 4)
                out[r0:r1, j] = torch.randn((r1 - r0, shape[2]), generator=gh, device=dev, dtype=torch.float32).to(dtype)
        return out

    k_pool, v_pool = rnd((P, Hkv, D), kv_ids, 1), rnd((P, Hkv, D), kv_ids, 2)
    k_pool[:page_size] = 0
    v_pool[:page_size] = 0
    if fp8:
        def quant(x):
            out = torch.empty(x.shape, dtype=torch.uint8, device=dev)
            step = 1 << 18
            for r0 in range(0, x.shape[0], step):
                y = x[r0:r0 + step]
                if kv_scale != 1.0:
                    y = (y.float() / kv_scale).to(dtype)
                out[r0:r0 + step] = y.to(torch.float8_e4m3fn).view(torch.uint8)
            return out
        k_pool, v_pool = quant(k_pool), quant(v_pool)
    q = rnd((max(T, 1), Hq, D), q_ids, 3)[:T]
    k_new, v_new = rnd((max(T, 1), Hkv, D), kv_ids, 4)[:T], rnd((max(T, 1), Hkv, D), kv_ids, 5)[:T]
    prefix = seq - ext
    start = torch.cumsum(ext, 0) - ext
    loc = torch.cat([req_to_token[req_rows[i], int(prefix[i]):int(seq[i])].to(torch.int64) for i in range(bs)]) \
        if bs else torch.zeros(0, dtype=torch.int64)
    return types.SimpleNamespace(
        bs=bs, Hq=Hq, Hkv=Hkv, D=D, T=T, P=P, page_size=page_size, fp8=fp8, kv_scale=float(kv_scale), decode=decode,
        sm_scale=float(D) ** -0.5, dtype=dtype, q=q, k_new=k_new, v_new=v_new, k_pool=k_pool, v_pool=v_pool,
        req_to_token=req_to_token.to(dev), req_pool_indices=req_rows.to(idx_dtype).to(dev),
        seq_lens=seq.to(idx_dtype).to(dev), out_cache_loc=loc.to(dev), extend_seq_lens=ext.to(torch.int32).to(dev),
        extend_prefix_lens=prefix.to(torch.int32).to(dev), extend_start_loc=start.to(torch.int32).to(dev),
        max_extend_len=int(ext.max()) if bs else 0, seq_lens_sum=int(seq.sum()),
        Hkv_global=hkv_global if head_shard is not None else Hkv)


def c5_lengths(seed: int = 5, scale: int = 1) -> tuple:
    """BASELINE configs[4]: one MIXED forward of 8 chunked-prefill rows (prefix 16 * U[0, 1536), chunk
    1024) followed by 248 running decode rows with L ~ U[1024, 32768].  `scale` divides counts and lengths for tests.
    Returns (seq_lens, extend_lens, n_extend_rows)."""
    g = torch.Generator().manual_seed(seed)
    n_dec, n_ext, E = 248 // scale, max(1, 8 // scale), 1024 // scale
    ext_prefix = (16 * torch.randint(0, max(1, 1536 // scale), (n_ext,), generator=g)).tolist()
    dec_lens = torch.randint(1024 // scale, 32768 // scale + 1, (n_dec,), generator=g).tolist()
    return [p + E for p in ext_prefix] + dec_lens, [E] * n_ext + [1] * n_dec, n_ext


def extract_subcase(case: types.SimpleNamespace, req_ids: Sequence[int], k_pool=None, v_pool=None) -> types.SimpleNamespace:
    """A compact CPU copy of the listed requests of a (device) case: their pool rows are gathered and the page
    table re-indexed, so a CPU oracle can check a sample of a BASELINE-size run.  `k_pool`/`v_pool` override the
    pools to read from (e.g. the pools after the forward's KV write)."""
    kp = case.k_pool if k_pool is None else k_pool
    vp = case.v_pool if v_pool is None else v_pool
    ids = list(req_ids)
    seq = case.seq_lens[ids].cpu().to(torch.int64)
    ext = case.extend_seq_lens[ids].cpu().to(torch.int64)
    rows = case.req_pool_indices[ids].cpu().to(torch.int64)
    ctx = int(seq.max())
    tab = case.req_to_token[rows.to(case.req_to_token.device)][:, :ctx].cpu()
    new_tab = torch.zeros((len(ids), ctx), dtype=torch.int32)
    slots = []
    nxt = 1
    for j in range(len(ids)):
        L = int(seq[j])
        slots.append(tab[j, :L].to(torch.int64))
        new_tab[j, :L] = torch.arange(nxt, nxt + L, dtype=torch.int32)
        nxt += L
    all_slots = torch.cat(slots).to(kp.device)
    sub_k = torch.zeros((nxt,) + tuple(kp.shape[1:]), dtype=kp.dtype)
    sub_v = torch.zeros_like(sub_k)
    sub_k[1:] = kp[all_slots].cpu()
    sub_v[1:] = vp[all_slots].cpu()
    start_full = case.extend_start_loc[ids].cpu().to(torch.int64)
    tok = torch.cat([torch.arange(int(start_full[j]), int(start_full[j] + ext[j])) for j in range(len(ids))])
    prefix = seq - ext
    start = torch.cumsum(ext, 0) - ext
    loc = torch.cat([new_tab[j, int(prefix[j]):int(seq[j])].to(torch.int64) for j in range(len(ids))])
    tokd = tok.to(case.q.device)
    return types.SimpleNamespace(
        bs=len(ids), Hq=case.Hq, Hkv=case.Hkv, D=case.D, T=int(ext.sum()), P=nxt, page_size=1, fp8=case.fp8,
        kv_scale=case.kv_scale, decode=case.decode, sm_scale=case.sm_scale, dtype=case.dtype,
        q=case.q[tokd].cpu(), k_new=case.k_new[tokd].cpu(), v_new=case.v_new[tokd].cpu(), k_pool=sub_k, v_pool=sub_v,
        req_to_token=new_tab, req_pool_indices=torch.arange(len(ids), dtype=torch.int64), seq_lens=seq,
        out_cache_loc=loc, extend_seq_lens=ext.to(torch.int32), extend_prefix_lens=prefix.to(torch.int32),
        extend_start_loc=start.to(torch.int32), max_extend_len=int(ext.max()), seq_lens_sum=int(seq.sum()),
        token_rows=tok)
