"""CPU tests of the host-side logic: the synthetic workload generator, the SGLang interface stand-ins,
the algorithmic byte/FLOP formulas bench.py reports, and the N>1 head-partition path (gloo, world_size 2)."""
from __future__ import annotations

import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers as H
from oracle import torch_native as tn
from sglang_custom_b200 import interfaces as itf
from sglang_custom_b200 import synth, tp


def test_make_case_invariants():
    c = synth.make_case(5, 8, 2, 64, [17, 1, 64, 33, 9], extend_lens=[4, 1, 64, 10, 9], page_size=4, seed=5, max_reqs=9)
    assert c.T == 88 and c.q.shape == (88, 8, 64) and c.k_new.shape == (88, 2, 64)
    assert c.req_to_token.dtype == torch.int32 and c.out_cache_loc.dtype == torch.int64
    assert torch.equal(c.extend_prefix_lens + c.extend_seq_lens, c.seq_lens.to(torch.int32))
    # the dummy page is zero and never referenced by a live token
    assert not c.k_pool[:4].any()
    live = torch.cat([c.req_to_token[c.req_pool_indices[i], :c.seq_lens[i]] for i in range(5)])
    assert live.min() >= 4 and live.max() < c.P and live.unique().numel() == live.numel()
    # slots inside a page are consecutive (allocator.py:444-447)
    row = c.req_to_token[c.req_pool_indices[2], :64].view(-1, 4)
    assert torch.equal(row[:, 1:] - row[:, :-1], torch.ones(16, 3, dtype=torch.int32))
    assert (row[:, 0] % 4 == 0).all()
    # out_cache_loc = the table entries of the new tokens
    i = 3
    start = int(c.extend_start_loc[i])
    assert torch.equal(c.out_cache_loc[start:start + 10],
                       c.req_to_token[c.req_pool_indices[i], 23:33].to(torch.int64))


def test_make_case_is_deterministic_and_seed_sensitive():
    a = synth.make_case(2, 4, 2, 64, [10, 20], seed=9)
    b = synth.make_case(2, 4, 2, 64, [10, 20], seed=9)
    c = synth.make_case(2, 4, 2, 64, [10, 20], seed=10)
    assert torch.equal(a.q, b.q) and torch.equal(a.req_to_token, b.req_to_token) and torch.equal(a.k_pool, b.k_pool)
    assert not torch.equal(a.q, c.q)


def test_fp8_case_pool_is_quantised_with_scale():
    c = synth.make_case(1, 4, 2, 64, [12], fp8=True, kv_scale=0.1135, seed=2)
    assert c.k_pool.dtype == torch.uint8
    ref = synth.make_case(1, 4, 2, 64, [12], fp8=False, seed=2)
    want = (ref.k_pool.float() / 0.1135).to(torch.bfloat16).to(torch.float8_e4m3fn).view(torch.uint8)
    assert torch.equal(c.k_pool, want)


def test_algorithmic_formulas_match_survey_figures():
    # SURVEY 8d / BASELINE.md section 3: C2 = 4 305 453 056 B per layer, C3 = 1.6495e12 FLOP per layer
    assert synth.algorithmic_decode_bytes([4096] * 256, 32, 8, 128, 2) == 4_305_453_056
    fl = synth.algorithmic_extend_flops([2048] * 16, [2048] * 16, 32, 128)
    assert abs(fl / 1.6495e12 - 1) < 1e-3
    assert synth.config_shapes(2)["bs"] == 256 and synth.config_shapes(4)["fp8"]


def test_forward_mode_predicates_follow_the_reference():
    M = itf.ForwardMode
    assert M.DECODE.is_decode() and not M.MIXED.is_decode()
    assert all(m.is_extend() for m in (M.EXTEND, M.MIXED, M.DRAFT_EXTEND, M.TARGET_VERIFY))
    assert M.IDLE.is_decode_or_idle() and M.DECODE.is_decode_or_idle() and not M.EXTEND.is_decode_or_idle()


def test_model_config_kv_head_replication():
    cfg = itf.ModelConfig(context_len=8, num_attention_heads=64, num_key_value_heads=8, head_dim=128)
    assert [cfg.get_num_kv_heads(t) for t in (1, 2, 4, 8, 16)] == [8, 4, 2, 1, 1]


def test_shard_case_partitions_heads_like_sglang_tp():
    c = synth.make_case(2, 8, 4, 64, [20, 31], seed=3)
    full = H.oracle_decode(c)
    for world in (2, 4):
        parts = []
        for r in range(world):
            s = tp.shard_case(c, r, world)
            assert s.Hq == 8 // world and s.Hkv == 4 // world
            assert torch.equal(s.req_to_token, c.req_to_token)  # the page table is replicated
            parts.append(H.oracle_decode(s))
        assert torch.equal(torch.cat(parts, dim=1), full)  # heads are independent: bit-identical
    # world > Hkv: KV heads replicated, Q heads split
    s = tp.shard_case(c, 5, 8)
    assert s.Hq == 1 and s.Hkv == 1 and torch.equal(s.k_pool[:, 0], c.k_pool[:, 2]) and torch.equal(s.q[:, 0], c.q[:, 5])


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _tp_worker(rank: int, world: int, port: int, q):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    torch.set_num_threads(1)
    r, w, _ = tp.init("gloo")
    c = synth.make_case(3, 8, 2, 64, [40, 7, 25], extend_lens=[5, 7, 1], seed=11)
    local = tp.shard_case(c, r, w)
    kp, vp = H.pool_after_write(local)
    out_local = tn.extend(local.q, kp, vp, local.req_to_token, local.req_pool_indices, local.seq_lens,
                          local.extend_prefix_lens, local.extend_seq_lens, local.sm_scale)
    tp.barrier(w)
    gathered = tp.gather_heads(out_local, w)
    t_max = tp.max_over_ranks(1.0 + r, w)
    t_sum = tp.sum_over_ranks(1.0 + r, w)
    if r == 0:
        want = H.oracle_extend(c)
        q.put((bool(torch.equal(gathered, want)), t_max, t_sum))
    tp.barrier(w)
    dist.destroy_process_group()


def test_tp2_gloo_partition_gather_and_max_timing():
    """world_size 2 over gloo: every rank runs its KV-head slice, outputs are all-gathered and equal the
    global oracle; the timing reduction is the MAX over ranks."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_tp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, t_max, t_sum = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok and t_max == 2.0 and t_sum == 3.0


def _tp_subcase_worker(rank: int, world: int, port: int, q):
    """The bench.py / test_gpu_tp.py parity path with the per-rank attention replaced by the oracle on the rank's
    slice: head-sharded generation, extract_subcase on every rank, all-gather of the sample into the GLOBAL sub-case."""
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    torch.set_num_threads(1)
    r, w, _ = tp.init("gloo")
    lens, ext = [40, 7, 25, 33], [5, 7, 1, 1]
    local = synth.make_case_device(4, 8, 2, 64, lens, "cpu", extend_lens=ext, seed=21, head_shard=(r, w))
    assert local.Hq == 8 // w and local.Hkv == max(1, 2 // w) and local.Hkv_global == 2
    kp, vp = H.pool_after_write(local)
    out_local = tn.extend(local.q, kp, vp, local.req_to_token, local.req_pool_indices, local.seq_lens,
                          local.extend_prefix_lens, local.extend_seq_lens, local.sm_scale)
    sample = [0, 2, 3]
    sub = synth.extract_subcase(local, sample, kp, vp)
    gsub = tp.gather_subcase(sub, w, kv_replicas=max(1, w // 2))
    got = tp.gather_cat(out_local[sub.token_rows], w)
    ok_all = tp.all_true(r != 1, w)  # rank 1 votes False
    if r == 0:
        full = synth.make_case_device(4, 8, 2, 64, lens, "cpu", extend_lens=ext, seed=21, head_shard=(0, 1))
        fk, fv = H.pool_after_write(full)
        fsub = synth.extract_subcase(full, sample, fk, fv)
        same = all(torch.equal(getattr(gsub, n), getattr(fsub, n)) for n in ("q", "k_new", "v_new", "k_pool", "v_pool"))
        want = H.oracle_extend(gsub, gsub.k_pool, gsub.v_pool)
        q.put((same, bool(torch.equal(got, want)), gsub.Hq, gsub.Hkv, ok_all))
    tp.barrier(w)
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_tp_gloo_head_sharded_generation_and_global_subcase(world):
    """world 2 (KV heads split) and 4 (world > Hkv: KV heads replicated, Q heads split): the ranks' slices are one
    global case whatever the world size, and the gathered sample equals the world-1 sample bit for bit."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_tp_subcase_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    same, ok, hq, hkv, all_ok = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert same and ok and (hq, hkv) == (8, 2) and all_ok is False


def test_choose_kv_splits_cost_model():
    """The backend's split factor for prefill rows (host-side launch simulation, backend._choose_kv_splits): big grids
    are never split, a long-prefix chunk on few heads is, and the answer is a pure function of the host length lists."""
    from sglang_custom_b200 import synth
    from sglang_custom_b200.backend import B200AttnBackend

    class Stub:
        extend_kv_splits = 0
        n_ctas = 148

    def choose(Hq, Hkv, ext, pre):
        s = Stub()
        s.num_head, s.num_kv_head = Hq, Hkv
        return B200AttnBackend._choose_kv_splits(s, ext, pre)

    assert choose(32, 8, [2048] * 16, [2048] * 16) == 1        # config 3: 4096 items
    assert choose(32, 8, [300] * 3, [100] * 3) == 1            # short rows
    assert choose(8, 1, [2048], [30000]) > 1                   # one long chunked-prefill row, one KV head per rank
    assert choose(8, 1, [1024] * 2, [8192] * 2) > 1
    lens, ext, n_ext = synth.c5_lengths()
    pre = [l - e for l, e in zip(lens[:n_ext], ext[:n_ext])]
    assert choose(64, 8, ext[:n_ext], pre) == 1                # config 5 on one GPU: 2048 items
    assert choose(8, 1, ext[:n_ext], pre) > 1                  # config 5 at TP 8: 256 skewed items
    s = Stub()
    s.num_head, s.num_kv_head, s.extend_kv_splits = 8, 1, 4
    assert B200AttnBackend._choose_kv_splits(s, [128], [0]) == 4  # forced
