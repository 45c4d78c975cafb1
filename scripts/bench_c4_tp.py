#!/usr/bin/env python
"""BASELINE config 4 on N GPUs of one node: Llama-3-70B-shaped decode (64 Q / 8 KV heads, D 128), FP8-E4M3 KV cache,
bs 512, ctx 8192, KV heads partitioned over the ranks exactly as SGLang tensor parallelism does (rank r owns KV heads
[r * 8 / N, (r + 1) * 8 / N) and their 8 Q heads each; every rank holds the same page table).  Strong scaling: the
whole model's attention of one layer is split over the ranks; no collective inside the path.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_c4_tp.py

Each rank times plan + 32 x (stage 1 + merge) with CUDA events after warm-up, rotating over two pool copies; the job
time is the MAX over ranks; rank 0 prints one JSON line with the aggregate algorithmic GB/s (per-rank bytes include
the replicated page table) and tokens/s."""
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.stdout.flush()
_OUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)
import torch

from sglang_custom_b200 import ops, synth, tp


def main():
    rank, world, local = tp.init()
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    bs, ctx, HQ, HKV, D, layers = 512, 8192, 64, 8, 128, 32
    assert HKV % world == 0
    Hq, Hkv = HQ // world, HKV // world
    c = synth.make_case_device(bs, Hq, Hkv, D, [ctx] * bs, dev, fp8=True, kv_scale=0.1135, seed=1241)
    pools = [(c.k_pool, c.v_pool), (c.k_pool.clone(), c.v_pool.clone())]
    n_ctas = 148
    plan = torch.zeros(ops.decode_plan_bytes(bs, n_ctas) // 4, dtype=torch.int32, device=dev)
    ws = torch.empty(ops.decode_workspace_bytes(n_ctas, Hq, D) // 4, dtype=torch.float32, device=dev)
    out = torch.empty_like(c.q)

    def step():
        ops.decode_plan(c.seq_lens, n_ctas, plan, 4)
        for l in range(layers):
            kp, vp = pools[l & 1]
            ops.decode(c.q, out, kp, vp, c.req_to_token, c.req_pool_indices, c.seq_lens, plan, n_ctas, ws, c.sm_scale,
                       0.0, 0.1135, 0.1135, kv_dtype=torch.float8_e4m3fn)
    for _ in range(3):
        step()
    torch.cuda.synchronize(dev)
    tp.barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 10
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize(dev)
    tp.barrier(world)
    ms = tp.max_over_ranks(e0.elapsed_time(e1), world, dev) / steps
    per_rank = synth.algorithmic_decode_bytes([ctx] * bs, Hq, Hkv, D, 1)
    if rank == 0:
        print(json.dumps({
            "workload": "BASELINE configs[3]: decode bs 512, ctx 8192, 64 Q / 8 KV heads, D 128, FP8-E4M3 KV, page_size 1, "
                        "permuted page table, KV heads partitioned over the ranks", "n_gpus": world, "scaling": "strong",
            "heads_per_rank": f"{Hq}/{Hkv}", "ms_per_layer": round(ms / layers, 4),
            "aggregate_gbs": round(world * per_rank * layers / (ms * 1e-3) / 1e9, 1),
            "per_gpu_gbs": round(per_rank * layers / (ms * 1e-3) / 1e9, 1),
            "attention_tokens_per_s": round(bs / (ms * 1e-3), 1), "bytes_per_rank_per_layer": per_rank}), file=_OUT, flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
