#!/usr/bin/env python
"""Time the decode path (plan + stage 1 + merge) at the BASELINE shapes: C2 (bf16, 8 KV heads), its ragged variant,
and the per-rank shapes of C4 (FP8 KV, bs 512, ctx 8192, TP 2/4/8 -> 4/2/1 KV heads).  Prints algorithmic GB/s."""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from sglang_custom_b200 import ops, synth


def time_ms(fn, reps=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def run(name, bs, Hq, Hkv, D, lens, fp8, n_ctas_list=(148,), page_size=1, table="permuted"):
    dev = torch.device("cuda:0")
    case = synth.make_case_device(bs, Hq, Hkv, D, lens, dev, seed=1240, fp8=fp8, page_size=page_size, table=table)
    kvb = 1 if fp8 else 2
    nbytes = synth.algorithmic_decode_bytes(lens, Hq, Hkv, D, kvb)
    out = torch.empty_like(case.q)
    res = []
    for n_ctas in n_ctas_list:
        plan = torch.zeros(ops.decode_plan_bytes(bs, n_ctas) // 4, dtype=torch.int32, device=dev)
        ws = torch.empty(ops.decode_workspace_bytes(n_ctas, Hq, D) // 4, dtype=torch.float32, device=dev)
        ops.decode_plan(case.seq_lens, n_ctas, plan, 4)
        fn = lambda: ops.decode(case.q, out, case.k_pool, case.v_pool, case.req_to_token, case.req_pool_indices,
                                case.seq_lens, plan, n_ctas, ws, case.sm_scale,
                                kv_dtype=torch.float8_e4m3fn if fp8 else None)
        ms = time_ms(fn)
        # the same call replayed from a CUDA graph of 10 (device time without the host's launch path)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for _ in range(10):
                fn()
        ms_g = time_ms(g.replay, reps=5, warm=2) / 10
        del g
        res.append(dict(name=name, n_ctas=n_ctas, us=round(ms * 1e3, 1), gbs=round(nbytes / ms / 1e6, 1),
                        graph_us=round(ms_g * 1e3, 1), graph_gbs=round(nbytes / ms_g / 1e6, 1)))
        print(json.dumps(res[-1]), flush=True)
    del case
    torch.cuda.empty_cache()
    return res


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    g = torch.Generator().manual_seed(7)
    if which in ("all", "c2"):
        run("C2 bf16 32/8 bs256 ctx4096", 256, 32, 8, 128, [4096] * 256, False, (148, 296))
        ragged = torch.randint(2048, 4097, (256,), generator=g).tolist()
        run("C2 ragged U[2048,4096]", 256, 32, 8, 128, ragged, False, (148,))
        run("C2 page16", 256, 32, 8, 128, [4096] * 256, False, (148,), page_size=16)
        run("C2 arange table (sequential slots)", 256, 32, 8, 128, [4096] * 256, False, (148,), table="arange")
    if which in ("all", "c4"):
        for tp, (hq, hkv) in ((1, (64, 8)), (2, (32, 4)), (4, (16, 2)), (8, (8, 1))):
            run(f"C4 fp8 tp{tp} {hq}/{hkv} bs512 ctx8192", 512, hq, hkv, 128, [8192] * 512, True, (148, 296))
    if which in ("all", "c4layout"):  # how much of the FP8 / 1-KV-head gap is DRAM row locality (128-byte random rows)
        run("C4 fp8 tp8 8/1 arange table", 512, 8, 1, 128, [8192] * 512, True, (148,), table="arange")
        run("C4 fp8 tp8 8/1 page16 permuted", 512, 8, 1, 128, [8192] * 512, True, (148,), page_size=16)
        run("C4 fp8 tp2 32/4 arange table", 512, 32, 4, 128, [8192] * 512, True, (148,), table="arange")
        run("bf16 8/1 bs512 ctx8192 permuted (256-byte rows)", 512, 8, 1, 128, [8192] * 512, False, (148,))
        run("bf16 8/1 bs512 ctx8192 arange", 512, 8, 1, 128, [8192] * 512, False, (148,), table="arange")
    if which in ("all", "small"):
        run("bs16 ctx1024 32/8 bf16", 16, 32, 8, 128, [1024] * 16, False, (148,))
        run("bs1 ctx32768 32/8 bf16", 1, 32, 8, 128, [32768], False, (148,))
        run("bs64 ctx512 32/8 bf16", 64, 32, 8, 128, [512] * 64, False, (148,))
    if which in ("lowbatch",):
        # the low-batch regime: few requests, long contexts -- split-KV has to spread ONE request over all SMs
        for bs, ctx in ((1, 4096), (1, 32768), (1, 131072), (4, 32768), (8, 8192), (16, 4096), (32, 4096), (64, 4096)):
            run(f"bf16 32/8 bs{bs} ctx{ctx}", bs, 32, 8, 128, [ctx] * bs, False, (148, 296))
        for bs, ctx in ((1, 32768), (8, 8192), (32, 8192)):
            run(f"fp8 8/1 bs{bs} ctx{ctx}", bs, 8, 1, 128, [ctx] * bs, True, (148, 296))


if __name__ == "__main__":
    main()
