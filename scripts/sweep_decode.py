#!/usr/bin/env python
"""Time the decode stage-1 kernel at BASELINE config 2 under different page-table layouts / grid sizes."""
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


def main():
    dev = torch.device("cuda:0")
    bs, ctx, Hq, Hkv, D = 256, 4096, 32, 8, 128
    for table in ("permuted", "arange"):
        for fp8 in (False, True):
            case = synth.make_case_device(bs, Hq, Hkv, D, [ctx] * bs, dev, seed=1236, table=table, fp8=fp8)
            kvb = 1 if fp8 else 2
            nbytes = bs * ctx * Hkv * 2 * D * kvb + bs * ctx * 4 + 2 * bs * Hq * D * 2
            for n_ctas in (148, 296, 444):
                plan = torch.zeros(ops.decode_plan_bytes(bs, n_ctas) // 4, dtype=torch.int32, device=dev)
                ws = torch.empty(ops.decode_workspace_bytes(n_ctas, Hq, D) // 4, dtype=torch.float32, device=dev)
                ops.decode_plan(case.seq_lens, n_ctas, plan, 4)
                out = torch.empty_like(case.q)
                fn = lambda: ops.decode(case.q, out, case.k_pool, case.v_pool, case.req_to_token, case.req_pool_indices,
                                        case.seq_lens, plan, n_ctas, ws, case.sm_scale, stage="partial",
                                        kv_dtype=torch.float8_e4m3fn if fp8 else None)
                ms = time_ms(fn)
                print(f"table={table:9s} fp8={fp8!s:5s} n_ctas={n_ctas:4d}  {ms * 1e3:8.1f} us  {nbytes / ms / 1e6:8.1f} GB/s", flush=True)
            # memory-system reference on the same access pattern: gather 2 KiB rows -> contiguous copy (read + write)
            idx = case.req_to_token[case.req_pool_indices].reshape(-1)[: bs * ctx // 4].contiguous()
            dst = torch.empty((idx.numel(),) + tuple(case.k_pool.shape[1:]), dtype=case.k_pool.dtype, device=dev)
            ms = time_ms(lambda: ops.gather_kv(case.k_pool, idx, dst))
            gb = 2 * dst.numel() * dst.element_size()
            print(f"table={table:9s} fp8={fp8!s:5s} gather_kv (read+write)   {ms * 1e3:8.1f} us  {gb / ms / 1e6:8.1f} GB/s", flush=True)
            del case, dst
            torch.cuda.empty_cache()
    a = torch.empty(1 << 30, dtype=torch.bfloat16, device=dev)
    b = torch.empty_like(a)
    ms = time_ms(lambda: b.copy_(a), reps=10)
    print(f"torch copy 2 GiB read + 2 GiB write: {4 * (1 << 30) / ms / 1e6:8.1f} GB/s")
    ms = time_ms(lambda: a.sum(), reps=10)
    print(f"torch sum over 2 GiB (read only):     {2 * (1 << 30) / ms / 1e6:8.1f} GB/s")


if __name__ == "__main__":
    main()
