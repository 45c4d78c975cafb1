#!/usr/bin/env python
"""BASELINE config 3 in the three schedules SURVEY.md 8(d) lists -- same FLOPs (1.6495e12 per layer), CUDA events:
  (A) one forward_extend with all 16 requests (prefix 2048 + 2048 new each)
  (B) chunked as the scheduler does with chunked_prefill_size = 8192: 4 forwards x 4 requests
  (C) per-request chunking 2 x 1024: forward 1 = 16 x (prefix 2048, 1024 new), forward 2 = 16 x (prefix 3072, 1024 new)
Each forward = KV write of its new tokens + the extend kernel (what B200AttnBackend.forward_extend launches)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from sglang_custom_b200 import ops, synth


def main():
    dev = torch.device("cuda:0")
    bs, P, E, Hq, Hkv, D = 16, 2048, 2048, 32, 8, 128
    c = synth.make_case_device(bs, Hq, Hkv, D, [P + E] * bs, dev, extend_lens=[E] * bs, seed=1237)
    out = torch.empty_like(c.q)
    fl = synth.algorithmic_extend_flops([E] * bs, [P] * bs, Hq, D)
    i32 = dict(dtype=torch.int32, device=dev)

    def forward(req_lo, req_hi, prefix, new, tok_off):
        """requests [req_lo, req_hi), each with `prefix` cached tokens and `new` new ones; their q rows are the
        slice of the full q that holds positions [tok_off, tok_off + new) of every request's 2048 new tokens"""
        n = req_hi - req_lo
        rows = torch.cat([torch.arange(r * E + tok_off, r * E + tok_off + new, device=dev) for r in range(req_lo, req_hi)])
        q, k, v = c.q[rows], c.k_new[rows], c.v_new[rows]
        loc = c.out_cache_loc[rows]
        o = torch.empty_like(q)
        ext = torch.full((n,), new, **i32)
        pre = torch.full((n,), prefix, **i32)
        start = torch.arange(0, n * new, new, **i32)
        seq = (pre + ext).to(c.seq_lens.dtype)
        rpi = c.req_pool_indices[req_lo:req_hi]

        def fn():
            ops.set_kv_buffer(c.k_pool, c.v_pool, loc, k, v, c.k_pool.dtype)
            ops.extend(q, o, c.k_pool, c.v_pool, c.req_to_token, rpi, seq, ext, pre, start, new, c.sm_scale)
        return fn

    schedules = {
        "A: 1 forward x 16 requests (2048 + 2048)": [forward(0, 16, P, E, 0)],
        "B: 4 forwards x 4 requests (2048 + 2048)": [forward(4 * i, 4 * i + 4, P, E, 0) for i in range(4)],
        "C: 2 forwards x 16 requests (2048 + 1024, then 3072 + 1024)": [forward(0, 16, P, 1024, 0),
                                                                       forward(0, 16, P + 1024, 1024, 1024)],
    }
    for name, fns in schedules.items():
        for _ in range(3):
            for f in fns:
                f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            for f in fns:
                f()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"C3 {name}: {ms:.3f} ms per layer  {fl / ms / 1e9:.0f} TFLOP/s", flush=True)


if __name__ == "__main__":
    main()
