#!/usr/bin/env python
"""Time the sliding-window and verify (custom-mask) paths at BASELINE-sized shapes; CUDA events, 20 launches.

  window decode  : C2 shape (bs 256, ctx 4096, 32/8, D 128, bf16), W = 1023 -> algorithmic GB/s over the window bytes
  window extend  : C3 shape (bs 16, prefix 2048 + 2048 new), W in {1023, 4095}: tcgen05 vs register-MMA kernel;
                   TFLOP/s over the FLOPs of the VISIBLE (query, key) pairs
  target verify  : bs 256, ctx 4096, 8 draft tokens per request, tree mask: GB/s over the KV bytes one pass reads
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from sglang_custom_b200 import _lib, ops, synth


def timed(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def window_decode(dev):
    bs, L, Hq, Hkv, D, W = 256, 4096, 32, 8, 128, 1023
    c = synth.make_case_device(bs, Hq, Hkv, D, [L] * bs, dev, seed=1236)
    n_ctas = 148
    plan = torch.zeros(ops.decode_plan_bytes(bs, n_ctas) // 4, dtype=torch.int32, device=dev)
    ws = torch.empty(ops.decode_workspace_bytes(n_ctas, Hq, D) // 4, dtype=torch.float32, device=dev)
    out = torch.empty_like(c.q)
    for w in (-1, W, 255):
        def fn():
            ops.decode_plan(c.seq_lens, n_ctas, plan, 4, sliding_window_size=w)
            ops.decode(c.q, out, c.k_pool, c.v_pool, c.req_to_token, c.req_pool_indices, c.seq_lens, plan, n_ctas, ws,
                       c.sm_scale)
        ms = timed(fn)
        keys = L if w < 0 else min(L, w + 1)
        by = synth.algorithmic_decode_bytes([keys] * bs, Hq, Hkv, D, 2)
        print(f"window decode C2 W={w}: {ms * 1e3:.1f} us  {by / ms / 1e6:.0f} GB/s (plan + stage 1 + merge)", flush=True)


def visible_flops(P, E, W, Hq, D, bs):
    t = torch.arange(E, dtype=torch.int64)
    if W < 0:
        pre = torch.full((E,), P, dtype=torch.int64)
    else:
        kept = min(P, W + 1)
        pre = torch.clamp(kept - torch.clamp(t - W, min=0), min=0) if W > 0 else torch.full((E,), kept)
    pairs = int(pre.sum()) + E * (E + 1) // 2
    return 4 * D * pairs * Hq * bs


def window_extend(dev):
    bs, P, E, Hq, Hkv, D = 16, 2048, 2048, 32, 8, 128
    c = synth.make_case_device(bs, Hq, Hkv, D, [P + E] * bs, dev, extend_lens=[E] * bs, seed=1237)
    out = torch.empty_like(c.q)
    for W in (-1, 4095, 1023):
        for impl, name in ((_lib.EXTEND_TCGEN05, "tcgen05"), (_lib.EXTEND_MMA, "register-MMA")):
            fn = lambda: ops.extend(c.q, out, c.k_pool, c.v_pool, c.req_to_token, c.req_pool_indices, c.seq_lens,
                                    c.extend_seq_lens, c.extend_prefix_lens, c.extend_start_loc, E, c.sm_scale,
                                    impl=impl, sliding_window_size=W)
            ms = timed(fn, n=10)
            fl = visible_flops(P, E, W, Hq, D, bs)
            print(f"window extend C3 W={W} {name}: {ms:.3f} ms  {fl / ms / 1e9:.0f} TFLOP/s (visible pairs)", flush=True)


def target_verify(dev):
    bs, L, Hq, Hkv, D, nd = 256, 4096, 32, 8, 128, 8
    c = synth.make_case_device(bs, Hq, Hkv, D, [L + nd] * bs, dev, extend_lens=[nd] * bs, seed=1238)
    out = torch.empty_like(c.q)
    blk = torch.cat([torch.ones(nd, L, dtype=torch.bool), torch.tril(torch.ones(nd, nd)).bool()], 1)
    blk[5, L + 3] = False  # not the plain causal mask
    mask = blk.reshape(-1).repeat(bs).to(dev)
    indptr = (torch.arange(bs + 1, dtype=torch.int64) * blk.numel()).to(dev)
    fn = lambda: ops.extend(c.q, out, c.k_pool, c.v_pool, c.req_to_token, c.req_pool_indices, c.seq_lens,
                            c.extend_seq_lens, c.extend_prefix_lens, c.extend_start_loc, nd, c.sm_scale,
                            custom_mask=mask, mask_indptr=indptr)
    ms = timed(fn)
    by = bs * (L + nd) * Hkv * D * 2 * 2
    print(f"target verify bs={bs} ctx={L} draft={nd}, one masked extend (register-MMA): {ms * 1e3:.1f} us  "
          f"{by / ms / 1e6:.0f} GB/s of KV (one pass counted; the kernel re-reads K/V once per q head of a group)", flush=True)
    # the backend's route: prefix part on the split-KV decode kernel (16 (token, head) rows per pass), masked draft
    # part on the register-MMA kernel, LSE merge
    n_ctas, per_pass = 148, 16 // (Hq // Hkv)
    plan = torch.zeros(ops.decode_plan_bytes(bs, n_ctas) // 4, dtype=torch.int32, device=dev)
    ws = torch.empty(ops.decode_workspace_bytes(n_ctas, Hkv * 16, D) // 4, dtype=torch.float32, device=dev)
    pre = c.extend_prefix_lens.to(torch.int64)
    q4, o4 = c.q.unflatten(0, (bs, nd)), out.unflatten(0, (bs, nd))
    lse_p = torch.empty((bs, nd, Hq), dtype=torch.float32, device=dev)
    o_d = torch.empty_like(out)
    lse_d = torch.empty((bs * nd, Hq), dtype=torch.float32, device=dev)

    def route():
        ops.decode_plan(pre, n_ctas, plan, 4)
        for t0 in range(0, nd, per_pass):
            ops.decode_multiq(q4[:, t0:t0 + per_pass], o4[:, t0:t0 + per_pass], c.k_pool, c.v_pool, c.req_to_token,
                              c.req_pool_indices, pre, plan, n_ctas, ws, c.sm_scale, lse_out=lse_p[:, t0:t0 + per_pass])
        ops.extend(c.q, o_d, c.k_pool, c.v_pool, c.req_to_token, c.req_pool_indices, c.seq_lens, c.extend_seq_lens,
                   c.extend_prefix_lens, c.extend_start_loc, nd, c.sm_scale, lse_out=lse_d,
                   sliding_window_size=_lib.WINDOW_NEW_TOKENS_ONLY, custom_mask=mask, mask_indptr=indptr)
        ops.merge_state(out, lse_p.view(bs * nd, Hq), o_d, lse_d, out=out)
    ms = timed(route)
    passes = -(-nd // per_pass)
    print(f"target verify, decode_multiq x{passes} + masked draft part + merge: {ms * 1e3:.1f} us  "
          f"{by / ms / 1e6:.0f} GB/s of KV (one pass counted), {by * passes / ms / 1e6:.0f} GB/s actually streamed", flush=True)


if __name__ == "__main__":
    dev = torch.device("cuda:0")
    which = sys.argv[1:] or ["decode", "extend", "verify"]
    if "decode" in which:
        window_decode(dev)
    if "extend" in which:
        window_extend(dev)
    if "verify" in which:
        target_verify(dev)
