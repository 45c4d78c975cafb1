#!/usr/bin/env python
"""BASELINE config 5 at full size, per-rank TP 8 shape: Qwen2-72B heads 64/8 -> 8 Q / 1 KV head per rank, D 128,
FP8-E4M3 KV, page_size 16, one MIXED forward of 8 chunked-prefill rows (prefix 16*U[0,1536], chunk 1024) + 248
running decode rows with L ~ U[1024, 32768].  Times B200AttnBackend.forward (KV write + extend kernel
for the prefill rows + decode kernel for the tail) and each part alone."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from sglang_custom_b200 import harness, synth


def t_ms(fn, reps=10, warm=3):
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
    g = torch.Generator().manual_seed(5)
    n_dec, n_ext, E = 248, 8, 1024
    ext_prefix = (16 * torch.randint(0, 1536, (n_ext,), generator=g)).tolist()
    dec_lens = torch.randint(1024, 32768 + 1, (n_dec,), generator=g).tolist()
    lens = [p + E for p in ext_prefix] + dec_lens
    ext = [E] * n_ext + [1] * n_dec
    Hq, Hkv, D = 8, 1, 128
    case = synth.make_case_device(len(lens), Hq, Hkv, D, lens, dev, extend_lens=ext, fp8=True, kv_scale=0.1135,
                                  page_size=16, seed=1239)
    r = harness.Runner(case, dev)
    T = case.T
    q = case.q.reshape(T, -1); k = case.k_new.reshape(T, -1); v = case.v_new.reshape(T, -1)
    dec_bytes = synth.algorithmic_decode_bytes(dec_lens, Hq, Hkv, D, 1)
    ext_flops = synth.algorithmic_extend_flops([E] * n_ext, ext_prefix, Hq, D)
    res = {}
    for route in (True, False):
        r.backend.route_decode_tail = route
        fb = r.batch(r.itf.ForwardMode.MIXED)
        r.backend.init_forward_metadata(fb)
        res[route] = t_ms(lambda: r.layer.forward(q, k, v, fb))
    r.backend.route_decode_tail = True
    r.backend.overlap_mixed = True
    fb = r.batch(r.itf.ForwardMode.MIXED)
    r.backend.init_forward_metadata(fb)
    ref_out = None
    r.backend.overlap_mixed = False
    ref_out = r.layer.forward(q, k, v, fb).clone()
    r.backend.overlap_mixed = True
    ov_out = r.layer.forward(q, k, v, fb)
    torch.cuda.synchronize()
    same = torch.equal(ref_out, ov_out)
    res["overlap"] = t_ms(lambda: r.layer.forward(q, k, v, fb))
    print(f"  decode tail on a side stream : {res['overlap']:8.3f} ms per layer   (bit-identical to the serial route: {same})")
    r.backend.overlap_mixed = 2
    ov2 = r.layer.forward(q, k, v, fb)
    torch.cuda.synchronize()
    same2 = torch.equal(ref_out, ov2)
    res["overlap2"] = t_ms(lambda: r.layer.forward(q, k, v, fb))
    print(f"  same, prefill launched first : {res['overlap2']:8.3f} ms per layer   (bit-identical: {same2})")
    print(f"C5 per-rank mixed forward: sum L = {sum(lens)} tokens, decode rows {n_dec}, prefill rows {n_ext} x {E}")
    print(f"  decode tail -> decode kernel : {res[True]:8.3f} ms per layer")
    print(f"  all rows -> extend kernels   : {res[False]:8.3f} ms per layer")
    print(f"  decode part alone would move {dec_bytes / 1e9:.3f} GB ({dec_bytes / res[True] / 1e6:.0f} GB/s if it were the whole forward); "
          f"prefill part is {ext_flops / 1e9:.1f} GFLOP")


if __name__ == "__main__":
    main()
