#!/usr/bin/env python
"""Time one forward_extend at BASELINE config 3 (or a variant); prints TFLOP/s."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from sglang_custom_b200 import _lib, ops, synth


def main():
    import os
    dev = torch.device("cuda:0")
    prof = None
    if os.environ.get("B200ATTN_PROF") == "1":
        prof = torch.zeros(128, dtype=torch.int64, device=dev)
        os.environ["B200ATTN_PROF_PTR"] = hex(prof.data_ptr())
    bs, P, E, Hq, Hkv, D = 16, 2048, 2048, 32, 8, 128
    if len(sys.argv) > 1:
        bs, P, E = (int(x) for x in sys.argv[1:4])
    fp8 = "fp8" in sys.argv
    page = 16 if "page16" in sys.argv else 1
    c = synth.make_case_device(bs, Hq, Hkv, D, [P + E] * bs, dev, extend_lens=[E] * bs, seed=1237, fp8=fp8, page_size=page)
    out = torch.empty_like(c.q)
    fl = synth.algorithmic_extend_flops([E] * bs, [P] * bs, Hq, D)
    fn = lambda: ops.extend(c.q, out, c.k_pool, c.v_pool, c.req_to_token, c.req_pool_indices, c.seq_lens,
                            c.extend_seq_lens, c.extend_prefix_lens, c.extend_start_loc, E, c.sm_scale,
                            kv_dtype=torch.float8_e4m3fn if fp8 else None, impl=_lib.EXTEND_TCGEN05)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    # bit-level checksum of the output: lets two runs with different dev knobs (B200ATTN_EXP_EMU=4 = legacy two-pass
    # softmax schedule, B200ATTN_EXTEND_TC=64) be compared for bit-identity
    chk = int(out.view(torch.int16).to(torch.int64).sum().item())
    print(f"extend bs={bs} P={P} E={E} fp8={fp8} page={page}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s  checksum {chk}", flush=True)
    if prof is not None:
        pr = prof.cpu().view(16, 8)
        n_it = (P + E + 127) // 128
        names = {0: "softmax t0 w0 [wait S, ld S, max, rescale, exp+st+arrive, wait O, epilogue]",
                 4: "softmax t1 w4", 8: "mma [wait QK0, wait P0, wait P1, wait V, wait K, -, -, issue+other]",
                 9: "producer [idx load, wait K_EMPTY, issue K, wait V_EMPTY, issue V, arrive K, arrive V]", 10: "producer 2", 11: "producer 3"}
        for w, nm in names.items():
            print(f"  warp {w}: {nm}\n     total cycles {pr[w].tolist()}  per-iteration {[int(x) // n_it for x in pr[w].tolist()]}")


if __name__ == "__main__":
    main()
