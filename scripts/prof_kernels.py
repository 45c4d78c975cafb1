#!/usr/bin/env python
"""Run each hot kernel a few times at its BASELINE shape (for `ncu -k regex:... -s 2 -c 1` captures).

  python scripts/prof_kernels.py decode|decode_fp8|decode_fp8_tp1|decode_fp8_arange|extend [reps]
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from sglang_custom_b200 import harness, synth


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "decode"
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    dev = torch.device("cuda:0")
    if what == "decode":
        case = synth.make_case_device(256, 32, 8, 128, [4096] * 256, dev, seed=1236)
    elif what == "decode_fp8":  # C4 at TP 8: 8 Q / 1 KV head per rank
        case = synth.make_case_device(512, 8, 1, 128, [8192] * 512, dev, fp8=True, seed=1238)
    elif what == "decode_fp8_tp1":  # C4 on one GPU: 64 Q / 8 KV heads
        case = synth.make_case_device(512, 64, 8, 128, [8192] * 512, dev, fp8=True, seed=1238)
    elif what == "decode_fp8_arange":  # C4 at TP 8 with sequential slots (separates DRAM locality from the SM side)
        case = synth.make_case_device(512, 8, 1, 128, [8192] * 512, dev, fp8=True, seed=1238, table="arange")
    elif what == "extend":
        case = synth.make_case_device(16, 32, 8, 128, [4096] * 16, dev, extend_lens=[2048] * 16, seed=1237)
    else:
        raise SystemExit(f"unknown workload {what}")
    r = harness.Runner(case, dev)
    for _ in range(reps):
        r.forward()
    torch.cuda.synchronize()
    print("ok", what, reps)


if __name__ == "__main__":
    main()
