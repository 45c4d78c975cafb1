#!/usr/bin/env python
"""Time the HiCache KV transfer kernel: host (pinned) -> device, device -> host, device -> device; GB/s of payload
(K + V bytes moved once).  Llama-3-8B token items: 8 heads x 128 x bf16 = 2 KiB per token per K/V per layer."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from sglang_custom_b200 import kvcacheio as io


def t_ms(fn, reps=10):
    for _ in range(2):
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
    L, total, item = 8, 65536, 1024  # bf16 elements: 2 KiB items, 128 MiB per layer per buffer
    n = 32768
    ib = item * 2
    hk = torch.zeros((L, total, item), dtype=torch.bfloat16).pin_memory()
    hv = torch.zeros((L, total, item), dtype=torch.bfloat16).pin_memory()
    dk = torch.zeros((L, total, item), dtype=torch.bfloat16, device=dev)
    dv = torch.zeros((L, total, item), dtype=torch.bfloat16, device=dev)
    dk2, dv2 = torch.zeros_like(dk), torch.zeros_like(dv)
    si = torch.randperm(total, device=dev)[:n]
    di = torch.randperm(total, device=dev)[:n]
    tbl = lambda p: torch.tensor([p[l].data_ptr() for l in range(L)], dtype=torch.uint64, device=dev)
    payload_layer = 2 * n * ib
    for quota in (2, 8, 0):
        ms = t_ms(lambda: io.transfer_kv_per_layer(hk[0], dk[0], hv[0], dv[0], si, di, "kernel", 1, item, block_quota=quota))
        print(f"host->device one layer   quota {quota}: {ms:8.3f} ms  {payload_layer / ms / 1e6:8.1f} GB/s")
    for quota in (2, 8, 0):
        ms = t_ms(lambda: io.transfer_kv_all_layer(tbl(dk), tbl(hk), tbl(dv), tbl(hv), si, di, "kernel", ib, L, block_quota=quota), reps=3)
        print(f"device->host {L} layers     quota {quota}: {ms:8.3f} ms  {L * payload_layer / ms / 1e6:8.1f} GB/s")
    for quota in (2, 0):
        ms = t_ms(lambda: io.transfer_kv_all_layer(tbl(dk), tbl(dk2), tbl(dv), tbl(dv2), si, di, "kernel", ib, L, block_quota=quota))
        print(f"device->device {L} layers   quota {quota}: {ms:8.3f} ms  {L * payload_layer / ms / 1e6:8.1f} GB/s (payload; HBM traffic is 2x)")


if __name__ == "__main__":
    main()
