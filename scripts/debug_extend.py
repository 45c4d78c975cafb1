"""Debug helper: the tcgen05 extend kernel against the register-MMA kernel on one shape.
usage: python scripts/debug_extend.py bs P E [Hq Hkv] [fp8] [page16] [ragged]
With B200ATTN_LIB=sglang_custom_b200/libb200attn_dbg.so the timed-out mbarrier waits are printed (bar index = (addr -
first) / 8 in the Bar enum of extend_tc.cu)."""
import ctypes
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from sglang_custom_b200 import _lib, ops, synth  # noqa: E402

nums = [int(x) for x in sys.argv[1:] if x.lstrip("-").isdigit()]
bs, P, E = nums[:3]
Hq, Hkv = (nums[3], nums[4]) if len(nums) >= 5 else (32, 8)
fp8, page = "fp8" in sys.argv, 16 if "page16" in sys.argv else 1
dev = torch.device("cuda", 0)
if "ragged" in sys.argv:
    g = torch.Generator().manual_seed(3)
    ext = [int(x) for x in torch.randint(1, E + 1, (bs,), generator=g)]
    pre = [int(x) for x in torch.randint(0, P + 1, (bs,), generator=g)]
else:
    ext, pre = [E] * bs, [P] * bs
lens = [a + b for a, b in zip(pre, ext)]
c = synth.make_case_device(bs, Hq, Hkv, 128, lens, dev, extend_lens=ext, seed=1237, fp8=fp8, page_size=page,
                           kv_scale=0.1135 if fp8 else 1.0)
lib = _lib.load()
dump = getattr(lib, "b200attn_debug_wait_dump_extend", None)  # exported by the development builds of the extend kernel only
outs = {}
for name, impl in (("mma", _lib.EXTEND_MMA), ("tc", _lib.EXTEND_TCGEN05)):
    out = torch.zeros_like(c.q)
    lse = torch.zeros((c.q.shape[0], Hq), dtype=torch.float32, device=dev)
    ops.extend(c.q, out, c.k_pool, c.v_pool, c.req_to_token, c.req_pool_indices, c.seq_lens, c.extend_seq_lens,
               c.extend_prefix_lens, c.extend_start_loc, max(ext), c.sm_scale, k_scale=c.kv_scale, v_scale=c.kv_scale,
               kv_dtype=torch.float8_e4m3fn if fp8 else None, impl=impl, lse_out=lse, page_size=page)
    torch.cuda.synchronize()
    outs[name] = (out, lse)
    if dump is not None and name == "tc":
        buf = (ctypes.c_uint * 257)()
        n = dump(buf, 1)
        print(f"{n} timed-out waits")
        for i in range(min(n, 24)):
            b, t, bar, par = buf[1 + 4 * i:5 + 4 * i]
            print(f"   block ({b & 0xffff},{b >> 16}) thread {t} (warp {t // 32}) bar 0x{bar:x} (#{(bar & 0xff) // 8}) parity {par}")
for a in sys.argv:
    if a.startswith("splits="):
        S = int(a.split("=")[1])
        ws = torch.empty(ops.extend_split_workspace_bytes(c.q.shape[0], Hq, 128, S), dtype=torch.uint8, device=dev)
        out = torch.zeros_like(c.q)
        lse = torch.zeros((c.q.shape[0], Hq), dtype=torch.float32, device=dev)
        ops.extend(c.q, out, c.k_pool, c.v_pool, c.req_to_token, c.req_pool_indices, c.seq_lens, c.extend_seq_lens,
                   c.extend_prefix_lens, c.extend_start_loc, max(ext), c.sm_scale, k_scale=c.kv_scale, v_scale=c.kv_scale,
                   kv_dtype=torch.float8_e4m3fn if fp8 else None, lse_out=lse, page_size=page, kv_splits=S, workspace=ws)
        torch.cuda.synchronize()
        d = (out.float() - outs["tc"][0].float()).abs()
        dl = (lse - outs["tc"][1]).abs()
        print(f"splits={S} vs tc: out max {float(d.max()):.5f} mean {float(d.mean()):.6f}   lse max {float(dl.max()):.5f}  nan {int(torch.isnan(out.float()).sum())}")
        import time
        for nm, kw in (("unsplit", {}), (f"splits={S}", dict(kv_splits=S, workspace=ws))):
            fn = lambda: ops.extend(c.q, out, c.k_pool, c.v_pool, c.req_to_token, c.req_pool_indices, c.seq_lens,  # noqa: E731
                                    c.extend_seq_lens, c.extend_prefix_lens, c.extend_start_loc, max(ext), c.sm_scale,
                                    k_scale=c.kv_scale, v_scale=c.kv_scale, kv_dtype=torch.float8_e4m3fn if fp8 else None,
                                    page_size=page, impl=_lib.EXTEND_TCGEN05 if not kw else _lib.EXTEND_AUTO, **kw)
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                fn()
            e1.record()
            torch.cuda.synchronize()
            print(f"   {nm}: {e0.elapsed_time(e1) / 10 * 1e3:.1f} us")
d = (outs["tc"][0].float() - outs["mma"][0].float()).abs()
dl = (outs["tc"][1] - outs["mma"][1]).abs()
print(f"tc vs mma: out max {float(d.max()):.5f} mean {float(d.mean()):.6f}   lse max {float(dl.max()):.5f}  nan {int(torch.isnan(outs['tc'][0].float()).sum())}", flush=True)
