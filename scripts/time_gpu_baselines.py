#!/usr/bin/env python
"""Same-box GPU baselines for the attention path (test infrastructure; nothing here is on the product path).

  python scripts/time_gpu_baselines.py            # every baseline, each in its own subprocess with a time limit
  python scripts/time_gpu_baselines.py --one NAME # one baseline in this process, one JSON line on stdout

What can and cannot be timed on the GPU box:
  * The reference's own Triton kernels (triton_ops/decode_attention.py:677-728, extend_attention.py:306-438) live in
    /root/reference, which does not exist on the GPU box and whose sources must not be copied into the repo -- so they
    cannot be run there.  `ref_triton_*` reports "unavailable" unless B200ATTN_REFERENCE points at a reference tree.
  * `vllm_triton_decode`: the image's vLLM ships `vllm/v1/attention/ops/triton_decode_attention.py`, which its header
    says is adapted from SGLang's `triton_ops/decode_attention.py` (same two-stage grouped split-KV kernels, reading
    req_to_token directly).  It is the closest runnable stand-in for the reference's Triton decode design.
  * `flashinfer_*`: FlashInfer is the reference's DEFAULT GPU backend (flashinfer_backend.py:155-160 picks
    fmha_backend on sm_100); the image has flashinfer 0.6.11.  Paged decode / prefill wrappers with page_size 1 and
    flattened kv_indices, exactly how FlashInferAttnBackend drives them (flashinfer_backend.py:560-567).
  * `b200_*`: this repo's kernels at the same shapes, same process, same CUDA-event method.
Shapes: C2 (Llama-3-8B decode bs 256 ctx 4096 bf16), C4/TP8 (8 Q / 1 KV head, FP8 KV, bs 512 ctx 8192), C3 (extend
bs 16, prefix 2048 + new 2048, 32/8 heads).  Timing: 3 warm-ups, 10 launches between CUDA events, inputs >> L2.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

NAMES = ["b200_decode_c2", "vllm_triton_decode_c2", "flashinfer_decode_c2", "flashinfer_decode_c2_tc",
         "b200_decode_c4tp8", "flashinfer_decode_c4tp8", "flashinfer_decode_c4tp8_tc",
         "b200_extend_c3", "flashinfer_prefill_c3", "flash_attn_varlen_c3", "ref_triton_decode_c2"]


def _t_ms(fn, reps=10, warm=3):
    import torch
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


def _decode_case(which):
    import torch
    from sglang_custom_b200 import synth
    dev = torch.device("cuda:0")
    if which == "c2":
        bs, ctx, Hq, Hkv, fp8 = 256, 4096, 32, 8, False
    else:
        bs, ctx, Hq, Hkv, fp8 = 512, 8192, 8, 1, True
    case = synth.make_case_device(bs, Hq, Hkv, 128, [ctx] * bs, dev, fp8=fp8, kv_scale=0.1135 if fp8 else 1.0, seed=1236)
    kvb = 1 if fp8 else 2
    nbytes = bs * ctx * Hkv * 2 * 128 * kvb + bs * ctx * 4 + 2 * bs * Hq * 128 * 2
    return case, dev, nbytes


def _flat_indices(case):
    import torch
    rows = case.req_to_token[case.req_pool_indices]  # [bs, ctx], full rows (every request has ctx tokens)
    L = int(case.seq_lens[0])
    idx = rows[:, :L].reshape(-1).contiguous().to(torch.int32)
    indptr = torch.arange(0, case.bs + 1, device=idx.device, dtype=torch.int32) * L
    return indptr, idx


def run_one(name: str) -> dict:
    import torch
    if name.startswith("ref_triton"):
        ref = os.environ.get("B200ATTN_REFERENCE", "/root/reference")
        if not os.path.isdir(os.path.join(ref, "python", "sglang")):
            return {"name": name, "unavailable": "the reference sources are not on the GPU box (they may not be copied "
                                                 "into the repo); see vllm_triton_decode_c2 for the vendored descendant"}
        return {"name": name, "unavailable": "reference present but no GPU loader is wired for it in this script"}

    if "decode" in name:
        which = "c2" if name.endswith("c2") or name.endswith("c2_tc") else "c4tp8"
        case, dev, nbytes = _decode_case(which)
        bs, Hq, Hkv, D = case.bs, case.Hq, case.Hkv, case.D
        out = torch.empty_like(case.q)
        if name.startswith("b200"):
            from sglang_custom_b200 import _lib, ops
            n_ctas = _lib.device_info()["sm_count"]
            plan = torch.zeros(ops.decode_plan_bytes(bs, n_ctas) // 4, dtype=torch.int32, device=dev)
            ws = torch.empty(ops.decode_workspace_bytes(n_ctas, Hq, D) // 4, dtype=torch.float32, device=dev)
            s = case.kv_scale
            kvd = torch.float8_e4m3fn if case.fp8 else None

            def fn():
                ops.decode_plan(case.seq_lens, n_ctas, plan, 2)
                ops.decode(case.q, out, case.k_pool, case.v_pool, case.req_to_token, case.req_pool_indices,
                           case.seq_lens, plan, n_ctas, ws, case.sm_scale, 0.0, s, s, kv_dtype=kvd)
            what = "plan + split-KV decode + merge (libb200attn)"
        elif name.startswith("vllm_triton"):
            from vllm.v1.attention.ops.triton_decode_attention import decode_attention_fwd
            splits = 8  # SGLang's default triton_attention_num_kv_splits (server_args.py)
            # this kernel indexes req_to_token by batch position: give it the rows in batch order
            table = case.req_to_token[case.req_pool_indices].contiguous()
            logits = torch.empty((bs, Hq, splits, D + 1), dtype=torch.float32, device=dev)
            lse = torch.empty((bs, Hq), dtype=torch.float32, device=dev)
            kp, vp = case.k_pool.unsqueeze(1), case.v_pool.unsqueeze(1)  # [P, page_size = 1, Hkv, D]
            sl = case.seq_lens.to(torch.int32)

            def fn():
                decode_attention_fwd(case.q, kp, vp, out, lse, table, sl, logits, splits, case.sm_scale, 1)
            what = "vLLM's vendored descendant of SGLang triton_ops/decode_attention.py, 8 kv splits"
        else:
            import flashinfer
            tc = name.endswith("_tc")
            wsb = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
            w = flashinfer.BatchDecodeWithPagedKVCacheWrapper(wsb, "NHD", use_tensor_cores=tc)
            indptr, idx = _flat_indices(case)
            last = torch.ones(bs, dtype=torch.int32, device=dev)
            kvd = torch.float8_e4m3fn if case.fp8 else torch.bfloat16
            w.plan(indptr, idx, last, Hq, Hkv, D, 1, q_data_type=torch.bfloat16, kv_data_type=kvd, sm_scale=case.sm_scale)
            kp = case.k_pool.view(kvd).unsqueeze(1) if case.fp8 else case.k_pool.unsqueeze(1)
            vp = case.v_pool.view(kvd).unsqueeze(1) if case.fp8 else case.v_pool.unsqueeze(1)
            kw = dict(k_scale=case.kv_scale, v_scale=case.kv_scale) if case.fp8 else {}

            def fn():
                w.run(case.q, (kp, vp), out=out, **kw)
            what = f"flashinfer {flashinfer.__version__} BatchDecodeWithPagedKVCacheWrapper, page_size 1, use_tensor_cores={tc} (plan outside the timed region)"
        ms = _t_ms(fn)
        return {"name": name, "us": round(ms * 1e3, 1), "gbs": round(nbytes / ms / 1e6, 1), "bytes": nbytes, "what": what}

    # ---- extend C3
    from sglang_custom_b200 import synth
    dev = torch.device("cuda:0")
    bs, P, E, Hq, Hkv, D = 16, 2048, 2048, 32, 8, 128
    case = synth.make_case_device(bs, Hq, Hkv, D, [P + E] * bs, dev, extend_lens=[E] * bs, seed=1237)
    flops = synth.algorithmic_extend_flops([E] * bs, [P] * bs, Hq, D)
    out = torch.empty_like(case.q)
    if name.startswith("b200"):
        from sglang_custom_b200 import ops

        def fn():
            ops.extend(case.q, out, case.k_pool, case.v_pool, case.req_to_token, case.req_pool_indices, case.seq_lens,
                       case.extend_seq_lens, case.extend_prefix_lens, case.extend_start_loc, E, case.sm_scale)
        what = "tcgen05 extend kernel (libb200attn), page table read in place"
    elif name.startswith("flashinfer"):
        import flashinfer
        wsb = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
        w = flashinfer.BatchPrefillWithPagedKVCacheWrapper(wsb, "NHD")
        indptr, idx = _flat_indices(case)
        qo = torch.arange(0, bs + 1, device=dev, dtype=torch.int32) * E
        last = torch.ones(bs, dtype=torch.int32, device=dev)
        w.plan(qo, indptr, idx, last, Hq, Hkv, D, 1, causal=True, sm_scale=case.sm_scale, q_data_type=torch.bfloat16,
               kv_data_type=torch.bfloat16)
        kp, vp = case.k_pool.unsqueeze(1), case.v_pool.unsqueeze(1)

        def fn():
            w.run(case.q, (kp, vp), out=out)
        what = f"flashinfer {flashinfer.__version__} BatchPrefillWithPagedKVCacheWrapper (backend auto: {getattr(w, '_backend', '?')}), page_size 1, causal"
    else:
        from flash_attn import flash_attn_varlen_func
        # contiguous K/V per request (no paging): an upper bound for a gather-free FA2 kernel on this part
        rows = case.req_to_token[case.req_pool_indices][:, :P + E].reshape(-1).long()
        k = case.k_pool[rows].contiguous()
        v = case.v_pool[rows].contiguous()
        cq = torch.arange(0, bs + 1, device=dev, dtype=torch.int32) * E
        ck = torch.arange(0, bs + 1, device=dev, dtype=torch.int32) * (P + E)

        def fn():
            flash_attn_varlen_func(case.q, k, v, cq, ck, E, P + E, softmax_scale=case.sm_scale, causal=True)
        what = "flash_attn 2.8.3 varlen (mma.sync FA2 kernel, CONTIGUOUS K/V: no page gather)"
    ms = _t_ms(fn)
    return {"name": name, "ms": round(ms, 4), "tflops": round(flops / ms / 1e9, 1), "flops": flops, "what": what}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--one", default=None)
    ap.add_argument("--timeout", type=int, default=420)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    if a.one:
        try:
            res = run_one(a.one)
        except Exception as e:
            res = {"name": a.one, "error": f"{type(e).__name__}: {e}"[:400]}
        print(json.dumps(res), flush=True)
        return
    lines = []
    for n in NAMES:
        try:
            p = subprocess.run([sys.executable, __file__, "--one", n], capture_output=True, text=True, timeout=a.timeout)
            js = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
            res = json.loads(js[-1]) if js else {"name": n, "error": (p.stderr or "no output")[-300:]}
        except subprocess.TimeoutExpired:
            res = {"name": n, "error": f"timeout after {a.timeout} s (JIT build?)"}
        lines.append(res)
        print(json.dumps(res), flush=True)
    if a.out:
        Path(a.out).write_text("\n".join(json.dumps(x) for x in lines) + "\n")


if __name__ == "__main__":
    main()
