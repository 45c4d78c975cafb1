"""Generate tests/golden/masked_*.pt by running the UNMODIFIED reference Triton kernels on the CPU.

Only runs in the build container (it loads files from /root/reference through oracle/ref_triton.py, which turns on
Triton's interpreter -- TRITON_INTERPRET=1 -- before triton is imported):
  python tests/golden/make_golden_masked.py
Pins the sliding-window and custom-mask (speculative decoding TARGET_VERIFY) semantics of
  triton_ops/extend_attention.py:42-438, triton_ops/decode_attention.py:677-728, triton_backend.py:927-955
for which the reference keeps no golden vectors and its CPU backends have no argument.  Fixtures are fp16 / fp32:
the interpreter mishandles bf16 (probed).  Each fixture holds the synthetic inputs (sglang_custom_b200.synth), the pool
after the KV write and the reference's output; tests/test_oracle_pin.py replays them through
oracle/triton_masked.py, tests/test_gpu_masked.py through the CUDA path.

Every case stays inside the domain where the reference kernel is well defined: with a window, each query must see a
key in its first prefix block (E - 1 <= min(64, kept prefix) - 1 + W), otherwise the reference returns NaN
(extend_attention.py:178-180, exp(-inf - -inf)).
"""
from __future__ import annotations

import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle import ref_triton  # noqa: E402  (first: sets TRITON_INTERPRET before triton loads)

import torch  # noqa: E402

from oracle import torch_native as tn  # noqa: E402
from oracle import triton_masked as tm  # noqa: E402
from sglang_custom_b200 import synth  # noqa: E402

OUT = Path(__file__).resolve().parent

F16, F32 = torch.float16, torch.float32
# name -> (make_case kwargs, extras)
EXTEND_CASES = {
    "masked_window_extend": (dict(bs=3, Hq=8, Hkv=2, D=128, seq_lens=[90, 42, 25], extend_lens=[20, 30, 25],
                                  dtype=F16, seed=301), dict(window=32)),
    "masked_window_small": (dict(bs=2, Hq=4, Hkv=4, D=64, seq_lens=[52, 133], extend_lens=[12, 5],
                                 dtype=F16, seed=302), dict(window=8)),
    "masked_window_zero": (dict(bs=2, Hq=4, Hkv=1, D=128, seq_lens=[40, 9], extend_lens=[7, 9],
                                dtype=F16, seed=303), dict(window=0)),
    "masked_window_f32": (dict(bs=2, Hq=4, Hkv=2, D=64, seq_lens=[100, 31], extend_lens=[33, 8],
                               dtype=F32, seed=304), dict(window=40)),
    "masked_window_page16": (dict(bs=2, Hq=8, Hkv=1, D=128, seq_lens=[80, 48], extend_lens=[16, 16], page_size=16,
                                  dtype=F16, seed=305), dict(window=24)),
    "masked_tree_verify": (dict(bs=3, Hq=8, Hkv=2, D=128, seq_lens=[58, 25, 41], extend_lens=[8, 8, 8],
                                dtype=F16, seed=311), dict(tree=True)),
    "masked_tree_verify_d64": (dict(bs=2, Hq=4, Hkv=4, D=64, seq_lens=[200, 20], extend_lens=[16, 16],
                                    dtype=F16, seed=312), dict(tree=True)),
    "masked_tree_f32": (dict(bs=2, Hq=4, Hkv=2, D=128, seq_lens=[77, 12], extend_lens=[4, 4],
                             dtype=F32, seed=313), dict(tree=True)),
    "masked_tree_prefixmask": (dict(bs=2, Hq=4, Hkv=2, D=128, seq_lens=[70, 30], extend_lens=[6, 6],
                                    dtype=F16, seed=314), dict(tree=True, prefix_mask=True)),
}
DECODE_CASES = {
    "masked_window_decode": (dict(bs=4, Hq=8, Hkv=2, D=128, seq_lens=[100, 7, 64, 33], dtype=F16, seed=321),
                             dict(window=32)),
    "masked_window_decode_mha": (dict(bs=3, Hq=4, Hkv=4, D=64, seq_lens=[150, 17, 16], dtype=F16, seed=322),
                                 dict(window=15)),
    "masked_window_decode_zero": (dict(bs=2, Hq=4, Hkv=1, D=128, seq_lens=[40, 1], dtype=F16, seed=323),
                                  dict(window=0)),
    "masked_window_decode_f32": (dict(bs=2, Hq=4, Hkv=2, D=64, seq_lens=[90, 50], dtype=F32, seed=324),
                                 dict(window=63)),
}


def _fixture(case, extra: dict) -> dict:
    fx = {k: v for k, v in vars(case).items() if isinstance(v, (torch.Tensor, int, float, bool))}
    fx["dtype_name"] = str(case.dtype).replace("torch.", "")
    fx.update(extra)
    return fx


def build_masks(case, ex: dict, gen: torch.Generator):
    if not ex.get("tree"):
        return None, None
    blocks = []
    for i in range(case.bs):
        P, E = int(case.extend_prefix_lens[i]), int(case.extend_seq_lens[i])
        m = tm.tree_mask(P, E, gen)
        if ex.get("prefix_mask"):
            keep = torch.rand((E, P), generator=gen) < 0.6
            keep[:, 0] = True  # every query keeps a key in its first prefix block (reference NaN domain)
            m[:, :P] = keep
        blocks.append(m)
    return tm.flatten_masks(blocks)


def run_extend(kw: dict, ex: dict) -> dict:
    case = synth.make_case(**kw)
    kp, vp = case.k_pool.clone(), case.v_pool.clone()
    tn.set_kv_buffer(kp, vp, case.out_cache_loc, case.k_new.clone(), case.v_new.clone(), case.dtype)
    gen = torch.Generator().manual_seed(kw["seed"] + 7000)
    mask, indptr = build_masks(case, ex, gen)
    window = ex.get("window", -1)
    skip = not ex.get("prefix_mask", False)
    out = ref_triton.ref_extend(case.q, case.k_new, case.v_new, kp, vp, case.req_to_token, case.req_pool_indices,
                                case.extend_prefix_lens, case.extend_seq_lens, case.sm_scale, True, window, mask,
                                indptr, skip)
    assert not bool(torch.isnan(out).any()), "case left the reference's well-defined domain"
    extra = dict(ref_k_pool_after=kp, ref_v_pool_after=vp, ref_out=out, window=int(window),
                 skip_prefix_custom_mask=bool(skip))
    if mask is not None:
        extra.update(custom_mask=mask, mask_indptr=indptr)
    return _fixture(case, extra)


def run_decode(kw: dict, ex: dict) -> dict:
    case = synth.make_case(**kw)
    kp, vp = case.k_pool.clone(), case.v_pool.clone()
    tn.set_kv_buffer(kp, vp, case.out_cache_loc, case.k_new.clone(), case.v_new.clone(), case.dtype)
    out = ref_triton.ref_decode(case.q, kp, vp, case.req_to_token, case.req_pool_indices, case.seq_lens,
                                case.sm_scale, ex["window"])
    assert not bool(torch.isnan(out).any())
    wi, wx = None, None
    ns = ref_triton.load()
    wi = torch.zeros(case.bs + 1, dtype=torch.int32)
    wi, wx, _ = ns.tb.update_sliding_window_buffer(wi, case.req_to_token, ex["window"], case.seq_lens.to(torch.int32),
                                                   case.req_pool_indices.to(torch.int32), case.bs, "cpu")
    return _fixture(case, dict(ref_k_pool_after=kp, ref_v_pool_after=vp, ref_out=out, window=int(ex["window"]),
                               ref_window_kv_indptr=wi.clone(), ref_window_kv_indices=wx.clone()))


def main() -> None:
    assert ref_triton.available(), "the reference is not mounted"
    for name, (kw, ex) in EXTEND_CASES.items():
        fx = run_extend(kw, ex)
        torch.save(fx, OUT / f"{name}.pt")
        print(f"{name}: out {tuple(fx['ref_out'].shape)} mean|o| {fx['ref_out'].float().abs().mean():.4f}")
    for name, (kw, ex) in DECODE_CASES.items():
        fx = run_decode(kw, ex)
        torch.save(fx, OUT / f"{name}.pt")
        print(f"{name}: out {tuple(fx['ref_out'].shape)} mean|o| {fx['ref_out'].float().abs().mean():.4f}")


if __name__ == "__main__":
    main()
