"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference.

Only runs in the build container (it imports files from /root/reference through oracle/ref_loader.py):
  python tests/golden/make_golden.py
The reference stores no golden vectors for the attention path, so these
fixtures -- inputs plus the outputs of the reference's own code on them -- are the pin:
  MHATokenToKVPool.set_kv_buffer             python/sglang/srt/mem_cache/memory_pool.py:420-458
  TorchNativeAttnBackend._run_sdpa_forward_decode   layers/attention/torch_native_backend.py:112-180
  TorchNativeAttnBackend._run_sdpa_forward_extend   layers/attention/torch_native_backend.py:27-110
Each fixture is a small torch.save()d dict of CPU tensors (inputs, the pool after the reference's KV
write, the reference's attention output).  tests/test_oracle_pin.py replays them through oracle/ and
tests/test_gpu_golden.py through the CUDA path.
"""
from __future__ import annotations

import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle import ref_loader  # noqa: E402
from sglang_custom_b200 import synth  # noqa: E402

OUT = Path(__file__).resolve().parent

# name -> make_case kwargs
CASES = {
    # decode
    "decode_c1_mini": dict(bs=4, Hq=8, Hkv=2, D=64, seq_lens=[33, 17, 48, 1], seed=101),
    "decode_gqa4_d128": dict(bs=3, Hq=8, Hkv=2, D=128, seq_lens=[100, 7, 64], seed=102),
    "decode_mha_d128": dict(bs=2, Hq=4, Hkv=4, D=128, seq_lens=[19, 35], seed=103),
    "decode_gqa8_page16": dict(bs=3, Hq=8, Hkv=1, D=128, seq_lens=[40, 16, 33], page_size=16, seed=104),
    "decode_arange_table": dict(bs=2, Hq=4, Hkv=2, D=64, seq_lens=[50, 21], table="arange", seed=105),
    "decode_f16": dict(bs=2, Hq=4, Hkv=2, D=128, seq_lens=[30, 45], dtype=torch.float16, seed=106),
    # extend (prefix + new tokens, causal)
    "extend_prefix": dict(bs=2, Hq=8, Hkv=2, D=128, seq_lens=[40, 25], extend_lens=[16, 25], seed=201),
    "extend_no_prefix": dict(bs=3, Hq=4, Hkv=2, D=64, seq_lens=[20, 1, 33], extend_lens=[20, 1, 33], seed=202),
    "extend_page16": dict(bs=2, Hq=8, Hkv=1, D=128, seq_lens=[48, 37], extend_lens=[16, 5], page_size=16, seed=203),
    "extend_mixed": dict(bs=4, Hq=8, Hkv=2, D=128, seq_lens=[70, 31, 55, 12], extend_lens=[1, 31, 1, 7], seed=204),
}


def run_case(kw: dict) -> dict:
    case = synth.make_case(**kw)
    kp, vp = case.k_pool.clone(), case.v_pool.clone()
    ref_loader.ref_set_kv_buffer(kp, vp, case.out_cache_loc, case.k_new.clone(), case.v_new.clone(), case.dtype)
    if case.decode:
        out = ref_loader.ref_decode(case.q, kp, vp, case.req_to_token, case.req_pool_indices, case.seq_lens,
                                    case.sm_scale)
    else:
        out = ref_loader.ref_extend(case.q, kp, vp, case.req_to_token, case.req_pool_indices, case.seq_lens,
                                    case.extend_prefix_lens, case.extend_seq_lens, case.sm_scale, causal=True)
    fx = {k: v for k, v in vars(case).items() if isinstance(v, (torch.Tensor, int, float, bool))}
    fx["dtype_name"] = str(case.dtype).replace("torch.", "")
    fx["ref_k_pool_after"] = kp
    fx["ref_v_pool_after"] = vp
    fx["ref_out"] = out
    return fx


def fp8_write_case(seed: int, scale) -> dict:
    """set_kv_buffer into an fp8 pool, with values that exercise every branch of the cast."""
    g = torch.Generator().manual_seed(seed)
    T, H, D, P = 24, 2, 64, 64
    k = (torch.randn((T, H, D), generator=g) * 3.0).to(torch.bfloat16)
    v = (torch.randn((T, H, D), generator=g) * 0.02).to(torch.bfloat16)
    special = torch.tensor([0.0, -0.0, 448.0, 449.0, 464.0, 465.0, 479.0, 480.0, 1e4, -1e4, float("inf"), float("-inf"),
                            float("nan"), 2.0 ** -9, 2.0 ** -10, 1.5 * 2.0 ** -10, 2.0 ** -6, 0.9 * 2.0 ** -6, 1e-8, -1e-8,
                            0.013671875, 0.015625, 0.017578125, 240.0, 247.9, 256.0, 1.0625, 1.1875, -1.0625, 3.25],
                           dtype=torch.float32).to(torch.bfloat16)
    k.view(-1)[: special.numel()] = special
    v.view(-1)[: special.numel()] = special * (scale if scale else 1.0)
    loc = torch.randperm(P - 1, generator=g)[:T].to(torch.int64) + 1
    loc[-1] = 0  # padded CUDA-graph rows write the dummy slot 0
    kp = torch.zeros((P, H, D), dtype=torch.uint8)
    vp = torch.zeros((P, H, D), dtype=torch.uint8)
    k_in, v_in = k.clone(), v.clone()
    ref_loader.ref_set_kv_buffer(kp, vp, loc, k_in, v_in, torch.float8_e4m3fn, scale, scale)
    return dict(k=k, v=v, loc=loc, scale=0.0 if scale is None else float(scale), ref_k_pool_after=kp,
                ref_v_pool_after=vp, ref_k_after_call=k_in, ref_v_after_call=v_in)


LLAMA3_SCALING = {"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                  "original_max_position_embeddings": 8192}
# name -> head_size, rotary_dim, max_position, base, is_neox, dtype, rope_scaling, Hq, Hkv, T
ROPE_CASES = {
    "rope_neox_llama3": (128, 128, 16384, 500000, True, torch.bfloat16, LLAMA3_SCALING, 8, 2, 19),
    "rope_neox_d64": (64, 64, 2048, 10000, True, torch.bfloat16, None, 4, 4, 11),
    "rope_gptj_partial": (128, 64, 2048, 10000, False, torch.bfloat16, None, 4, 1, 13),
    "rope_neox_f16": (128, 128, 4096, 10000, True, torch.float16, None, 2, 2, 7),
}


def rope_case(seed: int, hs, rd, max_pos, base, neox, dtype, scaling, Hq, Hkv, T) -> dict:
    """The reference's RotaryEmbedding.forward_native (layers/rotary_embedding.py:138-166) on random q/k, with the
    table the reference module itself built (in `dtype`, as it keeps it off CUDA, :101-102)."""
    g = torch.Generator().manual_seed(seed)
    ref = ref_loader.ref_rotary_embedding(hs, rd, max_pos, base, neox, dtype, scaling)
    q = torch.randn((T, Hq * hs), generator=g).to(dtype)
    k = torch.randn((T, Hkv * hs), generator=g).to(dtype)
    pos = torch.randint(0, max_pos, (T,), generator=g)
    pos[0], pos[-1] = 0, max_pos - 1
    rq, rk = ref.forward_native(pos, q.clone(), k.clone())
    return dict(head_size=hs, rotary_dim=rd, max_position=max_pos, base=float(base), is_neox=bool(neox),
                dtype_name=str(dtype).replace("torch.", ""), llama3=scaling is not None, Hq=Hq, Hkv=Hkv,
                q=q, k=k, positions=pos, ref_cache_rows=ref.cos_sin_cache.index_select(0, pos), ref_q=rq, ref_k=rk)


def main() -> None:
    torch.set_num_threads(1)
    for i, (name, args) in enumerate(ROPE_CASES.items()):
        torch.save(rope_case(401 + i, *args), OUT / f"{name}.pt")
        print("wrote", name)
    for name, kw in CASES.items():
        torch.save(run_case(kw), OUT / f"{name}.pt")
        print("wrote", name)
    for name, scale in (("kvwrite_fp8_noscale", None), ("kvwrite_fp8_scale", 0.1135), ("kvwrite_fp8_scale2", 2.5)):
        torch.save(fp8_write_case(301, scale), OUT / f"{name}.pt")
        print("wrote", name)


if __name__ == "__main__":
    main()
