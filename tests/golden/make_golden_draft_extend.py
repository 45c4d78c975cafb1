"""Generate tests/golden/draft_extend_*.pt: what the UNMODIFIED reference computes for a DRAFT_EXTEND batch, both ways.

Only runs in the build container (loads /root/reference through oracle/ref_triton.py, Triton interpreter on the CPU):
  python tests/golden/make_golden_draft_extend.py

After a verify step the draft model extends request i by its accept_length[i] accepted tokens; forward_batch.seq_lens
already INCLUDES them (speculative/eagle_utils.py:107-133).  The reference has two answers for the attention of that
batch, depending on the backend:

  ref_out_causal  -- the default (FlashInfer) backend: qo_indptr = cumsum(accept_length), paged_kernel_lens = seq_lens,
                     causal (eagle_utils.py:135-164 + flashinfer_backend.py prefill updater): query t of request i sees
                     the keys at positions <= seq_lens[i] - accept_length[i] + t.  A plain causal extend whose prefix is
                     seq_lens - accept_length; made here with the reference's own extend_attention_fwd on that prefix.
  ref_out_triton  -- the Triton backend (triton_backend.py:265-281): generate_attn_arg_prefill flattens ALL seq_lens[i]
                     slots -- the accepted tokens included -- into kv_indices, which extend_attention_fwd treats as an
                     unmasked prefix, and then adds k_extend / v_extend (the same accepted tokens) under the causal
                     mask: every query sees all accepted tokens once unmasked and the ones up to itself a second time.

B200AttnBackend implements the first; tests/test_oracle_pin_masked.py and tests/test_gpu_draft.py pin
both numbers, so the difference is on record.  fp16 / fp32 only: the interpreter mishandles bf16.
"""
from __future__ import annotations

import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle import ref_triton  # noqa: E402  (first: sets TRITON_INTERPRET before triton loads)

import torch  # noqa: E402

from oracle import torch_native as tn  # noqa: E402
from sglang_custom_b200 import synth  # noqa: E402

OUT = Path(__file__).resolve().parent

CASES = {
    # seq_lens include the accepted tokens; extend_lens = accept_length
    "draft_extend_f16": dict(bs=4, Hq=8, Hkv=2, D=128, seq_lens=[70, 9, 133, 6], extend_lens=[4, 1, 6, 6],
                             dtype=torch.float16, seed=801),
    "draft_extend_page16_f32": dict(bs=3, Hq=4, Hkv=1, D=128, seq_lens=[50, 17, 96], extend_lens=[3, 5, 2],
                                    page_size=16, dtype=torch.float32, seed=802),
    "draft_extend_d64": dict(bs=2, Hq=4, Hkv=4, D=64, seq_lens=[40, 200], extend_lens=[2, 4],
                             dtype=torch.float16, seed=803),
}


def run(kw: dict) -> dict:
    case = synth.make_case(**kw)
    kp, vp = case.k_pool.clone(), case.v_pool.clone()
    tn.set_kv_buffer(kp, vp, case.out_cache_loc, case.k_new.clone(), case.v_new.clone(), case.dtype)
    accept = case.extend_seq_lens
    args = (case.q, case.k_new, case.v_new, kp, vp, case.req_to_token, case.req_pool_indices)
    causal = ref_triton.ref_extend(*args, case.seq_lens - accept, accept, case.sm_scale, True)
    triton = ref_triton.ref_extend(*args, case.seq_lens, accept, case.sm_scale, True)
    assert not bool(torch.isnan(causal).any()) and not bool(torch.isnan(triton).any())
    fx = {k: v for k, v in vars(case).items() if isinstance(v, (torch.Tensor, int, float, bool))}
    fx["dtype_name"] = str(case.dtype).replace("torch.", "")
    fx.update(ref_k_pool_after=kp, ref_v_pool_after=vp, ref_out_causal=causal, ref_out_triton=triton,
              accept_length=accept.clone())
    return fx


def main() -> None:
    assert ref_triton.available(), "the reference is not mounted"
    for name, kw in CASES.items():
        fx = run(kw)
        torch.save(fx, OUT / f"{name}.pt")
        d = (fx["ref_out_causal"].float() - fx["ref_out_triton"].float()).abs()
        print(f"{name}: out {tuple(fx['ref_out_causal'].shape)}  max|causal - triton| {float(d.max()):.4f} "
              f"mean {float(d.mean()):.5f}")


if __name__ == "__main__":
    main()
