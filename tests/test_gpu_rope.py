"""GPU parity of the fused rotary-embedding + KV-write kernel (b200attn_rope_kv_write) through the host mirror of the
reference's RotaryEmbedding (sglang_custom_b200/rotary.py) -> C ABI.

  NATIVE math mode + table in the tensor dtype : bit-exact with the reference's forward_native (golden fixtures made
      from the unmodified reference, and oracle/rope.py on fresh cases)
  FP32 math mode + fp32 table (the CUDA op's semantics) : bit-exact with the fp32 restatement, and within the
      reference test's tolerance of forward_native
  pool after the fused call : torch.equal with set_kv_buffer(rotated k, v) of oracle #1, bf16 and FP8 (scaled) pools
"""
from __future__ import annotations

import pytest
import torch

import golden_util as gu
import helpers as H
from oracle import rope
from oracle import torch_native as tn
from sglang_custom_b200 import _lib, ops, rotary

pytestmark = pytest.mark.gpu
LLAMA3 = {"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
          "original_max_position_embeddings": 8192}


def _module(hs, rd, max_pos, base, neox, dtype, llama3, device, native):
    kw = dict(device=device, cache_dtype=dtype if native else torch.float32,
              math_mode=_lib.ROPE_NATIVE if native else _lib.ROPE_FP32)
    if llama3:
        return rotary.Llama3RotaryEmbedding(hs, rd, max_pos, base, neox, dtype, 8.0, 1.0, 4.0, 8192, **kw)
    return rotary.RotaryEmbedding(hs, rd, max_pos, base, neox, dtype, **kw)


@pytest.mark.parametrize("name", gu.ROPE_FIXTURES)
def test_rope_golden_fixtures_bit_exact(name, device):
    fx = gu.load(name)
    m = _module(fx.head_size, fx.rotary_dim, fx.max_position, fx.base, fx.is_neox, fx.dtype, fx.llama3, device, True)
    assert torch.equal(m.cos_sin_cache.index_select(0, fx.positions.to(device)).cpu(), fx.ref_cache_rows)
    q, k = fx.q.to(device), fx.k.to(device)
    q2, k2 = m(fx.positions.to(device), q, k)
    assert q2.data_ptr() == q.data_ptr()  # in place, as the reference's CUDA op
    assert torch.equal(q.cpu(), fx.ref_q) and torch.equal(k.cpu(), fx.ref_k)
    # fp32 mode: within the reference test's tolerance of forward_native
    m32 = _module(fx.head_size, fx.rotary_dim, fx.max_position, fx.base, fx.is_neox, fx.dtype, fx.llama3, device, False)
    q, k = fx.q.to(device), fx.k.to(device)
    m32(fx.positions.to(device), q, k)
    assert torch.allclose(q.cpu().float(), fx.ref_q.float(), atol=3e-2, rtol=1e-2)
    assert torch.allclose(k.cpu().float(), fx.ref_k.float(), atol=3e-2, rtol=1e-2)


@pytest.mark.parametrize("hs,rd,neox,llama3,dtype", [(128, 128, True, True, torch.bfloat16), (128, 64, False, False, torch.bfloat16),
                                                     (64, 64, True, False, torch.float16), (128, 32, True, False, torch.bfloat16),
                                                     (128, 128, False, False, torch.float16)])
@pytest.mark.parametrize("native", [True, False])
@pytest.mark.parametrize("pool", ["none", "same", "fp8", "fp8_scaled"])
def test_fused_rope_kv_write(hs, rd, neox, llama3, dtype, native, pool, device):
    g = torch.Generator().manual_seed(hs * 7 + rd + int(neox))
    T, Hq, Hkv, P, max_pos = 67, 8, 2, 200, 8192
    base = 500000 if llama3 else 10000
    # q, k, v as strided views of one qkv buffer, as LlamaAttention hands them over (models/llama.py:182-193)
    qkv = torch.randn((T, (Hq + 2 * Hkv) * hs), generator=g).to(dtype)
    pos = torch.randint(0, max_pos, (T,), generator=g)
    loc = (torch.randperm(P - 1, generator=g)[:T] + 1).to(torch.int64)
    q, k, v = qkv.split([Hq * hs, Hkv * hs, Hkv * hs], dim=1)
    table = rope.cos_sin_cache(rd, max_pos, base, LLAMA3 if llama3 else None)
    if native:
        want_q, want_k = rope.rope_native(pos, q.clone(), k.clone(), table.to(dtype), hs, neox)
    else:
        want_q, want_k = rope.rope_fp32(pos, q.clone(), k.clone(), table, hs, neox)
    m = _module(hs, rd, max_pos, base, neox, dtype, llama3, device, native)
    d_qkv = qkv.to(device)
    dq, dk, dv = d_qkv.split([Hq * hs, Hkv * hs, Hkv * hs], dim=1)
    if pool == "none":
        m(pos.to(device), dq, dk)
    else:
        fp8 = pool.startswith("fp8")
        scale = 0.1135 if pool == "fp8_scaled" else None
        pool_dtype = torch.float8_e4m3fn if fp8 else dtype
        store = torch.uint8 if fp8 else dtype
        kp = torch.zeros((P, Hkv, hs), dtype=store)
        vp = torch.zeros((P, Hkv, hs), dtype=store)
        tn.set_kv_buffer(kp, vp, loc, want_k.clone().view(T, Hkv, hs), v.clone().view(T, Hkv, hs), pool_dtype, scale, scale)
        dkp, dvp = torch.zeros_like(kp, device=device), torch.zeros_like(vp, device=device)
        ops.rope_kv_write(pos.to(device), dq.view(T, Hq, hs), dk.view(T, Hkv, hs), m.cos_sin_cache, neox,
                          v=dv.view(T, Hkv, hs), loc=loc.to(device), k_pool=dkp, v_pool=dvp, k_scale=scale, v_scale=scale,
                          pool_dtype=pool_dtype, math_mode=m.math_mode)
        assert H.pools_equal(dkp, kp) and H.pools_equal(dvp, vp), "pool after the fused write"
    torch.cuda.synchronize()
    assert torch.equal(dq.cpu(), want_q) and torch.equal(dk.cpu(), want_k), "rotated q/k"
    assert torch.equal(dv.cpu(), v), "v must be untouched"


def test_fused_path_through_the_backend_equals_two_step_path(device):
    """forward_fused_kv + attention(save_kv_cache=False) == rotary_emb + attention(save_kv_cache=True)."""
    from sglang_custom_b200 import synth
    case = synth.make_case(3, 8, 2, 128, [300, 45, 129], seed=55)
    pos = (case.seq_lens - 1).to(torch.int64)
    outs = []
    for fused in (False, True):
        r = H.Runner(case, device)
        m = rotary.RotaryEmbedding(128, 128, 4096, 10000, True, torch.bfloat16, device=device)
        fb = r.batch()
        r.backend.init_forward_metadata(fb)
        q = case.q.reshape(case.T, -1).to(device).clone()
        k = case.k_new.reshape(case.T, -1).to(device).clone()
        v = case.v_new.reshape(case.T, -1).to(device).clone()
        if fused:
            m.forward_fused_kv(pos.to(device), q, k, v, r.layer, fb)
            out = r.layer.forward(q, k, v, fb, save_kv_cache=False)
        else:
            m(pos.to(device), q, k)
            out = r.layer.forward(q, k, v, fb, save_kv_cache=True)
        torch.cuda.synchronize()
        outs.append((out.cpu(), r.pools()[0].cpu().clone(), r.pools()[1].cpu().clone()))
    assert torch.equal(outs[0][1], outs[1][1]) and torch.equal(outs[0][2], outs[1][2])
    assert torch.equal(outs[0][0], outs[1][0])


def test_rope_argument_errors(device):
    q = torch.zeros((4, 2, 120), dtype=torch.bfloat16, device=device)
    k = torch.zeros((4, 1, 120), dtype=torch.bfloat16, device=device)
    cache = torch.zeros((16, 120), dtype=torch.float32, device=device)
    pos = torch.zeros(4, dtype=torch.int64, device=device)
    with pytest.raises(_lib.B200AttnError, match="multiples of 16"):
        ops.rope_kv_write(pos, q, k, cache, True)
