"""GPU parity of the KV-write kernel (b200attn_kv_write) against the reference's set_kv_buffer
(memory_pool.py:420-458): the WHOLE pool is compared bit for bit, untouched rows included."""
from __future__ import annotations

import pytest
import torch

import golden_util as gu
import helpers as H
from oracle import exact
from oracle import torch_native as tn
from sglang_custom_b200 import _lib, ops, synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", gu.KVWRITE_FIXTURES)
def test_fp8_write_matches_reference_golden(name, device):
    fx = gu.load(name)
    scale = None if fx.scale == 0.0 else fx.scale
    kp = torch.zeros_like(fx.ref_k_pool_after, device=device)
    vp = torch.zeros_like(fx.ref_v_pool_after, device=device)
    # slot 0 receives the last token in the fixture (one writer), so the result is deterministic
    k, v = fx.k.to(device), fx.v.to(device)
    ops.set_kv_buffer(kp, vp, fx.loc.to(device), k, v, torch.float8_e4m3fn, scale, scale, writeback_scaled=True)
    torch.cuda.synchronize()
    assert H.pools_equal(kp, fx.ref_k_pool_after) and H.pools_equal(vp, fx.ref_v_pool_after)
    # the reference's in-place div_ side effect on the caller's tensors (memory_pool.py:437-440)
    assert torch.equal(H.canon_16(k), H.canon_16(fx.ref_k_after_call))
    assert torch.equal(H.canon_16(v), H.canon_16(fx.ref_v_after_call))


def test_fp8_cast_exhaustive_all_bf16_patterns(device):
    bits = torch.arange(65536, dtype=torch.int32).to(torch.int16)
    x = bits.view(torch.bfloat16).reshape(1024, 1, 64)
    loc = torch.arange(1024, dtype=torch.int64)
    for scale in (None, 0.1135, 2.5, 0.04):
        y = x.clone()
        if scale is not None:
            y.div_(scale)
        want = y.to(torch.float8_e4m3fn).view(torch.uint8)
        kp = torch.zeros((1024, 1, 64), dtype=torch.uint8, device=device)
        vp = torch.zeros_like(kp)
        ops.set_kv_buffer(kp, vp, loc.to(device), x.to(device), x.to(device), torch.float8_e4m3fn, scale, scale)
        assert H.pools_equal(kp, want) and H.pools_equal(vp, want), f"scale={scale}"


def test_fp8_cast_exhaustive_f16_input(device):
    bits = torch.arange(65536, dtype=torch.int32).to(torch.int16)
    x = bits.view(torch.float16).reshape(1024, 1, 64)
    want = x.to(torch.float8_e4m3fn).view(torch.uint8)
    kp = torch.zeros((1024, 1, 64), dtype=torch.uint8, device=device)
    vp = torch.zeros_like(kp)
    ops.set_kv_buffer(kp, vp, torch.arange(1024, device=device), x.to(device), x.to(device), torch.float8_e4m3fn)
    assert H.pools_equal(kp, want)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("Hkv,D,T,strided", [(8, 128, 256, True), (2, 64, 33, False), (1, 128, 1, True), (4, 128, 4096, True)])
def test_bf16_scatter_whole_pool_bit_exact(dtype, Hkv, D, T, strided, device):
    c = synth.make_case(T, Hkv * 4, Hkv, D, [3] * T, seed=T + Hkv, dtype=dtype, pool_slack=500)
    kp, vp = H.pool_after_write(c)
    dk, dv = c.k_pool.to(device), c.v_pool.to(device)
    if strided:  # k, v as views of one qkv buffer (models/llama.py:182-193)
        qkv = torch.cat([c.q.reshape(T, -1), c.k_new.reshape(T, -1), c.v_new.reshape(T, -1)], 1).to(device)
        _, k, v = qkv.split([c.Hq * D, Hkv * D, Hkv * D], dim=1)
        k, v = k.view(T, Hkv, D), v.view(T, Hkv, D)
        assert T == 1 or not k.is_contiguous()
    else:
        k, v = c.k_new.to(device), c.v_new.to(device)
    ops.set_kv_buffer(dk, dv, c.out_cache_loc.to(device), k, v, dtype)
    assert H.pools_equal(dk, kp) and H.pools_equal(dv, vp)


def test_per_head_device_scales_and_scalar_parameter(device):
    """k_scale as a 0-dim fp32 device tensor (layer.k_scale Parameter) and the per-head superset."""
    g = torch.Generator().manual_seed(5)
    T, Hkv, D = 64, 4, 128
    k = torch.randn((T, Hkv, D), generator=g).to(torch.bfloat16)
    v = torch.randn((T, Hkv, D), generator=g).to(torch.bfloat16)
    loc = (torch.randperm(200, generator=g)[:T] + 1).to(torch.int64)
    # scalar parameter
    s = torch.tensor(0.0625 * 1.7, dtype=torch.float32)
    kp, vp = torch.zeros((201, Hkv, D), dtype=torch.uint8), torch.zeros((201, Hkv, D), dtype=torch.uint8)
    tn.set_kv_buffer(kp, vp, loc, k.clone(), v.clone(), torch.float8_e4m3fn, s, s)
    dk, dv = torch.zeros_like(kp, device=device), torch.zeros_like(vp, device=device)
    ops.set_kv_buffer(dk, dv, loc.to(device), k.to(device), v.to(device), torch.float8_e4m3fn, s.to(device), s.to(device))
    assert H.pools_equal(dk, kp) and H.pools_equal(dv, vp)
    # per-head vector: checked against the plain-C restatement
    sv = torch.tensor([0.05, 0.11, 1.0, 3.0], dtype=torch.float32)
    kp2 = torch.zeros_like(kp)
    exact.kv_write(k.clone(), loc, kp2, sv)
    want = torch.zeros_like(kp)
    for h in range(Hkv):
        kk = k[:, h:h + 1].clone()
        kk.div_(float(sv[h]))
        want[loc, h:h + 1] = kk.to(torch.float8_e4m3fn).view(torch.uint8)
    assert torch.equal(kp2, want)
    dk.zero_(); dv.zero_()
    ops.set_kv_buffer(dk, dv, loc.to(device), k.to(device), v.to(device), torch.float8_e4m3fn, sv.to(device), sv.to(device))
    assert H.pools_equal(dk, want)


def test_rcp_scale_mode_matches_torch_cuda_div_by_python_float(device):
    """torch CUDA `div_` by a python scalar multiplies by the reciprocal (BinaryDivTrueKernel.cu); the kernel's
    SCALE_RCP mode reproduces that bit for bit, SCALE_DIV reproduces torch CPU / tensor division."""
    g = torch.Generator().manual_seed(6)
    k = (torch.randn((128, 2, 128), generator=g) * 2).to(torch.bfloat16).to(device)
    loc = torch.arange(1, 129, device=device)
    s = 0.1135
    ref = k.clone()
    ref.div_(s)  # torch CUDA semantics
    want = ref.to(torch.float8_e4m3fn).view(torch.uint8)
    kp = torch.zeros((129, 2, 128), dtype=torch.uint8, device=device)
    vp = torch.zeros_like(kp)
    ops.set_kv_buffer(kp, vp, loc, k.clone(), k.clone(), torch.float8_e4m3fn, s, s, scale_mode=_lib.SCALE_RCP)
    assert torch.equal(kp[1:], want)


def test_duplicate_loc_dummy_slot_and_empty(device):
    """Padded CUDA-graph rows all write slot 0 (memory_pool.py:238-243): any writer may win, nothing else moves."""
    T, Hkv, D = 16, 2, 64
    g = torch.Generator().manual_seed(8)
    k = torch.randn((T, Hkv, D), generator=g).to(torch.bfloat16).to(device)
    loc = torch.zeros(T, dtype=torch.int64, device=device)
    loc[:4] = torch.tensor([5, 6, 7, 8])
    kp = torch.zeros((10, Hkv, D), dtype=torch.bfloat16, device=device)
    vp = torch.zeros_like(kp)
    ops.set_kv_buffer(kp, vp, loc, k, k, torch.bfloat16)
    assert torch.equal(kp[5:9], k[:4]) and not kp[1:5].any() and not kp[9:].any()
    assert any(torch.equal(kp[0], k[t]) for t in range(4, T))
    ops.set_kv_buffer(kp, vp, loc[:0], k[:0], k[:0], torch.bfloat16)  # T == 0 is a no-op


def test_full_size_c2_and_c3_writes_round_trip(device):
    """BASELINE-size writes checked through a size-independent property: scatter then gather returns the input."""
    for T, Hkv, D in ((256, 8, 128), (32768, 8, 128)):
        g = torch.Generator(device=device).manual_seed(T)
        k = torch.randn((T, Hkv, D), generator=g, device=device, dtype=torch.bfloat16)
        v = torch.randn((T, Hkv, D), generator=g, device=device, dtype=torch.bfloat16)
        P = 2 * T + 1
        loc = (torch.randperm(P - 1, device=device)[:T] + 1).to(torch.int64)
        kp = torch.zeros((P, Hkv, D), dtype=torch.bfloat16, device=device)
        vp = torch.zeros_like(kp)
        ops.set_kv_buffer(kp, vp, loc, k, v, torch.bfloat16)
        back_k = ops.gather_kv(kp, loc.to(torch.int32))
        back_v = ops.gather_kv(vp, loc.to(torch.int32))
        assert torch.equal(back_k, k) and torch.equal(back_v, v)
        assert int((kp != 0).any(dim=(1, 2)).sum()) <= T  # nothing outside loc was touched
