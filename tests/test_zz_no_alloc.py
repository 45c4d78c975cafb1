"""b200attn_decode_static_ws: decode for embedders that cannot give the library a scratch buffer.

Some hosts of this library cannot allocate device memory at the point where they call it: a cgo / JNI / ctypes
binding that was handed q, out, the pools and the tables but does not link a CUDA allocator of its own, or a serving
process that has already turned every free byte of HBM into KV pool and must not cudaMalloc inside its step loop (the
call may fail, and while it runs it stalls every stream of the GPU).  For them the library carries its own 10 MiB array
for the split-KV partials and exposes `b200attn_decode_static_ws` (include/b200attn.h), which takes no workspace.

What is checked here, through the same ctypes binding every other GPU test uses:
  * two calls in a row, on two different batches, give what `b200attn_decode` gives with a caller-allocated workspace
    (the second call finds the first call's partials in the shared array and must not be disturbed by them);
  * the call neither takes nor returns device memory (free bytes as the driver reports them stay put);
  * a problem too large for the built-in array is refused with a WORKSPACE error instead of being run.

This file is named test_zz_* so that pytest collects it last: under `-x` a failure here must not hide other tests.
It does not depend on anything the other tests leave behind and passes on its own in a fresh process as well.
"""
from __future__ import annotations

import pytest
import torch

import helpers as H
from sglang_custom_b200 import ops, synth
from sglang_custom_b200._lib import B200AttnError

pytestmark = pytest.mark.gpu

N_CTAS = 148
HQ, HKV, D = 16, 4, 128


def _batch(seq_lens, seed, device, fp8=False):
    case = synth.make_case(len(seq_lens), HQ, HKV, D, seq_lens, seed=seed, fp8=fp8, kv_scale=0.5 if fp8 else 1.0)
    kp, vp = H.pool_after_write(case)
    return case, H.to_device(case, device), kp, vp, kp.to(device), vp.to(device)


@pytest.mark.parametrize("fp8", [False, True], ids=["bf16", "fp8"])
def test_static_workspace_matches_caller_workspace_twice(fp8, device):
    # Long and short requests together: the long ones are cut across CTAs, so partials really go through the array.
    batches = [_batch([300, 5000, 31, 2048], 101, device, fp8), _batch([7000, 1, 640, 4097], 202, device, fp8)]
    plan = torch.zeros(ops.decode_plan_bytes(4, N_CTAS) // 4, dtype=torch.int32, device=device)
    ws = torch.empty(ops.decode_workspace_bytes(N_CTAS, HQ, D) // 4, dtype=torch.float32, device=device)
    assert ops.decode_workspace_bytes(N_CTAS, HQ, D) <= ops.decode_static_ws_capacity()

    wants, gots, lses_want, lses_got = [], [], [], []
    for case, c, kp, vp, dk, dv in batches:  # the existing entry point first, workspace from torch
        kw = dict(k_scale=case.kv_scale, v_scale=case.kv_scale,
                  kv_dtype=torch.float8_e4m3fn if fp8 else None) if fp8 else {}
        ops.decode_plan(c.seq_lens, N_CTAS, plan, min_units_per_cta=1)
        out = torch.empty_like(c.q)
        lse = torch.empty((4, HQ), dtype=torch.float32, device=device)
        ops.decode(c.q, out, dk, dv, c.req_to_token, c.req_pool_indices, c.seq_lens, plan, N_CTAS, ws,
                   case.sm_scale, lse_out=lse, **kw)
        wants.append(out)
        lses_want.append(lse)
    torch.cuda.synchronize(device)

    for case, c, kp, vp, dk, dv in batches:  # then the new one, twice in a row on the one stream
        kw = dict(k_scale=case.kv_scale, v_scale=case.kv_scale,
                  kv_dtype=torch.float8_e4m3fn if fp8 else None) if fp8 else {}
        ops.decode_plan(c.seq_lens, N_CTAS, plan, min_units_per_cta=1)
        out = torch.full_like(c.q, float("nan"))
        lse = torch.full((4, HQ), float("nan"), dtype=torch.float32, device=device)
        ops.decode_static_ws(c.q, out, dk, dv, c.req_to_token, c.req_pool_indices, c.seq_lens, plan, N_CTAS,
                             case.sm_scale, lse_out=lse, **kw)
        gots.append(out)
        lses_got.append(lse)
    torch.cuda.synchronize(device)

    for i, (case, c, kp, vp, dk, dv) in enumerate(batches):
        H.assert_close(gots[i], wants[i], H.EXACT_MAX_ABS_TOL, H.EXACT_MEAN_ABS_TOL,
                       what=f"call {i + 1}: static workspace vs caller workspace", rel=True)
        assert torch.allclose(lses_got[i], lses_want[i], atol=2e-3, rtol=1e-4), f"call {i + 1}: lse"
        # and against the checker itself, so the comparison above cannot pass on two equally wrong results
        H.assert_close(gots[i], H.exact_decode(case, kp, vp), H.EXACT_MAX_ABS_TOL, H.EXACT_MEAN_ABS_TOL,
                       what=f"call {i + 1}: static workspace vs fp64 restatement", rel=True)


def test_static_workspace_call_takes_no_device_memory(device):
    case, c, kp, vp, dk, dv = _batch([300, 5000, 31, 2048], 303, device)
    plan = torch.zeros(ops.decode_plan_bytes(4, N_CTAS) // 4, dtype=torch.int32, device=device)
    ops.decode_plan(c.seq_lens, N_CTAS, plan, min_units_per_cta=1)
    out = torch.empty_like(c.q)
    args = (c.q, out, dk, dv, c.req_to_token, c.req_pool_indices, c.seq_lens, plan, N_CTAS, case.sm_scale)
    ops.decode_static_ws(*args)  # first call on this device: address lookup, kernels loaded
    torch.cuda.synchronize(device)
    free_before, _ = torch.cuda.mem_get_info(device)
    torch_before = torch.cuda.memory_allocated(device)
    for _ in range(3):
        ops.decode_static_ws(*args)
    torch.cuda.synchronize(device)
    free_after, _ = torch.cuda.mem_get_info(device)
    assert torch.cuda.memory_allocated(device) == torch_before
    assert free_after == free_before, f"device free memory moved by {free_before - free_after} bytes"
    H.assert_close(out, H.exact_decode(case, kp, vp), H.EXACT_MAX_ABS_TOL, H.EXACT_MEAN_ABS_TOL,
                   what="static workspace, repeated", rel=True)


def test_static_workspace_rejects_oversize(device):
    n_ctas = 1000  # 2 * 1000 * 16 * 129 * 4 bytes = 16.5 MB of partials: more than the 10 MiB array
    assert ops.decode_workspace_bytes(n_ctas, HQ, D) > ops.decode_static_ws_capacity()
    case, c, kp, vp, dk, dv = _batch([40, 900], 404, device)
    plan = torch.zeros(ops.decode_plan_bytes(2, n_ctas) // 4, dtype=torch.int32, device=device)
    ops.decode_plan(c.seq_lens, n_ctas, plan, min_units_per_cta=1)
    out = torch.full_like(c.q, 7.0)
    with pytest.raises(B200AttnError, match="WORKSPACE"):
        ops.decode_static_ws(c.q, out, dk, dv, c.req_to_token, c.req_pool_indices, c.seq_lens, plan, n_ctas,
                             case.sm_scale)
    torch.cuda.synchronize(device)
    assert bool((out == 7.0).all()), "a refused call must not have launched anything"
