"""GPU parity of sliding-window attention and of speculative-decoding verify batches (custom mask).

SURVEY 8f rank 4.  The CUDA path (b200attn_decode_plan_window + b200attn_decode, b200attn_extend_masked, and the
backend's window / TARGET_VERIFY routes) is compared with
  * tests/golden/masked_*.pt -- outputs of the reference's unmodified Triton kernels (extend_attention.py,
    decode_attention.py, triton_backend.update_sliding_window_buffer) run by Triton's CPU interpreter,
  * oracle/triton_masked.py, the restatement pinned against those kernels (tests/test_oracle_pin_masked.py),
at bf16 / FP8 / page 16 / multi-tile sizes, plus properties the reference's own test states (a mask equal to the causal
mask == no mask, test/srt/test_triton_attention_kernels.py:121-172) at larger sizes.
Tolerance: north_star's bf16 bound (max-abs 2e-2, mean-abs 2e-3); fp16 fixtures get a 4x tighter bound.
"""
from __future__ import annotations

import types

import pytest
import torch

import golden_util as gu
import helpers as H
from oracle import triton_masked as tm
from sglang_custom_b200 import _lib, ops, synth

pytestmark = pytest.mark.gpu

F16_MAX, F16_MEAN = 5e-3, 5e-4


def _as16(fx):
    """fp32 fixtures are replayed in bf16 (the kernels take 16-bit q); the oracle is then run on the rounded inputs."""
    if fx.dtype != torch.float32:
        return fx, False
    for k in ("q", "k_new", "v_new", "k_pool", "v_pool", "ref_k_pool_after", "ref_v_pool_after"):
        setattr(fx, k, getattr(fx, k).to(torch.bfloat16))
    fx.dtype = torch.bfloat16
    return fx, True


def _direct_extend(case, device, kp, vp, window=-1, mask=None, indptr=None, skip=True, causal=True, logit_cap=0.0,
                   impl=_lib.EXTEND_AUTO):
    c = H.to_device(case, device)
    out = torch.empty_like(c.q)
    s = case.kv_scale if case.fp8 else 1.0
    ops.extend(c.q, out, kp.to(device), vp.to(device), c.req_to_token, c.req_pool_indices, c.seq_lens,
               c.extend_seq_lens, c.extend_prefix_lens, c.extend_start_loc, case.max_extend_len, case.sm_scale,
               logit_cap, s, s, causal=causal, kv_dtype=torch.float8_e4m3fn if case.fp8 else None,
               impl=impl, sliding_window_size=window, custom_mask=None if mask is None else mask.to(device),
               mask_indptr=None if indptr is None else indptr.to(device), skip_prefix_custom_mask=skip)
    torch.cuda.synchronize(device)
    return out


def _direct_decode(case, device, kp, vp, window, n_ctas=148, logit_cap=0.0):
    c = H.to_device(case, device)
    out = torch.empty_like(c.q)
    plan = torch.zeros(ops.decode_plan_bytes(case.bs, n_ctas) // 4, dtype=torch.int32, device=device)
    ws = torch.empty(ops.decode_workspace_bytes(n_ctas, case.Hq, case.D) // 4, dtype=torch.float32, device=device)
    ops.decode_plan(c.seq_lens, n_ctas, plan, 1, sliding_window_size=window)
    s = case.kv_scale if case.fp8 else 1.0
    ops.decode(c.q, out, kp.to(device), vp.to(device), c.req_to_token, c.req_pool_indices, c.seq_lens, plan, n_ctas, ws,
               case.sm_scale, logit_cap, s, s, kv_dtype=torch.float8_e4m3fn if case.fp8 else None)
    torch.cuda.synchronize(device)
    return out


def _oracle_pools(case, kp, vp):
    return H.oracle_pools_for_attention(case, kp, vp)


# ------------------------------------------------------------------ golden fixtures (reference Triton kernels)
@pytest.mark.parametrize("name", gu.MASKED_EXTEND_FIXTURES)
def test_masked_extend_golden(name, device):
    fx, rounded = _as16(gu.load(name))
    fx.fp8 = False
    mask, indptr = getattr(fx, "custom_mask", None), getattr(fx, "mask_indptr", None)
    out = _direct_extend(fx, device, fx.ref_k_pool_after, fx.ref_v_pool_after, fx.window, mask, indptr,
                         fx.skip_prefix_custom_mask)
    if rounded:
        want = tm.extend(fx.q, fx.ref_k_pool_after, fx.ref_v_pool_after, fx.req_to_token, fx.req_pool_indices,
                         fx.seq_lens, fx.extend_prefix_lens, fx.extend_seq_lens, fx.sm_scale, True, fx.window, mask,
                         indptr, fx.skip_prefix_custom_mask)
        H.assert_close(out, want, what=name)
    else:
        H.assert_close(out, fx.ref_out, F16_MAX, F16_MEAN, what=name)


@pytest.mark.parametrize("name", gu.MASKED_DECODE_FIXTURES)
def test_window_decode_golden(name, device):
    fx, rounded = _as16(gu.load(name))
    fx.fp8 = False
    for n_ctas in (148, 3):
        out = _direct_decode(fx, device, fx.ref_k_pool_after, fx.ref_v_pool_after, fx.window, n_ctas)
        if rounded:
            want = tm.decode(fx.q, fx.ref_k_pool_after, fx.ref_v_pool_after, fx.req_to_token, fx.req_pool_indices,
                             fx.seq_lens, fx.sm_scale, fx.window)
            H.assert_close(out, want, what=name)
        else:
            H.assert_close(out, fx.ref_out, F16_MAX, F16_MEAN, what=f"{name} n_ctas={n_ctas}")


# ------------------------------------------------------------------ sliding window through the backend
WINDOW_EXTEND = [
    # bs, Hq, Hkv, D, seq_lens, extend_lens, window, kwargs
    (2, 32, 8, 128, [700, 300], [200, 300], 128, {}),                       # several key tiles, no-prefix request
    (3, 8, 2, 128, [1000, 129, 64], [100, 1, 64], 255, {}),                 # window cuts the prefix mid-tile
    (2, 8, 1, 128, [900, 450], [130, 64], 63, {"fp8": True, "kv_scale": 0.1135}),
    (2, 8, 1, 128, [640, 320], [128, 64], 100, {"fp8": True, "page_size": 16}),
    (3, 8, 2, 64, [300, 200, 100], [100, 20, 37], 31, {}),                  # D = 64; rows beyond window + tile
    (2, 4, 4, 128, [515, 260], [3, 260], 4096, {"dtype": torch.float16}),   # window wider than every prefix
    (2, 16, 2, 128, [400, 77], [150, 7], 16, {"idx_dtype": torch.int32}),   # rows whose prefix is fully masked
]


@pytest.mark.parametrize("bs,Hq,Hkv,D,lens,ext,W,kw", WINDOW_EXTEND)
def test_window_extend_backend_vs_oracle(bs, Hq, Hkv, D, lens, ext, W, kw, device):
    case = synth.make_case(bs, Hq, Hkv, D, lens, extend_lens=ext, seed=4000 + W, **kw)
    kp, vp = H.pool_after_write(case)
    r = H.Runner(case, device, sliding_window_size=W)
    out = r.forward()
    dk, dv = r.pools()
    assert H.pools_equal(dk, kp) and H.pools_equal(dv, vp), "KV write must be bit-exact"
    ok, ov = _oracle_pools(case, kp, vp)
    want = tm.extend(case.q, ok, ov, case.req_to_token, case.req_pool_indices, case.seq_lens, case.extend_prefix_lens,
                     case.extend_seq_lens, case.sm_scale, True, W)
    H.assert_close(out, want, what="window extend")


WINDOW_DECODE = [
    (8, 32, 8, 128, [4096, 100, 1, 513, 2048, 17, 800, 129], 511, {}),
    (5, 8, 1, 128, [3000, 33, 1024, 16, 700], 127, {"fp8": True, "kv_scale": 0.1135}),
    (4, 8, 1, 128, [640, 320, 48, 1000], 200, {"fp8": True, "page_size": 16}),
    (3, 8, 2, 64, [300, 200, 100], 15, {}),
    (3, 4, 4, 128, [50, 257, 128], 1000, {"dtype": torch.float16}),
    (6, 16, 2, 128, [900, 2, 64, 65, 1500, 31], 63, {"idx_dtype": torch.int32}),
]


@pytest.mark.parametrize("bs,Hq,Hkv,D,lens,W,kw", WINDOW_DECODE)
def test_window_decode_backend_vs_oracle(bs, Hq, Hkv, D, lens, W, kw, device):
    case = synth.make_case(bs, Hq, Hkv, D, lens, seed=5000 + W, **kw)
    kp, vp = H.pool_after_write(case)
    r = H.Runner(case, device, sliding_window_size=W)
    out = r.forward()
    ok, ov = _oracle_pools(case, kp, vp)
    want = tm.decode(case.q, ok, ov, case.req_to_token, case.req_pool_indices, case.seq_lens, case.sm_scale, W)
    H.assert_close(out, want, what="window decode")
    # a global layer of the same model (Gemma-2 style local/global interleave) keeps the full plan
    r.layer.sliding_window_size = -1
    full = r.forward(save_kv_cache=False)
    H.assert_close(full, H.oracle_decode(case, kp, vp), what="global layer of a window model")


def test_window_decode_split_invariance_and_wide_window(device):
    case = synth.make_case(6, 32, 8, 128, [5000, 3000, 100, 4097, 64, 2500], seed=77)
    kp, vp = H.pool_after_write(case)
    a = _direct_decode(case, device, kp, vp, 1023, n_ctas=148)
    b = _direct_decode(case, device, kp, vp, 1023, n_ctas=1)
    c = _direct_decode(case, device, kp, vp, 1023, n_ctas=296)
    H.assert_close(a, b, 4e-3, 4e-4, what="148 vs 1 CTA")
    H.assert_close(a, c, 4e-3, 4e-4, what="148 vs 296 CTAs")
    wide = _direct_decode(case, device, kp, vp, 1 << 20)
    none = _direct_decode(case, device, kp, vp, -1)
    assert torch.equal(wide, none), "a window wider than every request must not change a bit"


def test_window_decode_under_cuda_graph(device):
    """capture with one batch, replay with other lengths: both plans are refreshed in place."""
    W = 100
    case = synth.make_case(4, 8, 2, 128, [300, 40, 1000, 129], seed=91, idx_dtype=torch.int32, max_reqs=8)
    kp, vp = H.pool_after_write(case)
    r = H.Runner(case, device, sliding_window_size=W)
    be, c = r.backend, H.to_device(case, device)
    be.init_cuda_graph_state(8, 8)
    rpi, sl = c.req_pool_indices.clone(), c.seq_lens.clone()
    r.mr.token_to_kv_pool.k_buffer[0].copy_(kp.to(device))
    r.mr.token_to_kv_pool.v_buffer[0].copy_(vp.to(device))
    fb = r.batch()
    fb.req_pool_indices, fb.seq_lens = rpi, sl
    q = c.q.reshape(case.T, -1).contiguous()
    be.init_forward_metadata_capture_cuda_graph(4, 4, rpi, sl, None, fb.forward_mode, None)
    s = torch.cuda.Stream(device)
    s.wait_stream(torch.cuda.current_stream(device))
    with torch.cuda.stream(s):
        r.layer.forward(q, None, None, fb, save_kv_cache=False)
    torch.cuda.current_stream(device).wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        out_static = r.layer.forward(q, None, None, fb, save_kv_cache=False)
    ok, ov = _oracle_pools(case, kp, vp)
    for lens in ([300, 40, 1000, 129], [90, 40, 101, 7]):
        sl.copy_(torch.tensor(lens, dtype=torch.int32))
        be.init_forward_metadata_replay_cuda_graph(4, rpi, sl, sum(lens), None, fb.forward_mode, None, None)
        g.replay()
        torch.cuda.synchronize(device)
        want = tm.decode(case.q, ok, ov, case.req_to_token, case.req_pool_indices,
                         torch.tensor(lens, dtype=torch.int64), case.sm_scale, W)
        H.assert_close(out_static.view(case.T, case.Hq, case.D), want, what=f"graph replay {lens}")


# ------------------------------------------------------------------ speculative decoding: TARGET_VERIFY
def _tree_masks(case, seed, prefix_keep=None):
    g = torch.Generator().manual_seed(seed)
    blocks = []
    for i in range(case.bs):
        P, E = int(case.extend_prefix_lens[i]), int(case.extend_seq_lens[i])
        m = tm.tree_mask(P, E, g)
        if prefix_keep is not None:
            m[:, :P] = torch.rand((E, P), generator=g) < prefix_keep
        blocks.append(m)
    return tm.flatten_masks(blocks)


VERIFY = [
    # bs, Hq, Hkv, D, prefix lens, num_draft_tokens, kwargs
    (4, 32, 8, 128, [2000, 64, 513, 1], 8, {}),
    (3, 8, 1, 128, [1500, 333, 100], 16, {"fp8": True, "kv_scale": 0.1135}),
    (2, 8, 1, 128, [640, 304], 4, {"fp8": True, "page_size": 16}),
    (3, 8, 2, 64, [300, 200, 100], 64, {}),                        # a full 64-row query tile of draft tokens
    (2, 4, 4, 128, [50, 257], 5, {"dtype": torch.float16}),
    (5, 16, 2, 128, [90, 0, 64, 65, 700], 3, {"idx_dtype": torch.int32}),  # one request without any cached token
]


@pytest.mark.parametrize("via_decode", [True, False])
@pytest.mark.parametrize("bs,Hq,Hkv,D,prefix,nd,kw", VERIFY)
def test_target_verify_backend_vs_oracle(bs, Hq, Hkv, D, prefix, nd, kw, via_decode, device):
    lens = [p + nd for p in prefix]
    case = synth.make_case(bs, Hq, Hkv, D, lens, extend_lens=[nd] * bs, seed=6000 + nd, **kw)
    kp, vp = H.pool_after_write(case)
    mask, indptr = _tree_masks(case, 6100 + nd)
    r = H.Runner(case, device, num_draft_tokens=nd)
    # both routes: prefix part on the split-KV decode kernel + masked draft part + LSE merge, or one masked extend
    r.backend.verify_via_decode = via_decode
    out = r.forward(fb=r.verify_batch(mask))
    dk, dv = r.pools()
    assert H.pools_equal(dk, kp) and H.pools_equal(dv, vp), "KV write of the draft tokens must be bit-exact"
    ok, ov = _oracle_pools(case, kp, vp)
    want = tm.extend(case.q, ok, ov, case.req_to_token, case.req_pool_indices, case.seq_lens, case.extend_prefix_lens,
                     case.extend_seq_lens, case.sm_scale, True, -1, mask, indptr, True)
    H.assert_close(out, want, what="target verify")
    # the tree mask matters: without it the result is the plain causal extend
    if nd >= 4:
        plain = H.oracle_extend(case, kp, vp)
        assert float((want.float() - plain.float()).abs().max()) > 1e-2


def test_target_verify_under_cuda_graph(device):
    """TARGET_VERIFY is a CUDA-graph mode in the reference (forward_batch_info.py: is_cuda_graph; hooks
    triton_backend.py:456-474,596-610): lens / mask_indptr / mask live in fixed buffers refreshed before replay."""
    import types
    nd, bs = 4, 3
    case = synth.make_case(bs, 8, 2, 128, [300 + nd, 40 + nd, 129 + nd], extend_lens=[nd] * bs, seed=93,
                           idx_dtype=torch.int32, max_reqs=4)
    kp, vp = H.pool_after_write(case)
    r = H.Runner(case, device, num_draft_tokens=nd)
    be, c = r.backend, H.to_device(case, device)
    r.mr.token_to_kv_pool.k_buffer[0].copy_(kp.to(device))
    r.mr.token_to_kv_pool.v_buffer[0].copy_(vp.to(device))
    be.init_cuda_graph_state(4, 4 * nd)
    rpi = c.req_pool_indices.clone()
    sl = c.extend_prefix_lens.clone()  # int32: tokens already cached
    mask_a, indptr_a = _tree_masks(case, 1)
    fb = r.verify_batch(mask_a)
    fb.req_pool_indices, fb.seq_lens = rpi, sl
    q = c.q.reshape(case.T, -1).contiguous()
    mode = fb.forward_mode
    be.init_forward_metadata_capture_cuda_graph(bs, bs * nd, rpi, sl, None, mode, fb.spec_info)
    s = torch.cuda.Stream(device)
    s.wait_stream(torch.cuda.current_stream(device))
    with torch.cuda.stream(s):
        r.layer.forward(q, None, None, fb, save_kv_cache=False)
    torch.cuda.current_stream(device).wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        out_static = r.layer.forward(q, None, None, fb, save_kv_cache=False)
    ok, ov = _oracle_pools(case, kp, vp)
    for seed in (1, 2):
        mask, indptr = _tree_masks(case, seed)
        spec = types.SimpleNamespace(custom_mask=mask.to(device), draft_token_num=nd)
        be.init_forward_metadata_replay_cuda_graph(bs, rpi, sl, int(sl.sum()), None, mode, spec, None)
        g.replay()
        torch.cuda.synchronize(device)
        want = tm.extend(case.q, ok, ov, case.req_to_token, case.req_pool_indices, case.seq_lens,
                         case.extend_prefix_lens, case.extend_seq_lens, case.sm_scale, True, -1, mask, indptr, True)
        H.assert_close(out_static.view(case.T, case.Hq, case.D), want, what=f"verify graph replay, mask {seed}")


def test_draft_decode_without_the_multistep_backend_raises(device):
    """A decode batch that carries a spec_info is a draft step: its rows come from B200MultiStepDraftBackend (the
    reference reads spec_info.kv_indptr / kv_indices there, triton_backend.py:203-205)."""
    case = synth.make_case(2, 8, 2, 128, [40, 30], seed=95)
    r = H.Runner(case, device, num_draft_tokens=4)
    fb = r.batch()
    fb.spec_info = types.SimpleNamespace()
    with pytest.raises(ops.B200AttnError):
        r.backend.init_forward_metadata(fb)


# ------------------------------------------------------------------ properties at larger sizes
def test_causal_mask_equals_no_mask_bitwise(device):
    """The reference's own statement (test_triton_attention_kernels.py:121-172): a custom mask equal to the causal
    mask changes nothing.  Here: bit-identical to the unmasked register-MMA kernel; likewise a window wider than
    every prefix."""
    case = synth.make_case(4, 12, 4, 128, [1400, 333, 64, 700], extend_lens=[700, 33, 64, 1], seed=81)
    kp, vp = H.pool_after_write(case)
    blocks = []
    for i in range(case.bs):
        P, E = int(case.extend_prefix_lens[i]), int(case.extend_seq_lens[i])
        blocks.append(torch.cat([torch.ones(E, P, dtype=torch.bool), torch.tril(torch.ones(E, E)).bool()], 1))
    mask, indptr = tm.flatten_masks(blocks)
    c = H.to_device(case, device)
    plain = torch.empty_like(c.q)
    ops.extend(c.q, plain, kp.to(device), vp.to(device), c.req_to_token, c.req_pool_indices, c.seq_lens,
               c.extend_seq_lens, c.extend_prefix_lens, c.extend_start_loc, case.max_extend_len, case.sm_scale,
               impl=_lib.EXTEND_MMA)
    masked = _direct_extend(case, device, kp, vp, -1, mask, indptr)
    assert torch.equal(plain, masked)
    masked_all = _direct_extend(case, device, kp, vp, -1, mask, indptr, skip=False)
    assert torch.equal(plain, masked_all)
    wide = _direct_extend(case, device, kp, vp, 1 << 20, impl=_lib.EXTEND_MMA)
    assert torch.equal(plain, wide)
    # the tcgen05 kernel's window variant with a window wider than every prefix == its plain variant
    plain_tc = torch.empty_like(c.q)
    ops.extend(c.q, plain_tc, kp.to(device), vp.to(device), c.req_to_token, c.req_pool_indices, c.seq_lens,
               c.extend_seq_lens, c.extend_prefix_lens, c.extend_start_loc, case.max_extend_len, case.sm_scale,
               impl=_lib.EXTEND_TCGEN05)
    wide_tc = _direct_extend(case, device, kp, vp, 1 << 20, impl=_lib.EXTEND_TCGEN05)
    assert torch.equal(plain_tc, wide_tc)
    H.assert_close(plain, H.oracle_extend(case, kp, vp), what="vs torch_native")


def test_prefix_mask_noncausal_and_fully_masked_rows(device):
    """skip_prefix_custom_mask = False with a random prefix mask, a non-causal layer with a mask, logit cap, and
    queries with no visible key at all (the reference gives NaN there; this kernel gives 0, as documented)."""
    case = synth.make_case(3, 8, 2, 128, [200, 90, 150], extend_lens=[70, 10, 6], seed=83)
    kp, vp = H.pool_after_write(case)
    mask, indptr = _tree_masks(case, 9, prefix_keep=0.5)
    # request 1, query 3: nothing visible
    P1, E1 = int(case.extend_prefix_lens[1]), int(case.extend_seq_lens[1])
    blk = mask[int(indptr[1]):int(indptr[2])].view(E1, P1 + E1)
    blk[3, :] = False
    args = (case.q, kp, vp, case.req_to_token, case.req_pool_indices, case.seq_lens, case.extend_prefix_lens,
            case.extend_seq_lens, case.sm_scale)
    for causal, cap in ((True, 0.0), (False, 0.0), (True, 30.0)):
        out = _direct_extend(case, device, kp, vp, -1, mask, indptr, skip=False, causal=causal, logit_cap=cap)
        if causal:
            want = tm.extend(*args, True, -1, mask, indptr, False, cap)
        else:
            # non-causal + mask: every key the mask allows (the reference visits all new keys when IS_CAUSAL is off)
            want = tm.extend(*args, False, -1, mask, indptr, False, cap)
        H.assert_close(out, want, what=f"prefix mask causal={causal} cap={cap}")
        row = int(case.extend_start_loc[1]) + 3
        assert float(out[row].float().abs().max()) == 0.0
    # window + mask together (kernel level; the backend refuses the combination like the reference)
    out = _direct_extend(case, device, kp, vp, 40, mask, indptr, skip=True)
    want = tm.extend(*args, True, 40, mask, indptr, True)
    H.assert_close(out, want, what="window + mask")


def test_masked_extend_rejects_bad_args(device):
    case = synth.make_case(1, 8, 2, 128, [300], extend_lens=[200], seed=85)
    kp, vp = H.pool_after_write(case)
    c = H.to_device(case, device)
    out = torch.empty_like(c.q)
    with pytest.raises(_lib.B200AttnError):
        ops.extend(c.q, out, kp.to(device), vp.to(device), c.req_to_token, c.req_pool_indices, c.seq_lens,
                   c.extend_seq_lens, c.extend_prefix_lens, c.extend_start_loc, case.max_extend_len, case.sm_scale,
                   impl=_lib.EXTEND_TCGEN05, custom_mask=torch.ones(200 * 300, dtype=torch.bool, device=device),
                   mask_indptr=torch.tensor([0, 60000], dtype=torch.int64, device=device))
    with pytest.raises(_lib.B200AttnError):
        ops.extend(c.q, out, kp.to(device), vp.to(device), c.req_to_token, c.req_pool_indices, c.seq_lens,
                   c.extend_seq_lens, c.extend_prefix_lens, c.extend_start_loc, case.max_extend_len, case.sm_scale,
                   custom_mask=torch.ones(10, dtype=torch.bool, device=device), mask_indptr=None)
    r = H.Runner(case, device)  # a model without a window: a window layer is a configuration error
    r.layer.sliding_window_size = 64
    with pytest.raises(_lib.B200AttnError):
        r.forward()


# ------------------------------------------------------------------ sliding window on the tcgen05 kernel
TC_WINDOW = [
    # bs, Hq, Hkv, D, seq_lens, extend_lens, window, kwargs
    (2, 4, 2, 128, [3000, 1500], [1500, 700], 300, {}),          # window start mid-tile, rows with no visible prefix
    (2, 4, 2, 128, [2100, 900], [600, 900], 1000, {}),           # several fully visible prefix tiles
    (2, 3, 1, 128, [1300, 400], [300, 130], 127, {}),            # odd GQA group: two q tiles of one head per CTA
    (2, 4, 1, 128, [1500, 777], [520, 129], 200, {"fp8": True, "kv_scale": 0.1135}),
    (1, 4, 2, 128, [1029], [900], 0, {}),                        # window 0: one prefix key, no window mask
    (2, 8, 2, 128, [640, 384], [256, 128], 64, {"page_size": 16, "dtype": torch.float16}),
]


@pytest.mark.parametrize("bs,Hq,Hkv,D,lens,ext,W,kw", TC_WINDOW)
def test_window_extend_tcgen05_vs_oracle_and_mma(bs, Hq, Hkv, D, lens, ext, W, kw, device):
    case = synth.make_case(bs, Hq, Hkv, D, lens, extend_lens=ext, seed=7000 + W, **kw)
    kp, vp = H.pool_after_write(case)
    tc = _direct_extend(case, device, kp, vp, W, impl=_lib.EXTEND_TCGEN05)
    mma = _direct_extend(case, device, kp, vp, W, impl=_lib.EXTEND_MMA)
    ok, ov = _oracle_pools(case, kp, vp)
    want = tm.extend(case.q, ok, ov, case.req_to_token, case.req_pool_indices, case.seq_lens, case.extend_prefix_lens,
                     case.extend_seq_lens, case.sm_scale, True, W)
    H.assert_close(tc, want, what="tcgen05 window vs oracle")
    H.assert_close(mma, want, what="register-MMA window vs oracle")
    H.assert_close(tc, mma, 8e-3, 8e-4, what="tcgen05 vs register-MMA")


# ------------------------------------------------------------------ several query tokens per request on the decode kernel
@pytest.mark.parametrize("bs,Hq,Hkv,D,lens,nq,kw", [
    (5, 32, 8, 128, [3000, 0, 64, 513, 17], 4, {}),
    (3, 8, 1, 128, [2000, 333, 100], 2, {"fp8": True, "kv_scale": 0.1135}),
    (3, 8, 2, 64, [300, 200, 100], 3, {}),
    (2, 4, 4, 128, [50, 257], 16, {"dtype": torch.float16}),
    (4, 16, 2, 128, [900, 2, 64, 1500], 2, {"idx_dtype": torch.int32, "page_size": 16}),
])
def test_decode_multiq_vs_oracle(bs, Hq, Hkv, D, lens, nq, kw, device):
    """nq query tokens per request over the same keys == nq independent decode calls (oracle: torch_native decode),
    with the LSE of every (request, token, head); split-invariant."""
    case = synth.make_case(bs, Hq, Hkv, D, [max(x, 1) for x in lens], seed=8000 + nq, **kw)
    kp, vp = case.k_pool, case.v_pool  # no KV write: the keys are the pool as it is
    g = torch.Generator().manual_seed(1)
    q = torch.randn((bs, nq, Hq, D), generator=g).to(case.dtype)
    seq = torch.tensor(lens, dtype=case.seq_lens.dtype)
    ok, ov = _oracle_pools(case, kp, vp)
    want = torch.stack([H.tn.decode(q[:, t].contiguous(), ok, ov, case.req_to_token, case.req_pool_indices, seq,
                                    case.sm_scale) for t in range(nq)], 1)
    for i, L in enumerate(lens):
        if L == 0:
            want[i] = 0  # empty request: defined as zeros, LSE -inf
    c = H.to_device(case, device)
    s = case.kv_scale if case.fp8 else 1.0
    outs = []
    for n_ctas in (148, 2):
        plan = torch.zeros(ops.decode_plan_bytes(bs, n_ctas) // 4, dtype=torch.int32, device=device)
        ws = torch.empty(ops.decode_workspace_bytes(n_ctas, Hq * nq, D) // 4, dtype=torch.float32, device=device)
        ops.decode_plan(seq.to(device), n_ctas, plan, 1)
        qd = q.to(device)
        out = torch.full_like(qd, float("nan"))
        lse = torch.full((bs, nq, Hq), float("nan"), dtype=torch.float32, device=device)
        ops.decode_multiq(qd, out, kp.to(device), vp.to(device), c.req_to_token, c.req_pool_indices, seq.to(device), plan,
                          n_ctas, ws, case.sm_scale, 0.0, s, s, kv_dtype=torch.float8_e4m3fn if case.fp8 else None,
                          lse_out=lse)
        torch.cuda.synchronize(device)
        H.assert_close(out, want, what=f"decode_multiq n_ctas={n_ctas}")
        for i, L in enumerate(lens):
            assert bool(torch.isinf(lse[i]).all()) == (L == 0)
        outs.append((out, lse))
    H.assert_close(outs[0][0], outs[1][0], 4e-3, 4e-4, what="split invariance")
    fin = torch.isfinite(outs[0][1])
    assert float((outs[0][1][fin] - outs[1][1][fin]).abs().max()) < 1e-3


def test_target_verify_full_size_routes_agree(device):
    """Verify batch at the size the timing script reports (bs 256, ctx 4096, 8 draft tokens, 32/8 heads, P = 1M rows):
    the decode_multiq + masked draft part + LSE merge route against the single masked extend kernel (both checked
    against the oracle at small sizes above), and against the plain causal extend when the mask is the causal one."""
    from sglang_custom_b200 import harness
    bs, L, Hq, Hkv, D, nd = 256, 4096, 32, 8, 128, 8
    c = synth.make_case_device(bs, Hq, Hkv, D, [L + nd] * bs, device, extend_lens=[nd] * bs, seed=1238)
    g = torch.Generator().manual_seed(3)
    blk = tm.tree_mask(L, nd, g)
    mask = blk.reshape(-1).repeat(bs)
    r = harness.Runner(c, device, num_draft_tokens=nd)
    outs = {}
    for via in (True, False):
        r.backend.verify_via_decode = via
        outs[via] = r.forward(fb=r.verify_batch(mask), save_kv_cache=False).clone()
    H.assert_close(outs[True], outs[False], 8e-3, 6e-4, what="verify routes at full size")
    causal = torch.cat([torch.ones(nd, L, dtype=torch.bool), torch.tril(torch.ones(nd, nd)).bool()], 1).reshape(-1).repeat(bs)
    r.backend.verify_via_decode = True
    got = r.forward(fb=r.verify_batch(causal), save_kv_cache=False)
    plain = torch.empty_like(c.q)
    ops.extend(c.q, plain, c.k_pool, c.v_pool, c.req_to_token, c.req_pool_indices, c.seq_lens, c.extend_seq_lens,
               c.extend_prefix_lens, c.extend_start_loc, nd, c.sm_scale, impl=_lib.EXTEND_MMA)
    torch.cuda.synchronize(device)
    H.assert_close(got, plain, 8e-3, 6e-4, what="causal tree mask == plain causal extend")
