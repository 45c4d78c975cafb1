"""GPU parity of the small kernels: 2-way LSE merge, page-table flatten, page gather."""
from __future__ import annotations

import pytest
import torch

from oracle import exact
from oracle import torch_native as tn
from sglang_custom_b200 import ops
from sglang_custom_b200._lib import B200AttnError

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("T,Hh,D", [(256, 8, 128), (512, 16, 64), (37, 32, 256), (1, 1, 32)])
def test_merge_state_vs_reference_statement(dtype, T, Hh, D, device):
    """Grid after sgl-kernel/tests/test_merge_state_v2.py:100-140 (pure-torch merge reference)."""
    g = torch.Generator().manual_seed(T + D)
    a = torch.randn((T, Hh, D), generator=g).to(dtype)
    b = torch.randn((T, Hh, D), generator=g).to(dtype)
    la = torch.randn((T, Hh), generator=g) * 5
    lb = torch.randn((T, Hh), generator=g) * 5
    la[0, 0] = float("inf")   # +inf counts as -inf (merge_attn_states.cu:47-50)
    if T > 1:
        lb[1, 0] = float("-inf")
    want_o, want_l = exact.merge_state(a, la, b, lb)
    o, l = ops.merge_state(a.to(device), la.to(device), b.to(device), lb.to(device))
    tol = 1e-5 if dtype == torch.float32 else (8e-3 if dtype == torch.bfloat16 else 1e-3)
    assert (o.float().cpu() - want_o).abs().max() <= tol * max(1.0, float(want_o.abs().max()))
    assert torch.allclose(l.cpu(), want_l, atol=1e-4, rtol=1e-5)
    o2, l2 = tn.merge_state(a, la, b, lb)
    assert (o.float().cpu() - o2.float()).abs().max() <= 2 * tol * max(1.0, float(want_o.abs().max()))


@pytest.mark.parametrize("batch", [1, 37, 1786])
def test_flatten_kv_indices_exact(batch, device):
    """test/srt/test_create_kvindices.py:18-71: 4096 x 4096 table, batch in {1, 37, 1786}, exact equality."""
    g = torch.Generator().manual_seed(batch)
    max_b, max_ctx = 4096, 4096
    table = torch.randint(0, 1 << 20, (max_b, max_ctx), generator=g, dtype=torch.int32)
    rpi = torch.randperm(max_b, generator=g)[:batch]
    lens = torch.randint(1, max_ctx, (batch,), generator=g)
    start = torch.zeros(batch, dtype=torch.int64)
    start[::3] = torch.minimum(torch.tensor(5), max_ctx - lens[::3])
    indptr = torch.zeros(batch + 1, dtype=torch.int32)
    indptr[1:] = torch.cumsum(lens, 0)
    total = int(lens.sum())
    want = torch.cat([table[rpi[i], start[i]:start[i] + lens[i]] for i in range(batch)])
    for idx_dtype in (torch.int64, torch.int32):
        out = torch.full((total,), -1, dtype=torch.int32, device=device)
        ops.flatten_kv_indices(table.to(device), rpi.to(idx_dtype).to(device), lens.to(idx_dtype).to(device),
                               indptr.to(device), start.to(idx_dtype).to(device), out)
        assert torch.equal(out.cpu(), want)
    assert torch.equal(exact.flatten_kv_indices(table, rpi, lens, indptr, start, total), want)


@pytest.mark.parametrize("dtype,Hkv,D", [(torch.bfloat16, 8, 128), (torch.uint8, 1, 128), (torch.float16, 2, 64)])
def test_gather_kv_bit_exact(dtype, Hkv, D, device):
    g = torch.Generator().manual_seed(D + Hkv)
    pool = torch.randint(0, 255, (5000, Hkv, D * (1 if dtype == torch.uint8 else 2)), generator=g, dtype=torch.uint8)
    pool = pool if dtype == torch.uint8 else pool.view(dtype)
    idx = torch.randint(0, 5000, (12345,), generator=g, dtype=torch.int32)
    got = ops.gather_kv(pool.to(device), idx.to(device)).cpu()
    want = pool[idx.long()]
    assert torch.equal(got.view(torch.uint8), want.view(torch.uint8))
    assert torch.equal(exact.gather_rows(pool, idx).view(torch.uint8), want.view(torch.uint8))


@pytest.mark.parametrize("dtype,head_num,head_dim,layers,n", [
    (torch.bfloat16, 8, 128, 4, 37),          # 2 KiB rows
    (torch.float8_e4m3fn, 1, 128, 3, 300),    # 128-byte FP8 rows (C4 / C5 at TP 8), stored as uint8
    (torch.bfloat16, 2, 13, 2, 9),            # 52-byte rows: not a multiple of 16 -> byte path
    (torch.float16, 4, 64, 2, 5000),          # many rows: narrower column slices per CTA
])
def test_move_kv_cache_moves_rows_in_every_layer_like_the_reference(dtype, head_num, head_dim, layers, n, device):
    """MHATokenToKVPool.move_kv_cache (memory_pool.py:460-468,1150-1184), used by EAGLE after a verify step with
    page_size > 1 and topk > 1 (eagle_utils.py:575-583): buf[tgt] = buf[src] in all K and V buffers, every source row
    read as it was before the call -- also rows that are somebody else's target (the reference loads a whole column
    block of all rows before it stores)."""
    from sglang_custom_b200 import interfaces as itf
    size = 8192
    pool = itf.MHATokenToKVPool(size, 16, dtype, head_num, head_dim, layers, str(device))
    g = torch.Generator(device=device).manual_seed(n)
    for buf in pool.k_buffer + pool.v_buffer:
        buf.copy_(torch.randint(0, 255, buf.shape, generator=g, device=device, dtype=torch.uint8).view(torch.uint8)
                  if buf.dtype == torch.uint8 else torch.randn(buf.shape, generator=g, device=device).to(buf.dtype))
    before = [b.clone() for b in pool.k_buffer + pool.v_buffer]
    perm = torch.randperm(size, generator=g, device=device) + 16
    tgt = perm[:n].clone()
    src = perm[n // 2:n // 2 + n].clone()   # overlaps tgt: half of the sources are overwritten by this very call
    pool.move_kv_cache(tgt, src)
    torch.cuda.synchronize(device)
    for got, old in zip(pool.k_buffer + pool.v_buffer, before):
        want = old.clone()
        want[tgt] = old[src]
        assert torch.equal(got.view(torch.uint8), want.view(torch.uint8))
    # int32 locations, and an empty call
    pool.move_kv_cache(tgt.to(torch.int32)[:3], src.to(torch.int32)[:3])
    pool.move_kv_cache(tgt[:0], src[:0])


def test_move_kv_rejects_more_rows_than_one_cta_can_stage(device):
    from sglang_custom_b200 import interfaces as itf
    pool = itf.MHATokenToKVPool(20000, 1, torch.bfloat16, 1, 64, 1, str(device))
    loc = torch.arange(13000, device=device)
    with pytest.raises(B200AttnError, match="UNSUPPORTED"):
        pool.move_kv_cache(loc, loc)
