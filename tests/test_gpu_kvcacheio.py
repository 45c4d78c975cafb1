"""GPU parity of the HiCache KV transfer kernel (b200attn_transfer_kv) through sglang_custom_b200.kvcacheio, the
drop-in for the reference's sgl_kernel.kvcacheio.  Mirrors sgl-kernel/tests/test_kvcacheio.py:16-261 (pools as
[layers, items, item_size] tensors, pinned host source, random disjoint page sets, `dst[dst_idx] = src[src_idx]` as
the statement of truth) and adds the page-first layouts, device -> host backup and the plain-C oracle."""
from __future__ import annotations

import pytest
import torch

from oracle import exact
from sglang_custom_b200 import kvcacheio as io
from sglang_custom_b200._lib import B200AttnError

pytestmark = pytest.mark.gpu
L = 4


def _indices(total_items, n_items, page_size, seed):
    g = torch.Generator().manual_seed(seed)
    pages = torch.randperm(total_items // page_size, generator=g)
    n_pages = n_items // page_size
    cat = lambda ps: torch.cat([torch.arange(p * page_size, (p + 1) * page_size) for p in ps])
    return cat(pages[:n_pages]), cat(pages[n_pages:2 * n_pages])


def _tbl(pool, device):
    return torch.tensor([pool[l].data_ptr() for l in range(pool.shape[0])], dtype=torch.uint64, device=device)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("n_items", [1, 128, 1024])
@pytest.mark.parametrize("page_size", [1, 16, 64])
@pytest.mark.parametrize("is_mla", [False, True])
@pytest.mark.parametrize("all_layers", [False, True])
def test_transfer_kv_host_to_device(dtype, n_items, page_size, is_mla, all_layers, device):
    total, item = 10240, 256
    if n_items // page_size == 0:
        pytest.skip("fewer items than one page")
    si, di = _indices(total, n_items, page_size, 42)
    g = torch.Generator().manual_seed(1)
    mk = lambda: torch.randn((L, total, item), generator=g).to(dtype).pin_memory()
    src_k, src_v = mk(), (None if is_mla else mk())
    dst_k = torch.zeros((L, total, item), dtype=dtype, device=device)
    dst_v = None if is_mla else torch.zeros_like(dst_k)
    dst_k_direct = torch.zeros_like(dst_k)
    dst_v_direct = None if is_mla else torch.zeros_like(dst_k)
    sid, did = si.to(device), di.to(device)
    layers = range(L) if all_layers else [0]
    want_k = torch.zeros((L, total, item), dtype=dtype)
    want_v = torch.zeros_like(want_k)
    for l in layers:
        want_k[l][di] = src_k[l][si]
        if not is_mla:
            want_v[l][di] = src_v[l][si]
    if all_layers:
        if is_mla:
            io.transfer_kv_all_layer_mla(_tbl(src_k, device), _tbl(dst_k, device), sid, did, io_backend="kernel",
                                         item_size=item * dst_k.element_size(), num_layers=L)
        else:
            io.transfer_kv_all_layer(_tbl(src_k, device), _tbl(dst_k, device), _tbl(src_v, device), _tbl(dst_v, device),
                                     sid, did, io_backend="kernel", item_size=item * dst_k.element_size(), num_layers=L)
    else:
        if is_mla:
            io.transfer_kv_per_layer_mla(src_k[0], dst_k[0], sid, did, io_backend="kernel", page_size=page_size,
                                         item_size=item)
            io.transfer_kv_direct([src_k[0]], [dst_k_direct[0]], si, did, page_size=page_size)
        else:
            io.transfer_kv_per_layer(src_k[0], dst_k[0], src_v[0], dst_v[0], sid, did, io_backend="kernel",
                                     page_size=page_size, item_size=item)
            io.transfer_kv_per_layer(src_k[0], dst_k_direct[0], src_v[0], dst_v_direct[0], si, did, io_backend="direct",
                                     page_size=page_size, item_size=item)
    torch.cuda.synchronize()
    assert torch.equal(dst_k.cpu(), want_k)
    if not is_mla:
        assert torch.equal(dst_v.cpu(), want_v)
    if not all_layers:
        assert torch.equal(dst_k_direct.cpu(), want_k)
        if not is_mla:
            assert torch.equal(dst_v_direct.cpu(), want_v)


@pytest.mark.parametrize("is_mla", [False, True])
@pytest.mark.parametrize("item_bytes", [256, 1024, 8 * 9])  # the last one forces the 8-byte path (not a multiple of 16)
def test_page_first_layouts_round_trip(is_mla, item_bytes, device):
    """Backup device (layer-first, pointer table) -> host page-first, then load host page-first -> device layer-first per
    layer (memory_pool_host.py's page_first layout): the round trip restores the pages, and the host image equals
    the plain-C oracle's."""
    total, n_items, page_size = 2048, 256, 16
    item = item_bytes // 2
    si, di = _indices(total, n_items, page_size, 7)
    g = torch.Generator().manual_seed(3)
    dev_k = torch.randn((L, total, item), generator=g).to(torch.bfloat16).to(device)
    dev_v = None if is_mla else torch.randn((L, total, item), generator=g).to(torch.bfloat16).to(device)
    host_k = torch.zeros((total, L, item), dtype=torch.bfloat16).pin_memory()
    host_v = None if is_mla else torch.zeros((total, L, item), dtype=torch.bfloat16).pin_memory()
    sid, did = si.to(device), di.to(device)
    if is_mla:
        io.transfer_kv_all_layer_mla_lf_pf(_tbl(dev_k, device), host_k, sid, did, item_size=item_bytes,
                                           dst_layout_dim=L * item_bytes, num_layers=L)
    else:
        io.transfer_kv_all_layer_lf_pf(_tbl(dev_k, device), host_k, _tbl(dev_v, device), host_v, sid, did,
                                       item_size=item_bytes, dst_layout_dim=L * item_bytes, num_layers=L)
    torch.cuda.synchronize()
    # plain-C oracle on CPU copies
    ck = dev_k.cpu()
    want_host = torch.zeros((total, L, item), dtype=torch.bfloat16)
    exact.transfer_kv(torch.tensor([ck[l].data_ptr() for l in range(L)], dtype=torch.int64), want_host, 2, 1, 0,
                      L * item_bytes, si, di, item_bytes, L)
    assert torch.equal(host_k, want_host)
    # load back, one layer per call, into fresh device pools at the ORIGINAL slots
    back_k = torch.zeros_like(dev_k)
    back_v = None if is_mla else torch.zeros_like(dev_k)
    for l in range(L):
        hk = host_k.view(-1)[l * item:]
        if is_mla:
            io.transfer_kv_per_layer_mla_pf_lf(hk, back_k[l], did, sid, item_size=item_bytes, src_layout_dim=L * item_bytes)
        else:
            io.transfer_kv_per_layer_pf_lf(hk, back_k[l], host_v.view(-1)[l * item:], back_v[l], did, sid,
                                           item_size=item_bytes, src_layout_dim=L * item_bytes)
    torch.cuda.synchronize()
    assert torch.equal(back_k[:, si].cpu(), dev_k[:, si].cpu())
    mask = torch.ones(total, dtype=torch.bool); mask[si] = False
    assert not back_k[:, mask].any()
    if not is_mla:
        assert torch.equal(back_v[:, si].cpu(), dev_v[:, si].cpu())


def test_device_to_device_full_pool_and_block_quota(device):
    """Device -> device move of every page of a pool-sized buffer (HBM-bound), with and without a CTA quota."""
    total, item = 65536, 1024  # 128 MiB per buffer
    src_k = torch.randint(0, 255, (total, item * 2), dtype=torch.uint8, device=device).view(torch.bfloat16)
    src_v = torch.randint(0, 255, (total, item * 2), dtype=torch.uint8, device=device).view(torch.bfloat16)
    perm = torch.randperm(total, device=device)
    for quota in (2, 0):
        dst_k, dst_v = torch.zeros_like(src_k), torch.zeros_like(src_v)
        io._launch(src_k, dst_k, src_v, dst_v, io.LF, io.LF, 0, 0, perm, torch.arange(total, device=device),
                   item * 2, 1, quota)
        torch.cuda.synchronize()
        assert torch.equal(dst_k.view(torch.int16), src_k[perm].view(torch.int16))
        assert torch.equal(dst_v.view(torch.int16), src_v[perm].view(torch.int16))


def test_argument_errors(device):
    a = torch.zeros((4, 64), dtype=torch.bfloat16, device=device)
    idx = torch.zeros(2, dtype=torch.int64, device=device)
    with pytest.raises(B200AttnError, match="type long"):
        io.transfer_kv_per_layer_mla(a, a.clone(), idx.int(), idx, "kernel", 1, 64)
    with pytest.raises(B200AttnError, match="same length"):
        io.transfer_kv_per_layer_mla(a, a.clone(), idx, idx[:1], "kernel", 1, 64)
    with pytest.raises(B200AttnError, match="divisible by 8"):
        io.transfer_kv_per_layer_mla(a, a.clone(), idx, idx, "kernel", 1, 2)
    with pytest.raises(ValueError):
        io.transfer_kv_per_layer_mla(a, a.clone(), idx, idx, "nope", 1, 64)
    with pytest.raises(B200AttnError, match="CUDA tensor"):
        io.transfer_kv_per_layer_mla(a, a.clone(), idx.cpu(), idx, "kernel", 1, 64)
