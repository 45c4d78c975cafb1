"""ctypes caller of oracle/exact.c (built with gcc into oracle/_build/).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
from pathlib import Path
from typing import Optional

import torch

_DIR = Path(__file__).resolve().parent
SRC = _DIR / "exact.c"
LIB = _DIR / "_build" / "liboracle_exact.so"

_DT = {torch.bfloat16: 0, torch.float16: 1, torch.float8_e4m3fn: 2, torch.uint8: 2, torch.float32: 3}
_lib = None


def build(force: bool = False) -> Path:
    LIB.parent.mkdir(parents=True, exist_ok=True)
    if force or not LIB.exists() or LIB.stat().st_mtime < SRC.stat().st_mtime:
        gcc = os.environ.get("CC") or shutil.which("gcc") or "gcc"
        # -O2 without -ffast-math: the fp8/bf16 rounding code relies on IEEE semantics
        cmd = [gcc, "-O2", "-fPIC", "-shared", "-std=gnu11", "-o", str(LIB), str(SRC), "-lm"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"gcc failed for {SRC}:\n{res.stderr}")
    return LIB


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        L = C.CDLL(str(build()))
        L.orc_f32_to_e4m3.restype = C.c_uint8
        L.orc_f32_to_e4m3.argtypes = [C.c_float]
        L.orc_e4m3_to_f32.restype = C.c_float
        L.orc_e4m3_to_f32.argtypes = [C.c_uint8]
        L.orc_f32_to_bf16.restype = C.c_uint16
        L.orc_f32_to_bf16.argtypes = [C.c_float]
        L.orc_f32_to_f16.restype = C.c_uint16
        L.orc_f32_to_f16.argtypes = [C.c_float]
        L.orc_decode.restype = C.c_int
        L.orc_extend.restype = C.c_int
        _lib = L
    return _lib


def _p(t: Optional[torch.Tensor]):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _cpu(*ts):
    for t in ts:
        if t is not None and t.device.type != "cpu":
            raise ValueError("the exact oracle runs on CPU tensors")


def kv_write(x: torch.Tensor, loc: torch.Tensor, pool: torch.Tensor, scale=None, scale_mode: int = 0,
             writeback: bool = False) -> None:
    """x [T,H,D] bf16/f16 (last dim contiguous), loc int64 [T], pool [P,H,D] same dtype or uint8 (e4m3).
    scale: None | float | fp32 tensor with 1 or H entries."""
    _cpu(x, loc, pool)
    T, H, D = x.shape
    assert x.stride(2) == 1 and loc.dtype == torch.int64 and loc.is_contiguous()
    assert pool.stride(2) == 1 and pool.stride(1) == D
    sc = None
    if scale is not None:
        sc = (scale.detach().float().reshape(-1).contiguous() if isinstance(scale, torch.Tensor)
              else torch.tensor([float(scale)], dtype=torch.float32))
    lib().orc_kv_write(_p(x), C.c_int64(x.stride(0)), C.c_int64(x.stride(1)), _p(loc), C.c_int(T), C.c_int(H),
                       C.c_int(D), _p(pool), C.c_int64(pool.stride(0)), C.c_int(_DT[x.dtype]),
                       C.c_int(_DT[pool.dtype]), _p(sc), C.c_int(1 if sc is None else sc.numel()),
                       C.c_int(scale_mode), C.c_int(int(writeback)))


def gather_rows(pool: torch.Tensor, indices: torch.Tensor) -> torch.Tensor:
    _cpu(pool, indices)
    assert indices.dtype == torch.int32 and indices.is_contiguous() and pool[0].is_contiguous()
    out = torch.empty((indices.numel(),) + tuple(pool.shape[1:]), dtype=pool.dtype)
    rb = pool[0].numel() * pool.element_size()
    lib().orc_gather_rows(_p(pool), C.c_int64(pool.stride(0) * pool.element_size()), _p(indices),
                          C.c_int64(indices.numel()), C.c_int64(rb), _p(out))
    return out


def flatten_kv_indices(table: torch.Tensor, rpi: torch.Tensor, lens: torch.Tensor, kv_indptr: torch.Tensor,
                       start: Optional[torch.Tensor], total: int) -> torch.Tensor:
    _cpu(table, rpi, lens, kv_indptr, start)
    assert table.dtype == torch.int32 and kv_indptr.dtype == torch.int32 and rpi.dtype == lens.dtype
    out = torch.full((total,), -1, dtype=torch.int32)
    lib().orc_flatten_kv_indices(_p(table), C.c_int64(table.stride(0)), _p(rpi.contiguous()), _p(lens.contiguous()),
                                 _p(start), C.c_int(int(lens.dtype == torch.int64)), _p(kv_indptr), _p(out),
                                 C.c_int(lens.numel()))
    return out


def decode(q, k_pool, v_pool, table, rpi, seq_lens, sm_scale, logit_cap=0.0, k_scale=1.0, v_scale=1.0,
           return_lse=False):
    """q [bs,Hq,D] bf16/f16; pools [P,Hkv,D] bf16/f16/uint8(e4m3).  -> fp32 [bs,Hq,D] (, lse [bs,Hq])"""
    _cpu(q, k_pool, v_pool, table, rpi, seq_lens)
    bs, Hq, D = q.shape
    assert q.stride(2) == 1 and q.stride(1) == D and table.dtype == torch.int32 and rpi.dtype == seq_lens.dtype
    out = torch.empty((bs, Hq, D), dtype=torch.float32)
    lse = torch.empty((bs, Hq), dtype=torch.float32) if return_lse else None
    rc = lib().orc_decode(_p(q), C.c_int64(q.stride(0)), C.c_int(_DT[q.dtype]), _p(k_pool), _p(v_pool),
                          C.c_int64(k_pool.stride(0)), C.c_int(_DT[k_pool.dtype]), _p(table),
                          C.c_int64(table.stride(0)), _p(rpi.contiguous()), _p(seq_lens.contiguous()),
                          C.c_int(int(seq_lens.dtype == torch.int64)), C.c_int(bs), C.c_int(Hq),
                          C.c_int(k_pool.shape[1]), C.c_int(D), C.c_double(sm_scale * k_scale),
                          C.c_double(logit_cap), C.c_double(v_scale), _p(out), _p(lse))
    assert rc == 0
    return (out, lse) if return_lse else out


def extend(q, k_pool, v_pool, table, rpi, seq_lens, extend_seq_lens, extend_prefix_lens, extend_start_loc,
           sm_scale, logit_cap=0.0, k_scale=1.0, v_scale=1.0, causal=True, return_lse=False):
    _cpu(q, k_pool, v_pool, table, rpi, seq_lens, extend_seq_lens, extend_prefix_lens, extend_start_loc)
    T, Hq, D = q.shape
    assert q.stride(2) == 1 and q.stride(1) == D and table.dtype == torch.int32
    for t in (extend_seq_lens, extend_prefix_lens, extend_start_loc):
        assert t.dtype == torch.int32 and t.is_contiguous()
    out = torch.empty((T, Hq, D), dtype=torch.float32)
    lse = torch.empty((T, Hq), dtype=torch.float32) if return_lse else None
    rc = lib().orc_extend(_p(q), C.c_int64(q.stride(0)), C.c_int(_DT[q.dtype]), _p(k_pool), _p(v_pool),
                          C.c_int64(k_pool.stride(0)), C.c_int(_DT[k_pool.dtype]), _p(table),
                          C.c_int64(table.stride(0)), _p(rpi.contiguous()), _p(seq_lens.contiguous()),
                          C.c_int(int(seq_lens.dtype == torch.int64)), _p(extend_seq_lens), _p(extend_prefix_lens),
                          _p(extend_start_loc), C.c_int(seq_lens.numel()), C.c_int(Hq), C.c_int(k_pool.shape[1]),
                          C.c_int(D), C.c_double(sm_scale * k_scale), C.c_double(logit_cap), C.c_double(v_scale),
                          C.c_int(int(causal)), _p(out), _p(lse))
    assert rc == 0
    return (out, lse) if return_lse else out


def merge_state(a, lse_a, b, lse_b):
    _cpu(a, lse_a, b, lse_b)
    a, b = a.contiguous(), b.contiguous()
    T, H, D = a.shape
    out = torch.empty((T, H, D), dtype=torch.float32)
    lse = torch.empty((T, H), dtype=torch.float32)
    lib().orc_merge_state(_p(a), _p(lse_a.float().contiguous()), _p(b), _p(lse_b.float().contiguous()),
                          C.c_int64(T * H), C.c_int(D), C.c_int(_DT[a.dtype]), _p(out), _p(lse))
    return out, lse


def transfer_kv(src: torch.Tensor, dst: torch.Tensor, src_layout: int, dst_layout: int, src_dim: int, dst_dim: int,
                src_idx: torch.Tensor, dst_idx: torch.Tensor, item_bytes: int, num_layers: int, start_layer: int = 0):
    """Plain-C restatement of the reference's transfer kernel (transfer.cu:9-110) on CPU tensors; `src`/`dst` are the
    base buffers (layouts 0/1) or int64 tensors of per-layer data_ptr()s (layout 2)."""
    _cpu(src, dst, src_idx, dst_idx)
    L = lib()
    L.orc_transfer_kv.restype = None
    L.orc_transfer_kv.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                  C.c_int64, C.c_int64, C.c_int64, C.c_int64]
    si, di = src_idx.to(torch.int64).contiguous(), dst_idx.to(torch.int64).contiguous()
    L.orc_transfer_kv(src.data_ptr(), dst.data_ptr(), src_layout, dst_layout, src_dim, dst_dim, si.data_ptr(), di.data_ptr(),
                      si.numel(), item_bytes, start_layer, num_layers)
