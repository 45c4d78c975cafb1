"""CPU checks of the drop-in boundary: the C-ABI library builds, loads, and exports every symbol that
include/b200attn.h declares; argument validation answers with error codes (no device needed); the product
package never imports the oracle and has no CPU fallback."""
from __future__ import annotations

import ctypes as C
import re
import subprocess
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "b200attn.h"


def _declared_symbols() -> list:
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(b200attn_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_all_bound_and_exported(lib):
    from sglang_custom_b200 import _lib
    declared = _declared_symbols()
    assert len(declared) >= 12
    assert set(declared) == set(_lib.SIGNATURES), "ctypes SIGNATURES must list exactly what the header declares"
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in include/b200attn.h but not exported"


def test_exports_are_plain_c_and_nothing_else_leaks(lib):
    from sglang_custom_b200 import _lib
    out = subprocess.run(["nm", "-D", "--defined-only", str(_lib.LIB_PATH)], capture_output=True, text=True, check=True)
    names = [ln.split()[-1] for ln in out.stdout.splitlines() if " T " in ln]
    ours = [n for n in names if n.startswith("b200attn_")]
    assert set(ours) == set(_declared_symbols())
    assert not [n for n in names if n.startswith("_Z") and "b200attn" in n], "C++-mangled symbols leaked into the ABI"


def test_library_targets_sm100a_and_uses_blackwell_instructions(lib):
    """cuobjdump evidence: the cubin is sm_100a, the decode path stages with LDGSTS (cp.async), the extend path
    has tcgen05 (UTC*MMA / LDTM / STTM) and no legacy-only tensor path."""
    from sglang_custom_b200 import _lib
    res = subprocess.run(["cuobjdump", "-lelf", str(_lib.LIB_PATH)], capture_output=True, text=True)
    if res.returncode != 0:
        pytest.skip("cuobjdump not available")
    assert "sm_100a" in res.stdout
    assert "sm_90" not in res.stdout and "sm_80" not in res.stdout
    # the SASS itself (scripts/dump_sass.py writes the same counts to profiles/r02_sass_summary.txt)
    import sys
    sys.path.insert(0, str(ROOT / "scripts"))
    import dump_sass
    c = dump_sass.family_counts()
    tc = c["tc::extend_tc_kernel"]
    assert tc.get("UTCHMMA", 0) > 0, "extend: no tcgen05.mma (UTCHMMA) in the SASS"
    assert tc.get("LDTM", 0) > 0 and tc.get("STTM", 0) > 0, "extend: no tcgen05.ld / tcgen05.st (LDTM / STTM)"
    assert tc.get("UTCBAR", 0) > 0 and tc.get("SYNCS", 0) > 0, "extend: no tcgen05.commit / mbarrier"
    assert tc.get("LDGSTS", 0) > 0, "extend: the page-1 gather is LDGSTS"
    assert tc.get("HMMA", 0) == 0 and tc.get("HGMMA", 0) == 0, "extend: legacy tensor path in the tcgen05 kernel"
    dec = c["decode_kernel"]
    assert dec.get("LDGSTS", 0) > 0 and dec.get("SYNCS", 0) > 0, "decode: LDGSTS gather + mbarrier hand-off"
    assert dec.get("HMMA", 0) > 0 and dec.get("LDSM", 0) > 0, "decode: GQA-packed mma.sync from ldmatrix"
    assert dec.get("F2FP.E4M3", 0) > 0, "decode: FP8 dequantisation in registers"
    # paged pools (page_size % 16 == 0): K / V tiles arrive as bulk-tensor (TMA) boxes in both hot kernels
    assert dec.get("UTMALDG", 0) > 0, "decode: no cp.async.bulk.tensor (UTMALDG) in the paged-tile producers"
    assert tc.get("UTMALDG", 0) > 0, "extend: no cp.async.bulk.tensor (UTMALDG) in the paged-tile producers"
    for fam in c:
        assert c[fam].get("HGMMA", 0) == 0  # wgmma does not exist on sm_100a


def test_version_and_error_string(lib):
    from sglang_custom_b200 import _lib
    assert lib.b200attn_version() == 220
    assert isinstance(lib.b200attn_last_error(), bytes)
    assert lib.b200attn_decode_plan_bytes(256, 148) == 4 * (8 + 257 + 148)
    assert lib.b200attn_decode_workspace_bytes(148, 32, 128) == 2 * 148 * 32 * 129 * 4
    assert _lib.version() == 220


def test_argument_validation_returns_error_codes_without_a_device(lib):
    """Host-side checks run before any CUDA call (reference style: TORCH_CHECK -> RuntimeError)."""
    from sglang_custom_b200 import _lib
    buf = torch.zeros(4096, dtype=torch.uint8)
    p = buf.data_ptr()
    # D not a multiple of 8
    rc = lib.b200attn_kv_write(p, p, 8, 8, 8, 8, p, 1, 1, 12, p, p, 16, _lib.BF16, _lib.BF16, 0.0, 0.0, None, None, 1, 0, 0, None)
    assert rc == -1 and b"multiple of 8" in lib.b200attn_last_error()
    # fp32 input is not built
    rc = lib.b200attn_kv_write(p, p, 64, 64, 64, 64, p, 1, 1, 64, p, p, 64, _lib.F32, _lib.F32, 0.0, 0.0, None, None, 1, 0, 0, None)
    assert rc == -2
    # decode: head dim 96 unsupported, GQA mismatch invalid
    rc = lib.b200attn_decode(p, 96, p, 96, p, p, 96, p, 16, p, p, 1, 1, 1, 1, 96, 0, 0, 1.0, 0.0, 1.0, 1.0, p, 148, p, 1 << 30, None, None)
    assert rc == -2
    rc = lib.b200attn_decode(p, 128, p, 128, p, p, 128, p, 16, p, p, 1, 1, 3, 2, 128, 0, 0, 1.0, 0.0, 1.0, 1.0, p, 148, p, 1 << 30, None, None)
    assert rc == -1
    # workspace too small
    rc = lib.b200attn_decode(p, 512, p, 512, p, p, 256, p, 16, p, p, 1, 1, 4, 2, 128, 0, 0, 1.0, 0.0, 1.0, 1.0, p, 148, p, 16, None, None)
    assert rc == -4
    # plan buffer too small
    rc = lib.b200attn_decode_plan(p, 1, 100, 148, 4, p, 64, None)
    assert rc == -4
    with pytest.raises(_lib.B200AttnError):
        _lib.check(rc, "b200attn_decode_plan")
    # empty batches are no-ops that succeed
    assert lib.b200attn_kv_write(None, None, 0, 0, 0, 0, None, 0, 1, 64, None, None, 64, 0, 0, 0.0, 0.0, None, None, 1, 0, 0, None) == 0
    assert lib.b200attn_merge_state(None, None, None, None, None, None, 0, 1, 64, 0, None) == 0


def test_ops_refuse_cpu_tensors_no_fallback():
    from sglang_custom_b200 import ops
    from sglang_custom_b200._lib import B200AttnError
    k = torch.zeros((2, 1, 64), dtype=torch.bfloat16)
    pool = torch.zeros((8, 1, 64), dtype=torch.bfloat16)
    with pytest.raises(B200AttnError, match="CUDA tensors only"):
        ops.kv_write(k, k, torch.zeros(2, dtype=torch.int64), pool, pool)
    with pytest.raises(B200AttnError):
        ops.gather_kv(pool, torch.zeros(2, dtype=torch.int32))


def test_backend_refuses_cpu_device(lib):
    from sglang_custom_b200 import interfaces as itf
    from sglang_custom_b200._lib import B200AttnError
    from sglang_custom_b200.backend import B200AttnBackend
    cfg = itf.ModelConfig(context_len=64, num_attention_heads=4, num_key_value_heads=2, head_dim=64)
    mr = itf.ModelRunner(cfg, max_reqs=4, pool_size=64, device="cpu")
    with pytest.raises(B200AttnError, match="no CPU path"):
        B200AttnBackend(mr)


def test_product_never_imports_the_oracle():
    pkg = ROOT / "sglang_custom_b200"
    offenders = []
    for f in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")):
        txt = f.read_text()
        if re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M) or "oracle/" in txt and f.suffix != ".py":
            offenders.append(str(f))
    assert not offenders, offenders


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from sglang_custom_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", tmp_path / "nope.so")
    with pytest.raises(ImportError, match="no CPU or torch fallback"):
        _lib.load()
