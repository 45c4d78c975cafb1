"""Drift detector between the reference's REAL classes and what this repo assumes about them.

The CUDA backend can only ever be exercised on the GPU box, where /root/reference does not exist, and the reference can
only be loaded in the build container, which has no GPU -- so `B200AttnBackend` is always driven through the mirrors in
sglang_custom_b200/interfaces.py ("a drift between that mirror and the reference is invisible
to every test").  This file closes that hole from the other side, in the build container:

  * every attribute the backend reads from `forward_batch`, `layer`, `model_runner` and the KV pool is looked up in the
    UNMODIFIED reference sources (AST of forward_batch_info.py, radix_attention.py, triton_backend.py,
    memory_pool.py) -- a renamed or removed field fails here;
  * the mirrors (ForwardMode members / predicates, ForwardBatch fields, RadixAttention attributes, pool and
    req_to_token_pool methods and signatures) are compared member by member with the real classes;
  * the real `MHATokenToKVPool` (loaded by oracle/ref_loader.py, CPU tensors) and the mirror are built side by side:
    same buffer shapes / dtypes / page padding, same `set_kv_buffer` result for bf16 and FP8 with a scale.

Skipped where the reference is not mounted (the GPU box)."""
from __future__ import annotations

import ast
import inspect
import os
import re
from pathlib import Path

import pytest
import torch

from oracle import ref_loader
from sglang_custom_b200 import interfaces as itf

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="/root/reference is only mounted in the build container")

REF = Path(os.environ.get("B200ATTN_REFERENCE", "/root/reference")) / "python" / "sglang" / "srt"
PKG = Path(itf.__file__).resolve().parent


def _tree(path: Path) -> ast.Module:
    return ast.parse(path.read_text())


def _class(tree: ast.Module, name: str) -> ast.ClassDef:
    for node in ast.walk(tree):
        if isinstance(node, ast.ClassDef) and node.name == name:
            return node
    raise AssertionError(f"class {name} not found")


def _self_attrs(cls: ast.ClassDef) -> set:
    """names assigned as self.X anywhere in the class + annotated class-level fields + method names"""
    out = set()
    for node in ast.walk(cls):
        if isinstance(node, ast.Attribute) and isinstance(node.value, ast.Name) and node.value.id == "self" and \
                isinstance(node.ctx, ast.Store):
            out.add(node.attr)
    for node in cls.body:
        if isinstance(node, ast.AnnAssign) and isinstance(node.target, ast.Name):
            out.add(node.target.id)
        if isinstance(node, (ast.FunctionDef, ast.AsyncFunctionDef)):
            out.add(node.name)
        if isinstance(node, ast.Assign):
            out.update(t.id for t in node.targets if isinstance(t, ast.Name))
    return out


def _reads(src: str, var: str) -> set:
    """attribute names read as `<var>.name` in a source text (first level only)"""
    return set(re.findall(rf"\b{var}\.([A-Za-z_][A-Za-z0-9_]*)", src))


BACKEND_SRC = (PKG / "backend.py").read_text() + (PKG / "draft_backend.py").read_text()


def test_forward_batch_fields_the_backend_reads_exist_in_the_reference():
    real = _self_attrs(_class(_tree(REF / "model_executor" / "forward_batch_info.py"), "ForwardBatch"))
    used = _reads(BACKEND_SRC, "forward_batch") | _reads(BACKEND_SRC, "fb")
    used -= {"b200_kv_written"}  # this repo's own marker (rotary.forward_fused_kv), set with setattr on the batch object
    missing = sorted(a for a in used if a not in real)
    assert not missing, f"backend reads ForwardBatch attributes the reference does not have: {missing}"
    mirror = {f.name for f in itf.ForwardBatch.__dataclass_fields__.values()}
    extra = sorted(a for a in mirror if a not in real and a != "b200_kv_written")
    assert not extra, f"interfaces.ForwardBatch has fields the reference does not: {extra}"


def test_forward_mode_members_and_predicates_match_the_reference():
    cls = _class(_tree(REF / "model_executor" / "forward_batch_info.py"), "ForwardMode")
    real_members = [n.targets[0].id for n in cls.body if isinstance(n, ast.Assign) and isinstance(n.targets[0], ast.Name)]
    real_methods = {n.name for n in cls.body if isinstance(n, ast.FunctionDef)}
    mirror_members = [m.name for m in itf.ForwardMode]
    # members are auto(): order defines the values
    assert mirror_members == real_members[:len(mirror_members)] or set(mirror_members) <= set(real_members), \
        f"ForwardMode members differ: {mirror_members} vs {real_members}"
    for name in mirror_members:
        assert itf.ForwardMode[name].value == real_members.index(name) + 1, f"ForwardMode.{name} value drifted"
    mirror_methods = {n for n, f in inspect.getmembers(itf.ForwardMode, inspect.isfunction) if n.startswith("is_")}
    assert mirror_methods <= real_methods, f"predicates not in the reference: {sorted(mirror_methods - real_methods)}"
    # the predicates the backend calls
    called = set(re.findall(r"forward_mode\.(is_[a-z_]+)\(", BACKEND_SRC))
    assert called <= real_methods, f"backend calls predicates the reference lacks: {sorted(called - real_methods)}"


def test_forward_mode_predicates_have_the_reference_truth_table():
    """Evaluate the real predicate bodies (pure comparisons on self) against the mirror for every member."""
    src = (REF / "model_executor" / "forward_batch_info.py").read_text()
    cls = _class(ast.parse(src), "ForwardMode")
    import enum
    ns = {"IntEnum": enum.IntEnum, "auto": enum.auto}
    exec(compile(ast.Module(body=[cls], type_ignores=[]), "<ref ForwardMode>", "exec"), ns)  # noqa: S102 (reference enum only)
    Real = ns["ForwardMode"]
    for m in itf.ForwardMode:
        r = Real[m.name]
        for pred in ("is_extend", "is_decode", "is_mixed", "is_idle", "is_target_verify", "is_draft_extend",
                     "is_decode_or_idle", "is_cuda_graph"):
            if hasattr(m, pred):
                assert getattr(m, pred)() == getattr(r, pred)(), f"ForwardMode.{m.name}.{pred}() differs from the reference"


def test_layer_attributes_the_backend_reads_exist_on_radix_attention():
    real = _self_attrs(_class(_tree(REF / "layers" / "radix_attention.py"), "RadixAttention"))
    used = _reads(BACKEND_SRC, "layer")
    # read with getattr(..., default): optional by construction
    optional = set(re.findall(r'getattr\(layer, "([a-z_]+)"', BACKEND_SRC))
    missing = sorted(a for a in used if a not in real and a not in optional)
    assert not missing, f"backend reads RadixAttention attributes the reference does not set: {missing}"
    for a in sorted(optional):
        assert a in real, f"optional layer attribute {a!r} does not exist in the reference at all (typo?)"
    mirror = set(itf.RadixAttention.__dataclass_fields__) if hasattr(itf.RadixAttention, "__dataclass_fields__") else set()
    computed = {"tp_q_head_num", "tp_k_head_num", "tp_v_head_num", "qk_head_dim", "v_head_dim", "scaling", "layer_id",
                "logit_cap", "sliding_window_size", "k_scale", "v_scale", "k_scale_float", "v_scale_float", "head_dim",
                "is_cross_attention", "attn_type"}
    for a in sorted((mirror | computed) & used):
        assert a in real, f"interfaces.RadixAttention.{a} has no counterpart in the reference"


def test_model_runner_attributes_are_the_ones_the_triton_backend_reads():
    triton = (REF / "layers" / "attention" / "triton_backend.py").read_text()
    real_reads = set(re.findall(r"model_runner\.([A-Za-z_][A-Za-z0-9_.]*)", triton))
    ours = set(re.findall(r"model_runner\.([A-Za-z_][A-Za-z0-9_.]*)", BACKEND_SRC))
    # shape / size of a tensor attribute are tensor API, not SGLang API
    norm = lambda s: re.sub(r"\.(shape|size|device|dtype)$", "", s)  # noqa: E731
    real_n = {norm(x) for x in real_reads} | {"model_config", "req_to_token_pool", "token_to_kv_pool", "server_args",
                                              "page_size", "tp_size"}
    missing = sorted(x for x in {norm(o) for o in ours} if x not in real_n and x.split(".")[0] not in real_n)
    assert not missing, f"B200AttnBackend reads model_runner attributes TritonAttnBackend never touches: {missing}"


def test_pool_mirror_matches_the_real_pool_member_by_member():
    ns = ref_loader.load()
    Real = ns.MHATokenToKVPool
    for meth in ("get_key_buffer", "get_value_buffer", "get_kv_buffer", "set_kv_buffer"):
        rs, ms = inspect.signature(getattr(Real, meth)), inspect.signature(getattr(itf.MHATokenToKVPool, meth))
        rp = [p for p in rs.parameters if p != "self"]
        mp = [p for p in ms.parameters if p != "self"]
        assert mp == rp[:len(mp)], f"{meth}: mirror parameters {mp} vs reference {rp}"
    kw = dict(size=48, page_size=16, dtype=torch.bfloat16, head_num=2, head_dim=8, layer_num=2, device="cpu")
    try:
        real = Real(enable_memory_saver=False, **kw)
    except TypeError:
        real = Real(**kw)
    mirror = itf.MHATokenToKVPool(**kw)
    assert len(real.k_buffer) == len(mirror.k_buffer) == 2
    assert real.k_buffer[0].shape == mirror.k_buffer[0].shape and real.k_buffer[0].dtype == mirror.k_buffer[0].dtype, \
        "pool buffer layout drifted (rows = size + page_size, [rows, heads, dim])"
    assert real.get_key_buffer(1).shape == mirror.get_key_buffer(1).shape
    for name in ("page_size", "dtype", "store_dtype", "head_num", "head_dim", "layer_num", "size"):
        if hasattr(mirror, name):
            assert getattr(real, name) == getattr(mirror, name), f"pool.{name} differs"


@pytest.mark.parametrize("dtype,scale", [(torch.bfloat16, None), (torch.float8_e4m3fn, 0.25)])
def test_real_pool_set_kv_buffer_equals_the_oracle_write(dtype, scale):
    """The KV-write contract the CUDA kernel is tested against (oracle/torch_native.set_kv_buffer, pinned elsewhere
    against golden outputs) restated on the REAL pool object: same bytes in the same rows."""
    from oracle import torch_native as tn
    ns = ref_loader.load()
    kw = dict(size=32, page_size=1, dtype=dtype, head_num=2, head_dim=8, layer_num=1, device="cpu")
    try:
        real = ns.MHATokenToKVPool(enable_memory_saver=False, **kw)
    except TypeError:
        real = ns.MHATokenToKVPool(**kw)
    g = torch.Generator().manual_seed(5)
    k = torch.randn((5, 2, 8), generator=g).to(torch.bfloat16)
    v = torch.randn((5, 2, 8), generator=g).to(torch.bfloat16)
    loc = torch.tensor([3, 9, 1, 30, 17])
    layer = type("L", (), {"layer_id": 0})()
    real.set_kv_buffer(layer, loc, k.clone(), v.clone(), scale, scale)
    tk = torch.zeros_like(real.k_buffer[0])
    tv = torch.zeros_like(real.v_buffer[0])
    tn.set_kv_buffer(tk, tv, loc, k.clone(), v.clone(), dtype, scale, scale)
    assert torch.equal(real.k_buffer[0].view(torch.uint8), tk.view(torch.uint8))
    assert torch.equal(real.v_buffer[0].view(torch.uint8), tv.view(torch.uint8))
