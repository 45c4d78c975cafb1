"""Load the committed golden fixtures (tests/golden/*.pt, made by tests/golden/make_golden.py)."""
from __future__ import annotations

import types
from pathlib import Path

import torch

GOLDEN_DIR = Path(__file__).resolve().parent / "golden"

ATTN_FIXTURES = sorted(p.stem for p in GOLDEN_DIR.glob("*.pt") if not p.stem.startswith(("kvwrite", "rope", "masked", "draft")))
# sliding window / custom mask: outputs of the reference's Triton kernels (tests/golden/make_golden_masked.py)
MASKED_EXTEND_FIXTURES = sorted(p.stem for p in GOLDEN_DIR.glob("masked_*.pt") if "decode" not in p.stem)
MASKED_DECODE_FIXTURES = sorted(p.stem for p in GOLDEN_DIR.glob("masked_*decode*.pt"))
ROPE_FIXTURES = sorted(p.stem for p in GOLDEN_DIR.glob("rope*.pt"))
# DRAFT_EXTEND both ways (FlashInfer-style causal, Triton-style double count): tests/golden/make_golden_draft_extend.py
DRAFT_EXTEND_FIXTURES = sorted(p.stem for p in GOLDEN_DIR.glob("draft_extend_*.pt"))
KVWRITE_FIXTURES = sorted(p.stem for p in GOLDEN_DIR.glob("kvwrite*.pt"))


def load(name: str) -> types.SimpleNamespace:
    fx = torch.load(GOLDEN_DIR / f"{name}.pt", map_location="cpu", weights_only=True)
    ns = types.SimpleNamespace(**fx)
    if "dtype_name" in fx:
        ns.dtype = getattr(torch, fx["dtype_name"])
    return ns
