"""Generate tests/golden/draft_kv_indices.pt by running the UNMODIFIED reference kernel
`generate_draft_decode_kv_indices` (python/sglang/srt/speculative/eagle_utils.py:821-899) on the CPU through Triton's
interpreter.  Only runs in the build container:
  python tests/golden/make_golden_draft.py
eagle_utils.py imports half of SGLang, so the kernel's source text is cut out of the file (ast) and executed as it
stands in a namespace holding triton / tl; nothing of it is stored in the repo -- only its outputs are.
The fixture pins sglang_custom_b200/draft_backend.py::build_branch_table (tests/test_draft_cpu.py): for every step i
and row r = b * topk + t the reference list kv_indices[i][kv_indptr[i][r] : kv_indptr[i][r + 1]] must equal the first
seq_lens[b] + i + 1 entries of branch-table row r.
"""
from __future__ import annotations

import ast
import os
import sys
from pathlib import Path

os.environ["TRITON_INTERPRET"] = "1"
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402
import triton  # noqa: E402
import triton.language as tl  # noqa: E402

REF = Path("/root/reference/python/sglang/srt/speculative/eagle_utils.py")
OUT = Path(__file__).resolve().parent / "draft_kv_indices.pt"


def load_kernel():
    import importlib.util
    import tempfile
    text = REF.read_text()
    tree = ast.parse(text)
    node = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "generate_draft_decode_kv_indices")
    lines = text.splitlines()
    src = "import triton\nimport triton.language as tl\n\n\n" + "\n".join(lines[node.decorator_list[0].lineno - 1:node.end_lineno]) + "\n"
    with tempfile.NamedTemporaryFile("w", suffix="_ref_draft_kernel.py", delete=False) as f:  # the interpreter wants a file
        f.write(src)
        path = f.name
    spec = importlib.util.spec_from_file_location("ref_draft_kernel", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    os.unlink(path)
    return mod.generate_draft_decode_kv_indices


def next_pow2(n):
    return 1 << (max(int(n), 1) - 1).bit_length()


def run_case(kernel, seq_lens, topk, num_steps, page_size, width, seed):
    g = torch.Generator().manual_seed(seed)
    bs = len(seq_lens)
    max_reqs = bs + 3
    req_to_token = torch.randint(1, 1 << 20, (max_reqs, width), generator=g, dtype=torch.int32)
    req_pool_indices = torch.randperm(max_reqs, generator=g)[:bs].to(torch.int64)
    sl = torch.tensor(seq_lens, dtype=torch.int64)
    n = bs * topk
    # positions as eagle_worker.draft_forward passes them: seq_lens repeated per branch (eagle_utils.py:84-105)
    positions = sl.repeat_interleave(topk)
    kv_indices = torch.zeros((num_steps, n * (max(seq_lens) + num_steps) + 64), dtype=torch.int32)
    kv_indptr = torch.zeros((num_steps, n + 1), dtype=torch.int32)
    kernel[(num_steps, bs, topk)](req_pool_indices, req_to_token, sl, kv_indices, kv_indptr, positions, width,
                                  kv_indices.shape[1], kv_indptr.shape[1], next_pow2(bs), next_pow2(num_steps),
                                  next_pow2(n), page_size)
    return dict(seq_lens=sl, topk=topk, num_steps=num_steps, page_size=page_size, req_to_token=req_to_token,
                req_pool_indices=req_pool_indices, ref_kv_indices=kv_indices, ref_kv_indptr=kv_indptr)


def main():
    kernel = load_kernel()
    cases = [
        run_case(kernel, [37, 5, 129], 1, 3, 1, 256, 11),
        run_case(kernel, [37, 5, 129, 64], 4, 5, 1, 256, 12),
        run_case(kernel, [48, 17, 100], 2, 3, 16, 256, 13),   # page_size > 1 and topk > 1: page-aligned branch starts
        run_case(kernel, [31, 16], 8, 2, 16, 320, 14),
        run_case(kernel, [200], 1, 4, 16, 256, 15),           # topk == 1 with pages: draft tokens follow the prefix
    ]
    torch.save(cases, OUT)
    print(f"wrote {OUT} ({len(cases)} cases)")


if __name__ == "__main__":
    main()
