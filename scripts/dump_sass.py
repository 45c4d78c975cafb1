#!/usr/bin/env python
"""SASS evidence for libb200attn.so: which Blackwell instructions each kernel family contains.

  python scripts/dump_sass.py [--out profiles/r02_sass_summary.txt] [--lines profiles/r02_sass_lines.txt]

Runs `cuobjdump -sass` on the built library (no GPU needed), groups the functions by kernel name (template
arguments dropped) and counts the mnemonics that prove the design (profiling guide, "What proves a Blackwell-native
kernel"): UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UBLKCP = TMA, LDGSTS = cp.async,
SYNCS = mbarrier, HMMA = mma.sync, LDSM = ldmatrix, F2FP...E4M3 = the FP8 dequantisation.  `--lines` additionally
writes, per family, the first occurrences of every key mnemonic as they appear in the SASS (address + instruction).
`tests/test_cabi_cpu.py::test_library_targets_sm100a_and_uses_blackwell_instructions` asserts on `family_counts()`.
"""
from __future__ import annotations

import argparse
import re
import subprocess
import sys
from collections import Counter, defaultdict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "sglang_custom_b200" / "libb200attn.so"

KEY = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTMAPF", "LDGSTS", "SYNCS",
       "HMMA", "LDSM", "MOVM", "F2FP.E4M3", "MUFU.EX2", "FFMA2", "ATOMG", "REDG", "ACQBULK", "UCGABAR", "HGMMA"]
_FUNC = re.compile(r"^\s*Function : (\S+)")
_INSTR = re.compile(r"^\s+/\*([0-9a-f]{4,})\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*(?:\.[A-Z0-9_.]+)?)")


def _family(mangled: str) -> str:
    try:
        dem = subprocess.run(["c++filt", mangled], capture_output=True, text=True).stdout.strip() or mangled
    except OSError:
        dem = mangled
    dem = re.sub(r"^void\s+", "", dem)
    name = dem.split("(")[0]
    depth, out = 0, []
    for ch in name:  # drop template arguments
        if ch == "<":
            depth += 1
        elif ch == ">":
            depth -= 1
        elif depth == 0:
            out.append(ch)
    return "".join(out).replace("b200attn::", "")


def _key_of(op: str):
    if op.startswith("F2FP") and "E4M3" in op:
        return "F2FP.E4M3"
    if op.startswith("MUFU.EX2"):
        return "MUFU.EX2"
    base = op.split(".")[0]
    return base if base in KEY else None


def parse(lib: Path = LIB):
    """-> (counts[family][mnemonic], n_functions[family], first_lines[family][mnemonic] = [str, ...])"""
    res = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"cuobjdump failed: {res.stderr[:300]}")
    counts, nfunc, first = defaultdict(Counter), Counter(), defaultdict(lambda: defaultdict(list))
    fam, cache = None, {}
    for line in res.stdout.splitlines():
        m = _FUNC.match(line)
        if m:
            if m.group(1) not in cache:
                cache[m.group(1)] = _family(m.group(1))
            fam = cache[m.group(1)]
            nfunc[fam] += 1
            continue
        if fam is None:
            continue
        m = _INSTR.match(line)
        if not m:
            continue
        k = _key_of(m.group(2))
        if k:
            counts[fam][k] += 1
            if len(first[fam][k]) < 3:
                first[fam][k].append(line.split("/*", 2)[0].rstrip() + " " + line.strip()[:110])
    return counts, nfunc, first


def family_counts(lib: Path = LIB) -> dict:
    counts, _, _ = parse(lib)
    return {f: dict(c) for f, c in counts.items()}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--lines", default=None)
    a = ap.parse_args()
    counts, nfunc, first = parse()
    rows = [f"# cuobjdump -sass {LIB.relative_to(ROOT)}: key mnemonics per kernel family (all template instances summed)",
            "# UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UBLKCP = TMA, LDGSTS = cp.async, SYNCS = mbarrier,",
            "# HMMA = mma.sync, LDSM = ldmatrix, MOVM = movmatrix, F2FP.E4M3 = e4m3 <-> f16 conversion", ""]
    for fam in sorted(counts, key=lambda f: -sum(counts[f].values())):
        c = counts[fam]
        rows.append(f"{fam}  [{nfunc[fam]} instance(s)]")
        rows.append("    " + "  ".join(f"{k}={c[k]}" for k in KEY if c.get(k)))
    text = "\n".join(rows) + "\n"
    if a.out:
        Path(a.out).write_text(text)
    sys.stdout.write(text)
    if a.lines:
        out = ["# first occurrences of the key instructions, per kernel family (cuobjdump -sass)"]
        for fam in sorted(first):
            out.append(f"\n== {fam}")
            for k in KEY:
                for ln in first[fam].get(k, []):
                    out.append(f"  {ln}")
        Path(a.lines).write_text("\n".join(out) + "\n")


if __name__ == "__main__":
    main()
