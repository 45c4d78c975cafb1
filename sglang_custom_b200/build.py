"""Build the C-ABI CUDA library `libb200attn.so` in-tree with nvcc for sm_100a.

The library has no torch dependency (plain CUDA runtime, statically linked), so the
same file serves ctypes here and any cgo/JNI/N-API style binding a host would add.
Run as `python -m sglang_custom_b200.build [--force]`.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
INCLUDE = PKG_DIR.parent / "include"
OBJ_DIR = CSRC / "_build"
LIB_PATH = PKG_DIR / "libb200attn.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC",
    "-Xcompiler", "-fvisibility=hidden",
]


def _nvcc() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(cand).exists():
        raise RuntimeError("nvcc not found; set NVCC=/path/to/nvcc")
    return cand


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _deps_mtime() -> float:
    hdrs = list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))
    return max(p.stat().st_mtime for p in hdrs)


def _compile(src: Path, force: bool, verbose: bool, extra: tuple = (), obj_dir: Path = OBJ_DIR) -> Path:
    obj = obj_dir / (src.stem + ".o")
    newest = max(src.stat().st_mtime, _deps_mtime())
    if not force and obj.exists() and obj.stat().st_mtime >= newest:
        return obj
    cmd = [_nvcc(), *NVCC_FLAGS, *extra, "-I", str(INCLUDE), "-c", str(src), "-o", str(obj)]
    if verbose:
        print(" ".join(cmd), flush=True)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stdout}\n{res.stderr}")
    return obj


def build(force: bool = False, verbose: bool = False, debug_wait: bool = False) -> Path:
    """Compile every .cu under csrc/ and link libb200attn.so. Returns the library path.
    debug_wait: development variant libb200attn_dbg.so (-DB200ATTN_DEBUG_WAIT: timed-out mbarrier waits are recorded
    instead of trapping); load it with B200ATTN_LIB=<path>."""
    obj_dir = OBJ_DIR / "dbg" if debug_wait else OBJ_DIR
    lib_path = PKG_DIR / "libb200attn_dbg.so" if debug_wait else LIB_PATH
    extra = ("-DB200ATTN_DEBUG_WAIT",) if debug_wait else ()
    obj_dir.mkdir(parents=True, exist_ok=True)
    srcs = sources()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, force, verbose, extra, obj_dir), srcs))
    newest_obj = max(o.stat().st_mtime for o in objs)
    if force or not lib_path.exists() or lib_path.stat().st_mtime < newest_obj:
        cmd = [_nvcc(), "-shared", "-o", str(lib_path), *[str(o) for o in objs],
               "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static",
               "-Xcompiler", "-fPIC"]
        if verbose:
            print(" ".join(cmd), flush=True)
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    return lib_path


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose=True, debug_wait="--debug-wait" in sys.argv)
    print(f"built {path}")
