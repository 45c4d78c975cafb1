"""Compile the reference's own sgl-kernel CPU attention into oracle/_ref/.  TEST INFRASTRUCTURE.

Sources are compiled FROM WHERE THEY LIE under /root/reference (nothing is copied into the repo):
  sgl-kernel/csrc/cpu/decode.cpp   decode_attention_cpu  (:1375-1575)
  sgl-kernel/csrc/cpu/extend.cpp   extend_attention_cpu  (:579-723)
plus a 12-line pybind shim written to a scratch directory that forward-declares those two entry
points and exposes them.  The reference's own build system (CMake + FetchContent) is NOT run; the
two translation units only need the torch headers of this image and `-fopenmp`.

Two variants are built because the GPU box's host CPU may differ from the build container's:
  sgl_cpu_attn_avx512.so   -mavx512{f,bw,vl,dq,bf16} -mamx-{tile,bf16}: turns on the reference's
                           CPU_CAPABILITY_AVX512 paths (vec.h:3-5) -- its fast path on Xeon hosts
  sgl_cpu_attn_generic.so  -march=x86-64-v3: the reference's scalar fallbacks (decode.cpp:101-126)
oracle/sgl_cpu.py picks one from /proc/cpuinfo at load time.

Output: oracle/_ref/*.so (git-ignored; NOT gpurun-ignored, so the built files travel to the GPU
box, where /root/reference does not exist).  Run: `python -m oracle.build_ref`.
"""
from __future__ import annotations

import os
import shutil
import sys
import tempfile
from pathlib import Path

REF_ROOT = Path(os.environ.get("B200ATTN_REFERENCE", "/root/reference"))
CPU_DIR = REF_ROOT / "sgl-kernel" / "csrc" / "cpu"
OUT_DIR = Path(__file__).resolve().parent / "_ref"
VARIANTS = {
    "sgl_cpu_attn_avx512": ["-mavx512f", "-mavx512bw", "-mavx512vl", "-mavx512dq", "-mavx512bf16", "-mamx-tile",
                            "-mamx-bf16", "-mfma", "-mf16c", "-mavx2"],
    "sgl_cpu_attn_generic": ["-march=x86-64-v3"],
}

_BIND = r"""
#include <torch/extension.h>
void decode_attention_cpu(at::Tensor& query, at::Tensor& k_buffer, at::Tensor& v_buffer, at::Tensor& output,
                          at::Tensor& key, at::Tensor& value, at::Tensor& loc, at::Tensor& attn_logits,
                          at::Tensor& req_to_token, at::Tensor& req_pool_indices, at::Tensor& seq_lens,
                          double sm_scale, double logit_cap);
void extend_attention_cpu(at::Tensor& q_extend, at::Tensor& k_extend, at::Tensor& v_extend, at::Tensor& o_extend,
                          at::Tensor& k_buffer, at::Tensor& v_buffer, at::Tensor& req_to_token,
                          at::Tensor& req_pool_indices, at::Tensor& seq_lens, at::Tensor& extend_seq_lens,
                          at::Tensor& extend_start_loc, int64_t max_len_extend, double sm_scale, double logit_cap);
PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("decode_attention_cpu", &decode_attention_cpu);
  m.def("extend_attention_cpu", &extend_attention_cpu);
}
"""


def built_path(module: str) -> Path | None:
    p = OUT_DIR / f"{module}.so"
    return p if p.is_file() else None


def available_sources() -> bool:
    return (CPU_DIR / "decode.cpp").is_file() and (CPU_DIR / "extend.cpp").is_file()


def _build_one(module: str, isa_flags: list[str], verbose: bool) -> None:
    from torch.utils import cpp_extension

    # the image's $CXX wrapper (/opt/gcc/bin/g++) ships without libgomp.spec; the distro g++ has it
    if Path("/usr/bin/g++").exists():
        os.environ["CXX"] = "/usr/bin/g++"
    scratch = Path(tempfile.mkdtemp(prefix="b200attn_ref_"))
    try:
        bind = scratch / "bind.cpp"
        bind.write_text(_BIND)
        cpp_extension.load(
            name=module, sources=[str(bind), str(CPU_DIR / "decode.cpp"), str(CPU_DIR / "extend.cpp")],
            extra_include_paths=[str(CPU_DIR)],
            extra_cflags=["-O3", *isa_flags, "-fopenmp", "-Wno-unknown-pragmas", "-std=c++17"],
            extra_ldflags=["-fopenmp"], build_directory=str(scratch), verbose=verbose, is_python_module=False)
        shutil.copy2(scratch / f"{module}.so", OUT_DIR / f"{module}.so")
    finally:
        shutil.rmtree(scratch, ignore_errors=True)


def build(force: bool = False, verbose: bool = False) -> list[Path]:
    """Builds whatever is missing (when the reference sources are mounted); returns the .so files present."""
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    for module, flags in VARIANTS.items():
        if (force or built_path(module) is None) and available_sources():
            _build_one(module, flags, verbose)
    return [p for p in (built_path(m) for m in VARIANTS) if p is not None]


if __name__ == "__main__":
    got = build(force="--force" in sys.argv, verbose=True)
    print("built: " + ", ".join(str(p) for p in got) if got else "reference sources not available, nothing prebuilt")
