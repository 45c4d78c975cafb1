"""sglang_custom_b200 -- a B200 (sm_100a) paged-KV attention backend for SGLang.

Only what the attention hot path needs lives here:
  csrc/          hand-written CUDA kernels + the C ABI (include/b200attn.h) -> libb200attn.so
  _lib.py        ctypes binding of the C ABI (no fallback: a missing library raises)
  ops.py         tensor-level wrappers (torch supplies device memory and streams only)
  backend.py     B200AttnBackend: the SGLang AttentionBackend (same method signatures)
  interfaces.py  stand-ins for the SGLang objects the backend reads (tests / bench only)
  synth.py       seeded synthetic workloads of the BASELINE.json shapes
  build.py       nvcc build of the shared library, in-tree

Importing the package does not load the CUDA library; the first compute call does.
"""
from ._lib import B200AttnError, LIB_PATH  # noqa: F401

__version__ = "0.1.0"
__all__ = ["B200AttnError", "LIB_PATH", "__version__"]
