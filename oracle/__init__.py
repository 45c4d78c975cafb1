"""CPU oracles for the paged-KV attention path.  TEST INFRASTRUCTURE ONLY.

Nothing under `sglang_custom_b200/` imports this package.  It is used by `tests/`,
by `__graft_entry__.smoke()` and by the `cpu_baseline` / `--impl reference` legs of
`bench.py`, always as the checker or as the timed CPU baseline -- never as the
product path.

Contents
  torch_native.py  restatement of the reference's torch_native attention backend
                   (python/sglang/srt/layers/attention/torch_native_backend.py:27-267) and of
                   MHATokenToKVPool.set_kv_buffer (mem_cache/memory_pool.py:420-458).  Parity
                   oracle of record (north_star: "torch_native/sgl-kernel CPU attention").
  exact.c/exact.py plain-C restatement with fp64 accumulation: FP8-E4M3 quantisation (bit
                   exact), page gather, kv-indices flatten, decode/extend attention, LSE merge.
                   Used for byte-exact checks and to attribute the residual error between the
                   CUDA path (fp32 accumulate) and the bf16 SDPA oracle.
  ref_loader.py    loads the UNMODIFIED reference files from /root/reference under stub
                   modules (only possible in the build container) -- pins torch_native.py.
  build_ref.py     compiles the reference's own sgl-kernel CPU attention (decode.cpp/extend.cpp)
                   from where it lies under /root/reference into oracle/_ref/ (git-ignored,
                   travels to the GPU box); second oracle + `cpu_baseline.kind == "reference"`.
  sgl_cpu.py       thin caller of oracle/_ref/sgl_cpu_attn*.so.

Pinning status: torch_native.py is pinned bit-exact against the unmodified reference run in
the build container (tests/test_oracle_pin.py, fixtures under tests/golden/ made by
tests/golden/make_golden.py); exact.c's fp8 cast is pinned exhaustively against torch's own
cast (all 65536 bf16 bit patterns).  The reference stores no golden vectors of its own
(SURVEY.md section 8c), so "outputs of the reference itself run here" is the pin.
"""
