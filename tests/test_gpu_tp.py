"""Multi-GPU parity of the CUDA path under north_star's partition (SGLang tensor parallelism by KV head,
models/llama.py:118-131, configs/model_config.py:342-349): every rank runs ITS head slice of one global model through
B200AttnBackend on its own GPU, the outputs are all-gathered over the head dimension, and rank 0 compares a sample of
requests with the GLOBAL torch_native oracle; each rank's KV write is checked bit for bit over its whole pool.
Needs >= 2 CUDA devices (skipped otherwise); `gpurun --gpus 2 -- python -m pytest tests/test_gpu_tp.py -m gpu`."""
from __future__ import annotations

import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

CASES = {
    # name: (job, seq_lens, extend_lens, mode, sample)
    "decode_bf16": (dict(Hq=32, Hkv=8, D=128, fp8=False, kv_scale=1.0, page_size=1, seed=31),
                    [1500, 64, 300, 17, 4096, 2049], None, None, [0, 3, 4]),
    "decode_fp8_c4_heads": (dict(Hq=64, Hkv=8, D=128, fp8=True, kv_scale=0.1135, page_size=1, seed=32),
                            [900, 8192, 33, 2500], None, None, [1, 2, 3]),
    "extend_bf16": (dict(Hq=32, Hkv=8, D=128, fp8=False, kv_scale=1.0, page_size=1, seed=33),
                    [700, 384, 130], [300, 384, 129], None, [0, 1, 2]),
    "mixed_fp8_page16_c5_heads": (dict(Hq=64, Hkv=8, D=128, fp8=True, kv_scale=0.1135, page_size=16, seed=34),
                                  [16 * 20 + 256, 16 * 3 + 256, 700, 1999, 45], [256, 256, 1, 1, 1], "MIXED", [0, 1, 2, 3, 4]),
}


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, names, q):
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    sys.path.insert(0, str(root))
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import bench  # the parity leg of bench.py is the code under test (no timing here)
    from sglang_custom_b200 import tp
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    tp.init("nccl")
    results = {}
    for name in names:
        job, lens, ext, mode, sample = CASES[name]
        case, r, layers = bench._build(job, rank, world, dev, lens, extend_lens=ext, n_pools=1)
        _, (qq, kk, vv) = bench._qkv(case)
        fb = r.batch(getattr(r.itf.ForwardMode, mode) if mode else None)
        results[name] = bench._parity(job, case, r, layers[0], fb, qq, kk, vv, sample, rank, world, dev)
    if rank == 0:
        q.put(results)
    tp.barrier(world)
    torch.distributed.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_tp_partition_matches_global_oracle(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} CUDA devices")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    names = list(CASES)
    procs = [ctx.Process(target=_worker, args=(r, world, port, names, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = q.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for name, res in results.items():
        assert res["ranks"] == world
        assert res["pool_bit_exact"], f"{name}: KV write not bit-exact on some rank"
        assert res["ok"], f"{name}: max-abs {res['max_abs']} mean-abs {res['mean_abs']} vs the global oracle"
