import sys, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent)); sys.path.insert(0, str(Path(__file__).resolve().parent))
import torch
from sglang_custom_b200 import ops, synth
from sweep_shapes import time_ms
dev = torch.device("cuda:0")
for bs, ctx in ((1, 4096), (1, 32768), (4, 8192), (16, 4096)):
    case = synth.make_case_device(bs, 32, 8, 128, [ctx] * bs, dev, seed=1240)
    out = torch.empty_like(case.q)
    n_ctas = 148
    plan = torch.zeros(ops.decode_plan_bytes(bs, n_ctas) // 4, dtype=torch.int32, device=dev)
    ws = torch.empty(ops.decode_workspace_bytes(n_ctas, 32, 128) // 4, dtype=torch.float32, device=dev)
    for mu in (1, 2, 4, 8):
        ops.decode_plan(case.seq_lens, n_ctas, plan, mu)
        fn = lambda: ops.decode(case.q, out, case.k_pool, case.v_pool, case.req_to_token, case.req_pool_indices,
                                case.seq_lens, plan, n_ctas, ws, case.sm_scale)
        side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side): fn()
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for _ in range(10): fn()
        us = time_ms(g.replay, reps=5, warm=2) * 100
        print(f"bs{bs} ctx{ctx} min_units {mu}: {us:.1f} us", flush=True)
        del g
