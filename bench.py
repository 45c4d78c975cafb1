#!/usr/bin/env python
"""bench.py -- the attention hot path on B200, beside the reference's CPU attention.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Metric (BASELINE.json): decode-attention HBM GB/s (+ tokens/s/GPU) at Llama-3-8B shapes, bs 256, ctx 4 K,
bf16 KV, page_size 1 -- BASELINE configs[1] ("C2") -- with the other GPU configs as objects on the same line:
`extend` (configs[2], C3, TFLOP/s), `c4_tp` (configs[3], Llama-3-70B heads, FP8 KV, bs 512, ctx 8 K) and `c5_mixed`
(configs[4], Qwen2-72B heads, FP8 KV, page_size 16, one MIXED forward).

A STEP is one model forward's worth of decode attention for the batch: the metadata hook once, then `forward_decode`
(KV write + split-KV decode + LSE merge) for each of the 32 layers -- the calls SGLang makes per forward -- run the way
SGLang runs decode in production: captured once through `init_forward_metadata_capture_cuda_graph` and replayed after
`init_forward_metadata_replay_cuda_graph` refreshed the plan in place (cuda_graph_runner.py:456-716).  The eager figure
is reported beside it (`eager`).  Layers rotate over two distinct pool copies; one layer reads far more than the 126 MB
L2, so no L2 flush is needed between iterations (config.l2).

  value     whole-job algorithmic GB/s, every input resident in HBM (SURVEY.md 8d bytes / device time, MAX over ranks)
  e2e       the same through the public API (RadixAttention.forward -> B200AttnBackend) with HOST q/k/v: every layer's
            q, k, v are copied from pinned host memory and every layer's output is read back, inside the timed region
  roofline  the dominant kernel (decode stage 1) timed alone with CUDA events on its launch stream
  parity_n  before timing: one forward per rank, outputs all-gathered over the head dimension and a sample of requests
            checked on rank 0 against the GLOBAL torch_native oracle (checker only, outside every timed region); the
            KV write is checked bit for bit over each rank's whole pool
  cpu_baseline  the reference's own sgl-kernel CPU decode (oracle/_ref) on a bounded sample, on this host's cores

N > 1 (torchrun): STRONG scaling of ONE fixed model -- north_star's partition, exactly SGLang tensor parallelism
(models/llama.py:118-131, configs/model_config.py:342-349): rank r owns Hq/N query heads and max(1, Hkv/N) KV heads of
every config above; the page table is replicated; no collective inside the path (the harness's own barrier, MAX-reduce
of device times and the parity all-gather are the only ones).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

# stdout carries exactly one JSON line.  Native libraries write to file descriptor 1 behind Python's back (NCCL prints
# "NCCL version ..." there on boxes that export NCCL_DEBUG), so fd 1 is pointed at stderr for the whole run and the
# JSON line goes to a private duplicate of the original stdout.
sys.stdout.flush()
_JSON_OUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)

import torch  # noqa: E402

READ_STREAM_CEILING_GBS = 7272.3  # coalesced read-only stream, scripts/ubench/hbm_read.cu (profiles/r01_ubench_hbm_read.txt)

# The GPU configs of BASELINE.json as GLOBAL models; every rank takes its head slice (synth.make_case_device head_shard)
JOBS = {
    "c2": dict(tag="BASELINE configs[1]", model="Llama-3-8B", bs=256, ctx=4096, Hq=32, Hkv=8, D=128, fp8=False,
               kv_scale=1.0, page_size=1, layers=32, seed=1236, sample=[0, 17, 101, 255]),
    "c4": dict(tag="BASELINE configs[3]", model="Llama-3-70B", bs=512, ctx=8192, Hq=64, Hkv=8, D=128, fp8=True,
               kv_scale=0.1135, page_size=1, layers=80, seed=1241, sample=[0, 300, 511]),
}
# dram__bytes_read.sum + dram__bytes_write.sum of decode_kernel, one launch, from the committed `ncu --set full`
# captures: {(job, world): (bytes, source file)}; other shapes report null
NCU_TRAFFIC = {("c2", 1): (4_302_502_000 + 7_314_176, "profiles/r02_final_ncu_decode_c2.txt"),
               ("c4", 1): (8_617_694_000 + 16_515_328, "profiles/r02_final_ncu_decode_c4_fp8_tp1.txt")}


def _peaks() -> dict:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return dict(hbm=float(d["hbm_gbs"]), tf_burst=float(d["bf16_tflops"]),
                    tf_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w": statistics.median(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def _gpu_index(local: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local])
        except (ValueError, IndexError):
            pass
    return local


# ------------------------------------------------------------------------------------------ shared pieces
def _timed(fn, steps, warmup, stream, dev, world):
    """W untimed calls, then K calls between two CUDA events on the launch stream, bracketed by barrier + synchronize;
    milliseconds per call, MAX over ranks."""
    from sglang_custom_b200 import tp
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(dev)
    tp.barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        fn()
    e1.record(stream)
    torch.cuda.synchronize(dev)
    tp.barrier(world)
    return tp.max_over_ranks(e0.elapsed_time(e1), world, dev) / steps


def _kv_scale(job):
    return job["kv_scale"] if job.get("fp8") and job.get("kv_scale", 1.0) != 1.0 else None


def _build(job, rank, world, dev, seq_lens, extend_lens=None, n_pools=2):
    """Rank's head slice of the global job as the backend sees it: ModelRunner stand-in (global head counts + tp_size),
    B200AttnBackend, `n_pools` layer pools with identical content at distinct addresses, one RadixAttention per pool."""
    from sglang_custom_b200 import harness, synth
    case = synth.make_case_device(len(seq_lens), job["Hq"], job["Hkv"], job["D"], seq_lens, dev,
                                  extend_lens=extend_lens, page_size=job["page_size"], fp8=job["fp8"],
                                  kv_scale=job["kv_scale"], seed=job["seed"], head_shard=(rank, world))
    r = harness.Runner(case, dev, tp_size=world, tp_rank=rank)
    pool = r.mr.token_to_kv_pool
    for _ in range(n_pools - 1):
        pool.k_buffer.append(pool.k_buffer[0].clone())
        pool.v_buffer.append(pool.v_buffer[0].clone())
    s = _kv_scale(job)
    layers = [r.itf.RadixAttention(num_heads=case.Hq, head_dim=case.D, scaling=case.sm_scale, num_kv_heads=case.Hkv,
                                   layer_id=i, k_scale_float=s, v_scale_float=s) for i in range(n_pools)]
    return case, r, layers


def _qkv(case):
    T = case.T
    qkv = torch.cat([case.q.reshape(T, -1), case.k_new.reshape(T, -1), case.v_new.reshape(T, -1)], 1)
    return qkv, qkv.split([case.Hq * case.D, case.Hkv * case.D, case.Hkv * case.D], dim=1)


def _parity(job, case, r, layer, fb, q, k, v, sample, rank, world, dev, mode=None):
    """One forward through the backend on every rank; outputs of the sampled requests are all-gathered over the head
    dimension and compared on rank 0 with the GLOBAL torch_native oracle (the reference's CPU attention restated,
    oracle/torch_native.py -- the checker, never the thing measured); each rank's KV write is compared bit for bit with
    the oracle's cast over the rows it wrote, and every other row of the pool must be untouched."""
    sys.path.insert(0, str(ROOT / "tests"))
    import helpers as H
    from oracle import torch_native as tn
    from sglang_custom_b200 import synth, tp
    pool = r.mr.token_to_kv_pool
    lid = layer.layer_id
    before_k, before_v = pool.k_buffer[lid].clone(), pool.v_buffer[lid].clone()
    r.backend.init_forward_metadata(fb)
    out = layer.forward(q, k, v, fb).view(case.T, case.Hq, case.D)
    torch.cuda.synchronize(dev)
    after_k, after_v = pool.k_buffer[lid], pool.v_buffer[lid]
    # ---- KV write, whole pool: written rows == oracle cast, all other rows unchanged
    T = case.T
    tk = torch.zeros((T + 1, case.Hkv, case.D), dtype=after_k.dtype)
    tv = torch.zeros_like(tk)
    kn, vn = case.k_new.cpu().clone(), case.v_new.cpu().clone()
    s = _kv_scale(job)
    tn.set_kv_buffer(tk, tv, torch.arange(1, T + 1), kn, vn, torch.float8_e4m3fn if job["fp8"] else case.dtype, s, s)
    loc = case.out_cache_loc
    ok = H.pools_equal(after_k[loc], tk[1:]) and H.pools_equal(after_v[loc], tv[1:])
    before_k[loc] = after_k[loc]
    before_v[loc] = after_v[loc]
    ok = ok and torch.equal(before_k, after_k) and torch.equal(before_v, after_v)
    del before_k, before_v
    pool_ok = tp.all_true(ok, world, dev)
    # ---- attention: sampled requests, global heads
    sub = synth.extract_subcase(case, sample, after_k, after_v)
    kv_rep = max(1, world // job["Hkv"])
    gsub = tp.gather_subcase(sub, world, kv_rep)
    got = tp.gather_cat(out[sub.token_rows.to(dev)], world)
    res = None
    if rank == 0:
        gsub.fp8, gsub.kv_scale = job["fp8"], job["kv_scale"]
        want = (H.oracle_decode if case.decode else H.oracle_extend)(gsub, gsub.k_pool, gsub.v_pool)
        d = (got.float() - want.float()).abs()
        res = {"max_abs": round(float(d.max()), 6), "mean_abs": round(float(d.mean()), 7),
               "tol": {"max_abs": H.MAX_ABS_TOL, "mean_abs": H.MEAN_ABS_TOL},
               "ok": bool(d.max() <= H.MAX_ABS_TOL and d.mean() <= H.MEAN_ABS_TOL and torch.isfinite(got.float()).all()),
               "pool_bit_exact": bool(pool_ok), "ranks": world, "heads_checked": int(got.shape[1]),
               "requests_checked": list(sample), "tokens_checked": int(got.shape[0]),
               "oracle": "torch_native restatement (oracle/torch_native.py), global heads, CPU"}
    return res


# ------------------------------------------------------------------------------------------ decode jobs (C2, C4)
def _run_decode_job(name, args, rank, world, local, dev, peaks, headline):
    """headline=True: every leg (graph step, eager, e2e, roofline, clocks); False: graph step + kernel-only."""
    from sglang_custom_b200 import ops, synth, tp
    job = JOBS[name]
    bs, ctx, D = job["bs"], job["ctx"], job["D"]
    kvb = 1 if job["fp8"] else 2
    case, r, layers = _build(job, rank, world, dev, [ctx] * bs)
    be = r.backend
    Hq, Hkv, T = case.Hq, case.Hkv, case.T
    qkv, (q, k, v) = _qkv(case)
    fb = r.batch()
    LAY = job["layers"] if headline else min(job["layers"], 16)
    bytes_rank = synth.algorithmic_decode_bytes([ctx] * bs, Hq, Hkv, D, kvb)
    bytes_job = world * bytes_rank  # every rank reads its own replica of the page table
    stream = torch.cuda.current_stream(dev)
    steps, warmup = (args.steps, args.warmup) if headline else (max(5, args.steps // 2), 3)

    parity = _parity(job, case, r, layers[0], fb, q, k, v, job["sample"], rank, world, dev)

    def step_eager():
        be.init_forward_metadata(fb)
        for l in range(LAY):
            layers[l % len(layers)].forward(q, k, v, fb)

    # ---- the production decode step: captured once through the capture hook, replayed after the replay hook
    be.init_cuda_graph_state(bs, bs)
    g_rpi, g_sl = case.req_pool_indices.to(torch.int32), case.seq_lens.to(torch.int32)
    fbg = r.batch()
    fbg.req_pool_indices, fbg.seq_lens = g_rpi, g_sl
    be.init_forward_metadata_capture_cuda_graph(bs, bs, g_rpi, g_sl, None, fbg.forward_mode, None)
    cap = torch.cuda.Stream(dev)
    cap.wait_stream(stream)
    with torch.cuda.stream(cap):
        layers[0].forward(q, k, v, fbg)
    stream.wait_stream(cap)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=cap):
        for l in range(LAY):
            layers[l % len(layers)].forward(q, k, v, fbg)

    def step_graph():
        be.init_forward_metadata_replay_cuda_graph(bs, g_rpi, g_sl, int(bs * ctx), None, fbg.forward_mode, None, None)
        graph.replay()

    sampler = ClockSampler(_gpu_index(local))
    if rank == 0 and headline:
        sampler.start()
    ms_step = _timed(step_graph, steps, warmup, stream, dev, world)
    clocks = sampler.stop() if rank == 0 and headline else None
    del graph
    ms_eager = _timed(step_eager, max(3, steps // 2), 3, stream, dev, world)
    be.forward_metadata = None

    # ---- roofline leg: the dominant kernel (decode stage 1) alone, CUDA events on its stream
    pool = r.mr.token_to_kv_pool
    out = torch.empty((T, Hq, D), dtype=case.dtype, device=dev)
    q3 = q.view(T, Hq, D)
    ops.decode_plan(case.seq_lens, be.n_ctas, be.plan, be.min_units_per_cta)
    s = job["kv_scale"] if job["fp8"] else 1.0

    def kernel_only():
        for l in range(LAY):
            i = l % len(layers)
            ops.decode(q3, out, pool.k_buffer[i], pool.v_buffer[i], r.mr.req_to_token_pool.req_to_token,
                       case.req_pool_indices, case.seq_lens, be.plan, be.n_ctas, be.workspace, case.sm_scale,
                       0.0, s, s, kv_dtype=pool.dtype, stage="partial")
    ms_kernel = _timed(kernel_only, steps, warmup, stream, dev, world) / LAY
    kernel_bytes = bs * ctx * Hkv * 2 * D * kvb + bs * ctx * 4 + 2 * bs * Hq * D * 2  # KV + table + Q + O, per rank
    achieved = kernel_bytes / (ms_kernel * 1e-3) / 1e9
    traffic = NCU_TRAFFIC.get((name, world), (None, None))
    roofline = {"bound": "hbm", "kernel": "decode_kernel (split-KV stage 1), per GPU", "achieved": round(achieved, 1),
                "peak": peaks["hbm"], "unit": "GB/s", "frac": round(achieved / peaks["hbm"], 4),
                "peak_source": f"{peaks['source']} (MEASURED_PEAKS.json hbm_gbs, a COPY figure: read + write)",
                "frac_of_nominal_8TBs": round(achieved / 8000.0, 4),
                "frac_of_read_stream_ceiling": round(achieved / READ_STREAM_CEILING_GBS, 4),
                "read_stream_ceiling_gbs": READ_STREAM_CEILING_GBS,
                "us_per_launch": round(ms_kernel * 1e3, 2), "bytes_per_launch": kernel_bytes,
                "traffic": traffic[0], "traffic_source": traffic[1]}

    gbs = LAY * bytes_job / (ms_step * 1e-3) / 1e9
    gbs_eager = LAY * bytes_job / (ms_eager * 1e-3) / 1e9
    shape = (f"{job['tag']}: {job['model']}-shaped paged decode, bs {bs}, ctx {ctx}, {job['Hq']} Q / {job['Hkv']} KV heads"
             f" (per rank {Hq} / {Hkv}), D {D}, {'FP8-E4M3' if job['fp8'] else 'bf16'} KV, page_size {job['page_size']}, "
             f"permuted page table")
    if not headline:
        return {"workload": shape + f"; step = {LAY} layers x (KV write + split-KV decode + LSE merge) under CUDA-graph replay",
                "value": round(gbs, 1), "unit": "GB/s", "gbs_per_gpu": round(gbs / world, 1),
                "ms_per_layer": round(ms_step / LAY, 4), "tokens_per_s_per_layer_equiv": round(bs / (ms_step / LAY * 1e-3), 1),
                "eager": {"value": round(gbs_eager, 1), "ms_per_layer": round(ms_eager / LAY, 4)},
                "bytes_per_layer_per_rank": bytes_rank, "roofline": roofline, "parity_n": parity,
                "limits": "per-rank decode kernel on %d KV head(s) with %d-byte rows; ~3 launches per layer of fixed cost"
                          % (Hkv, Hkv * D * kvb)}, None

    # ---- e2e: host q/k/v per layer in pinned memory, output read back per layer, inside the timed region.  The copies
    # ride two side streams so that layer l+1's inputs and layer l-1's output cross PCIe while layer l computes (what
    # a host integration does with pinned buffers); the whole step -- copies, plan, kernels -- is one CUDA graph.
    hq = torch.empty((T, (Hq + 2 * Hkv) * D), dtype=case.dtype).pin_memory()
    hq.copy_(qkv.cpu())
    hout = torch.empty((T, Hq * D), dtype=case.dtype).pin_memory()
    h_idx = torch.stack([case.req_pool_indices.cpu(), case.seq_lens.cpu(), case.out_cache_loc.cpu()]).pin_memory()
    d_qkv = [torch.empty_like(qkv) for _ in range(2)]
    d_idx = torch.empty_like(h_idx, device=dev)
    d_out = [torch.empty((T, Hq * D), dtype=case.dtype, device=dev) for _ in range(2)]
    h2d = LAY * hq.numel() * 2 + h_idx.numel() * 8
    d2h = LAY * hout.numel() * 2
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    fb2 = r.batch()  # its index tensors are rows of d_idx: refreshed from the host every step
    fb2.req_pool_indices, fb2.seq_lens, fb2.out_cache_loc = d_idx[0], d_idx[1], d_idx[2]

    def e2e_body(main):
        in_ready = [torch.cuda.Event() for _ in range(2)]
        in_free = [torch.cuda.Event() for _ in range(2)]
        out_ready = [torch.cuda.Event() for _ in range(2)]
        out_free = [torch.cuda.Event() for _ in range(2)]
        d_idx.copy_(h_idx, non_blocking=True)
        be.init_forward_metadata(fb2)
        for b in range(2):
            in_free[b].record(main)
            out_free[b].record(main)
        for l in range(LAY):
            b = l & 1
            with torch.cuda.stream(s_in):
                s_in.wait_event(in_free[b])
                d_qkv[b].copy_(hq, non_blocking=True)
                in_ready[b].record(s_in)
            main.wait_event(in_ready[b])
            main.wait_event(out_free[b])
            dq, dk, dv = d_qkv[b].split([Hq * D, Hkv * D, Hkv * D], dim=1)
            o = layers[l % len(layers)].forward(dq, dk, dv, fb2)
            d_out[b].copy_(o)  # the layer output lives in a graph-private buffer: stage it where the copy-out reads
            in_free[b].record(main)
            out_ready[b].record(main)
            with torch.cuda.stream(s_out):
                s_out.wait_event(out_ready[b])
                hout.copy_(d_out[b], non_blocking=True)
                out_free[b].record(s_out)
        main.wait_stream(s_in)
        main.wait_stream(s_out)

    e2e_mode = "cuda_graph"
    try:
        cap2 = torch.cuda.Stream(dev)
        cap2.wait_stream(stream)
        with torch.cuda.stream(cap2):
            e2e_body(cap2)
        stream.wait_stream(cap2)
        torch.cuda.synchronize(dev)
        g2 = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g2, stream=cap2):
            e2e_body(cap2)

        def step_e2e():
            g2.replay()
            stream.synchronize()  # the host owns `hout` again: the result of the step has arrived
    except Exception as e:  # capture of the copy pipeline failed on this box: run the same pipeline eagerly
        print(f"[bench] e2e graph capture failed ({type(e).__name__}: {e}); eager copy pipeline", file=sys.stderr)
        torch.cuda.synchronize(dev)
        e2e_mode = "eager"

        def step_e2e():
            e2e_body(stream)
            stream.synchronize()
    ms_e2e = _timed(step_e2e, max(3, steps // 2), 3, stream, dev, world)
    gbs_e2e = LAY * bytes_job / (ms_e2e * 1e-3) / 1e9
    # the host copy of the last layer's output must be what the device computed
    e2e_ok = bool(torch.equal(hout, d_out[(LAY - 1) & 1].cpu()))

    line = {
        "metric": "decode-attn HBM GB/s (Llama-3-8B shapes, bs=256, ctx=4K, bf16 KV)", "value": round(gbs, 1),
        "unit": "GB/s", "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": shape + f"; step = replay hook (plan refresh) + one CUDA graph of {LAY} layers x "
                               "(KV write + split-KV decode + LSE merge), the production decode path",
                   "layers_per_step": LAY, "bytes_per_layer_per_rank": bytes_rank, "bytes_per_layer_job": bytes_job,
                   "l2": f"inputs larger than L2 ({bytes_rank / 1e9:.2f} GB read per layer per GPU vs 126 MB L2; layers "
                         "alternate between two pool copies); no flush",
                   "parallelism": f"tp{world}: ONE model split by KV head, {Hq} Q / {Hkv} KV heads per rank, page table "
                                  "replicated, no collective in the path", "n_ctas": be.n_ctas},
        "tokens_per_s_per_gpu": round(bs / (ms_step * 1e-3) / world, 1),
        "tokens_per_s": round(bs / (ms_step * 1e-3), 1),
        "gbs_per_gpu": round(gbs / world, 1),
        "eager": {"value": round(gbs_eager, 1), "unit": "GB/s", "ms_per_step": round(ms_eager, 4),
                  "what": "the same step launched eagerly (init_forward_metadata + per-layer forward from Python)"},
        "e2e": {"value": round(gbs_e2e, 1), "unit": "GB/s", "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h * world,
                "ms_per_step": round(ms_e2e, 4), "tokens_per_s": round(bs / (ms_e2e * 1e-3), 1), "mode": e2e_mode,
                "host_result_matches_device": e2e_ok,
                "what": "per layer: q,k,v copied from pinned host memory, output copied back to pinned host memory, all "
                        "inside the timed region; the host waits for the last output every step"},
        "gpu_launches": world * steps * (1 + 3 * LAY),
        "roofline": roofline, "parity_n": parity, "clocks": clocks,
    }
    if world > 1:
        line["limits"] = (f"strong scaling: per-rank KV rows shrink to {Hkv * D * kvb} bytes and the per-layer kernel to "
                          f"{ms_kernel * 1e3:.0f} us, against ~3 launches of fixed cost per layer; no collective in the path")
    return line, None


# ------------------------------------------------------------------------------------------ extend (C3)
def _run_extend(args, rank, world, local, dev, peaks):
    """BASELINE configs[2]: bs 16, 2048 cached prefix + 2048 new tokens, 32/8 heads, D 128, one forward_extend."""
    from sglang_custom_b200 import synth, tp
    job = dict(tag="BASELINE configs[2]", model="Llama-3-8B", Hq=32, Hkv=8, D=128, fp8=False, kv_scale=1.0, page_size=1,
               seed=1237)
    bs, P, E = 16, 2048, 2048
    case, r, layers = _build(job, rank, world, dev, [P + E] * bs, extend_lens=[E] * bs, n_pools=1)
    qkv, (q, k, v) = _qkv(case)
    fb = r.batch()
    parity = _parity(job, case, r, layers[0], fb, q, k, v, [5], rank, world, dev)
    r.backend.init_forward_metadata(fb)
    flops_rank = synth.algorithmic_extend_flops([E] * bs, [P] * bs, case.Hq, case.D)
    stream = torch.cuda.current_stream(dev)
    fwd = lambda: layers[0].forward(q, k, v, fb)  # noqa: E731
    # (1) the kernel timed alone, a short run (compare with the BURST tensor peak) ...
    steps = max(3, args.steps // 2)
    ms = _timed(fwd, steps, max(3, args.warmup // 2), stream, dev, world)
    tf = world * flops_rank / (ms * 1e-3) / 1e12
    # (2) ... and back to back for ~0.3 s with clocks sampled: at ~1 kW the part runs into its power cap and the SM
    # clock drops, exactly as it does for cuBLAS (MEASURED_PEAKS.json bf16_tflops_sustained) -- compare with that
    time.sleep(0.5)
    n_sus = 200
    sampler = ClockSampler(_gpu_index(local))
    if rank == 0:
        sampler.start()
    ms_sus = _timed(fwd, n_sus, 1, stream, dev, world)
    clocks = sampler.stop() if rank == 0 else None
    tf_sus = world * flops_rank / (ms_sus * 1e-3) / 1e12
    return {"workload": f"{job['tag']}: extend bs 16, prefix 2048 + new 2048, 32 Q / 8 KV heads (per rank {case.Hq} / "
                        f"{case.Hkv}), D 128, bf16; one forward_extend (KV write + attention) per step",
            "value": round(tf, 1), "unit": "TFLOP/s", "tflops_per_gpu": round(tf / world, 1),
            "ms_per_forward": round(ms, 4), "flops_per_forward_job": world * flops_rank, "forwards_timed": steps,
            "frac_of_measured_bf16_burst": round(tf / world / peaks["tf_burst"], 4),
            "frac_of_nominal_2250": round(tf / world / 2250.0, 4),
            "sustained": {"value": round(tf_sus, 1), "unit": "TFLOP/s", "forwards_timed": n_sus,
                          "ms_per_forward": round(ms_sus, 4),
                          "frac_of_measured_bf16_sustained": round(tf_sus / world / peaks["tf_sustained"], 4),
                          "clocks": clocks},
            "parity_n": parity}


# ------------------------------------------------------------------------------------------ mixed batch (C5)
def _run_c5(args, rank, world, local, dev, peaks):
    """BASELINE configs[4]: Qwen2-72B heads (64 Q / 8 KV), FP8 KV, page_size 16, one MIXED forward of 8 chunked-prefill
    rows (E = 1024, prefix <= 24 K) + 248 running decode rows with L ~ U[1 K, 32 K]; heads split over the ranks."""
    from sglang_custom_b200 import synth
    job = dict(tag="BASELINE configs[4]", model="Qwen2-72B", Hq=64, Hkv=8, D=128, fp8=True, kv_scale=0.1135, page_size=16,
               seed=1239)
    lens, ext, n_ext = synth.c5_lengths()
    case, r, layers = _build(job, rank, world, dev, lens, extend_lens=ext, n_pools=1)
    qkv, (q, k, v) = _qkv(case)
    fb = r.batch(r.itf.ForwardMode.MIXED)
    # sample: the prefill row with the shortest prefix (the CPU oracle is quadratic in it) + three decode rows
    i_ext = min(range(n_ext), key=lambda i: lens[i])
    parity = _parity(job, case, r, layers[0], fb, q, k, v, [i_ext, n_ext, n_ext + 100, len(lens) - 1], rank, world, dev)
    r.backend.init_forward_metadata(fb)
    stream = torch.cuda.current_stream(dev)
    ms = _timed(lambda: layers[0].forward(q, k, v, fb), max(5, args.steps // 2), 3, stream, dev, world)
    dec_bytes = synth.algorithmic_decode_bytes(lens[n_ext:], case.Hq, case.Hkv, case.D, 1)
    ext_flops = synth.algorithmic_extend_flops(ext[:n_ext], [l - e for l, e in zip(lens[:n_ext], ext[:n_ext])], case.Hq, case.D)
    # lower bound of one layer per GPU: the decode rows at the measured HBM peak, the prefill rows at the measured
    # sustained bf16 rate (the two overlap on the machine, so the bound is their max)
    t_hbm = dec_bytes / (peaks["hbm"] * 1e9) * 1e3
    t_tc = ext_flops / (peaks["tf_sustained"] * 1e12) * 1e3
    return {"workload": f"{job['tag']}: {job['model']}-shaped MIXED forward, {n_ext} prefill rows x 1024 new tokens (prefix "
                        f"16*U[0,1536)) + {len(lens) - n_ext} decode rows L~U[1K,32K], 64 Q / 8 KV heads (per rank "
                        f"{case.Hq} / {case.Hkv}), D 128, FP8-E4M3 KV, page_size 16; one layer = KV write + extend kernel "
                        "(prefill rows) + split-KV decode (tail rows) + merge",
            "ms_per_layer": round(ms, 4), "sum_kv_tokens": int(sum(lens)),
            "decode_part": {"bytes_per_rank": dec_bytes, "gbs_job_if_alone": round(world * dec_bytes / (ms * 1e-3) / 1e9, 1)},
            "extend_part": {"flops_per_rank": ext_flops, "tflops_job_if_alone": round(world * ext_flops / (ms * 1e-3) / 1e12, 1)},
            "roofline_ms_per_layer": {"hbm_bound_decode_rows": round(t_hbm, 4), "tensor_bound_prefill_rows": round(t_tc, 4),
                                      "frac_of_bound": round(max(t_hbm, t_tc) / ms, 4)},
            "parity_n": parity}


# ------------------------------------------------------------------------------------------ small-kernel legs
def _run_aux(dev, peaks):
    """Fused RoPE + KV write and the HiCache page transfer: algorithmic bytes / CUDA-event time (N = 1 only)."""
    from sglang_custom_b200 import _lib, kvcacheio, ops, rotary
    out = {}
    stream = torch.cuda.current_stream(dev)

    def t_us(fn, reps=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) / reps * 1e3

    # ---- RoPE + KV write: Llama-3-8B heads, a 16 K-token prefill chunk (bf16 pool) and a 256-token decode batch
    try:
        Hq, Hkv, D = 32, 8, 128
        rope = rotary.RotaryEmbedding(D, D, 8192, 500000.0, True, torch.bfloat16, device=dev)
        res = {}
        for T in (16384, 256):
            qkv = torch.randn((T, (Hq + 2 * Hkv) * D), device=dev, dtype=torch.bfloat16)
            q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=1)
            pos = torch.randint(0, 8192, (T,), device=dev)
            loc = (torch.randperm(T + 64, device=dev)[:T] + 1).to(torch.int64)
            kp = torch.zeros((T + 65, Hkv, D), dtype=torch.bfloat16, device=dev)
            vp = torch.zeros_like(kp)
            fn = lambda: ops.rope_kv_write(pos, q.view(T, Hq, D), k.view(T, Hkv, D), rope.cos_sin_cache, True,  # noqa: E731
                                           v=v.view(T, Hkv, D), loc=loc, k_pool=kp, v_pool=vp)
            us = t_us(fn)
            byt = T * (2 * (Hq + Hkv) * D * 2 + Hkv * D * 2 + 2 * Hkv * D * 2 + D * 4 + 16)
            res[f"T{T}"] = {"us": round(us, 2), "bytes": byt, "gbs": round(byt / us / 1e3, 1),
                            "frac_of_measured_hbm": round(byt / us / 1e3 / peaks["hbm"], 4)}
        out["rope_kv_write"] = {"workload": "fused RoPE (NeoX, fp32 table) + KV write, 32 Q / 8 KV heads, D 128, bf16 pool",
                                "bound": "hbm (T 16384) / launch latency (T 256)", **res}
    except Exception as e:
        out["rope_kv_write"] = {"error": f"{type(e).__name__}: {e}"[:200]}
    # ---- HiCache transfer: 4096 token items of 2 KiB (8 heads x 128 x bf16), host <-> device, K and V
    try:
        n, item = 4096, 2048
        dk = torch.zeros((n + 8, item), dtype=torch.uint8, device=dev)
        dv = torch.zeros_like(dk)
        hk = torch.zeros((n + 8, item), dtype=torch.uint8).pin_memory()
        hv = torch.zeros((n + 8, item), dtype=torch.uint8).pin_memory()
        si = torch.randperm(n, device=dev).to(torch.int64)
        di = torch.randperm(n, device=dev).to(torch.int64)
        byt = 2 * n * item
        res = {}
        for nm, (a, b, c, d) in {"h2d": (hk, dk, hv, dv), "d2h": (dk, hk, dv, hv), "d2d": (dk, dk.clone(), dv, dv.clone())}.items():
            us = t_us(lambda: kvcacheio.transfer_kv_per_layer(a, b, c, d, si, di, "kernel", 1, item, block_quota=0), reps=10)
            res[nm] = {"us": round(us, 1), "gbs": round(byt / us / 1e3, 1)}
        res["h2d"]["frac_of_pcie5_x16_64GBs"] = round(res["h2d"]["gbs"] / 64.0, 3)
        res["d2h"]["frac_of_pcie5_x16_64GBs"] = round(res["d2h"]["gbs"] / 64.0, 3)
        res["d2d"]["frac_of_measured_hbm_copy"] = round(2 * res["d2d"]["gbs"] / peaks["hbm"], 4)  # read + write
        out["transfer"] = {"workload": "transfer_kv_per_layer: 4096 items x 2 KiB, K and V, random src/dst indices",
                           "bytes_moved": byt, **res}
    except Exception as e:
        out["transfer"] = {"error": f"{type(e).__name__}: {e}"[:200]}
    return out


# ------------------------------------------------------------------------------------------ CPU legs
def _cpu_decode_sample(sample_bs: int, layers: int, reps: int):
    """The reference's own CPU decode attention (sgl-kernel/csrc/cpu/decode.cpp compiled into oracle/_ref) on a
    bounded sample of config 2: `sample_bs` requests x ctx 4096, `layers` layer calls per rep."""
    from oracle import sgl_cpu
    from oracle import torch_native as tn
    from sglang_custom_b200 import synth
    torch.set_num_threads(os.cpu_count() or 1)
    ctx, Hq, Hkv, D = 4096, 32, 8, 128
    case = synth.make_case(sample_bs, Hq, Hkv, D, [ctx] * sample_bs, seed=1236)
    kind = "reference" if sgl_cpu.available() else "port"
    logits = torch.zeros((sample_bs, Hq, sgl_cpu.NUM_KV_SPLITS, D + 1), dtype=torch.float32)
    out = torch.empty((sample_bs, Hq, D), dtype=torch.bfloat16)

    def one_layer():
        if kind == "reference":
            sgl_cpu.decode(case.q, case.k_pool, case.v_pool, case.k_new, case.v_new, case.out_cache_loc,
                           case.req_to_token, case.req_pool_indices, case.seq_lens, case.sm_scale, out=out,
                           attn_logits=logits)
        else:
            tn.set_kv_buffer(case.k_pool, case.v_pool, case.out_cache_loc, case.k_new, case.v_new, torch.bfloat16)
            tn.decode(case.q, case.k_pool, case.v_pool, case.req_to_token, case.req_pool_indices, case.seq_lens,
                      case.sm_scale, out=out)
    one_layer()  # warm-up
    bytes_layer = synth.algorithmic_decode_bytes([ctx] * sample_bs, Hq, Hkv, D, 2)
    times = []
    for _ in range(reps):
        t0 = time.perf_counter()
        for _ in range(layers):
            one_layer()
        times.append((time.perf_counter() - t0) / layers)
    return kind, bytes_layer, times


def _cpu_baseline(budget_s: float = 12.0) -> dict:
    kind, bytes_layer, t = _cpu_decode_sample(sample_bs=32, layers=1, reps=2)
    reps = max(2, min(40, int(budget_s / max(t[-1], 1e-4))))
    kind, bytes_layer, times = _cpu_decode_sample(sample_bs=32, layers=1, reps=reps)
    best = min(times)
    return {"value": round(bytes_layer / best / 1e9, 2), "unit": "GB/s", "cores": torch.get_num_threads(),
            "host_cpus": os.cpu_count(), "kind": kind,
            "sample": f"bs 32 of 256 requests (ctx 4096, 32/8 heads, D 128, bf16), one layer per call, best of {reps} "
                      f"calls; {'sgl-kernel decode_attention_cpu, num_kv_splits=8' if kind == 'reference' else 'torch_native restatement'}",
            "ms_per_layer_sample": round(best * 1e3, 3),
            "tokens_per_s_equiv": round(32 / (best * JOBS['c2']['layers']), 2)}


def _run_reference(args, rank, world):
    if rank != 0:
        return None
    layers_per_step = 2
    kind, bytes_layer, times = _cpu_decode_sample(sample_bs=16, layers=layers_per_step, reps=args.warmup + args.steps)
    times = times[args.warmup:]
    sec_layer = sum(times) / len(times)
    gbs = bytes_layer / sec_layer / 1e9
    sample = (f"bounded sample per step: bs 16 of 256 requests x ctx 4096, {layers_per_step} of 32 layer calls; "
              f"{'reference sgl-kernel CPU decode (oracle/_ref)' if kind == 'reference' else 'torch_native restatement'}")
    return {"impl": "reference", "metric": "decode-attn HBM GB/s (Llama-3-8B shapes, bs=256, ctx=4K, bf16 KV)",
            "value": round(gbs, 2), "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(sec_layer * layers_per_step * 1e3, 3), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "BASELINE configs[1] paged decode (bs 256, ctx 4096, 32/8 heads, D 128, bf16), "
                                   "CPU arm on a bounded sample", "sample": sample},
            "cpu_baseline": {"value": round(gbs, 2), "unit": "GB/s", "cores": torch.get_num_threads(), "kind": kind,
                             "sample": sample},
            "e2e": {"value": round(gbs, 2), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def _summary(line: dict) -> dict:
    """The handful of numbers a reader of the line's TAIL needs (the line is long; tools that keep only its last
    kilobyte lose `extend` / `c4_tp` otherwise).  Pure re-statement of values already on the line, whole job at N GPUs:
    c2 / c2_e2e GB/s, c3 extend TFLOP/s (+ sustained), c4 GB/s per GPU, c5 ms per layer, parity flags."""
    def get(obj, *path):
        for k in path:
            if not isinstance(obj, dict) or k not in obj:
                return None
            obj = obj[k]
        return obj
    parity = [get(line, "parity", "ok"), get(line, "parity_n", "ok"), get(line, "extend", "parity_n", "ok"),
              get(line, "extend", "parity", "ok"), get(line, "c4_tp", "parity_n", "ok"), get(line, "c4_tp", "parity", "ok"),
              get(line, "c5_mixed", "parity_n", "ok"), get(line, "c5_mixed", "parity", "ok")]
    seen = [p for p in parity if p is not None]
    return {"n_gpus": line.get("n_gpus"), "c2": line.get("value"), "c2_e2e": get(line, "e2e", "value"),
            "c2_roofline_frac": get(line, "roofline", "frac"), "c3_tflops": get(line, "extend", "value"),
            "c3_sustained": get(line, "extend", "sustained", "value"), "c4_gbs_per_gpu": get(line, "c4_tp", "gbs_per_gpu"),
            "c5_ms": get(line, "c5_mixed", "ms_per_layer"), "c5_frac_of_bound": get(line, "c5_mixed", "roofline_ms_per_layer", "frac_of_bound"),
            "parity_all_ok": bool(seen) and all(seen), "parity_checks": len(seen)}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extend", action="store_true", help="skip the configs[2] extend object")
    ap.add_argument("--no-c4", action="store_true", help="skip the configs[3] FP8 TP decode object")
    ap.add_argument("--no-c5", action="store_true", help="skip the configs[4] mixed-batch object")
    ap.add_argument("--no-aux", action="store_true", help="skip the RoPE+KV-write / HiCache-transfer objects")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--only", default=None, help="debug: run only this object (c2, c4, c5, extend, aux)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.only:
        args.no_extend, args.no_c4, args.no_c5, args.no_aux, args.no_cpu = (args.only != "extend", args.only != "c4",
                                                                           args.only != "c5", args.only != "aux", True)

    from sglang_custom_b200 import tp
    rank, world, local = tp.env_rank_world()
    if args.impl == "reference":
        line = _run_reference(args, rank, world)
        if line is not None:
            print(json.dumps(line), file=_JSON_OUT, flush=True)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the attention path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    tp.init("nccl")
    peaks = _peaks()

    def guarded(fn, *a):
        """The secondary objects must never cost the headline line; ranks fail together (same code, same shapes)."""
        try:
            return fn(*a)
        except Exception as e:
            import traceback
            traceback.print_exc()
            return {"error": f"{type(e).__name__}: {e}"[:300]}
        finally:
            torch.cuda.empty_cache()

    if args.only in (None, "c2"):
        line, _ = _run_decode_job("c2", args, rank, world, local, dev, peaks, True)
    else:
        line = {"metric": "debug run (--only)", "n_gpus": world}
    torch.cuda.empty_cache()
    if not args.no_extend:
        line["extend"] = guarded(_run_extend, args, rank, world, local, dev, peaks)
    if not args.no_c4:
        res = guarded(_run_decode_job, "c4", args, rank, world, local, dev, peaks, False)
        line["c4_tp"] = res[0] if isinstance(res, tuple) else res
    if not args.no_c5:
        line["c5_mixed"] = guarded(_run_c5, args, rank, world, local, dev, peaks)
    if rank == 0:
        if world == 1 and not args.no_aux:
            line.update(guarded(_run_aux, dev, peaks))
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = _cpu_baseline()
        line["summary"] = _summary(line)
        print(json.dumps(line), file=_JSON_OUT, flush=True)
    tp.barrier(world)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
