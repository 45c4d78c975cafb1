#!/usr/bin/env python
"""bench.py -- the attention hot path on B200, beside the reference's CPU attention.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload decode_c2]

Metric (BASELINE.json): decode-attention HBM GB/s (+ tokens/s/GPU) at Llama-3-8B shapes, bs 256, ctx 4 K,
bf16 KV, page_size 1 -- BASELINE config 2 -- and extend TFLOP/s at config 3 as a second object on the line.

A STEP is one model forward's worth of decode attention for the batch: `init_forward_metadata` once, then
`forward_decode` (KV write + split-KV decode + LSE merge) for each of the 32 layers, exactly the calls SGLang
makes per forward.  Layers rotate over two distinct pool copies; one layer reads 4.3 GB, far beyond the
126 MB L2, so no L2 flush is needed between iterations (stated in config.l2).
  value  whole-job algorithmic GB/s with every input already resident in HBM (SURVEY.md 8d bytes / device time)
  e2e    the same metric through the public API (B200AttnBackend.forward) with HOST q/k/v buffers: every layer
         call copies q,k,v from pinned host memory and reads the layer output back, inside the timed region
  roofline  the dominant kernel (decode stage 1) timed alone with CUDA events on its launch stream
  cpu_baseline  the reference's own sgl-kernel CPU decode (oracle/_ref, compiled from the reference sources) on a
         bounded sample of the same workload, on this host's cores
N > 1 (torchrun): weak scaling -- every rank owns 8 KV heads / 32 Q heads of a model with 8N KV heads,
partitioned exactly as SGLang tensor parallelism does; no collective inside the path (barrier + MAX only).
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

LAYERS = 32  # Llama-3-8B
NCU_DECODE_C2_TRAFFIC_BYTES = 4_302_531_000 + 7_630_336  # ncu --set full (profiles/r01b_ncu_decode_c2.txt), see profiles/README.md
READ_STREAM_CEILING_GBS = 7272.3  # coalesced read-only stream, scripts/ubench/hbm_read.cu (profiles/r01_ubench_hbm_read.txt)


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


# ------------------------------------------------------------------------------------------ workloads
def _decode_setup(dev, fp8=False, bs=256, ctx=4096, Hq=32, Hkv=8, D=128, seed=1236, n_pools=2):
    """BASELINE config 2 (or a sibling shape) as the backend sees it: ModelRunner stand-in with `n_pools` layer
    pools of P = bs*ctx + 1 rows, a permuted page table, int64 index tensors (eager mode)."""
    from sglang_custom_b200 import harness, synth
    case = synth.make_case_device(bs, Hq, Hkv, D, [ctx] * bs, dev, fp8=fp8, seed=seed)
    r = harness.Runner(case, dev)
    pool = r.mr.token_to_kv_pool
    for _ in range(n_pools - 1):  # further layer pools with the same content (distinct addresses)
        pool.k_buffer.append(pool.k_buffer[0].clone())
        pool.v_buffer.append(pool.v_buffer[0].clone())
    layers = [r.itf.RadixAttention(num_heads=Hq, head_dim=D, scaling=case.sm_scale, num_kv_heads=Hkv, layer_id=i)
              for i in range(n_pools)]
    return case, r, layers


def _run_decode(args, rank, world, dev, peaks):
    from sglang_custom_b200 import ops, synth, tp
    bs, ctx, Hq, Hkv, D = 256, 4096, 32, 8, 128
    case, r, layers = _decode_setup(dev, bs=bs, ctx=ctx, Hq=Hq, Hkv=Hkv, D=D, seed=1236 + rank)
    be = r.backend
    T = case.T
    qkv = torch.cat([case.q.reshape(T, -1), case.k_new.reshape(T, -1), case.v_new.reshape(T, -1)], 1)
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=1)
    fb = r.batch()
    bytes_layer = synth.algorithmic_decode_bytes([ctx] * bs, Hq, Hkv, D, 2)
    stream = torch.cuda.current_stream(dev)

    def step_resident():
        be.init_forward_metadata(fb)
        for l in range(LAYERS):
            layers[l % len(layers)].forward(q, k, v, fb)

    # ---- e2e: host q/k/v per layer in pinned memory, output read back per layer
    hq = torch.empty((T, (Hq + 2 * Hkv) * D), dtype=torch.bfloat16).pin_memory()
    hq.copy_(qkv.cpu())
    hout = torch.empty((T, Hq * D), dtype=torch.bfloat16).pin_memory()
    h_idx = torch.stack([case.req_pool_indices.cpu(), case.seq_lens.cpu(), case.out_cache_loc.cpu()]).pin_memory()
    d_qkv = [torch.empty_like(qkv) for _ in range(2)]
    d_idx = torch.empty_like(h_idx, device=dev)
    h2d = LAYERS * hq.numel() * 2 + h_idx.numel() * 8
    d2h = LAYERS * hout.numel() * 2
    # copies ride two side streams so that layer l+1's inputs and layer l-1's output cross PCIe while layer l
    # computes (what a host integration does with pinned buffers); every copy stays inside the timed region and
    # the launch stream waits for the last read-back before the closing event
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    in_ready = [torch.cuda.Event() for _ in range(2)]
    in_free = [torch.cuda.Event() for _ in range(2)]
    out_done = torch.cuda.Event()

    def step_e2e():
        d_idx.copy_(h_idx, non_blocking=True)
        fb2 = r.batch()
        fb2.req_pool_indices, fb2.seq_lens, fb2.out_cache_loc = d_idx[0], d_idx[1], d_idx[2]
        be.init_forward_metadata(fb2)
        for b in range(2):
            in_free[b].record(stream)
        for l in range(LAYERS):
            b = l & 1
            with torch.cuda.stream(s_in):
                s_in.wait_event(in_free[b])
                d_qkv[b].copy_(hq, non_blocking=True)
                in_ready[b].record(s_in)
            stream.wait_event(in_ready[b])
            dq, dk, dv = d_qkv[b].split([Hq * D, Hkv * D, Hkv * D], dim=1)
            o = layers[l % len(layers)].forward(dq, dk, dv, fb2)
            in_free[b].record(stream)
            out_done.record(stream)
            with torch.cuda.stream(s_out):
                s_out.wait_event(out_done)
                hout.copy_(o, non_blocking=True)
                o.record_stream(s_out)
        stream.wait_stream(s_out)
        stream.synchronize()

    def timed(fn, steps, warmup):
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

    sampler = ClockSampler(torch.cuda.current_device() if rank == 0 else 0)
    if rank == 0:
        sampler.start()
    ms_step = timed(step_resident, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e = timed(step_e2e, max(2, args.steps // 2), max(3, args.warmup // 2))

    # ---- CUDA-graph leg (how SGLang runs decode in production, cuda_graph_runner.py:456-593): the 32-layer step is
    # captured once through the backend's capture hook and replayed after the replay hook refreshed the plan in place
    graph_leg = None
    try:
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
            for l in range(LAYERS):
                layers[l % len(layers)].forward(q, k, v, fbg)

        def step_graph():
            be.init_forward_metadata_replay_cuda_graph(bs, g_rpi, g_sl, int(bs * ctx), None, fbg.forward_mode, None, None)
            graph.replay()
        ms_graph = timed(step_graph, args.steps, args.warmup)
        graph_leg = {"value": round(world * LAYERS * bytes_layer / (ms_graph * 1e-3) / 1e9, 1), "unit": "GB/s",
                     "ms_per_step": round(ms_graph, 4), "launches_per_step": 1 + 1,
                     "what": "replay hook (plan refresh) + one graph of 32 x (KV write + decode + merge)"}
        del graph
    except Exception as e:  # the graph leg must never cost the headline line
        graph_leg = {"error": f"{type(e).__name__}: {e}"[:200]}
    be.forward_metadata = None

    # ---- roofline leg: the dominant kernel (decode stage 1) alone, CUDA events on its stream
    pool = r.mr.token_to_kv_pool
    out = torch.empty((T, Hq, D), dtype=torch.bfloat16, device=dev)
    q3 = q.view(T, Hq, D)
    ops.decode_plan(case.seq_lens, be.n_ctas, be.plan, be.min_units_per_cta)

    def kernel_only():
        for l in range(LAYERS):
            i = l % len(layers)
            ops.decode(q3, out, pool.k_buffer[i], pool.v_buffer[i], r.mr.req_to_token_pool.req_to_token,
                       case.req_pool_indices, case.seq_lens, be.plan, be.n_ctas, be.workspace, case.sm_scale,
                       stage="partial")
    ms_kernel = timed(kernel_only, args.steps, args.warmup) / LAYERS
    kv_bytes_layer = bs * ctx * Hkv * 2 * D * 2 + bs * ctx * 4 + 2 * bs * Hq * D * 2  # KV + table + Q + O
    achieved = kv_bytes_layer / (ms_kernel * 1e-3) / 1e9

    gbs = world * LAYERS * bytes_layer / (ms_step * 1e-3) / 1e9
    gbs_e2e = world * LAYERS * bytes_layer / (ms_e2e * 1e-3) / 1e9
    line = {
        "metric": "decode-attn HBM GB/s (Llama-3-8B shapes, bs=256, ctx=4K, bf16 KV)", "value": round(gbs, 1),
        "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "BASELINE configs[1]: paged decode, bs 256, ctx 4096, 32 Q / 8 KV heads per GPU, D 128, "
                               "bf16 KV, page_size 1, permuted page table; step = init_forward_metadata + 32 layers x "
                               "(KV write + split-KV decode + LSE merge)",
                   "layers_per_step": LAYERS, "bytes_per_layer": bytes_layer,
                   "l2": "inputs larger than L2 (4.3 GB read per layer vs 126 MB L2); no flush",
                   "parallelism": f"tp{world} by KV head (weak: 8 KV heads per rank)", "n_ctas": be.n_ctas},
        "tokens_per_s_per_gpu": round(bs / (ms_step * 1e-3), 1),
        "e2e": {"value": round(gbs_e2e, 1), "unit": "GB/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": round(ms_e2e, 4), "tokens_per_s_per_gpu": round(bs / (ms_e2e * 1e-3), 1)},
        "cuda_graph": graph_leg,
        "gpu_launches": world * args.steps * (1 + 3 * LAYERS),
        "roofline": {"bound": "hbm", "kernel": "decode_kernel (split-KV stage 1)", "achieved": round(achieved, 1),
                     "peak": peaks["hbm"], "unit": "GB/s", "frac": round(achieved / peaks["hbm"], 4),
                     "peak_source": f"{peaks['source']} (MEASURED_PEAKS.json hbm_gbs)", "frac_of_nominal_8TBs": round(achieved / 8000.0, 4),
                     # `peak` is a COPY figure (read + write); a read-only stream reaches more, so frac can pass 1
                     "frac_of_read_stream_ceiling": round(achieved / READ_STREAM_CEILING_GBS, 4),
                     "read_stream_ceiling_gbs": READ_STREAM_CEILING_GBS,
                     "us_per_launch": round(ms_kernel * 1e3, 2), "bytes_per_launch": kv_bytes_layer,
                     # dram__bytes_read.sum + dram__bytes_write.sum of this kernel at this shape, one launch, from the
                     # committed `ncu --set full` capture profiles/r01b_ncu_decode_c2.txt (4.302531 GB + 7.630 MB)
                     "traffic": NCU_DECODE_C2_TRAFFIC_BYTES, "traffic_source": "profiles/r01b_ncu_decode_c2.txt"},
        "clocks": clocks,
    }
    return line, (case, r)


def _run_extend(args, rank, world, dev, peaks):
    """BASELINE config 3: bs 16, 2048 cached prefix + 2048 new tokens, 32/8 heads, D 128, one forward_extend."""
    from sglang_custom_b200 import harness, synth, tp
    bs, P, E, Hq, Hkv, D = 16, 2048, 2048, 32, 8, 128
    case = synth.make_case_device(bs, Hq, Hkv, D, [P + E] * bs, dev, extend_lens=[E] * bs, seed=1237 + rank)
    r = harness.Runner(case, dev)
    T = case.T
    qkv = torch.cat([case.q.reshape(T, -1), case.k_new.reshape(T, -1), case.v_new.reshape(T, -1)], 1)
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=1)
    fb = r.batch()
    r.backend.init_forward_metadata(fb)
    flops = synth.algorithmic_extend_flops([E] * bs, [P] * bs, Hq, D)
    stream = torch.cuda.current_stream(dev)
    warm = max(3, args.warmup // 2)
    for _ in range(warm):
        r.layer.forward(q, k, v, fb)

    def run(n):
        torch.cuda.synchronize(dev)
        tp.barrier(world)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(n):
            r.layer.forward(q, k, v, fb)
        e1.record(stream)
        torch.cuda.synchronize(dev)
        return tp.max_over_ranks(e0.elapsed_time(e1), world, dev) / n

    # (1) the kernel timed alone, a short run (compare with the BURST tensor peak) ...
    steps = max(3, args.steps // 2)
    ms = run(steps)
    tf = world * flops / (ms * 1e-3) / 1e12
    # (2) ... and back to back for ~0.3 s with clocks sampled: at ~1 kW the part runs into its power cap and the SM
    # clock drops, exactly as it does for cuBLAS (MEASURED_PEAKS.json bf16_tflops_sustained) -- compare with that
    time.sleep(0.5)
    n_sus = 200
    sampler = ClockSampler(torch.cuda.current_device() if rank == 0 else 0)
    if rank == 0:
        sampler.start()
    ms_sus = run(n_sus)
    clocks = sampler.stop() if rank == 0 else None
    tf_sus = world * flops / (ms_sus * 1e-3) / 1e12
    return {"workload": "BASELINE configs[2]: extend bs 16, prefix 2048 + new 2048, 32/8 heads, D 128, bf16; one "
                        "forward_extend (KV write + attention) per step", "value": round(tf, 1), "unit": "TFLOP/s",
            "ms_per_forward": round(ms, 4), "flops_per_forward": flops,
            "forwards_timed": steps,
            "frac_of_measured_bf16_burst": round(tf / world / peaks["tf_burst"], 4),
            "frac_of_nominal_2250": round(tf / world / 2250.0, 4),
            "sustained": {"value": round(tf_sus, 1), "unit": "TFLOP/s", "forwards_timed": n_sus,
                          "ms_per_forward": round(ms_sus, 4),
                          "frac_of_measured_bf16_sustained": round(tf_sus / world / peaks["tf_sustained"], 4),
                          "clocks": clocks}}


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
            "tokens_per_s_equiv": round(32 / (best * LAYERS), 2)}


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
            "ms_per_step": round(sec_layer * layers_per_step * 1e3, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "BASELINE configs[1] paged decode (bs 256, ctx 4096, 32/8 heads, D 128, bf16), "
                                   "CPU arm on a bounded sample", "sample": sample},
            "cpu_baseline": {"value": round(gbs, 2), "unit": "GB/s", "cores": torch.get_num_threads(), "kind": kind,
                             "sample": sample},
            "e2e": {"value": round(gbs, 2), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extend", action="store_true", help="skip the config-3 extend object")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

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
    line, keep = _run_decode(args, rank, world, dev, peaks)
    del keep
    torch.cuda.empty_cache()
    if not args.no_extend:
        line["extend"] = _run_extend(args, rank, world, dev, peaks)
    if rank == 0:
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = _cpu_baseline()
        print(json.dumps(line), file=_JSON_OUT, flush=True)
    tp.barrier(world)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
