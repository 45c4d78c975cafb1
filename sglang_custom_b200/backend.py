"""B200AttnBackend: the SGLang AttentionBackend whose hot path is libb200attn.so.

Mirrors the reference interface one to one (python/sglang/srt/layers/attention/...):
  AttentionBackend ABC                      base_attn_backend.py:14-115
  constructor / attributes read             triton_backend.py:41-122
  init_forward_metadata                     triton_backend.py:160-336   (here: one device-side plan kernel)
  forward_decode / forward_extend           triton_backend.py:632-732, torch_native_backend.py:182-267
  CUDA-graph hooks                          triton_backend.py:338-630, cuda_graph_runner.py:251-256,550-558,686-695
KV write semantics: mem_cache/memory_pool.py:420-458 (the backend owns that call, as
triton_backend.py:647-650,706-709 do), done by the b200attn KV-write kernel.

There is no fallback path: a missing library or an unsupported feature raises.
"""
from __future__ import annotations

import os

from abc import ABC, abstractmethod
from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib, ops
from ._lib import B200AttnError


class AttentionBackend(ABC):
    """Same surface as the reference's base class (base_attn_backend.py:14-115)."""

    @abstractmethod
    def init_forward_metadata(self, forward_batch):
        raise NotImplementedError()

    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int):
        raise NotImplementedError()

    def init_forward_metadata_capture_cuda_graph(self, bs, num_tokens, req_pool_indices, seq_lens, encoder_lens,
                                                 forward_mode, spec_info):
        raise NotImplementedError()

    def init_forward_metadata_replay_cuda_graph(self, bs, req_pool_indices, seq_lens, seq_lens_sum, encoder_lens,
                                                forward_mode, spec_info, seq_lens_cpu):
        raise NotImplementedError()

    def get_cuda_graph_seq_len_fill_value(self):
        raise NotImplementedError()

    def forward(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
        if forward_batch.forward_mode.is_decode():
            return self.forward_decode(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache, **kwargs)
        return self.forward_extend(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache, **kwargs)

    def forward_decode(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True):
        raise NotImplementedError()

    def forward_extend(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True):
        raise NotImplementedError()

    def support_triton(self):
        return True


@dataclass
class ForwardMetadata:
    is_decode: bool
    bs: int
    req_pool_indices: Optional[torch.Tensor] = None  # set only by the CUDA-graph hooks (fixed buffers)
    seq_lens: Optional[torch.Tensor] = None
    extend_start_loc: Optional[torch.Tensor] = None
    max_extend_len: int = 0
    # MIXED batches: the running (decode) requests are appended after the prefill requests with
    # extend_len = 1 (schedule_batch.py:1310-1338); that tail is served by the decode kernel
    n_decode_tail: int = 0
    # number of q rows that belong to the prefill requests (sum of their extend lens, from the host list): the tail
    # rows start there even when q carries padded rows behind them (DP attention / num_token_non_padded)
    n_extend_tokens: int = -1
    # KV splits of the prefill rows' extend kernel (chosen on the host from extend_*_lens_cpu, see _choose_kv_splits)
    kv_splits: int = 1
    # prefill rows re-ordered longest first: permuted copies of the per-request arrays (None: batch order)
    ext_order: Optional[dict] = None
    # TARGET_VERIFY (speculative decoding): every request contributes num_draft_tokens query rows over its
    # seq_lens cached tokens; built on the device, no host sync (triton_backend.py:227-264)
    extend_seq_lens: Optional[torch.Tensor] = None
    extend_prefix_lens: Optional[torch.Tensor] = None
    total_lens: Optional[torch.Tensor] = None
    custom_mask: Optional[torch.Tensor] = None
    mask_indptr: Optional[torch.Tensor] = None
    prefix_seq_lens: Optional[torch.Tensor] = None  # = forward_batch.seq_lens of a verify batch (its own dtype)
    # DRAFT_EXTEND (speculative decoding): accept_length[i] new tokens at the end of seq_lens[i], plain causal extend
    draft_extend: bool = False
    # a draft decode step (B200MultiStepDraftBackend): the rows read their keys through this table instead of the
    # pool's req_to_token (topk > 1: a per-branch copy with the branch's draft tokens behind the shared prefix)
    req_to_token: Optional[torch.Tensor] = None


class B200AttnBackend(AttentionBackend):
    def __init__(self, model_runner, skip_prefill: bool = False):
        super().__init__()
        lib = _lib.load()  # raises if libb200attn.so is missing: there is no other implementation
        self.device = model_runner.device
        if not str(self.device).startswith("cuda"):
            raise B200AttnError("B200AttnBackend needs a CUDA device (sm_100a); there is no CPU path")
        info = _lib.device_info()
        if info["cc_major"] != 10:
            raise B200AttnError(f"libb200attn.so is built for sm_100a only, found compute capability "
                                f"{info['cc_major']}.{info['cc_minor']}")
        self.skip_prefill = skip_prefill
        # sliding-window models (triton_backend.py:124): window layers read the last min(L, W + 1) tokens
        sw = getattr(model_runner, "sliding_window_size", None)
        self.sliding_window_size = int(sw) if sw is not None and sw > 0 else None
        sargs = getattr(model_runner, "server_args", None)
        self.num_draft_tokens = getattr(sargs, "speculative_num_draft_tokens", None)
        self.speculative_num_steps = getattr(sargs, "speculative_num_steps", None)
        if getattr(model_runner.model_config, "is_encoder_decoder", False):
            raise NotImplementedError("b200 attention backend: encoder-decoder (cross attention) is not built")

        self.req_to_token = model_runner.req_to_token_pool.req_to_token
        self.max_bs = model_runner.req_to_token_pool.size
        self.max_context_len = model_runner.model_config.context_len
        tp = getattr(model_runner, "tp_size", 1)
        self.num_head = model_runner.model_config.num_attention_heads // tp
        self.num_kv_head = model_runner.model_config.get_num_kv_heads(tp)
        pool = model_runner.token_to_kv_pool
        self.v_head_dim = pool.get_value_buffer(getattr(pool, "start_layer", 0)).shape[-1]
        self.kv_dtype = pool.dtype
        self.page_size = getattr(model_runner, "page_size", 1)

        # persistent decode grid: one CTA per SM (the decode kernel keeps ~190 KiB of gathers in flight per CTA)
        self.sm_count = info["sm_count"]
        self.n_ctas = self.sm_count
        self.min_units_per_cta = 2  # scripts/sweep_min_units.py: 12.5 vs 13.2 us at bs 1 / ctx 4K, no effect elsewhere
        self._alloc_decode_buffers(self.max_bs)
        self.forward_metadata: Optional[ForwardMetadata] = None
        self.scale_mode = _lib.SCALE_DIV
        self.writeback_scaled_kv = False
        # prefill rows: 0 = choose KV splits per batch (_choose_kv_splits), 1 = never, n = always n
        self.extend_kv_splits = int(os.environ.get("B200ATTN_EXTEND_KV_SPLITS", "0"))
        self._split_ws = None
        self.sort_extend_requests = os.environ.get("B200ATTN_SORT_EXTEND", "1") != "0"  # prefill rows: longest request first
        self.verify_via_decode = True  # TARGET_VERIFY: cached-prefix part on the split-KV decode kernel
        # ... as long as it takes at most this many passes of 16 (draft token, q head) rows: a pass streams the prefix once
        # (0.65 ms at bs 256 / ctx 4096, profiles/r01_time_masked_verify.txt), the masked extend kernel alone takes 6.9 ms
        self.verify_max_passes = 8
        # MIXED: decode tail on a side stream, concurrent with the prefill rows (False: serial; True: tail launched
        # first; 2: prefill launched first).  C5 per-rank shape: 0.823 -> 0.731 ms per layer, bit-identical.
        self.overlap_mixed = 2
        self._side_stream = None
        self.route_decode_tail = True  # MIXED batches: extend_len == 1 tail rows go to the decode kernel
        # (layer_id, sequence number) of the last fused RoPE + KV write (rotary.forward_fused_kv), also stamped on the
        # batch object it was done for: the only case in which forward_* may be called with k/v and
        # save_kv_cache=False (see _require_kv_in_pool)
        self._kv_written = None
        self._kv_written_seq = 0
        self.check_loc = os.environ.get("B200ATTN_CHECK_LOC", "0") not in ("", "0")

    # ------------------------------------------------------------------ buffers
    def _alloc_decode_buffers(self, max_bs: int) -> None:
        self.plan = torch.zeros(ops.decode_plan_bytes(max_bs, self.n_ctas) // 4, dtype=torch.int32, device=self.device)
        ws = ops.decode_workspace_bytes(self.n_ctas, self.num_head, self.v_head_dim)
        self.workspace = torch.empty(ws // 4, dtype=torch.float32, device=self.device)
        # second plan for sliding-window layers (the reference keeps window_kv_indptr / window_kv_indices beside
        # kv_indptr / kv_indices, triton_backend.py:124-135)
        self.window_plan = torch.zeros_like(self.plan) if self.sliding_window_size is not None else None
        # TARGET_VERIFY: up to 16 (draft token, q head) rows of a kv head share one pass over the cached prefix
        self.verify_workspace = None
        if self.num_draft_tokens:
            vws = ops.decode_workspace_bytes(self.n_ctas, self.num_kv_head * 16, self.v_head_dim)
            self.verify_workspace = torch.empty(vws // 4, dtype=torch.float32, device=self.device)

    # ------------------------------------------------------------------ metadata
    def init_forward_metadata(self, forward_batch):
        """One tiny device-side kernel (decode) or nothing at all (extend): no flattened kv_indices
        copy, no host sync (the reference syncs on max_extend_len, triton_backend.py:320)."""
        bs = forward_batch.batch_size
        mode = forward_batch.forward_mode
        spec_info = getattr(forward_batch, "spec_info", None)
        if spec_info is not None:
            if hasattr(mode, "is_target_verify") and mode.is_target_verify():
                self.forward_metadata = self._verify_metadata(len(forward_batch.req_pool_indices),
                                                              forward_batch.seq_lens, spec_info, None)
                return
            if hasattr(mode, "is_draft_extend") and mode.is_draft_extend():
                self.forward_metadata = self._draft_extend_metadata(len(forward_batch.req_pool_indices),
                                                                    forward_batch.seq_lens, spec_info, None)
                return
            if mode.is_decode_or_idle():
                if bs == 0 or (hasattr(mode, "is_idle") and mode.is_idle()):
                    # an idle batch that carries an idle EagleDraftInput / EagleVerifyInput (EAGLE with DP attention):
                    # the reference reads spec_info.kv_indptr and launches nothing for it
                    self.forward_metadata = ForwardMetadata(True, 0)
                    return
                # a draft decode step: rows, lengths and table come from B200MultiStepDraftBackend
                # (the reference hands kv_indptr / kv_indices over the same way, triton_backend.py:203-205)
                self.forward_metadata = self._draft_step_metadata(spec_info)
                return
            # EXTEND with a spec_info attached (the draft model's own prefill, eagle_worker.py forward_draft_extend):
            # attention-wise a plain extend batch -- the Triton backend takes its default branch for it too
        if mode.is_decode_or_idle():
            if bs > 0:
                self._plan_decode(forward_batch.seq_lens)
            self.forward_metadata = ForwardMetadata(True, bs)
            return
        ext_cpu = getattr(forward_batch, "extend_seq_lens_cpu", None)
        n_tail = 0
        t_ext = -1
        if ext_cpu is not None:
            ext_cpu = list(ext_cpu)
            if self.route_decode_tail and self.sliding_window_size is None:
                while n_tail < len(ext_cpu) and ext_cpu[len(ext_cpu) - 1 - n_tail] == 1:
                    n_tail += 1
            head = ext_cpu[:len(ext_cpu) - n_tail]
            max_extend = max(head) if head else 0
            t_ext = int(sum(head))
        else:  # the reference's own (syncing) way, triton_backend.py:320
            max_extend = int(forward_batch.extend_seq_lens.max().item()) if bs else 0
        start = getattr(forward_batch, "extend_start_loc", None)
        if start is None and bs:
            ext = forward_batch.extend_seq_lens
            start = (torch.cumsum(ext, 0) - ext).to(torch.int32)
        if n_tail:
            # a query with extend_len 1 at position L-1 sees all L keys: exactly a decode row
            self._plan_decode(forward_batch.seq_lens[bs - n_tail:])
        kv_splits, ext_order = 1, None
        pre_cpu = getattr(forward_batch, "extend_prefix_lens_cpu", None)
        if ext_cpu is not None and pre_cpu is not None:
            n_ext = len(ext_cpu) - n_tail
            head_ext, head_pre = ext_cpu[:n_ext], list(pre_cpu)[:n_ext]
            # Longest request first: the extend grid is a queue in block-index order (request slowest), and a long row
            # that starts last is the tail of the launch.  The kernels address q / out rows through extend_start_loc,
            # so permuting the per-request arrays is all it takes; results are unchanged bit for bit.
            if self.sort_extend_requests and n_ext > 1:
                order = sorted(range(n_ext), key=lambda i: -(head_pre[i] + head_ext[i]))
                if order != list(range(n_ext)):
                    idx = torch.tensor(order, dtype=torch.int64, device=forward_batch.seq_lens.device)
                    ext_order = dict(
                        req_pool_indices=forward_batch.req_pool_indices[:n_ext][idx],
                        seq_lens=forward_batch.seq_lens[:n_ext][idx],
                        extend_seq_lens=forward_batch.extend_seq_lens[:n_ext][idx],
                        extend_prefix_lens=forward_batch.extend_prefix_lens[:n_ext][idx],
                        extend_start_loc=start[:n_ext][idx])
                    head_ext, head_pre = [head_ext[i] for i in order], [head_pre[i] for i in order]
            if self.extend_kv_splits != 1:
                kv_splits = self._choose_kv_splits(head_ext, head_pre)
        self.forward_metadata = ForwardMetadata(False, bs, extend_start_loc=start, max_extend_len=max_extend,
                                                n_decode_tail=n_tail, n_extend_tokens=t_ext, kv_splits=kv_splits,
                                                ext_order=ext_order)

    def _choose_kv_splits(self, ext, pre) -> int:
        """How many shares the K/V range of every prefill item is cut into (b200attn_extend_split).  The tcgen05 extend
        kernel has one CTA per (request, head pair, 128 q rows); a few long rows -- chunked prefill over a long prefix,
        few heads per rank -- leave SMs idle for the length of the longest item.  When the grid is small (< 4 CTAs per
        SM) the launch is simulated on the host for a few split factors: CTAs in block-index order onto the first free
        SM, cost of a CTA = its 128-key tiles + 3 (launch, Q load, fill, drain, epilogue; one tile = 1.84 us measured),
        plus the merge pass.  Host lists only, no sync; 0 = auto (default), n > 1 forces n."""
        if self.extend_kv_splits > 1:
            return self.extend_kv_splits
        if not ext or max(ext) < 128:
            return 1
        G = self.num_head // max(1, self.num_kv_head)
        hp = self.num_head // 2 if G % 2 == 0 else self.num_head
        rows_per_item = 128 if G % 2 == 0 else 256
        sms = self.n_ctas
        groups, items, longest, T = [], 0, 0, 0  # groups: tiles of the hp identical items of (request, q tile), launch order
        for e, p_ in zip(ext, pre):
            if e <= 0:
                continue
            nq = -(-e // rows_per_item)
            items += hp * nq
            if items >= 4 * sms:
                return 1  # the GPU is filled several times over: nothing to gain
            per_req = [(p_ + min((z + 1) * rows_per_item, e) + 127) // 128 for z in range(nq - 1, -1, -1)]  # heavy first
            groups.append(per_req)
            longest = max(longest, per_req[0])
            T += e
        if items == 0 or longest < 16:
            return 1
        import heapq
        c = 3.0
        merge_per_split = T * self.num_head * 256 / 5e6 / 1.84  # bytes of one partial / (5 TB/s), in tile units

        def makespan(S: int) -> float:
            free = [0.0] * sms
            for per_req in groups:
                for _ in range(S):
                    for n in per_req:
                        cost = n / S + c
                        for _ in range(hp):
                            heapq.heapreplace(free, free[0] + cost)
            return max(free) + (merge_per_split * S + 2.0 if S > 1 else 0.0)

        best, best_t = 1, makespan(1)
        t1 = best_t
        for S in (2, 3, 4, 6, 8):
            if longest / S < 8 or items * S > 6 * sms:
                break
            t = makespan(S)
            if t < 0.92 * t1 and t < best_t:
                best, best_t = S, t
        return best

    def _plan_decode(self, seq_lens) -> None:
        ops.decode_plan(seq_lens, self.n_ctas, self.plan, self.min_units_per_cta)
        if self.window_plan is not None:
            ops.decode_plan(seq_lens, self.n_ctas, self.window_plan, self.min_units_per_cta,
                            sliding_window_size=self.sliding_window_size)

    def _draft_step_metadata(self, spec_info) -> ForwardMetadata:
        step = getattr(spec_info, "b200_draft_step", None)
        if step is None:
            raise B200AttnError("decode batch with a spec_info but no draft-step rows: draft decoding runs through "
                                "B200MultiStepDraftBackend (draft_backend.py), as TritonMultiStepDraftBackend does")
        rows, lens, table = step
        if rows.numel() > 0:
            self._plan_decode(lens)
        return ForwardMetadata(True, rows.numel(), req_pool_indices=rows, seq_lens=lens, req_to_token=table)

    def _draft_extend_metadata(self, bs: int, seq_lens, spec_info, bufs) -> ForwardMetadata:
        """DRAFT_EXTEND (eagle_utils.py:107-133): after a verify step the draft model extends every request by its
        accept_length[i] (already + 1) accepted tokens; forward_batch.seq_lens INCLUDES them.  Query t of request i
        sits at position seq_lens[i] - accept_length[i] + t and sees the keys up to itself -- the semantics of the
        reference's default (FlashInfer) backend: qo_indptr = cumsum(accept_length), paged_kernel_lens = seq_lens,
        causal (flashinfer_backend.py prefill updater; eagle_utils.py:135-164).  The reference's Triton backend hands
        the SAME key range to its extend kernel as an unmasked prefix and then adds the new tokens a second time
        under the causal mask (triton_backend.py:265-281 + extend_attention.py stage 1 / stage 2), i.e. it counts the
        accepted tokens twice; that quirk is not reproduced.
        No host sync: the grid is sized by speculative_num_steps + 1 (the reference's own bound under CUDA graphs,
        triton_backend.py:476-499) instead of max(accept_length).item()."""
        acc = spec_info.accept_length[:bs].to(torch.int32)
        sl = seq_lens[:bs]
        steps = self.speculative_num_steps
        max_ext = int(steps) + 1 if steps else (int(acc.max().item()) if bs else 0)
        if bufs is None:
            ext = acc.contiguous()
            prefix = (sl.to(torch.int32) - acc).contiguous()
            start = (torch.cumsum(acc, 0, dtype=torch.int32) - acc).contiguous()
        else:
            ext, prefix, start = bufs[0][:bs], bufs[1][:bs], bufs[2][:bs]
            ext.copy_(acc)
            prefix.copy_(sl.to(torch.int32) - acc)
            start.copy_(torch.cumsum(acc, 0, dtype=torch.int32) - acc)
        return ForwardMetadata(False, bs, extend_start_loc=start, max_extend_len=max_ext, extend_seq_lens=ext,
                               extend_prefix_lens=prefix, total_lens=sl, draft_extend=True)

    def _verify_metadata(self, bs: int, seq_lens, spec_info, bufs) -> ForwardMetadata:
        """TARGET_VERIFY (triton_backend.py:227-264): qo_indptr = arange * num_draft_tokens, the prefix is seq_lens
        (the draft tokens' slots sit behind it in req_to_token: eagle_utils.py:268-276), mask block i is
        [num_draft_tokens, seq_lens[i] + num_draft_tokens].  `bufs` = fixed CUDA-graph buffers to fill in place."""
        nd = self.num_draft_tokens or getattr(spec_info, "draft_token_num", None)
        if not nd:
            raise B200AttnError("TARGET_VERIFY needs server_args.speculative_num_draft_tokens")
        nd = int(nd)
        sl = seq_lens[:bs]
        prefix = sl.to(torch.int32)
        if bufs is None:
            ext = torch.full((bs,), nd, dtype=torch.int32, device=self.device)
            start = torch.arange(0, bs * nd, nd, dtype=torch.int32, device=self.device)
            total = sl + nd
            mask_indptr = torch.zeros(bs + 1, dtype=torch.int64, device=self.device)
            mask_indptr[1:] = torch.cumsum(nd * (sl.to(torch.int64) + nd), 0)
            mask = spec_info.custom_mask
        else:
            ext, pre, start, total, mask_indptr, mask = (b[:n] for b, n in zip(
                bufs, (bs, bs, bs, bs, bs + 1, bufs[5].numel())))
            ext.fill_(nd)
            pre.copy_(prefix)
            prefix = pre
            start.copy_(torch.arange(0, bs * nd, nd, dtype=torch.int32, device=self.device))
            total.copy_(sl + nd)
            mask_indptr[1:] = torch.cumsum(nd * (sl.to(torch.int64) + nd), 0)
            n = spec_info.custom_mask.numel()
            mask[:n] = spec_info.custom_mask.view(torch.uint8) if spec_info.custom_mask.dtype == torch.bool \
                else spec_info.custom_mask
        if bs > 0:
            self._plan_decode(sl)  # the cached-prefix part runs on the split-KV decode kernel
        return ForwardMetadata(False, bs, extend_start_loc=start, max_extend_len=nd, extend_seq_lens=ext,
                               extend_prefix_lens=prefix, total_lens=total, custom_mask=mask, mask_indptr=mask_indptr,
                               prefix_seq_lens=sl)

    # ------------------------------------------------------------------ CUDA graphs
    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int):
        """Fixed-address plan/workspace buffers for capture + replay (triton_backend.py:338-388)."""
        if max_bs > self.max_bs:
            self.max_bs = max_bs
            self._alloc_decode_buffers(max_bs)
        if self.speculative_num_steps and not self.skip_prefill:
            i32 = dict(dtype=torch.int32, device=self.device)
            self.cg_draft_extend = (torch.zeros(max_bs, **i32), torch.zeros(max_bs, **i32), torch.zeros(max_bs, **i32))
        if self.num_draft_tokens and not self.skip_prefill:
            # TARGET_VERIFY graphs: fixed-address lens / mask buffers (triton_backend.py:366-371)
            i32 = dict(dtype=torch.int32, device=self.device)
            self.cg_verify = (torch.zeros(max_bs, **i32), torch.zeros(max_bs, **i32), torch.zeros(max_bs, **i32),
                              torch.zeros(max_bs, **i32), torch.zeros(max_bs + 1, dtype=torch.int64, device=self.device),
                              torch.zeros(max_num_tokens * self.max_context_len, dtype=torch.uint8, device=self.device))

    def init_forward_metadata_capture_cuda_graph(self, bs, num_tokens, req_pool_indices, seq_lens, encoder_lens,
                                                 forward_mode, spec_info):
        if encoder_lens is not None:
            raise NotImplementedError("b200 attention backend: encoder_lens is not supported")
        if forward_mode.is_decode_or_idle() and spec_info is None:
            self._plan_decode(seq_lens[:bs])
            self.forward_metadata = ForwardMetadata(True, bs, req_pool_indices=req_pool_indices[:bs],
                                                    seq_lens=seq_lens[:bs])
            return
        if hasattr(forward_mode, "is_target_verify") and forward_mode.is_target_verify() and spec_info is not None:
            if getattr(self, "cg_verify", None) is None:
                raise B200AttnError("TARGET_VERIFY capture before init_cuda_graph_state / without num_draft_tokens")
            md = self._verify_metadata(bs, seq_lens, spec_info, self.cg_verify)
            md.req_pool_indices = req_pool_indices[:bs]
            self.forward_metadata = md
            return
        if hasattr(forward_mode, "is_draft_extend") and forward_mode.is_draft_extend() and spec_info is not None:
            # triton_backend.py:476-499: captured with speculative_num_steps + 1 rows per request; the replay hook
            # refills the same buffers from accept_length (:609-625)
            if getattr(self, "cg_draft_extend", None) is None:
                raise B200AttnError("DRAFT_EXTEND capture before init_cuda_graph_state / without speculative_num_steps")
            md = self._draft_extend_metadata(bs, seq_lens, spec_info, self.cg_draft_extend)
            md.req_pool_indices = req_pool_indices[:bs]
            md.total_lens = seq_lens[:bs]
            self.forward_metadata = md
            return
        if forward_mode.is_decode_or_idle() and spec_info is not None:
            if bs == 0 or (hasattr(forward_mode, "is_idle") and forward_mode.is_idle()):
                self.forward_metadata = ForwardMetadata(True, 0)  # idle batch with an idle spec input: nothing to launch
                return
            self.forward_metadata = self._draft_step_metadata(spec_info)  # fixed buffers owned by the multi-step backend
            return
        raise NotImplementedError(f"b200 attention backend: {forward_mode!r} does not run under CUDA graphs "
                                  "(the reference raises for it too, triton_backend.py:503-506)")

    def init_forward_metadata_replay_cuda_graph(self, bs, req_pool_indices, seq_lens, seq_lens_sum, encoder_lens,
                                                forward_mode, spec_info, seq_lens_cpu):
        """Runs right before graph.replay(): refresh the plan in place from the (device) seq_lens.
        Padded rows arrive as seq_len = 1 / req_pool_idx = 0 (cuda_graph_runner.py:686-695)."""
        if forward_mode.is_decode_or_idle() and spec_info is None:
            self._plan_decode(seq_lens[:bs])
            return
        if hasattr(forward_mode, "is_target_verify") and forward_mode.is_target_verify() and spec_info is not None:
            self._verify_metadata(bs, seq_lens, spec_info, self.cg_verify)  # refills the captured buffers in place
            return
        if hasattr(forward_mode, "is_draft_extend") and forward_mode.is_draft_extend() and spec_info is not None:
            self._draft_extend_metadata(bs, seq_lens, spec_info, self.cg_draft_extend)
            return
        if forward_mode.is_decode_or_idle() and spec_info is not None:
            if bs == 0 or (hasattr(forward_mode, "is_idle") and forward_mode.is_idle()):
                return
            self._draft_step_metadata(spec_info)  # re-plans over the step's (in place refreshed) lengths
            return
        raise NotImplementedError(f"b200 attention backend: {forward_mode!r} does not run under CUDA graphs "
                                  "(the reference raises for it too, triton_backend.py:627-630)")

    def get_cuda_graph_seq_len_fill_value(self):
        return 1

    # ------------------------------------------------------------------ helpers
    def _check_layer(self, layer) -> None:
        if getattr(layer, "is_cross_attention", False):
            raise NotImplementedError("b200 attention backend: cross attention is not built")
        if layer.qk_head_dim != layer.v_head_dim:
            raise NotImplementedError("b200 attention backend: qk_head_dim != v_head_dim (MLA) is not built")

    def _layer_window(self, layer) -> int:
        """-1 for a global layer, else the layer's window -- which must be the one the window plan was built for
        (the reference builds window_kv_indices from the backend's size and masks with the layer's,
        triton_backend.py:187-201,658-665; a model has one window size)."""
        sw = getattr(layer, "sliding_window_size", None)
        if sw is None or sw <= -1:
            return -1
        if self.sliding_window_size is None or int(sw) != self.sliding_window_size:
            raise B200AttnError(f"layer.sliding_window_size={sw} but the backend was built for "
                                f"model_runner.sliding_window_size={self.sliding_window_size}")
        return int(sw)

    @staticmethod
    def _kv_scales(layer):
        ks = getattr(layer, "k_scale_float", None)
        vs = getattr(layer, "v_scale_float", None)
        return (1.0 if ks is None else float(ks)), (1.0 if vs is None else float(vs))

    def note_kv_written(self, layer, forward_batch) -> None:
        """Called by rotary.forward_fused_kv: this layer's new K/V of this batch are already in the pool."""
        # The mark lives ON the batch object: a (layer, id(batch)) pair kept here could be satisfied by a later batch
        # that CPython happened to give the same id after this one was collected.
        self._kv_written_seq += 1
        self._kv_written = (layer.layer_id, self._kv_written_seq)
        forward_batch.b200_kv_written = (id(self),) + self._kv_written

    def _require_kv_in_pool(self, layer, forward_batch, k) -> None:
        """save_kv_cache=False contract.  Every key -- cached prefix and new tokens alike -- is read from the paged
        pool (torch_native_backend.py:196-199 does the same), so with save_kv_cache=False the new tokens must already
        be there: either the caller passes k=None (the reference's own convention for "KV is in the cache",
        radix_attention.py:91-99), or rotary.forward_fused_kv wrote them for exactly this layer and batch.  The Triton
        backend would attend over the passed k/v instead (extend_attention_fwd reads K_Extend/V_Extend); silently
        reading stale pool rows in that situation is the one thing this must never do."""
        if k is None:
            return
        mark = getattr(forward_batch, "b200_kv_written", None)
        if mark is not None and self._kv_written is not None and mark == (id(self),) + self._kv_written \
                and mark[1] == layer.layer_id:
            return
        raise B200AttnError(
            "save_kv_cache=False with k/v given, but the new tokens of this layer/batch were not written by "
            "rotary.forward_fused_kv: the b200 backend reads every key from the paged pool. Pass save_kv_cache=True, "
            "or k=None if the pool already holds them.")

    def _save_kv(self, layer, forward_batch, k, v) -> None:
        pool = forward_batch.token_to_kv_pool
        lid = layer.layer_id - getattr(pool, "start_layer", 0)
        k3 = k.view(-1, layer.tp_k_head_num, layer.qk_head_dim)
        v3 = v.view(-1, layer.tp_v_head_num, layer.v_head_dim)
        if self.check_loc:  # B200ATTN_CHECK_LOC=1: debug validation, synchronises
            loc = forward_batch.out_cache_loc
            dup, oob = ops.check_loc(loc.to(torch.int64).contiguous(), self.page_size, pool.k_buffer[lid].shape[0])
            if dup or oob:
                raise B200AttnError(f"out_cache_loc: {dup} duplicate pair(s) outside the dummy page [0, {self.page_size}) "
                                    f"and {oob} slot(s) out of range -- the reference's last-writer-wins "
                                    "(memory_pool.py:457-458) is not reproduced for write races")
        # scales are only meaningful when the pool is narrower than k/v (memory_pool.py:436-442);
        # like FlashInfer (flashinfer_backend.py:470-476) the layer's calibrated scales are applied
        ks = getattr(layer, "k_scale", None) if getattr(layer, "k_scale", None) is not None \
            else getattr(layer, "k_scale_float", None)
        vs = getattr(layer, "v_scale", None) if getattr(layer, "v_scale", None) is not None \
            else getattr(layer, "v_scale_float", None)
        ops.set_kv_buffer(pool.k_buffer[lid], pool.v_buffer[lid], forward_batch.out_cache_loc, k3, v3, pool.dtype,
                          ks, vs, self.scale_mode, self.writeback_scaled_kv)

    # ------------------------------------------------------------------ forward
    def forward_decode(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True):
        self._check_layer(layer)
        q = q.reshape(-1, layer.tp_q_head_num * layer.qk_head_dim)
        o = torch.empty_like(q)
        if save_kv_cache:
            self._save_kv(layer, forward_batch, k, v)
        else:
            self._require_kv_in_pool(layer, forward_batch, k)
        if q.shape[0] == 0:
            return o
        md = self.forward_metadata
        rpi = md.req_pool_indices if md.req_pool_indices is not None else forward_batch.req_pool_indices
        sl = md.seq_lens if md.seq_lens is not None else forward_batch.seq_lens
        pool = forward_batch.token_to_kv_pool
        lid = layer.layer_id - getattr(pool, "start_layer", 0)
        # the layer_transfer_counter wait (HiCache) lives in get_key_buffer (memory_pool.py:401-402)
        pool.get_key_buffer(layer.layer_id)
        k_scale, v_scale = self._kv_scales(layer)
        # sliding-window layers: same kernel, the plan that covers the last min(L, W + 1) tokens
        plan = self.window_plan if self._layer_window(layer) > -1 else self.plan
        ops.decode(q.view(-1, layer.tp_q_head_num, layer.qk_head_dim),
                   o.view(-1, layer.tp_q_head_num, layer.v_head_dim), pool.k_buffer[lid], pool.v_buffer[lid],
                   md.req_to_token if md.req_to_token is not None else forward_batch.req_to_token_pool.req_to_token,
                   rpi, sl, plan, self.n_ctas, self.workspace,
                   layer.scaling, getattr(layer, "logit_cap", 0.0) or 0.0, k_scale, v_scale, kv_dtype=pool.dtype,
                   # paged pools: 16-token units are TMA boxes (a window plan starts requests off their page boundary)
                   # a draft step's branch table is a copy with the draft tokens spliced in: row gathers
                   page_size=self.page_size if (plan is self.plan and md.req_to_token is None) else 1)
        return o

    def forward_extend(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True):
        self._check_layer(layer)
        o = torch.empty_like(q)
        if save_kv_cache:
            self._save_kv(layer, forward_batch, k, v)
        else:
            self._require_kv_in_pool(layer, forward_batch, k)
        md = self.forward_metadata
        if md.bs == 0 or q.shape[0] == 0:
            return o
        from .interfaces import AttentionType  # value-compatible with the reference enum
        attn_type = getattr(layer, "attn_type", None)
        causal = not (attn_type is not None and getattr(attn_type, "value", attn_type) == AttentionType.ENCODER_ONLY.value)
        pool = forward_batch.token_to_kv_pool
        lid = layer.layer_id - getattr(pool, "start_layer", 0)
        pool.get_key_buffer(layer.layer_id)
        k_scale, v_scale = self._kv_scales(layer)
        q3 = q.view(-1, layer.tp_q_head_num, layer.qk_head_dim)
        o3 = o.view(-1, layer.tp_q_head_num, layer.v_head_dim)
        table = forward_batch.req_to_token_pool.req_to_token
        cap = getattr(layer, "logit_cap", 0.0) or 0.0
        window = self._layer_window(layer)
        if md.custom_mask is not None:  # TARGET_VERIFY
            if window > -1:  # the reference has no window in spec inference either (triton_backend.py:228)
                raise NotImplementedError("b200 attention backend: sliding window with speculative decoding")
            rpi = md.req_pool_indices if md.req_pool_indices is not None else forward_batch.req_pool_indices
            rpi = rpi[:md.bs]
            nd, Hq = md.max_extend_len, layer.tp_q_head_num
            per_pass = 16 // (Hq // layer.tp_k_head_num) if Hq // layer.tp_k_head_num <= 16 else 0
            if self.verify_via_decode and causal and per_pass > 0 and -(-nd // per_pass) <= self.verify_max_passes \
                    and self.verify_workspace is not None and o3.is_contiguous():
                # HBM-bound like decode: every draft token sees the whole cached prefix, so that part is ONE
                # split-KV pass per 16 (token, head) rows over the page table (b200attn_decode_multiq); the few
                # draft-token keys take the tree mask in the register-MMA kernel; an LSE merge joins the two.
                q4, o4 = q3.unflatten(0, (md.bs, nd)), o3.unflatten(0, (md.bs, nd))
                lse_p = torch.empty((md.bs, nd, Hq), dtype=torch.float32, device=q.device)
                for t0 in range(0, nd, per_pass):
                    ops.decode_multiq(q4[:, t0:t0 + per_pass], o4[:, t0:t0 + per_pass], pool.k_buffer[lid],
                                      pool.v_buffer[lid], table, rpi, md.prefix_seq_lens, self.plan, self.n_ctas,
                                      self.verify_workspace, layer.scaling, cap, k_scale, v_scale, kv_dtype=pool.dtype,
                                      lse_out=lse_p[:, t0:t0 + per_pass], page_size=self.page_size)
                o_d = torch.empty_like(o3)
                lse_d = torch.empty((md.bs * nd, Hq), dtype=torch.float32, device=q.device)
                ops.extend(q3, o_d, pool.k_buffer[lid], pool.v_buffer[lid], table, rpi, md.total_lens,
                           md.extend_seq_lens, md.extend_prefix_lens, md.extend_start_loc, nd, layer.scaling, cap,
                           k_scale, v_scale, causal=False, kv_dtype=pool.dtype, lse_out=lse_d,
                           sliding_window_size=_lib.WINDOW_NEW_TOKENS_ONLY, custom_mask=md.custom_mask,
                           mask_indptr=md.mask_indptr, page_size=self.page_size)
                ops.merge_state(o3, lse_p.view(md.bs * nd, Hq), o_d, lse_d, out=o3)
                return o
            ops.extend(q3, o3, pool.k_buffer[lid], pool.v_buffer[lid], table, rpi, md.total_lens,
                       md.extend_seq_lens, md.extend_prefix_lens, md.extend_start_loc, md.max_extend_len,
                       layer.scaling, cap, k_scale, v_scale, causal=False, kv_dtype=pool.dtype,
                       custom_mask=md.custom_mask, mask_indptr=md.mask_indptr, page_size=self.page_size)
            return o
        if md.draft_extend:
            if window > -1:
                raise NotImplementedError("b200 attention backend: sliding window with speculative decoding")
            rpi = md.req_pool_indices if md.req_pool_indices is not None else forward_batch.req_pool_indices
            ops.extend(q3, o3, pool.k_buffer[lid], pool.v_buffer[lid], table, rpi[:md.bs], md.total_lens,
                       md.extend_seq_lens, md.extend_prefix_lens, md.extend_start_loc, md.max_extend_len,
                       layer.scaling, cap, k_scale, v_scale, causal=causal, kv_dtype=pool.dtype,
                       page_size=self.page_size)
            return o
        n_tail = md.n_decode_tail  # a 1-token row sees all L keys with or without the causal mask
        n_ext = md.bs - n_tail
        t_ext = md.n_extend_tokens if md.n_extend_tokens >= 0 else q3.shape[0] - n_tail
        t_end = t_ext + n_tail  # q may carry padded rows behind the tail
        side = None
        if n_tail > 0 and n_ext > 0 and t_ext > 0 and self.overlap_mixed \
                and not torch.cuda.is_current_stream_capturing():
            # MIXED: the prefill rows are tensor-bound and fill the SMs unevenly (a few hundred CTAs), the decode tail
            # is HBM-bound: run the tail's split-KV kernel on a side stream so the two share the machine.  Both read
            # the pool after this layer's KV write (ordered by the event) and write disjoint rows of `o`.
            cur = torch.cuda.current_stream(q.device)
            if self._side_stream is None:
                self._side_stream = torch.cuda.Stream(q.device)
            side = self._side_stream
            side.wait_stream(cur)

            def tail_on_side():
                with torch.cuda.stream(side):
                    ops.decode(q3[t_ext:t_end], o3[t_ext:t_end], pool.k_buffer[lid], pool.v_buffer[lid], table,
                               forward_batch.req_pool_indices[n_ext:], forward_batch.seq_lens[n_ext:], self.plan,
                               self.n_ctas, self.workspace, layer.scaling, cap, k_scale, v_scale, kv_dtype=pool.dtype,
                               page_size=self.page_size)
            if self.overlap_mixed != 2:
                tail_on_side()
        if n_ext > 0 and t_ext > 0:
            S, ws = 1, None
            if md.kv_splits > 1 and window < 0 and not cap and layer.qk_head_dim == 128 and o3.is_contiguous():
                S = md.kv_splits
                need = ops.extend_split_workspace_bytes(t_ext, layer.tp_q_head_num, 128, S)
                if self._split_ws is None or self._split_ws.numel() < need:
                    self._split_ws = torch.empty(need, dtype=torch.uint8, device=q.device)
                ws = self._split_ws
            eo = md.ext_order
            ops.extend(q3[:t_ext], o3[:t_ext], pool.k_buffer[lid], pool.v_buffer[lid], table,
                       eo["req_pool_indices"] if eo else forward_batch.req_pool_indices[:n_ext],
                       eo["seq_lens"] if eo else forward_batch.seq_lens[:n_ext],
                       eo["extend_seq_lens"] if eo else forward_batch.extend_seq_lens[:n_ext],
                       eo["extend_prefix_lens"] if eo else forward_batch.extend_prefix_lens[:n_ext],
                       eo["extend_start_loc"] if eo else md.extend_start_loc[:n_ext], md.max_extend_len, layer.scaling,
                       cap, k_scale, v_scale, causal=causal, kv_dtype=pool.dtype, sliding_window_size=window,
                       page_size=self.page_size, kv_splits=S, workspace=ws)
        if side is not None:
            if self.overlap_mixed == 2:  # prefill rows launched first, the tail fills the SMs they leave
                tail_on_side()
            torch.cuda.current_stream(q.device).wait_stream(side)
            return o
        if n_tail > 0:
            # (window layers: a 1-token row's prefix range is its last min(L - 1, W + 1) cached tokens plus itself --
            #  the decode window plan keeps min(L, W + 1); the two differ by one key, so window layers keep the tail
            #  on the extend kernel: route_decode_tail is ignored for them in init_forward_metadata)
            ops.decode(q3[t_ext:t_end], o3[t_ext:t_end], pool.k_buffer[lid], pool.v_buffer[lid], table,
                       forward_batch.req_pool_indices[n_ext:], forward_batch.seq_lens[n_ext:], self.plan, self.n_ctas,
                       self.workspace, layer.scaling, cap, k_scale, v_scale, kv_dtype=pool.dtype,
                       page_size=self.page_size)
        return o
