"""B200MultiStepDraftBackend: the draft model's consecutive decode steps of EAGLE speculative decoding.

Mirrors TritonMultiStepDraftBackend (python/sglang/srt/layers/attention/triton_backend.py:735-873): one attention
backend per draft step, `init_forward_metadata` / the three CUDA-graph hooks fan out over them, and
`attn_backends[i]` is what eagle_worker.py installs as `forward_batch.attn_backend` for step i.

Where the keys of a draft row live (generate_draft_decode_kv_indices, speculative/eagle_utils.py:821-905): row
(request b, branch t) of step i reads the request's seq_lens[b] cached tokens and the i + 1 draft tokens of its own
branch, whose slots sit in the request's req_to_token row at
    start(b, t) = seq_lens[b] + t * num_steps                                              page_size == 1 or topk == 1
    start(b, t) = seq_lens[b] // page * page + t * ceil((seq_lens[b] % page + num_steps) / page) * page
                  + seq_lens[b] % page                                                     otherwise.
The reference flattens prefix + draft slots into a kv_indices list per step.  Here the decode kernel reads a page
table directly, so:
  * topk == 1: the draft tokens follow the prefix in the request's own req_to_token row -- step i is a plain decode
    over seq_lens + i + 1 tokens, nothing is copied;
  * topk > 1: ONE branch table [bs * topk, width] is built per draft forward (a row = the request's row with the
    branch's draft slots moved behind the prefix); step i reads its first seq_lens + i + 1 entries.  The reference
    writes num_steps such lists.
"""
from __future__ import annotations

from typing import Optional

import torch

from ._lib import B200AttnError
from .backend import B200AttnBackend


def draft_token_start(seq_lens: torch.Tensor, topk: int, num_steps: int, page_size: int) -> torch.Tensor:
    """[bs, topk] position in the request's req_to_token row of the first draft token of every branch
    (eagle_utils.py:868-887)."""
    L = seq_lens.to(torch.int64)[:, None]
    t = torch.arange(topk, device=seq_lens.device, dtype=torch.int64)[None, :]
    if page_size == 1 or topk == 1:
        return L + t * num_steps
    last = L % page_size
    new_pages = (last + num_steps + page_size - 1) // page_size
    return L // page_size * page_size + t * new_pages * page_size + last


def build_branch_table(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, seq_lens: torch.Tensor, topk: int,
                       num_steps: int, page_size: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Branch table [bs * topk, width]: row b * topk + t = req_to_token[req_pool_indices[b]] with the num_steps draft
    slots of branch t copied to positions [seq_lens[b], seq_lens[b] + num_steps).  Device-side torch ops, no sync."""
    bs = seq_lens.numel()
    src = req_to_token.index_select(0, req_pool_indices.to(torch.int64))            # [bs, width]
    table = src.repeat_interleave(topk, dim=0) if out is None else out[:bs * topk]
    if out is not None:
        table.view(bs, topk, -1).copy_(src[:, None, :])
    width = table.shape[1]
    start = draft_token_start(seq_lens, topk, num_steps, page_size).reshape(-1)      # [bs * topk]
    j = torch.arange(num_steps, device=table.device, dtype=torch.int64)[None, :]
    L = seq_lens.to(torch.int64).repeat_interleave(topk)[:, None]
    src_pos = start[:, None] + j
    dst_pos = L + j
    # A draft slot whose source or destination lies past the row (seq_len + num_steps > width) is dropped, the way the
    # reference masks its out-of-range loads (eagle_utils.py:889-905) -- it is never aliased onto the last column.
    valid = (src_pos < width) & (dst_pos < width)
    dst_c = dst_pos.clamp(max=width - 1)
    rows = src.repeat_interleave(topk, dim=0)
    vals = torch.where(valid, rows.gather(1, src_pos.clamp(max=width - 1)), rows.gather(1, dst_c))
    # Lanes past the row all land on column width - 1 after the clamp.  scatter_ with duplicate indices keeps an
    # unspecified one, so they are given the value of the lane that really owns that column (or the column's content
    # when none does): every duplicate writes the same number and the result is deterministic.
    j_last = (width - 1) - L
    owned = (j_last >= 0) & (j_last < num_steps)
    last_val = torch.where(owned, vals.gather(1, j_last.clamp(0, num_steps - 1)), rows[:, width - 1:width])
    vals = torch.where(dst_pos > width - 1, last_val, vals)
    table.scatter_(1, dst_c, vals)
    return table


class B200MultiStepDraftBackend:
    def __init__(self, model_runner, topk: int, speculative_num_steps: int):
        self.topk = int(topk)
        self.speculative_num_steps = int(speculative_num_steps)
        self.device = model_runner.device
        self.page_size = getattr(getattr(model_runner, "server_args", None), "page_size", None) or \
            getattr(model_runner, "page_size", 1)
        self.attn_backends = [B200AttnBackend(model_runner, skip_prefill=True) for _ in range(self.speculative_num_steps)]
        self.max_context_len = self.attn_backends[0].max_context_len
        self.num_head = self.attn_backends[0].num_head
        self.pool_len = model_runner.req_to_token_pool.req_to_token.shape[1]
        max_rows = model_runner.req_to_token_pool.size * self.topk
        for be in self.attn_backends:  # a draft step has bs * topk rows
            if max_rows > be.max_bs:
                be.max_bs = max_rows
                be._alloc_decode_buffers(max_rows)
        self.cg_table = None
        self.cg_rows = None
        self.cg_lens = None

    # ------------------------------------------------------------------ the shared walk over the steps
    def common_template(self, forward_batch, call_fn, fixed: bool):
        """Builds the rows of every step and calls call_fn(i, forward_batch) with
        forward_batch.spec_info.b200_draft_step = (rows, lens_i, table)."""
        num_seqs = forward_batch.batch_size
        rpi = forward_batch.req_pool_indices[:num_seqs]
        sl = forward_batch.seq_lens[:num_seqs]
        req_to_token = forward_batch.req_to_token_pool.req_to_token
        n = num_seqs * self.topk
        if self.topk == 1:
            table, rows = None, rpi
        elif fixed:
            table = build_branch_table(req_to_token, rpi, sl, self.topk, self.speculative_num_steps, self.page_size,
                                       out=self.cg_table)
            rows = self.cg_rows[:n]
        else:
            table = build_branch_table(req_to_token, rpi, sl, self.topk, self.speculative_num_steps, self.page_size)
            rows = torch.arange(n, device=self.device, dtype=sl.dtype)
        base = sl if self.topk == 1 else sl.repeat_interleave(self.topk)
        for i in range(self.speculative_num_steps):
            if fixed:
                lens = self.cg_lens[i][:n]
                lens.copy_(base + (i + 1))
            else:
                lens = base + (i + 1)
            forward_batch.spec_info.b200_draft_step = (rows if rows.dtype == lens.dtype else rows.to(lens.dtype), lens, table)
            call_fn(i, forward_batch)

    def init_forward_metadata(self, forward_batch):
        self.common_template(forward_batch, lambda i, fb: self.attn_backends[i].init_forward_metadata(fb), fixed=False)

    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int):
        """max_num_tokens = max_bs * topk rows per step (triton_backend.py:829-838)."""
        rows = max(max_num_tokens, max_bs * self.topk)
        if self.topk > 1:
            self.cg_table = torch.zeros((rows, self.pool_len), dtype=torch.int32, device=self.device)
        self.cg_rows = torch.arange(rows, device=self.device, dtype=torch.int64)
        self.cg_lens = [torch.ones(rows, dtype=torch.int64, device=self.device) for _ in range(self.speculative_num_steps)]
        for be in self.attn_backends:
            be.init_cuda_graph_state(rows, max_num_tokens)

    def _graph_dtype(self, forward_batch):
        dt = forward_batch.seq_lens.dtype
        if self.cg_rows is None:
            raise B200AttnError("B200MultiStepDraftBackend: CUDA-graph hook before init_cuda_graph_state")
        if self.cg_rows.dtype != dt:  # the graph runner's index dtype (int32 or int64), fixed from the first capture on
            self.cg_rows = self.cg_rows.to(dt)
            self.cg_lens = [t.to(dt) for t in self.cg_lens]

    def init_forward_metadata_capture_cuda_graph(self, forward_batch):
        self._graph_dtype(forward_batch)

        def call_fn(i, fb):
            self.attn_backends[i].init_forward_metadata_capture_cuda_graph(
                fb.batch_size, fb.batch_size * self.topk, fb.req_pool_indices, fb.seq_lens, encoder_lens=None,
                forward_mode=_DECODE, spec_info=fb.spec_info)
        self.common_template(forward_batch, call_fn, fixed=True)

    def init_forward_metadata_replay_cuda_graph(self, forward_batch, bs: int):
        def call_fn(i, fb):
            self.attn_backends[i].init_forward_metadata_replay_cuda_graph(
                bs, fb.req_pool_indices, fb.seq_lens, seq_lens_sum=-1, encoder_lens=None, forward_mode=_DECODE,
                spec_info=fb.spec_info, seq_lens_cpu=None)
        self.common_template(forward_batch, call_fn, fixed=True)


class _DecodeMode:
    """What the per-step hooks are told (ForwardMode.DECODE in the reference, triton_backend.py:846,862): only the
    predicates the backend asks are needed, so no import of the host's enum."""

    @staticmethod
    def is_decode_or_idle() -> bool:
        return True

    @staticmethod
    def is_decode() -> bool:
        return True

    @staticmethod
    def is_target_verify() -> bool:
        return False

    @staticmethod
    def is_draft_extend() -> bool:
        return False


_DECODE = _DecodeMode()
