"""Continuous batching behind the reference's queue protocol.

The reference serialises requests: `launch_thread_safe_queue` (fish_speech/models/text2semantic/inference.py:748-799) starts
ONE worker thread that pops a `GenerateRequest`, runs it to completion and puts `WrappedGenerateResponse` objects on the
request's own response queue.  On a B200 a decode step costs the full weight stream for any number of rows up to ~64
(DESIGN.md §5.5), so the worker here keeps up to `max_rows` requests in flight on one model: every request owns one KV slot
(= one row of the many-row step), new requests are admitted between chunks of frames (their prompt pass fills the slot),
rows retire at `<|im_end|>` or at their `max_new_tokens`, and a freed row is handed to the next waiting request at once.

Same names and message flow as the reference: `GenerateRequest(request, response_queue)` in, `WrappedGenerateResponse(status,
response)` out, with `GenerateResponse(action="sample", codes=...)` followed by `GenerateResponse(action="next")` per request
(inference.py:709-733), or `status="error"` with the exception (inference.py:789-793).  The request dict carries the
arguments of `generate()` (inference.py:243-253): `prompt` (encoded [C+1, T] tokens - text encoding is the tokenizer's job,
outside this path), `max_new_tokens`, `temperature`, `top_p`, `top_k`.

Rows of one step share the sampling parameters (one `fsb_frame_args`), so requests are grouped: the batch in flight has one
(temperature, top_p, top_k); a request with other values waits until the batch has drained, and nothing that arrived after it
is admitted before it (no starvation).
"""
from __future__ import annotations

import queue
import threading
import traceback
from collections import deque
from dataclasses import dataclass, field
from typing import Any, Deque, Dict, List, Literal, Optional, Tuple, Union

import torch
from torch import Tensor

from .dual_ar import _as_bf16_float, make_semantic_logit_bias


@dataclass
class GenerateResponse:            # inference.py:52-56
    action: Literal["sample", "next"]
    codes: Optional[Tensor] = None
    text: Optional[str] = None


@dataclass
class WrappedGenerateResponse:     # inference.py:736-739
    status: Literal["success", "error"]
    response: Optional[Union[GenerateResponse, Exception]] = None


@dataclass
class GenerateRequest:             # inference.py:742-745
    request: dict
    response_queue: queue.Queue


@dataclass
class _Row:
    item: GenerateRequest
    prompt_len: int
    budget: int                    # frames this request may still produce (generate(): max_new_tokens, clipped to the cache)
    frames: List[Tensor] = field(default_factory=list)   # [n, C+1] int32 pieces, on the device


class ContinuousBatcher:
    """The scheduling core, free of threads: `submit()` queues a request, `step()` admits what fits, runs one chunk of
    frames for all rows in flight and returns the requests that finished.  `model` is a DualARTransformerB200 built with
    max_rows >= `max_rows` (or anything with the same `prefill` / `session` / `config` surface: the CPU tests use a stand-in)."""

    def __init__(self, model, max_rows: Optional[int] = None, chunk_frames: int = 8, use_ras: bool = True):
        self.model = model
        self.rows = int(max_rows if max_rows is not None else model.max_rows)
        if self.rows < 1 or self.rows > model.max_rows:
            raise ValueError(f"max_rows {self.rows} must be in [1, {model.max_rows}] (the model's row capacity)")
        self.chunk = int(chunk_frames)
        if self.chunk < 1:
            raise ValueError("chunk_frames must be >= 1")
        self.use_ras = use_ras
        c = model.config
        if not getattr(model, "_cache_setup_done", False) or model.max_batch_size < self.rows:
            model.setup_caches(self.rows, c.max_seq_len)
        self.sess = model.session(self.rows, self.chunk)
        self.sess.bias = make_semantic_logit_bias(model).reshape(-1).to(torch.bfloat16).contiguous()
        self.sess.restrict_scan = True
        self.sess.done.fill_(1)                     # every row idle
        self.waiting: Deque[GenerateRequest] = deque()
        self.active: Dict[int, _Row] = {}
        self.key: Optional[Tuple[float, float, int]] = None
        self.steps = 0
        self.frames_run = 0                         # row-frames computed, idle rows included
        self.frames_kept = 0                        # frames delivered to requests

    # ------------------------------------------------------------------ queueing
    def submit(self, item: GenerateRequest):
        self.waiting.append(item)

    def idle(self) -> bool:
        return not self.waiting and not self.active

    @staticmethod
    def _key(req: dict) -> Tuple[float, float, int]:
        return (_as_bf16_float(req.get("temperature", 1.0)), _as_bf16_float(req.get("top_p", 0.9)), int(req.get("top_k", 30)))

    def _admit(self) -> List[Tuple[GenerateRequest, Exception]]:
        """Moves waiting requests into free rows, in arrival order, while they share the batch's sampling parameters."""
        c = self.model.config
        C1 = c.num_codebooks + 1
        failed = []
        s = self.sess
        while self.waiting and len(self.active) < self.rows:
            item = self.waiting[0]
            try:
                req = item.request
                prompt = req["prompt"]
                if prompt.dim() != 2 or prompt.shape[0] != C1:
                    raise ValueError(f"prompt must be [{C1}, T] tokens, got {tuple(prompt.shape)}")
                T = int(prompt.shape[1])
                if T < 1:
                    raise ValueError("empty prompt")
                if T >= c.max_seq_len:   # inference.py:262-265
                    raise ValueError(f"Input sequence length {T} exceeds max_seq_len {c.max_seq_len}")
                key = self._key(req)
                if self.active and key != self.key:
                    break                # other sampling parameters: wait for the batch to drain, keep the order
                max_new = int(req.get("max_new_tokens", 0) or 0)
                budget = min(max_new, c.max_seq_len - T) if max_new else c.max_seq_len - T   # inference.py:267-274
            except Exception as e:  # noqa: BLE001 - reported to the requester like the reference's worker does
                self.waiting.popleft()
                failed.append((item, e))
                continue
            self.waiting.popleft()
            row = next(r for r in range(self.rows) if r not in self.active)
            self.key = key
            self.model.prefill(prompt[:, :-1], 0, row)       # fills KV slot `row` (slot = row)
            s.cur_tokens[row].copy_(prompt[:, -1].to(device=s.cur_tokens.device, dtype=torch.int32))
            s.pos[row] = T - 1
            s.ras_window[row].zero_()
            s.done[row] = 0
            self.active[row] = _Row(item=item, prompt_len=T, budget=budget)
        return failed

    # ------------------------------------------------------------------ one chunk
    def step(self) -> List[Tuple[GenerateRequest, Union[Tensor, Exception]]]:
        """Admit, run up to `chunk_frames` frames for every row in flight, retire.  Returns [(request, codes | exception)];
        codes are [C, n] like `y[1:, prompt_length:-1]` (inference.py:708): the frames before the last one produced."""
        out: List[Tuple[GenerateRequest, Union[Tensor, Exception]]] = list(self._admit())
        if not self.active:
            return out
        c = self.model.config
        s = self.sess
        n = min(self.chunk, max(r.budget for r in self.active.values()))
        s.frame_counter.zero_()
        T_, P_, K_ = self.key
        hi = max(r.prompt_len - 1 + sum(int(p.shape[0]) for p in r.frames) for r in self.active.values())
        s.run(n, T_, P_, K_, self.use_ras, pos_bound=hi)
        self.steps += 1
        self.frames_run += n * self.rows
        toks = s.out_tokens[:n]                                  # [n][rows][C+1] on the device
        main = toks[:, :, 0].to("cpu")                           # one small synchronising copy per chunk
        if hasattr(s, "check_device"):
            s.check_device()
        for row in sorted(self.active):
            r = self.active[row]
            take = min(n, r.budget)
            hit = (main[:take, row] == c.im_end_id).nonzero()
            ended = hit.numel() > 0
            if ended:
                take = int(hit[0]) + 1                           # the <|im_end|> frame is the last one produced
            r.frames.append(toks[:take, row].clone())
            r.budget -= take
            if ended or r.budget <= 0:
                seq = torch.cat(r.frames, 0)                     # [n_total, C+1]
                codes = seq[:-1, 1:].t().contiguous()            # y[1:, prompt_length:-1]
                self.frames_kept += int(seq.shape[0])
                out.append((r.item, codes))
                del self.active[row]
                s.done[row] = 1
                s.pos[row] = 0                                   # an idle row keeps stepping: park it at the cache's start
            # rows still in flight simply continue from the state the kernel left (cur_tokens, pos, RAS window)
        for row in range(self.rows):                             # idle rows advanced too: keep them parked
            if row not in self.active:
                s.pos[row] = 0
        if not self.active:
            self.key = None
        return out

    def run_until_idle(self) -> List[Tuple[GenerateRequest, Union[Tensor, Exception]]]:
        done = []
        while not self.idle():
            done.extend(self.step())
        return done


def _deliver(item: GenerateRequest, result):
    if isinstance(result, Exception):
        item.response_queue.put(WrappedGenerateResponse(status="error", response=result))
        return
    item.response_queue.put(WrappedGenerateResponse(status="success", response=GenerateResponse(action="sample", codes=result)))
    item.response_queue.put(WrappedGenerateResponse(status="success", response=GenerateResponse(action="next")))


def launch_thread_safe_queue(checkpoint_path=None, device="cuda", precision=torch.bfloat16, compile: bool = False, *,
                             model=None, max_rows: int = 8, chunk_frames: int = 8, max_length: Optional[int] = None) -> queue.Queue:
    """inference.py:748-799 with continuous batching: returns the input queue; put `GenerateRequest`s on it, `None` to stop
    the worker.  Either `checkpoint_path` (loaded with checkpoint.init_model, rows = max_rows) or a ready `model`."""
    input_queue: queue.Queue = queue.Queue()
    init_event = threading.Event()
    init_error: List[BaseException] = []

    def worker():
        try:
            m = model
            if m is None:
                from .checkpoint import init_model
                m, _ = init_model(checkpoint_path, device, precision, compile, max_rows=max_rows, max_length=max_length)
            dev = getattr(m, "device", None)
            if dev is not None and torch.device(dev).type == "cuda":
                torch.cuda.set_device(dev)
            batcher = ContinuousBatcher(m, max_rows=min(max_rows, m.max_rows), chunk_frames=chunk_frames)
        except BaseException as e:  # noqa: BLE001
            init_error.append(e)
            init_event.set()
            return
        init_event.set()
        stop = False
        while not (stop and batcher.idle()):
            # block for work only when nothing is in flight; otherwise take what has arrived and keep stepping
            try:
                item = input_queue.get(block=batcher.idle() and not stop, timeout=None)
                while True:
                    if item is None:
                        stop = True
                    else:
                        batcher.submit(item)
                    item = input_queue.get_nowait()
            except queue.Empty:
                pass
            try:
                for it, res in batcher.step():
                    _deliver(it, res)
            except Exception as e:  # noqa: BLE001 - a failed step fails every request in flight, the worker lives on
                traceback.print_exc()
                for r in list(batcher.active.values()):
                    _deliver(r.item, e)
                batcher.active.clear()
                batcher.sess.done.fill_(1)
                batcher.key = None

    threading.Thread(target=worker, daemon=True).start()
    init_event.wait()
    if init_error:
        raise init_error[0]
    return input_queue
