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

Prefix sharing: the scheduler remembers, per row, which prompt columns the prompt pass left in that KV slot.  A request whose
prompt starts with `share_min` or more of the columns resident in some row (the same reference voice and system prompt, the
voice-clone serving case) copies those keys and values into its own slot (`copy_kv_prefix`) and runs the prompt pass over the
rest only - bit-identical to the full pass, like §5.7 of DESIGN.md.

Sampling draws are keyed by (seed, frame index, row): which row a request lands on and when it joins decide its draws, so
two runs of the same request agree in distribution, not token for token (exact ties between top candidates included).

Sampling parameters are per request: `fsb_frame_args.row_temperature / row_top_p / row_top_k` carry one entry per row, so
requests with different values decode side by side.  (With a session that has no such arrays - the CPU stand-in of the tests -
requests are grouped instead: the batch in flight has one (temperature, top_p, top_k), a request with other values waits until
the batch has drained, and nothing that arrived after it is admitted before it.)
"""
from __future__ import annotations

import queue
import threading
import traceback
from collections import deque
from dataclasses import dataclass, field
from typing import Any, Callable, Deque, Dict, List, Literal, Optional, Tuple, Union

import torch
from torch import Tensor

from .dual_ar import _as_bf16_float, make_semantic_logit_bias


@dataclass
class GenerateResponse:            # inference.py:52-56
    action: Literal["sample", "next", "partial"]   # "partial" is this scheduler's addition: see ContinuousBatcher.on_partial
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
    produced: int = 0              # frames kept so far: the row's next input position is prompt_len - 1 + produced
    hold: Optional[Tensor] = None  # streaming: the newest frame, held back until another one follows (the last is dropped)


@dataclass
class _Fill:
    """A request whose prompt pass is under way (chunked prefill): columns [0, at) of `cols` are in its KV slot."""
    item: GenerateRequest
    prompt: Tensor
    cols: Tensor                   # prompt[:, :-1]
    cpu: Tensor                    # the same on the host, int32 (the record for prefix sharing)
    at: int                        # -1: not started (the shared prefix is looked up when the first piece runs)
    budget: int


class ContinuousBatcher:
    """The scheduling core, free of threads: `submit()` queues a request, `step()` admits what fits, runs one chunk of
    frames for all rows in flight and returns the requests that finished.  `model` is a DualARTransformerB200 built with
    max_rows >= `max_rows` (or anything with the same `prefill` / `session` / `config` surface: the CPU tests use a stand-in)."""

    def __init__(self, model, max_rows: Optional[int] = None, chunk_frames: int = 8, use_ras: bool = True,
                 share_min: int = 32, ring_frames: int = 1024, prefill_budget: Optional[int] = None):
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
        # the frame counter runs on across chunks (it is part of the sampler's Philox counter: restarting it every chunk would
        # replay the same draws); the token record is a ring of `ring_frames` frames, and when it wraps the seed moves on
        self.ring = max(int(ring_frames), self.chunk)
        self.sess = model.session(self.rows, self.ring)
        self.sess.seed = int(torch.randint(0, 2 ** 62, (1,)).item())   # like generate(): keyed by torch's CPU generator
        self.sess.frame_counter.zero_()
        self.frame0 = 0
        self.sess.bias = make_semantic_logit_bias(model).reshape(-1).to(torch.bfloat16).contiguous()
        self.sess.restrict_scan = True
        self.sess.done.fill_(1)                     # every row idle
        # per-request (temperature, top_p, top_k): fsb_frame_args.row_* arrays, one entry per row, so requests with different
        # values share a batch; a session without them (older engine, the CPU stand-in) groups requests by their values instead
        self.per_row = hasattr(self.sess, "use_row_parameters")
        if self.per_row:
            self.sess.use_row_parameters()
            self.row_params: List[Tuple[float, float, int]] = [(_as_bf16_float(1.0), _as_bf16_float(0.9), 30)] * self.rows
            self.row_params_dirty = True
        # chunked prefill: at most `prefill_budget` prompt columns are computed between two chunks of frames while rows are
        # decoding (None: a prompt pass always runs in one piece), so a long prompt does not stall the requests in flight
        self.prefill_budget = None if prefill_budget is None else max(1, int(prefill_budget))
        self.filling: Dict[int, _Fill] = {}
        self.waiting: Deque[GenerateRequest] = deque()
        self.active: Dict[int, _Row] = {}
        self.key: Optional[Tuple[float, float, int]] = None
        self.share_min = int(share_min)             # shortest common prefix worth copying (0 = never share)
        self.resident: Dict[int, Tensor] = {}       # row -> CPU int32 [C+1, n]: columns whose prompt-pass KV sit in that slot
        self.park: Dict[int, int] = {}              # idle row -> position it idles at (behind its resident prompt columns)
        self.prompt_cols = 0                        # prompt columns admitted / of those, taken from another row or kept in place
        self.shared_cols = 0
        self.steps = 0
        self.frames_run = 0                         # row-frames computed, idle rows included
        self.frames_kept = 0                        # frames delivered to requests
        # streaming: a request with request["stream"] truthy gets its code frames as they are produced, chunk by chunk:
        # on_partial(request, codes [C, n]) is called from step(); the portions concatenated are exactly the final codes (the
        # newest frame is held back each time, because the last frame produced is not part of the result, inference.py:708)
        self.on_partial: Optional[Callable[[GenerateRequest, Tensor], None]] = None

    # ------------------------------------------------------------------ queueing
    def submit(self, item: GenerateRequest):
        self.waiting.append(item)

    def idle(self) -> bool:
        return not self.waiting and not self.active and not self.filling

    @staticmethod
    def _key(req: dict) -> Tuple[float, float, int]:
        return (_as_bf16_float(req.get("temperature", 1.0)), _as_bf16_float(req.get("top_p", 0.9)), int(req.get("top_k", 30)))

    def _admit(self) -> List[Tuple[GenerateRequest, Exception]]:
        """Moves waiting requests into free rows, in arrival order, while they share the batch's sampling parameters."""
        c = self.model.config
        C1 = c.num_codebooks + 1
        failed = []
        s = self.sess
        while self.waiting and len(self.active) + len(self.filling) < self.rows:
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
                if not 1 <= key[2] <= 64:
                    raise ValueError(f"top_k must be in [1, 64], got {key[2]}")
                if not self.per_row and (self.active or self.filling) and key != self.key:
                    break                # other sampling parameters: wait for the batch to drain, keep the order
                max_new = int(req.get("max_new_tokens", 0) or 0)
                budget = min(max_new, c.max_seq_len - T) if max_new else c.max_seq_len - T   # inference.py:267-274
            except Exception as e:  # noqa: BLE001 - reported to the requester like the reference's worker does
                self.waiting.popleft()
                failed.append((item, e))
                continue
            self.waiting.popleft()
            row = next(r for r in range(self.rows) if r not in self.active and r not in self.filling)
            self.key = key
            if self.per_row:
                self.row_params[row] = key
                self.row_params_dirty = True
            cols = prompt[:, :-1]                             # KV slot `row` (slot = row) <- the prompt but its last column
            cpu = cols.detach().to(device="cpu", dtype=torch.int32)
            # until its prompt is in, the row idles with the batch and writes K/V where it stands: behind the prompt columns
            # (what the slot held before survives in front of that spot only)
            self.park[row] = T - 1
            if row in self.resident:
                self.resident[row] = self.resident[row][:, : T - 1]
            self.filling[row] = _Fill(item=item, prompt=prompt, cols=cols, cpu=cpu, at=-1, budget=budget)
            if self.prefill_budget is None or T - 1 + self.chunk > c.max_seq_len:
                self._advance_fill(row, None)                 # in one piece (always, when there is no room to idle behind it)
        self._advance_fills()
        return failed

    def _advance_fills(self):
        """Spends the prompt-column budget of this chunk boundary on the requests being filled, oldest first; while nothing
        is decoding there is nobody to stall, so the oldest prompt is computed in one piece."""
        spent = 0
        for row in list(self.filling):
            if self.prefill_budget is None or not self.active:
                self._advance_fill(row, None)                 # (the request it lets in is decoding from here on: budget next)
                continue
            if spent >= self.prefill_budget:
                break
            spent = self.prefill_budget - self._advance_fill(row, self.prefill_budget - spent)

    def _advance_fill(self, row: int, left: Optional[int]) -> Optional[int]:
        f = self.filling[row]
        s = self.sess
        total = int(f.cols.shape[1])
        if f.at < 0:                                          # first piece: start behind what some slot already holds
            f.at = self._share_prefix(row, f.cpu)
        n = total - f.at if left is None else min(total - f.at, left)
        if n > 0 or total == 0:
            self.model.prefill(f.cols[:, f.at: f.at + n], f.at, row)
        f.at += n
        self.resident[row] = f.cpu[:, : f.at]                 # what is in the slot so far can already be shared
        if f.at >= total:                                     # the prompt is in: the row joins the batch
            s.cur_tokens[row].copy_(f.prompt[:, -1].to(device=s.cur_tokens.device, dtype=torch.int32))
            s.ras_window[row].zero_()
            self.park.pop(row, None)
            self.active[row] = _Row(item=f.item, prompt_len=total + 1, budget=f.budget)
            del self.filling[row]
        return None if left is None else left - n

    def _share_prefix(self, row: int, cpu: Tensor) -> int:
        """Copies into KV slot `row` the longest prefix of the prompt columns `cpu` that is already resident in some slot;
        returns its length (the prompt pass starts there)."""
        n = int(cpu.shape[1])
        best_row, best = -1, 0
        if self.share_min > 0 and hasattr(self.model, "copy_kv_prefix"):
            for r, have in self.resident.items():
                m = min(n, int(have.shape[1]))
                if m <= best or m < self.share_min or not torch.equal(have[:, :self.share_min], cpu[:, :self.share_min]):
                    continue
                same = (have[:, :m] == cpu[:, :m]).all(0)
                L = m if bool(same.all()) else int((~same).nonzero()[0])
                if L > best or (L == best and r == row):
                    best_row, best = r, L
        if best < self.share_min:
            best_row, best = -1, 0
        if best_row >= 0 and best_row != row:
            self.model.copy_kv_prefix(best_row, row, best)
        self.prompt_cols += n
        self.shared_cols += best
        return best

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
        if self.frame0 + n > self.ring:
            s.frame_counter.zero_()
            self.frame0 = 0
            s.seed = (int(s.seed) + 1) & (2 ** 62 - 1)
        f0 = self.frame0
        self.frame0 += n
        T_, P_, K_ = self.key
        # the host knows every row's position: rows in flight continue where their kept frames end, idle rows (they step along
        # with the batch, their output is dropped) idle behind their resident prompt columns - one small upload per chunk
        pos = [self.active[r].prompt_len - 1 + self.active[r].produced if r in self.active else self.park.get(r, 0)
               for r in range(self.rows)]
        s.pos.copy_(torch.tensor(pos, dtype=torch.int32))
        s.done.copy_(torch.tensor([0 if r in self.active else 1 for r in range(self.rows)], dtype=torch.int32))
        if self.per_row and self.row_params_dirty:
            s.row_temperature.copy_(torch.tensor([p[0] for p in self.row_params], dtype=torch.float32))
            s.row_top_p.copy_(torch.tensor([p[1] for p in self.row_params], dtype=torch.float32))
            s.row_top_k.copy_(torch.tensor([p[2] for p in self.row_params], dtype=torch.int32))
            self.row_params_dirty = False
        s.run(n, T_, P_, K_, self.use_ras, pos_bound=max(pos))
        self.steps += 1
        self.frames_run += n * self.rows
        toks = s.out_tokens[f0: f0 + n]                          # [n][rows][C+1] on the device
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
            r.produced += take
            if self.on_partial is not None and r.item.request.get("stream"):
                new = r.frames[-1] if r.hold is None else torch.cat([r.hold, r.frames[-1]], 0)
                r.hold = new[-1:]
                if new.shape[0] > 1:
                    self.on_partial(r.item, new[:-1, 1:].t().contiguous())
            if ended or r.budget <= 0:
                seq = torch.cat(r.frames, 0)                     # [n_total, C+1]
                codes = seq[:-1, 1:].t().contiguous()            # y[1:, prompt_length:-1]
                self.frames_kept += int(seq.shape[0])
                out.append((r.item, codes))
                del self.active[row]
                # the idle row keeps stepping with the batch and writes KV where it stands: let it stand behind its prompt
                # columns, which stay valid for requests that start the same way; too close to the end of the cache, give
                # the columns up and idle at the start
                end = r.prompt_len - 1 + r.produced
                spot = min(end, c.max_seq_len - self.chunk)
                if spot >= r.prompt_len - 1:
                    self.park[row] = spot
                else:
                    self.park.pop(row, None)
                    self.resident.pop(row, None)
            # rows still in flight continue from the state the kernel left (cur_tokens, RAS window)
        if not self.active and not self.filling:
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


def _deliver_partial(item: GenerateRequest, codes: Tensor):
    item.response_queue.put(WrappedGenerateResponse(status="success", response=GenerateResponse(action="partial", codes=codes)))


def launch_thread_safe_queue(checkpoint_path=None, device="cuda", precision=torch.bfloat16, compile: bool = False, *,
                             model=None, max_rows: int = 8, chunk_frames: int = 8, max_length: Optional[int] = None,
                             prefill_budget: Optional[int] = None) -> queue.Queue:
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
            batcher = ContinuousBatcher(m, max_rows=min(max_rows, m.max_rows), chunk_frames=chunk_frames,
                                        prefill_budget=prefill_budget)
            batcher.on_partial = _deliver_partial
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
                for r in list(batcher.active.values()) + list(batcher.filling.values()):
                    _deliver(r.item, e)
                batcher.active.clear()
                batcher.filling.clear()
                batcher.sess.done.fill_(1)
                batcher.key = None

    threading.Thread(target=worker, daemon=True).start()
    init_event.wait()
    if init_error:
        raise init_error[0]
    return input_queue
