"""Host logic of the continuous-batching worker (fish_speech_b200/scheduler.py) against a CPU stand-in for the engine: the
stand-in produces, for every KV slot, a token stream that depends only on the prompt in that slot and the position, so the
expected output of a request is known whatever it was scheduled with.  (The GPU tests run the same scheduler on the engine.)"""
import queue
import types

import pytest
import torch

from fish_speech_b200.scheduler import (ContinuousBatcher, GenerateRequest, GenerateResponse, WrappedGenerateResponse,
                                        launch_thread_safe_queue)

C = 3            # codebooks
SEM0, CB, IM_END, S = 100, 16, 7, 64


class FakeSession:
    def __init__(self, model, batch, max_frames):
        self.model, self.batch, self.max_frames = model, batch, max_frames
        self.cur_tokens = torch.zeros((batch, C + 1), dtype=torch.int32)
        self.pos = torch.zeros((batch,), dtype=torch.int32)
        self.ras_window = torch.zeros((batch, 10), dtype=torch.int32)
        self.out_tokens = torch.zeros((max_frames, batch, C + 1), dtype=torch.int32)
        self.frame_counter = torch.zeros((1,), dtype=torch.int32)
        self.done = torch.zeros((batch,), dtype=torch.int32)
        self.bias = None
        self.restrict_scan = True
        self.calls = []

    def run(self, n, temperature, top_p, top_k, use_ras, pos_bound=None):
        m = self.model
        for r in range(self.batch):           # a decode frame writes K/V at its position: model it on the stand-in's slots
            if r in m.kv:
                p0 = int(self.pos[r])
                kv = m.kv[r]
                for p in range(p0, p0 + n):
                    if p < len(kv):
                        kv[p] = -1            # overwritten: no longer the prompt pass's
                    else:
                        kv.append(-1)
        assert int(self.frame_counter) + n <= self.max_frames, "the token record would overflow"
        assert int(self.pos.max()) + n <= m.max_seq_len, "a row would step past the cache"
        self.calls.append((n, temperature, top_p, top_k, [int(d) for d in self.done]))
        for _ in range(n):
            f = int(self.frame_counter)         # the record is indexed by the running frame counter (frame_end_kernel)
            for r in range(self.batch):
                p = int(self.pos[r])
                tok = m.stream(r, p + 1)
                self.cur_tokens[r] = torch.tensor(tok, dtype=torch.int32)
                if tok[0] == IM_END:
                    self.done[r] = 1
                if f < self.max_frames:
                    self.out_tokens[f, r] = self.cur_tokens[r]
                self.pos[r] = p + 1
            self.frame_counter += 1


class RowParamSession(FakeSession):
    """A session with per-request sampling parameters (fsb_frame_args.row_*): records the arrays of every step."""

    def use_row_parameters(self):
        self.row_temperature = torch.ones((self.batch,), dtype=torch.float32)
        self.row_top_p = torch.full((self.batch,), 0.9, dtype=torch.float32)
        self.row_top_k = torch.full((self.batch,), 30, dtype=torch.int32)
        self.row_calls = []

    def run(self, n, temperature, top_p, top_k, use_ras, pos_bound=None):
        self.row_calls.append((self.row_temperature.clone(), self.row_top_p.clone(), self.row_top_k.clone(),
                               [int(d) for d in self.done]))
        super().run(n, temperature, top_p, top_k, use_ras, pos_bound)


class FakeModel:
    """Token at position p of the request whose prompt sums to h: main = SEM0 + (h + p) % CB, codes = (h + p + i) % CB;
    <|im_end|> at position eos_at(h) = prompt_len + 3 + h % 9."""

    def __init__(self, max_rows=3, row_params=False):
        self.max_rows = max_rows
        self.row_params = row_params
        self.config = types.SimpleNamespace(num_codebooks=C, max_seq_len=S, im_end_id=IM_END, vocab_size=SEM0 + CB,
                                            semantic_begin_id=SEM0, semantic_end_id=SEM0 + CB - 1, codebook_size=CB)
        self.device = torch.device("cpu")
        self.tokenizer = types.SimpleNamespace(get_token_id=lambda _n: IM_END)
        self._cache_setup_done = False
        self.max_batch_size = 0
        self.slots = {}
        self.kv = {}
        self.computed = self.copied = 0
        self.admitted = []

    def setup_caches(self, rows, max_seq_len, dtype=torch.bfloat16):
        self.max_batch_size, self.max_seq_len, self._cache_setup_done = rows, max_seq_len, True

    def session(self, batch, max_frames):
        self._s = (RowParamSession if self.row_params else FakeSession)(self, batch, max_frames)
        return self._s

    def prefill(self, cols, pos0, slot):
        kv = self.kv.setdefault(slot, [])
        assert pos0 <= len(kv), "a prompt pass from pos0 needs the keys before it"
        del kv[pos0:]
        kv.extend(int(x) for x in cols.sum(0))          # one number per column stands for its K/V
        self.computed += cols.shape[-1]
        T = len(kv) + 1
        h = sum(kv) % 1000
        self.slots[slot] = (h, T)
        self.admitted.append(h)

    def copy_kv_prefix(self, src, dst, n):
        self.kv[dst] = list(self.kv[src][:n])
        self.copied += n

    def stream(self, slot, p):
        if slot not in self.slots:
            return [SEM0] * (C + 1)
        h, T = self.slots[slot]
        h2 = h + int(self._s.cur_tokens[slot].sum()) * 0      # the stream depends on the slot's prompt and position only
        if p >= eos_at(h, T):
            return [IM_END] + [0] * C
        return [SEM0 + (h2 + p) % CB] + [(h2 + p + i) % CB for i in range(C)]


def eos_at(h, T):
    return T + 3 + h % 9


def prompt(T, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 50, (C + 1, T), generator=g)


def expected(p, max_new):
    T = p.shape[1]
    h = int(p[:, :-1].sum()) % 1000
    budget = min(max_new, S - T) if max_new else S - T
    frames = []
    for k in range(budget):
        pos = T + k
        if pos >= eos_at(h, T):
            frames.append([IM_END] + [0] * C)
            break
        frames.append([SEM0 + (h + pos) % CB] + [(h + pos + i) % CB for i in range(C)])
    seq = torch.tensor(frames, dtype=torch.int32)
    return seq[:-1, 1:].t().contiguous()


def req(p, max_new=0, **kw):
    return GenerateRequest(request=dict(prompt=p, max_new_tokens=max_new, **kw), response_queue=queue.Queue())


def drain(q):
    out = []
    while not q.empty():
        out.append(q.get_nowait())
    return out


def test_every_request_gets_its_own_stream_whatever_it_shares_the_batch_with():
    m = FakeModel(max_rows=3)
    b = ContinuousBatcher(m, chunk_frames=4)
    items = [req(prompt(5 + i, i), max_new=(0 if i % 3 else 6)) for i in range(8)]
    for it in items:
        b.submit(it)
    done = b.run_until_idle()
    assert len(done) == len(items) and b.idle()
    by_id = {id(it): res for it, res in done}
    for it in items:
        want = expected(it.request["prompt"], it.request["max_new_tokens"])
        got = by_id[id(it)]
        assert got.dtype == torch.int32 and torch.equal(got, want), (got, want)
    assert max(len([d for d in c[4] if d == 0]) for c in m._s.calls) == 3       # the rows were used together
    assert b.frames_kept <= b.frames_run


def test_sampling_parameters_group_batches_without_starvation():
    m = FakeModel(max_rows=3)
    b = ContinuousBatcher(m, chunk_frames=4)
    a1, hot, a2 = req(prompt(5, 1), temperature=0.7), req(prompt(6, 2), temperature=1.0), req(prompt(7, 3), temperature=0.7)
    for it in (a1, hot, a2):
        b.submit(it)
    done = b.run_until_idle()
    assert [id(x[0]) for x in done].index(id(hot)) < [id(x[0]) for x in done].index(id(a2))   # arrival order kept across groups
    hs = [int(it.request["prompt"][:, :-1].sum()) % 1000 for it in (a1, hot, a2)]
    assert m.admitted == hs
    temps = sorted({round(c[1], 3) for c in m._s.calls})
    assert temps == [round(float(torch.tensor(0.7, dtype=torch.bfloat16)), 3), 1.0]
    for c in m._s.calls:                                                          # one value per step
        assert c[1] in (float(torch.tensor(0.7, dtype=torch.bfloat16)), 1.0)


def test_requests_with_different_sampling_parameters_share_a_batch():
    """With per-request parameter arrays nothing waits for the batch to drain: the three requests run together from the first
    step, each row carrying its own (temperature, top_p, top_k) as generate() would round them; an out-of-range top_k fails
    its request alone; every request still gets its own stream."""
    m = FakeModel(max_rows=3, row_params=True)
    b = ContinuousBatcher(m, chunk_frames=4)
    assert b.per_row
    items = [req(prompt(5, 1), temperature=0.7, top_p=0.8, top_k=20), req(prompt(6, 2), temperature=1.0),
             req(prompt(7, 3), temperature=0.3, top_k=64), req(prompt(5, 4), top_k=65), req(prompt(5, 5), top_k=0)]
    for it in items:
        b.submit(it)
    done = b.run_until_idle()
    by_id = {id(it): res for it, res in done}
    assert isinstance(by_id[id(items[3])], ValueError) and isinstance(by_id[id(items[4])], ValueError)
    for it in items[:3]:
        assert torch.equal(by_id[id(it)], expected(it.request["prompt"], 0))
    t, p, k, idle = m._s.row_calls[0]
    assert idle == [0, 0, 0]                                   # all three in flight in the very first step
    bf = lambda v: float(torch.tensor(v, dtype=torch.bfloat16))
    assert [float(x) for x in t] == [bf(0.7), 1.0, bf(0.3)]
    assert [float(x) for x in p] == [bf(0.8), bf(0.9), bf(0.9)]
    assert [int(x) for x in k] == [20, 30, 64]


def test_long_prompts_are_filled_in_pieces_between_chunks():
    """prefill_budget: while rows are decoding, at most that many prompt columns are computed per chunk boundary, so the
    request in flight keeps producing frames while a long prompt comes in; every request still gets its own stream, and
    with nothing in flight the prompt pass runs in one piece."""
    m = FakeModel(max_rows=3)
    sizes = []
    real = m.prefill

    def spy(cols, pos0, slot):
        sizes.append((int(cols.shape[-1]), pos0, slot, decoding[0]))
        real(cols, pos0, slot)
    m.prefill = spy
    decoding = [0]                                             # rows in flight when the chunk boundary began
    b = ContinuousBatcher(m, chunk_frames=2, prefill_budget=5, share_min=0)
    first = req(prompt(9, 50))
    b.submit(first)
    assert b.step() == [] and sizes == [(8, 0, 0, 0)]          # nobody to stall: one piece
    long1, long2 = req(prompt(23, 51)), req(prompt(14, 52))
    b.submit(long1)
    b.submit(long2)
    done, steps_while_filling = [], 0
    while not b.idle():
        filling_before = bool(b.filling) or bool(b.waiting)
        decoding[0] = len(b.active)
        done.extend(b.step())
        steps_while_filling += int(filling_before and bool(b.filling))
    assert steps_while_filling >= 2                            # decode chunks ran while prompts were still coming in
    with_rows = [n for n, _, _, act in sizes[1:] if act > 0]
    assert with_rows and max(with_rows) <= 5                   # the budget held whenever somebody was decoding
    assert sum(n for n, _, slot, _ in sizes if slot == 1) == 22 and sum(n for n, _, slot, _ in sizes if slot == 2) == 13
    by_id = {id(it): res for it, res in done}
    for it in (first, long1, long2):
        assert torch.equal(by_id[id(it)], expected(it.request["prompt"], 0))


def test_bad_requests_fail_alone():
    m = FakeModel(max_rows=2)
    b = ContinuousBatcher(m, chunk_frames=3)
    good, long_, shape = req(prompt(5, 1)), req(prompt(S, 2)), req(torch.zeros((2, 4), dtype=torch.long))
    for it in (long_, good, shape):
        b.submit(it)
    res = {id(it): r for it, r in b.run_until_idle()}
    assert isinstance(res[id(long_)], ValueError) and "exceeds max_seq_len" in str(res[id(long_)])   # inference.py:262-265
    assert isinstance(res[id(shape)], ValueError)
    assert torch.equal(res[id(good)], expected(good.request["prompt"], 0))


def test_idle_rows_never_run_past_the_cache():
    m = FakeModel(max_rows=3)
    b = ContinuousBatcher(m, chunk_frames=5)
    p = prompt(4, 9)
    m.stream = lambda slot, pos: [SEM0 + pos % CB] + [pos % CB] * C          # never ends: runs to the end of the cache
    b.submit(req(p))
    (it, codes), = b.run_until_idle()
    assert codes.shape == (C, S - 4 - 1)
    with pytest.raises(ValueError):
        ContinuousBatcher(m, max_rows=4)


def test_requests_that_start_alike_share_the_prompt_pass():
    """Voice-clone serving: many requests start with the same long prefix (system prompt + reference audio).  The first one
    computes it, later ones copy it from a resident row (or keep it when they land on the same row) and compute only their
    own columns; outputs are what they are without sharing."""
    voice = prompt(40, 123)
    tails = [prompt(3 + i % 4, 200 + i) for i in range(7)]
    items = [req(torch.cat([voice, t], dim=1), max_new=5) for t in tails] + [req(prompt(12, 999), max_new=5)]
    m = FakeModel(max_rows=3)
    b = ContinuousBatcher(m, chunk_frames=4, share_min=16)
    for it in items:
        b.submit(it)
    got = {id(it): res for it, res in b.run_until_idle()}
    for it in items:
        assert torch.equal(got[id(it)], expected(it.request["prompt"], 5))
    total = sum(it.request["prompt"].shape[1] - 1 for it in items)
    assert b.prompt_cols == total and m.computed == total - b.shared_cols
    assert b.shared_cols >= 6 * 40 and m.copied > 0        # all but the first took the voice prefix
    # the same with the prompt passes cut into pieces: requests that arrive together still share (the prefix is looked up when
    # a request's first piece runs, and what a slot holds so far can already be copied)
    m2 = FakeModel(max_rows=3)
    b2 = ContinuousBatcher(m2, chunk_frames=4, share_min=16, prefill_budget=16)
    items2 = [req(it.request["prompt"], max_new=5) for it in items]
    for it in items2:
        b2.submit(it)
    got2 = {id(it): res for it, res in b2.run_until_idle()}
    for it in items2:
        assert torch.equal(got2[id(it)], expected(it.request["prompt"], 5))
    assert b2.prompt_cols == total and m2.computed == total - b2.shared_cols
    assert b2.shared_cols >= 6 * 16 and m2.copied > 0
    # every resident record still describes what is in the slot: the prompt-pass columns were not overwritten by idle rows
    for row, cols in b.resident.items():
        assert m.kv[row][: cols.shape[1]] == [int(x) for x in cols.sum(0)]
    m2 = FakeModel(max_rows=3)
    b2 = ContinuousBatcher(m2, chunk_frames=4, share_min=0)
    for it in items:
        b2.submit(it)
    b2.run_until_idle()
    assert b2.shared_cols == 0 and m2.computed == total


def test_queue_protocol_of_the_worker_thread():
    m = FakeModel(max_rows=2)
    q_in = launch_thread_safe_queue(model=m, max_rows=2, chunk_frames=4)
    items = [req(prompt(5 + i, 20 + i)) for i in range(5)] + [req(prompt(S + 1, 3))]
    for it in items:
        q_in.put(it)
    for it in items[:-1]:
        first = it.response_queue.get(timeout=30)
        second = it.response_queue.get(timeout=30)
        assert isinstance(first, WrappedGenerateResponse) and first.status == "success"
        assert isinstance(first.response, GenerateResponse) and first.response.action == "sample"
        assert torch.equal(first.response.codes, expected(it.request["prompt"], 0))
        assert second.status == "success" and second.response.action == "next"
    err = items[-1].response_queue.get(timeout=30)
    assert err.status == "error" and isinstance(err.response, ValueError)
    q_in.put(None)


def test_streamed_portions_add_up_to_the_final_codes():
    """request["stream"]: code frames arrive chunk by chunk as "partial" responses before the usual "sample" / "next" pair; the
    portions concatenated equal the final codes (the newest frame is held back: the last one produced is not part of the
    result), requests without the flag see no partial responses."""
    m = FakeModel(max_rows=2)
    q_in = launch_thread_safe_queue(model=m, max_rows=2, chunk_frames=3)
    items = [req(prompt(6, 40), stream=True), req(prompt(7, 41)), req(prompt(5, 42), max_new=5, stream=True),
             req(prompt(5, 43), max_new=1, stream=True)]
    for it in items:
        q_in.put(it)
    for it in items:
        got = []
        while True:
            r = it.response_queue.get(timeout=30)
            assert r.status == "success"
            got.append(r.response)
            if r.response.action == "next":
                break
        want = expected(it.request["prompt"], it.request["max_new_tokens"])
        assert [g.action for g in got[-2:]] == ["sample", "next"] and torch.equal(got[-2].codes, want)
        parts = [g.codes for g in got[:-2]]
        assert all(g.action == "partial" for g in got[:-2])
        if it.request.get("stream") and want.shape[1] > 0:
            assert len(parts) >= 1 and all(p.shape[0] == C and p.shape[1] >= 1 for p in parts)
            assert torch.equal(torch.cat(parts, 1), want)
            if want.shape[1] > 3:
                assert len(parts) > 1                         # really chunk by chunk, not once at the end
        else:
            assert parts == []
    q_in.put(None)
