"""Continuous batching on the engine (fish_speech_b200/scheduler.py): requests that join and leave a running batch must get
exactly the frames they get when they run alone through the same many-row kernels (greedy sampling: top_k = 1, RAS off, so the
draws do not matter), through the reference's queue protocol (inference.py:736-799)."""
import queue

import pytest
import torch

from oracle import dual_ar_ref as R
from tests.util import cfg_from_golden, load_golden

pytestmark = pytest.mark.gpu


def _model(max_rows):
    from fish_speech_b200 import DualARTransformerB200
    g = load_golden("dual_ar_tiny_greedy")
    cfg = cfg_from_golden(g)
    m = DualARTransformerB200(cfg.__dict__, R.make_state_dict(cfg, seed=g["wseed"]), device="cuda:0", max_rows=max_rows)
    m.setup_caches(max_rows, cfg.max_seq_len)
    return cfg, m


def _same_draws_everywhere(b, cfg):
    """Exact ties between top candidates are resolved by the draws, which are keyed by frame and row: give every frame and
    row the same record so that a request's tokens do not depend on where and when it runs."""
    n = 2 * cfg.vocab_size + (cfg.num_codebooks - 1) * cfg.codebook_size
    one = (torch.randint(1, 256, (n,), generator=torch.Generator().manual_seed(4)).float() / 256).to(torch.bfloat16)
    b.sess.uniforms = one.view(1, 1, n).expand(b.ring, b.rows, n).contiguous().cuda()


def _requests(cfg, n):
    from fish_speech_b200.scheduler import GenerateRequest
    out = []
    for i in range(n):
        T = 6 + 3 * (i % 5)
        p = R.make_prompt(cfg, T, seed=700 + i, voice_span=3 if i % 2 else 0)
        out.append(GenerateRequest(request=dict(prompt=p, max_new_tokens=5 + 4 * (i % 4), temperature=0.8, top_p=0.8, top_k=1),
                                   response_queue=queue.Queue()))
    return out


def test_joining_and_leaving_rows_get_the_frames_they_get_alone():
    from fish_speech_b200.scheduler import ContinuousBatcher
    cfg, m = _model(4)
    b = ContinuousBatcher(m, chunk_frames=4, use_ras=False, ring_frames=16)   # a short ring: the record wraps during the run
    _same_draws_everywhere(b, cfg)
    b.sess.bias[cfg.im_end_id] = 1.5             # some requests end on <|im_end|> before their budget
    items = _requests(cfg, 9)
    for it in items:
        b.submit(it)
    got = {id(it): res for it, res in b.run_until_idle()}
    assert len(got) == len(items)
    b.sess.check_device()
    n_eos = 0
    for it in items:
        solo = ContinuousBatcher(m, chunk_frames=4, use_ras=False, ring_frames=16)
        _same_draws_everywhere(solo, cfg)
        solo.sess.bias[cfg.im_end_id] = 1.5
        solo.submit(it)
        (_, want), = solo.run_until_idle()
        codes = got[id(it)]
        assert codes.shape[0] == cfg.num_codebooks and codes.dtype == torch.int32
        assert torch.equal(codes.cpu(), want.cpu()), (codes, want)
        assert bool((codes >= 0).all()) and bool((codes < cfg.codebook_size).all())
        n_eos += int(codes.shape[1] < it.request["max_new_tokens"] - 1)
    print(f"[sched] 9 requests over 4 rows: {b.steps} steps, {b.frames_kept} frames kept of {b.frames_run} row-frames run, "
          f"{n_eos} ended on <|im_end|>")
    assert b.frames_kept <= b.frames_run


def test_worker_thread_serves_the_queue_protocol():
    from fish_speech_b200.scheduler import launch_thread_safe_queue
    cfg, m = _model(4)
    q_in = launch_thread_safe_queue(model=m, max_rows=4, chunk_frames=4)
    items = _requests(cfg, 6)
    bad = _requests(cfg, 1)[0]
    bad.request["prompt"] = R.make_prompt(cfg, cfg.max_seq_len, seed=1)
    for it in items + [bad]:
        q_in.put(it)
    for it in items:
        a, b_ = it.response_queue.get(timeout=120), it.response_queue.get(timeout=120)
        assert a.status == "success" and a.response.action == "sample" and b_.response.action == "next"
        n = a.response.codes.shape[1]
        assert a.response.codes.shape[0] == cfg.num_codebooks and 0 <= n <= it.request["max_new_tokens"] - 1
    e = bad.response_queue.get(timeout=120)
    assert e.status == "error" and isinstance(e.response, ValueError)
    q_in.put(None)


def test_shared_prompt_prefix_is_copied_not_recomputed_and_changes_nothing():
    """Requests that start with the same 40 columns (a reference voice): the first computes them, the others copy the keys
    and values from a resident slot and run the prompt pass over their own columns only; every request still gets the frames
    it gets alone without sharing, bit for bit."""
    from fish_speech_b200.scheduler import ContinuousBatcher, GenerateRequest
    cfg, m = _model(4)
    voice = R.make_prompt(cfg, 40, seed=900, voice_span=30)
    items = []
    for i in range(7):
        p = torch.cat([voice, R.make_prompt(cfg, 4 + i % 3, seed=910 + i)], dim=1)
        items.append(GenerateRequest(request=dict(prompt=p, max_new_tokens=6 + i % 3, temperature=0.8, top_p=0.8, top_k=1),
                                     response_queue=queue.Queue()))
    b = ContinuousBatcher(m, chunk_frames=4, use_ras=False, share_min=16, ring_frames=16)
    _same_draws_everywhere(b, cfg)
    for it in items:
        b.submit(it)
    got = {id(it): res for it, res in b.run_until_idle()}
    assert b.shared_cols >= 6 * 40
    print(f"[sched] shared prefix: {b.shared_cols} of {b.prompt_cols} prompt columns taken from resident slots")
    for it in items:
        solo = ContinuousBatcher(m, chunk_frames=4, use_ras=False, share_min=0, ring_frames=16)
        _same_draws_everywhere(solo, cfg)
        solo.submit(it)
        (_, want), = solo.run_until_idle()
        assert solo.shared_cols == 0
        assert torch.equal(got[id(it)].cpu(), want.cpu())


def test_requests_with_their_own_sampling_parameters_share_a_batch():
    """fsb_frame_args.row_temperature / row_top_p / row_top_k: three requests with different (temperature, top_p, top_k) decode
    side by side and each gets, bit for bit, the frames it gets alone through the scalar parameters of the same call (real
    sampling: top_k > 1, RAS on; the draws are the same record for every frame and row, so placement does not matter)."""
    from fish_speech_b200.scheduler import ContinuousBatcher, GenerateRequest
    cfg, m = _model(4)
    params = [(0.7, 0.8, 20), (1.0, 0.9, 30), (0.3, 0.5, 3), (1.2, 0.95, 64)]
    items = []
    for i, (t, p, k) in enumerate(params):
        pr = R.make_prompt(cfg, 7 + 2 * i, seed=800 + i, voice_span=3 if i % 2 else 0)
        items.append(GenerateRequest(request=dict(prompt=pr, max_new_tokens=14, temperature=t, top_p=p, top_k=k),
                                     response_queue=queue.Queue()))
    b = ContinuousBatcher(m, chunk_frames=4, use_ras=True, ring_frames=32)
    assert b.per_row
    _same_draws_everywhere(b, cfg)
    for it in items:
        b.submit(it)
    got = {id(it): res for it, res in b.run_until_idle()}
    assert b.steps <= 5                                        # together, not one after the other
    b.sess.check_device()
    outs = []
    for it in items:
        solo = ContinuousBatcher(m, chunk_frames=4, use_ras=True, ring_frames=32)
        solo.per_row = False                                   # the scalar parameters of fsb_frame_args
        solo.sess.row_temperature = solo.sess.row_top_p = solo.sess.row_top_k = None
        _same_draws_everywhere(solo, cfg)
        solo.submit(it)
        (_, want), = solo.run_until_idle()
        assert torch.equal(got[id(it)].cpu(), want.cpu()), (it.request["temperature"], got[id(it)], want)
        outs.append(want.cpu())
    assert not torch.equal(outs[0][:, :8], outs[2][:, :8]) or not torch.equal(outs[1][:, :8], outs[3][:, :8])


def test_prompts_filled_in_pieces_give_the_same_frames():
    """prefill_budget: prompt passes cut into pieces of at most 5 columns between chunks of frames, idle-while-filling rows
    parked behind their prompt: every request gets bit for bit the frames of the one-piece prompt pass."""
    from fish_speech_b200.scheduler import ContinuousBatcher
    cfg, m = _model(4)
    items = _requests(cfg, 8)
    outs = []
    for budget in (None, 5):
        b = ContinuousBatcher(m, chunk_frames=4, use_ras=False, ring_frames=16, prefill_budget=budget, share_min=0)
        _same_draws_everywhere(b, cfg)
        for it in items:
            b.submit(it)
        outs.append({id(it): res.cpu() for it, res in b.run_until_idle()})
        b.sess.check_device()
        steps = b.steps
    assert steps >= 1
    for it in items:
        assert torch.equal(outs[0][id(it)], outs[1][id(it)]), (outs[0][id(it)], outs[1][id(it)])
