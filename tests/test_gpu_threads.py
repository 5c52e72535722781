"""Two host threads, two CUDA streams, one process: a Dual-AR generate() and a DAC decode running at the same time (what
the reference's worker thread + codec thread do, inference.py:447-520 / inference_engine/__init__.py) must give the bytes
they give alone.  A second case runs two independent Dual-AR handles side by side (one cooperative frame kernel each)."""
import threading

import pytest
import torch

from oracle import dac_ref as D
from oracle import dual_ar_ref as R
from tests.test_gpu_dual_ar import _mk, _run_generate
from tests.util import cfg_from_golden, load_golden

pytestmark = pytest.mark.gpu


def _in_threads(*jobs):
    """Run the callables at the same time, each under its own non-default stream; re-raise the first failure."""
    out, errs = [None] * len(jobs), []
    gate = threading.Barrier(len(jobs))

    def run(i, fn):
        try:
            with torch.cuda.stream(torch.cuda.Stream()):
                gate.wait(timeout=60)
                out[i] = fn()
                torch.cuda.current_stream().synchronize()
        except BaseException as e:  # noqa: BLE001
            errs.append(e)

    ts = [threading.Thread(target=run, args=(i, fn)) for i, fn in enumerate(jobs)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=300)
    assert not any(t.is_alive() for t in ts), "a worker thread did not finish"
    if errs:
        raise errs[0]
    return out


def test_dual_ar_and_dac_on_two_threads_match_their_serial_results():
    from fish_speech_b200.dac import DACB200
    g = load_golden("dual_ar_tiny_greedy")
    cfg = cfg_from_golden(g)
    m = _mk(cfg, R.make_state_dict(cfg, seed=g["wseed"]))
    c = D.small_config()
    dac = DACB200(c.__dict__, D.make_state_dict(c, seed=3), device="cuda:0", max_batch=2, max_frames=64)
    codes = D.make_codes(c, 2, 60).cuda()
    want_seq = _run_generate(m, g, 64)
    want_wav = dac.from_indices(codes.clone()).cpu()
    torch.cuda.synchronize()

    def lm():
        return [_run_generate(m, g, 64) for _ in range(3)]

    def codec():
        return [dac.from_indices(codes.clone()).cpu() for _ in range(40)]

    seqs, wavs = _in_threads(lm, codec)
    assert all(torch.equal(s, want_seq) for s in seqs)
    assert all(torch.equal(w, want_wav) for w in wavs)
    m.session(1, 65).check_device()


def test_two_dual_ar_handles_on_two_threads():
    g = load_golden("dual_ar_tiny_greedy")
    cfg = cfg_from_golden(g)
    sd = R.make_state_dict(cfg, seed=g["wseed"])
    a, b = _mk(cfg, sd), _mk(cfg, sd)
    want = _run_generate(a, g, 48)
    assert torch.equal(_run_generate(b, g, 48), want)
    ra, rb = _in_threads(lambda: [_run_generate(a, g, 48) for _ in range(2)], lambda: [_run_generate(b, g, 48) for _ in range(2)])
    assert all(torch.equal(s, want) for s in ra + rb)
    a.session(1, 49).check_device()
    b.session(1, 49).check_device()
