"""GPU parity of the DAC codec decode (fsb_dac_decode through the host mirror) against the golden waveforms the
reference classes produced (oracle/gen_golden_dac.py) and the fp32 CPU oracle.  Bar: SNR >= 40 dB (north star);
the fp32 round-1 path is expected far above it, and the number is printed."""
import math

import pytest
import torch

from oracle import dac_ref as D
from tests.util import load_golden

pytestmark = pytest.mark.gpu


def snr_db(ref, y):
    return 10 * math.log10(float(ref.double().pow(2).sum()) / max(float((ref.double() - y.double()).pow(2).sum()), 1e-30))


def _model(c, sd, max_batch=2, max_frames=32):
    from fish_speech_b200.dac import DACB200
    return DACB200(c.__dict__, sd, device="cuda:0", max_batch=max_batch, max_frames=max_frames)


@pytest.mark.parametrize("tf32", [1, 0])
@pytest.mark.parametrize("name", ["dac_small", "dac_full_T12"])
def test_from_indices_vs_reference_golden(name, tf32):
    """tf32=1: the default tensor-core path (TF32 operands, fp32 accumulate), bar SNR >= 40 dB (north star; the
    reference's own bf16-autocast path is at 43 dB); tf32=0: the fp32 CUDA-core parity anchor, bar 90 dB."""
    g = load_golden(name)
    c = D.DACConfig(**g["cfg"])
    sd = D.make_state_dict(c, seed=g["seed"])
    m = _model(c, sd)
    m.set_option("tf32", tf32)
    codes = g["codes"].clone().cuda()
    y = m.from_indices(codes).cpu()
    assert y.shape == g["wav"].shape and y.dtype == torch.float32
    s = snr_db(g["wav"], y)
    print(f"[parity] {name} tf32={tf32}: SNR vs reference fp32 waveform {s:.1f} dB, max-abs {float((y - g['wav']).abs().max()):.2e}")
    assert s >= (40.0 if tf32 else 90.0)
    assert float(y.abs().max()) <= 1.0
    # the reference clamps the caller's indices in place (rvq.py:354-359); so does the engine
    exp = g["codes"].clone()
    exp[:, 0].clamp_(max=c.semantic_codebook_size - 1)
    exp[:, 1:].clamp_(max=c.codebook_size - 1)
    assert torch.equal(codes.cpu(), exp)


def test_causality_and_batch_chunking():
    """rvq.py:374-399 self-check (prefix decode == decode prefix) and batch > max_batch goes through in chunks."""
    c = D.small_config()
    sd = D.make_state_dict(c, seed=3)
    m = _model(c, sd, max_batch=2, max_frames=24)
    codes = D.make_codes(c, 5, 20)
    full = m.from_indices(codes.clone().cuda()).cpu()
    pre = m.from_indices(codes[:, :, :12].clone().cuda()).cpu()
    assert torch.allclose(full[..., : pre.shape[-1]], pre, atol=1e-5)   # causal: a prefix decodes to the prefix
    ref = D.from_indices(c, sd, codes.clone())
    assert snr_db(ref, full) >= 40.0
    m.set_option("tf32", 0)
    assert snr_db(ref, m.from_indices(codes.clone().cuda()).cpu()) >= 60.0
    m.set_option("tf32", 1)
    one = m.from_indices(codes[3:4].clone().cuda()).cpu()
    assert torch.equal(one[0], full[3])


def test_errors():
    c = D.small_config()
    m = _model(c, D.make_state_dict(c, seed=3), max_frames=8)
    with pytest.raises(ValueError):
        m.from_indices(torch.zeros((1, 10, 9), dtype=torch.long).cuda())
    with pytest.raises(ValueError):
        m.from_indices(torch.zeros((1, 3, 4), dtype=torch.long).cuda())
