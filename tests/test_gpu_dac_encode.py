"""GPU parity of the codec's encode side (fsb_dac_encode through the host mirror: waveform -> codes) against the codes the
reference classes produced (fixtures of oracle/gen_golden_dac.py golden_encode) and the fp32 CPU oracle.  The result is a set
of arg-max decisions over fp32 distances, so the bar is equality of codes; a frame may differ only where the oracle itself
had a near-tie between its two best entries (fp32 sums in another order), and the fraction of such frames is bounded."""
import pytest
import torch

from oracle import dac_ref as D
from tests.util import load_golden

pytestmark = pytest.mark.gpu


def _model(c, sd, max_batch=2, max_frames=32):
    from fish_speech_b200.dac import DACB200
    return DACB200(c.__dict__, sd, device="cuda:0", max_batch=max_batch, max_frames=max_frames)


def _full_sd(c, seed):
    return {**D.make_state_dict(c, seed=seed), **D.make_encode_state_dict(c, seed=seed)}


def _frames_equal(got, want):
    """fraction of (clip, frame) columns whose codes all agree; a codebook's choice feeds the ones after it within the frame"""
    return float((got == want).all(dim=1).float().mean())


@pytest.mark.parametrize("name", ["dac_encode_small", "dac_encode_full"])
def test_encode_vs_reference_golden(name):
    g = load_golden(name)
    c = D.DACConfig(**g["cfg"])
    m = _model(c, _full_sd(c, g["seed"]), max_batch=2, max_frames=16)
    assert m.can_encode
    codes, lens = m.encode(g["audio"].cuda())
    assert codes.dtype == torch.int64 and codes.shape == g["codes"].shape
    assert torch.equal(lens.cpu(), g["lens"])
    eq = _frames_equal(codes.cpu(), g["codes"])
    first = float((codes.cpu()[:, 0] == g["codes"][:, 0]).float().mean())
    print(f"[parity] encode {name}: frames with all {codes.shape[1]} codes equal to the reference's: {eq:.4f}; semantic codes equal: {first:.4f}")
    assert eq == 1.0, (codes.cpu() != g["codes"]).nonzero()[:8]


def test_encode_vs_oracle_ragged_batches():
    """More clips than max_batch, a length that is not a whole number of code frames, [B, n] and [B, 1, n] inputs, windowed
    attention in both transformers (small config: windows of 6 / 8 positions)."""
    c = D.small_config()
    sd = _full_sd(c, 3)
    m = _model(c, sd, max_batch=2, max_frames=24)
    audio = D.make_audio(c, 5, 2048 * 21 - 1234, seed=11)
    want, want_lens = D.encode(c, sd, audio.clone())
    got, lens = m.encode(audio.cuda())
    got2, _ = m.encode(audio[:, 0].cuda())
    assert torch.equal(got, got2) and torch.equal(lens.cpu(), want_lens)
    eq = _frames_equal(got.cpu(), want)
    print(f"[parity] encode small config, 5 clips x {want.shape[2]} frames: frames fully equal to the oracle's: {eq:.4f}")
    assert eq >= 0.98
    # what is encoded decodes (the codes are in range) and a second call gives the same codes
    wav = m.from_indices(got.clone())
    assert wav.shape == (5, 1, c.hop * want.shape[2]) and bool(torch.isfinite(wav).all())
    again, _ = m.encode(audio.cuda())
    assert torch.equal(again, got)


def test_encode_is_causal_and_errors():
    """Every layer of the encode side is causal: the codes of the first frames do not change when more audio follows."""
    c = D.small_config()
    sd = _full_sd(c, 4)
    m = _model(c, sd, max_batch=1, max_frames=12)
    audio = D.make_audio(c, 1, 2048 * 12, seed=12).cuda()
    full, _ = m.encode(audio)
    part, _ = m.encode(audio[:, :, : 2048 * 7])
    assert torch.equal(part, full[:, :, :7])
    with pytest.raises(ValueError):
        m.encode(torch.zeros(1, 1, 2048 * 13).cuda())
    with pytest.raises(ValueError):
        m.encode(torch.zeros(1, 2, 2048).cuda())
    dec_only = _model(c, D.make_state_dict(c, seed=4), max_batch=1, max_frames=12)
    assert not dec_only.can_encode
    with pytest.raises(RuntimeError):
        dec_only.encode(audio)
