"""Cost of streaming the codec: time per push of `n` new code frames (full DAC config, synthetic weights) for a few halo sizes,
and how far a short halo's samples are from the exact ones.  Usage: python tools/dac_stream_bench.py [n_frames_per_push] [batch]"""
import sys

import torch

sys.path.insert(0, ".")
from fish_speech_b200 import DACB200, DACConfig  # noqa: E402
from fish_speech_b200.dac import synthetic_state_dict  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 21
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    c = DACConfig()
    m = DACB200(c, synthetic_state_dict(c, seed=0), max_batch=batch, max_frames=1536)
    T = 1280
    g = torch.Generator().manual_seed(1)
    codes = torch.randint(0, c.codebook_size, (batch, c.n_codebooks + 1, T), generator=g).cuda()
    want = m.from_indices(codes.clone())
    hop = m.hop_length
    for halo in (None, 512, 256, 128, 64):
        s = m.stream(batch=batch, halo=halo)
        outs, times = [], []
        for t in range(0, T, n):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            outs.append(s.push(codes[:, :, t: t + n].clone()))
            e1.record()
            e1.synchronize()
            times.append(e0.elapsed_time(e1))
        got = torch.cat(outs, dim=2)
        tail = times[len(times) // 2:]
        err = (got - want)[:, :, hop * s.halo:] if s.halo else (got - want)
        ref = want[:, :, hop * s.halo:] if s.halo else want
        snr = 10 * torch.log10(ref.pow(2).sum() / err.pow(2).sum().clamp_min(1e-30))
        print(f"halo {s.halo:5d}: {sum(tail) / len(tail):7.2f} ms per push of {n} frames x {batch} clips "
              f"({n * hop / c.sample_rate * 1e3:.0f} ms of audio each), exact={bool(torch.equal(got, want))}, "
              f"SNR vs one pass {float(snr):.1f} dB")


if __name__ == "__main__":
    main()
