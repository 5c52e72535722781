"""oracle/ — TEST INFRASTRUCTURE ONLY.

CPU restatement (plain PyTorch-on-CPU, fp32/bf16 op-by-op) of the reference's
algorithm for the one hot path this repo accelerates: the Dual-AR decode step
(`fish_speech/models/text2semantic/{llama,inference}.py`) and the DAC codec
decode (`fish_speech/models/dac/{modded_dac,rvq}.py` + the un-vendored
`descript-audio-codec==1.0.0` symbols it calls).

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` /
`--impl reference` legs may import this package, and only as the checker /
the timed CPU baseline.  Nothing under `fish_speech_b200/` imports it; the
product path has no CPU fallback.

`oracle/make_ref.py` is the recipe for `oracle/_ref/` (git-ignored, travels to the GPU box): the reference's own ten files for
the path, copied unmodified from the read-only checkout, so that the CPU arm of `bench.py` times the reference itself
(`kind: "reference"`); `oracle/refshim/` stands in for the absent descript packages those files import.

Pinning status
--------------
* Dual-AR: the reference has no tests / golden vectors, so the
  restatement is pinned against THE REFERENCE ITSELF, imported unmodified from
  `/root/reference` on CPU by `oracle/gen_golden.py`, which (a) asserts the
  restatement is bit-identical to the reference on the same seeded weights and
  (b) writes the fixtures under `tests/golden/`.
* DAC: the reference classes import only through a shim for the absent
  third-party `dac` / `audiotools` packages (`oracle/refshim/`), so the DAC
  oracle is "reference classes + shim" — **parity unpinned** with respect to the
  real `descript-audio-codec==1.0.0` wheel (formulas cross-checked against the
  local Hugging Face `transformers.models.dac` port in tests).
"""
