"""fish_speech_b200 — B200-native (sm_100a) hot path of fish-speech: Dual-AR decode + DAC codec decode.

Only what the path needs lives here: `csrc/` (CUDA kernels + the C ABI of include/fish_b200.h) and the
host-side mirrors of the reference interfaces (`dual_ar`, `dac`).  Importing the package does not touch
CUDA; constructing a model does, and raises if the extension or a B200 is missing (no CPU fallback).
"""
from .dual_ar import (DualARConfig, DualARTransformerB200, decode_n_tokens, decode_one_token_ar, generate,  # noqa: F401
                      init_model_from_state_dict, make_semantic_logit_bias, s2_pro_config)

from .dac import DACB200, DACConfig, DACStream  # noqa: F401
from . import partition  # noqa: F401

__all__ = ["DACB200", "DACConfig", "DACStream", "partition", "DualARConfig", "DualARTransformerB200", "decode_one_token_ar", "decode_n_tokens", "generate",
           "init_model_from_state_dict", "make_semantic_logit_bias", "s2_pro_config"]

from .checkpoint import init_model, load_codec, load_config, load_weights  # noqa: E402,F401
from . import scheduler  # noqa: E402,F401
