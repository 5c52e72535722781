"""ctypes binding of libfish_b200.so (include/fish_b200.h).  Fails loudly: there is no CPU fallback."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "libfish_b200.so"

_lib = None


class DualARDesc(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "dim", "n_layer", "n_fast_layer", "n_head", "n_kv_head", "head_dim", "ffn", "vocab",
        "num_codebooks", "codebook_size", "semantic_begin_id", "semantic_end_id", "im_end_id",
        "qk_norm", "scale_codebook_embeddings", "norm_fastlayer_input", "tie_word_embeddings", "max_rows")] + [
        ("norm_eps", C.c_float)]


class FrameArgs(C.Structure):
    _fields_ = [
        ("batch", C.c_int32),
        ("cur_tokens", C.c_void_p), ("pos", C.c_void_p), ("slots", C.c_void_p), ("ras_window", C.c_void_p),
        ("out_tokens", C.c_void_p), ("max_frames", C.c_int32), ("frame_counter", C.c_void_p), ("done", C.c_void_p),
        ("logit_bias", C.c_void_p), ("restrict_scan", C.c_int32),
        ("temperature", C.c_float), ("top_p", C.c_float), ("top_k", C.c_int32),
        ("uniforms", C.c_void_p), ("seed", C.c_uint64),
        ("row_temperature", C.c_void_p), ("row_top_p", C.c_void_p), ("row_top_k", C.c_void_p),
    ]


class DACDesc(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "latent_dim", "decoder_dim", "n_rates", "rate0", "rate1", "rate2", "rate3", "n_codebooks", "codebook_size",
        "semantic_codebook_size", "codebook_dim", "n_up", "up0", "up1", "tf_layers", "tf_heads", "tf_dim", "tf_ffn",
        "tf_head_dim", "tf_window", "convnext_ratio", "max_batch", "max_frames")] + [("tf_norm_eps", C.c_float)]


class TcConvArgs(C.Structure):
    _fields_ = [
        ("A16", C.c_void_p), ("W16", C.c_void_p),
        ("B", C.c_int32), ("T", C.c_int32), ("Cin", C.c_int32), ("N", C.c_int32), ("taps", C.c_int32), ("dil", C.c_int32),
        ("f16", C.c_int32), ("use_simt", C.c_int32),
        ("bias", C.c_void_p), ("bias_mod", C.c_int32), ("gamma", C.c_void_p),
        ("res32", C.c_void_p), ("res16", C.c_void_p), ("ldres", C.c_int32),
        ("alpha", C.c_void_p), ("alpha_mod", C.c_int32), ("act", C.c_int32), ("round_lin", C.c_int32),
        ("out16", C.c_void_p), ("ld16", C.c_int32), ("out32", C.c_void_p), ("ld32", C.c_int32),
        ("out16r", C.c_void_p), ("ld16r", C.c_int32),
    ]


EXPORTS = {
    # name: (restype, argtypes)
    "fsb_abi_version": (C.c_int, []),
    "fsb_dual_ar_workspace_bytes": (C.c_size_t, [C.POINTER(DualARDesc), C.c_int32]),
    "fsb_dual_ar_create": (C.c_int, [C.POINTER(DualARDesc), C.c_int32, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "fsb_dual_ar_destroy": (C.c_int, [C.c_void_p]),
    "fsb_dual_ar_last_error": (C.c_char_p, [C.c_void_p]),
    "fsb_dual_ar_bind_weight": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int64, C.c_int64]),
    "fsb_dual_ar_bind_rope": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32]),
    "fsb_dual_ar_bind_kv": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32]),
    "fsb_dual_ar_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "fsb_dual_ar_prefill": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "fsb_dual_ar_slow_step": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fsb_dual_ar_fast_step": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "fsb_dual_ar_buffer": (C.c_int, [C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "fsb_dual_ar_decode_frames": (C.c_int, [C.c_void_p, C.POINTER(FrameArgs), C.c_int32, C.c_void_p]),
    "fsb_dual_ar_device_error": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.c_void_p]),
    "fsb_dual_ar_frame_bytes": (C.c_double, [C.c_void_p, C.c_int32, C.c_int64, C.c_int32]),
    "fsb_dual_ar_launches_per_frame": (C.c_int32, [C.c_void_p, C.c_int32]),
    "fsb_gemv": (C.c_int, [C.c_void_p] * 6 + [C.c_int32] * 5 + [C.c_float, C.c_void_p]),
    "fsb_rows_gemm_scratch_bytes": (C.c_size_t, []),
    "fsb_rows_gemm": (C.c_int, [C.c_void_p] * 5 + [C.c_int32] * 4 + [C.c_void_p, C.c_size_t, C.c_void_p]),
    "fsb_rows_gemm_narrow": (C.c_int, [C.c_void_p] * 3 + [C.c_int32] * 4 + [C.c_void_p]),
    "fsb_rows_gemm_static": (C.c_int, [C.c_void_p] * 4 + [C.c_int32] * 4 + [C.c_void_p, C.c_char_p, C.c_size_t]),
    "fsb_sample": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_int32,
                             C.c_float, C.c_float, C.c_void_p, C.c_uint64, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                             C.c_void_p, C.c_void_p]),
}

DAC_EXPORTS = {
    "fsb_dac_workspace_bytes": (C.c_size_t, [C.POINTER(DACDesc)]),
    "fsb_dac_create": (C.c_int, [C.POINTER(DACDesc), C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "fsb_dac_destroy": (C.c_int, [C.c_void_p]),
    "fsb_dac_last_error": (C.c_char_p, [C.c_void_p]),
    "fsb_dac_bind_weight": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int64]),
    "fsb_dac_decode": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "fsb_dac_encode": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p]),
    "fsb_dac_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "fsb_dac_stage": (C.c_int, [C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "fsb_tc_conv": (C.c_int, [C.POINTER(TcConvArgs), C.c_void_p]),
    "fsb_dac_flops": (C.c_double, [C.c_void_p, C.c_int32, C.c_int32]),
}


if (PKG / "csrc" / "dac.cu").exists():
    EXPORTS.update(DAC_EXPORTS)


def header_symbols() -> list[str]:
    """Function names declared in include/fish_b200.h (used by the CPU-side export test)."""
    import re
    txt = (PKG.parent / "include" / "fish_b200.h").read_text()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(fsb_[a-z0-9_]+)\s*\(", txt)))


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load the shared library; build it with nvcc if it is not there.  Raises if neither works."""
    global _lib
    if _lib is not None:
        return _lib
    if build_if_missing:
        from . import build as _build
        _build.build()
    if not LIB_PATH.exists():
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -m fish_speech_b200.build` "
                           "(there is no CPU fallback)")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in EXPORTS.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.fsb_abi_version() != 1:
        raise RuntimeError("libfish_b200.so ABI version mismatch")
    _lib = lib
    return lib


def require_cuda():
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("fish_speech_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    major, _ = torch.cuda.get_device_capability()
    if major < 10:
        raise RuntimeError("fish_speech_b200 kernels are compiled for sm_100a only")
