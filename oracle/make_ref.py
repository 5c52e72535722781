"""Recipe for `oracle/_ref/`: the reference's OWN Python files for the Dual-AR path, taken unmodified from the read-only
checkout where they lie and placed in a git-ignored directory that still travels to the GPU box (like the built .so files).
Nothing under `oracle/_ref/` is ever committed, and no product code imports it: `bench.py --impl reference` and
`cpu_baseline` time the reference's `generate()` through it on the host cores (kind "reference"); when the directory is
missing they fall back to the oracle port (kind "port").

    python -m oracle.make_ref            # run by __graft_entry__.build() when /root/reference exists

TEST / MEASUREMENT INFRASTRUCTURE — see oracle/__init__.py.
"""
from __future__ import annotations

import os
import shutil
import sys
import types
from pathlib import Path

HERE = Path(__file__).resolve().parent
DEST = HERE / "_ref"
REF = Path(os.environ.get("FISH_REFERENCE", "/root/reference"))

# what `fish_speech.models.text2semantic.inference` imports at module level (checked by importing it from DEST alone)
FILES = [
    "fish_speech/content_sequence.py",
    "fish_speech/conversation.py",
    "fish_speech/tokenizer.py",
    "fish_speech/models/text2semantic/__init__.py",
    "fish_speech/models/text2semantic/llama.py",
    "fish_speech/models/text2semantic/inference.py",
    "fish_speech/models/text2semantic/lora.py",
    # codec decode side (descript's dac / audiotools packages are absent: oracle/refshim restates the few classes they lend)
    "fish_speech/models/dac/__init__.py",
    "fish_speech/models/dac/modded_dac.py",
    "fish_speech/models/dac/rvq.py",
]
MARK = DEST / "fish_speech" / "models" / "text2semantic" / "inference.py"


def available() -> bool:
    return MARK.exists()


def make(verbose: bool = True) -> bool:
    """Copies FILES from the reference checkout into oracle/_ref/; False (and no change) when there is no checkout."""
    if not (REF / "fish_speech/models/text2semantic/inference.py").exists():
        return False
    for rel in FILES:
        src, dst = REF / rel, DEST / rel
        dst.parent.mkdir(parents=True, exist_ok=True)
        if src.exists():
            shutil.copyfile(src, dst)
        else:
            dst.write_text("")
    for pkg in ("fish_speech", "fish_speech/models"):   # namespace packages in the checkout: plain packages here
        init = DEST / pkg / "__init__.py"
        if not init.exists():
            init.write_text("")
    if verbose:
        print(f"[oracle] {len(FILES)} reference files -> {DEST}")
    return True


def import_vendored():
    """(llama, inference) modules of the vendored reference; stubs as in oracle/gen_golden.py (loralib is imported by lora.py and
    never used at inference; oracle/refshim stands in for descript's dac / audiotools, which the codec modules import)."""
    if not available():
        raise RuntimeError("oracle/_ref is missing: run `python -m oracle.make_ref` where /root/reference exists")
    for name in [n for n in sys.modules if n == "fish_speech" or n.startswith("fish_speech.")]:
        mod = sys.modules[name]
        if not str(getattr(mod, "__file__", "") or "").startswith(str(DEST)):
            del sys.modules[name]      # a copy imported from the checkout (gen_golden) must not shadow the vendored one
    for p in (str(HERE / "refshim"), str(DEST)):
        if p not in sys.path:
            sys.path.insert(0, p)
    sys.modules.setdefault("loralib", types.ModuleType("loralib"))
    import fish_speech.models.text2semantic.inference as inf
    import fish_speech.models.text2semantic.llama as llama
    inf.tqdm = lambda it, *a, **k: it
    return llama, inf


if __name__ == "__main__":
    ok = make()
    print("ok" if ok else f"no reference checkout at {REF}: nothing done")
