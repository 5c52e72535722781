"""SASS evidence per kernel from the built library: python tools/sass_excerpts.py > profiles/r2_sass_excerpts.txt
For each kernel of interest: instruction count and every tensor-core / TMA / tensor-memory / mixed-precision mnemonic with its
count and two sample lines (tcgen05.mma -> UTC*MMA, tcgen05.commit -> UTCBAR, tcgen05.ld -> LDTM, cp.async.bulk.tensor -> UTMALDG,
cp.async.bulk -> UBLKCP, cp.async -> LDGSTS, mma.sync -> HMMA, fma.rn.f32.bf16 -> FHFMA.BF16)."""
import re
import subprocess
import sys
from collections import OrderedDict

LIB = "fish_speech_b200/libfish_b200.so"
KERNELS = OrderedDict([
    ("frame_mega_kernelILi128ELi4ELb0", "frame_mega_kernel<128,4> (persistent frame kernel, batch 1)"),
    ("rows_gemm_kernelILi64E", "rows_gemm_kernel<64> (many-row Linear, tcgen05 stream-K)"),
    ("attn_tc_kernelILi128ELi4E", "attn_tc_kernel<128,4> (many-row decode attention)"),
    ("tcv_kernelILi256ELb0", "tcv_kernel<256, bf16> (prompt-pass Linears)"),
    ("fa_prefill_kernel", "fa_prefill_kernel (causal flash attention of the prompt pass)"),
    ("tcv_kernelILi192ELb1", "tcv_kernel<192, fp16> (tcgen05 implicit-GEMM conv: DAC decoder)"),
    ("tcv_kernelILi96ELb1", "tcv_kernel<96, fp16> (the 96-channel decoder block)"),
    ("gemv_tma_kernelILi1ELi1ELi2E", "gemv_tma_kernel<1,RMSNORM,SWIGLU> (kernel chain GEMV)"),
])
PAT = re.compile(r"\b(UTC[A-Z]*MMA[.\w]*|UTCBAR[.\w]*|LDTM[.\w]*|STTM[.\w]*|UTMALDG[.\w]*|UTMASTG[.\w]*|UBLKCP[.\w]*|UBLKPF[.\w]*|LDGSTS[.\w]*|HMMA[.\w]*|"
                 r"FHFMA[.\w]*|SYNCS[.\w]*|USETMAXREG[.\w]*)")
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
funcs, cur = OrderedDict(), None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        funcs[cur] = []
    elif cur is not None and re.search(r"/\*[0-9a-f]{4,}\*/", line):
        funcs[cur].append(line.rstrip())
print("# SASS evidence: `cuobjdump -sass fish_speech_b200/libfish_b200.so` (nvcc 12.9, -gencode arch=compute_100a,code=sm_100a), tools/sass_excerpts.py")
print("# tcgen05.mma -> UTC*MMA, tcgen05.commit -> UTCBAR, tcgen05.ld -> LDTM, cp.async.bulk.tensor -> UTMALDG, cp.async.bulk -> UBLKCP,")
print("# cp.async -> LDGSTS, mbarrier -> SYNCS, mma.sync -> HMMA, fma.rn.f32.bf16 -> FHFMA.BF16")
for key, title in KERNELS.items():
    names = [n for n in funcs if key in n]
    if not names:
        print(f"\n## {title}\n   (not in this build)")
        continue
    name = names[0]
    lines = funcs[name]
    ins = [l for l in lines if re.search(r"/\*[0-9a-f]{4,}\*/\s+\S", l) and not re.search(r"^\s*/\*[0-9a-f]+\*/\s*$", l)]
    n_ins = sum(1 for l in ins if re.search(r"/\*[0-9a-f]{4,}\*/\s+[@A-Z]", l))
    found = OrderedDict()
    for l in ins:
        m = PAT.search(l)
        if m:
            found.setdefault(m.group(1), []).append(l.strip().split("/* 0x")[0].rstrip())
    print(f"\n## {title}\n   symbol {name}: {n_ins} instructions")
    print("   " + ", ".join(f"{k} x{len(v)}" for k, v in found.items()))
    for k, v in found.items():
        if k.startswith("SYNCS"):
            continue
        for l in v[:2]:
            print("     " + l)
