"""Instruction counts per kernel from `cuobjdump -sass` of the built library -> profiles/rNN_sass_summary.txt.
    python tools/sass_summary.py > profiles/r02_sass_summary.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "flexible-llm-sharding_b200", "csrc", "libfls_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
names = [l.split("Function : ")[1].strip() for l in sass.splitlines() if "Function : " in l]
demangled = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines() if names else []
if len(demangled) != len(names):
    demangled = names
short = {}
for n, d in zip(names, demangled):
    d = re.sub(r"^void ", "", d)
    d = re.sub(r"fls::\(anonymous namespace\)::|fls::<unnamed>::|fls::", "", d)
    short[n] = d.split("(CUtensorMap")[0].split("(const")[0].split("(__half")[0].split("(unsigned")[0].split("(float")[0]
cols = [("2CTA", r"UTCHMMA\.2CTA"), ("UTCHMMA", r"UTCHMMA"), ("UTMALDG", r"UTMALDG"), ("UTMASTG", r"UTMASTG"),
        ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"), ("UTCBAR", r"UTCBAR"), ("HMMA.16816", r"HMMA\.16816"),
        ("SYNCS", r"\bSYNCS"), ("MUFU.EX2", r"MUFU\.EX2")]
counts = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    if "Function : " in line:
        cur = short[line.split("Function : ")[1].strip()]
        counts[cur] = collections.Counter()
    elif cur and "/*" in line and ";" in line:
        for name, pat in cols:
            if re.search(pat, line):
                counts[cur][name] += 1
print("cuobjdump -sass flexible-llm-sharding_b200/csrc/libfls_b200.so (nvcc 12.9, -gencode arch=compute_100a,code=sm_100a):")
print("instruction counts per kernel.  UTCHMMA = tcgen05.mma (2CTA column: how many of them are cta_group::2), UTMALDG /")
print("UTMASTG = TMA tensor load / store, LDTM / STTM = tcgen05.ld / st (TMEM), UTCBAR = tcgen05.commit, SYNCS = mbarrier")
print("operations, HMMA.16816 = mma.sync (only the comparison attention kernel, used for head_dim 64).")
print("gemm_tcgen05_kernel<EPI, CG, TS>: EPI 0 store, 1 +residual, 2 SwiGLU, 3 lse partial, 4 probabilities, 5 store+RoPE; CG =")
print("cta_group; TS = staged TMA-store epilogue.  gemm_wide_kernel / gemm_wide_skew_kernel<EPI>: 256 x 512 tiles per CTA pair.")
print("attention_tcgen05_kernel<NK>: one CTA per (query tile, head), NK K buffers; attention_persistent_kernel: persistent CTAs.")
print()
w = max(len(k) for k in counts) + 2
print("kernel".ljust(w) + "".join(n.rjust(11) for n, _ in cols))
for k in sorted(counts):
    if any(counts[k][n] for n, _ in cols):
        print(k.ljust(w) + "".join(str(counts[k][n]).rjust(11) for n, _ in cols))
