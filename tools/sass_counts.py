#!/usr/bin/env python3
"""Per-kernel SASS instruction counts of libdau_conv_b200.so (cuobjdump -sass): the mnemonics that show how the
kernels move and multiply -- UTMALDG (TMA tensor load), UBLKCP (bulk copy), SYNCS (mbarrier), FFMA2 (packed fp32
FMA), FFMA, LDS.64 / LDS.128, LDG / STG, ATOM / RED, HMMA / UTC*MMA (tensor cores: none expected).
usage: python tools/sass_counts.py [path to .so] > profiles/rN_sass_counts.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "dau-convnet_b200", "dau_conv", "libdau_conv_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", txt)), capture_output=True, text=True).stdout.split("\n")
pats = [("UTMALDG", r"\bUTMALDG"), ("UBLKCP", r"\bUBLKCP"), ("SYNCS", r"\bSYNCS"), ("FFMA2", r"\bFFMA2\b"), ("FFMA", r"\bFFMA\b"),
        ("FMUL2", r"\bFMUL2\b"), ("LDS.64", r"\bLDS(\.U)?\.64"), ("LDS.128", r"\bLDS(\.U)?\.128"), ("LDS", r"\bLDS\b"),
        ("STS", r"\bSTS"), ("LDG", r"\bLDG"), ("STG", r"\bSTG"), ("ATOM/RED", r"\b(ATOM|ATOMG|RED)\b"),
        ("HMMA/UTCMMA", r"\b(HMMA|UTC\w*MMA)"), ("LDL/STL", r"\b(LDL|STL)\b"), ("SETMAXREG", r"SETMAXREG")]
print("file: %s (%d bytes)" % (os.path.relpath(so, ROOT), os.path.getsize(so)))
print("%-64s %s" % ("kernel", " ".join("%8s" % p[0] for p in pats)))
blocks = re.split(r"\n\s*Function : ", txt)[1:]
tot = collections.Counter()
for name, blk in zip(names, blocks):
    short = re.sub(r"\(anonymous namespace\)::|dau::|\(CUtensorMap_st.*|\(dau::.*|\(float.*|\(int.*", "", name).replace("void ", "")
    c = [len(re.findall(p, blk)) for _, p in pats]
    for (k, _), v in zip(pats, c):
        tot[k] += v
    print("%-64s %s" % (short[:64], " ".join("%8d" % v for v in c)))
print("%-64s %s" % ("TOTAL", " ".join("%8d" % tot[k] for k, _ in pats)))
