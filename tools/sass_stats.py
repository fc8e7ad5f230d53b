#!/usr/bin/env python
"""Static SASS statistics of the trace kernels: instructions, code bytes and opcode-class mix per function.
usage: sass_stats.py [path/to/trace.o]   (default: gaussian-ray-tracing_b200/csrc/trace.o)"""
import collections
import re
import subprocess
import sys

obj = sys.argv[1] if len(sys.argv) > 1 else "gaussian-ray-tracing_b200/csrc/trace.o"
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
CLASSES = [("MUFU", r"^MUFU"), ("LDS/STS", r"^(LDS|STS|LDSM)"), ("LDG/LD", r"^(LDG|LD\b|LDC|ULDC)"), ("LDL/STL", r"^(LDL|STL)"), ("STG", r"^(STG|ST\b|RED|ATOM)"),
           ("SHFL", r"^SHFL"), ("VOTE/REDUX/MATCH", r"^(VOTE|REDUX|MATCH|CREDUX)"), ("BAR/WARPSYNC", r"^(BAR|WARPSYNC|BSYNC|BSSY|BREAK|NANOSLEEP)"),
           ("BRA/CALL/RET", r"^(BRA|BRX|CALL|RET|EXIT|JMP)"), ("FP32", r"^(FADD|FMUL|FFMA|FMNMX|FSEL|FSET|FSETP|FCHK|FRND)"), ("FP64", r"^D(ADD|MUL|FMA|SETP)"),
           ("CVT", r"^(F2F|F2I|I2F|I2I|F2FP|I2FP)"), ("INT/LOGIC", r"^(IADD|IMAD|LOP|SHF|LEA|ISETP|SEL|IABS|POPC|FLO|BREV|PRMT|VIMNMX|VIADD|IMNMX|PLOP|P2R|R2P|SGXT|BMSK)"),
           ("MOV/S2R", r"^(MOV|S2R|S2UR|CS2R|R2UR|UMOV)"), ("UNIFORM", r"^U")]
fn = None
stats = collections.OrderedDict()
for line in txt.split("\n"):
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1)
        stats[fn] = collections.Counter()
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and fn:
        op = m.group(3)
        stats[fn]["_n"] += 1
        if m.group(2):
            stats[fn]["_pred"] += 1
        for name, pat in CLASSES:
            if re.match(pat, op):
                stats[fn][name] += 1
                break
        else:
            stats[fn]["other"] += 1
for fn, c in stats.items():
    n = c["_n"]
    if n < 50:
        continue
    short = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip()[:70]
    print(f"{short}: {n} instructions, {n * 16 / 1024:.1f} KB, predicated {c['_pred'] / n * 100:.0f} %")
    print("   " + "  ".join(f"{k} {v / n * 100:.1f}%" for k, v in c.most_common() if not k.startswith("_")))
