#!/usr/bin/env python
"""Aggregate an ncu report's per-SASS-instruction counters by CUDA source line / function region.
usage: ncu_lines.py report.ncu-rep [top_n]"""
import csv, subprocess, sys, re, io
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr = None; out = []
for r in rows:
    if len(r) > 5 and r[0] == 'Line No': hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0] != '':
        try:
            out.append((int(r[0]), int(r[hdr.index('Instructions Executed')]), int(r[hdr.index('# Samples')]), r[1][:130]))
        except ValueError:
            continue
tot = sum(o[1] for o in out); tots = sum(o[2] for o in out)
print('total inst', tot, 'samples', tots)
# regions = device functions, found from the source text
src = open('gaussian-ray-tracing_b200/csrc/trace.cu').read().split('\n')
starts = []
for i, l in enumerate(src, 1):
    m = re.match(r'(?:template <[^>]*>\s*)?(?:__device__|__global__).*?\b(\w+)\s*\(', l)
    if m and not l.startswith(' '): starts.append((i, m.group(1)))
starts.append((10**9, 'end'))
reg = {}
for line, inst, samp, _ in out:
    name = 'pre'
    for (a, n), (b, _) in zip(starts, starts[1:]):
        if a <= line < b: name = n; break
    d = reg.setdefault(name, [0, 0]); d[0] += inst; d[1] += samp
for n, (i, s) in sorted(reg.items(), key=lambda kv: -kv[1][0]):
    print(f'{n:24s} inst {i/tot*100:5.1f}%  samples {s/tots*100:5.1f}%')
print()
for o in sorted(out, key=lambda o: -o[1])[:top]:
    print(o[0], f'{o[1]/tot*100:.2f}%', f'{o[2]/tots*100:.2f}%', o[3])
