#!/usr/bin/env python
"""Static code size of one kernel instantiation, attributed to the source function each SASS instruction's line belongs to
(nvdisasm --print-line-info of the cubin inside trace.o).  Shows which inlined helpers make a kernel large.
usage: sass_by_function.py [mangled-substring, default ILb0ELb1E = trace_kernel<false,true>]"""
import os, re, subprocess, sys, tempfile, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
want = sys.argv[1] if len(sys.argv) > 1 else "trace_kernelILb0ELb1E"
src = open(os.path.join(ROOT, "gaussian-ray-tracing_b200/csrc/trace.cu")).read().split("\n")
starts = []
for i, l in enumerate(src, 1):
    m = re.match(r'(?:template <[^>]*>\s*)?(?:__device__|__global__).*?\b(\w+)\s*\(', l)
    if m and not l.startswith(' '):
        starts.append((i, m.group(1)))
starts.append((10**9, 'end'))
def fn_of(line):
    name = 'pre'
    for (a, n), (b, _) in zip(starts, starts[1:]):
        if a <= line < b:
            return n
    return name
with tempfile.TemporaryDirectory() as d:
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "gaussian-ray-tracing_b200/csrc/trace.o")], cwd=d, capture_output=True)
    cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(d, cub)], capture_output=True, text=True).stdout
cur, on, fname = None, False, None
cnt = collections.Counter(); lines = collections.Counter()
for l in txt.split("\n"):
    m = re.match(r'\.text\.(\S+):', l)
    if m:
        on = want in m.group(1); continue
    if l.startswith("//-----") or l.startswith("\t.section"):
        if ".text." in l and want not in l: on = False
        continue
    if not on: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        f = os.path.basename(m.group(1))
        cur = (f, int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/', l):
        if cur and cur[0] == "trace.cu":
            fn = fn_of(cur[1]); lines[cur[1]] += 1
        else:
            fn = "<" + (cur[0] if cur else "?") + ">"
        cnt[fn] += 1
tot = sum(cnt.values())
print(f"{want}: {tot} instructions, {tot * 16 / 1024:.1f} KB")
for n, c in cnt.most_common(24):
    print(f"  {n:28s} {c:6d}  {c / tot * 100:5.1f} %")
print("largest source lines:")
for ln, c in lines.most_common(14):
    print(f"  {ln:5d} {c:5d}  {src[ln - 1].strip()[:110]}")
