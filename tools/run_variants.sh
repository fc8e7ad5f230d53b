#!/bin/bash
# Run tools/prof_trace.py for every tuning build present; one JSON line per variant into gpurun_out/variants.jsonl
mkdir -p gpurun_out
: > gpurun_out/variants.jsonl
for so in gaussian-ray-tracing_b200/csrc/libgrt.so gaussian-ray-tracing_b200/csrc/variants/libgrt_*.so; do
  echo -n "{\"lib\": \"$(basename $so)\", \"r\": " >> gpurun_out/variants.jsonl
  GRT_LIBGRT=$PWD/$so python tools/prof_trace.py "$@" 2>/dev/null | tail -1 | tr -d "\n" >> gpurun_out/variants.jsonl
  echo "}" >> gpurun_out/variants.jsonl
done
python - <<'PY'
import json
for l in open('gpurun_out/variants.jsonl'):
    try:
        d=json.loads(l); r=d['r']; c=r.get('counters',{}); pw=r.get('per_warp',{})
        print(f"{d['lib']:32s} ms {min(r['launch_ms']):7.3f}  nodes/w {pw.get('warp_nodes',0):6.1f} prims/w {pw.get('warp_leaf_prims',0):6.1f} hit/w {c.get('warp_prims_hit',0)/64800:6.1f} iters/w {c.get('warp_consume_iters',0)/64800:6.1f} repass {c.get('lanes_repass',0)} dfs {c.get('warp_dfs_fallbacks',0)}")
    except Exception as ex:
        print('bad line', l[:100], ex)
PY
