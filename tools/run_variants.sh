#!/bin/bash
# Run tools/prof_trace.py for every tuning build present; one JSON line per variant into gpurun_out/variants<tag>.jsonl
# usage: tools/run_variants.sh [tag] [prof_trace args...]
tag=$1; shift
mkdir -p gpurun_out
out=gpurun_out/variants$tag.jsonl
: > $out
for so in gaussian-ray-tracing_b200/csrc/libgrt.so gaussian-ray-tracing_b200/csrc/variants/libgrt_*.so; do
  [ -f $so ] || continue
  echo -n "{\"lib\": \"$(basename $so)\", \"r\": " >> $out
  GRT_LIBGRT=$PWD/$so timeout 300 python tools/prof_trace.py --cache "$@" 2>gpurun_out/err_$(basename $so).txt | tail -1 | tr -d "\n" >> $out
  echo "}" >> $out
done
python - $out <<'PY'
import json, sys
for l in open(sys.argv[1]):
    try:
        d=json.loads(l); r=d['r']; c=r.get('counters',{}); pw=r.get('per_warp',{}); pr=r.get('per_ray',{})
        print(f"{d['lib']:30s} ms {min(r['launch_ms']):7.3f} nodes/w {pw.get('warp_nodes',0):6.1f} recs/w {pw.get('warp_leaf_prims',0):6.1f} loop/w {pw.get('warp_loop_iters',0):6.1f} "
              f"cand_it/w {pw.get('warp_candidate_iters',0):6.1f} cand/ray {pr.get('candidates',0):5.1f} rounds/w {pw.get('warp_composite_rounds',0):5.1f} shade_it/w {pw.get('warp_consume_iters',0):5.1f} "
              f"hg {pr.get('hits_gathered',0):5.2f} shifts/ins {pr.get('shift_steps',0)/max(pr.get('inserts',1),1e-9):4.2f} spills/ray {pr.get('spills',0):4.2f} forced/w {pw.get('warp_forced_flushes',0):4.2f} listfull/w {pw.get('warp_list_full',0):4.2f} repass {c.get('lanes_repass',0)} dfs {c.get('warp_dfs_fallbacks',0)} ovf {c.get('stack_overflow',0)} "
              f"h8 {r.get('rgb8_sha1')} hf {r.get('rgb32f_sha1')}")
    except Exception as ex:
        print('bad line', l[:100], ex)
PY
