#!/bin/bash
# runs tools/bench_trace.py against every kernel-variant library under tools/probes/variants/
for f in tools/probes/variants/lib_*.so; do
  echo "== $f"
  FF_LIB_OVERRIDE=$PWD/$f timeout 200 python tools/bench_trace.py --cases ${CASES:-soupA,soupB,c2} --reps 3 ${EXTRA} 2>&1 | grep -v soup_build | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(d['case'], d.get('integrator',''), round(d['mrays_s']), 'n/r %.2f t/r %.2f' % (d.get('nodes_per_ray',0), d.get('tris_per_ray',0)), ('gpu_ms %.2f' % d['gpu_ms']) if 'gpu_ms' in d else '')
"
done
