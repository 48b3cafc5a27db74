summ() { python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    if d['case']=='soup_build': continue
    print(d['case'], d.get('integrator',''), round(d.get('mrays_s',0)))
"; }
for r in 12 16 20 24 28 31; do echo "== refill $r"; python tools/bench_trace.py --cases soupA,c2 --reps 3 --builder 1 --tune 0,$r,0 2>&1 | summ; done
