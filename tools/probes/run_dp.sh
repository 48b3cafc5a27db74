python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_dp.log 2>&1; tail -3 gpurun_out/pytest_gpu_dp.log
python tools/probes/node_stats.py 2>&1 | cut -c1-250
summ() { python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(d['case'], d.get('integrator',''), round(d.get('mrays_s',0)), 'n/r', round(d.get('nodes_per_ray',0),2), 't/r', round(d.get('tris_per_ray',0),2), d.get('nodes',''), d.get('build_ms',''))
"; }
for nc in 1.5 1.0 2.5; do for scst in 0.25 0.5; do
echo "== node_cost $nc slot_cost $scst"
FF_NODE_COST=$nc FF_SLOT_COST=$scst python tools/bench_trace.py --cases soupA,soupB,c2,c3 --reps 3 2>&1 | summ
done; done
echo "== SAH builder default costs"
python tools/bench_trace.py --cases soupA,soupB,c2,c3 --reps 3 --builder 1 2>&1 | summ
