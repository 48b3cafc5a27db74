#!/bin/bash
TAG=${1:?usage: run_evidence.sh TAG   (e.g. r01l) -- GPU tests, both bench arms, ncu launch list, full captures of the trace and shade kernels, config table}
# Evidence run of one round: writes gpurun_out/${TAG}_* (copy what is to be judged into profiles/, then tools/summarise_launches.py)
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; tail -3 gpurun_out/${TAG}_pytest_gpu.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref_c2.json 2> gpurun_out/${TAG}_bench_ref.err
python bench.py > gpurun_out/${TAG}_bench_c2.json 2> gpurun_out/${TAG}_bench.err; rc=$?
cat gpurun_out/${TAG}_bench_c2.json | cut -c1-600
if [ $rc -eq 0 ]; then
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1200 --csv \
      --log-file gpurun_out/${TAG}_launches_bench_c2.csv python bench.py --steps 2 --warmup 1 > gpurun_out/${TAG}_ncu_launch.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:trace_closest_persistent -s 10 -c 1 -f -o gpurun_out/${TAG}_trace_c2_full \
      python bench.py --steps 1 --warmup 1 > gpurun_out/${TAG}_ncu_full_c2.log 2>&1
fi
python tools/bench_trace.py --cases soupA,soupB,c2,c3 --reps 5 > gpurun_out/${TAG}_bench_trace.jsonl 2> gpurun_out/${TAG}_bench_trace.err
python tools/bench_trace.py --cases soupA,soupB,c2,c3 --reps 5 --builder 1 >> gpurun_out/${TAG}_bench_trace.jsonl 2>> gpurun_out/${TAG}_bench_trace.err
ncu --set full --clock-control none --import-source on -k regex:trace_closest_persistent_kernel -s 2 -c 1 -f -o gpurun_out/${TAG}_trace_full \
    python tools/bench_trace.py --cases soupA --reps 1 > gpurun_out/${TAG}_ncu_full.log 2>&1
cut -c1-300 gpurun_out/${TAG}_bench_trace.jsonl
ls -la gpurun_out/*.ncu-rep
ncu --set full --clock-control none --import-source on -k regex:wf_shade_kernel -s 3 -c 1 -f -o gpurun_out/${TAG}_shade_c2_full \
    python bench.py --steps 1 --warmup 1 > gpurun_out/${TAG}_ncu_shade.log 2>&1
python tools/run_configs.py > gpurun_out/${TAG}_configs.jsonl 2> gpurun_out/${TAG}_configs.err; cut -c1-250 gpurun_out/${TAG}_configs.jsonl
