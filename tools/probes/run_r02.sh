#!/bin/bash
TAG=${1:?usage: run_r02.sh TAG  -- one-GPU evidence run of round 2: GPU tests, both bench arms, ncu launch list + full capture of the C4 trace kernel}
set -x
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/${TAG}_pytest_no_alloc_alone.log 2>&1; tail -3 gpurun_out/${TAG}_pytest_no_alloc_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; tail -5 gpurun_out/${TAG}_pytest_gpu.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; cut -c1-400 gpurun_out/${TAG}_bench_ref.json
python bench.py ${BENCH_ARGS} > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; rc=$?
tail -5 gpurun_out/${TAG}_bench.err; cut -c1-1500 gpurun_out/${TAG}_bench.json
if [ $rc -eq 0 ] && [ -z "$NO_NCU" ]; then
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1500 --csv \
      --log-file gpurun_out/${TAG}_launches_bench.csv env FF_TRACE_STREAM=0 python bench.py --steps 2 --warmup 1 --skip c2,c3,c5,cpu > gpurun_out/${TAG}_ncu_launch.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:trace_closest_persistent -s 1 -c 1 -f -o gpurun_out/${TAG}_trace_c4_full \
      python tools/bench_trace.py --cases soupA --reps 1 > gpurun_out/${TAG}_ncu_full.log 2>&1
  ls -la gpurun_out/*.ncu-rep
fi
python tools/bench_trace.py --cases soupA,soupB,c2,c3 --reps 5 > gpurun_out/${TAG}_bench_trace.jsonl 2> gpurun_out/${TAG}_bench_trace.err
cut -c1-300 gpurun_out/${TAG}_bench_trace.jsonl
