#!/bin/bash
# usage: run_scale.sh N TAG  -- C2 weak-scaling bench + C5 strong-scaling frame + multi-GPU parity test on N GPUs
N=$1; TAG=$2; SPP=${3:-$((128*N))}
if [ "$N" -gt 1 ]; then
  python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -2 | tee gpurun_out/${TAG}_pytest_multi_n$N.log
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_c2_n$N.json 2> gpurun_out/${TAG}_bench_c2_n$N.err
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 tools/bench_c5.py --spp $SPP > gpurun_out/${TAG}_c5_n$N.json 2> gpurun_out/${TAG}_c5_n$N.err
else
  python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_c2_n1.json 2> gpurun_out/${TAG}_bench_c2_n1.err
  python tools/bench_c5.py --spp $SPP > gpurun_out/${TAG}_c5_n1.json 2> gpurun_out/${TAG}_c5_n1.err
fi
cut -c1-330 gpurun_out/${TAG}_bench_c2_n$N.json; echo; cat gpurun_out/${TAG}_c5_n$N.json | cut -c1-700
