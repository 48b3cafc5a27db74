#!/bin/bash
# usage: run_r02_multi.sh N TAG [STEPS]  -- multi-GPU parity tests + the C5 strong-split bench on N GPUs (both reduce paths)
N=$1; TAG=$2; STEPS=${3:-5}
set -x
nvidia-smi topo -m > gpurun_out/${TAG}_topo_n$N.txt 2>&1
python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/${TAG}_pytest_multi_n$N.log 2>&1; tail -15 gpurun_out/${TAG}_pytest_multi_n$N.log
for R in peer nccl; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps $STEPS --warmup 3 --reduce $R \
   > gpurun_out/${TAG}_bench_c5_n${N}_$R.json 2> gpurun_out/${TAG}_bench_c5_n${N}_$R.err
tail -4 gpurun_out/${TAG}_bench_c5_n${N}_$R.err; cut -c1-2500 gpurun_out/${TAG}_bench_c5_n${N}_$R.json
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref_n$N.json 2> gpurun_out/${TAG}_bench_ref_n$N.err
cut -c1-600 gpurun_out/${TAG}_bench_ref_n$N.json
