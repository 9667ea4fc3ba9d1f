#!/usr/bin/env bash
# s66: the bench line at N = 2 on the final build, launched as the driver launches it (configs block incl. the sequence-sharded step
# with peer-memory and NCCL exchanges across two GPUs)
set -u
mkdir -p gpurun_out
( time timeout 60 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 30 --warmup 5 --no-cpu-baseline ) > gpurun_out/r2s66_n2.json 2> gpurun_out/r2s66_n2.err; echo "bench N=2 rc=$?"
tail -c 400 gpurun_out/r2s66_n2.json; grep -c "NCCL" gpurun_out/r2s66_n2.err; tail -3 gpurun_out/r2s66_n2.err
