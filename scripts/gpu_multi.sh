#!/bin/bash
# N-GPU bench the way the driver launches it, plus the NCCL sharding parity test
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 4 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"
tail -c 1200 gpurun_out/bench_n$N.json; tail -5 gpurun_out/bench_n$N.err
timeout 900 python -m pytest tests/test_sharding_gpu.py -x -q > gpurun_out/shard_n$N.log 2>&1; echo "shard test rc=$?"; tail -5 gpurun_out/shard_n$N.log
