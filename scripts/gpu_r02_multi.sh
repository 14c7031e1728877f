#!/bin/bash
# N-GPU evidence: NCCL sharding parity (pytest -s keeps the per-world summary lines) and the bench
# line at N ranks (boundary exchange + e2e legs inside).  SKIP_SHARD=1 runs the bench only.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/gpus_n$N.txt
[ -n "$SKIP_SHARD" ] || timeout 1500 python -m pytest tests/test_sharding_gpu.py -m gpu -q -s --timeout 900 > gpurun_out/shard_nccl_n$N.log 2>&1; echo "shard pytest rc=$?" | tee -a gpurun_out/shard_nccl_n$N.log
[ -n "$SKIP_SHARD" ] || grep -E "sharding parity|passed|skipped|failed" gpurun_out/shard_nccl_n$N.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
  bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_c5_n$N.json 2> gpurun_out/bench_c5_n$N.err; echo "bench n=$N rc=$?"
tail -3 gpurun_out/bench_c5_n$N.err | cut -c1-300
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_c5_n$N.json') if l.startswith('{')][-1])
print('n',d['n_gpus'],'ms/step',round(d['ms_per_step'],2),'value',d['value'])
for k in ('e2e','e2e_u8f16','h2d_ceiling','boundary','e2e_api'): print(k, json.dumps(d.get(k))[:700])
PY
