#!/bin/bash
# 2-GPU run exactly as the driver launches it (torchrun, NCCL): weak-scaling line + strong_c4 leg
mkdir -p gpurun_out
nvidia-smi --query-gpu=name --format=csv > gpurun_out/gpu2.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "rc=$?" >> gpurun_out/bench_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err
tail -5 gpurun_out/bench_n2.err; head -c 2500 gpurun_out/bench_n2.json; echo; python - <<'PY'
import json
b=json.loads(open('gpurun_out/bench_n2.json').read().strip().splitlines()[-1])
print(json.dumps(b.get('strong_c4'),indent=1))
print(b['value'], b['e2e'])
PY
