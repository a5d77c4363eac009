#!/bin/bash
# N-GPU bench exactly as the driver launches it (torchrun, NCCL): weak-scaling line + strong_c4 leg
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo$N.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "rc=$?" >> gpurun_out/bench_n$N.err
tail -2 gpurun_out/bench_n$N.err
python - <<PY
import json
b=json.loads(open("gpurun_out/bench_n$N.json").read().strip().splitlines()[-1])
print(json.dumps(b.get("strong_c4"),indent=1))
print(b["value"], b["ms_per_step"], json.dumps(b["e2e"]), json.dumps(b.get("per_rank")))
PY
