#!/bin/bash
# N-GPU run exactly as the driver launches it (torchrun, NCCL): weak-scaling line + strong_c4 leg
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name --format=csv > gpurun_out/gpu$N.txt
nvidia-smi topo -m > gpurun_out/topo$N.txt 2>&1
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 $BENCH_EXTRA > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "rc=$?" >> gpurun_out/bench_n$N.err
tail -3 gpurun_out/bench_n$N.err; python - <<PY
import json
b=json.loads(open('gpurun_out/bench_n$N.json').read().strip().splitlines()[-1])
print(json.dumps(b.get('strong_c4'),indent=1))
print(b['value'], b['ms_per_step'], json.dumps(b['e2e']))
PY
# the same weak-scaling line with the upload buffers in write-combined page-locked memory
VO_B200_PINNED=wc timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --no-configs --no-cpu > gpurun_out/bench_n${N}_wc.json 2> gpurun_out/bench_n${N}_wc.err; echo "rc=$?" >> gpurun_out/bench_n${N}_wc.err
tail -2 gpurun_out/bench_n${N}_wc.err; python - <<PY
import json
for f in ('gpurun_out/bench_n$N.json', 'gpurun_out/bench_n${N}_wc.json'):
    b=json.loads(open(f).read().strip().splitlines()[-1])
    print(f, b['value'], b['ms_per_step'], json.dumps(b['e2e']), json.dumps(b.get('per_rank')))
PY
