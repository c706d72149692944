#!/bin/bash
# N = 8 bench line of the final code (driver's launch style), bounded by timeout
mkdir -p gpurun_out
N=${1:-8}
CRT_BENCH_WATCHDOG_S=120 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/r2g_bench_n$N.json 2> gpurun_out/r2g_bench_n$N.err; echo rc=$?; tail -3 gpurun_out/r2g_bench_n$N.err | cut -c1-300; python -c "
import json; d=json.load(open('gpurun_out/r2g_bench_n$N.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], json.dumps(d['check'])[:900])"
