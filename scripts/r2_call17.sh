#!/bin/bash
# two GPUs after the tail fix: the 2-GPU tests, then the N=2 bench line (both arms' launch style)
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_multi.py tests/test_cli.py -m gpu -q 2>&1 | tail -3 )
CRT_BENCH_WATCHDOG_S=120 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2f_bench_n2.json 2> gpurun_out/r2f_bench_n2.err; echo rc=$?; tail -3 gpurun_out/r2f_bench_n2.err | cut -c1-300; python -c "
import json; d=json.load(open('gpurun_out/r2f_bench_n2.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], json.dumps(d['check'])[:900])"
