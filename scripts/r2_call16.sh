#!/bin/bash
# after the tail-stack fix: GPU suite, full bench (10 frames), 14-frame soak, reference arm
mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 )
CRT_BENCH_WATCHDOG_S=120 timeout 400 python bench.py --steps 3 --warmup 3 > gpurun_out/r2f_bench_n1.json 2> gpurun_out/r2f_bench_n1.err; echo rc=$?; tail -3 gpurun_out/r2f_bench_n1.err; cut -c1-500 gpurun_out/r2f_bench_n1.json
timeout 120 python scripts/hang_repro.py 14 > gpurun_out/d4_soak.log 2>&1; echo rc=$?; tail -4 gpurun_out/d4_soak.log
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2f_bench_ref.json 2>/dev/null; cut -c1-300 gpurun_out/r2f_bench_ref.json
