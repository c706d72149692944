#!/bin/bash
# no-alloc entry point: its test alone in a fresh process, then the whole GPU suite (it is collected last), smoke, short bench
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -5 ) > gpurun_out/d1_noalloc_alone.log; cat gpurun_out/d1_noalloc_alone.log
( timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 2>&1 | tail -16 ) > gpurun_out/d1_gpu_tests.log; cat gpurun_out/d1_gpu_tests.log
python __graft_entry__.py smoke 2>&1 | tail -2
python bench.py --steps 1 --warmup 3 > gpurun_out/d1_bench_n1.json 2> gpurun_out/d1_bench_n1.err; echo rc=$?; tail -2 gpurun_out/d1_bench_n1.err; cat gpurun_out/d1_bench_n1.json
