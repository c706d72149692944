#!/bin/bash
# end of round: C1 / C3 / C4 / C5 lines of the final code (this library only), C4 bench line
mkdir -p gpurun_out
timeout 200 python scripts/run_configs.py --ours-only > gpurun_out/r2h_configs.jsonl 2> gpurun_out/r2h_configs.err; echo rc=$?; cut -c1-330 gpurun_out/r2h_configs.jsonl
timeout 150 python bench.py --scene terrain --spp 256 --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/r2h_bench_c4.json 2> gpurun_out/r2h_bench_c4.err; echo rc=$?; cut -c1-300 gpurun_out/r2h_bench_c4.json
timeout 60 python scripts/c5.py | cut -c1-500
