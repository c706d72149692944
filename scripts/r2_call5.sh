#!/bin/bash
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -q -s 2>&1 | tail -30 ) > gpurun_out/c5_pytest.log 2>&1
tail -8 gpurun_out/c5_pytest.log
timeout 900 python scripts/parity_report.py > gpurun_out/r2_parity_report.jsonl 2> gpurun_out/parity.err; tail -2 gpurun_out/parity.err; cut -c1-330 gpurun_out/r2_parity_report.jsonl
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo rc=$?; tail -2 gpurun_out/r2_bench_n1.err; cut -c1-700 gpurun_out/r2_bench_n1.json
bash scripts/profile_round.sh r2 > gpurun_out/r2_profile.log 2>&1; tail -12 gpurun_out/r2_profile.log
