#!/bin/bash
# final single-GPU numbers of the round: tests, bench, all configs, parity report, ncu captures
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -4 ) > gpurun_out/c9_pytest.log 2>&1; tail -3 gpurun_out/c9_pytest.log
python __graft_entry__.py smoke 2>&1 | tail -3
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo rc=$?; tail -2 gpurun_out/r2_bench_n1.err; cut -c1-400 gpurun_out/r2_bench_n1.json
timeout 900 python bench.py --scene terrain --spp 256 --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/r2_bench_c4.json 2> gpurun_out/r2_bench_c4.err; cut -c1-300 gpurun_out/r2_bench_c4.json
timeout 900 python scripts/run_configs.py --ours-only > gpurun_out/r2_configs.jsonl 2> gpurun_out/r2_configs.err; cut -c1-260 gpurun_out/r2_configs.jsonl
timeout 900 python scripts/parity_report.py > gpurun_out/r2_parity_report.jsonl 2> gpurun_out/parity.err; cut -c1-330 gpurun_out/r2_parity_report.jsonl | grep -c '"radiance_bit_exact": 1.0'
bash scripts/profile_round.sh r2 > gpurun_out/r2_profile.log 2>&1; tail -10 gpurun_out/r2_profile.log
CRT_SHADE_BULK=1 CRT_WF_GRAPH=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade_bulk -s 100 -c 1 -f -o gpurun_out/r2_shade_bulk python scripts/prof_render.py --spp 64 > gpurun_out/ncu_sb.log 2>&1
