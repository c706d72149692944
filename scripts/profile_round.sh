#!/bin/bash
# ncu captures for profiles/ (run under gpurun, 1 GPU): usage: bash scripts/profile_round.sh <tag>
TAG=${1:-r1}
set -x
B="python bench.py --spp 32 --steps 2 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/${TAG}_bench_spp32.json 2> gpurun_out/${TAG}_bench_spp32.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/${TAG}_launches.csv $B > gpurun_out/ncu_l.log 2>&1
P="python scripts/prof_render.py --spp 64"
$P > gpurun_out/${TAG}_plain_wf.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_wf_extend -s 100 -c 1 -o gpurun_out/${TAG}_extend $P > gpurun_out/ncu_e.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 100 -c 1 -o gpurun_out/${TAG}_shade $P > gpurun_out/ncu_s.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wf_tail -c 1 -o gpurun_out/${TAG}_tail $P > gpurun_out/ncu_t.log 2>&1
P="python scripts/prof_render.py --spp 8 --mode megakernel"
$P > gpurun_out/${TAG}_plain_mega.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_render_mega -c 1 -o gpurun_out/${TAG}_mega $P > gpurun_out/ncu_m.log 2>&1
cat gpurun_out/${TAG}_plain_wf.log gpurun_out/${TAG}_plain_mega.log
