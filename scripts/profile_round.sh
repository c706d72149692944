#!/bin/bash
# ncu captures for profiles/ (run under gpurun, 1 GPU): usage: bash scripts/profile_round.sh <tag>
# 1. launch list of the bench command (per-launch times, for kernel SHARES)   2. one ncu --set full capture per kernel
# (steady state: launch #100 of a 64-spp frame) with source import, on the headline scene and on the 5 M-triangle terrain
TAG=${1:-r2}
mkdir -p gpurun_out
set -x
B="python bench.py --spp 32 --steps 2 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/${TAG}_bench_spp32.json 2> gpurun_out/${TAG}_bench_spp32.err && \
CRT_WF_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/${TAG}_launches.csv $B > gpurun_out/ncu_l.log 2>&1
P="python scripts/prof_render.py --spp 64"
$P > gpurun_out/${TAG}_plain_wf.log 2>&1
for K in extend shade tail begin compact_scatter; do
  SKIP=100; [ $K = tail ] && SKIP=0; [ $K = begin ] && SKIP=0; [ $K = compact_scatter ] && SKIP=2
  CRT_WF_GRAPH=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_$K\|k_$K -s $SKIP -c 1 -f -o gpurun_out/${TAG}_$K $P > gpurun_out/ncu_$K.log 2>&1
done
CRT_WF_GRAPH=0 timeout 600 ncu --set full --clock-control none -k regex:k_rng_init -c 1 -f -o gpurun_out/${TAG}_rng_init $P > gpurun_out/ncu_rng.log 2>&1
PT="python scripts/prof_render.py --scene terrain --spp 16"
for K in extend shade; do
  CRT_WF_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_$K -s 60 -c 1 -f -o gpurun_out/${TAG}_terrain_$K $PT > gpurun_out/ncu_t_$K.log 2>&1
done
P="python scripts/prof_render.py --spp 8 --mode megakernel"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render_mega -c 1 -f -o gpurun_out/${TAG}_mega $P > gpurun_out/ncu_m.log 2>&1
ls -la gpurun_out/${TAG}_*.ncu-rep
