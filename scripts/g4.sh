set -x
P="python scripts/prof_render.py --spp 4"
$P > gpurun_out/plain_wf.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_wf_extend -s 12 -c 1 -o gpurun_out/r1b_extend $P > gpurun_out/ncu_e.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 12 -c 1 -o gpurun_out/r1b_shade $P > gpurun_out/ncu_s.log 2>&1
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r1b_bench_2000.json 2> gpurun_out/r1b_bench_2000.err; cat gpurun_out/r1b_bench_2000.json
