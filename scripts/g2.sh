set -x
B="python bench.py --spp 16 --steps 2 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r1_bench_wf16.json 2> gpurun_out/r1_bench_wf16.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r1_launches_wf16.csv $B > gpurun_out/ncu_l.log 2>&1
cat gpurun_out/r1_bench_wf16.json
P="python scripts/prof_render.py --spp 4"
$P > gpurun_out/plain_wf.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_wf_extend -s 12 -c 2 -o gpurun_out/r1_extend $P > gpurun_out/ncu_e.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 12 -c 1 -o gpurun_out/r1_shade $P > gpurun_out/ncu_s.log 2>&1
cat gpurun_out/plain_wf.log
P="python scripts/prof_render.py --spp 4 --mode megakernel"
$P > gpurun_out/plain_mega.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_render_mega -c 1 -o gpurun_out/r1_mega $P > gpurun_out/ncu_m.log 2>&1
cat gpurun_out/plain_mega.log
python scripts/prof_render.py --spp 8 --counted --reps 2
python scripts/prof_render.py --spp 64 --reps 2
python scripts/prof_render.py --spp 64 --reps 2 --mode megakernel
oracle/_ref/ref_render_b20 --scene cornell_blob --width 2560 --height 1440 --spp 64 --frames 2 --warmup 1 --twin
ls -la gpurun_out
