set -x
P="python scripts/prof_render.py --spp 64"
$P > gpurun_out/plain_wf.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 100 -c 1 -o gpurun_out/r1f_shade $P > gpurun_out/ncu_s.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wf_extend -s 100 -c 1 -o gpurun_out/r1f_extend $P > gpurun_out/ncu_e.log 2>&1
