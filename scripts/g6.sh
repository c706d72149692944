set -x
export CRT_COOP=1
P="python scripts/prof_render.py --spp 64"
$P > gpurun_out/plain_wf.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_wf_extend -s 100 -c 1 -o gpurun_out/r1e_extend_coop $P > gpurun_out/ncu_e.log 2>&1
cat gpurun_out/plain_wf.log
