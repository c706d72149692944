#!/bin/bash
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests -m gpu -x -q -s 2>&1 | tail -25 ) > gpurun_out/c3_pytest.log 2>&1
tail -6 gpurun_out/c3_pytest.log
V=raytracer-cuda_b200/build/var; L=raytracer-cuda_b200/libcrt_b200.so
timeout 900 bash scripts/ab2.sh "bulk1@5;;$L" "nobulk@5;CRT_SHADE_BULK=0;$L" "bulk1@4;;$V/libsh4.so" "nobulk@4;CRT_SHADE_BULK=0;$V/libsh4.so" "bulk1@6;;$V/libsh6.so" > gpurun_out/c3_ab.log 2>&1
cat gpurun_out/c3_ab.log
timeout 600 python bench.py --spp 64 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/c3_bench64.json 2> gpurun_out/c3_bench64.err; echo rc=$?; tail -3 gpurun_out/c3_bench64.err; cut -c1-1500 gpurun_out/c3_bench64.json
P="python scripts/prof_render.py --spp 64"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade_bulk -s 100 -c 1 -f -o gpurun_out/c3_shade_bulk $P > gpurun_out/ncu_sb.log 2>&1
CRT_SHADE_BULK=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 100 -c 1 -f -o gpurun_out/c3_shade_plain $P > gpurun_out/ncu_sp.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_extend -s 100 -c 1 -f -o gpurun_out/c3_extend $P > gpurun_out/ncu_e.log 2>&1
ls -la gpurun_out/*.ncu-rep
