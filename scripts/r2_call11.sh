#!/bin/bash
mkdir -p gpurun_out
L=raytracer-cuda_b200/libcrt_b200.so
timeout 900 bash scripts/ab2.sh "default;;$L" "parts1;CRT_WF_PARTS=1;$L" "parts3;CRT_WF_PARTS=3;$L" "drain2;CRT_DRAIN_ROUNDS=2;$L" "drain8;CRT_DRAIN_ROUNDS=8;$L" "batch8;CRT_WF_BATCH=8;$L" > gpurun_out/c11_ab.log 2>&1
cat gpurun_out/c11_ab.log
echo "== megakernel C2 128 spp"; python scripts/prof_render.py --spp 128 --reps 2 --mode megakernel | grep "rep 1"
echo "== compute-sanitizer memcheck (small frames, wavefront incl. tail, bulk shade, megakernel)"
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python scripts/prof_render.py --width 320 --height 180 --spp 3 > gpurun_out/c11_memcheck_wf.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/c11_memcheck_wf.log
CRT_SHADE_BULK=1 CRT_TAIL_LIVE=0 timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python scripts/prof_render.py --width 320 --height 180 --spp 3 > gpurun_out/c11_memcheck_bulk.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/c11_memcheck_bulk.log
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python scripts/prof_render.py --width 160 --height 90 --spp 2 --mode megakernel --scene terrain --param 100 > gpurun_out/c11_memcheck_mega.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/c11_memcheck_mega.log
timeout 600 compute-sanitizer --tool racecheck --error-exitcode 9 python scripts/prof_render.py --width 160 --height 90 --spp 2 > gpurun_out/c11_racecheck_wf.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/c11_racecheck_wf.log
