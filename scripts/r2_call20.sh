#!/bin/bash
# L1 carve-out preference of extend / shade, and ld.global.cg for the path state (L1 bypass)
mkdir -p gpurun_out
bash scripts/ab2.sh "base;;raytracer-cuda_b200/libcrt_b200.so" "carve16;CRT_EXTEND_CARVEOUT=16;raytracer-cuda_b200/libcrt_b200.so" "carve28;CRT_EXTEND_CARVEOUT=28;raytracer-cuda_b200/libcrt_b200.so" "carve44;CRT_EXTEND_CARVEOUT=44;raytracer-cuda_b200/libcrt_b200.so" "carve100;CRT_EXTEND_CARVEOUT=100;raytracer-cuda_b200/libcrt_b200.so" "ldcg1;;raytracer-cuda_b200/build/var/libldcg1.so" "ldcg2;;raytracer-cuda_b200/build/var/libldcg2.so" "ldcg1+carve16;CRT_EXTEND_CARVEOUT=16;raytracer-cuda_b200/build/var/libldcg1.so" > gpurun_out/d8_ab.log 2>&1
cat gpurun_out/d8_ab.log
