#!/bin/bash
# BVH quality knobs: PLOC search radius, primitive / node cost ratio of the collapse
mkdir -p gpurun_out
bash scripts/ab2.sh "base;;raytracer-cuda_b200/libcrt_b200.so" "r8;;raytracer-cuda_b200/build/var/libr8.so" "r48;;raytracer-cuda_b200/build/var/libr48.so" "r100;;raytracer-cuda_b200/build/var/libr100.so" "cp04;;raytracer-cuda_b200/build/var/libcp04.so" "cp10;;raytracer-cuda_b200/build/var/libcp10.so" "cp15;;raytracer-cuda_b200/build/var/libcp15.so" > gpurun_out/d10_ab.log 2>&1
for l in raytracer-cuda_b200/libcrt_b200.so raytracer-cuda_b200/build/var/libr8.so raytracer-cuda_b200/build/var/libr48.so raytracer-cuda_b200/build/var/libr100.so raytracer-cuda_b200/build/var/libcp04.so raytracer-cuda_b200/build/var/libcp10.so raytracer-cuda_b200/build/var/libcp15.so; do echo "== counted $l"; CRT_LIB=$PWD/$l python scripts/prof_render.py --spp 4 --counted | grep -E "rep 0|bvh:" ; CRT_LIB=$PWD/$l python scripts/prof_render.py --scene spheres --width 1200 --height 800 --spp 100 --reps 2 --mode megakernel | grep -E "rep 1"; done >> gpurun_out/d10_ab.log 2>&1
cat gpurun_out/d10_ab.log
