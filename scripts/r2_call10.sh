#!/bin/bash
mkdir -p gpurun_out
V=raytracer-cuda_b200/build/var
timeout 900 bash scripts/ab2.sh "cur;;$V/libcur2.so" "prim-pipe;;$V/libpp1.so" "shade-noprefetch;;$V/libnopf.so" > gpurun_out/c10_ab.log 2>&1
cat gpurun_out/c10_ab.log
for lib in $V/libcur2.so $V/libpp1.so; do echo "== spheres mega $lib"; CRT_LIB=$PWD/$lib python scripts/prof_render.py --scene spheres --width 1200 --height 800 --spp 500 --reps 2 --mode megakernel | grep -E "rep 1"; done
