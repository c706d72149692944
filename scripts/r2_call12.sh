#!/bin/bash
mkdir -p gpurun_out
V=raytracer-cuda_b200/build/var; L=raytracer-cuda_b200/libcrt_b200.so
( timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -3 )
for lib in $L $V/libmstk0.so $L $V/libmstk0.so; do echo "== $lib"
  CRT_LIB=$PWD/$lib python scripts/prof_render.py --scene spheres --width 1200 --height 800 --spp 500 --reps 2 --mode megakernel | grep -E "rep 1"
  CRT_LIB=$PWD/$lib python scripts/prof_render.py --scene cornell --width 256 --height 256 --spp 16 --bounces 8 --reps 5 --mode megakernel | grep -E "rep 4"
  CRT_LIB=$PWD/$lib python scripts/prof_render.py --spp 64 --reps 2 --mode megakernel | grep -E "rep 1"
  CRT_LIB=$PWD/$lib python scripts/prof_render.py --scene terrain --spp 16 --reps 2 --mode megakernel | grep -E "rep 1"
done
for lib in $L $V/libcur2.so $L $V/libcur2.so; do echo "== tail (2000 spp frame) $lib"; CRT_LIB=$PWD/$lib python scripts/prof_render.py --spp 2000 --reps 1 --time-kernels | grep -E "rep 0|extend"; done
