#!/bin/bash
# megakernel: traversal stack in local memory (-DCRT_MEGA_SMEM_STACK=0) against 4 entries in shared memory, every config
for lib in raytracer-cuda_b200/libcrt_b200.so raytracer-cuda_b200/build/var/libmstk0.so raytracer-cuda_b200/libcrt_b200.so raytracer-cuda_b200/build/var/libmstk0.so; do echo "== $lib"
  CRT_LIB=$PWD/$lib python scripts/prof_render.py --scene spheres --width 1200 --height 800 --spp 500 --reps 2 --mode megakernel | grep -E "rep 1" | cut -c1-60
  CRT_LIB=$PWD/$lib python scripts/prof_render.py --scene cornell --width 256 --height 256 --spp 16 --bounces 8 --reps 5 --mode megakernel | grep -E "rep 4" | cut -c1-60
  CRT_LIB=$PWD/$lib python scripts/prof_render.py --width 1920 --height 1080 --spp 1 --bounces 8 --reps 20 --mode megakernel | grep -E "rep 19" | cut -c1-60
  CRT_LIB=$PWD/$lib python scripts/prof_render.py --scene terrain --spp 16 --reps 2 --mode megakernel | grep -E "rep 1" | cut -c1-60
done
