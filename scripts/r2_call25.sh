#!/bin/bash
# BVH build temporaries from one arena: GPU suite, build times with / without (CRT_BUILD_NO_ARENA=1), host time inside cudaMalloc / cudaFree
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 )
for s in cornell_blob terrain spheres cornell; do for e in CRT_NOP=1 CRT_BUILD_NO_ARENA=1; do echo "== $s $e"; for i in 1 2; do env $e CRT_BUILD_TRACE=1 python scripts/prof_render.py --scene $s --spp 2 --width 640 --height 360 2>&1 | grep -E "bvh:|bvh_build_trace" | cut -c1-220; done; done; done
