#!/bin/bash
# PLOC's last rounds in one launch (k_ploc_small): GPU suite (ids bit-exact on every scene), build times before / after
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q --durations=4 2>&1 | tail -8 )
for s in cornell_blob terrain spheres cornell; do echo "== $s"; for i in 1 2; do CRT_BUILD_TRACE_ROUNDS=1 python scripts/prof_render.py --scene $s --spp 2 --width 640 --height 360 2>&1 | grep -E "bvh:|one launch"; done; done
