#!/bin/bash
# explicit __syncwarp at the tail loop's rendezvous: does the shared-stack tail still stall? perf of the default build unchanged?
mkdir -p gpurun_out
V=$PWD/raytracer-cuda_b200/build/var
echo "== tail4sync (shared-stack tail + explicit syncwarp)"; CRT_LIB=$V/libtail4sync.so timeout 60 python scripts/hang_repro.py 8 > gpurun_out/d6_tail4sync.log 2>&1; echo rc=$?; tail -3 gpurun_out/d6_tail4sync.log
echo "== default"; timeout 60 python scripts/hang_repro.py 8 > gpurun_out/d6_default.log 2>&1; echo rc=$?; tail -3 gpurun_out/d6_default.log
( timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 )
