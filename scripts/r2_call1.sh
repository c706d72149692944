#!/bin/bash
# round 2, GPU call 1: parity of the restructured shade step + new full-size tests, A/B of the short-stack variants
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
( timeout 1200 python -m pytest tests -m gpu -x -q -s 2>&1 | tail -25 ) > gpurun_out/c1_pytest.log 2>&1
tail -8 gpurun_out/c1_pytest.log
V=raytracer-cuda_b200/build/var
bash scripts/ab.sh $V/libr1.so raytracer-cuda_b200/libcrt_b200.so $V/libstk4.so $V/libstk8.so > gpurun_out/c1_ab.log 2>&1
cat gpurun_out/c1_ab.log
