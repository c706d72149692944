#!/bin/bash
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests -m gpu -x -q -s 2>&1 | tail -25 ) > gpurun_out/c2_pytest.log 2>&1
tail -6 gpurun_out/c2_pytest.log
V=raytracer-cuda_b200/build/var; L=raytracer-cuda_b200/libcrt_b200.so
timeout 900 bash scripts/ab2.sh "main(bulk,stk4);;$L" "nobulk;CRT_SHADE_BULK=0;$L" "stk2;;$V/libstk2.so" "stk3;;$V/libstk3.so" "stk6;;$V/libstk6.so" "sh4;;$V/libsh4.so" "sh6;;$V/libsh6.so" > gpurun_out/c2_ab.log 2>&1
cat gpurun_out/c2_ab.log
