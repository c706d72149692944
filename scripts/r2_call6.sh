#!/bin/bash
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -6 ) > gpurun_out/c6_pytest.log 2>&1
tail -4 gpurun_out/c6_pytest.log
V=raytracer-cuda_b200/build/var; L=raytracer-cuda_b200/libcrt_b200.so
timeout 900 bash scripts/ab2.sh "main(dp-collapse);;$L" "greedy-collapse;;$V/libcur.so" "greedy+prefetch-next;;$V/libpf1.so" "octant-sorted;CRT_EXTEND_OCTANT=1;$L" > gpurun_out/c6_ab.log 2>&1
cat gpurun_out/c6_ab.log
for lib in $L $V/libcur.so; do echo "== spheres $lib"; CRT_LIB=$PWD/$lib python scripts/prof_render.py --scene spheres --width 1200 --height 800 --spp 500 --reps 2 --mode megakernel | grep -E "rep 1|bvh"; CRT_LIB=$PWD/$lib python scripts/prof_render.py --scene spheres --width 1200 --height 800 --spp 500 --reps 2 | grep -E "rep 1"; done > gpurun_out/c6_spheres.log 2>&1; cat gpurun_out/c6_spheres.log
CRT_BUILD_TRACE=1 python scripts/prof_render.py --scene terrain --spp 4 > gpurun_out/c6_buildtrace.log 2>&1; grep -E "trace|bvh" gpurun_out/c6_buildtrace.log
CRT_WF_TRACE=1 python scripts/prof_render.py --spp 2000 > gpurun_out/c6_trace2000.log 2>&1; grep -c wf_trace gpurun_out/c6_trace2000.log; tail -45 gpurun_out/c6_trace2000.log | cut -c1-140
timeout 600 python bench.py --impl reference --ref-spp 2000 --steps 1 --warmup 0 > gpurun_out/r2_bench_ref_full.json 2> gpurun_out/r2_bench_ref_full.err; cut -c1-900 gpurun_out/r2_bench_ref_full.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_ref.json 2>/dev/null; cut -c1-500 gpurun_out/r2_bench_ref.json
