#!/bin/bash
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests -m gpu -x -q -s 2>&1 | tail -25 ) > gpurun_out/c4_pytest.log 2>&1
tail -6 gpurun_out/c4_pytest.log
L=raytracer-cuda_b200/libcrt_b200.so
timeout 900 bash scripts/ab2.sh "main(graph);;$L" "nograph;CRT_WF_GRAPH=0;$L" "fetch8;CRT_FETCH_LANES=8;$L" "fetch16;CRT_FETCH_LANES=16;$L" "blocked4;CRT_WF_BLOCKED=1 CRT_WF_PARTS=4;$L" "blocked8;CRT_WF_BLOCKED=1 CRT_WF_PARTS=4 CRT_TAIL_LIVE=40000;$L" > gpurun_out/c4_ab.log 2>&1
cat gpurun_out/c4_ab.log
for lm in 1 2 3; do echo "== leaf_max $lm"; python scripts/prof_render.py --spp 128 --reps 3 --leaf-max $lm | grep -E "rep 2|bvh"; python scripts/prof_render.py --scene terrain --spp 32 --reps 2 --leaf-max $lm | grep -E "rep 1|bvh"; done > gpurun_out/c4_leaf.log 2>&1; cat gpurun_out/c4_leaf.log
python scripts/c5.py > gpurun_out/c4_c5_graph.json 2>&1; CRT_WF_GRAPH=0 python scripts/c5.py > gpurun_out/c4_c5_nograph.json 2>&1; cat gpurun_out/c4_c5_graph.json gpurun_out/c4_c5_nograph.json
timeout 600 python bench.py --spp 64 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/c4_bench64.json 2> gpurun_out/c4_bench64.err; echo rc=$?; tail -3 gpurun_out/c4_bench64.err; cut -c1-1200 gpurun_out/c4_bench64.json
