timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
echo C5; for c in 0 4; do CRT_DRAIN_ROUNDS=$c python scripts/prof_render.py --reps 40 --spp 1 --bounces 8 --width 1920 --height 1080| grep "rep 3[89]"; done
python bench.py > gpurun_out/r1_bench_n1.json 2> gpurun_out/r1_bench_n1.err; tail -1 gpurun_out/r1_bench_n1.json
python bench.py --impl reference > gpurun_out/r1_bench_ref.json 2> gpurun_out/r1_bench_ref.err; tail -1 gpurun_out/r1_bench_ref.json
