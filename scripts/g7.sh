set -x
python bench.py > gpurun_out/r1_bench_n1.json 2> gpurun_out/r1_bench_n1.err; echo rc=$?; tail -3 gpurun_out/r1_bench_n1.err; cat gpurun_out/r1_bench_n1.json
python bench.py --impl reference > gpurun_out/r1_bench_ref.json 2> gpurun_out/r1_bench_ref.err; echo rc=$?; cat gpurun_out/r1_bench_ref.json
python bench.py --mode megakernel --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r1_bench_mega.json 2> gpurun_out/r1_bench_mega.err; echo rc=$?; cat gpurun_out/r1_bench_mega.json
