set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu1.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu1.log
tail -30 gpurun_out/pytest_gpu1.log
timeout 600 bash tests/golden/make_golden.sh gpurun_out/golden > gpurun_out/golden.log 2>&1; echo "golden rc=$?"
tail -5 gpurun_out/golden.log
timeout 300 python bench.py --spp 16 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_wf16.json 2> gpurun_out/bench_wf16.err; echo rc=$?; cat gpurun_out/bench_wf16.json
timeout 300 python bench.py --spp 16 --steps 2 --warmup 1 --no-cpu-baseline --mode megakernel > gpurun_out/bench_mega16.json 2> gpurun_out/bench_mega16.err; echo rc=$?; cat gpurun_out/bench_mega16.json
timeout 600 python bench.py --impl reference --ref-spp 8 --steps 2 --warmup 1 > gpurun_out/bench_ref8.json 2> gpurun_out/bench_ref8.err; echo rc=$?; cat gpurun_out/bench_ref8.json
