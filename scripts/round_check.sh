# Round-end check on one B200 (under gpurun): gpu tests, smoke, reference arm, full bench, ncu captures for profiles/
set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python __graft_entry__.py smoke 2>&1 | tail -3
python bench.py --impl reference --steps 2 --warmup 3 > gpurun_out/r1h_bench_ref.json 2> gpurun_out/r1h_bench_ref.err; tail -c 600 gpurun_out/r1h_bench_ref.json
python bench.py --steps 2 --warmup 3 > gpurun_out/r1h_bench_n1.json 2> gpurun_out/r1h_bench_n1.err; echo rc=$?; tail -3 gpurun_out/r1h_bench_n1.err; cat gpurun_out/r1h_bench_n1.json
bash scripts/profile_round.sh r1h > gpurun_out/r1h_profile.log 2>&1; tail -8 gpurun_out/r1h_profile.log
