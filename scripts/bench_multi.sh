# N-GPU bench + reference arm under gpurun --gpus N: bash scripts/bench_multi.sh N
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 2 --warmup 3 > gpurun_out/r1_bench_n$N.json 2> gpurun_out/r1_bench_n$N.err; echo rc=$?; tail -5 gpurun_out/r1_bench_n$N.err; cat gpurun_out/r1_bench_n$N.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 1 --warmup 1 --ref-spp 16
