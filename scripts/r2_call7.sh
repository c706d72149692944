#!/bin/bash
# 2 GPUs: multi-GPU tests through the C-ABI, bench at N=2, crt_cli --gpus 2, tail threshold sweep on GPU 0
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader
( timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -s 2>&1 | tail -8 ) > gpurun_out/c7_pytest_multi.log 2>&1
cat gpurun_out/c7_pytest_multi.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo rc=$?; tail -3 gpurun_out/r2_bench_n2.err; python -c "
import json; j=json.load(open('gpurun_out/r2_bench_n2.json')); print({k:j[k] for k in ['value','ms_per_step','n_gpus','e2e','check']})"
raytracer-cuda_b200/crt_cli --scene cornell_blob --width 2560 --height 1440 --spp 256 --gpus 2 --ppm gpurun_out/c7_cli_g2.ppm
raytracer-cuda_b200/crt_cli --scene cornell_blob --width 2560 --height 1440 --spp 256 --gpus 2 --nccl
raytracer-cuda_b200/crt_cli --scene cornell_blob --width 2560 --height 1440 --spp 256 --ppm gpurun_out/c7_cli_g1.ppm
cmp gpurun_out/c7_cli_g1.ppm gpurun_out/c7_cli_g2.ppm | head -2; ls -la gpurun_out/*.ppm; rm -f gpurun_out/*.ppm
for tl in 150000 50000 20000 8000 3000; do echo "== tail_live $tl"; CRT_TAIL_LIVE=$tl python scripts/prof_render.py --spp 2000 --reps 1 | grep "rep 0"; done
for ti in 256 1024; do echo "== tail_live 8000 tail_iters $ti"; CRT_TAIL_LIVE=8000 CRT_TAIL_ITERS=$ti python scripts/prof_render.py --spp 2000 --reps 1 | grep "rep 0"; done
