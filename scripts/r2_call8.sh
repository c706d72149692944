#!/bin/bash
# 8 GPUs: C2 at N=8 and N=4 (the driver's launch), C4 (5 M triangles, 256 spp) at N=1,2,4,8, crt_cli --gpus 8
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | wc -l
run() { N=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N "$@"; }
for N in 8 4; do
  NCCL_DEBUG=INFO run $N --steps 3 --warmup 3 > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err; echo "C2 N=$N rc=$?"
  grep -E "NVLS|Connected all rings|comm 0x.* rank 0 nranks" gpurun_out/r2_bench_n$N.err | head -4
  python -c "
import json; j=json.load(open('gpurun_out/r2_bench_n$N.json')); print({k:j[k] for k in ['value','ms_per_step','n_gpus','e2e','check']})"
done
python bench.py --scene terrain --spp 256 --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/r2_bench_c4_n1.json 2> gpurun_out/r2_bench_c4_n1.err; echo "C4 N=1 rc=$?"
for N in 2 4 8; do run $N --scene terrain --spp 256 --steps 2 --warmup 2 > gpurun_out/r2_bench_c4_n$N.json 2> gpurun_out/r2_bench_c4_n$N.err; echo "C4 N=$N rc=$?"; done
for N in 1 2 4 8; do python -c "
import json; j=json.load(open('gpurun_out/r2_bench_c4_n$N.json')); print($N, {k:j[k] for k in ['value','ms_per_step','e2e']}, j['config']['bvh_build_ms'], j['check'])"; done
raytracer-cuda_b200/crt_cli --scene cornell_blob --width 2560 --height 1440 --spp 2000 --gpus 8
raytracer-cuda_b200/crt_cli --scene terrain --width 2560 --height 1440 --spp 256 --gpus 8
( timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q 2>&1 | tail -3 )
