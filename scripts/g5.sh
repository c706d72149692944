P="python bench.py --spp 128 --steps 1 --warmup 1 --no-cpu-baseline"
for v in 0 1; do echo "SORT=$v"; CRT_SORT=$v $P | python -c "
import json,sys; j=json.loads(sys.stdin.read()); r=j['roofline']; print(j['value'], j['ms_per_step'], 'ext ms/launch', r['kernel_ms_per_launch'], 'ext total', r['kernel_ms_per_launch']*r['kernel_launches_per_step'], 'shade ms', r['shade_ms_per_step'])"; done
