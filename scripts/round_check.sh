# Round-end check on one B200 (under gpurun): gpu tests, smoke, reference arm, full bench
TAG=${1:-r2g}
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 )
python __graft_entry__.py smoke 2>&1 | tail -2
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; cut -c1-260 gpurun_out/${TAG}_bench_ref.json
CRT_BENCH_WATCHDOG_S=120 timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err; echo rc=$?; tail -3 gpurun_out/${TAG}_bench_n1.err; python -c "
import json; d=json.load(open('gpurun_out/${TAG}_bench_n1.json')); r=d['roofline']
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['check'], r['frac'], r['issue_slots'], d['clocks'])
for k in r['kernels']: print(k['kernel'], k.get('ms'), k.get('frac'), (k.get('ncu_counters') or {}).get('l1_hit_pct'))"
