#!/bin/bash
# C5 (1 spp per frame, 1920x1080, 8 bounces): which wavefront settings suit a 9-iteration frame?
mkdir -p gpurun_out
run() { echo "== $*"; env "$@" python scripts/c5.py 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); w = d['wavefront']; print('wavefront median %.3f p99 %.3f min %.3f ms, %d launches %d graphs %d syncs %d iterations; mega %.3f' % (w['median_ms'], w['p99_ms'], w['min_ms'], w['launches_per_frame'], w['graph_launches'], w['host_syncs'], w['iterations'], d['megakernel']['median_ms']))
"; }
run CRT_NOP=1
run CRT_DRAIN_ROUNDS=0
run CRT_WF_PARTS=1
run CRT_WF_PARTS=1 CRT_DRAIN_ROUNDS=0
run CRT_TAIL_LIVE=400000
run CRT_TAIL_LIVE=1000000
run CRT_DRAIN_ROUNDS=0 CRT_TAIL_LIVE=400000
run CRT_WF_PARTS=1 CRT_DRAIN_ROUNDS=0 CRT_TAIL_LIVE=800000
run CRT_DRAIN_ROUNDS=8
run CRT_NOP=2
