set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
P="python scripts/prof_render.py --reps 2 --spp 128"
for f in 2 4 6 8 12; do echo "FETCH=$f"; CRT_FETCH_LANES=$f $P | grep "rep 1"; done
CRT_COOP=1 $P | grep "rep 1"
