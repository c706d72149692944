timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
P="python scripts/prof_render.py --reps 3 --spp 64"
for c in 0 2 4 6 8; do echo "DRAIN=$c"; CRT_DRAIN_ROUNDS=$c $P | grep "rep [12]"; done
