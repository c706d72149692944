timeout 800 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
P="python scripts/prof_render.py --reps 2 --spp 128"
for c in 32 64 96 128 256 512; do echo "CHUNK=$c"; CRT_FETCH_CHUNK=$c $P | grep "rep 1"; done
