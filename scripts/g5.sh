P="python scripts/prof_render.py --reps 3 --spp 64"
for c in 0 4 8 16 32 64; do echo "DRAIN=$c"; CRT_POSTPONE_LANES=$c $P | grep "rep [12]"; done
