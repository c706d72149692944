set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
P="python scripts/prof_render.py --reps 2 --spp 128"
$P | grep -E "rep 1|bvh"
CRT_COOP=1 $P | grep "rep 1"
$P --mode megakernel | grep "rep 1"
for s in spheres terrain; do python scripts/prof_render.py --reps 2 --spp 32 --scene $s | grep -E "rep 1|bvh"; done
