set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
P="python scripts/prof_render.py --reps 2"
for c in 1 0; do echo "COOP=$c"; CRT_COOP=$c $P --spp 32 | grep "rep 1";  CRT_COOP=$c $P --spp 256 | grep "rep 1"; done
$P --spp 256 --mode megakernel | grep "rep 1"
