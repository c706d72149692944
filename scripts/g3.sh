set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
P="python scripts/prof_render.py --spp 32 --reps 2"
$P | grep -E "rep 1|bvh"
$P --counted | grep "rep 1"
$P --mode megakernel | grep "rep 1"
for f in 1 4 8 12 16 24 32; do echo "FETCH=$f"; CRT_FETCH_LANES=$f $P | grep "rep 1"; done
for p in 0 4 8 12 16; do echo "POSTPONE=$p"; CRT_POSTPONE_LANES=$p $P | grep "rep 1"; done
for b in 2 3 4 5 6 8; do echo "BLOCKS=$b"; CRT_EXTEND_BLOCKS=$b $P | grep "rep 1"; done
