timeout 800 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
P="python scripts/prof_render.py --reps 2 --spp 128"
$P | grep "rep 1"; $P --mode megakernel | grep "rep 1"
python scripts/prof_render.py --reps 2 --spp 32 --scene terrain | grep "rep 1"
