# A/B of tuning builds on one box: bash scripts/ab.sh libX.so libY.so ...   (paths relative to the repo root; 128-spp frames of the headline scene)
for rep in 1 2; do for L in "$@"; do echo "== $L"; CRT_LIB=$PWD/$L python scripts/prof_render.py --spp 128 --reps 3 2>&1 | grep "rep 2"; done; done
for L in "$@"; do echo "== time-kernels $L"; CRT_LIB=$PWD/$L python scripts/prof_render.py --spp 128 --reps 2 --time-kernels 2>&1 | tail -2 | head -1; done
for L in "$@"; do echo "== terrain $L"; CRT_LIB=$PWD/$L python scripts/prof_render.py --scene terrain --spp 32 --reps 2 2>&1 | grep "rep 1"; done
