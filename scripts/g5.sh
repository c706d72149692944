P="python scripts/prof_render.py --reps 2 --spp 128"
for v in b200 sb4 sb5 sb6 sb8; do echo "LIB=$v"; CRT_LIB=$PWD/raytracer-cuda_b200/libcrt_$v.so $P | grep "rep 1"; done
