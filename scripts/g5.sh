P="python scripts/prof_render.py --reps 3 --spp 128"
for lib in b200 pf8192 pf32768 b200 pf8192 pf32768; do echo "LIB=$lib"; CRT_LIB=/root/repo/raytracer-cuda_b200/libcrt_$lib.so $P | grep "rep [2]"; done
