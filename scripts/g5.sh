P="python scripts/prof_render.py --reps 3 --spp 128"
for lib in b200 t128 t64; do for c in 1 2; do echo "LIB=$lib PARTS=$c"; CRT_LIB=/root/repo/raytracer-cuda_b200/libcrt_$lib.so CRT_WF_PARTS=$c $P | grep "rep [12]"; done; done
