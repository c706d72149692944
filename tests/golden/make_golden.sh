#!/bin/bash
# Generates the golden fixtures of tests/golden/ by running the UNMODIFIED reference kernels
# (oracle/_ref/ref_render_b*, built by `make -C oracle ref` from /root/reference) on a B200.
# Run on the GPU box:   gpurun -- 'bash tests/golden/make_golden.sh gpurun_out/golden'
# then copy gpurun_out/golden/* into tests/golden/ and commit.
set -e
OUT=${1:-gpurun_out/golden}
mkdir -p "$OUT"
REF=oracle/_ref
SEED=1804289383
# C1: Cornell box 256x256, 16 spp, 8 bounces (BASELINE.json configs[0])
$REF/ref_render_b8  --scene cornell      --width 256 --height 256 --spp 16 --seed $SEED --twin --primary --rngkat --dump "$OUT/c1"  | tee "$OUT/c1.json"
# C2 small: Cornell + blob 320x180, 4 spp, 20 bounces
$REF/ref_render_b20 --scene cornell_blob --width 320 --height 180 --spp 4  --seed $SEED --twin --primary --dump "$OUT/c2s" | tee "$OUT/c2s.json"
# C3 small: random spheres 300x200, 4 spp, 20 bounces
$REF/ref_render_b20 --scene spheres      --width 300 --height 200 --spp 4  --seed $SEED --twin --primary --dump "$OUT/c3s" | tee "$OUT/c3s.json"
mv "$OUT/c1_rngkat.txt" "$OUT/rngkat.txt"
ls -la "$OUT"
