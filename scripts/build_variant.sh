#!/bin/bash
# Tuning build of the library for A/B runs: bash scripts/build_variant.sh NAME "-DMACRO=VALUE ..."  ->  raytracer-cuda_b200/build/var/libNAME.so
set -e
NAME=$1; FLAGS=$2
PKG=$(cd "$(dirname "$0")/../raytracer-cuda_b200" && pwd)
D=$PKG/build/var/$NAME
mkdir -p "$D/csrc" "$PKG/build/var/include"
cp "$PKG"/../include/crt_api.h "$PKG"/../include/crt_host.hpp "$PKG/build/var/include/"
cp "$PKG"/csrc/* "$D/csrc/"
make -C "$PKG" -j4 SRC="$D/csrc" BUILD="$D/build" OUT="$PKG/build/var/lib$NAME.so" CLI="$PKG/build/var/cli_$NAME" EXTRA="$FLAGS" 2>&1 | grep -E " error |Error 1" | grep -v cli_ || true
grep -E "Compiling|Used|spill" "$D/build/crt_wavefront.ptxas.log" | grep -A2 "k_wf_shade\|k_wf_extendILb0ELb0" | grep -E "Used|spill" || true
ls -la "$PKG/build/var/lib$NAME.so"
