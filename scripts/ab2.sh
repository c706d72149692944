#!/bin/bash
# A/B of tuning builds / env settings on one box. Each argument: "label;ENV=VAL ENV2=VAL2;path/to/lib.so" (env part may be empty).
# 128-spp frames of the headline scene (two rounds, interleaved), one per-kernel timing pass, one 32-spp terrain frame.
run() { local envs="$1" lib="$2"; shift 2; env $envs CRT_LIB=$PWD/$lib python scripts/prof_render.py "$@" 2>&1; }
for rep in 1 2; do for A in "$@"; do IFS=';' read -r label envs lib <<< "$A"; echo "== $label"; run "$envs" "$lib" --spp 128 --reps 3 | grep "rep 2"; done; done
for A in "$@"; do IFS=';' read -r label envs lib <<< "$A"; echo "== time-kernels $label"; run "$envs" "$lib" --spp 128 --reps 2 --time-kernels | tail -2 | head -1; done
for A in "$@"; do IFS=';' read -r label envs lib <<< "$A"; echo "== terrain $label"; run "$envs" "$lib" --scene terrain --spp 32 --reps 2 | grep "rep 1"; done
