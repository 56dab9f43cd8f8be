#!/bin/bash
# Developer tool: same-box A/B of prebuilt libnerf_b200.so variants (scripts/_<name>_libnerf_b200.so), interleaved.
# usage: scripts/ab.sh <rounds> <name> [<name> ...]
rounds=$1; shift
for r in $(seq 1 $rounds); do
  for name in "$@"; do
    cp scripts/_${name}_libnerf_b200.so nerf_rs_b200/libnerf_b200.so
    out=$(timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu 2>&1 | tail -1)
    echo "$name $(echo "$out" | grep -o '"mlp_coarse": [0-9.]*, "mlp_fine": [0-9.]*') $(echo "$out" | grep -o '"sm_mhz": [0-9.]*')"
  done
done
