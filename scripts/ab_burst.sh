#!/bin/bash
# Developer tool: per-sample fine-pass time at several frame sizes for prebuilt library variants.
# usage: scripts/ab_burst.sh <name> [<name> ...]: scripts/burst_ab.py for each prebuilt library variant
for name in "$@"; do
  cp scripts/_${name}_libnerf_b200.so nerf_rs_b200/libnerf_b200.so
  echo "== $name"; timeout 300 python scripts/burst_ab.py 2>&1 | tail -4
done
