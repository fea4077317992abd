#!/bin/bash
# fill kernel ring depth on the long-read workloads (device-resident stage times)
cd "$(dirname "$0")/.."
for wl in "map-pb-ul-20k 1500" "asm5-50mb 8" "ava-ont-50k 4000" "map-ont-100k 100000"; do
  set -- $wl
  for ring in 64 -1 128 256; do   # -1: 64-anchor ring + batched deep loads (the long-read default)
    echo "== $1 reads=$2 MM2B200_RING=$ring"
    MM2B200_RING=$ring timeout 300 python tools/profile_step.py --workload $1 --reads $2 --steps 2 2>&1 | tail -1
  done
done
