#!/bin/bash
# sweep of the tile kernel's shape on map-ont-100k (fill stage, device-resident):
# runs per warp x warps per CTA x chunk-0 passes per straight-line block
cd "$(dirname "$0")/.."
for cfg in ${CFGS:-"0,4,4" "8,4,2" "8,4,4" "8,4,8" "16,2,2" "16,2,4" "16,2,8" "16,4,4" "16,4,8" "32,2,4" "32,2,8"}; do
  IFS=, read r nw g <<< "$cfg"
  echo "== MM2B200_TILE=$r MM2B200_TILE_WARPS=$nw MM2B200_TILE_GROUP=$g"
  MM2B200_TILE=$r MM2B200_TILE_WARPS=$nw MM2B200_TILE_GROUP=$g timeout 300 python tools/profile_step.py --reads ${READS:-100000} --steps 3 2>&1 | tail -1
done
