#!/bin/bash
# level-1 drop-in rate from C callers (run on the GPU box): builds the harness, dumps 16 Ki map-ont reads, runs both modes
cd "$(dirname "$0")/.."
set -e
gcc -O2 -I include tools/dropin_rate.c -o /tmp/dropin_rate -L minimap2_chaindp_b200 -lmm2chain_b200 -Wl,-rpath,$PWD/minimap2_chaindp_b200 -lpthread
python tools/dropin_rate.py --dump /tmp/anchors.bin
/tmp/dropin_rate /tmp/anchors.bin chain
/tmp/dropin_rate /tmp/anchors.bin split
