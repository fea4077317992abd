#!/bin/bash
# make_seed_golden.sh -- regenerates tests/golden/seed_*.npz from the UNMODIFIED reference at /root/reference: for every
# fixture run, everything the fork's device receives (the four index streams B/H/V/P of index.c:603-720, the parameters of
# fpga_set_params, every request packet with its minimizers, map.c:286-324) and the anchors the reference's own
# collect_seed_hits (map.c:187-236) produced for every read.  Same build as make_golden.sh (loop-back stub + dump hook).
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
W=$(mktemp -d /tmp/mm2seed.XXXXXX)
cleanup() { [ -n "${W:-}" ] && [ -d "$W" ] && case "$W" in /tmp/mm2seed.*) rm -rf -- "$W";; esac; }
trap cleanup EXIT
CF="-g -w -O2 -std=gnu99 -DHAVE_KALLOC -I$REF"
for f in fpga_chaindp kthread kalloc misc bseq sketch sdust options index align hit map format pe esterr main getopt; do
	gcc -c $CF -msse4.1 "$REF/$f.c" -o "$W/$f.o"
done
gcc -c $CF -Dmm_chain_dp_fpga=ref_real_fpga -Dmm_chain_dp_bottom=ref_real_bottom "$REF/chain.c" -o "$W/chain.o"
gcc -c $CF -msse2 "$REF/ksw2_ll_sse.c" -o "$W/ksw2_ll_sse.o"
gcc -c $CF -msse4.1 -DKSW_CPU_DISPATCH "$REF/ksw2_dispatch.c" -o "$W/ksw2_dispatch.o"
for f in extz2 extd2 exts2; do
	gcc -c $CF -msse4.1 -DKSW_CPU_DISPATCH "$REF/ksw2_${f}_sse.c" -o "$W/ksw2_${f}_sse41.o"
	gcc -c $CF -msse2 -mno-sse4.1 -DKSW_CPU_DISPATCH -DKSW_SSE2_ONLY "$REF/ksw2_${f}_sse.c" -o "$W/ksw2_${f}_sse2.o"
done
gcc -c -g -O2 -w "$HERE/gen/loopback_stub.c" -o "$W/stub.o"
gcc -c -g -O2 -w "$HERE/gen/dump_hook.c" -o "$W/hook.o"
gcc -o "$W/minimap2" "$W"/*.o -lm -lz -lpthread
T=$REF/test
cat "$T/MT-human.fa" "$T/MT-orang.fa" > "$W/both.fa"
python "$HERE/gen/sim_reads.py" "$W/gseq_ref.fa" "$W/gseq_reads.fa"
run() { # name, args...
	local name=$1; shift
	: > "$W/$name.dump"
	( cd "$W" && MM2_SEED_DUMP="$W/$name.dump" ./minimap2 -t 12 "$@" 2> "$W/$name.err" > "$W/$name.paf" ) || { tail "$W/$name.err"; exit 1; }
	python "$HERE/gen/pack_seed_golden.py" "$W/$name.dump" "$HERE/seed_$name.npz"
}
run map-ont_MT -x map-ont "$T/MT-human.fa" "$T/MT-orang.fa"
run ava-ont_MT -x ava-ont "$W/both.fa" "$W/both.fa"
run default_inv "$T/t-inv.fa" "$T/q-inv.fa"
run map-ont_gseq -x map-ont "$W/gseq_ref.fa" "$W/gseq_reads.fa"
[ "${KEEP:-0}" = 1 ] && cp "$W"/*.dump /tmp/ || true
echo "seeding fixtures written to $HERE"
