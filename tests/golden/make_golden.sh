#!/bin/bash
# make_golden.sh -- regenerates tests/golden/*.npz and paf.txt from the UNMODIFIED reference at /root/reference.
#
# Builds the reference's own C files (in a scratch directory, never copied into the repo) with
#   * gen/loopback_stub.c in place of the closed libfpgadrv.a (runs the fork's software device model fpga_work), and
#   * gen/dump_hook.c wrapped around mm_chain_dp_fpga / mm_chain_dp_bottom,
# runs the fixture commands of SURVEY.md section 4 (BASELINE.json configs[0] is the first one), and converts the dump
# into one .npz per command with gen/pack_golden.py.  Needs /root/reference, gcc, zlib; run from the repo root.
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
W=$(mktemp -d /tmp/mm2golden.XXXXXX)
cleanup() { [ -n "${W:-}" ] && [ -d "$W" ] && case "$W" in /tmp/mm2golden.*) rm -rf -- "$W";; esac; }
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
: > "$HERE/paf.txt"
run() { # name, args...
	local name=$1; shift
	: > "$W/$name.dump"
	echo "## $name: minimap2 -t 12 $*" | sed "s#$REF/##g; s#$W/##g" >> "$HERE/paf.txt"
	( cd "$W" && MM2_GOLDEN_DUMP="$W/$name.dump" ./minimap2 -t 12 "$@" 2> "$W/$name.err" | cut -f1-18 >> "$HERE/paf.txt" ) || { tail "$W/$name.err"; exit 1; }
	python "$HERE/gen/pack_golden.py" "$W/$name.dump" "$HERE/$name.npz"
}
run map-ont_MT -x map-ont "$T/MT-human.fa" "$T/MT-orang.fa"
run default_inv "$T/t-inv.fa" "$T/q-inv.fa"
run default_q2 "$T/t2.fa" "$T/q2.fa"
run ava-ont_MT -x ava-ont "$W/both.fa" "$W/both.fa"
run asm5_MT -cx asm5 "$T/MT-human.fa" "$T/MT-orang.fa"
echo "golden vectors written to $HERE"
