#!/bin/bash
# tools/r02_capture.sh -- the evidence run of round 2 on one B200: GPU tests, seeded leg, launch list and one full ncu
# capture of the benchmarked fill kernel (after the same command exited 0 without ncu).
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -n 15 > gpurun_out/r02_gpu_tests.log
timeout 600 python bench.py --only-seeded --steps 3 > gpurun_out/r02_seeded.json 2> gpurun_out/r02_seeded.err
CMD="python tools/profile_step.py --workload map-ont-100k --reads 100000 --steps 2"
if timeout 300 $CMD > gpurun_out/r02_plain.log 2>&1; then
	timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_100k.csv $CMD > gpurun_out/r02_ncu_list.log 2>&1
	MM2B200_SPLIT=0 timeout 300 $CMD > gpurun_out/r02_plain_nosplit.log 2>&1 &&
	MM2B200_SPLIT=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fill_fast -s 1 -c 1 -o gpurun_out/r02_fill $CMD > gpurun_out/r02_ncu_full.log 2>&1
fi
ls -la gpurun_out | tail -n 20
