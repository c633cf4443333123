#!/bin/bash
# Run on the GPU box (gpurun): the partition-first path (PTE_PATH=partition) on BASELINE config 3 -- plain run first,
# then one ncu --set full capture of its three kernels.  Outputs land in gpurun_out/; summaries go to profiles/ by hand.
set -u
mkdir -p gpurun_out
export PTE_PATH=partition
CMD="python tools/run_once.py 1e9 1e7 2"
$CMD > gpurun_out/part_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/part_plain.log; exit 1; }
tail -2 gpurun_out/part_plain.log
# matched launches per run: hist, scatter, replay (sample), replay; the first run is warm-up
ncu --set full --clock-control none --import-source on -k regex:"part_hist|part_scatter|part_replay" -s 4 -c 4 -o gpurun_out/prof_part $CMD > gpurun_out/part_ncu.log 2>&1
tail -1 gpurun_out/part_ncu.log
