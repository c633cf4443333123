#!/bin/bash
# Run on the GPU box (gpurun): plain bench first, then the ncu launch list and one full capture of the
# same command line.  Outputs land in gpurun_out/ and are summarised into profiles/ by tools/summarise_profiles.py.
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
tail -c 400 gpurun_out/plain.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
tail -1 gpurun_out/ncu_list.log
# matched launches: 3 radix passes of the generator + 11 of the warm-up step; then capture the first timed step in order
ncu --set full --clock-control none --import-source on -k regex:"intern_pack|radix_pass|diff_flag|diff_apply|diff_sample|bitmap_init|bitmap_count|tile_scan|assemble_bitmap" -s 14 -c 11 -o gpurun_out/prof_full $CMD > gpurun_out/ncu_full.log 2>&1
tail -1 gpurun_out/ncu_full.log
