#!/bin/bash
# Round evidence, run on the GPU box under gpurun (outputs into gpurun_out/, summarised into profiles/ afterwards):
#   1. plain short bench, then the ncu launch list of the same command
#   2. ncu --set full of the dominant kernels (k_map_fused; k_align_band on the c5 line)
# usage: bash scripts/run_profiles.sh <tag>
set -u
tag=${1:-r02}
out=gpurun_out
mkdir -p $out
cmd="python bench.py --steps 2 --warmup 1 --no-cpu --reads 200000"
$cmd > $out/plain_$tag.log 2>&1 || { echo "plain run failed"; tail -5 $out/plain_$tag.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/launches_$tag.csv $cmd > $out/ncu_l_$tag.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_map_fused -s 2 -c 1 -o $out/map_fused_$tag $cmd > $out/ncu_f_$tag.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_locate_idx -s 2 -c 1 -o $out/locate_idx_$tag $cmd > $out/ncu_i_$tag.log 2>&1
cmd5="python bench.py --config c5 --steps 2 --warmup 1 --no-cpu --reads 200000"
$cmd5 > $out/plain_c5_$tag.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_align_band -s 1 -c 1 -o $out/align_band_$tag $cmd5 > $out/ncu_a_$tag.log 2>&1
ls -la $out | tail -12
