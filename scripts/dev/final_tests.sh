set -u
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_r02d.log 2>&1; tail -5 gpurun_out/pytest_gpu_r02d.log
bash scripts/run_configs.sh r02d c1
cmd="python bench.py --steps 2 --warmup 1 --no-cpu --reads 200000"
$cmd > gpurun_out/plain_r02.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_map_fused -s 2 -c 1 -f -o gpurun_out/map_fused_r02 $cmd > gpurun_out/ncu_f_r02.log 2>&1
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
