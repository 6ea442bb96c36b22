# final bench lines of the round (one GPU): default line with the reference arm, then the other configurations
set -u
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1_r02d.json 2> gpurun_out/bench_n1_r02d.err || tail -5 gpurun_out/bench_n1_r02d.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_n1_r02d_reference.json 2> gpurun_out/bench_n1_r02d_reference.err || tail -5 gpurun_out/bench_n1_r02d_reference.err
python - <<P
import json
d=json.loads(open("gpurun_out/bench_n1_r02d.json").read().strip().splitlines()[-1])
print("c2 value %.4g ms %.3f e2e %.4g one-call %.4g roofline frac %.3f traffic_frac %s parity %s cpu %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e_one_call_at_a_time"]["value"], d["roofline"]["frac"], d["roofline"].get("traffic_frac"), d.get("parity",{}).get("sam_fields_equal"), d.get("cpu_baseline",{}).get("value")))
print(d["kernels_ms_per_step"], d["accuracy"])
r=json.loads(open("gpurun_out/bench_n1_r02d_reference.json").read().strip().splitlines()[-1]); print("reference", r.get("value"), r.get("unavailable"))
P
bash scripts/run_configs.sh r02d c2r c2x c5 c2f025 c1 c4-262144-k10 c4-16384
