set -u
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_r02f.log 2>&1; tail -4 gpurun_out/pytest_gpu_r02f.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1_r02f.json 2> gpurun_out/bench_n1_r02f.err || tail -5 gpurun_out/bench_n1_r02f.err
python - <<P
import json
d=json.loads(open("gpurun_out/bench_n1_r02f.json").read().strip().splitlines()[-1])
r=d["roofline"]
print("c2 value %.4g ms %.3f e2e %.4g one-call %.4g frac %.3f traffic_frac %s model_frac %.3f parity %s cpu %.4g launches %d" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e_one_call_at_a_time"]["value"], r["frac"], r.get("traffic_frac"), r["traffic_model_frac"], d.get("parity",{}).get("sam_fields_equal"), d.get("cpu_baseline",{}).get("value"), d["gpu_launches"]))
print(d["clocks"])
P
