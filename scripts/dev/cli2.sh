for thp in 1 0 1 0; do
BM_HOST_THP=$thp BM_PIPE_TRACE=1 BM_HOST_TRACE=1 python scripts/bench_cli.py > gpurun_out/cli_trace_$thp.json 2> gpurun_out/cli_trace.err
python - <<E
import json
d=json.loads(open("gpurun_out/cli_trace_$thp.json").read().strip().splitlines()[-1])
print("THP=$thp", d["map_wall_s"], d["index_wall_s"])
for l in d["tool_log"]:
    if "fastq batch" in l or "loop entered" in l or "Host pipeline" in l or "Wall clock" in l: print(l[:230])
E
done
