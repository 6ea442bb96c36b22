N=$1
if [ "$N" = "1" ]; then python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/bench_n1_dev.json 2> gpurun_out/bench_n1_dev.err
else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n${N}_dev.json 2> gpurun_out/bench_n${N}_dev.err; fi
tail -3 gpurun_out/bench_n${N}_dev.err
python - <<E
import json
d=json.loads(open("gpurun_out/bench_n${N}_dev.json").read().strip().splitlines()[-1])
print("N=$N value", d["value"], "ms", d["ms_per_step"], "gather_verified", d.get("gather_verified"))
print("e2e", d["e2e"]["value"], d["e2e"].get("lanes_give_the_same_records"), "one call", d["e2e_one_call_at_a_time"]["value"], "gathered", (d.get("e2e_gathered_to_rank0") or {}).get("value"), "ascii", (d.get("e2e_ascii") or {}).get("value"))
print("index ms: build", d["index_build_ms"], "bcast", d["index_bcast_ms"], "sharded", d.get("index_build_sharded_ms"), d.get("index_build_sharded_verified"))
E
