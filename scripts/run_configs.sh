#!/bin/bash
# One bench line per BASELINE configuration, run on the GPU box under gpurun; lines go to gpurun_out/bench_<cfg>_<tag>.json
# usage: bash scripts/run_configs.sh <tag> [cfg ...]
set -u
tag=${1:-r02}; shift
out=gpurun_out
mkdir -p $out
run() { name=$1; shift; echo "== $name: $*"; python bench.py "$@" > $out/bench_${name}_$tag.json 2> $out/bench_${name}_$tag.err || tail -3 $out/bench_${name}_$tag.err; python - <<P
import json
try:
    d = json.load(open("$out/bench_${name}_$tag.json"))
    print("   value %.4g reads/s, %.3f ms/step, e2e %.4g, roofline %s %s, parity %s" % (d["value"], d["ms_per_step"], (d.get("e2e") or {}).get("value", 0), d["roofline"].get("achieved"), d["roofline"].get("unit"), (d.get("parity") or {}).get("sam_fields_equal")))
except Exception as e:
    print("   no line:", e)
P
}
cfgs=${*:-"c2f025 c2r c5 c1 c3 c4-262144-k10 c4-262144 c4-16384"}
for c in $cfgs; do
  case $c in
    c2f025) run c2_f025 --config c2 --frac 0.25 --steps 10 --warmup 3 ;;
    c2r) run c2r --config c2r --steps 10 --warmup 3 --no-cpu ;;
    c2x) run c2x --config c2x --steps 10 --warmup 3 --no-cpu ;;
    c5) run c5 --config c5 --steps 5 --warmup 2 ;;
    c1) run c1 --config c1 --frac 0.25 --steps 20 --warmup 3 ;;
    c3) run c3 --config c3 --steps 3 --warmup 2 --no-cpu ;;
    c4-262144-k10) run c4_262144_k10 --config c4-262144-k10 --steps 10 --warmup 3 --no-cpu ;;
    c4-262144) run c4_262144 --config c4-262144 --steps 10 --warmup 3 ;;
    c4-16384) run c4_16384 --config c4-16384 --steps 5 --warmup 2 --no-cpu ;;
  esac
done
