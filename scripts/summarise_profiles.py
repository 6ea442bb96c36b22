"""Turns the ncu reports scripts/run_profiles.sh brought back in gpurun_out/ into the tracked summaries under profiles/:
raw-page CSV + details head + per-source-line table of each captured kernel, the launch list, and roofline_traffic.json
(DRAM bytes of k_map_fused per launch + the hash of the kernel sources the capture belongs to: bench.py refuses it once the
sources change).  usage: python scripts/summarise_profiles.py <tag> [reads per launch]"""
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
reads = int(sys.argv[2]) if len(sys.argv) > 2 else 200000
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")


def sh(cmd, out):
    with open(out, "w") as f:
        subprocess.run(cmd, stdout=f, stderr=subprocess.DEVNULL, check=False)


def metric(raw_csv, name):
    rows = list(csv.reader(open(raw_csv)))
    hdr, unit, val = rows[0], rows[1], rows[2]
    i = hdr.index(name)
    v = float(val[i].replace(",", ""))
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1e-3, "us": 1e-6, "ns": 1e-9}.get(unit[i], 1)
    return v * scale


for kern in ("map_fused", "locate_idx", "align_band"):
    rep = os.path.join(G, f"{kern}_{tag}.ncu-rep")
    if not os.path.exists(rep):
        continue
    raw = os.path.join(P, f"{kern}_full_{tag}.csv")
    sh(["ncu", "-i", rep, "--page", "raw", "--csv"], raw)
    det = os.path.join(P, f"{kern}_details_{tag}.txt")
    sh(["ncu", "-i", rep, "--page", "details"], det)
    lines = open(det).read().splitlines()[:160]
    open(det, "w").write("\n".join(lines) + "\n")
    src = os.path.join("/tmp", f"{kern}_src_{tag}.csv")
    sh(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], src)
    with open(os.path.join(P, f"{kern}_lines_{tag}.txt"), "w") as f:
        subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_lines.py"), src, "45"], stdout=f, check=False)
    print(kern, "time %.1f us  dram read %.3f GB  write %.3f GB" % (
        metric(raw, "gpu__time_duration.sum") * 1e6, metric(raw, "dram__bytes_read.sum") / 1e9, metric(raw, "dram__bytes_write.sum") / 1e9))

ll = os.path.join(G, f"launches_{tag}.csv")
if os.path.exists(ll):
    shutil.copy(ll, os.path.join(P, f"launches_{tag}.csv"))

raw = os.path.join(P, f"map_fused_full_{tag}.csv")
if os.path.exists(raw):
    import bench
    d = {"workload": "c2", "frac": 1.0, "kernel": "k_map_fused", "kmer_table": True,
         "capture": f"profiles/map_fused_full_{tag}.csv (ncu --set full --clock-control none, bench.py --steps 2 --warmup 1 --no-cpu --reads {reads})",
         "captured_reads_per_launch": reads,
         "dram_bytes_read_per_launch": metric(raw, "dram__bytes_read.sum"),
         "dram_bytes_write_per_launch": metric(raw, "dram__bytes_write.sum"),
         "gpu_time_s": metric(raw, "gpu__time_duration.sum"),
         "source_sha256_16": bench.kernel_source_hash(), "sources": bench.TRAFFIC_SOURCES}
    json.dump(d, open(os.path.join(P, "roofline_traffic.json"), "w"), indent=1)
    print("roofline_traffic.json:", d["dram_bytes_read_per_launch"] / reads, "B read per read, hash", d["source_sha256_16"])
