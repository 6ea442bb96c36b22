"""cuobjdump -sass excerpts of the hot kernels for profiles/: mnemonic histogram + every global-memory / prefetch / vote /
shuffle instruction, so that the memory path of each kernel can be read without a GPU (no UTMALDG / UBLKCP: the gathers are
plain vector loads).  usage: python scripts/sass_excerpts.py <tag>"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
KERNELS = {"bm_mapfused.o": ["k_map_fusedILi4ELi12E"], "bm_map.o": ["k_map_voteILi4ELi3E"], "bm_align.o": ["k_align_bandILi5E"],
           "bm_locidx.o": ["k_locate_idxILi256E"], "bm_reads.o": ["k_pack_reads"]}
for obj, names in KERNELS.items():
    text = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "build", obj)], stdout=subprocess.PIPE, text=True).stdout
    funcs = re.split(r"\n\s*Function : ", text)
    for name in names:
        body = next((f for f in funcs if name in f.split("\n", 1)[0]), None)
        if body is None:
            continue
        ins = re.findall(r"/\*[0-9a-f]{4}\*/\s+(.*?);", body)
        hist = collections.Counter()
        for i in ins:
            m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", i.strip())
            if m:
                hist[m.group(2).split(".")[0]] += 1
        out = os.path.join(ROOT, "profiles", f"sass_{name.split('IL')[0]}_{tag}.txt")
        with open(out, "w") as f:
            f.write(f"# {body.splitlines()[0]}\n# {len(ins)} SASS instructions (static); cuobjdump -sass build/{obj}\n# mnemonic histogram:\n")
            for k, v in hist.most_common():
                f.write(f"#   {k:10s} {v}\n")
            f.write("# global loads / stores / prefetches / atomics / warp collectives:\n")
            for i in ins:
                if re.search(r"\b(LDG|STG|CCTL|ATOM|RED|VOTE|SHFL|REDUX|LDGSTS|UTMA|UBLKCP|LDS|STS)\b|PREFETCH", i):
                    f.write("    " + i.strip() + "\n")
        print(out, len(ins), "instructions;", {k: hist[k] for k in ("LDG", "STG", "LDS", "CCTL", "VOTE", "SHFL", "LOP3") if k in hist})
