#!/usr/bin/env python
"""bench.py -- reads/sec mapped through the candidate-bucket hot path (BASELINE.json metric) on N B200s.

A step is one pass of the hot path (mapper K2f = filter + sampler + coarse vote + exact recount, locator K3i + gather,
best-location merge K3b) over one batch of synthetic reads.  Workload at any N: BASELINE config 1 of `configs` ("synthetic 1.7 Gbp reference, BUCKET_LEN=65536,
1M x 300bp reads, 0.2% sub, 0.025% ins/del") per GPU -- reads are sharded, the index is replicated (weak scaling).

  value     whole-job reads/s with the read batch resident in HBM (CUDA events on the library's stream, max over ranks)
  e2e       the same through bm_map_locate() with pinned HOST buffers: H2D of bases/qualities/offsets and D2H of the
            location records inside the timed region
  roofline  the dominant kernel of the step, the fused mapper kernel (k_map_fused): algorithmic bytes = the bytes its
            loads ask for, counted in the kernel; time = CUDA events around each launch inside the timed region.
            roofline_qgram_path: the same for k_map_vote, the full-width q-gram row gather used when the coarse k-mer
            table is off or does not fit (measured in the same run, outside the timed region)
  cpu_baseline / --impl reference: the reference's CPU implementation of the path on the host cores, on a bounded
            sample of the same workload: oracle/_ref (the reference's own main.cpp and headers, built from
            /root/reference against the test-only seqan3 shim; "reference") as one single-threaded process per core,
            else the plain C++ restatement under oracle/ with all host threads ("port")
  N > 1     reads sharded over the ranks, index replicated; each step ends with the gather of the location records
            to rank 0's device (bucket-map_b200/dist.py); wall clock between device-synchronised barriers, max over
            ranks
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (genome bases, bucket_len, read_len, sub, indel, frac)
    "c1": dict(n_bases=4_600_000, L=65536, read_len=150, sub=0.002, indel=0.00025, frac=1.0,
               desc="4.6 Mbp synthetic reference, L=65536, 150bp reads"),
    "c2": dict(n_bases=1_700_000_000, L=65536, read_len=300, sub=0.002, indel=0.00025, frac=1.0,
               desc="1.7 Gbp synthetic reference, L=65536, 300bp reads (0.2% sub, 0.025% ins/del), -f 1"),
    "c3": dict(n_bases=3_100_000_000, L=65536, read_len=150, sub=0.002, indel=0.00025, frac=1.0,
               desc="3.1 Gbp synthetic reference, L=65536, 150bp reads"),
}
FMH_X, FMH_Y = 40503, 7919
GENOME_SEED, READ_SEED = 3, 4


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    ap.add_argument("--reads", type=int, default=1_000_000, help="reads per GPU per step")
    ap.add_argument("--cpu-sample", type=int, default=0, help="reads in the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--lanes", type=int, default=0,
                    help="bm_map_locate runs batches of at least this many reads on two pipeline lanes (0: library default)")
    ap.add_argument("--lidx", type=int, default=1, choices=[0, 1],
                    help="1: locator searches the genome k-mer position index; 0: scans each candidate bucket's text")
    ap.add_argument("--ktable", type=int, default=1, choices=[0, 1, 8, 16, 32],
                    help="1: mapper votes over the coarse k-mer table (derived on the device when 4^k rows fit; 16 "
                         "buckets per coarse bit; 8/16/32 pick the group size); 0: gathers the k-q+1 q-gram rows per "
                         "sample over the full row width")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """samples SM clock and throttle reasons of one GPU while the timed region runs (pynvml)"""

    def __init__(self, index, period=0.05):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.sm, self.reasons, self.sm_max = [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._halt.is_set():
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._halt.wait(self.period)

    def stop(self):
        self._halt.set()
        if self.is_alive():
            self.join()
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons)}


def make_workload(cfg, n_reads, seed, region=None):
    """genome (packed) + buckets + simulated reads.  region=(lo,hi): reads start inside that genome range."""
    from bucketmap_b200 import datagen
    packed = datagen.random_genome_packed(cfg["n_bases"], GENOME_SEED)
    NB, b_start, b_len, _ = datagen.bucketize(np.array([cfg["n_bases"]], np.uint64), cfg["L"], 300)
    return packed, NB, b_start, b_len


def simulate(cfg, packed, n_reads, seed, n_span=None):
    from bucketmap_b200 import datagen
    return datagen.simulate_reads(packed, n_span or cfg["n_bases"], n_reads, cfg["read_len"], cfg["sub"],
                                  cfg["indel"], cfg["indel"], seed)[:3]


def bind_to_gpu_numa_node(device):
    """run this rank (and first-touch its pinned buffers) on the CPUs next to its GPU: on a two-socket box a rank that
    lands on the far socket pushes every host<->device byte across the socket link"""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(device)
        try:
            pynvml.nvmlDeviceSetCpuAffinity(h)
        except Exception:
            words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
            cpus = [64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1]
            if cpus:
                os.sched_setaffinity(0, cpus)
        return sorted(os.sched_getaffinity(0))
    except Exception:
        return None


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(config_name, n_reads, ktable):
    """per-launch DRAM bytes (read+write) of k_map_vote from the committed ncu --set full capture of this workload,
    scaled from the captured launch size to this launch's reads (traffic is linear in reads: no reuse across reads)"""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            if d.get("workload") == config_name and bool(d.get("kmer_table", False)) == bool(ktable):
                per = (d["dram_bytes_read_per_launch"] + d["dram_bytes_write_per_launch"]) / d["captured_reads_per_launch"]
                return per * n_reads
        except Exception:
            pass
    return None


# ------------------------------------------------------------------------------------------------- CPU (oracle) leg
def cpu_leg(cfg, packed, NB, b_start, b_len, k2i, rows, n_sample, full_reads, threads):
    """times the CPU restatement (map + locate + best-location merge) on a bounded sample of the workload.
    The sample reads start inside the first n_sample/full_reads of the genome, so the number of reads per bucket (the
    amortisation of the reference's per-bucket k-mer index build) matches the full workload."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import bm_oracle
    bm_oracle.build()
    o = bm_oracle.Oracle(threads)
    P = o.params(cfg["L"])
    span = max(int(cfg["n_bases"] * n_sample / full_reads), 4 * cfg["L"])
    span = min(span, cfg["n_bases"])
    seq, qual, off = simulate(cfg, packed, n_sample, READ_SEED + 1000, n_span=span)
    segs = o.segments(P, off)
    ix = o.index_open(P, NB, k2i, rows)
    t0 = time.perf_counter()
    coff, cand, stats = o.map_segments_h(ix, P, seq, qual, off, segs)
    t1 = time.perf_counter()
    loff, locs = o.locate(P, packed, b_start, b_len, seq, qual, off, segs, coff, cand)
    foff, best = o.filter_best(P, off, loff, locs)
    t2 = time.perf_counter()
    o.index_close(ix)
    return {"reads": n_sample, "seconds": t2 - t0, "map_s": t1 - t0, "locate_s": t2 - t1,
            "reads_per_s": n_sample / (t2 - t0), "threads": o.threads, "n_locs": int(len(best)),
            "span": span}


# ----------------------------------------------------------------------------------- CPU (reference itself) leg
def reference_leg(cfg, packed, NB, b_start, k2i, rows, n_per_proc, full_reads, procs):
    """times oracle/_ref/bucketmap_<NB>_<L> -- the reference's own main.cpp built against oracle/shim -- on a bounded
    sample of the workload: `procs` independent single-threaded processes (the reference has no threads), each mapping
    its own n_per_proc reads drawn from its own genome span sized so that reads per bucket match the full batch.
    Timed: the reference's own [BENCHMARK] phase timers for bucket mapping + per-bucket k-mer index builds + offset
    finding (the scope of one bench step); FASTA/index loading and SAM output are outside, as they are for the GPU arm.
    The index files are handed over ready-made (they are byte-identical to what `bucketmap -x` writes: tests/golden)."""
    import shutil
    import subprocess
    import tempfile
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import bm_ref
    from bucketmap_b200 import datagen
    L, n_bases = cfg["L"], cfg["n_bases"]
    work = tempfile.mkdtemp(prefix="bm_ref_bench_")
    try:
        fa = os.path.join(work, "genome.fa")
        datagen.write_fasta_fast(fa, packed, n_bases, "chr1")
        np.ascontiguousarray(rows).tofile(os.path.join(work, "idx.qgram"))
        with open(os.path.join(work, "idx.kmers_index"), "wb") as f:
            f.write(("\n".join(map(str, k2i.tolist())) + "\n").encode())
        with open(os.path.join(work, "idx.bucket_id"), "w") as f:
            f.write("chr1\n" * len(b_start))
        span = max(int(n_bases * n_per_proc / full_reads), 4 * L)
        span -= span % 4
        procs = max(1, min(procs, n_bases // span))
        runs = []
        env = dict(os.environ, BM_GENOME=fa, BM_RAND=f"{FMH_X - 1},{FMH_Y}")
        for pi in range(procs):
            d = os.path.join(work, f"p{pi}")
            os.makedirs(d)
            for ext in ("qgram", "kmers_index", "bucket_id"):
                os.symlink(os.path.join(work, "idx." + ext), os.path.join(d, "idx." + ext))
            lo = (n_bases - span) // max(1, procs - 1) * pi if procs > 1 else 0
            lo -= lo % 4
            seq, qual, off = datagen.simulate_reads(packed[lo // 4:], span, n_per_proc, cfg["read_len"], cfg["sub"],
                                                    cfg["indel"], cfg["indel"], READ_SEED + 1000 + pi)[:3]
            fq = os.path.join(d, "reads.fq")
            datagen.write_fastq_fast(fq, seq, qual, off)
            cmd = [bm_ref.binary(NB, L), "-i", "idx", "-q", fq, "-o", os.path.join(d, "out.sam"), "-f", repr(float(cfg["frac"]))]
            runs.append(subprocess.Popen(cmd, cwd=d, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True))
        t0 = time.perf_counter()
        logs = [r.communicate()[1] for r in runs]
        wall = time.perf_counter() - t0
        for r, lg in zip(runs, logs):
            if r.returncode != 0:
                raise RuntimeError("reference binary failed: " + lg[-1500:])
        ph = [bm_ref.parse_log(lg) for lg in logs]
        per = [x["map_s"] + x["kmer_index_s"] + x["find_offset_s"] for x in ph]
        t = max(per)
        return {"reads": procs * n_per_proc, "seconds": t, "procs": procs, "reads_per_s": procs * n_per_proc / t,
                "map_s": max(x["map_s"] for x in ph), "kmer_index_s": max(x["kmer_index_s"] for x in ph),
                "find_offset_s": max(x["find_offset_s"] for x in ph), "load_s": max(x["load_s"] for x in ph),
                "wall_s": wall, "span": span, "n_locs": sum(x["locations"] for x in ph)}
    finally:
        shutil.rmtree(work, ignore_errors=True)


def ref_procs():
    ncores = os.cpu_count() or 1
    try:
        import psutil
        by_mem = int(psutil.virtual_memory().available // (6 << 30))      # ~4 GB peak per reference process at C2
    except Exception:
        by_mem = 8
    return max(1, min(ncores, 16, by_mem))


def ref_sample_text(c):
    return (f"{c['procs']} single-threaded reference processes (oracle/_ref, the reference's own main.cpp) x "
            f"{c['reads'] // c['procs']} reads, each drawn from its own {c['span']}-base span (same reads per bucket as "
            f"the full batch), full index; slowest process: bucket mapping {c['map_s']:.2f}s + per-bucket k-mer index "
            f"{c['kmer_index_s']:.2f}s + offset vote {c['find_offset_s']:.2f}s (its own [BENCHMARK] timers; index load "
            f"{c['load_s']:.0f}s and FASTA parse not counted)")


def main():
    a = parse()
    cfg = CONFIGS[a.config]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    workload = f"{a.config}: {cfg['desc']}; {a.reads} reads per GPU per step"
    metric, unit = "reads/sec mapped", "reads/s"

    if a.impl == "reference":
        if rank != 0:
            return 0
        run_reference(a, cfg, workload, metric, unit)
        return 0

    import torch
    import torch.distributed as dist
    import bucketmap_b200 as bm
    from bucketmap_b200 import dist as bmd

    # libraries (NCCL) print banners on stdout: keep fd 1 for the one JSON line
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    all_cpus = os.sched_getaffinity(0)
    bind_to_gpu_numa_node(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = bm.Context(local)
    ctx.index_kmer_table(a.ktable)
    ctx.locator_index(a.lidx)
    ctx.pipeline_lanes(a.lanes)
    P = bm.default_params(ctx.lib, cfg["L"])
    q = 9

    # ---- reference + index: every rank uploads the (seeded) genome; rank 0 builds the matrix, NCCL replicates it
    t0 = time.perf_counter()
    packed, NB, b_start, b_len = make_workload(cfg, a.reads, GENOME_SEED)
    k2i, n_rows = fracminhash(q, FMH_X, FMH_Y, cfg["frac"])
    ctx.ref_upload(packed, cfg["n_bases"], b_start, b_len, NB)
    ctx.kernel_timing(True)
    if rank == 0:
        ctx.index_build(k2i, q, n_rows, download=False)
        ctx.sync()
    index_ms = ctx.kernel_stats()["index_build"][0]
    bcast_ms = None
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        tb = time.perf_counter()
        nbytes = bmd.broadcast_index(ctx, k2i, q, n_rows, NB, local)
        dist.barrier()
        bcast_ms = (time.perf_counter() - tb) * 1e3
    W = (NB + 7) >> 3

    # ---- reads of this rank (its contiguous shard of the job's N*reads reads)
    seq, qual, off = simulate(cfg, packed, a.reads, READ_SEED + rank)
    n_reads = len(off) - 1
    setup_s = time.perf_counter() - t0

    gatherer = None

    def barrier():
        if gatherer is not None:
            gatherer.wait()                      # collectives are issued by one thread at a time
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # resident steps leave their results on the device (as at N=1): gathered records stay on rank 0's device; the
    # end-to-end steps bring them to rank 0's host
    gatherer = bmd.LocationGatherer(host_copy=False) if world > 1 else None

    def gather():
        """N>1: location records of every rank -> rank 0 (NCCL gather; pinned host copy on a side stream)"""
        if world > 1:
            counts, locs = bmd.device_results(ctx, local)
            gatherer.gather(counts, locs)
        return None

    # ---- device-resident steps ------------------------------------------------------------------------------------
    ctx.reads_upload(P, seq, qual, off)

    trace = {} if os.environ.get("BM_GATHER_TRACE") else None

    def step_resident():
        if trace is None:
            ctx.map_device(P)
            ctx.locate_device(P, True)
            gather()
            return
        t0 = time.perf_counter(); ctx.map_device(P)
        t1 = time.perf_counter(); ctx.locate_device(P, True)
        t2 = time.perf_counter(); gather()
        t3 = time.perf_counter()
        for k, v in (("map", t1 - t0), ("locate", t2 - t1), ("gather", t3 - t2)):
            trace[k] = trace.get(k, 0.0) + v * 1e3

    for _ in range(a.warmup):
        step_resident()
    ctx.sync()
    ktable_ms = ctx.kernel_stats()["kmer_table"][0]     # one-off, inside the first (untimed) mapper call
    lidx_ms = ctx.kernel_stats()["locator_index"][0]    # one-off, inside the first (untimed) locator call
    ktable_k, ktable_bytes = ctx.index_kmer_table_info()
    ctx.kernel_stats_reset()
    if trace is not None:
        trace.clear(); bmd.TRACE.clear()
    barrier()
    clocks = ClockSampler(local)
    clocks.start()
    t_wall = time.perf_counter()
    ctx.timer_start()
    for _ in range(a.steps):
        step_resident()
    ms = ctx.timer_stop()                # CUDA events on the library's stream: every kernel of the path
    if world > 1:
        # the NCCL gather and the pinned copy of the gathered records run on other streams: time the region between
        # the two device-synchronised barriers instead
        gatherer.wait()
        barrier()
        ms = (time.perf_counter() - t_wall) * 1e3
    else:
        barrier()
    clk = clocks.stop()
    if trace is not None:
        print("[trace] rank %d host ms (warmup+timed): %s | gather phases: %s" % (rank, trace, bmd.TRACE), file=sys.stderr)
    kst = ctx.kernel_stats()
    _, _, mstats = ctx.map_download()
    tms = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_max = float(tms.item())
    value = world * n_reads * a.steps / (ms_max * 1e-3)

    # ---- end to end: pinned host buffers -> bm_map_locate -> host records --------------------------------------------
    hseq = ctx.pinned(len(seq), np.uint8); hseq[:] = seq
    hqual = ctx.pinned(len(qual), np.uint8); hqual[:] = qual
    hoff = ctx.pinned(len(off), np.uint64); hoff[:] = off
    ooff = ctx.pinned(n_reads + 1, np.uint32)
    olocs = ctx.pinned(max(4 * n_reads, 1024), bm.LOC_DTYPE)       # merged best locations: ~1 per read
    t_ooff = torch.from_numpy(ooff.view(np.int32))                   # the same pinned memory as torch tensors
    t_olocs = torch.from_numpy(olocs.view(np.uint8).reshape(-1, bm.LOC_DTYPE.itemsize))
    d2h = 0

    def gather_host(n_locs):
        """N>1: the location records of every rank go to rank 0 over NCCL (from the host arrays the API filled)"""
        if world > 1:
            counts = (t_ooff[1:] - t_ooff[:-1]).to(torch.int32).cuda(non_blocking=True)
            lb = t_olocs[:n_locs].cuda(non_blocking=True)
            gatherer.gather(counts, lb)
        return None

    if world > 1:
        gatherer = bmd.LocationGatherer(host_copy=True)
    for _ in range(max(1, min(a.warmup, 3))):
        n_locs, _ = ctx.map_locate_into(P, hseq, hqual, hoff, ooff, olocs, True)
    barrier()
    te = time.perf_counter()
    for _ in range(a.steps):
        n_locs, _ = ctx.map_locate_into(P, hseq, hqual, hoff, ooff, olocs, True)
        d2h = ooff.nbytes + n_locs * bm.LOC_DTYPE.itemsize + 64
        gather_host(n_locs)
    if world > 1:
        gatherer.wait()
    barrier()
    e2e_s = time.perf_counter() - te
    t2 = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
    e2e_value = world * n_reads * a.steps / float(t2.item())
    h2d = hseq.nbytes + hqual.nbytes + hoff.nbytes

    # ---- roofline of the dominant kernel -------------------------------------------------------------------------------
    vote_ms, vote_n = kst["map_vote"]

    # the full-width q-gram gather (k_map_vote) on the same batch, outside the timed region: the mapper when the coarse
    # table is off / does not fit, and the kernel the HBM roofline target of BASELINE.json was written for
    alt = None
    if ktable_k and rank == 0:
        ctx.index_kmer_table(0)
        ctx.reads_upload(P, seq, qual, off)              # the two-lane e2e calls leave no batch resident
        ctx.map_device(P)                                # warm-up (frees the table)
        ctx.kernel_stats_reset()
        for _ in range(3):
            ctx.map_device(P)
        ctx.sync()
        k2 = ctx.kernel_stats()
        _, _, ms2 = ctx.map_download()
        t_ms = k2["map_vote"][0] / max(1, k2["map_vote"][1])
        alt = {"kernel": "k_map_vote", "bound": "hbm", "achieved": int(ms2[4]) / (t_ms * 1e-3) / 1e9,
               "unit": "GB/s", "alg_bytes_per_launch": int(ms2[4]), "ms_per_launch": t_ms,
               "map_sample_ms_per_launch": k2["map_sample"][0] / max(1, k2["map_sample"][1]),
               "alg_bytes_definition": "16-B q-gram row loads issued by k_map_vote (counted in-kernel) until all 4096 "
                                       "buckets of a 512-B column tile have missed F samples"}
        ctx.index_kmer_table(a.ktable)
    rows_gathered = int(mstats[3])          # rows the reference's cascade ANDs (R of SURVEY 8d)
    alg_bytes = int(mstats[4])              # bytes of those rows K2's algorithm needs: 16-B loads it issued, counted
                                            # in the kernel (tiles whose buckets all reached F misses stop early)
    peak, peak_src = measured_peak()
    achieved = alg_bytes / (vote_ms / vote_n * 1e-3) / 1e9 if vote_n else None
    traffic = ncu_traffic(a.config, n_reads, ktable_k)
    rows_per_sample = 1 if ktable_k else 4
    launches = int(sum(v[1] for v in kst.values()))
    out = {
        "metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": ms_max / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32 bitwise", "data": "synthetic",
        "config": {"workload": workload, "genome_bases": cfg["n_bases"], "bucket_len": cfg["L"], "NB": NB,
                   "row_bytes": W, "matrix_bytes": int(n_rows) * W, "kmer_table_bytes": int(ktable_bytes),
                   "kmer_table": ("4^%d coarse k-mer rows (AND of each k-mer's q-gram rows, OR-reduced over groups of "
                                  "buckets), derived on the device from the matrix in %.1f ms; the mapper gathers 1 "
                                  "coarse row per sample and recounts passing groups exactly" % (ktable_k, ktable_ms))
                                 if ktable_k else "off: the mapper gathers the k-q+1 q-gram rows per sample",
                   "reads_per_gpu_per_step": n_reads,
                   "read_len": cfg["read_len"], "flags": "-k 9 -l 12 -s 15 -e 0.4 -p 10 -f %g" % cfg["frac"],
                   "parallelism": f"reads sharded x{world}, index replicated",
                   "timing": "CUDA events on the library stream" if world == 1 else "wall clock between device-synchronised barriers (NCCL send/recv of each batch's records to rank 0's device inside the step), max over ranks",
                   "l2": "inputs larger than L2 (bit matrix %d MB vs 126 MB L2)" % (n_rows * W >> 20),
                   "locator": "genome k-mer position index (K3i)" if a.lidx else "bucket text scan (K3)",
                   "step": "bm_map_device + bm_locate_device(filter_best=1)" + (" + NCCL gather" if world > 1 else "")},
        "e2e": {"value": e2e_value, "unit": unit, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "api": "bm_map_locate (pinned host buffers in, host records out; H2D chunks pipelined against the mapper inside the call)"},
        "gpu_launches": launches,
        "clocks": clk,
        "roofline": {"kernel": "k_map_fused" if ktable_k else "k_map_vote", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": (achieved / peak) if achieved else None, "traffic": traffic,
                     "traffic_gbs": (traffic / (vote_ms / vote_n * 1e-3) / 1e9) if (traffic and vote_n) else None,
                     "traffic_frac": (traffic / (vote_ms / vote_n * 1e-3) / 1e9 / peak) if (traffic and vote_n) else None,
                     "peak_source": peak_src, "alg_bytes_per_launch": alg_bytes,
                     "ms_per_launch": vote_ms / vote_n if vote_n else None,
                     "rows_gathered_per_read": rows_gathered / max(1, n_reads),
                     "reference_cascade_bytes_per_launch": rows_gathered * W,
                     "rows_per_sample": rows_per_sample,
                     "alg_bytes_definition": "16-B row loads issued by k_map_vote (counted in-kernel): %s of "
                                             "every 512-B column tile until all 4096 buckets of the tile have missed F "
                                             "samples; reference_cascade_bytes = all S*(k-q+1) rows of both strands x W"
                                             % ("the one k-mer-table row per sample" if ktable_k else "the q-gram rows")
                                             if not ktable_k else
                                             "bytes k_map_fused's loads ask for (counted in-kernel): the read's bases and "
                                             "qualities + S coarse k-mer rows per strand (row bytes in use, not the "
                                             "padded stride) + 4 B per .qgram word read for the exact recount of "
                                             "passing groups; after these reductions the kernel sits at ~0.7 of both "
                                             "limits at once (ncu: IPC 2.7 of 4 issue slots; DRAM traffic 19.6 kB per "
                                             "read = traffic_frac of the measured copy peak: scattered 4-B words and "
                                             "204-B rows cost whole sectors), see DESIGN.md section 4",
                     "reference_cascade_equivalent_gbs": (rows_gathered * W) / (vote_ms / vote_n * 1e-3) / 1e9 if vote_n else None},
        "roofline_qgram_path": dict(alt, peak=peak, frac=alt["achieved"] / peak) if alt else None,
        "kernels_ms_per_step": {k: v[0] / a.steps for k, v in kst.items() if v[1]},
        "index_build_ms": index_ms, "kmer_table_build_ms": ktable_ms, "locator_index_build_ms": lidx_ms,
        "index_bcast_ms": bcast_ms, "setup_s": setup_s,
        "mapped": {"reads_with_candidate": int(mstats[0]), "cand_fwd": int(mstats[1]), "cand_rev": int(mstats[2])},
    }
    if rank == 0 and world == 1 and not a.no_cpu:
        os.sched_setaffinity(0, all_cpus)            # the CPU baseline may use every host core
        rows = ctx.index_download()
        ncores = os.cpu_count() or 1
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import bm_ref
        if bm_ref.available(NB, cfg["L"]):
            c = reference_leg(cfg, packed, NB, b_start, k2i, rows, a.cpu_sample or 12_000, a.reads, ref_procs())
            out["cpu_baseline"] = {"value": c["reads_per_s"], "unit": unit, "cores": c["procs"], "kind": "reference",
                                   "sample": ref_sample_text(c)}
        else:
            n_sample = a.cpu_sample or max(2000, min(50_000, 600 * ncores))
            c = cpu_leg(cfg, packed, NB, b_start, b_len, k2i, rows, n_sample, a.reads, ncores)
            out["cpu_baseline"] = {"value": c["reads_per_s"], "unit": unit, "cores": c["threads"], "kind": "port",
                                   "sample": f"{c['reads']} reads drawn from the first {c['span']} bases (same reads per "
                                             f"bucket as the full batch), full index; map {c['map_s']:.2f}s + locate/merge "
                                             f"{c['locate_s']:.2f}s"}
    if rank == 0:
        os.write(real_stdout, (json.dumps(out) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()
    return 0


def fracminhash(q, x, y, frac):
    """tools/hash_function_generator.h:95-117 + indexer/bucket_indexer.h:147-159 (host logic of the product)"""
    from bucketmap_b200 import hostlogic
    return hostlogic.fracminhash(q, x, y, frac)


def run_reference(a, cfg, workload, metric, unit):
    """reference arm: the reference's own CPU implementation of the path on the host cores.  oracle/_ref (the
    reference's main.cpp built from /root/reference against oracle/shim) when the binary for this bucket count was
    prebuilt, else the oracle port with all host threads."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import bm_oracle
    import bm_ref
    bm_oracle.build()
    ncores = os.cpu_count() or 1
    o = bm_oracle.Oracle(ncores)
    packed, NB, b_start, b_len = make_workload(cfg, a.reads, GENOME_SEED)
    k2i, n_rows = o.fracminhash(9, FMH_X, FMH_Y, cfg["frac"])
    rows = o.index_build(packed, b_start, b_len, NB, 9, k2i, n_rows)   # setup only: == `bucketmap -x` (tests/golden)
    steps = max(1, a.steps)
    if bm_ref.available(NB, cfg["L"]):
        n_per = a.cpu_sample or max(2000, min(12_000, 36_000 // steps))
        c = reference_leg(cfg, packed, NB, b_start, k2i, rows, n_per * steps, a.reads, ref_procs())
        v = c["reads_per_s"]
        print(json.dumps({
            "impl": "reference", "metric": metric, "value": v, "unit": unit, "n_gpus": a.gpus, "steps": steps,
            "warmup": a.warmup, "ms_per_step": c["seconds"] / steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64 bitwise (std::bitset)", "data": "synthetic",
            "config": {"workload": workload, "NB": NB,
                       "note": "unmodified reference sources (oracle/_ref); a step = %d reads per process; one launch "
                               "covers all steps, warm-up reads are not separable inside the reference process" % n_per},
            "cpu_baseline": {"value": v, "unit": unit, "cores": c["procs"], "kind": "reference", "sample": ref_sample_text(c)},
            "e2e": {"value": v, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }), flush=True)
        return
    P = o.params(cfg["L"])
    n_sample = a.cpu_sample or max(2000, min(50_000, 600 * ncores))
    span = min(max(int(cfg["n_bases"] * n_sample / a.reads), 4 * cfg["L"]), cfg["n_bases"])
    seq, qual, off = simulate(cfg, packed, n_sample, READ_SEED + 1000, n_span=span)
    segs = o.segments(P, off)
    ix = o.index_open(P, NB, k2i, rows)

    def step():
        coff, cand, _ = o.map_segments_h(ix, P, seq, qual, off, segs)
        loff, locs = o.locate(P, packed, b_start, b_len, seq, qual, off, segs, coff, cand)
        return o.filter_best(P, off, loff, locs)

    for _ in range(min(a.warmup, 1)):
        step()
    steps = max(1, min(a.steps, 3))
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = time.perf_counter() - t0
    v = n_sample * steps / dt
    sample = (f"{n_sample} reads per step drawn from the first {span} bases (same reads per bucket as the full "
              f"batch), full index, {steps} timed steps")
    print(json.dumps({
        "impl": "reference", "metric": metric, "value": v, "unit": unit, "n_gpus": a.gpus, "steps": steps,
        "warmup": min(a.warmup, 1), "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64 bitwise", "data": "synthetic",
        "config": {"workload": workload, "NB": NB, "note": "CPU restatement of the reference (oracle/), all host threads"},
        "cpu_baseline": {"value": v, "unit": unit, "cores": o.threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


if __name__ == "__main__":
    sys.exit(main())
