#!/usr/bin/env python
"""bench.py -- overlapInCore hot path on N B200s of one node (contract: see the task statement / DESIGN.md section 6).

Workload (BASELINE.json's metric config: synthetic 100 Mbp genome, 30x corrected reads 5-15 kb, 2 % error, k=22):
one STEP = one overlapInCore job of that sweep = build the k-mer index of one 100 Mbp hash block (1x of the genome)
and overlap one 100 Mbp query block per GPU against it (every query ID is below the hash IDs, as in the
triangular sweep).  N > 1: weak scaling -- rank 0 builds the hash block, the table / occurrence array / flags are
replicated with an NCCL broadcast over NVLink, every rank scans and extends its own query block; no other
data-path collective.  `value` = query Mbp/s of the whole job (all ranks) with the packed reads resident in HBM;
`e2e` = the same job through the C ABI from HOST buffers (ASCII reads in pinned memory -> H2D + pack -> build ->
scan -> extend -> overlap records back on the host), every step.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--scale S]
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

GENOME = 100_000_000
BLOCK_BASES = 100_000_000      # canu's utg/obt hash block (-bl 100000000)
K = 22
ERATE = "0.06"
READ_ERR = 0.02
LEN_MIN, LEN_MAX = 5000, 15000
SEED = 1003
MODE = "blocks"                # "blocks": disjoint query block(s) vs one hash block; "self": -h 1-N -r 1-N on one read set
PARTIAL, MIN_OLAP, SUB_ONLY, HASH_BITS, REF_SCALE_MBP = 0, 500, False, 23, 20
REF_FULL = True                # reference arm / cpu_baseline run the FULL job of the step (False: a REF_SCALE_MBP-genome sample, c2 only)
REPEATS = None                 # c5: dict(frac, families, unit_min, unit_max, div) -- repeat model of tools/gen_reads.c
SKIP_T = 0                     # c5: K-mers seen >= SKIP_T times in the loaded reads (both strands) go to the -k skip list

# BASELINE.json's named shapes.  The driver's bench line is c3 (the config the metric is quoted on); the others are run by
# hand (`--workload`) for the per-shape table of DESIGN.md section 7 and are parity-test cases at reduced size in tests/.
WORKLOADS = {
    "c1": dict(desc="c1: synthetic 4.6 Mbp genome, 30x corrected reads 5-15 kb, 2% error, k=22, --maxerate 0.06, one hash block vs one "
                    "query block (the same 13.8 k reads: -h 1-N -r 1-N)",
               GENOME=4_600_000, BLOCK_BASES=138_000_000, ERATE="0.06", READ_ERR=0.02, SEED=1001, MODE="self", REF_SCALE_MBP=0.3),
    "c2": dict(desc="c2: synthetic 4.6 Mbp genome, 30x raw-noise reads 5-15 kb, 6% per read = 12% pairwise, k=22, --maxerate 0.15 "
                    "(wide-band extension), -h 1-N -r 1-N",
               GENOME=4_600_000, BLOCK_BASES=138_000_000, ERATE="0.15", READ_ERR=0.06, SEED=1002, MODE="self", REF_SCALE_MBP=0.1,
               REF_FULL=False),   # the full c2 job takes the reference ~20 min on 16 cores
    "c3": dict(desc="c3: synthetic 100 Mbp genome, 30x corrected reads 5-15 kb, 2% error, k=22, --maxerate 0.06, "
                    "100 Mbp hash block (1x) replicated, one 100 Mbp query block per GPU"),
    "c4": dict(desc="c4: synthetic 100 Mbp genome, 150 bp reads (substitutions only, 0.5%), -G --minlength 40, k=22, --maxerate 0.06, "
                    "100 Mbp hash block (667 k reads, 1x) vs one 100 Mbp query block per GPU",
               LEN_MIN=150, LEN_MAX=150, READ_ERR=0.005, SUB_ONLY=True, PARTIAL=1, MIN_OLAP=40, HASH_BITS=24, SEED=1004, REF_SCALE_MBP=5),
    "c5": dict(desc="c5: synthetic 1 Gbp repeat-rich genome (30% of it copies of 500 repeat families, 1-6 kb units, 2% divergence), corrected "
                    "reads 5-15 kb, 2% error, k=22, --maxerate 0.06, frequent k-mers (>= 4 occurrences in the 0.2x of reads loaded, i.e. copy "
                    "number >~ 30) on a -k skip list, 100 Mbp hash block (0.1x) vs one 100 Mbp query block per GPU",
               GENOME=1_000_000_000, SEED=1005, REPEATS=dict(frac=0.3, families=500, unit_min=1000, unit_max=6000, div=0.02), SKIP_T=4,
               REF_SCALE_MBP=300),
}


_DEFAULTS = None


def set_workload(name):
    """make `name` (a key of WORKLOADS) the module's current shape: the generator / flag globals above"""
    global _DEFAULTS
    g = globals()
    if _DEFAULTS is None:
        _DEFAULTS = {k: g[k] for k in ("GENOME", "BLOCK_BASES", "ERATE", "READ_ERR", "LEN_MIN", "LEN_MAX", "SEED", "MODE", "PARTIAL", "MIN_OLAP",
                                       "SUB_ONLY", "HASH_BITS", "REF_SCALE_MBP", "REF_FULL", "REPEATS", "SKIP_T")}
    g.update(_DEFAULTS)
    g.update({k: v for k, v in WORKLOADS[name].items() if k != "desc"})
    return WORKLOADS[name]


def params_for(n_hash_reads, n_total, threads=1):
    import helpers as H
    common = dict(kmer_len=K, max_erate=H.strtof(ERATE), min_olap_len=MIN_OLAP, partial=PARTIAL, hash_bits=HASH_BITS, hash_load=0.75,
                  max_hash_strings=n_hash_reads, max_hash_data_len=2 * BLOCK_BASES, num_threads=threads)
    if MODE == "self":
        return H.default_params(bgn_hash_id=1, end_hash_id=n_total, bgn_ref_id=1, end_ref_id=n_total, **common)
    return H.default_params(bgn_hash_id=n_total - n_hash_reads + 1, end_hash_id=n_total, bgn_ref_id=1,
                            end_ref_id=n_total - n_hash_reads, **common)


def gen_blocks(genome_len, n_blocks, block_bases, seed):
    """reads for n_blocks blocks of ~block_bases each, sampled from one genome (each block = block_bases/genome coverage)"""
    import helpers as H
    cov = n_blocks * block_bases / genome_len
    rep = {}
    if REPEATS:   # the number of families follows the genome so that copies per family stay the same on a scaled-down sample
        rep = dict(repeat_frac=REPEATS["frac"], repeat_families=max(1, int(round(REPEATS["families"] * genome_len / GENOME))),
                   repeat_min=REPEATS["unit_min"], repeat_max=REPEATS["unit_max"], repeat_div=REPEATS["div"])
    return H.gen_reads(genome_len, cov, LEN_MIN, LEN_MAX, READ_ERR, seed, sub_only=SUB_ONLY, **rep)


def gen_job(genome_len, world, block_bases, seed):
    """(reads, blocks) of the step at `world` GPUs: `world` query blocks + the hash block, IDs laid out [Q0 .. Q(world-1) | H] (every
    query ID below every hash ID).  The generator makes read i from (seed, i) alone, so the job is the SAME at every N where it can
    be: H = generated reads [0, per), Q_r = generated reads [(r+1) per, (r+2) per), per = the reads of one block -- rank 0 of any
    run searches the job of the 1-GPU run, rank r the same block at every N > r.  (Until r2p the hash block was the LAST block of an
    (N+1)-block set: a different hash block and a different block 0 at every N, 12.8 to 13.6 G DP cells on rank 0, which the
    driver's N x (1-GPU rate) comparison booked as scaling loss or gain.)"""
    from canu_b200 import shard
    g = gen_blocks(genome_len, world + 1, block_bases, seed)
    per = int(block_bases / (0.5 * (LEN_MIN + LEN_MAX)) + 0.5)
    n = min(g.n, per * (world + 1))
    cut, end = int(g.offsets[per]), int(g.offsets[n])
    import helpers as H
    lengths = np.concatenate([g.lengths[per:n], g.lengths[:per]])
    offsets = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(lengths, out=offsets[1:])
    bases = np.empty(end, dtype=np.uint8)
    bases[:end - cut] = g.bases[cut:end]
    bases[end - cut:] = g.bases[:cut]
    blocks = [(r * per + 1, min((r + 1) * per, n - per)) for r in range(world)] + [(n - per + 1, n)]
    return H.Reads(bases, offsets, lengths), blocks


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons of one GPU DURING the timed region (the recipe's clocks line).  Read through NVML inside this
    process (nvidia_ml_py: one nvmlInit, then two cheap queries per sample): forking `nvidia-smi` 5x a second attaches a new NVML
    client to every GPU of the box each time, and the step of a run next to it showed gaps of 5-35 ms between kernels (r2q: extend
    stage 75.6 / 100.4 ms around the same 65.9 ms of kernels).  OIC_CLOCK_SAMPLER=smi|nvml|off selects; `smi` is the fallback
    when NVML cannot be loaded."""

    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.stop_flag = False
        self.rows = []          # (sm MHz, max MHz, [hw_slowdown, hw_thermal, sw_thermal, sw_power_cap])
        self.mode = os.environ.get("OIC_CLOCK_SAMPLER", "nvml")
        self.nv, self.h = None, None
        if self.mode == "nvml":
            try:
                import pynvml
                pynvml.nvmlInit()
                # NVML indexes the physical GPUs: follow CUDA_VISIBLE_DEVICES when it is a plain list of indices
                vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
                phys = gpu_index
                if vis and all(x.strip().isdigit() for x in vis.split(",")):
                    phys = int(vis.split(",")[gpu_index])
                self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
                self.nv = pynvml
                self.gpu = phys
            except Exception:
                self.mode = "smi"

    def sample_nvml(self):
        nv = self.nv
        sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
        mx = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        bits = [nv.nvmlClocksEventReasonHwSlowdown, nv.nvmlClocksEventReasonHwThermalSlowdown, nv.nvmlClocksEventReasonSwThermalSlowdown,
                nv.nvmlClocksEventReasonSwPowerCap]
        self.rows.append((float(sm), float(mx), [bool(r & b) for b in bits]))

    def sample_smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        out = subprocess.check_output(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                      stderr=subprocess.DEVNULL, timeout=5).decode().strip()
        r = [x.strip() for x in out.split(",")]
        self.rows.append((float(r[0]), float(r[1]), [x.lower().startswith("active") for x in r[2:6]]))

    def run(self):
        if self.mode == "off":
            return
        while not self.stop_flag:
            try:
                self.sample_nvml() if self.mode == "nvml" else self.sample_smi()
            except Exception:
                if self.mode == "nvml" and not self.rows:
                    self.mode = "smi"    # an NVML build without these queries: fall back to the command-line tool
            time.sleep(0.05 if self.mode == "nvml" else 0.2)

    def summary(self):
        sm = [r[0] for r in self.rows]
        mx = [r[1] for r in self.rows]
        reasons = sorted({self.NAMES[i] for r in self.rows for i in range(4) if r[2][i]})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm), "source": {"nvml": "NVML in-process, 20 Hz", "smi": "nvidia-smi, 5 Hz", "off": "disabled"}[self.mode]}


def reference_records(store, ovb):
    """sorted `overlapConvert -raw` lines of a reference .ovb (the parity comparator of SURVEY.md 8(c))"""
    import helpers as H
    return H.dump_ovb_sorted(store, ovb)


def executable_leg(exe, store, wd, args, cores, qbases, runs=2):
    """SURVEY.md 8(d)'s own definition of the metric, through the DROP-IN EXECUTABLE: wall time of one `overlapInCore` process from
    opening the gkpStore to `.ovb` / `.counts` / `.stats` closed (CUDA start-up, store read, 2-bit reads to the GPU, build, search,
    snappy-framed output), on the store the reference leg just wrote.  The first run also pays the cold start of the driver; both
    are reported, the rate is of the last."""
    import helpers as H
    walls, ovb, st = [], None, None
    for i in range(runs):
        t0 = time.time()
        ovb, st = H.run_overlapper(exe, store, os.path.join(wd, "exe%d" % i), args, cores)
        walls.append(time.time() - t0)
    n = int(H.parse_stats(st).get("Total overlaps produced", 0))
    w = walls[-1]
    return {"what": "canu_b200/bin/overlapInCore, one process, gkpStore in -> .ovb/.counts/.stats closed (wall)", "wall_s": w,
            "wall_s_all_runs": walls, "overlaps": n, "overlaps_per_sec": n / w, "query_mbp_per_sec": qbases / 1e6 / w,
            "ovb_bytes": os.path.getsize(ovb), "ovb": ovb}


def reference_leg(world=1, steps=2, warmup=1, cores=None, budget_s=None, keep=None, exe=None):
    """The reference's own CPU overlapInCore (oracle/_ref, built from /root/reference/src by oracle/build_ref.sh) on the SAME job
    the GPU arm times at `world` GPUs: same generator seed, the full hash block, and the same `world` query blocks searched
    by ONE reference process (-h <hash block> -r 1-<hash block - 1>), -t <all host cores>; wall time of the process (store open
    + hash build + search + .ovb/.counts written).  The steps/warm-ups actually run are bounded by `budget_s` of host time
    (default 300 s, OIC_REF_BUDGET_S) and REPORTED: the first run always counts unless two more fit the budget, in which case it
    is the warm-up.  Workloads with REF_FULL = False (c2) run a REF_SCALE_MBP-genome sample of the same read model instead.
    keep: optional dict that receives store / ovb paths of the last run (bench.py --check); the work directory then stays."""
    import helpers as H
    cores = cores or os.cpu_count()
    budget_s = float(os.environ.get("OIC_REF_BUDGET_S", "300")) if budget_s is None else budget_s
    from canu_b200 import shard
    t_start = time.time()
    full = bool(REF_FULL)
    g = GENOME if full else int(REF_SCALE_MBP * 1_000_000)
    blk = BLOCK_BASES if full else int(g * BLOCK_BASES / GENOME)
    seed = SEED if full else SEED + 7
    if MODE == "self":
        reads = gen_blocks(g, 1, blk, seed)
        qbases = int(reads.lengths.sum())
        P = params_for(reads.n, reads.n, threads=cores)
        sample = ("reference overlapInCore -t %d, %s: %d reads (%.1f Mbp) against themselves (-h 1-N -r 1-N)"
                  % (cores, "the full job of the step" if full else "genome %.1f Mbp at the same coverage, same read model/flags" % (g / 1e6),
                     reads.n, qbases / 1e6))
    else:
        reads, blocks = gen_job(g, world, blk, seed)
        hb, he = blocks[-1]
        qbases = int(reads.lengths[:hb - 1].sum())
        P = params_for(he - hb + 1, reads.n, threads=cores)
        sample = ("reference overlapInCore -t %d, %s: hash block %d reads (%.1f Mbp) x %d query block(s) %d reads (%.1f Mbp) in one process "
                  "(-h %d-%d -r 1-%d)" % (cores, "the full job of the step (same seed, same reads as the GPU arm)" if full else
                                          "genome %.1f Mbp sample, same read model/flags" % (g / 1e6),
                                          he - hb + 1, reads.lengths[hb - 1:he].sum() / 1e6, world, hb - 1, qbases / 1e6, hb, he, hb - 1))
    kind = "reference" if H.have_ref() else "port"
    times, n_ovl, n_warm = [], 0, 0
    wd = tempfile.mkdtemp(prefix="oic_ref_")
    try:
        if kind == "reference":
            store = H.make_store(reads, wd)
            args = H.oic_args(P, ERATE)
            if SKIP_T:
                import cases
                sf = os.path.join(wd, "skip.fasta")
                cases.write_skip_file(H.frequent_kmers_fast(reads, K, SKIP_T), sf)
                args += ["-k", sf]
            prep_s = time.time() - t_start

            def one(i):
                t0 = time.time()
                ovb, st = H.run_overlapper(os.path.join(H.REF_DIR, "overlapInCore"), store, os.path.join(wd, "ref%d" % i), args, cores)
                return time.time() - t0, ovb, st
            dt, ovb, st = one(0)
            times.append(dt)
            # how many more runs fit the budget (projected from the first)
            more = int(max(0.0, budget_s - (time.time() - t_start)) // max(dt, 1e-3))
            more = min(more, max(0, steps - 1) + (1 if warmup > 0 else 0))
            for i in range(more):
                dt, ovb, st = one(i + 1)
                times.append(dt)
            if warmup > 0 and len(times) >= 3:
                n_warm, times = 1, times[1:]
            times = times[:max(1, steps)]
            n_ovl = H.parse_stats(st).get("Total overlaps produced", 0)
            exe_res = None
            if exe and os.access(exe, os.X_OK):
                try:
                    exe_res = executable_leg(exe, store, wd, args, cores, qbases)
                    # its records against the reference's, every stored field (the files are small next to the search)
                    exe_res["equals_reference_ovb"] = H.dump_ovb_sorted(store, exe_res.pop("ovb")) == H.dump_ovb_sorted(store, ovb)
                except Exception as ex:   # the line must not fail because of the extra leg
                    exe_res = {"error": "%s: %s" % (type(ex).__name__, ex)}
            if keep is not None:
                keep.update(store=store, ovb=ovb, stats=st, workdir=wd)
        else:
            cores, prep_s = 1, 0.0
            t0 = time.time()
            res = H.run_oracle(reads, P, skip_kmers=H.frequent_kmers_fast(reads, K, SKIP_T) if SKIP_T else None)
            times.append(time.time() - t0)
            n_ovl = len(res["overlaps"])
    finally:
        if keep is None:
            subprocess.call(["rm", "-rf", wd])
    t = float(np.mean(times))
    sample += "; %d timed run(s) after %d warm-up, %.1f s each (store written in %.1f s, not timed)" % (len(times), n_warm, t, prep_s)
    return dict(value=qbases / 1e6 / t, unit="query Mbp/s", cores=cores, kind=kind, sample=sample, full_job=full,
                seconds_per_step=t, overlaps_per_sec=n_ovl / t, overlaps=int(n_ovl), steps=len(times), warmup=n_warm,
                executable=exe_res if kind == "reference" else None)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--scale", type=float, default=1.0, help="shrink genome and blocks (debug only; 1.0 = the named config)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-input", default="2bit", choices=["2bit", "ascii"],
                    help="host buffers of the end-to-end leg: the gkpStore's 2-bit read encoding (oic_load_reads_2bit) or ASCII (oic_load_reads)")
    ap.add_argument("--replicate", default="sharded", choices=["sharded", "broadcast", "rebuild"],
                    help="N > 1: every rank builds 1/N of the hash table and the slices are exchanged by ncclAllGather (default); rank 0 "
                         "builds and ncclBroadcasts the block; or every rank builds its own copy (no collective)")
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS), help="BASELINE.json shape; c3 is the bench line")
    ap.add_argument("--shard", default="interleaved", choices=["interleaved", "blocks"],
                    help="N > 1: how the N query blocks are dealt to the ranks: read by read (default: every rank gets the same mix -- the step "
                         "is the max over ranks and whole blocks differ by up to 6 per cent in work; 78.1 vs 79.3 ms at 4 GPUs) or one "
                         "contiguous block per rank (the reference's own partition).  The job and its record set are the same either way")
    ap.add_argument("--as-rank", default=None, metavar="R/W",
                    help="diagnostic: ONE GPU runs the job that rank R of a W-rank run would run (its query block, the same hash block; no "
                         "collective) -- separates the spread between the ranks' query blocks from anything the box does to a multi-GPU run")
    ap.add_argument("--check", action="store_true",
                    help="after the timed runs: gather every rank's records on the host of rank 0 and compare the multiset (all stored fields) "
                         "with the reference binary run on the same job (exit 1 on any difference)")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    as_rank = None
    if a.as_rank:
        assert world == 1, "--as-rank is a single-GPU diagnostic"
        as_rank = tuple(int(x) for x in a.as_rank.split("/"))

    wl = set_workload(a.workload)
    if MODE == "self" and world > 1:
        raise SystemExit("workload %s is a single-GPU shape (one hash block vs one query block)" % a.workload)
    metric = "overlapInCore query Mbp/sec (+overlaps/sec), 30x 100 Mbp genome, one hash block x one query block per GPU per step"
    if a.workload != "c3":
        metric = "overlapInCore query Mbp/sec (+overlaps/sec), workload %s" % a.workload
    config = {"workload": wl["desc"],
              "hash_block_mbp": BLOCK_BASES * a.scale / 1e6, "query_block_mbp_per_gpu": BLOCK_BASES * a.scale / 1e6,
              "parallelism": "query blocks sharded over %d GPU(s)%s, hash block replicated%s" % (
                  world, "" if world == 1 else (" read by read (rank r takes every N-th query read)" if a.shard == "interleaved" else " one block per rank"),
                  "" if world == 1 else {"sharded": " by NCCL: every rank builds 1/N of the table, ncclAllGather of the slices (oic_build_hash_index_sharded)",
                                                "broadcast": " by ncclBroadcast from rank 0 (oic_hash_block_replicate)",
                                                "rebuild": " by rebuilding it on every rank (no collective)"}[a.replicate]),
              "l2": "inputs larger than L2 (k-mer table + occurrence array > 1 GB per block); no flush needed",
              "e2e_input": ("gkpStore 2-bit encoded reads (2SEQ chunks, ST/gkStoreEncode.C) in pinned host memory -> oic_load_reads_2bit"
                            if a.e2e_input == "2bit" else "ASCII reads in pinned host memory -> oic_load_reads")}
    if a.scale != 1.0:
        config["scale"] = a.scale

    if a.impl == "reference":
        # the reference's own CPU program on the SAME job (same seed, full blocks, the same a.gpus query blocks in one process);
        # rank 0 alone runs it.  A step is ~30-250 s of host time, so the runs that fit OIC_REF_BUDGET_S are made and the counts
        # actually run are what the line reports (steps_requested / warmup_requested keep the command line's).
        if rank != 0:
            return
        if a.scale != 1.0:
            globals().update(GENOME=int(GENOME * a.scale), BLOCK_BASES=int(BLOCK_BASES * a.scale))
        r = reference_leg(max(a.gpus, 1), max(1, min(a.steps, 2)), min(a.warmup, 1))
        line = {"impl": "reference", "metric": metric, "value": r["value"], "unit": r["unit"], "n_gpus": a.gpus, "steps": r["steps"],
                "warmup": r["warmup"], "steps_requested": a.steps, "warmup_requested": a.warmup,
                "ms_per_step": 1000 * r["seconds_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u8/int32 (2-bit packed bases, int32 edit rows, f64 branch score)", "data": "synthetic",
                "config": config, "same_config": bool(r["full_job"]),
                "overlaps_per_sec": r["overlaps_per_sec"], "overlaps_per_step": r["overlaps"],
                "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
                "e2e": {"value": r["value"], "unit": r["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    import torch
    import helpers as H
    from canu_b200 import api
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path)"
    torch.cuda.set_device(local)
    dist, host_group = None, None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        host_group = dist.new_group(backend="gloo")   # host-side gather of the overlap records (no GPU collective on them)

    genome = int(GENOME * a.scale)
    block = int(BLOCK_BASES * a.scale)
    from canu_b200 import shard
    if MODE == "self":
        # one read set against itself (-h 1-N -r 1-N): pairs with query ID < hash ID
        reads = gen_blocks(genome, 1, block, SEED)
        n_q, n_h = reads.n, reads.n
        q_bases_rank = q_bases_all = int(reads.lengths.sum())
        LP = params_for(reads.n, reads.n)
        total = q_bases_all
        pinned = torch.empty(total, dtype=torch.uint8).pin_memory()
        pinned.numpy()[:] = reads.bases[:total]
        offs, lens = reads.offsets[:-1].astype(np.uint64), reads.lengths.astype(np.uint32)
        rj = None
        hash_range, query_range = (1, reads.n), (1, reads.n)
    else:
        # reads: `world` query blocks + 1 hash block, all sampled from the same genome (every rank generates the same set)
        gen_world, gen_rank = (as_rank[1], as_rank[0]) if as_rank else (world, rank)
        reads, blocks = gen_job(genome, gen_world, block, SEED)
        rj = shard.RankJob(blocks, gen_rank, interleave=(a.shard == "interleaved"))
        hb, he, qb, qe, n_q, n_h = rj.hb, rj.he, rj.qb, rj.qe, rj.n_q, rj.n_h
        q_bases_rank = int(reads.lengths[rj.q_ids - 1].sum())
        q_bases_all = q_bases_rank if as_rank else int(reads.lengths[:hb - 1].sum())
        if as_rank:
            config["as_rank"] = "the job of rank %d of a %d-rank run, alone on one GPU (diagnostic)" % as_rank
        P = params_for(he - hb + 1, reads.n)
        # this rank needs its query block and the hash block, gathered into one pinned host buffer in a local ID space
        # (queries 1..n_q, hash n_q+1..n_q+n_h: query ID < hash ID, as in the sweep); canu_b200/shard.py maps records back
        total = int(reads.lengths[rj.ids - 1].sum())
        pinned = torch.empty(total, dtype=torch.uint8).pin_memory()
        _, offs, lens = rj.select(reads.bases, reads.offsets[:-1], reads.lengths, out=pinned.numpy())
        LP = rj.local_params(P)
        hash_range, query_range = (n_q + 1, n_q + n_h), (1, n_q)

    # host buffers of the end-to-end leg: the reads as the gkpStore holds them (2SEQ chunks: 4 bases per byte), pinned
    lens64 = np.asarray(lens, dtype=np.int64)
    pinned2 = torch.empty(int(((lens64 + 3) // 4).sum()) + 8, dtype=torch.uint8).pin_memory()
    _, boffs = H.encode_2bit(pinned.numpy(), offs, lens, out=pinned2.numpy())

    ov = api.Overlapper(LP, device=local)
    n_skip, skip_info = 0, {}
    if world > 1 and a.replicate != "rebuild":
        # the library's own NCCL communicator (oic_comm_init_rank): rank 0 draws the id, torch.distributed only ships its 128 bytes
        idt = torch.zeros(api.COMM_ID_BYTES, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(api.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        ov.comm_init_rank(bytes(idt.cpu().numpy().tobytes()), world, rank)

    trace = os.environ.get("OIC_BENCH_TRACE")
    last_records = [None]   # rank 0: the gathered records of the last end-to-end step (bench.py --check)

    def job(load):
        """one step; returns number of overlap records of this rank"""
        t = [time.perf_counter()]
        if load:
            if a.e2e_input == "2bit":
                ov.load_reads_2bit(pinned2.data_ptr(), boffs, lens, first_id=1)
            else:
                ov.load_reads_ptr(pinned.data_ptr(), offs, lens, first_id=1)
        ov.clear_overlaps()
        t.append(time.perf_counter())
        if world == 1 or a.replicate == "rebuild":
            ov.build_hash_index(*hash_range)
        elif a.replicate == "sharded":
            ov.build_hash_index_sharded(*hash_range)      # every rank builds 1/N of the table; ncclAllGather inside the library
        else:
            if rank == 0:
                ov.build_hash_index(*hash_range)
            t.append(time.perf_counter())
            ov.hash_block_replicate(0)                    # ncclBroadcast of the built block inside the library
        t.append(time.perf_counter())
        ov.process_overlaps(*query_range)
        n = ov.num_overlaps()
        t.append(time.perf_counter())
        if load:
            rec = ov.overlaps()   # records back in a host buffer (D2H inside the library)
            if world > 1:         # per-GPU records merged on the HOST of rank 0, in store IDs (gloo; nothing goes back to a GPU)
                last_records[0] = shard.gather_records(rj.to_global(rec), dist, group=host_group, dst=0)
            else:
                last_records[0] = rj.to_global(rec) if rj is not None else rec
        t.append(time.perf_counter())
        if trace:
            sys.stderr.write("[rank %d] step phases ms: %s\n" % (rank, " ".join("%.1f" % (1000 * (b - a_)) for a_, b in zip(t, t[1:]))))
        return n

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(nsteps, load):
        barrier()
        t0 = time.perf_counter()
        n = 0
        for _ in range(nsteps):
            n = job(load)
        barrier()
        dt = time.perf_counter() - t0
        if dist is not None:
            tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = float(tt.item())
            nn = torch.tensor([n], dtype=torch.int64, device="cuda")
            dist.all_reduce(nn)
            n = int(nn.item())
        return dt, n

    ov.load_reads_ptr(pinned.data_ptr(), offs, lens, first_id=1)
    if SKIP_T:   # the -k skip list from the reads this rank holds (its query block + the hash block), counted on the GPU
        keys, _, kinfo = ov.count_kmers(K, SKIP_T)   # what `meryl -B -C` + `meryl -Dt -n T` give the pipeline (oic_count_kmers)
        skip = H.kmer_strings(keys, K)
        n_skip = len(skip)
        skip_info = {"skip_kmers": n_skip,
                     "skip_list": "oic_count_kmers on the resident reads: %.1f ms on the GPU, %d pass(es)" % (kinfo["device_ms"], kinfo["passes"])}
        ov.set_skip_kmers(skip)
    for _ in range(a.warmup):
        job(False)
    sampler = ClockSampler(local)
    sampler.start()
    ov.reset_timings()
    st0 = ov.stats()
    dt, n_ovl = timed(a.steps, False)
    tm = ov.timings()
    st1 = ov.stats()
    job(True)   # warm the end-to-end path once
    ov.reset_timings()
    dt_e2e, n_ovl_e2e = timed(a.steps, True)
    tm_e2e = ov.timings()
    sampler.stop_flag = True
    sampler.join(timeout=2)
    ov.close()

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    d = {k: st1[k] - st0[k] for k in st0 if isinstance(st0[k], int)}
    steps = a.steps
    # algorithmic bytes: B_scan = 0.25 N_qb + 4 N_win + 8 N_found + 8 N_ref ; B_build = 0.25 N_b + 16 N_ins
    b_scan = (0.25 * d["n_query_bases"] + 4 * d["n_windows"] + 8 * d["n_found"] + 8 * d["n_refs"]) / steps
    b_build = (0.25 * d["n_hash_bases"] + 16 * d["n_inserts"]) / steps
    i_ext = (6 * d["ext_cells"] + 4 * d["ext_words"]) / steps
    ms_scan, ms_build, ms_ext, ms_group = tm["ms_scan"] / steps, tm["ms_build"] / steps, tm["ms_extend"] / steps, tm["ms_group"] / steps
    scan_gbs = b_scan / (ms_scan * 1e-3) / 1e9 if ms_scan > 0 else 0.0
    build_gbs = b_build / (ms_build * 1e-3) / 1e9 if ms_build > 0 else 0.0
    ext_gops = i_ext / (ms_ext * 1e-3) / 1e9 if ms_ext > 0 else 0.0
    # the dominant kernel is the speculative extension (k_extend_tasks<true>, ~70 % of the step): INT32-issue bound.
    # achieved = algorithmic integer ops (6 per DP cell + 4 per 16-base compare word) / its own launch time;
    # peak = INT32 ALU rate measured on this GPU by oic_measure_int32_peak (LOP3+IADD3 chains), nominal 148*64*clock as fallback
    try:
        alu_peak, imad_peak = api.measure_int32_peak(local)
        int_src = "measured (oic_measure_int32_peak: LOP3+IADD3 chains; IMAD chains %.0f Gop/s)" % imad_peak
    except Exception:
        alu_peak, int_src = 148 * 64 * 1.965, "nominal 148 SM x 64 lanes x 1.965 GHz"
    ms_tasks = tm["ms_ext_tasks"] / steps
    ms_first, ms_commit = tm["ms_ext_first"] / steps, tm["ms_ext_commit"] / steps
    # cells / words of the first extensions run in k_process_pairs; the task kernel does the rest (ext_calls - n_pairs of ext_calls)
    share = 1.0 - (d["n_pairs"] / max(1, d["ext_calls"]))
    i_tasks = i_ext * share
    tasks_gops = i_tasks / (ms_tasks * 1e-3) / 1e9 if ms_tasks > 0 else 0.0
    # dram read+write bytes per launch and pipe utilisation: from the committed summary of the newest `ncu --set full`
    # capture of this command (profiles/*_kernels.json, written by tools/ncu_summary.py)
    traffic, ncu_k, ncu_src = None, {}, None
    narrow = not (H.strtof(ERATE) > 0.06) and os.environ.get("OIC_LEAN_WIDE") is None
    lean_prefix = "k_extend_tasks<"
    lean_name = None
    try:
        import glob
        # the newest committed summary (files are named per round: r1e_, r2a_, ... sort in that order) that CONTAINS the lean task kernel
        for cand in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_kernels.json")), reverse=True):
            ks = {}
            for k in json.load(open(cand)).get("kernels", []):
                ks.setdefault(k["kernel"], k)
            names = [n for n in ks if n.startswith(lean_prefix) and not n.startswith("k_extend_tasks<0")]
            if names:
                ncu_k, ncu_src = ks, cand
                lean_name = max(names, key=lambda n: ks[n].get("ms", 0.0))
                traffic = ks[lean_name]["dram_bytes_per_launch"]
                break
    except Exception:
        pass
    kernel_label = lean_name or ("k_extend_tasks (lean row loop, %s staging instance)" % ("narrow" if narrow else "wide"))

    def ncu_of(name):
        k = ncu_k.get(name)
        if not k:   # template instances: "k_insert" finds "k_insert<0>"
            k = next((v for n, v in sorted(ncu_k.items()) if n.startswith(name + "<")), None)
        if not k:
            return None
        return {x: k.get(x) for x in ("ms", "dram_bytes_per_launch", "dram_GBps", "issue_active_pct", "alu_pipe_pct", "lsu_pipe_pct",
                                      "warps_active_pct", "l1_hit_pct", "l2_hit_pct", "sectors_per_ld_request")}
    roof = {"kernel": kernel_label + " (lean row loop: rows staged in shared memory, template args = staging width, blocks per SM)", "bound": "int32", "achieved": tasks_gops, "peak": alu_peak, "unit": "Gop/s",
            "frac": tasks_gops / alu_peak if alu_peak else None, "traffic": traffic, "peak_source": int_src,
            "algorithmic_ops_per_launch": i_tasks, "ms_per_launch": ms_tasks,
            "formula": "6*C + 4*W (C = DP cells, W = 16-base compare words) x share of extension calls run by this kernel; "
                       "not an HBM- or tensor-bound kernel: rows live in shared memory, sequences in L1/L2",
            "ncu": ncu_of(lean_name), "ncu_source": os.path.relpath(ncu_src, ROOT) if ncu_src else None,
            "hbm_side": {"kernel": "k_seed_scan", "bound": "hbm", "achieved": scan_gbs, "peak": hbm_peak, "unit": "GB/s",
                         "frac": scan_gbs / hbm_peak, "peak_source": peak_src, "algorithmic_bytes_per_launch": b_scan,
                         "ms_per_launch": ms_scan, "traffic": (ncu_k.get("k_seed_scan") or {}).get("dram_bytes_per_launch"),
                         "ncu": ncu_of("k_seed_scan"),
                         "formula": "0.25*N_query_bases + 4*N_windows + 8*N_found + 8*N_refs (SURVEY.md 8(d)), rank 0's block"}}
    stages = {"build": {"ms": ms_build, "algorithmic_GBps": build_gbs, "frac_hbm": build_gbs / hbm_peak,
                        "formula": "0.25*N_hash_bases + 16*N_inserts; whole build (table init + insert + list kernels + occupancy bitmap)",
                        "ncu_k_insert": ncu_of("k_insert")},
              "seed_scan": {"ms": ms_scan, "algorithmic_GBps": scan_gbs, "frac_hbm": scan_gbs / hbm_peak},
              "group_runs": {"ms": ms_group},
              "extend": {"ms": ms_ext, "ms_first_extensions": ms_first, "ms_speculative_tasks": ms_tasks, "ms_commit": ms_commit, "ms_retry": tm["ms_ext_retry"] / steps,
                         "int32_Gops": ext_gops, "formula": "6*C + 4*W (cells, 16-base words)",
                         "frac_int32": ext_gops / alu_peak if alu_peak else None}}
    value = q_bases_all / 1e6 * steps / dt
    e2e_v = q_bases_all / 1e6 * steps / dt_e2e
    e2e_stages = {k: tm_e2e[k] / steps for k in ("ms_pack", "ms_build", "ms_scan", "ms_group", "ms_extend")}
    line = {"metric": metric, "value": value, "unit": "query Mbp/s", "n_gpus": world, "steps": steps, "warmup": a.warmup,
            "ms_per_step": 1000 * dt / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8/int32 (2-bit packed bases, int32 edit rows, f64 branch score)", "data": "synthetic",
            "config": config, "overlaps_per_sec": n_ovl * 1.0 / (dt / steps), "overlaps_per_step": n_ovl,
            "clocks": sampler.summary(),
            "e2e": {"value": e2e_v, "unit": "query Mbp/s", "ms_per_step": 1000 * dt_e2e / steps,
                    "overlaps_per_sec": n_ovl_e2e * 1.0 / (dt_e2e / steps),
                    "h2d_bytes_per_step": tm_e2e["h2d_bytes"] // steps, "d2h_bytes_per_step": tm_e2e["d2h_bytes"] // steps,
                    "stages_ms": e2e_stages},
            "device_ms_per_step": (tm["ms_build"] + tm["ms_total_search"]) / steps,
            "gpu_launches": int(tm["launches"]), "roofline": roof, "stages": stages,
            "work_per_step_rank0": {k: d[k] // steps for k in ("n_inserts", "n_distinct", "n_windows", "n_found", "n_refs", "n_add_ref",
                                                                 "n_runs", "n_pairs", "ext_calls", "ext_cells", "ext_words")}}
    line.update(skip_info)
    keep = {} if a.check else None
    if (world == 1 and not a.no_cpu) or a.check:
        if a.scale != 1.0:
            globals().update(GENOME=genome, BLOCK_BASES=block)
        # the same job, one run of the reference binary on this box's host cores; at one GPU the drop-in executable then runs on the same store
        exe = os.environ.get("OIC_BENCH_EXE", os.path.join(ROOT, "canu_b200", "bin", "overlapInCore")) if world == 1 else None
        r = reference_leg(world, 1, 0, keep=keep, exe=exe)
        if r.get("executable"):
            line["executable"] = r["executable"]
        line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
        line["cpu_baseline"]["overlaps_per_sec"] = r["overlaps_per_sec"]
        line["cpu_baseline"]["overlaps"] = r["overlaps"]
    if a.check:
        # record-for-record: the gathered records of the last end-to-end step (all ranks, store IDs) against the reference's .ovb
        from canu_b200 import api as _api
        dd = _api.decode_records(last_records[0])
        mine = sorted(H.format_raw({k: int(dd[k][i]) for k in dd}) for i in range(len(dd["a_iid"])))
        want = reference_records(keep["store"], keep["ovb"])
        ok = mine == want
        line["check"] = {"records": len(mine), "reference_records": len(want), "equal": ok,
                         "what": "every stored field of every overlap gathered from %d rank(s) vs oracle/_ref/overlapInCore on the same job" % world}
        subprocess.call(["rm", "-rf", keep["workdir"]])
    print(json.dumps(line))
    if a.check and not line["check"]["equal"]:
        if dist is not None:
            dist.destroy_process_group()
        raise SystemExit("bench.py --check: records differ from the reference (%d vs %d)" % (line["check"]["records"], line["check"]["reference_records"]))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
