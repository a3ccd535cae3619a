"""Seeded parity cases shared by the golden generator, the oracle tests and the GPU parity tests.

Each case is (reads spec, overlapper flags).  Sizes are chosen so the CPU oracle finishes in
seconds; the BASELINE.json shapes (C1..C5) are represented at reduced genome size.
"""
import helpers as H

# name -> dict(gen=kwargs for gen_reads, erate="string as passed to --maxerate", params=overrides, threads=T, skip=bool)
CASES = {
    # C1 shape: corrected long reads, one hash block vs one query block
    "c1_small": dict(gen=dict(genome_len=50000, coverage=20, len_min=1000, len_max=3000, err=0.02, seed=7),
                     erate="0.06", params=dict(hash_bits=18, hash_load=0.75), threads=4),
    # C2 shape: raw-noise reads, wide-band extension, hopeless check auto-off
    "c2_noisy": dict(gen=dict(genome_len=30000, coverage=15, len_min=1500, len_max=3500, err=0.06, seed=1002),
                     erate="0.15", params=dict(hash_bits=18, hash_load=0.75, kmer_len=17), threads=3),
    # C4 shape: short reads, partial overlaps (-G), min length 40
    "c4_short": dict(gen=dict(genome_len=20000, coverage=25, len_min=150, len_max=150, err=0.005, seed=1004, sub_only=True),
                     erate="0.06", params=dict(hash_bits=16, hash_load=0.75, partial=1, min_olap_len=40), threads=2),
    # several hash sub-blocks in one invocation (--hashstrings smaller than the -h range) + a -r sub-range
    "multiblock": dict(gen=dict(genome_len=30000, coverage=12, len_min=800, len_max=2000, err=0.015, seed=21),
                       erate="0.045", params=dict(hash_bits=17, hash_load=0.75, max_hash_strings=100,
                                                  bgn_hash_id=20, end_hash_id=230, bgn_ref_id=5, end_ref_id=200), threads=1),
    # hash-load cutoff trips (tiny table) -> Build_Hash_Index loads fewer reads than asked, several times
    "loadlimit": dict(gen=dict(genome_len=20000, coverage=10, len_min=800, len_max=1500, err=0.01, seed=33),
                      erate="0.06", params=dict(hash_bits=10, hash_load=0.8), threads=2),
    # scheduler quirk: (end-bgn) % perThread == 0 drops the last query read; -m keeps multiple overlaps per pair
    "quirk_m": dict(gen=dict(genome_len=15000, coverage=16, len_min=700, len_max=1600, err=0.02, seed=5),
                    erate="0.06", params=dict(hash_bits=16, hash_load=0.75, unique_olap=0, bgn_ref_id=1, end_ref_id=161),
                    threads=2),
    # repeat-rich genome with a frequent-k-mer skip list (C5 shape)
    "repeats_skip": dict(gen=dict(genome_len=60000, coverage=12, len_min=1000, len_max=2500, err=0.015, seed=1005,
                                  repeat_frac=0.3, repeat_families=4, repeat_min=300, repeat_max=900, repeat_div=0.02),
                         erate="0.06", params=dict(hash_bits=18, hash_load=0.75), threads=4, skip=True),
    # -z: hopeless check off at low error rate
    "nohopeless": dict(gen=dict(genome_len=20000, coverage=12, len_min=900, len_max=2000, err=0.025, seed=77),
                       erate="0.06", params=dict(hash_bits=16, hash_load=0.75, use_hopeless_check=0), threads=2),
    # wide band at scale: 8-12 kb reads at 12 % pairwise error -> rows wider than the 1024 diagonals staged in shared memory
    # (generic-row rerun of the task kernel) and histories that overflow the first-pass scratch (retry path)
    "long_wide": dict(gen=dict(genome_len=26000, coverage=4, len_min=8000, len_max=12000, err=0.06, seed=4041),
                      erate="0.15", params=dict(hash_bits=16, hash_load=0.75, kmer_len=16), threads=2),
    # rows wider than the 1024 diagonals staged in shared memory: 17-20 kb reads at 12 % pairwise error keep a band of
    # ~0.56 e diagonals alive for e > 1800 errors -> the generic-row instance of the task kernel
    "very_wide": dict(gen=dict(genome_len=22000, coverage=3.2, len_min=17000, len_max=20000, err=0.06, seed=9100),
                      erate="0.15", params=dict(hash_bits=16, hash_load=0.75, kmer_len=16), threads=1),
    # --minkmers: pairs with fewer k-mer hits than exp(-K erate) x (diagonal span - K + 1) are skipped (computeMinimumKmers,
    # Process_String_Overlaps.C:75-86; the flag is evaluated with the K / erate / --minlength seen so far, overlapInCore.C:392-393)
    "minkmers": dict(gen=dict(genome_len=40000, coverage=14, len_min=1200, len_max=3200, err=0.03, seed=314),
                     erate="0.09", params=dict(hash_bits=17, hash_load=0.75, kmer_len=14, min_olap_len=400, minkmers=True), threads=2),
    # error-free reads of very mixed length: exact-match extensions (row 0 reaches the end), containments, duplicates of
    # whole reads inside longer ones
    "exact_contained": dict(gen=dict(genome_len=12000, coverage=14, len_min=300, len_max=4000, err=0.0, seed=616),
                            erate="0.03", params=dict(hash_bits=15, hash_load=0.75, min_olap_len=100), threads=3),
    # ragged input: many reads shorter than K or than --minlength (never loaded / never queried), a few long ones
    "ragged_tiny": dict(gen=dict(genome_len=6000, coverage=12, len_min=10, len_max=700, err=0.01, seed=8),
                        erate="0.06", params=dict(hash_bits=14, hash_load=0.75, min_olap_len=60), threads=2),
    # reads with N: the hash side skips windows with a non-acgt base (Build_Hash_Index.C:376-379), the forward query strand treats
    # 'n' as a wildcard in the edit distance (prefixEditDistance-forward.C:172), the REVERSE strand is reverse-complemented with
    # non-ACGT -> NUL (AS_UTL_reverseComplement.C:39-72): its seed scan stops at the first NUL and NUL mismatches everything
    "n_bases": dict(gen=dict(genome_len=30000, coverage=14, len_min=900, len_max=2600, err=0.02, seed=2024),
                    erate="0.06", params=dict(hash_bits=16, hash_load=0.75), threads=2, inject_n=dict(rate=0.0015, runs=25, seed=99)),
    # the same in -G mode on short reads (C4 shape with N)
    "n_bases_partial": dict(gen=dict(genome_len=12000, coverage=25, len_min=150, len_max=150, err=0.005, seed=1044, sub_only=True),
                            erate="0.06", params=dict(hash_bits=15, hash_load=0.75, partial=1, min_olap_len=40), threads=2,
                            inject_n=dict(rate=0.003, runs=10, seed=5)),
    # -l: per-end cap on the overlaps of a query strand (Process_String_Overlaps.C:482-490,743-797): partners sorted by mean diagonal
    # (ties in array order: glibc's qsort is a merge sort), walked outward from diagonal zero until each end is full; seeds that would
    # add to a full end are skipped.  Deep coverage so that most strands have more partners than the limit.
    "olap_limit": dict(gen=dict(genome_len=30000, coverage=28, len_min=900, len_max=2600, err=0.02, seed=4101),
                       erate="0.06", params=dict(hash_bits=17, hash_load=0.75, frag_olap_limit=2), threads=3),
    # -l 1 with -m (several overlaps of one pair count separately towards the ends)
    "olap_limit_m": dict(gen=dict(genome_len=24000, coverage=24, len_min=800, len_max=2200, err=0.025, seed=4102),
                         erate="0.06", params=dict(hash_bits=16, hash_load=0.75, frag_olap_limit=1, unique_olap=0), threads=2),
    # a limit that only some strands reach (both branches of Process_String_Olaps in one job), hash sub-blocks, -z
    "olap_limit_blocks": dict(gen=dict(genome_len=36000, coverage=18, len_min=700, len_max=3000, err=0.015, seed=4103),
                              erate="0.045", params=dict(hash_bits=17, hash_load=0.75, frag_olap_limit=7, max_hash_strings=120,
                                                         use_hopeless_check=0), threads=4),
    # -w: window filter (Process_String_Overlaps.C:365-388,560-620).  Reads at 3 % error so that a good share of the overlaps has a
    # 50-base window with >= 8 errors.  The reference's gatekeeperCreate stores one quality per read (its library's default QV, 20);
    # one thread: the two "Rejected by ... window" counters are unsynchronised globals in the reference
    "window_filter": dict(gen=dict(genome_len=30000, coverage=14, len_min=800, len_max=2500, err=0.03, seed=4201),
                          erate="0.06", params=dict(hash_bits=17, hash_load=0.75, use_window_filter=1), threads=1),
    # four libraries with default QVs 12 / 31 / 17 / 25 (min(query, hash read) and the cap at 20 both matter), -m
    "window_libs": dict(gen=dict(genome_len=30000, coverage=16, len_min=800, len_max=2500, err=0.035, seed=4202),
                        erate="0.06", params=dict(hash_bits=17, hash_load=0.75, use_window_filter=1, unique_olap=0), threads=1,
                        lib_qv=[12, 31, 17, 25]),
    # -w together with -l, on reads with N (the filter's base comparison has the same wildcard / NUL rules as the aligner)
    "window_limit_n": dict(gen=dict(genome_len=26000, coverage=18, len_min=900, len_max=2400, err=0.03, seed=4203),
                           erate="0.06", params=dict(hash_bits=16, hash_load=0.75, use_window_filter=1, frag_olap_limit=4), threads=1,
                           lib_qv=[25, 15], inject_n=dict(rate=0.001, runs=12, seed=7)),
    # empty output: sparse reads of a large genome share no k-mer runs worth an overlap
    "no_overlaps": dict(gen=dict(genome_len=2000000, coverage=0.02, len_min=900, len_max=1500, err=0.02, seed=12),
                        erate="0.06", params=dict(hash_bits=14, hash_load=0.75), threads=2),
}


def build_case(name):
    """-> (reads, params dict for the oracle, erate string, threads, skip k-mer list or None)"""
    c = CASES[name]
    reads = H.gen_reads(**c["gen"])
    if c.get("inject_n"):
        inject_n(reads, **c["inject_n"])
    if c.get("lib_qv"):   # consecutive libraries of (almost) equal read count, each with its own default QV
        import numpy as np
        k = len(c["lib_qv"])
        reads.read_qv = np.array([c["lib_qv"][min(k - 1, i * k // reads.n)] for i in range(reads.n)], dtype=np.uint8)
    over = dict(c["params"])
    if over.pop("minkmers", False):
        import math
        over["filter_by_kmer_count"] = int(math.floor(math.exp(-1.0 * over.get("kmer_len", 22) * H.strtof(c["erate"])) *
                                                      (over.get("min_olap_len", 500) - over.get("kmer_len", 22) + 1)))
    over.setdefault("max_hash_strings", reads.n)
    over.setdefault("max_hash_data_len", 100000000)
    over.setdefault("end_hash_id", reads.n)
    over.setdefault("end_ref_id", reads.n)
    P = H.default_params(max_erate=H.strtof(c["erate"]), num_threads=c["threads"], **over)
    skip = None
    if c.get("skip"):
        skip = frequent_kmers(reads, P["kmer_len"], 40)
    return reads, P, c["erate"], c["threads"], skip


def inject_n(reads, rate, runs, seed):
    """Overwrite single bases (fraction `rate`) and `runs` short stretches (2-12 bases) of the reads with 'N', seeded."""
    import numpy as np
    rng = np.random.default_rng(seed)
    total = int(reads.lengths.sum())
    pos = rng.choice(total, size=max(1, int(total * rate)), replace=False)
    reads.bases[pos] = ord("N")
    for _ in range(runs):
        rid = int(rng.integers(0, reads.n))
        ln = int(reads.lengths[rid])
        k = int(rng.integers(2, 13))
        if ln > k + 2:
            o = int(reads.offsets[rid]) + int(rng.integers(0, ln - k))
            reads.bases[o:o + k] = ord("N")


def frequent_kmers(reads, K, threshold):
    """Canonical-agnostic frequent k-mer list (what `meryl -Dt -n T` would feed `-k file`)."""
    from collections import Counter
    cnt = Counter()
    for rid in range(1, reads.n + 1):
        s = reads.seq(rid).lower()
        for i in range(0, len(s) - K + 1):
            cnt[s[i:i + K]] += 1
    comp = str.maketrans("acgt", "tgca")
    out = []
    seen = set()
    for k, v in sorted(cnt.items()):
        rc = k.translate(comp)[::-1]
        tot = v + (cnt.get(rc, 0) if rc != k else 0)
        canon = min(k, rc)
        if tot >= threshold and canon not in seen:
            seen.add(canon)
            out.append(canon)
    return out


def write_skip_file(kmers, path):
    with open(path, "w") as f:
        for k in kmers:
            f.write(">1\n%s\n" % k)


# (case, K, min count) of the frequent-k-mer fixtures (tests/golden/make_golden_kmers.py: reference meryl output)
KMER_CASES = [("repeats_skip", 22, 40), ("repeats_skip", 17, 14), ("c1_small", 31, 14), ("n_bases", 22, 7), ("c4_short", 14, 12)]


# ----------------------------------------------------------------------------- seeded random cases
ERATES = ["0.03", "0.045", "0.06", "0.1", "0.15"]
N_FUZZ = 80    # jobs 64.. also draw a -l limit (the first 64 are the round-1 set, unchanged)


def fuzz_case(i):
    """Seeded random job i -> (reads, params, erate string, threads, None): random read shape / error rate, K, --maxerate, -G,
    -m/-u, -z, -l (jobs 64..), --minlength, hash sub-blocks (--hashstrings, tiny tables that trip the load cut-off), -h/-r sub-ranges,
    thread count (scheduler quirk) and N bases.  The oracle is held to the LIVE reference binary on these (test_oracle.py,
    where the reference is built) and the CUDA path to the oracle (test_gpu_parity.py)."""
    import numpy as np
    rng = np.random.default_rng(7000 + i)
    short = rng.random() < 0.25
    if short:
        lmin = int(rng.integers(100, 200)); lmax = lmin + int(rng.integers(0, 120))
        glen = int(rng.integers(4000, 12000)); cov = float(rng.uniform(10, 22))
    else:
        lmin = int(rng.integers(300, 1500)); lmax = lmin + int(rng.integers(0, 2500))
        glen = int(rng.integers(8000, 30000)); cov = float(rng.uniform(6, 14))
    err = float(rng.choice([0.0, 0.005, 0.01, 0.02, 0.04, 0.06]))
    reads = H.gen_reads(genome_len=glen, coverage=cov, len_min=lmin, len_max=lmax, err=err, seed=int(9000 + i),
                        sub_only=bool(rng.random() < 0.2))
    if rng.random() < 0.35:
        inject_n(reads, rate=float(rng.choice([0.0005, 0.002, 0.01])), runs=int(rng.integers(0, 12)), seed=int(i))
    erate = str(rng.choice(ERATES))
    over = dict(kmer_len=int(rng.choice([12, 14, 16, 17, 20, 22, 25, 31])),
                hash_bits=int(rng.choice([10, 12, 14, 16])), hash_load=float(rng.choice([0.3, 0.6, 0.75, 0.9])),
                partial=int(short or rng.random() < 0.2), unique_olap=int(rng.random() < 0.7),
                use_hopeless_check=int(rng.random() < 0.7),
                min_olap_len=int(rng.choice([40, 60]) if short else rng.choice([100, 250, 500])))
    n = reads.n
    over["max_hash_strings"] = int(rng.choice([n, max(2, n // 2), max(2, n // 5), 17]))
    over["max_hash_data_len"] = int(rng.choice([100000000, max(2000, int(reads.lengths.sum()) // 3)]))
    if rng.random() < 0.4:
        a, b = sorted(int(x) for x in rng.integers(1, n + 1, size=2))
        over["bgn_hash_id"], over["end_hash_id"] = a, max(b, a + 1)
    else:
        over["end_hash_id"] = n
    if rng.random() < 0.4:
        a, b = sorted(int(x) for x in rng.integers(1, n + 1, size=2))
        over["bgn_ref_id"], over["end_ref_id"] = a, b
    else:
        over["end_ref_id"] = n
    threads = int(rng.choice([1, 2, 3, 4, 7]))
    if i >= 64 and not over["partial"]:   # (-G with -l makes the reference read uninitialised alignment bounds: not testable)
        over["frag_olap_limit"] = int(rng.choice([1, 2, 3, 5, 9]))
    P = H.default_params(max_erate=H.strtof(erate), num_threads=threads, **over)
    return reads, P, erate, threads, None


N_FUZZ_W = 24


def fuzz_case_w(i):
    """Seeded random -w job i (CPU only: the oracle against the live reference binary, tests/test_oracle.py): long reads at 2-4 % error
    so that 50-base windows with many errors occur, --maxerate <= 0.06 (above it the reference turns -w off), one thread (the two
    window counters are unsynchronised globals in the reference), 1-4 libraries with their own default QV (the window score uses
    min(quality, 20) of both reads), sometimes -m, -l, N bases, hash sub-blocks and -h / -r sub-ranges."""
    import numpy as np
    rng = np.random.default_rng(7700 + i)
    lmin = int(rng.integers(500, 1400)); lmax = lmin + int(rng.integers(200, 2000))
    glen = int(rng.integers(12000, 30000)); cov = float(rng.uniform(8, 16))
    err = float(rng.choice([0.02, 0.03, 0.035, 0.04]))
    reads = H.gen_reads(genome_len=glen, coverage=cov, len_min=lmin, len_max=lmax, err=err, seed=int(9700 + i), sub_only=bool(rng.random() < 0.2))
    if rng.random() < 0.3:
        inject_n(reads, rate=float(rng.choice([0.0005, 0.002])), runs=int(rng.integers(0, 10)), seed=int(i))
    k = int(rng.integers(1, 5))
    qv = [int(x) for x in rng.choice([5, 10, 12, 15, 17, 19, 20, 21, 25, 31, 40], size=k)]
    reads.read_qv = np.array([qv[min(k - 1, j * k // reads.n)] for j in range(reads.n)], dtype=np.uint8)
    erate = str(rng.choice(["0.045", "0.06"]))
    over = dict(kmer_len=int(rng.choice([14, 16, 17, 20, 22])), hash_bits=int(rng.choice([12, 14, 16])), hash_load=float(rng.choice([0.6, 0.75])),
                partial=0, unique_olap=int(rng.random() < 0.6), use_hopeless_check=int(rng.random() < 0.7),
                min_olap_len=int(rng.choice([100, 250, 500])), use_window_filter=1)
    n = reads.n
    over["max_hash_strings"] = int(rng.choice([n, max(2, n // 2), max(2, n // 4)]))
    over["end_hash_id"] = n
    if rng.random() < 0.3:
        a, b = sorted(int(x) for x in rng.integers(1, n + 1, size=2))
        over["bgn_ref_id"], over["end_ref_id"] = a, b
    else:
        over["end_ref_id"] = n
    if rng.random() < 0.35:
        over["frag_olap_limit"] = int(rng.choice([2, 3, 5, 9]))
    P = H.default_params(max_erate=H.strtof(erate), num_threads=1, **over)
    return reads, P, erate, 1, None

