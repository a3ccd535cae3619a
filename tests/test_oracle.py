"""The CPU restatement (oracle/) against outputs of the reference binary.

* golden: committed fixtures produced by tests/golden/make_golden.py (reference binary run in the build
  container) -- always checked;
* live: the reference binary itself (oracle/_ref/overlapInCore) when it is present.
Bit-exact: every stored field of every overlap record, and the .stats counters.
"""
import gzip
import json
import os
import tempfile

import pytest

import cases
import helpers as H

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

STAT_MAP = {"Kmer hits without olaps": "kmer_hits_without_olap", "Kmer hits with olaps": "kmer_hits_with_olap",
            "Multiple overlaps/pair": "multi_overlap", "Total overlaps produced": "total_overlaps",
            "Contained overlaps": "contained", "Dovetail overlaps": "dovetail"}


def check_stats(ref_stats, st, partial):
    for k, v in STAT_MAP.items():
        if partial and k == "Total overlaps produced":
            continue  # unsynchronised global counter in -G mode (OIC/overlapInCore-Output.C:253)
        assert ref_stats[k] == st[v], k


@pytest.mark.parametrize("name", list(cases.CASES))
def test_oracle_vs_golden(name):
    reads, P, erate, threads, skip = cases.build_case(name)
    meta = json.load(open(os.path.join(GOLDEN, name + ".json")))
    assert meta["n_reads"] == reads.n and meta["total_bases"] == int(reads.lengths.sum())
    gold = gzip.open(os.path.join(GOLDEN, name + ".txt.gz")).read().decode().splitlines()
    res = H.run_oracle(reads, P, skip_kmers=skip)
    mine = sorted(H.format_raw(o) for o in res["overlaps"])
    assert len(mine) == len(gold)
    assert mine == gold
    check_stats(meta["stats"], res["stats"], P["partial"])


@pytest.mark.skipif(not H.have_ref(), reason="reference binary not built (oracle/build_ref.sh)")
@pytest.mark.parametrize("seed,threads", [(101, 1), (202, 5)])
def test_oracle_vs_live_reference(seed, threads):
    reads = H.gen_reads(25000, 14, 900, 2200, 0.02, seed)
    P = H.default_params(hash_bits=17, hash_load=0.75, max_hash_strings=reads.n, end_hash_id=reads.n,
                         end_ref_id=reads.n, num_threads=threads)
    with tempfile.TemporaryDirectory() as wd:
        store = H.make_store(reads, wd)
        ovb, st = H.run_overlapper(os.path.join(H.REF_DIR, "overlapInCore"), store, os.path.join(wd, "ref"),
                                   H.oic_args(P, "0.06"), threads)
        ref = H.dump_ovb_sorted(store, ovb)
        ref_stats = H.parse_stats(st)
    res = H.run_oracle(reads, P)
    assert sorted(H.format_raw(o) for o in res["overlaps"]) == ref
    check_stats(ref_stats, res["stats"], False)


@pytest.mark.skipif(not H.have_ref(), reason="reference binary not built (oracle/build_ref.sh)")
@pytest.mark.parametrize("i", list(range(cases.N_FUZZ)))
def test_oracle_vs_live_reference_fuzz(i):
    """Seeded random jobs (tests/cases.py fuzz_case): the oracle against the reference binary run on the spot."""
    reads, P, erate, threads, _ = cases.fuzz_case(i)
    with tempfile.TemporaryDirectory() as wd:
        store = H.make_store(reads, wd)
        ovb, st = H.run_overlapper(os.path.join(H.REF_DIR, "overlapInCore"), store, os.path.join(wd, "ref"),
                                   H.oic_args(P, erate), threads)
        ref = H.dump_ovb_sorted(store, ovb)
        ref_stats = H.parse_stats(st)
    res = H.run_oracle(reads, P)
    assert sorted(H.format_raw(o) for o in res["overlaps"]) == ref
    check_stats(ref_stats, res["stats"], P["partial"])


@pytest.mark.skipif(not H.have_ref(), reason="reference binary not built (oracle/build_ref.sh)")
@pytest.mark.parametrize("i", list(range(cases.N_FUZZ_W)))
def test_oracle_vs_live_reference_fuzz_window_filter(i):
    """Seeded random -w jobs (tests/cases.py fuzz_case_w): records and all -s counters, the two window rejections included."""
    reads, P, erate, threads, _ = cases.fuzz_case_w(i)
    with tempfile.TemporaryDirectory() as wd:
        store = H.make_store(reads, wd)
        ovb, st = H.run_overlapper(os.path.join(H.REF_DIR, "overlapInCore"), store, os.path.join(wd, "ref"),
                                   H.oic_args(P, erate), threads)
        ref = H.dump_ovb_sorted(store, ovb)
        ref_stats = H.parse_stats(st)
    res = H.run_oracle(reads, P)
    assert sorted(H.format_raw(o) for o in res["overlaps"]) == ref
    check_stats(ref_stats, res["stats"], P["partial"])
    for k, v in {"Rejected by short window": "bad_short_window", "Rejected by long window": "bad_long_window"}.items():
        assert ref_stats[k] == res["stats"][v], k


_EXTRA = int(os.environ.get("OIC_FUZZ_EXTRA", "0"))


@pytest.mark.skipif(not (H.have_ref() and _EXTRA), reason="set OIC_FUZZ_EXTRA=<n> to hold the oracle to the reference on n more random jobs")
@pytest.mark.parametrize("i", list(range(cases.N_FUZZ, cases.N_FUZZ + _EXTRA)))
def test_oracle_vs_live_reference_fuzz_extra(i):
    """More of the same sweep on demand (jobs N_FUZZ .. N_FUZZ + n - 1; run with n = 240 at the end of round 2: all equal)."""
    test_oracle_vs_live_reference_fuzz(i)


def _kmer_golden(name, K, T):
    return gzip.open(os.path.join(GOLDEN, "kmers_%s_%d_%d.txt.gz" % (name, K, T))).read().decode()


@pytest.mark.parametrize("name,K,T", cases.KMER_CASES)
def test_kmer_counter_vs_meryl_golden(name, K, T):
    """frequent-k-mer row: the CPU restatement of `meryl -B -C` + `meryl -Dt -n T` (tools/gen_reads.c) writes
    the bytes the reference's meryl wrote (tests/golden/make_golden_kmers.py)."""
    reads = cases.build_case(name)[0]
    keys, counts = H.count_kmers_cpu(reads, K, T)
    assert H.meryl_dump_text(keys, counts, K) == _kmer_golden(name, K, T)


@pytest.mark.parametrize("name,K", sorted({(n, k) for n, k, _ in cases.KMER_CASES}))
def test_kmer_histogram_vs_meryl_golden(name, K):
    """the count histogram of the CPU restatement prints as the reference's `meryl -Dh` did (stdout and the four stderr lines)"""
    reads = cases.build_case(name)[0]
    out, info = H.meryl_hist_text(H.kmer_histogram_cpu(reads, K))
    assert out + "--info--\n" + info == gzip.open(os.path.join(GOLDEN, "khist_%s_%d.txt.gz" % (name, K))).read().decode()


@pytest.mark.skipif(not H.have_ref(), reason="reference binary not built (oracle/build_ref.sh)")
@pytest.mark.parametrize("i", [3, 9, 19])
def test_kmer_counter_vs_live_meryl(i):
    reads = cases.fuzz_case(i)[0]
    K, T = [(15, 2), (22, 2), (27, 3)][i % 3]
    with tempfile.TemporaryDirectory() as wd:
        want = H.run_meryl(H.make_store(reads, wd), K, T, wd)
    keys, counts = H.count_kmers_cpu(reads, K, T)
    assert H.meryl_dump_text(keys, counts, K) == want


def test_host_tables_monotone():
    lib = H.oracle_lib()
    import numpy as np
    e = H.strtof("0.06")
    n = lib.oracle_max_errors(e)
    ml = np.zeros(n + 1, dtype=np.int32)
    lib.oracle_match_limit(e, n, ml.ctypes.data)
    assert ml[0] == 0 and ml[1] == 0
    assert (np.diff(ml[:n]) >= 0).all()
    # Error_Bound uses the float-parsed rate: 0.045f > 0.045 so 1000 bases allow 46 errors
    assert lib.oracle_error_bound(H.strtof("0.045"), 1000) == 46
