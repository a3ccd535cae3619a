"""GPU parity tests: the CUDA path, called through the C ABI (include/oic_b200.h), against

* the committed golden fixtures (outputs of the reference binary, tests/golden/), and
* the CPU oracle (oracle/) on the same seeded inputs, stage by stage: k-mer index, per-pair seed lists,
  overlap records, .stats counters and the roofline work counters.

Bit-exact: integer / byte / index work throughout; the only floating point on the path (branch-point score,
overlap quality -> evalue) is IEEE double without FMA and must give identical integers.
"""
import gzip
import json
import os

import numpy as np
import pytest

import cases
import helpers as H

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
STAT_KEYS = ["kmer_hits_without_olap", "kmer_hits_with_olap", "kmer_hits_skipped", "multi_overlap", "contained", "dovetail",
             "bad_short_window", "bad_long_window"]
WORK_KEYS = ["n_hash_bases", "n_inserts", "n_distinct", "n_query_bases", "n_windows", "n_found", "n_refs", "n_add_ref",
             "ext_calls", "ext_cells", "ext_words", "n_hash_blocks", "last_end_hash_id"]


def make_overlapper(reads, P, skip=None, debug=0):
    from canu_b200 import api
    ov = api.Overlapper(P)
    ov.load_reads(reads.bases, reads.offsets[:-1], reads.lengths)
    if P.get("use_window_filter"):
        ov.load_qualities(reads.qualities(), reads.offsets[:-1])
    if skip:
        ov.set_skip_kmers(skip)
    ov.set_debug(debug)
    return ov


def formatted(rec):
    from canu_b200 import api
    d = api.decode_records(rec)
    n = len(d["a_iid"])
    return sorted(H.format_raw({k: int(d[k][i]) for k in d}) for i in range(n))


@pytest.mark.parametrize("name", list(cases.CASES))
def test_overlaps_vs_golden_and_oracle(name):
    reads, P, erate, threads, skip = cases.build_case(name)
    ov = make_overlapper(reads, P, skip)
    ov.overlap_driver()
    mine = formatted(ov.overlaps())
    st = ov.stats()
    ov.close()
    gold = gzip.open(os.path.join(GOLDEN, name + ".txt.gz")).read().decode().splitlines()
    meta = json.load(open(os.path.join(GOLDEN, name + ".json")))
    assert len(mine) == len(gold), (len(mine), len(gold))
    assert mine == gold
    # .counts parity (ST/ovStoreHistogram.C:230-241): one count per read ID for both ends of every overlap
    res = H.run_oracle(reads, P, skip_kmers=skip)
    for k in STAT_KEYS + ([] if P["partial"] else ["total_overlaps"]):
        assert st[k] == res["stats"][k], k
    for k in WORK_KEYS:
        assert st[k] == res["stats"][k], k
    ref_map = {"Kmer hits without olaps": "kmer_hits_without_olap", "Kmer hits with olaps": "kmer_hits_with_olap",
               "Multiple overlaps/pair": "multi_overlap", "Contained overlaps": "contained", "Dovetail overlaps": "dovetail"}
    if P.get("use_window_filter"):   # (these cases run the reference with one thread: its two window counters are racy otherwise)
        ref_map.update({"Rejected by short window": "bad_short_window", "Rejected by long window": "bad_long_window"})
        assert st["bad_short_window"] + st["bad_long_window"] > 0, "the case is meant to reject something"
    for k, v in ref_map.items():
        assert meta["stats"][k] == st[v], k


@pytest.mark.parametrize("name", ["c1_small", "repeats_skip", "loadlimit", "c4_short", "n_bases"])
def test_index_vs_oracle(name):
    """k-mer index of the LAST hash block: key set, occurrence lists in chain order, Empty flags."""
    reads, P, erate, threads, skip = cases.build_case(name)
    res = H.run_oracle(reads, P, dump=4, skip_kmers=skip)
    ov = make_overlapper(reads, P, skip)
    # replay the driver's hash sub-block loop without searching
    n = reads.n
    g_bgn, g_end = max(P["bgn_hash_id"], 1), min(P["end_hash_id"], n)
    bgn, last = g_bgn, 0
    while bgn < g_end:
        end = min(bgn + P["max_hash_strings"] - 1, g_end)
        last = ov.build_hash_index(bgn, end)
        bgn = last + 1
    assert last == res["stats"]["last_end_hash_id"]
    idx = ov.dump_index()
    ov.close()
    ref = res["index"]
    assert len(idx["keys"]) == len(ref["keys"])
    assert (idx["keys"] == ref["keys"]).all()
    assert (idx["flags"] == ref["flags"]).all()
    # screened-absent k-mers carry a placeholder occurrence in the oracle (an "extra string"); compare real lists only
    mine_len = np.diff(idx["occ_start"].astype(np.int64))
    ref_len = np.diff(ref["occ_start"].astype(np.int64))
    absent = (ref["flags"] == 1) & (mine_len == 0) & (ref_len == 1)
    assert (mine_len[~absent] == ref_len[~absent]).all()
    if absent.any():
        keep = np.ones(len(ref["occ"]), dtype=bool)
        keep[ref["occ_start"][:-1][absent].astype(np.int64)] = False
        ref_occ = ref["occ"][keep]
    else:
        ref_occ = ref["occ"]
    assert len(idx["occ"]) == len(ref_occ)
    assert (idx["occ"] == ref_occ).all()


@pytest.mark.parametrize("name", ["c1_small", "c2_noisy", "repeats_skip", "c4_short", "n_bases", "n_bases_partial"])
def test_seed_lists_vs_oracle(name):
    """Per (query, strand, hash read): seed list in list order, `consistent`, diag_ct/bgn/end."""
    reads, P, erate, threads, skip = cases.build_case(name)
    res = H.run_oracle(reads, P, dump=1, skip_kmers=skip)
    if res["stats"]["n_hash_blocks"] != 1:
        pytest.skip("single hash block cases only")
    ov = make_overlapper(reads, P, skip, debug=1)
    last = ov.build_hash_index(max(P["bgn_hash_id"], 1), min(P["end_hash_id"], reads.n))
    assert last == res["stats"]["last_end_hash_id"]
    bgn_ref, end_ref = max(P["bgn_ref_id"], 1), min(P["end_ref_id"], reads.n)
    ov.process_overlaps(bgn_ref, ov.last_processed_ref(bgn_ref, end_ref, P["num_threads"]))
    pairs, matches = ov.dump_pairs()
    ov.close()

    def as_dict(pr, mt, get):
        d = {}
        for p in pr:
            q, h, s, cons, ct, b, e, n, first = get(p)
            d[(q, h, s)] = (cons, ct, b, e, [tuple(int(x) for x in mt[first + i]) for i in range(n)])
        return d
    mine = as_dict(pairs, matches, lambda p: tuple(int(x) for x in p))
    om = np.stack([res["matches"]["offset"], res["matches"]["len"], res["matches"]["start"]], axis=1)
    ref = as_dict(res["pairs"], om, lambda p: (int(p["q_id"]), int(p["h_id"]), int(p["strand"]), int(p["consistent"]), int(p["diag_ct"]),
                                               int(p["diag_bgn"]), int(p["diag_end"]), int(p["n_matches"]), int(p["first_match"])))
    assert len(mine) == len(ref)
    assert set(mine) == set(ref)
    bad = [k for k in ref if mine[k] != ref[k]]
    assert not bad, (len(bad), bad[0], mine[bad[0]], ref[bad[0]])


def test_no_device_fallback_symbols():
    """The library exports the whole ABI and nothing routes through the oracle."""
    from canu_b200 import api
    lib = api.load_library()
    for s in api.ABI_SYMBOLS:
        assert hasattr(lib, s)


@pytest.mark.parametrize("name", ["c1_small", "c4_short", "multiblock"])
def test_two_bit_load_equals_ascii_load(name):
    """oic_load_reads_2bit (the gkpStore's 2SEQ bytes, no host decode) gives the same overlaps and counters as oic_load_reads"""
    from canu_b200 import api
    reads, P, erate, threads, skip = cases.build_case(name)
    ov = make_overlapper(reads, P, skip)
    ov.overlap_driver()
    want, st_want = formatted(ov.overlaps()), ov.stats()
    ov.close()
    packed, boff = H.encode_2bit(reads.bases, reads.offsets, reads.lengths)
    ov = api.Overlapper(P)
    ov.load_reads_2bit(packed, boff, reads.lengths)
    if skip:
        ov.set_skip_kmers(skip)
    ov.overlap_driver()
    got, st_got = formatted(ov.overlaps()), ov.stats()
    ov.close()
    assert got == want
    for k in STAT_KEYS + WORK_KEYS:
        assert st_got[k] == st_want[k], k


@pytest.mark.parametrize("name,env", [("c2_noisy", {"OIC_VALS_BUDGET_MB": "48"}), ("c1_small", {"OIC_DPOOL_MB": "0"}),
                                      ("long_wide", {"OIC_VALS_BUDGET_MB": "256", "OIC_DPOOL_MB": "0"})])
def test_overflow_retry_path(name, env, monkeypatch):
    """Edit-row history or delta pool too small for the speculative pass: the affected pairs are rerun serially with worst-case
    scratch (oic_b200.cu run_batch) and the records / counters must not change."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    reads, P, erate, threads, skip = cases.build_case(name)
    ov = make_overlapper(reads, P, skip)
    ov.overlap_driver()
    mine, st, tm = formatted(ov.overlaps()), ov.stats(), ov.timings()
    ov.close()
    gold = gzip.open(os.path.join(GOLDEN, name + ".txt.gz")).read().decode().splitlines()
    assert mine == gold
    assert tm["ms_ext_retry"] > 0, "the knob did not trigger the retry path"
    res = H.run_oracle(reads, P, skip_kmers=skip)
    for k in ["ext_calls", "ext_cells", "ext_words"] + STAT_KEYS:
        assert st[k] == res["stats"][k], k


@pytest.mark.parametrize("i", list(range(cases.N_FUZZ)))
def test_fuzz_vs_oracle(i):
    """Seeded random jobs (tests/cases.py fuzz_case: K, --maxerate, -G, -m/-u, -z, --minlength, hash sub-blocks, load cut-off,
    -h/-r sub-ranges, thread counts, N bases).  The oracle is held to the live reference binary on the same jobs by
    test_oracle.py::test_oracle_vs_live_reference_fuzz; here the CUDA path is held to the oracle: records, -s counters, work counters."""
    reads, P, erate, threads, _ = cases.fuzz_case(i)
    ov = make_overlapper(reads, P)
    ov.overlap_driver()
    mine = formatted(ov.overlaps())
    st = ov.stats()
    ov.close()
    res = H.run_oracle(reads, P)
    want = sorted(H.format_raw(o) for o in res["overlaps"])
    assert len(mine) == len(want), (len(mine), len(want))
    assert mine == want
    for k in STAT_KEYS + WORK_KEYS:
        assert st[k] == res["stats"][k], k


@pytest.mark.parametrize("name,K,T", cases.KMER_CASES)
@pytest.mark.parametrize("budget_mb", [None, "1"])
def test_kmer_counts_vs_meryl(name, K, T, budget_mb, monkeypatch):
    """frequent-k-mer row: oic_count_kmers on the GPU writes the bytes the reference's meryl wrote
    (`meryl -B -C` + `meryl -Dt -n T`, fixtures of tests/golden/make_golden_kmers.py), in one pass and -- with the table budget
    forced down to 1 MB -- in many hash-partition passes; the CPU restatement agrees key by key."""
    if budget_mb:
        monkeypatch.setenv("OIC_COUNT_BUDGET_MB", budget_mb)
    reads = cases.build_case(name)[0]
    P = H.default_params(kmer_len=K, max_hash_strings=reads.n, end_hash_id=reads.n, end_ref_id=reads.n)
    ov = make_overlapper(reads, P)
    keys, counts, info = ov.count_kmers(K, T)
    ov.close()
    assert (info["passes"] > 1) == bool(budget_mb)
    gold = gzip.open(os.path.join(GOLDEN, "kmers_%s_%d_%d.txt.gz" % (name, K, T))).read().decode()
    assert H.meryl_dump_text(keys, counts, K) == gold
    ck, cc = H.count_kmers_cpu(reads, K, T)
    assert (keys == ck).all() and (counts == cc).all()
    # the count histogram = what `meryl -Dh` printed
    out, hinfo = H.meryl_hist_text(info["histogram"])
    assert out + "--info--\n" + hinfo == gzip.open(os.path.join(GOLDEN, "khist_%s_%d.txt.gz" % (name, K))).read().decode()


def test_kmer_counts_feed_skip_list():
    """the GPU-counted list, handed back as the -k skip list, gives the overlaps of the golden repeats_skip job
    (whose list came from the pure-Python counter of tests/cases.py)"""
    reads, P, erate, threads, skip = cases.build_case("repeats_skip")
    ov = make_overlapper(reads, P)
    keys, counts, _ = ov.count_kmers(P["kmer_len"], 40)
    ov.set_skip_kmers(H.kmer_strings(keys, P["kmer_len"]))
    ov.overlap_driver()
    mine = formatted(ov.overlaps())
    ov.close()
    gold = gzip.open(os.path.join(GOLDEN, "repeats_skip.txt.gz")).read().decode().splitlines()
    assert mine == gold
