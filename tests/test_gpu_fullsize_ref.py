"""Full-size RECORD parity against the reference binary itself.

The small seeded cases of test_gpu_parity.py never reach a 2^28-slot table, multi-chunk seed-run chains, the block-wide
occurrence sort or 10^4-read blocks.  Here the jobs BASELINE.json names are run at their full size by BOTH programs on the
SAME gkpStore (written by the reference's gatekeeperCreate):

* `oracle/_ref/overlapInCore -t <cores>`  -- the reference, compiled from /root/reference/src by oracle/build_ref.sh (the
  binary travels to the GPU box; /root/reference itself is not needed at run time), about a minute of host time per job;
* `canu_b200/bin/overlapInCore -t <cores>` -- the drop-in executable over the C ABI / CUDA path.

Compared: every field of every overlap (`overlapConvert -raw | sort`, OIC/overlapInCore.C:190-300 end to end), the `.counts`
side file byte for byte (ST/ovStoreHistogram.C:230-241) and the `-s` counters (OIC/overlapInCore.C:569-588).  `-t` is the same
for both, so the scheduler quirk (Process_Overlaps.C:86) drops the same last read, if any.

Also here: the hash-block RECEIVE path (oic_hash_block_reserve -> copy -> oic_hash_block_commit) checked record for record, on one
GPU (two contexts, device-to-device copy) and, when the box has two GPUs, across ranks with the NCCL replication of the C ABI.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu

BIN = os.path.join(H.ROOT, "canu_b200", "bin", "overlapInCore")
needs_tools = pytest.mark.skipif(not (H.have_ref() and os.access(BIN, os.X_OK)),
                                 reason="oracle/_ref tools or canu_b200/bin/overlapInCore not built")


def _threads():
    return max(1, min(os.cpu_count() or 1, 32))


def _job(workload):
    """(reads, P, erate string) of one bench step of a BASELINE.json shape (bench.py WORKLOADS)"""
    import bench
    from canu_b200 import shard
    bench.set_workload(workload)
    if bench.MODE == "self":
        reads = bench.gen_blocks(bench.GENOME, 1, bench.BLOCK_BASES, bench.SEED)
        P = bench.params_for(reads.n, reads.n, threads=_threads())
    else:
        reads, ((qb, qe), (hb, he)) = bench.gen_job(bench.GENOME, 1, bench.BLOCK_BASES, bench.SEED)
        P = bench.params_for(he - hb + 1, reads.n, threads=_threads())
    return reads, P, bench.ERATE


def _both(workload, wd, min_records):
    reads, P, erate = _job(workload)
    store = H.make_store(reads, wd)
    args = H.oic_args(P, erate)
    t = _threads()
    out = {}
    for name, binary in (("ref", os.path.join(H.REF_DIR, "overlapInCore")), ("mine", BIN)):
        ovb, st = H.run_overlapper(binary, store, os.path.join(wd, name), args, t)
        out[name] = (H.dump_ovb_sorted(store, ovb), H.parse_stats(st), open(os.path.join(wd, name + ".counts"), "rb").read())
    want, got = out["ref"], out["mine"]
    assert len(want[0]) >= min_records, "the reference found only %d overlaps" % len(want[0])
    assert len(got[0]) == len(want[0])
    assert got[0] == want[0]
    assert got[2] == want[2], ".counts differ"
    s_mine, s_ref = dict(got[1]), dict(want[1])
    if P["partial"]:
        # -G: the reference bumps an unsynchronised global from every thread (OIC/overlapInCore-Output.C:253): racy, <= records
        assert s_mine.pop("Total overlaps produced") == len(want[0]) >= s_ref.pop("Total overlaps produced")
    assert s_mine == s_ref
    return len(want[0])


@needs_tools
def test_c3_bench_job_records_equal_reference(tmp_path):
    """the bench line's job (100 Mbp hash block x 100 Mbp query block, seed 1003): 19 024 overlaps"""
    n = _both("c3", str(tmp_path), 10000)
    assert n == 19024


@needs_tools
def test_c1_full_size_records_equal_reference(tmp_path):
    """BASELINE.json configs[0]: 4.6 Mbp genome 30x, 13.8 k reads against themselves (-h 1-N -r 1-N), ~3.9e5 overlaps"""
    _both("c1", str(tmp_path), 300000)


@needs_tools
def test_c4_block_pair_records_equal_reference(tmp_path):
    """BASELINE.json configs[3]: one full 667 k-read hash block x one 667 k-read query block of 150 bp reads, -G"""
    _both("c4", str(tmp_path), 500000)


# ------------------------------------------------------------------------------------------------ receive path

def _sorted(rec):
    r = np.ascontiguousarray(np.asarray(rec, dtype=np.uint32).reshape(-1, 6))
    return r[np.lexsort([r[:, i] for i in range(5, -1, -1)])]


def _oracle_records(reads, P):
    res = H.run_oracle(reads, P)
    return sorted(H.format_raw(o) for o in res["overlaps"])


def _formatted(rec):
    from canu_b200 import api
    d = api.decode_records(rec)
    return sorted(H.format_raw({k: int(d[k][i]) for k in d}) for i in range(len(d["a_iid"])))


def test_received_hash_block_gives_oracle_records_one_gpu():
    """A context that never builds: it RESERVES the block from another context's meta, receives the three buffers by a
    device-to-device copy, COMMITS (rebuilding its occupancy bitmap) and searches.  Records, -s counters and work counters
    must equal the oracle's, as on the context that built the block."""
    import torch
    from canu_b200 import api
    reads = H.gen_reads(60000, 14, 1000, 3000, 0.02, 555)
    n_h = reads.n // 2
    P = H.default_params(hash_bits=18, hash_load=0.75, max_hash_strings=n_h, bgn_hash_id=reads.n - n_h + 1, end_hash_id=reads.n,
                         bgn_ref_id=1, end_ref_id=reads.n - n_h, num_threads=3)
    want = _oracle_records(reads, P)
    assert len(want) > 200
    a, b = api.Overlapper(P), api.Overlapper(P)
    for ov in (a, b):
        ov.load_reads(reads.bases, reads.offsets[:-1], reads.lengths)
    a.build_hash_index(P["bgn_hash_id"], P["end_hash_id"])
    meta = a.hash_block_meta()
    b.hash_block_reserve(meta)

    class View:
        def __init__(self, ptr, nbytes):
            self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}
    for (pa, na), (pb, _) in zip(a.hash_block_buffers(), b.hash_block_buffers()):
        torch.as_tensor(View(pb, na), device="cuda").copy_(torch.as_tensor(View(pa, na), device="cuda"))
    torch.cuda.synchronize()
    b.hash_block_commit(meta)
    last = a.last_processed_ref(P["bgn_ref_id"], P["end_ref_id"], P["num_threads"])
    for ov in (a, b):
        ov.process_overlaps(P["bgn_ref_id"], last)
    ra, rb = _sorted(a.overlaps()), _sorted(b.overlaps())
    sa, sb = a.stats(), b.stats()
    a.close()
    b.close()
    assert np.array_equal(ra, rb)
    assert _formatted(rb) == want
    for k in ("kmer_hits_without_olap", "kmer_hits_with_olap", "multi_overlap", "total_overlaps", "contained", "dovetail",
              "n_windows", "n_found", "n_refs", "n_add_ref", "ext_calls", "ext_cells", "ext_words"):
        assert sa[k] == sb[k], k


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def test_nccl_replicated_hash_block_two_ranks(tmp_path):
    """2 ranks (needs 2 GPUs): rank 0 builds, oic_hash_block_replicate (ncclBroadcast inside the C ABI) ships the block, rank 1
    searches a table it only RECEIVED; each rank's records equal the oracle's for its query block, and `bench.py --check`
    asserts the gathered record multiset at 2 ranks against the reference binary."""
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    script = os.path.join(H.ROOT, "tests", "nccl_receive_worker.py")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29631", script], capture_output=True, timeout=900)
    assert r.returncode == 0, r.stdout.decode()[-2000:] + r.stderr.decode()[-2000:]
    assert r.stdout.decode().count("RECEIVE_OK") == 2
