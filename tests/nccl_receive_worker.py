"""torchrun worker of tests/test_gpu_fullsize_ref.py::test_nccl_replicated_hash_block_two_ranks (one process per GPU).

Every rank loads the same reads.  Two ways of replicating the hash block through the library's own NCCL communicator are
checked record for record against the CPU oracle (test infrastructure), each rank searching its own half of the queries:

* broadcast: rank 0 builds (oic_build_hash_index), oic_hash_block_replicate ships it, rank 1 searches a table it only RECEIVED;
* sharded:   oic_build_hash_index_sharded -- every rank builds the slices it owns, ncclAllGather exchanges them; the gathered
             index (keys, occurrence lists in chain order, Empty flags) must equal the oracle's index on EVERY rank, with and
             without a skip-k-mer list, and also when the hash-load limit trips (fallback: build on rank 0 + broadcast).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))

import helpers as H  # noqa: E402


def formatted(rec):
    from canu_b200 import api
    d = api.decode_records(rec)
    return sorted(H.format_raw({k: int(d[k][i]) for k in d}) for i in range(len(d["a_iid"])))


def check_index(ov, want):
    got = ov.dump_index()
    assert np.array_equal(got["keys"], want["keys"]), "key set"
    assert np.array_equal(got["occ_start"], want["occ_start"]), "list lengths"
    assert np.array_equal(got["occ"], want["occ"]), "occurrence order"
    assert np.array_equal(got["flags"], want["flags"]), "Empty flags"


def main():
    import torch
    import torch.distributed as dist
    from canu_b200 import api
    import cases
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def fresh_uid():   # one ncclUniqueId per communicator: rank 0 draws it, torch.distributed ships the 128 bytes
        idt = torch.zeros(api.COMM_ID_BYTES, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(api.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        return idt.cpu().numpy().tobytes()

    for mode, case, hash_bits in (("broadcast", "plain", 18), ("sharded", "plain", 18), ("sharded", "skip", 18), ("sharded", "loadlimit", 11)):
        if case == "skip":
            reads, P, erate, threads, skip = cases.build_case("repeats_skip")
            n_h = reads.n // 2
            P = dict(P, max_hash_strings=n_h, bgn_hash_id=reads.n - n_h + 1, end_hash_id=reads.n, bgn_ref_id=1, end_ref_id=reads.n - n_h, num_threads=3)
        else:
            reads = H.gen_reads(60000, 14, 1000, 3000, 0.02, 556)
            skip = None
            n_h = reads.n // 2
            P = H.default_params(hash_bits=hash_bits, hash_load=0.75, max_hash_strings=n_h, bgn_hash_id=reads.n - n_h + 1, end_hash_id=reads.n,
                                 bgn_ref_id=1, end_ref_id=reads.n - n_h, num_threads=3)
        ov = api.Overlapper(P, device=local)
        ov.comm_init_rank(fresh_uid(), world, rank)
        assert ov.comm_info() == (world, rank)
        ov.load_reads(reads.bases, reads.offsets[:-1], reads.lengths)
        if skip:
            ov.set_skip_kmers(skip)
        if mode == "broadcast":
            if rank == 0:
                ov.build_hash_index(P["bgn_hash_id"], P["end_hash_id"])
            ov.hash_block_replicate(0)
            last = ov.hash_block_meta().last_id
        else:
            last = ov.build_hash_index_sharded(P["bgn_hash_id"], P["end_hash_id"])
        # the oracle's first hash block (the load limit may end it early) and this rank's half of the queries against it
        PO = dict(P, end_hash_id=last)
        if case == "loadlimit":
            assert last < P["end_hash_id"], "the load limit was meant to trip"
        q_lo, q_hi = 1, P["end_ref_id"]
        mid = (q_lo + q_hi) // 2
        mine = (q_lo, mid) if rank == 0 else (mid + 1, q_hi)
        # a -t for which the reference's scheduler processes the whole range (Process_Overlaps.C:86 may drop the last read)
        t_full = next(t for t in range(1, 200) if ov.last_processed_ref(mine[0], mine[1], t) == mine[1])
        want = H.run_oracle(reads, dict(PO, bgn_ref_id=mine[0], end_ref_id=mine[1], num_threads=t_full), dump=4, skip_kmers=skip)
        assert want["stats"]["n_hash_blocks"] == 1 and want["stats"]["last_end_hash_id"] == last
        check_index(ov, want["index"])
        ov.process_overlaps(*mine)
        got = formatted(ov.overlaps())
        exp = sorted(H.format_raw(o) for o in want["overlaps"])
        assert len(exp) > 50, (mode, case, len(exp))
        assert got == exp, (mode, case, rank, len(got), len(exp))
        ov.close()
        dist.barrier()
    print("RECEIVE_OK rank %d" % rank, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
