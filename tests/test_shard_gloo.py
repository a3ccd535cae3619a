"""world_size-2 gloo test (CPU) of the multi-GPU host logic in canu_b200/shard.py: two ranks each take one query block and
the shared hash block in a LOCAL ID space, their records are mapped back to store IDs and gathered; the union must equal one
unsharded run over the same hash block.  The per-rank "overlapper" here is the CPU oracle (this is a test: the product path
has no CPU implementation) -- what is under test is the split, the ID mapping and the gather."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers as H
from canu_b200 import shard


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _pack(ov):
    from test_host_stores import pack_records
    return pack_records(ov)


def _worker(rank, world, port, outdir, interleave=False):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    reads = H.gen_reads(30000, 3 * 6, 800, 2000, 0.02, 808)
    blocks = shard.split_blocks(reads.n, world + 1)
    job = shard.RankJob(blocks, rank, interleave=interleave)
    bases, offs, lens = job.select(reads.bases, reads.offsets[:-1], reads.lengths)
    P = H.default_params(hash_bits=16, hash_load=0.75, end_hash_id=reads.n, end_ref_id=reads.n, max_hash_strings=reads.n)
    LP = job.local_params(P)
    sub = H.Reads(bases, np.concatenate([offs, [offs[-1] + lens[-1]]]).astype(np.uint64), lens)
    rec = job.to_global(_pack(H.run_oracle(sub, LP)["overlaps"]))
    allrec = shard.gather_records(rec, dist)
    if rank == 0:
        np.save(os.path.join(outdir, "gathered.npy"), allrec)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("interleave", [False, True], ids=["blocks", "interleaved"])
def test_two_rank_split_equals_unsharded(tmp_path, interleave):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path), interleave), nprocs=world, join=True)
    got = np.load(str(tmp_path / "gathered.npy"))
    reads = H.gen_reads(30000, 3 * 6, 800, 2000, 0.02, 808)
    blocks = shard.split_blocks(reads.n, world + 1)
    hb, he = blocks[-1]
    P = H.default_params(hash_bits=16, hash_load=0.75, max_hash_strings=reads.n, bgn_hash_id=hb, end_hash_id=he,
                         bgn_ref_id=1, end_ref_id=hb - 1)
    want = _pack(H.run_oracle(reads, P)["overlaps"])
    assert len(want) > 100
    key = lambda r: sorted(map(tuple, r.tolist()))  # noqa: E731
    assert key(got) == key(want)


def test_split_blocks_and_id_mapping():
    blocks = shard.split_blocks(1003, 4)
    assert blocks == [(1, 250), (251, 500), (501, 750), (751, 1003)]
    job = shard.RankJob(blocks, 1)
    assert (job.n_q, job.n_h) == (250, 253)
    rec = np.zeros((2, 6), dtype=np.uint32)
    rec[0, :2] = (1, 251)       # first query, first hash read
    rec[1, :2] = (250, 503)     # last query, last hash read
    g = job.to_global(rec)
    assert g[:, :2].tolist() == [[251, 751], [500, 1003]]
    LP = job.local_params(dict(max_hash_strings=1))
    assert (LP["bgn_ref_id"], LP["end_ref_id"], LP["bgn_hash_id"], LP["end_hash_id"]) == (1, 250, 251, 503)


def test_interleaved_id_mapping():
    """rank r of N takes query reads r+1, r+1+N, ... of ALL query blocks; local IDs map back to those store IDs"""
    blocks = shard.split_blocks(1003, 4)            # hash block 751-1003, query reads 1-750
    jobs = [shard.RankJob(blocks, r, interleave=True) for r in range(3)]
    assert sorted(int(i) for j in jobs for i in j.q_ids) == list(range(1, 751))
    assert [j.n_q for j in jobs] == [250, 250, 250]
    job = jobs[1]
    assert job.q_ids[:3].tolist() == [2, 5, 8] and not job.contiguous
    rec = np.zeros((3, 6), dtype=np.uint32)
    rec[0, :2] = (1, 251)        # first query of the rank, first hash read
    rec[1, :2] = (503, 250)      # last hash read (local 250 + 253), last query of the rank -- either column may hold either
    rec[2, :2] = (3, 252)
    g = job.to_global(rec)
    assert g[:, :2].tolist() == [[2, 751], [1003, 749], [8, 752]]
    # select() lays the strided reads out back to back, in rank order
    lengths = np.arange(1, 1004, dtype=np.uint32)
    offsets = np.zeros(1003, dtype=np.uint64)
    np.cumsum(lengths[:-1], out=offsets[1:])
    bases = np.repeat((np.arange(1003) % 251).astype(np.uint8), lengths)
    out, offs, lens = job.select(bases, offsets, lengths)
    assert lens[:3].tolist() == [2, 5, 8] and lens[250] == 751
    assert out[int(offs[1])] == (5 - 1) % 251 and out[int(offs[250])] == 750 % 251
