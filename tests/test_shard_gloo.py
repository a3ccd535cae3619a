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


def _worker(rank, world, port, outdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    reads = H.gen_reads(30000, 3 * 6, 800, 2000, 0.02, 808)
    blocks = shard.split_blocks(reads.n, world + 1)
    job = shard.RankJob(blocks, rank)
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


def test_two_rank_split_equals_unsharded(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
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
