"""Host-side logic of the multi-GPU split: which reads a rank loads, the local ID space it runs in,
the mapping of its overlap records back to store IDs, and the host gather of per-rank records.

The split is the one the reference pipeline already makes between processes (overlapInCorePartition: independent
(hash range, query range) jobs, OIC/overlapInCorePartition.C:388-) applied between the GPUs of one box: the hash block is the
same on every rank, query blocks are disjoint.  No collective touches the search itself.
"""
import numpy as np


def split_blocks(n_reads, n_blocks):
    """contiguous 1-based read-ID ranges of equal read count; blocks 0..n-2 are query blocks, the last is the hash block"""
    per = n_reads // n_blocks
    return [(i * per + 1, (i + 1) * per if i < n_blocks - 1 else n_reads) for i in range(n_blocks)]


class RankJob:
    """Rank `rank` of `world` = len(blocks) - 1: its share of the query reads against the hash block (the last block).

    interleave = False: query block `rank`, a contiguous ID range.
    interleave = True : every world-th read of ALL query blocks (IDs rank+1, rank+1+world, ...).  The step of a multi-GPU run is the
        maximum over the ranks, and whole blocks differ by several per cent in extension work (measured: 12.8 to 13.6 G DP cells for
        100 Mbp blocks of the same read model); dealt out read by read -- as the reference deals small blocks of reads to its threads,
        OIC/overlapInCore-Process_Overlaps.C:86-172 -- every rank gets the same mix.  The job (all query blocks against the hash block)
        and its record set are the same either way.

    Local IDs: queries 1..n_q, hash reads n_q+1..n_q+n_h (every query ID below every hash ID, as in the triangular sweep,
    so the `query < hash` rule of OIC/overlapInCore-Find_Overlaps.C:316-318 keeps every pair)."""

    def __init__(self, blocks, rank, interleave=False):
        self.hb, self.he = blocks[-1]
        world = len(blocks) - 1
        if interleave:
            self.q_ids = np.arange(1 + rank, self.hb, world, dtype=np.int64)
        else:
            self.q_ids = np.arange(blocks[rank][0], blocks[rank][1] + 1, dtype=np.int64)
        self.qb, self.qe = int(self.q_ids[0]), int(self.q_ids[-1])
        self.n_q = len(self.q_ids)
        self.n_h = self.he - self.hb + 1
        self.contiguous = bool(self.qe - self.qb + 1 == self.n_q)
        self.ids = np.concatenate([self.q_ids, np.arange(self.hb, self.he + 1, dtype=np.int64)])

    def local_params(self, P):
        LP = dict(P)
        LP.update(bgn_hash_id=self.n_q + 1, end_hash_id=self.n_q + self.n_h, bgn_ref_id=1, end_ref_id=self.n_q,
                  max_hash_strings=self.n_h)
        return LP

    def select(self, bases, offsets, lengths, out=None):
        """this rank's reads, concatenated: (bases uint8, offsets uint64[n], lengths uint32[n]); `out` may be a pinned buffer"""
        lens = np.asarray(lengths)[self.ids - 1].astype(np.uint32)
        offs = np.zeros(len(lens), dtype=np.uint64)
        np.cumsum(lens[:-1], out=offs[1:])
        total = int(lens.sum())
        if out is None:
            out = np.empty(total, dtype=np.uint8)
        w = 0
        if self.contiguous:
            spans = [(self.qb, self.qe)]
        else:
            spans = [(int(i), int(i)) for i in self.q_ids]
        for (b, e) in spans + [(self.hb, self.he)]:
            o0, o1 = int(offsets[b - 1]), int(offsets[e - 1]) + int(lengths[e - 1])
            out[w:w + (o1 - o0)] = bases[o0:o1]
            w += o1 - o0
        assert w == total
        return out[:total], offs, lens

    def to_global(self, rec):
        """(n, 6) uint32 .ovb words with local IDs -> store IDs (words 0 and 1 are a_iid, b_iid: ST/ovStoreFile.C:196-221)"""
        if self.contiguous:
            # one offset per side, applied in place (`rec` is the caller's fresh copy from Overlapper.overlaps()): a short-read block
            # returns a million records per step and this runs inside the end-to-end timing
            rec = np.asarray(rec, dtype=np.uint32).reshape(-1, 6)
            dq = np.uint32(int(self.q_ids[0]) - 1 if self.n_q else 0)
            dh = np.uint32(self.hb - self.n_q - 1)
            if dq or dh:
                if not rec.flags.writeable:
                    rec = rec.copy()
                for col in (0, 1):
                    v = rec[:, col]
                    v += np.where(v <= np.uint32(self.n_q), dq, dh).astype(np.uint32)
            return rec
        rec = np.array(rec, dtype=np.uint32).reshape(-1, 6)
        for col in (0, 1):
            v = rec[:, col].astype(np.int64)
            is_q = v <= self.n_q
            g = np.where(is_q, self.q_ids[np.clip(v - 1, 0, self.n_q - 1)], v - self.n_q - 1 + self.hb)
            rec[:, col] = g.astype(np.uint32)
        return rec


def gather_records(rec, dist, group=None, dst=0, device="cpu"):
    """Host-side merge of the per-rank overlap records ("gathered on host"): every rank calls with its (n, 6)
    uint32 records already in host memory; rank `dst` gets the concatenation in rank order, the others None.  `group` should be
    a gloo group (CPU tensors: nothing goes back through a GPU); with the default group of an NCCL job pass device='cuda'."""
    import torch
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    rec = np.ascontiguousarray(rec, dtype=np.uint32).reshape(-1, 6)
    n = torch.tensor([rec.shape[0]], dtype=torch.int64, device=device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    cap = max(max(sizes), 1)
    pad = np.zeros((cap, 6), dtype=np.int32)
    pad[:rec.shape[0]] = rec.view(np.int32)
    mine = torch.from_numpy(pad).to(device)
    parts = [torch.zeros_like(mine) for _ in range(world)] if rank == dst else None
    dist.gather(mine, parts, dst=dst, group=group)
    if rank != dst:
        return None
    out = [p.cpu().numpy().view(np.uint32)[:sizes[r]] for r, p in enumerate(parts)]
    return np.concatenate(out, axis=0)
