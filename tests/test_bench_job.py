"""bench.gen_job: the job of a multi-GPU bench step is the SAME at every N where it can be: the hash block and
rank r's query block do not change with the number of ranks, IDs are laid out [Q0 .. Q(N-1) | H] with every query ID below every
hash ID (OIC/overlapInCore-Find_Overlaps.C:316-318), and nothing of the generated set is lost or duplicated."""
import hashlib

import numpy as np

import helpers as H  # noqa: F401  (puts the repo root on sys.path)


def _block_digest(reads, blk):
    o0, o1 = int(reads.offsets[blk[0] - 1]), int(reads.offsets[blk[1]])
    return hashlib.md5(reads.bases[o0:o1].tobytes()).hexdigest(), blk[1] - blk[0] + 1


def test_job_is_the_same_at_every_rank_count():
    import bench
    bench.set_workload("c3")
    scale = 0.01
    genome, block = int(bench.GENOME * scale), int(bench.BLOCK_BASES * scale)
    seen = {}
    for world in (1, 2, 3, 5):
        reads, blocks = bench.gen_job(genome, world, block, bench.SEED)
        assert len(blocks) == world + 1
        # contiguous, ascending, the hash block last and the same size as a query block
        assert blocks[0][0] == 1 and blocks[-1][1] == reads.n
        for a, b in zip(blocks, blocks[1:]):
            assert b[0] == a[1] + 1
        assert int(reads.offsets[-1]) == len(reads.bases) and len(reads.lengths) == reads.n
        assert np.array_equal(np.diff(reads.offsets.astype(np.int64)), reads.lengths.astype(np.int64))
        d = [_block_digest(reads, b) for b in blocks]
        seen.setdefault("H", d[-1])
        assert d[-1] == seen["H"], "the hash block changed with the number of ranks"
        for r in range(world):
            seen.setdefault(r, d[r])
            assert d[r] == seen[r], "query block %d changed with the number of ranks" % r
        assert len({x[0] for x in d}) == world + 1, "two blocks hold the same reads"


def test_job_blocks_cover_the_generated_reads_once():
    import bench
    bench.set_workload("c3")
    scale = 0.01
    genome, block = int(bench.GENOME * scale), int(bench.BLOCK_BASES * scale)
    g = bench.gen_blocks(genome, 3, block, bench.SEED)
    reads, blocks = bench.gen_job(genome, 2, block, bench.SEED)
    per = blocks[-1][1] - blocks[-1][0] + 1
    # [Q0 | Q1 | H] = generated reads [per, 3 per) followed by [0, per)
    want = np.concatenate([g.lengths[per:reads.n], g.lengths[:per]])
    assert np.array_equal(reads.lengths, want)
    assert reads.seq(blocks[-1][0]) == g.seq(1) and reads.seq(1) == g.seq(per + 1)
