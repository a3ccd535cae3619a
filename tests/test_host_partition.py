"""canu_b200/bin/overlapInCorePartition (canu_b200/host/partition.cpp) against the reference's overlapInCorePartition
(OIC/overlapInCorePartition.C, built by oracle/build_ref.sh) on the same gkpStore: the three job files must be identical
byte for byte, for every way of sizing hash and query blocks the pipeline uses (OverlapInCore.pm:123-134)."""
import os
import subprocess

import pytest

import helpers as H

MINE = os.path.join(H.ROOT, "canu_b200", "bin", "overlapInCorePartition")
REF = os.path.join(H.REF_DIR, "overlapInCorePartition")

OPTION_SETS = [
    ["-bl", "60000", "-rs", "40", "-ol", "500"],          # canu's utg/obt shape: hash blocks by bases, query blocks by reads
    ["-bl", "25000", "-rl", "30000", "-ol", "900"],
    ["-bs", "37", "-rs", "50", "-ol", "0"],
    ["-bs", "64", "-rl", "45000", "-ol", "1200"],
    ["-bl", "10000000", "-rs", "2000000", "-ol", "500"],  # one job
    ["-bl", "40000", "-rs", "30", "-ol", "500", "-H", "1", "-R", "1"],   # same library on both sides: full square, not the triangle
    ["-bs", "50", "-rs", "50", "-H", "1"],
]


@pytest.mark.skipif(not (H.have_ref() and os.access(REF, os.X_OK)), reason="reference tools not built (oracle/build_ref.sh)")
@pytest.mark.parametrize("opts", OPTION_SETS, ids=[" ".join(o) for o in OPTION_SETS])
def test_partition_files_equal_reference(opts, tmp_path):
    if not os.access(MINE, os.X_OK):
        subprocess.check_call(["make", "-C", os.path.join(H.ROOT, "canu_b200", "host"), "../bin/overlapInCorePartition"])
    reads = H.gen_reads(genome_len=30000, coverage=9, len_min=200, len_max=2600, err=0.01, seed=4)
    store = H.make_store(reads, str(tmp_path))
    out = {}
    for who, exe in (("ref", REF), ("mine", MINE)):
        prefix = str(tmp_path / who)
        subprocess.check_call([exe, "-g", store] + opts + ["-o", prefix], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        out[who] = [open(prefix + "." + t).read() for t in ("ovlbat", "ovljob", "ovlopt")]
    assert out["ref"][2].count("\n") >= 1
    assert out["mine"] == out["ref"]
