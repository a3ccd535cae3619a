"""GPU end-to-end test of the drop-in executable canu_b200/bin/overlapInCore: gkpStore in, .ovb/.counts/.stats out,
compared byte-for-byte (after the reference's own overlapConvert | sort) with the reference binary's outputs in
tests/golden/.  The store is written by the reference's gatekeeperCreate (oracle/_ref, prebuilt, travels to the box).
"""
import gzip
import hashlib
import json
import os
import subprocess

import pytest

import cases
import helpers as H

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BIN = os.path.join(H.ROOT, "canu_b200", "bin", "overlapInCore")

needs_tools = pytest.mark.skipif(not (H.have_ref() and os.access(BIN, os.X_OK)),
                                 reason="oracle/_ref tools or canu_b200/bin/overlapInCore not built")


def run_case(name, wd, extra=()):
    reads, P, erate, threads, skip = cases.build_case(name)
    store = H.make_store(reads, wd)
    args = H.oic_args(P, erate)
    if skip:
        sf = os.path.join(wd, "skip.fasta")
        cases.write_skip_file(skip, sf)
        args += ["-k", sf]
    ovb, st = H.run_overlapper(BIN, store, os.path.join(wd, "mine"), args + list(extra), threads)
    return store, ovb, st


@needs_tools
@pytest.mark.parametrize("name", list(cases.CASES))
def test_executable_vs_golden(name, tmp_path):
    store, ovb, st = run_case(name, str(tmp_path))
    got = H.dump_ovb_sorted(store, ovb)
    want = gzip.open(os.path.join(GOLDEN, name + ".txt.gz")).read().decode().splitlines()
    assert got == want
    meta = json.load(open(os.path.join(GOLDEN, name + ".json")))
    mine, ref = H.parse_stats(st), dict(meta["stats"])
    if "-G" in H.oic_args(cases.build_case(name)[1], "0.06"):
        # -G: the reference bumps an unsynchronised global from every thread (OIC/overlapInCore-Output.C:253), so its
        # own total is racy (<= the record count); ours is the record count
        assert mine.pop("Total overlaps produced") == len(want) >= ref.pop("Total overlaps produced")
    assert mine == ref
    assert open(st).read().count("\n") == 8
    assert hashlib.md5(open(os.path.join(str(tmp_path), "mine.counts"), "rb").read()).hexdigest() == meta["counts_md5"]


@needs_tools
def test_executable_usage_errors(tmp_path):
    """no -k / no -o: usage on stderr and exit status 1, like the reference (OIC/overlapInCore.C:423-470)"""
    r = subprocess.run([BIN, "-o", str(tmp_path / "x.ovb"), "nostore"], capture_output=True)
    assert r.returncode == 1 and b"No kmer length supplied" in r.stderr
    r = subprocess.run([BIN, "-k", "22", "nostore"], capture_output=True)
    assert r.returncode == 1 and b"No output file name" in r.stderr


@needs_tools
def test_executable_two_shards_same_gpu_count(tmp_path):
    """--gpus larger than the box has must fail loudly, not fall back"""
    import torch
    n = torch.cuda.device_count()
    reads, P, erate, threads, skip = cases.build_case("c4_short")
    store = H.make_store(reads, str(tmp_path))
    r = subprocess.run([BIN, "--gpus", str(n + 1)] + H.oic_args(P, erate) + ["-o", str(tmp_path / "x.ovb"), store], capture_output=True)
    assert r.returncode != 0


@needs_tools
@pytest.mark.parametrize("part_opts", [["-bl", "40000", "-rs", "60", "-ol", "500"], ["-bs", "45", "-rl", "70000", "-ol", "500"]],
                         ids=["bl_rs", "bs_rl"])
def test_sweep_equals_reference_jobs(part_opts, tmp_path):
    """--sweep: every job of the reference's own overlapInCorePartition run inside ONE process (reads resident, hash blocks
    reused across the query blocks of a hash block) writes, per job, the same overlaps / .counts / .stats as a separate
    process of the REFERENCE overlapInCore started with that job's options (OverlapInCore.pm:184-233)."""
    wd = str(tmp_path)
    reads = H.gen_reads(genome_len=40000, coverage=9, len_min=600, len_max=2600, err=0.015, seed=77)
    store = H.make_store(reads, wd)
    prefix = os.path.join(wd, "asm.partition")
    subprocess.check_call([os.path.join(H.REF_DIR, "overlapInCorePartition"), "-g", store] + part_opts + ["-o", prefix],
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    bats = open(prefix + ".ovlbat").read().split()
    jobs = open(prefix + ".ovljob").read().split()
    opts = open(prefix + ".ovlopt").read().splitlines()
    assert len(jobs) >= 4
    common = ["-k", "22", "--hashbits", "17", "--hashload", "0.75", "--maxerate", "0.06", "--minlength", "500"]
    out = os.path.join(wd, "sweep")
    with open(os.path.join(wd, "sweep.err"), "w") as ef:
        subprocess.check_call([BIN, "-t", "3"] + common + ["--sweep", prefix, "--sweep-out", out, store], stderr=ef, stdout=ef)
    total = 0
    for bat, job, opt in zip(bats, jobs, opts):
        ref_prefix = os.path.join(wd, "ref_" + job)
        ovb, st = H.run_overlapper(os.path.join(H.REF_DIR, "overlapInCore"), store, ref_prefix, common + opt.split(), 3)
        want = H.dump_ovb_sorted(store, ovb)
        base = os.path.join(out, bat, job)
        assert os.path.exists(base + ".ovb") and not os.path.exists(base + ".ovb.WORKING")
        got = H.dump_ovb_sorted(store, base + ".ovb")
        assert got == want, "job %s (%s)" % (job, opt)
        assert H.parse_stats(base + ".stats") == H.parse_stats(st)
        assert open(base + ".counts", "rb").read() == open(ref_prefix + ".counts", "rb").read()
        total += len(want)
    assert total > 1000


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@needs_tools
def test_multi_gpu_executable_equals_single(tmp_path):
    """N > 1 (needs a box with >= 2 GPUs): `--gpus 2` splits the query range of a plain job over two GPUs, and in --sweep mode
    hands WHOLE jobs to the GPUs from a shared list; both must write the same overlaps / .counts / .stats as one GPU."""
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    wd = str(tmp_path)
    reads = H.gen_reads(genome_len=40000, coverage=9, len_min=600, len_max=2600, err=0.015, seed=78)
    store = H.make_store(reads, wd)
    common = ["-k", "22", "--hashbits", "17", "--hashload", "0.75", "--maxerate", "0.06", "--minlength", "500"]
    # plain job, query range sharded
    res = {}
    for g in (1, 2):
        pre = os.path.join(wd, "plain%d" % g)
        subprocess.check_call([BIN, "-t", "3", "--gpus", str(g)] + common + ["-h", "1-%d" % reads.n, "-r", "1-%d" % reads.n,
                               "--hashstrings", "70", "--hashdatalen", "100000000", "-o", pre + ".ovb", "-s", pre + ".stats", store],
                              stderr=subprocess.DEVNULL, stdout=subprocess.DEVNULL)
        res[g] = (H.dump_ovb_sorted(store, pre + ".ovb"), H.parse_stats(pre + ".stats"), open(pre + ".counts", "rb").read())
    assert len(res[1][0]) > 500
    assert res[1] == res[2]
    # sweep, whole jobs per GPU
    prefix = os.path.join(wd, "asm.partition")
    subprocess.check_call([os.path.join(H.REF_DIR, "overlapInCorePartition"), "-g", store, "-bl", "40000", "-rs", "60", "-ol", "500", "-o", prefix],
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    bats = open(prefix + ".ovlbat").read().split()
    jobs = open(prefix + ".ovljob").read().split()
    # one GPU; two GPUs (one host thread per GPU, each taking whole jobs)
    for name, g, env in (("sw1", 1, {}), ("sw2", 2, {})):
        subprocess.check_call([BIN, "-t", "3", "--gpus", str(g)] + common + ["--sweep", prefix, "--sweep-out", os.path.join(wd, name), store],
                              stderr=subprocess.DEVNULL, stdout=subprocess.DEVNULL, env=dict(os.environ, **env))
    for bat, job in zip(bats, jobs):
        a = os.path.join(wd, "sw1", bat, job)
        for other in ("sw2",):
            b = os.path.join(wd, other, bat, job)
            assert os.path.exists(b + ".ovb") and not os.path.exists(b + ".ovb.WORKING")
            assert H.dump_ovb_sorted(store, a + ".ovb") == H.dump_ovb_sorted(store, b + ".ovb"), job
            assert H.parse_stats(a + ".stats") == H.parse_stats(b + ".stats")
            assert open(a + ".counts", "rb").read() == open(b + ".counts", "rb").read()
    # a sweep process that cannot get its GPU fails the whole sweep
    r = subprocess.run([BIN, "-t", "3", "--gpus", str(_n_gpus() + 1)] + common + ["--sweep", prefix, "--sweep-out", os.path.join(wd, "swbad"), store],
                       capture_output=True)
    assert r.returncode != 0


@needs_tools
@pytest.mark.parametrize("name,K,T", [("n_bases", 22, 7), ("c4_short", 14, 12)])
def test_frequent_mers_tool_vs_meryl(name, K, T, tmp_path):
    """canu_b200/bin/frequentMers on a gkpStore writes the bytes of the reference's `meryl -Dt -n T` (golden fixture), and of the
    reference's meryl run on the spot when it is there"""
    import gzip
    reads = cases.build_case(name)[0]
    store = H.make_store(reads, str(tmp_path))
    out = str(tmp_path / "freq.fasta")
    hist, info = str(tmp_path / "hist"), str(tmp_path / "hist.info")
    subprocess.check_call([os.path.join(os.path.dirname(BIN), "frequentMers"), "-m", str(K), "-n", str(T), "-s", store, "-o", out,
                           "--histogram", hist, "--info", info], stderr=subprocess.DEVNULL)
    gold_h = gzip.open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "khist_%s_%d.txt.gz" % (name, K))).read().decode()
    assert open(hist).read() + "--info--\n" + open(info).read() == gold_h
    got = open(out).read()
    gold = gzip.open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kmers_%s_%d_%d.txt.gz" % (name, K, T))).read().decode()
    assert got == gold
    assert got == H.run_meryl(store, K, T, str(tmp_path))


# ------------------------------------------------------------------------------------------------ overlap store
OVSB = os.path.join(os.path.dirname(BIN), "ovStoreBuild")
needs_store_tools = pytest.mark.skipif(not (H.have_ref() and os.access(OVSB, os.X_OK) and os.access(os.path.join(H.REF_DIR, "ovStoreBuild"), os.X_OK)),
                                       reason="ovStoreBuild (ours or the reference's) not built")
STORE_FILES = ["info", "index", "0001", "evalueLen", "overlapScores"]


@needs_store_tools
@pytest.mark.parametrize("name,extra", [("c1_small", []), ("c4_short", []), ("quirk_m", []), ("c2_noisy", ["-e", "0.1"])])
def test_ovStoreBuild_equals_reference(name, extra, tmp_path):
    """canu_b200/bin/ovStoreBuild (filter + mirror + sort on the GPU, oic_ovlstore_sort) on the .ovb our overlapInCore wrote vs the
    reference's ovStoreBuild on the reference overlapInCore's .ovb of the same job: every file of the two stores is identical,
    and so is the reference's own `ovStoreDump -d` of them."""
    wd = str(tmp_path)
    store, ovb, _ = run_case(name, wd)
    reads, P, erate, threads, skip = cases.build_case(name)
    args = H.oic_args(P, erate)
    if skip:
        args += ["-k", os.path.join(wd, "skip.fasta")]
    ref_ovb, _ = H.run_overlapper(os.path.join(H.REF_DIR, "overlapInCore"), store, os.path.join(wd, "ref"), args, threads)
    mine, ref = os.path.join(wd, "mine.ovlStore"), os.path.join(wd, "ref.ovlStore")
    subprocess.check_call([OVSB, "-O", mine, "-G", store] + extra + [ovb], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    subprocess.check_call([os.path.join(H.REF_DIR, "ovStoreBuild"), "-O", ref, "-G", store] + extra + [ref_ovb],
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    assert sorted(os.listdir(mine)) == sorted(os.listdir(ref)) == sorted(STORE_FILES)
    for f in STORE_FILES:
        assert open(os.path.join(mine, f), "rb").read() == open(os.path.join(ref, f), "rb").read(), f
    dump = [subprocess.check_output([os.path.join(H.REF_DIR, "ovStoreDump"), "-G", store, "-O", s, "-d", "-raw"], stderr=subprocess.DEVNULL)
            for s in (mine, ref)]
    assert dump[0] == dump[1] and dump[0].count(b"\n") > 100
    # ... and our tool reads the reference's snappy-compressed .ovb as well as its own
    mine2 = os.path.join(wd, "mine2.ovlStore")
    subprocess.check_call([OVSB, "-O", mine2, "-G", store] + extra + [ref_ovb], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    for f in STORE_FILES:
        assert open(os.path.join(mine2, f), "rb").read() == open(os.path.join(ref, f), "rb").read(), f


def test_ovlstore_sort_vs_numpy_restatement():
    """oic_ovlstore_sort through the C ABI on synthetic records that exercise every filter branch (flipped / unflipped, OBT and DUP
    flags, short and near-identical overlaps, two libraries, evalue cut, duplicate (a, b) pairs whose order the data words decide)"""
    import numpy as np
    from canu_b200 import api
    from test_host_stores import mirror_filter_sort, pack_records
    rng = np.random.default_rng(2024)
    n, n_reads = 200000, 3000
    length = np.zeros(n_reads + 1, dtype=np.uint32)
    length[1:] = rng.integers(150, 4000, n_reads)
    library = np.zeros(n_reads + 1, dtype=np.uint8)
    library[1:] = rng.integers(1, 3, n_reads)
    skip_dup = np.zeros(n_reads + 1, dtype=np.uint8)
    skip_dup[library == 2] = 1
    ov = np.zeros(n, dtype=H.OVERLAP_DT)
    ov["a_iid"] = rng.integers(1, n_reads + 1, n)
    ov["b_iid"] = rng.integers(1, 40, n) + (ov["a_iid"] % 7)       # many equal (a, b) pairs
    ov["b_iid"] = np.clip(ov["b_iid"], 1, n_reads)
    la, lb = length[ov["a_iid"]], length[ov["b_iid"]]
    ov["ahg5"] = rng.integers(0, 120, n) % la
    ov["ahg3"] = rng.integers(0, 60, n) % (la - ov["ahg5"])
    ov["bhg5"] = rng.integers(0, 120, n) % lb
    ov["bhg3"] = rng.integers(0, 60, n) % (lb - ov["bhg5"])
    ov["span"] = rng.integers(0, 4000, n)
    ov["evalue"] = rng.integers(0, 1500, n)
    ov["flipped"] = rng.integers(0, 2, n)
    ov["for_obt"] = rng.integers(0, 2, n)
    ov["for_dup"] = rng.integers(0, 2, n)
    ov["for_utg"] = rng.integers(0, 2, n)
    rec = pack_records(ov)
    rec[-500:] = rec[:500]                                          # exact duplicates too
    ovl = api.Overlapper(api.default_params(min_olap_len=500))
    for max_evalue in (4095, 1000):
        got, counts, ms = ovl.ovlstore_sort(rec, length, skip_dup, library, max_evalue)
        want = mirror_filter_sort(rec, length, skip_dup, library, max_evalue)
        assert len(got) == len(want) > n // 2
        assert np.array_equal(got, want)
        assert counts[0] == 0 and (counts[1] > 0) == (max_evalue < 1499)
    # an ID out of range is fatal, as in the reference (ST/ovStoreBuild.C:594-603)
    bad = rec[:10].copy()
    bad[3, 1] = n_reads + 5
    with pytest.raises(api.OicError):
        ovl.ovlstore_sort(bad, length, skip_dup, library)
    ovl.close()
