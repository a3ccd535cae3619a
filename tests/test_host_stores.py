"""The data formats either side of the path (SURVEY.md 8(b) "next" rows), on the CPU:
  * canu_b200/host GkStore reads a gkpStore written by the reference's gatekeeperCreate and returns the generated reads;
  * canu_b200/host OvbWriter writes an .ovb + .counts that the reference's overlapConvert / golden md5 accept.
The GPU end-to-end test of the executable is in test_gpu_parity.py.
"""
import ctypes as C
import gzip
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

import cases
import helpers as H

HERE = os.path.dirname(os.path.abspath(__file__))
HOST = os.path.join(H.ROOT, "canu_b200", "host")


def stores_lib():
    src = [os.path.join(HERE, "stores_host.cpp"), os.path.join(HOST, "stores.cpp"), os.path.join(HOST, "ovstore.cpp")]
    dep = src + [os.path.join(HOST, "stores.h"), os.path.join(HOST, "ovstore.h"), os.path.join(HOST, "snappy_lite.h")]
    so = os.path.join(HERE, "libstores_host.so")
    if (not os.path.exists(so)) or max(os.path.getmtime(s) for s in dep) > os.path.getmtime(so):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC"] + src + ["-o", so])
    lib = C.CDLL(so)
    lib.sh_store_info.restype = C.c_uint32
    lib.sh_store_info.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_uint32]
    lib.sh_store_bases.argtypes = [C.c_char_p, C.c_void_p]
    lib.sh_write_ovb.argtypes = [C.c_char_p, C.c_void_p, C.c_uint64, C.c_uint32]
    lib.sh_counts_name.argtypes = [C.c_char_p, C.c_char_p, C.c_uint32]
    lib.sh_snappy_compress.restype = C.c_uint64
    lib.sh_snappy_compress.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64]
    lib.sh_snappy_uncompress.restype = C.c_int64
    lib.sh_snappy_uncompress.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64]
    lib.sh_read_ovb.restype = C.c_uint64
    lib.sh_read_ovb.argtypes = [C.c_char_p, C.c_void_p, C.c_uint64]
    lib.sh_write_ovlstore.argtypes = [C.c_char_p, C.c_char_p, C.c_void_p, C.c_uint64]
    lib.sh_store_read_info.restype = C.c_uint32
    lib.sh_store_read_info.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p]
    return lib


def pack_records(ov):
    """inverse of canu_b200.api.decode_records: the 6 file words of ovFile::writeOverlap (ST/ovStoreFile.C:196-221)"""
    u = lambda n: ov[n].astype(np.uint64)  # noqa: E731
    d0 = u("ahg5") | (u("ahg3") << np.uint64(21)) | (u("evalue") << np.uint64(42)) | (u("flipped") << np.uint64(54)) | \
        (u("for_obt") << np.uint64(55)) | (u("for_dup") << np.uint64(56)) | (u("for_utg") << np.uint64(57))
    d1 = u("bhg5") | (u("bhg3") << np.uint64(21)) | (u("span") << np.uint64(42))
    rec = np.zeros((len(ov), 6), dtype=np.uint32)
    rec[:, 0], rec[:, 1] = ov["a_iid"], ov["b_iid"]
    rec[:, 2], rec[:, 3] = (d0 >> np.uint64(32)).astype(np.uint32), (d0 & np.uint64(0xffffffff)).astype(np.uint32)
    rec[:, 4], rec[:, 5] = (d1 >> np.uint64(32)).astype(np.uint32), (d1 & np.uint64(0xffffffff)).astype(np.uint32)
    return rec


def test_counts_name():
    lib = stores_lib()
    buf = C.create_string_buffer(512)
    for name, want in [("/a/b.c/001/000001.ovb", "/a/b.c/001/000001.counts"), ("x.WORKING.ovb", "x.counts"),
                       ("plain", "plain.counts"), ("d.1/f", "d.1/f.counts")]:
        lib.sh_counts_name(name.encode(), buf, 512)
        assert buf.value.decode() == want


@pytest.mark.skipif(not H.have_ref(), reason="reference tools not built (oracle/build_ref.sh)")
def test_gkstore_reader(tmp_path):
    lib = stores_lib()
    reads = H.gen_reads(genome_len=20000, coverage=6, len_min=37, len_max=2100, err=0.02, seed=99)
    store = H.make_store(reads, str(tmp_path))
    n = lib.sh_store_info(store.encode(), None, None, 0)
    assert n == reads.n
    lens = np.zeros(n, dtype=np.uint32)
    libs = np.zeros(n, dtype=np.uint32)
    lib.sh_store_info(store.encode(), lens.ctypes.data, libs.ctypes.data, n)
    assert (lens == reads.lengths).all()
    assert (libs == 1).all()
    out = np.zeros(int(lens.sum()) + 1, dtype=np.uint8)
    lib.sh_store_bases(store.encode(), out.ctypes.data)
    assert (out[:-1] == reads.bases).all()
    # the store's own 2SEQ chunks (what oic_load_reads_2bit takes) == the harness encoder used by the tests and bench.py
    # (tools/gen_reads.c oic_encode_2bit) == the numpy encoder of canu_b200.api: pins the 2-bit input format to the reference's
    # gatekeeperCreate (ST/gkStoreEncode.C:45-78)
    from canu_b200 import api
    packed, boff = H.encode_2bit(reads.bases, reads.offsets, reads.lengths)
    packed2, boff2 = api.encode_2bit(reads.bases, reads.offsets, reads.lengths)
    assert np.array_equal(packed, packed2) and np.array_equal(boff, boff2)
    raw = np.zeros(len(packed) + 8, dtype=np.uint8)
    lib.sh_store_2bit.restype = C.c_int
    lib.sh_store_2bit.argtypes = [C.c_char_p, C.c_void_p]
    assert lib.sh_store_2bit(store.encode(), raw.ctypes.data) == 1
    assert np.array_equal(raw[:len(packed)], packed)


@pytest.mark.skipif(not H.have_ref(), reason="reference tools not built (oracle/build_ref.sh)")
@pytest.mark.parametrize("name,pieces", [("c1_small", 1), ("c4_short", 7)])
def test_ovb_writer_accepted_by_reference(name, pieces, tmp_path):
    """oracle overlaps -> OvbWriter -> reference overlapConvert == golden dump; .counts md5 == golden"""
    lib = stores_lib()
    reads, P, erate, threads, skip = cases.build_case(name)
    ov = H.run_oracle(reads, P, skip_kmers=skip)["overlaps"]
    rec = pack_records(ov)
    store = H.make_store(reads, str(tmp_path))
    ovb = str(tmp_path / "mine.ovb")
    lib.sh_write_ovb(ovb.encode(), rec.ctypes.data, len(rec), pieces)
    got = H.dump_ovb_sorted(store, ovb)
    want = gzip.open(os.path.join(HERE, "golden", name + ".txt.gz")).read().decode().splitlines()
    assert got == want
    meta = json.load(open(os.path.join(HERE, "golden", name + ".json")))
    assert hashlib.md5(open(str(tmp_path / "mine.counts"), "rb").read()).hexdigest() == meta["counts_md5"]


@pytest.mark.skipif(not H.have_ref(), reason="reference tools not built (oracle/build_ref.sh)")
def test_ovb_writer_block_boundary(tmp_path):
    """more records than one 262080-word block holds (43680 records): every block must decode"""
    lib = stores_lib()
    reads = H.gen_reads(genome_len=3000, coverage=3, len_min=500, len_max=900, err=0.0, seed=3)
    store = H.make_store(reads, str(tmp_path))
    n = 100003
    rng = np.random.default_rng(5)
    ov = np.zeros(n, dtype=H.OVERLAP_DT)
    ov["a_iid"] = rng.integers(1, reads.n + 1, n)
    ov["b_iid"] = rng.integers(1, reads.n + 1, n)
    for f in ("ahg5", "ahg3", "bhg5", "bhg3"):
        ov[f] = rng.integers(0, 400, n)
    ov["span"] = rng.integers(0, 1 << 21, n)
    ov["evalue"] = rng.integers(0, 4096, n)
    ov["flipped"] = rng.integers(0, 2, n)
    ov["for_utg"] = 1
    rec = pack_records(ov)
    ovb = str(tmp_path / "big.ovb")
    lib.sh_write_ovb(ovb.encode(), rec.ctypes.data, n, 3)
    out = subprocess.check_output([os.path.join(H.REF_DIR, "overlapConvert"), "-G", store, "-raw", ovb], stderr=subprocess.DEVNULL)
    got = out.decode().splitlines()
    assert len(got) == n
    want = [H.format_raw(o) for o in ov[:50]] + [H.format_raw(o) for o in ov[-50:]]
    assert got[:50] + got[-50:] == want
    counts = np.fromfile(str(tmp_path / "big.counts"), dtype=np.uint32)
    assert counts[0] == len(counts) - 1
    assert counts[1:].sum() == 2 * n


def test_snappy_round_trip():
    """the raw-format compressor / decompressor of canu_b200/host/snappy_lite.h: records (repetitive), noise, runs, tiny inputs"""
    lib = stores_lib()
    rng = np.random.default_rng(11)
    rec = np.zeros((5000, 6), dtype=np.uint32)
    rec[:, 0] = np.sort(rng.integers(1, 300, 5000))
    rec[:, 1] = rng.integers(1, 300, 5000)
    rec[:, 2:] = rng.integers(0, 1 << 20, (5000, 4))
    for data in (rec.tobytes(), rng.integers(0, 256, 70001, dtype=np.uint8).tobytes(), b"a" * 100000, b"abcdefgh" * 9000 + b"xyz",
                 b"", b"q", b"0123456789abcde"):
        src = np.frombuffer(data, dtype=np.uint8)
        out = np.zeros(len(src) + len(src) // 6 + 64, dtype=np.uint8)
        n = lib.sh_snappy_compress(src.ctypes.data if len(src) else None, len(src), out.ctypes.data, len(out))
        assert 0 < n <= len(out)
        back = np.zeros(len(src) + 8, dtype=np.uint8)
        m = lib.sh_snappy_uncompress(out.ctypes.data, n, back.ctypes.data, len(back))
        assert m == len(src) and back[:m].tobytes() == data
    # records compress (the literal-only writer of round 1 made files 1.25-2x the reference's)
    src = np.frombuffer(rec.tobytes(), dtype=np.uint8)
    out = np.zeros(len(src) * 2, dtype=np.uint8)
    assert lib.sh_snappy_compress(src.ctypes.data, len(src), out.ctypes.data, len(out)) < 0.9 * len(src)
    # a corrupt stream is refused
    out[5] ^= 0xff
    assert lib.sh_snappy_uncompress(out.ctypes.data, 200, np.zeros(len(src) + 8, dtype=np.uint8).ctypes.data, len(src) + 8) == -1


def mirror_filter_sort(rec, length, skip_dup, library, max_evalue=4095):
    """numpy restatement (TEST ORACLE) of ovStoreBuild's bucketize + sort: ovStoreFilter::filterOverlap (ST/ovStoreFilter.C:118-260),
    ovOverlap::swapIDs (ST/ovOverlap.C:114-144), ovOverlap::operator< (ST/ovOverlap.H:327-339)"""
    rec = np.asarray(rec, dtype=np.uint32).reshape(-1, 6)
    a, b = rec[:, 0].astype(np.int64), rec[:, 1].astype(np.int64)
    d0 = (rec[:, 2].astype(np.uint64) << np.uint64(32)) | rec[:, 3].astype(np.uint64)
    d1 = (rec[:, 4].astype(np.uint64) << np.uint64(32)) | rec[:, 5].astype(np.uint64)
    m21, u = np.uint64((1 << 21) - 1), np.uint64
    ahg5, ahg3, bhg5, bhg3 = d0 & m21, (d0 >> u(21)) & m21, d1 & m21, (d1 >> u(21)) & m21
    flipped = ((d0 >> u(54)) & u(1)).astype(bool)
    rest0, rest1 = d0 & ~u((1 << 42) - 1), d1 & ~u((1 << 42) - 1)
    r0 = np.where(flipped, rest0 | bhg3 | (bhg5 << u(21)), rest0 | bhg5 | (bhg3 << u(21)))
    r1 = np.where(flipped, rest1 | ahg3 | (ahg5 << u(21)), rest1 | ahg5 | (ahg3 << u(21)))
    f0 = d0.copy()
    flags = u(7 << 55)
    bad = ((d0 >> u(42)) & u(4095)) > u(max_evalue)
    f0[bad] &= ~flags
    r0[bad] &= ~flags
    alen, blen = length[a].astype(np.int64), length[b].astype(np.int64)
    ab, ae = ahg5.astype(np.int64), alen - ahg3.astype(np.int64)
    bb = np.where(flipped, blen - bhg5.astype(np.int64), bhg5.astype(np.int64))
    be = np.where(flipped, bhg3.astype(np.int64), blen - bhg3.astype(np.int64))
    is_diff = flipped | (np.abs(ab - bb) > 29)
    is_long = ((ae - ab) > 75) & (np.abs(be - bb) > 75)
    obt = u(1 << 55)
    dup = u(1 << 56)
    drop_obt = ~is_diff | ~is_long
    f0[drop_obt] &= ~obt
    r0[drop_obt] &= ~obt
    sd = skip_dup[a].astype(bool)
    f0[sd] &= ~dup
    r0[sd] &= ~dup
    cross = (((f0 & dup) != 0) | ((r0 & dup) != 0)) & (library[a] != library[b])
    f0[cross] &= ~dup
    r0[cross] &= ~dup
    keep_f, keep_r = (f0 & flags) != 0, (r0 & flags) != 0
    A = np.concatenate([a[keep_f], b[keep_r]])
    B = np.concatenate([b[keep_f], a[keep_r]])
    D0 = np.concatenate([f0[keep_f], r0[keep_r]])
    D1 = np.concatenate([d1[keep_f], r1[keep_r]])
    order = np.lexsort([D1, D0, B, A])
    out = np.zeros((len(A), 6), dtype=np.uint32)
    out[:, 0], out[:, 1] = A[order], B[order]
    out[:, 2], out[:, 3] = (D0[order] >> u(32)).astype(np.uint32), (D0[order] & u(0xffffffff)).astype(np.uint32)
    out[:, 4], out[:, 5] = (D1[order] >> u(32)).astype(np.uint32), (D1[order] & u(0xffffffff)).astype(np.uint32)
    return out


STORE_FILES = ["info", "index", "0001", "evalueLen", "overlapScores"]


def reference_store(store, ovb, wd, extra=()):
    out = os.path.join(wd, "ref.ovlStore")
    subprocess.check_call([os.path.join(H.REF_DIR, "ovStoreBuild"), "-O", out, "-G", store] + list(extra) + [ovb],
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return out


@pytest.mark.skipif(not (H.have_ref() and os.access(os.path.join(H.REF_DIR, "ovStoreBuild"), os.X_OK)), reason="reference ovStoreBuild not built")
@pytest.mark.parametrize("name", ["c1_small", "c4_short", "quirk_m"])
def test_ovlstore_writer_vs_reference_ovStoreBuild(name, tmp_path):
    """The host half of the store construction on the CPU: the reference's own .ovb is read back by
    read_ovb_file, filtered / mirrored / sorted by the numpy restatement above, written by OvlStoreWriter -- every file of the
    store must equal, byte for byte, the store the reference's ovStoreBuild makes from the same .ovb."""
    lib = stores_lib()
    wd = str(tmp_path)
    reads, P, erate, threads, skip = cases.build_case(name)
    store = H.make_store(reads, wd)
    args = H.oic_args(P, erate)
    if skip:
        sf = os.path.join(wd, "skip.fasta")
        cases.write_skip_file(skip, sf)
        args += ["-k", sf]
    ovb, _ = H.run_overlapper(os.path.join(H.REF_DIR, "overlapInCore"), store, os.path.join(wd, "ref"), args, threads)
    n = lib.sh_read_ovb(ovb.encode(), None, 0)
    assert n > 100
    rec = np.zeros((n, 6), dtype=np.uint32)
    assert lib.sh_read_ovb(ovb.encode(), rec.ctypes.data, n) == n
    max_id = lib.sh_store_read_info(store.encode(), None, None, None)
    assert max_id == reads.n + 1
    length, skip_dup, library = np.zeros(max_id, dtype=np.uint32), np.zeros(max_id, dtype=np.uint8), np.zeros(max_id, dtype=np.uint8)
    lib.sh_store_read_info(store.encode(), length.ctypes.data, skip_dup.ctypes.data, library.ctypes.data)
    assert (length[1:] == reads.lengths).all()
    srt = np.ascontiguousarray(mirror_filter_sort(rec, length, skip_dup, library))
    mine = os.path.join(wd, "mine.ovlStore")
    lib.sh_write_ovlstore(mine.encode(), store.encode(), srt.ctypes.data, len(srt))
    ref = reference_store(store, ovb, wd)
    assert sorted(os.listdir(ref)) == sorted(os.listdir(mine)) == sorted(STORE_FILES)
    for f in STORE_FILES:
        assert open(os.path.join(mine, f), "rb").read() == open(os.path.join(ref, f), "rb").read(), f
