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
    src = [os.path.join(HERE, "stores_host.cpp"), os.path.join(HOST, "stores.cpp")]
    dep = src + [os.path.join(HOST, "stores.h")]
    so = os.path.join(HERE, "libstores_host.so")
    if (not os.path.exists(so)) or max(os.path.getmtime(s) for s in dep) > os.path.getmtime(so):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC"] + src + ["-o", so])
    lib = C.CDLL(so)
    lib.sh_store_info.restype = C.c_uint32
    lib.sh_store_info.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_uint32]
    lib.sh_store_bases.argtypes = [C.c_char_p, C.c_void_p]
    lib.sh_write_ovb.argtypes = [C.c_char_p, C.c_void_p, C.c_uint64, C.c_uint32]
    lib.sh_counts_name.argtypes = [C.c_char_p, C.c_char_p, C.c_uint32]
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
