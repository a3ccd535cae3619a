"""CPU unit tests of the scalar logic shared by the kernels (canu_b200/csrc/oic_core.cuh compiled for the host).

The Add_Match replay (run granularity) is checked against the oracle's per-pair seed lists: the runs are the
oracle's final nodes re-sorted into arrival order (query offset ascending, hash offset descending); the replay
must reproduce the reference's list order and its `consistent` flag.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import cases
import helpers as H

HERE = os.path.dirname(os.path.abspath(__file__))


def core_lib():
    src = os.path.join(HERE, "core_host.cpp")
    so = os.path.join(HERE, "libcore_host.so")
    hdr = os.path.join(H.ROOT, "canu_b200", "csrc", "oic_core.cuh")
    if (not os.path.exists(so)) or max(os.path.getmtime(src), os.path.getmtime(hdr)) > os.path.getmtime(so):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-DOIC_HOST_ONLY", "-shared", "-fPIC", src, "-o", so])
    lib = C.CDLL(so)
    lib.core_replay.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    lib.core_pack_record.argtypes = [C.c_void_p, C.c_void_p]
    return lib


@pytest.mark.parametrize("name", ["c1_small", "c2_noisy", "repeats_skip", "c4_short"])
def test_replay_matches_oracle_lists(name):
    lib = core_lib()
    reads, P, erate, threads, skip = cases.build_case(name)
    res = H.run_oracle(reads, P, dump=1, skip_kmers=skip)
    pairs, matches = res["pairs"], res["matches"]
    assert len(pairs) > 0
    K = P["kmer_len"]
    multi = 0
    for p in pairs:
        n = int(p["n_matches"])
        m = matches[int(p["first_match"]):int(p["first_match"]) + n]
        order = np.lexsort((-m["offset"].astype(np.int64), m["start"]))
        runs = np.ascontiguousarray(np.stack([m["start"][order], m["offset"][order], m["len"][order]], axis=1), dtype=np.int32)
        out = np.zeros((n, 3), dtype=np.int32)
        info = np.zeros(4, dtype=np.int32)
        lib.core_replay(runs.ctypes.data, n, K, out.ctypes.data, info.ctypes.data)
        want = np.stack([m["offset"], m["len"], m["start"]], axis=1)
        assert (out == want).all(), (name, p, runs, out, want)
        assert info[0] == p["consistent"] and info[1] == p["diag_ct"] and info[2] == p["diag_bgn"] and info[3] == p["diag_end"], (name, p, info)
        multi += n > 1
    assert multi > 0


def test_pack_record_layout():
    """ovOverlapDAT bit layout for AS_MAX_READLEN_BITS = 21 (ST/ovOverlap.H:100-115) and the word order of
    ovFile::writeOverlap (ST/ovStoreFile.C:196-221), against the Python decoder used by the GPU tests."""
    from canu_b200 import api
    lib = core_lib()
    f = np.array([7, 9, 123, 0, 0, 456, 1999, 345, 1, 0, 0, 1], dtype=np.uint32)
    w = np.zeros(6, dtype=np.uint32)
    lib.core_pack_record(w.ctypes.data, f.ctypes.data)
    d = api.decode_records(w.reshape(1, 6))
    names = ["a_iid", "b_iid", "ahg5", "ahg3", "bhg5", "bhg3", "span", "evalue", "flipped", "for_obt", "for_dup", "for_utg"]
    for i, n in enumerate(names):
        assert int(d[n][0]) == int(f[i]), n
