"""oic_ovlstore_sort_noalloc: the allocation-free variant of the overlap-store sort (include/oic_b200.h).

it has no counterpart in the reference.  Callers that may not call cudaMalloc
(code under CUDA graph capture, tight serving loops, harnesses that forbid it) get the same filter + mirror + sort as
oic_ovlstore_sort with every intermediate buffer in a static __device__ array of the library.  Inputs and outputs are DEVICE
pointers the caller owns (torch tensors here: test plumbing).

Checked: the output of the new entry point against the existing one's on the same records -- bit-exact, the tolerance of every
integer test in this suite -- on TWO calls in a row with different inputs (the scratch keeps the last call's contents: nothing of
call 1 may leak into call 2), plus the size limit and the error path.

This file sorts last on purpose (pytest collects files in name order and the round's end runs `pytest -x`): a failure here hides no
other test.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def make_records(seed, n, n_reads):
    """synthetic overlaps that take every branch of the store filter (see test_ovlstore_sort_vs_numpy_restatement)"""
    from test_host_stores import pack_records
    rng = np.random.default_rng(seed)
    length = np.zeros(n_reads + 1, dtype=np.uint32)
    length[1:] = rng.integers(150, 4000, n_reads)
    library = np.zeros(n_reads + 1, dtype=np.uint8)
    library[1:] = rng.integers(1, 3, n_reads)
    skip_dup = np.zeros(n_reads + 1, dtype=np.uint8)
    skip_dup[library == 2] = 1
    ov = np.zeros(n, dtype=H.OVERLAP_DT)
    ov["a_iid"] = rng.integers(1, n_reads + 1, n)
    ov["b_iid"] = np.clip(rng.integers(1, 40, n) + (ov["a_iid"] % 7), 1, n_reads)   # many equal (a, b) pairs
    la, lb = length[ov["a_iid"]], length[ov["b_iid"]]
    ov["ahg5"] = rng.integers(0, 120, n) % la
    ov["ahg3"] = rng.integers(0, 60, n) % (la - ov["ahg5"])
    ov["bhg5"] = rng.integers(0, 120, n) % lb
    ov["bhg3"] = rng.integers(0, 60, n) % (lb - ov["bhg5"])
    ov["span"] = rng.integers(0, 4000, n)
    ov["evalue"] = rng.integers(0, 1500, n)
    for f in ("flipped", "for_obt", "for_dup", "for_utg"):
        ov[f] = rng.integers(0, 2, n)
    rec = pack_records(ov)
    rec[-200:] = rec[:200]   # exact duplicates: their order is decided by nothing, their bytes are equal
    return np.ascontiguousarray(rec), length, skip_dup, library


def noalloc_sort(lib, rec, length, skip_dup, library, max_evalue):
    """call the new entry point through the C ABI with caller-owned device buffers"""
    import torch
    dev = torch.device("cuda", 0)
    d_rec = torch.from_numpy(rec.view(np.int32)).to(dev)
    d_len = torch.from_numpy(length.view(np.int32)).to(dev)
    d_skip = torch.from_numpy(skip_dup).to(dev)
    d_lib = torch.from_numpy(library).to(dev)
    d_out = torch.full((2 * len(rec), 6), -1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    m = C.c_uint64(0)
    counts = (C.c_uint64 * 9)()
    rc = lib.oic_ovlstore_sort_noalloc(0, d_rec.data_ptr(), len(rec), len(length), d_len.data_ptr(), d_skip.data_ptr(), d_lib.data_ptr(),
                                       max_evalue, d_out.data_ptr(), C.byref(m), counts)
    return rc, d_out[:m.value].cpu().numpy().view(np.uint32), np.array(list(counts), dtype=np.uint64)


def test_noalloc_sort_equals_allocating_sort_twice_in_a_row():
    from canu_b200 import api
    lib = api.load_library()
    limit = lib.oic_ovlstore_sort_noalloc_max()
    assert 1024 <= limit <= (1 << 20), "a stated, modest maximum"
    ovl = api.Overlapper(api.default_params(min_olap_len=500))
    # call 1: a large job (more than half the limit, both halves of the scratch in use); call 2: a smaller, different one right behind it
    for seed, n, n_reads, max_evalue in ((31, min(limit, 60000), 3000, 4095), (32, 7001, 900, 1000)):
        rec, length, skip_dup, library = make_records(seed, n, n_reads)
        want, want_counts, _ = ovl.ovlstore_sort(rec, length, skip_dup, library, max_evalue)
        rc, got, counts = noalloc_sort(lib, rec, length, skip_dup, library, max_evalue)
        assert rc == 0, lib.oic_last_error(None).decode()
        assert len(got) == len(want) > n // 2
        assert np.array_equal(got, want)              # bit-exact: integer / byte work
        assert np.array_equal(counts, want_counts)
    ovl.close()


def test_noalloc_sort_rejects_sizes_beyond_the_maximum_and_bad_ids():
    import torch
    from canu_b200 import api
    lib = api.load_library()
    limit = lib.oic_ovlstore_sort_noalloc_max()
    rec, length, skip_dup, library = make_records(33, 64, 50)
    dev = torch.device("cuda", 0)
    d = torch.zeros(4096, dtype=torch.int32, device=dev)
    m = C.c_uint64(7)
    # one record too many: refused before anything runs, with a message that names the limit
    rc = lib.oic_ovlstore_sort_noalloc(0, d.data_ptr(), limit + 1, len(length), d.data_ptr(), d.data_ptr(), d.data_ptr(), 4095, d.data_ptr(), C.byref(m), None)
    assert rc == api.OIC_ERR_CAPACITY and str(limit).encode() in lib.oic_last_error(None) and m.value == 0
    rc = lib.oic_ovlstore_sort_noalloc(0, None, 5, len(length), d.data_ptr(), d.data_ptr(), d.data_ptr(), 4095, d.data_ptr(), C.byref(m), None)
    assert rc == api.OIC_ERR_ARG
    # an ID out of range is fatal, as in the allocating entry point (ST/ovStoreBuild.C:594-603) ...
    bad = rec.copy()
    bad[3, 1] = 51 + 5
    rc, _, _ = noalloc_sort(lib, bad, length, skip_dup, library, 4095)
    assert rc == api.OIC_ERR_ARG
    # ... and leaves the scratch usable: the next call is right again
    ovl = api.Overlapper(api.default_params(min_olap_len=500))
    want, _, _ = ovl.ovlstore_sort(rec, length, skip_dup, library, 4095)
    ovl.close()
    rc, got, _ = noalloc_sort(lib, rec, length, skip_dup, library, 4095)
    assert rc == 0 and np.array_equal(got, want)
    # zero records: nothing to do, nothing written
    rc = lib.oic_ovlstore_sort_noalloc(0, d.data_ptr(), 0, len(length), d.data_ptr(), d.data_ptr(), d.data_ptr(), 4095, d.data_ptr(), C.byref(m), None)
    assert rc == 0 and m.value == 0
