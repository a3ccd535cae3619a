"""Random -w jobs on the GPU: the CUDA path against the CPU oracle on the 24 seeded jobs of cases.fuzz_case_w (several libraries with
their own default quality, -m, -l, N bases, hash sub-blocks) -- records, the -s counters incl. the two window rejections, and the
extension work counters.  The oracle is held to the live reference binary on the same jobs by tests/test_oracle.py.
(Collected late on purpose, like test_zz_no_alloc.py.)"""
import pytest

import cases
import helpers as H
from test_gpu_parity import STAT_KEYS, formatted, make_overlapper

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("lo", [0, 12])
def test_window_fuzz_vs_oracle(lo):
    for i in range(lo, min(lo + 12, cases.N_FUZZ_W)):
        reads, P, erate, threads, _ = cases.fuzz_case_w(i)
        ov = make_overlapper(reads, P)
        ov.overlap_driver()
        mine, st = formatted(ov.overlaps()), ov.stats()
        ov.close()
        res = H.run_oracle(reads, P)
        assert mine == sorted(H.format_raw(o) for o in res["overlaps"]), "job %d" % i
        for k in STAT_KEYS + ["ext_calls", "ext_cells", "ext_words"]:
            assert st[k] == res["stats"][k], "job %d: %s" % (i, k)


@pytest.mark.parametrize("extra", [["-H", "2-3", "-R", "1-2"], ["-H", "3-4", "-R", "1"], ["-R", "2"]])
def test_executable_library_filters_vs_live_reference(extra, tmp_path):
    """-H / -R (hash-side and query-side library ranges, OIC/overlapInCore-Build_Hash_Index.C:555-557, -Process_Overlaps.C:108-110)
    through the drop-in executable on the four-library store of the `window_libs` case, against the reference binary run on the spot."""
    import os
    from test_gpu_executable import BIN
    if not (H.have_ref() and os.access(BIN, os.X_OK)):
        pytest.skip("oracle/_ref tools or canu_b200/bin/overlapInCore not built")
    reads, P, erate, threads, skip = cases.build_case("window_libs")
    wd = str(tmp_path)
    store = H.make_store(reads, wd)
    args = H.oic_args(P, erate) + extra
    got, want = [], []
    for name, binary, out in (("mine", BIN, got), ("ref", os.path.join(H.REF_DIR, "overlapInCore"), want)):
        ovb, st = H.run_overlapper(binary, store, os.path.join(wd, name), args, threads)
        out.append(H.dump_ovb_sorted(store, ovb))
        out.append(H.parse_stats(st))
    assert got[0] == want[0] and len(want[0]) > 20
    assert got[1] == want[1]
    assert open(os.path.join(wd, "mine.counts"), "rb").read() == open(os.path.join(wd, "ref.counts"), "rb").read()
