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
