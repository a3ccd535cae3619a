"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/oic_b200.h
declares, and refuses to run without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

import helpers as H


def declared_symbols():
    hdr = open(os.path.join(H.ROOT, "include", "oic_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(oic_[a-z_0-9]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    from canu_b200 import api
    lib = api.load_library()
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), s
    assert sorted(api.ABI_SYMBOLS) == syms


def test_last_processed_ref_quirk():
    """Scheduler quirk (OIC/overlapInCore.C:249, Process_Overlaps.C:86): pure host arithmetic, no device needed.
    Verified empirically in the survey: `-r 1-9 -t 1` never processes read 9."""
    from canu_b200 import api
    lib = api.load_library()
    assert lib.oic_last_processed_ref(1, 9, 1) == 8        # perThread = 2, (9-1) % 2 == 0
    assert lib.oic_last_processed_ref(1, 10, 1) == 10      # perThread = 2, 9 % 2 == 1
    assert lib.oic_last_processed_ref(1, 25, 1) == 24      # perThread = 4, 24 % 4 == 0
    assert lib.oic_last_processed_ref(1, 161, 2) == 161    # perThread = 11, 160 % 11 != 0
    assert lib.oic_last_processed_ref(5, 4, 1) == 0


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from canu_b200 import api
    with pytest.raises(api.OicError):
        api.Overlapper(api.default_params(min_olap_len=500))


def test_unsupported_options_fail_loudly():
    from canu_b200 import api
    with pytest.raises(api.OicError):
        api.Overlapper(api.default_params(kmer_len=32))


def test_tools_fail_without_device(tmp_path):
    """the executables have no CPU path either: without a CUDA device they exit 1 with a message, they do not compute"""
    import subprocess
    import torch
    import cases
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    if not H.have_ref():
        pytest.skip("reference tools not built (no gatekeeperCreate to make a store)")
    bindir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "canu_b200", "bin")
    reads = cases.build_case("ragged_tiny")[0]
    store = H.make_store(reads, str(tmp_path))
    r = subprocess.run([os.path.join(bindir, "frequentMers"), "-m", "22", "-n", "3", "-s", store, "-o", str(tmp_path / "f.fasta")], capture_output=True)
    assert r.returncode == 1 and b"no CUDA device" in r.stderr and not os.path.exists(tmp_path / "f.fasta")
    r = subprocess.run([os.path.join(bindir, "overlapInCore"), "-k", "22", "--hashstrings", "100", "-o", str(tmp_path / "x.ovb"), store], capture_output=True)
    assert r.returncode == 1 and b"no CUDA device" in r.stderr
    # argument errors are caught before anything else (OIC/overlapInCore.C:423-483)
    r = subprocess.run([os.path.join(bindir, "frequentMers"), "-m", "40", "-n", "3", "-s", store], capture_output=True)
    assert r.returncode == 1 and b"usage" in r.stderr
