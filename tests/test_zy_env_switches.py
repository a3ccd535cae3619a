"""The library's environment switches select between code paths that must give the same records and the same
counters: each switch is set for a fresh process (several are read once per process) that runs a seeded case through the C ABI, and
its record digest and counters are compared with the run without any switch -- which test_gpu_parity.py holds to the goldens.
(Collected late on purpose, like test_zz_no_alloc.py.)"""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
SWITCHES = [
    {"OIC_NO_SPECULATION": "1"},                             # whole serial loop per pair in k_process_pairs, no task kernel
    {"OIC_NO_FIRST_LEAN": "1"},                              # first extensions through the generic row loop
    {"OIC_LEAN_WIDE": "1"},                                  # the wide task-kernel instance
    {"OIC_FILTER": "0"},                                     # seed scan without the Bloom filter
    {"OIC_TABLE_FACTOR": "3", "OIC_FILTER_BITS": "8"},       # another table / filter geometry
    {"OIC_REC_CAP": "65536", "OIC_BATCH_BASES": "20000"},    # tiny seed-run buffer, many scan batches (overflow -> smaller batch)
    {"OIC_MEMINFO_ALWAYS": "1", "OIC_PHASE_CLOCKS": "1"},    # diagnostics only
]


def run(case, env):
    e = dict(os.environ)
    e.update(env)
    e["PYTHONPATH"] = HERE + os.pathsep + os.path.dirname(HERE) + os.pathsep + e.get("PYTHONPATH", "")
    out = subprocess.run([sys.executable, os.path.join(HERE, "env_switch_worker.py"), case], env=e, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    return json.loads(out.stdout.strip().splitlines()[-1])


@pytest.mark.parametrize("case", ["c1_small", "n_bases"])
def test_switches_do_not_change_results(case):
    base = run(case, {})
    assert base["n"] > 0 and base["ms_ext_tasks"] > 0, "the default run must take the speculative task path"
    for env in SWITCHES:
        got = run(case, env)
        assert (got["n"], got["md5"]) == (base["n"], base["md5"]), env
        assert got["stats"] == base["stats"], env
        if "OIC_NO_SPECULATION" in env:
            assert got["ms_ext_tasks"] == 0, "the switch did not take the serial path"
