"""Worker of test_zy_env_switches.py: runs one seeded case through the C ABI in THIS process (some switches are read once per
process) and prints one JSON line: md5 of the sorted records + the counters that must not depend on any switch."""
import hashlib
import json
import sys

import cases
from test_gpu_parity import STAT_KEYS, formatted, make_overlapper

# (n_found / n_refs / n_windows count what the seed scan looked at: OIC_FILTER=0 changes how a miss is detected, not what is found)
KEYS = STAT_KEYS + ["n_inserts", "n_distinct", "n_found", "n_refs", "n_add_ref", "ext_calls", "ext_cells", "ext_words", "last_end_hash_id"]

name = sys.argv[1]
reads, P, erate, threads, skip = cases.build_case(name)
ov = make_overlapper(reads, P, skip)
ov.overlap_driver()
rec, st, tm = formatted(ov.overlaps()), ov.stats(), ov.timings()
ov.close()
print(json.dumps({"n": len(rec), "md5": hashlib.md5("\n".join(rec).encode()).hexdigest(), "stats": {k: st[k] for k in KEYS},
                  "ms_ext_tasks": tm["ms_ext_tasks"], "ms_ext_retry": tm["ms_ext_retry"]}))
