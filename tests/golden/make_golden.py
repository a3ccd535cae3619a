"""Generate the committed golden fixtures by RUNNING THE REFERENCE BINARY (oracle/_ref/overlapInCore,
built by oracle/build_ref.sh from /root/reference/src) on the seeded cases of tests/cases.py.

    python tests/golden/make_golden.py [case ...]

For each case: <name>.txt.gz = `overlapConvert -G store -raw x.ovb | LC_ALL=C sort` (every stored field
of every overlap record), <name>.json = the 8 counters of the -s stats file + the .counts file md5.
The reference ships no fixtures of its own for this path; these are outputs of
the reference itself, which is what pins the oracle.
"""
import gzip
import hashlib
import json
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers as H  # noqa: E402
import cases  # noqa: E402


def main():
    assert H.have_ref(), "run oracle/build_ref.sh first"
    for name in (sys.argv[1:] or cases.CASES):
        reads, P, erate, threads, skip = cases.build_case(name)
        with tempfile.TemporaryDirectory() as wd:
            store = H.make_store(reads, wd)
            args = H.oic_args(P, erate)
            if skip:
                sf = os.path.join(wd, "skip.fasta")
                cases.write_skip_file(skip, sf)
                args += ["-k", sf]
            ovb, st = H.run_overlapper(os.path.join(H.REF_DIR, "overlapInCore"), store, os.path.join(wd, "ref"), args, threads)
            lines = H.dump_ovb_sorted(store, ovb)
            stats = H.parse_stats(st)
            counts = open(os.path.join(wd, "ref.counts"), "rb").read()
        with gzip.GzipFile(os.path.join(HERE, name + ".txt.gz"), "wb", mtime=0) as f:
            f.write(("".join(x + "\n" for x in lines)).encode())
        meta = dict(stats=stats, n_overlaps=len(lines), n_reads=reads.n, total_bases=int(reads.lengths.sum()),
                    counts_md5=hashlib.md5(counts).hexdigest(), threads=threads, erate=erate)
        json.dump(meta, open(os.path.join(HERE, name + ".json"), "w"), indent=1, sort_keys=True)
        print(name, len(lines), stats)


if __name__ == "__main__":
    main()
