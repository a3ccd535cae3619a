"""Golden fixtures of the frequent-k-mer row, produced by RUNNING THE REFERENCE's meryl
(oracle/_ref/meryl, built by oracle/build_ref.sh) the way the pipeline does (Meryl.pm:434-450, 634):

    python tests/golden/make_golden_kmers.py

For each (case, K, T) of cases.KMER_CASES: kmers_<case>_<K>_<T>.txt.gz = the bytes of `meryl -Dt -n T -s <db>`;
khist_<case>_<K>.txt.gz = stdout of `meryl -Dh -s <db>`, a line `--info--`, then its stderr.
"""
import gzip
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers as H  # noqa: E402
import cases  # noqa: E402


def main():
    assert H.have_ref(), "run oracle/build_ref.sh first"
    for name, K, T in cases.KMER_CASES:
        reads = cases.build_case(name)[0]
        with tempfile.TemporaryDirectory() as wd:
            store = H.make_store(reads, wd)
            text = H.run_meryl(store, K, T, wd)
            hout, herr = H.run_meryl_hist(store, K, wd)
        with gzip.GzipFile(os.path.join(HERE, "khist_%s_%d.txt.gz" % (name, K)), "wb", mtime=0) as f:
            f.write((hout + "--info--\n" + herr).encode())
        with gzip.GzipFile(os.path.join(HERE, "kmers_%s_%d_%d.txt.gz" % (name, K, T)), "wb", mtime=0) as f:
            f.write(text.encode())
        print(name, K, T, text.count(">"))


if __name__ == "__main__":
    main()
