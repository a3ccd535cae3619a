"""Test/bench harness helpers: seeded read generator, reference-tool wrappers, oracle bindings.

Nothing here is on the product path.  The oracle (oracle/liboic_oracle.so) and the reference
binaries (oracle/_ref/*) are only ever used as checkers / CPU baselines.
"""
import ctypes as C
import os
import subprocess
import struct

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
TOOLS_DIR = os.path.join(ROOT, "tools")


def _newer(src, dst):
    return (not os.path.exists(dst)) or os.path.getmtime(src) > os.path.getmtime(dst)


def build_harness_libs():
    """Compile the generator and the oracle restatement (gcc/g++, seconds)."""
    gen_src = os.path.join(TOOLS_DIR, "gen_reads.c")
    gen_so = os.path.join(TOOLS_DIR, "libgenreads.so")
    if _newer(gen_src, gen_so):
        subprocess.check_call(["gcc", "-O2", "-fopenmp", "-shared", "-fPIC", gen_src, "-o", gen_so])
    gen_bin = os.path.join(TOOLS_DIR, "gen_reads")
    if _newer(gen_src, gen_bin):
        subprocess.check_call(["gcc", "-O2", "-fopenmp", "-DGEN_MAIN", gen_src, "-o", gen_bin])
    o_src = os.path.join(ORACLE_DIR, "oic_oracle.cpp")
    o_so = os.path.join(ORACLE_DIR, "liboic_oracle.so")
    if _newer(o_src, o_so) or _newer(os.path.join(ORACLE_DIR, "oic_oracle.h"), o_so):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-shared", "-fPIC", o_src, "-o", o_so])
    return gen_so, o_so


def have_ref():
    return all(os.access(os.path.join(REF_DIR, t), os.X_OK)
               for t in ("overlapInCore", "gatekeeperCreate", "overlapConvert"))


# ----------------------------------------------------------------------------- generator

class GenParams(C.Structure):
    _fields_ = [("genome_len", C.c_uint64), ("coverage", C.c_double), ("len_min", C.c_uint32),
                ("len_max", C.c_uint32), ("err", C.c_double), ("sub_only", C.c_int), ("seed", C.c_uint64),
                ("repeat_frac", C.c_double), ("repeat_families", C.c_uint32), ("repeat_min", C.c_uint32),
                ("repeat_max", C.c_uint32), ("repeat_div", C.c_double)]


class Reads:
    """Reads in memory: ASCII bases concatenated (uint8), offsets[n+1], lengths[n]. IDs are 1-based."""

    def __init__(self, bases, offsets, lengths, read_qv=None):
        self.bases, self.offsets, self.lengths = bases, offsets, lengths
        self.n = len(lengths)
        # optional: one quality value per READ.  The reference's gatekeeperCreate is built with DO_NOT_STORE_QVs (ST/gatekeeperCreate.C:
        # 49-54): it stores no per-base qualities, every read gets a QVAL chunk = its library's default QV (`qv N` in the .gkp file,
        # :663-664; 20 unless set).  make_store writes one library per run of equal values.
        self.read_qv = read_qv

    def qualities(self, default_qv=20):
        """what gkReadData_getQualities gives the overlapper (ST/gkStore.C:140-145): the read's QVAL constant for every base"""
        q = np.full(int(self.offsets[-1]), default_qv, dtype=np.uint8)
        if self.read_qv is not None:
            q[:] = np.repeat(np.asarray(self.read_qv, dtype=np.uint8), np.asarray(self.lengths, dtype=np.int64))
        return q

    def seq(self, rid):
        o = int(self.offsets[rid - 1])
        return self.bases[o:o + int(self.lengths[rid - 1])].tobytes().decode()

    def write_fasta(self, path, first=0, last=None):
        with open(path, "wb") as f:
            for i in range(first, self.n if last is None else last):
                f.write(b">read%d\n" % (i + 1))
                o = int(self.offsets[i])
                f.write(self.bases[o:o + int(self.lengths[i])].tobytes())
                f.write(b"\n")


def gen_reads(genome_len, coverage, len_min, len_max, err, seed, sub_only=False,
              repeat_frac=0.0, repeat_families=0, repeat_min=1000, repeat_max=6000, repeat_div=0.02):
    gen_so, _ = build_harness_libs()
    lib = C.CDLL(gen_so)
    lib.oic_gen_genome.restype = C.c_void_p
    lib.oic_gen_genome.argtypes = [C.POINTER(GenParams)]
    lib.oic_gen_num_reads.restype = C.c_uint64
    lib.oic_gen_num_reads.argtypes = [C.POINTER(GenParams)]
    lib.oic_gen_lengths.restype = C.c_uint64
    lib.oic_gen_lengths.argtypes = [C.POINTER(GenParams), C.c_void_p, C.c_uint64, C.c_void_p]
    lib.oic_gen_fill.argtypes = [C.POINTER(GenParams), C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    lib.oic_gen_free.argtypes = [C.c_void_p]
    p = GenParams(genome_len, coverage, len_min, len_max, err, int(sub_only), seed,
                  repeat_frac, repeat_families, repeat_min, repeat_max, repeat_div)
    g = lib.oic_gen_genome(C.byref(p))
    n = lib.oic_gen_num_reads(C.byref(p))
    lengths = np.zeros(n, dtype=np.uint32)
    total = lib.oic_gen_lengths(C.byref(p), g, n, lengths.ctypes.data)
    offsets = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(lengths, out=offsets[1:])
    bases = np.zeros(total + 1, dtype=np.uint8)
    lib.oic_gen_fill(C.byref(p), g, n, offsets.ctypes.data, bases.ctypes.data)
    lib.oic_gen_free(g)
    return Reads(bases[:total], offsets, lengths)


def encode_2bit(reads_bases, offsets, lengths, out=None):
    """ASCII reads -> gkpStore 2-bit encoding (tools/gen_reads.c oic_encode_2bit); returns (packed uint8, byte_offsets uint64[n]).
    `out`: optional preallocated uint8 buffer (e.g. the numpy view of a pinned tensor) of at least sum(ceil(len/4)) bytes."""
    gen_so, _ = build_harness_libs()
    lib = C.CDLL(gen_so)
    lib.oic_encode_2bit.restype = C.c_uint64
    lib.oic_encode_2bit.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    bases = np.ascontiguousarray(reads_bases, dtype=np.uint8)
    lengths = np.ascontiguousarray(lengths, dtype=np.uint32)
    offsets = np.ascontiguousarray(np.asarray(offsets)[:len(lengths)], dtype=np.uint64)
    total = int(((lengths.astype(np.int64) + 3) // 4).sum())
    if out is None:
        out = np.zeros(total + 8, dtype=np.uint8)
    assert out.dtype == np.uint8 and out.size >= total
    boff = np.zeros(len(lengths), dtype=np.uint64)
    w = lib.oic_encode_2bit(bases.ctypes.data, offsets.ctypes.data, lengths.ctypes.data, len(lengths), out.ctypes.data, boff.ctypes.data)
    if w == 2 ** 64 - 1:
        raise ValueError("encode_2bit: non-ACGT base")
    return out[:total], boff


def count_kmers_cpu(reads, K, threshold):
    """CPU restatement of `meryl -B -C` + `meryl -Dt -n threshold` (tools/gen_reads.c oic_frequent_kmers): canonical K-mers seen
    >= threshold times in the read set -> (keys uint64 ascending, first base in the HIGH bits; counts uint32)."""
    gen_so, _ = build_harness_libs()
    lib = C.CDLL(gen_so)
    lib.oic_frequent_kmers.restype = C.c_uint64
    lib.oic_frequent_kmers.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint64]
    bases = np.ascontiguousarray(reads.bases, dtype=np.uint8)
    offsets = np.ascontiguousarray(reads.offsets[:reads.n], dtype=np.uint64)
    lengths = np.ascontiguousarray(reads.lengths, dtype=np.uint32)
    cap = 1 << 22
    while True:
        keys = np.zeros(cap, dtype=np.uint64)
        counts = np.zeros(cap, dtype=np.uint32)
        n = lib.oic_frequent_kmers(bases.ctypes.data, offsets.ctypes.data, lengths.ctypes.data, reads.n, K, threshold,
                                   keys.ctypes.data, counts.ctypes.data, cap)
        if n == 2 ** 64 - 1:
            raise MemoryError("oic_frequent_kmers")
        if n <= cap:
            break
        cap = int(n)
    order = np.argsort(keys[:n], kind="stable")
    return keys[:n][order], counts[:n][order]


def kmer_strings(keys, K, upper=False):
    """uint64 keys (first base in the high bits) -> list of K-letter strings"""
    keys = np.asarray(keys, dtype=np.uint64)
    codes = np.zeros((len(keys), K), dtype=np.uint8)
    for i in range(K):
        codes[:, i] = (keys >> np.uint64(2 * (K - 1 - i))) & np.uint64(3)
    letters = np.frombuffer(b"ACGT" if upper else b"acgt", dtype=np.uint8)[codes]
    return [row.tobytes().decode() for row in letters]


def meryl_dump_text(keys, counts, K):
    """the text `meryl -Dt -n T` prints (meryl-dump.C): ">count\nKMER\n" per K-mer, ascending"""
    return "".join(">%d\n%s\n" % (int(c), k) for c, k in zip(counts, kmer_strings(keys, K, upper=True)))


def meryl_hist_text(hist):
    """the text of `meryl -Dh` (meryl-dump.C:120-152) from hist[i] = distinct K-mers seen exactly i times -> (stdout, stderr)"""
    hist = [int(x) for x in hist]
    n_distinct = sum(hist[1:])
    n_total = sum(i * h for i, h in enumerate(hist))
    out, distinct, total = [], 0, 0
    for i in range(1, len(hist)):
        if hist[i]:
            distinct += hist[i]
            total += hist[i] * i
            out.append("%d\t%d\t%.4f\t%.4f\n" % (i, hist[i], distinct / float(n_distinct), total / float(n_total)))
    info = "Found %d mers.\nFound %d distinct mers.\nFound %d unique mers.\nLargest mercount is %d.\n" % (
        n_total, n_distinct, hist[1] if len(hist) > 1 else 0, len(hist) - 1)
    return "".join(out), info


def kmer_histogram_cpu(reads, K):
    """hist[i] = distinct canonical K-mers seen exactly i times, from the CPU restatement"""
    _, counts = count_kmers_cpu(reads, K, 1)
    return np.bincount(counts.astype(np.int64)) if len(counts) else np.zeros(1, dtype=np.int64)


def run_meryl_hist(store, K, workdir, threads=4):
    """`meryl -Dh` on a database built as the pipeline builds it (Meryl.pm:434-455) -> (stdout, stderr)"""
    db = os.path.join(workdir, "refh.ms%d" % K)
    with open(os.path.join(workdir, "merylh.err"), "w") as ef:
        subprocess.check_call([os.path.join(REF_DIR, "meryl"), "-B", "-C", "-L", "2", "-v", "-m", str(K), "-threads", str(threads),
                               "-memory", "1024", "-s", store, "-o", db], stdout=ef, stderr=ef)
    r = subprocess.run([os.path.join(REF_DIR, "meryl"), "-Dh", "-s", db], capture_output=True, check=True)
    return r.stdout.decode(), r.stderr.decode()


def frequent_kmers_fast(reads, K, threshold):
    """Frequent canonical K-mers of a read set as strings: the harness's stand-in for the pipeline's `meryl -Dt -n T` skip list
    at sizes where the pure-Python counter of tests/cases.py is too slow."""
    return kmer_strings(count_kmers_cpu(reads, K, threshold)[0], K)


def run_meryl(store, K, threshold, workdir, threads=4):
    """reference meryl as the pipeline runs it (Meryl.pm:434-450, 634): build the canonical count database, dump mers >= threshold"""
    db = os.path.join(workdir, "ref.ms%d" % K)
    with open(os.path.join(workdir, "meryl.err"), "w") as ef:
        subprocess.check_call([os.path.join(REF_DIR, "meryl"), "-B", "-C", "-L", "2", "-v", "-m", str(K), "-threads", str(threads),
                               "-memory", "1024", "-s", store, "-o", db], stdout=ef, stderr=ef)
        return subprocess.check_output([os.path.join(REF_DIR, "meryl"), "-Dt", "-n", str(threshold), "-s", db], stderr=ef).decode()


# ----------------------------------------------------------------------------- reference tools

def make_store(reads, workdir, minlength=0):
    """reads -> FASTA -> reference gatekeeperCreate -> <workdir>/t.gkpStore"""
    os.makedirs(workdir, exist_ok=True)
    gkp = os.path.join(workdir, "in.gkp")
    qv = getattr(reads, "read_qv", None)
    if qv is None:
        fa = os.path.join(workdir, "reads.fasta")
        reads.write_fasta(fa)
        with open(gkp, "w") as f:
            f.write("name reads\npreset pacbio-corrected\n%s\n" % fa)
    else:   # one library (with its own default QV) per run of reads with the same value; read IDs stay in order
        with open(gkp, "w") as f:
            i = 0
            while i < reads.n:
                j = i
                while j < reads.n and qv[j] == qv[i]:
                    j += 1
                fa = os.path.join(workdir, "reads_%d.fasta" % i)
                reads.write_fasta(fa, i, j)
                f.write("name lib%d\npreset pacbio-corrected\nqv %d\n%s\n" % (i, int(qv[i]), fa))
                i = j
    store = os.path.join(workdir, "t.gkpStore")
    subprocess.check_call(["rm", "-rf", store])
    subprocess.check_call([os.path.join(REF_DIR, "gatekeeperCreate"), "-minlength", str(minlength), "-o", store, gkp],
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return store


def run_overlapper(binary, store, out_prefix, args, threads=4):
    """Run an overlapInCore-compatible binary; returns (ovb path, stats path)."""
    ovb = out_prefix + ".ovb"
    stats = out_prefix + ".stats"
    cmd = [binary, "-t", str(threads)] + list(args) + ["-o", ovb, "-s", stats, store]
    with open(out_prefix + ".err", "w") as ef:
        subprocess.check_call(cmd, stderr=ef, stdout=ef)
    return ovb, stats


def dump_ovb_sorted(store, ovb):
    """reference overlapConvert -raw | sort  (the parity comparator of SURVEY.md 8(c))"""
    out = subprocess.check_output([os.path.join(REF_DIR, "overlapConvert"), "-G", store, "-raw", ovb],
                                  stderr=subprocess.DEVNULL)
    lines = out.decode().splitlines()
    lines.sort()
    return lines


def format_raw(o):
    """Render decoded overlap fields the way ovOverlap::toString(ovOverlapAsRaw) does (ST/ovOverlap.C:67-79)."""
    return "%10u %10u  %c  %6d  %6u %6u  %6u %6u  %7.6f %s %s %s" % (
        o["a_iid"], o["b_iid"], "I" if o["flipped"] else "N", o["span"],
        o["ahg5"], o["ahg3"], o["bhg5"], o["bhg3"], o["evalue"] / 10000.0,
        "OBT" if o["for_obt"] else "   ", "DUP" if o["for_dup"] else "   ", "UTG" if o["for_utg"] else "   ")


def parse_stats(path):
    d = {}
    for line in open(path):
        if "=" in line:
            k, v = line.split("=")
            d[k.strip()] = int(v)
    return d


def decode_ovb(path):
    """Native .ovb decoder (no reference tool): needs a snappy decompressor -> provided by canu_b200.ovb"""
    from canu_b200 import ovb
    return ovb.read_ovb(path)


# ----------------------------------------------------------------------------- oracle

class OracleParams(C.Structure):
    _fields_ = [("kmer_len", C.c_int32), ("max_erate", C.c_double), ("min_olap_len", C.c_int32),
                ("partial", C.c_int32), ("unique_olap", C.c_int32), ("use_hopeless_check", C.c_int32),
                ("hash_bits", C.c_uint32), ("hash_load", C.c_double), ("max_hash_strings", C.c_uint32),
                ("max_hash_data_len", C.c_uint64), ("frag_olap_limit", C.c_uint64),
                ("filter_by_kmer_count", C.c_uint64), ("num_threads", C.c_uint32),
                ("bgn_hash_id", C.c_uint32), ("end_hash_id", C.c_uint32),
                ("bgn_ref_id", C.c_uint32), ("end_ref_id", C.c_uint32)]


OVERLAP_DT = np.dtype([(n, np.uint32) for n in
                       ("a_iid", "b_iid", "ahg5", "ahg3", "bhg5", "bhg3", "span", "evalue",
                        "flipped", "for_obt", "for_dup", "for_utg")])
MATCH_DT = np.dtype([("offset", np.int32), ("len", np.int32), ("start", np.int32)])
PAIR_DT = np.dtype([("q_id", np.uint32), ("h_id", np.uint32), ("strand", np.uint32), ("consistent", np.uint32),
                    ("diag_ct", np.int32), ("diag_bgn", np.int32), ("diag_end", np.int32), ("n_matches", np.uint32),
                    ("first_match", np.uint64)])
EXT_DT = np.dtype([("q_id", np.uint32), ("h_id", np.uint32), ("strand", np.uint32),
                   ("m_offset", np.int32), ("m_len", np.int32), ("m_start", np.int32),
                   ("s_lo", np.int32), ("s_hi", np.int32), ("t_lo", np.int32), ("t_hi", np.int32),
                   ("errors", np.int32), ("kind", np.int32), ("delta_len", np.int32), ("delta_hash", np.int32)])
WINDOW_FIELDS = ["bad_short_window", "bad_long_window"]
STATS_FIELDS = ["kmer_hits_without_olap", "kmer_hits_with_olap", "kmer_hits_skipped", "multi_overlap",
                "total_overlaps", "contained", "dovetail", "n_hash_bases", "n_inserts", "n_distinct",
                "n_query_bases", "n_windows", "n_found", "n_refs", "n_add_ref", "ext_calls", "ext_cells",
                "ext_words"]


class OracleStats(C.Structure):
    _fields_ = ([(n, C.c_uint64) for n in STATS_FIELDS] + [("n_hash_blocks", C.c_uint32), ("last_end_hash_id", C.c_uint32)] +
                [(n, C.c_uint64) for n in WINDOW_FIELDS])


def strtof(s):
    """--maxerate is parsed with strtof and widened to double (OIC/overlapInCore.C:395)."""
    return float(np.float32(float(s)))


def oracle_lib():
    _, o_so = build_harness_libs()
    lib = C.CDLL(o_so)
    lib.oracle_create.restype = C.c_void_p
    lib.oracle_create.argtypes = [C.POINTER(OracleParams), C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_void_p]
    lib.oracle_set_skip_kmers.argtypes = [C.c_void_p, C.c_uint64, C.c_char_p]
    lib.oracle_set_dump.argtypes = [C.c_void_p, C.c_uint32]
    lib.oracle_set_qualities.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.oracle_run.argtypes = [C.c_void_p]
    for name in ("overlaps", "pairs", "matches", "exts"):
        getattr(lib, "oracle_num_" + name).restype = C.c_uint64
        getattr(lib, "oracle_num_" + name).argtypes = [C.c_void_p]
        getattr(lib, "oracle_" + name).restype = C.c_void_p
        getattr(lib, "oracle_" + name).argtypes = [C.c_void_p]
    lib.oracle_stats.restype = C.POINTER(OracleStats)
    lib.oracle_stats.argtypes = [C.c_void_p]
    lib.oracle_index_num_keys.restype = C.c_uint64
    lib.oracle_index_num_keys.argtypes = [C.c_void_p]
    for name in ("keys", "occ_start", "occ", "flags"):
        getattr(lib, "oracle_index_" + name).restype = C.c_void_p
        getattr(lib, "oracle_index_" + name).argtypes = [C.c_void_p]
    lib.oracle_destroy.argtypes = [C.c_void_p]
    lib.oracle_max_errors.restype = C.c_int32
    lib.oracle_max_errors.argtypes = [C.c_double]
    lib.oracle_match_limit.argtypes = [C.c_double, C.c_int32, C.c_void_p]
    lib.oracle_error_bound.restype = C.c_int32
    lib.oracle_error_bound.argtypes = [C.c_double, C.c_int32]
    lib.oracle_extend.restype = C.c_int32
    lib.oracle_extend.argtypes = [C.c_double, C.c_int32, C.c_char_p, C.c_int32, C.c_char_p, C.c_int32, C.c_int32,
                                  C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
                                  C.c_void_p, C.c_void_p]
    return lib


def _copy(ptr, n, dt):
    if n == 0:
        return np.zeros(0, dtype=dt)
    buf = (C.c_char * (n * dt.itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dt).copy()


def default_params(**kw):
    d = dict(kmer_len=22, max_erate=strtof("0.06"), min_olap_len=500, partial=0, unique_olap=1,
             use_hopeless_check=1, hash_bits=22, hash_load=0.6, max_hash_strings=10000,
             max_hash_data_len=100000000, frag_olap_limit=2 ** 64 - 1, filter_by_kmer_count=0,
             num_threads=1, bgn_hash_id=1, end_hash_id=2 ** 32 - 1, bgn_ref_id=1, end_ref_id=2 ** 32 - 1, use_window_filter=0)
    d.update(kw)
    if d["max_erate"] > 0.06:
        d["use_hopeless_check"] = 0
        d["use_window_filter"] = 0   # OIC/overlapInCore.C:416-421
    return d


def run_oracle(reads, params, dump=0, skip_kmers=None):
    """Run the CPU restatement. Returns dict(overlaps, stats, pairs, matches, exts, index).  params["use_window_filter"]: -w, with the
    qualities of `reads` (Reads.qualities())."""
    lib = oracle_lib()
    p = OracleParams(**{k: v for k, v in params.items() if k != "use_window_filter"})
    offs = np.ascontiguousarray(reads.offsets[:-1], dtype=np.uint64)
    lens = np.ascontiguousarray(reads.lengths, dtype=np.uint32)
    bases = np.ascontiguousarray(reads.bases)
    ctx = lib.oracle_create(C.byref(p), reads.n, offs.ctypes.data, lens.ctypes.data, bases.ctypes.data, None, None)
    if skip_kmers:
        lib.oracle_set_skip_kmers(ctx, len(skip_kmers), "".join(skip_kmers).encode())
    if params.get("use_window_filter"):
        q = reads.qualities()
        lib.oracle_set_qualities(ctx, offs.ctypes.data, q.ctypes.data)
    lib.oracle_set_dump(ctx, dump)
    lib.oracle_run(ctx)
    res = {}
    res["overlaps"] = _copy(lib.oracle_overlaps(ctx), lib.oracle_num_overlaps(ctx), OVERLAP_DT)
    st = lib.oracle_stats(ctx).contents
    res["stats"] = {n: getattr(st, n) for n in STATS_FIELDS + ["n_hash_blocks", "last_end_hash_id"] + WINDOW_FIELDS}
    res["pairs"] = _copy(lib.oracle_pairs(ctx), lib.oracle_num_pairs(ctx), PAIR_DT)
    res["matches"] = _copy(lib.oracle_matches(ctx), lib.oracle_num_matches(ctx), MATCH_DT)
    res["exts"] = _copy(lib.oracle_exts(ctx), lib.oracle_num_exts(ctx), EXT_DT)
    nk = lib.oracle_index_num_keys(ctx)
    if nk:
        u64 = np.dtype(np.uint64)
        res["index"] = dict(keys=_copy(lib.oracle_index_keys(ctx), nk, u64),
                            occ_start=_copy(lib.oracle_index_occ_start(ctx), nk + 1, u64),
                            flags=_copy(lib.oracle_index_flags(ctx), nk, np.dtype(np.uint8)))
        res["index"]["occ"] = _copy(lib.oracle_index_occ(ctx), int(res["index"]["occ_start"][-1]), u64)
    lib.oracle_destroy(ctx)
    return res


def oic_args(params, erate_str):
    """argv (minus -t/-o/-s/store) for an overlapInCore-compatible binary equivalent to an oracle param dict."""
    a = []
    if params["partial"]:
        a.append("-G")
    if not params["unique_olap"]:
        a.append("-m")
    if not params["use_hopeless_check"] and float(erate_str) <= 0.06:
        a.append("-z")
    if params.get("use_window_filter"):
        a.append("-w")
    a += ["-k", str(params["kmer_len"]), "--hashbits", str(params["hash_bits"]),
          "--hashload", repr(params["hash_load"]), "--maxerate", erate_str,
          "--minlength", str(params["min_olap_len"])] + (["--minkmers"] if params.get("filter_by_kmer_count") else []) + [
          "--hashstrings", str(params["max_hash_strings"]), "--hashdatalen", str(params["max_hash_data_len"])]
    if params.get("frag_olap_limit", 2 ** 64 - 1) != 2 ** 64 - 1:
        a += ["-l", str(params["frag_olap_limit"])]
    if params["end_hash_id"] != 2 ** 32 - 1 or params["bgn_hash_id"] != 1:
        a += ["-h", "%d-%d" % (params["bgn_hash_id"], params["end_hash_id"])]
    if params["end_ref_id"] != 2 ** 32 - 1 or params["bgn_ref_id"] != 1:
        a += ["-r", "%d-%d" % (params["bgn_ref_id"], params["end_ref_id"])]
    return a
