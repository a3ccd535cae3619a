"""ctypes binding of the C ABI (include/oic_b200.h) -- the host-side mirror used by tests and bench.py.

The names follow the reference's call stack for this path (OIC/ = src/overlapInCore/ in the reference):
``Overlapper.build_hash_index`` = Build_Hash_Index (OIC/overlapInCore-Build_Hash_Index.C:443),
``Overlapper.process_overlaps`` = Process_Overlaps (OIC/overlapInCore-Process_Overlaps.C:78),
``Overlapper.overlap_driver`` = OverlapDriver (OIC/overlapInCore.C:190).

There is no CPU path: loading fails loudly if liboic_b200.so is missing, and creating an Overlapper fails
if there is no CUDA device.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboic_b200.so")

ABI_SYMBOLS = [
    "oic_create", "oic_destroy", "oic_last_error", "oic_load_reads", "oic_load_reads_2bit", "oic_set_skip_kmers", "oic_build_hash_index",
    "oic_process_overlaps", "oic_overlap_driver", "oic_last_processed_ref", "oic_num_overlaps", "oic_copy_overlaps",
    "oic_clear_overlaps", "oic_get_stats", "oic_get_timings", "oic_reset_timings", "oic_hash_block_meta",
    "oic_hash_block_reserve", "oic_hash_block_buffers", "oic_hash_block_commit", "oic_dump_index", "oic_set_debug",
    "oic_dump_pairs", "oic_measure_int32_peak", "oic_set_job", "oic_reset_stats", "oic_count_kmers", "oic_copy_frequent_kmers",
    "oic_copy_kmer_histogram", "oic_comm_unique_id", "oic_comm_init_rank", "oic_comm_init_all", "oic_comm_info",
    "oic_hash_block_replicate", "oic_build_hash_index_sharded", "oic_ovlstore_sort", "oic_ovlstore_sort_noalloc",
    "oic_ovlstore_sort_noalloc_max", "oic_load_qualities",
]

# status codes of include/oic_b200.h
OIC_OK, OIC_ERR_CUDA, OIC_ERR_ARG, OIC_ERR_UNSUPPORTED, OIC_ERR_BASES, OIC_ERR_CAPACITY, OIC_ERR_STATE = 0, -1, -2, -3, -4, -5, -6


class OicError(RuntimeError):
    pass


class Params(C.Structure):
    _fields_ = [("kmer_len", C.c_int32), ("max_erate", C.c_double), ("min_olap_len", C.c_int32),
                ("partial", C.c_int32), ("unique_olap", C.c_int32), ("use_hopeless_check", C.c_int32),
                ("hash_bits", C.c_uint32), ("hash_load", C.c_double), ("max_hash_strings", C.c_uint32),
                ("max_hash_data_len", C.c_uint64), ("frag_olap_limit", C.c_uint64),
                ("filter_by_kmer_count", C.c_uint64), ("num_threads", C.c_uint32),
                ("bgn_hash_id", C.c_uint32), ("end_hash_id", C.c_uint32),
                ("bgn_ref_id", C.c_uint32), ("end_ref_id", C.c_uint32), ("use_window_filter", C.c_int32)]


STATS_FIELDS = ["kmer_hits_without_olap", "kmer_hits_with_olap", "kmer_hits_skipped", "multi_overlap",
                "total_overlaps", "contained", "dovetail", "n_hash_bases", "n_inserts", "n_distinct",
                "n_query_bases", "n_windows", "n_found", "n_refs", "n_add_ref", "n_runs", "n_pairs",
                "ext_calls", "ext_cells", "ext_words"]


class Stats(C.Structure):
    _fields_ = ([(n, C.c_uint64) for n in STATS_FIELDS] + [("n_hash_blocks", C.c_uint32), ("last_end_hash_id", C.c_uint32)] +
                [("bad_short_window", C.c_uint64), ("bad_long_window", C.c_uint64)])


class Timings(C.Structure):
    _fields_ = [("ms_pack", C.c_float), ("ms_build", C.c_float), ("ms_scan", C.c_float), ("ms_group", C.c_float),
                ("ms_extend", C.c_float), ("ms_total_search", C.c_float), ("launches", C.c_uint64),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("ms_ext_first", C.c_float), ("ms_ext_tasks", C.c_float), ("ms_ext_commit", C.c_float), ("ms_ext_retry", C.c_float)]


class HashMeta(C.Structure):
    _fields_ = [("table_slots", C.c_uint64), ("n_occ", C.c_uint64), ("n_strings", C.c_uint32), ("bgn_id", C.c_uint32),
                ("last_id", C.c_uint32), ("has_empty", C.c_uint32), ("n_distinct", C.c_uint64),
                ("n_slices", C.c_uint32), ("slice_bits", C.c_uint32), ("filter_words", C.c_uint64)]


HASH_BLOCK_NBUF = 4
COMM_ID_BYTES = 128


class DevBuffer(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("bytes", C.c_uint64)]


_lib = None


def load_library():
    """dlopen liboic_b200.so (built in tree by canu_b200/csrc/Makefile) and declare the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise OicError("%s is missing: build it with `make -C canu_b200/csrc` (there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(os.environ.get("OIC_LIB", LIB_PATH))   # OIC_LIB: A/B builds of the same library during kernel work
    vp = C.c_void_p
    lib.oic_create.argtypes = [C.POINTER(Params), C.c_int, C.POINTER(vp)]
    lib.oic_destroy.argtypes = [vp]
    lib.oic_last_error.restype = C.c_char_p
    lib.oic_last_error.argtypes = [vp]
    lib.oic_load_reads.argtypes = [vp, C.c_uint32, C.c_uint32, vp, vp, vp, vp, vp]
    lib.oic_load_reads_2bit.argtypes = [vp, C.c_uint32, C.c_uint32, vp, vp, vp, vp, vp]
    lib.oic_load_qualities.argtypes = [vp, vp, vp]
    lib.oic_set_skip_kmers.argtypes = [vp, C.c_uint64, C.c_char_p]
    lib.oic_set_job.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64]
    lib.oic_reset_stats.argtypes = [vp]
    lib.oic_build_hash_index.argtypes = [vp, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]
    lib.oic_count_kmers.argtypes = [vp, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.POINTER(C.c_double)]
    lib.oic_copy_frequent_kmers.argtypes = [vp, vp, vp, C.c_uint64, C.POINTER(C.c_uint64)]
    lib.oic_copy_kmer_histogram.argtypes = [vp, vp, C.c_uint64, C.POINTER(C.c_uint64)]
    lib.oic_process_overlaps.argtypes = [vp, C.c_uint32, C.c_uint32]
    lib.oic_overlap_driver.argtypes = [vp]
    lib.oic_last_processed_ref.restype = C.c_uint32
    lib.oic_last_processed_ref.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32]
    lib.oic_num_overlaps.restype = C.c_uint64
    lib.oic_num_overlaps.argtypes = [vp]
    lib.oic_copy_overlaps.argtypes = [vp, vp, C.c_uint64, C.POINTER(C.c_uint64)]
    lib.oic_clear_overlaps.argtypes = [vp]
    lib.oic_get_stats.argtypes = [vp, C.POINTER(Stats)]
    lib.oic_get_timings.argtypes = [vp, C.POINTER(Timings)]
    lib.oic_reset_timings.argtypes = [vp]
    lib.oic_hash_block_meta.argtypes = [vp, C.POINTER(HashMeta)]
    lib.oic_hash_block_reserve.argtypes = [vp, C.POINTER(HashMeta)]
    lib.oic_hash_block_buffers.argtypes = [vp, C.POINTER(DevBuffer)]
    lib.oic_hash_block_commit.argtypes = [vp, C.POINTER(HashMeta)]
    lib.oic_comm_unique_id.argtypes = [vp]
    lib.oic_comm_init_rank.argtypes = [vp, vp, C.c_int, C.c_int]
    lib.oic_comm_init_all.argtypes = [C.POINTER(vp), C.c_int]
    lib.oic_comm_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.oic_hash_block_replicate.argtypes = [vp, vp, C.c_int]
    lib.oic_build_hash_index_sharded.argtypes = [vp, vp, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]
    lib.oic_ovlstore_sort.argtypes = [vp, vp, C.c_uint64, C.c_uint32, vp, vp, vp, C.c_uint32, vp, C.POINTER(C.c_uint64), vp, C.POINTER(C.c_double)]
    # allocation-free variant of the store sort: DEVICE pointers in and out, scratch in a static __device__ array of the library
    lib.oic_ovlstore_sort_noalloc.argtypes = [C.c_int, vp, C.c_uint64, C.c_uint32, vp, vp, vp, C.c_uint32, vp, C.POINTER(C.c_uint64), vp]
    lib.oic_ovlstore_sort_noalloc_max.restype = C.c_uint64
    lib.oic_ovlstore_sort_noalloc_max.argtypes = []
    lib.oic_dump_index.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), vp, vp, vp, vp]
    lib.oic_set_debug.argtypes = [vp, C.c_uint32]
    lib.oic_dump_pairs.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), vp, vp]
    _lib = lib
    return lib


def strtof(s):
    """--maxerate is parsed with strtof and widened to double (OIC/overlapInCore.C:395)."""
    return float(np.float32(float(s)))


def default_params(**kw):
    """oicParameters::initialize() defaults (OIC/overlapInCore.H:425-464)."""
    d = dict(kmer_len=22, max_erate=strtof("0.06"), min_olap_len=0, partial=0, unique_olap=1,
             use_hopeless_check=1, hash_bits=22, hash_load=0.6, max_hash_strings=10000,
             max_hash_data_len=100000000, frag_olap_limit=2 ** 64 - 1, filter_by_kmer_count=0,
             num_threads=1, bgn_hash_id=1, end_hash_id=2 ** 32 - 1, bgn_ref_id=1, end_ref_id=2 ** 32 - 1, use_window_filter=0)
    d.update(kw)
    if d["max_erate"] > 0.06:
        d["use_hopeless_check"] = 0
        d["use_window_filter"] = 0
    return d


def encode_2bit(bases, offsets, lengths):
    """ASCII reads -> the gkpStore 2-bit encoding (ST/gkStoreEncode.C:45-78): 4 bases per byte, first base in bits 7-6,
    every read starting on a byte boundary.  Returns (packed uint8, byte_offsets uint64[n]).  Raises on a non-ACGT base
    (the store keeps such reads unencoded)."""
    bases = np.asarray(bases, dtype=np.uint8)
    lengths = np.asarray(lengths, dtype=np.int64)
    offsets = np.asarray(offsets, dtype=np.int64)[:len(lengths)]
    lut = np.full(256, 255, dtype=np.uint8)
    for ch, v in (("a", 0), ("c", 1), ("g", 2), ("t", 3)):
        lut[ord(ch)] = lut[ord(ch.upper())] = v
    nbytes = (lengths + 3) // 4
    boff = np.zeros(len(lengths), dtype=np.uint64)
    if len(lengths) > 1:
        np.cumsum(nbytes[:-1], out=boff[1:])
    total = int(nbytes.sum())
    # position of every base in a 4-bases-per-byte padded layout
    padded = np.zeros(total * 4, dtype=np.uint8)
    read_of_base = np.repeat(np.arange(len(lengths)), lengths)
    within = np.arange(int(lengths.sum()), dtype=np.int64) - np.repeat(np.cumsum(lengths) - lengths, lengths)
    src = np.repeat(offsets, lengths) + within
    codes = lut[bases[src]]
    if (codes == 255).any():
        raise OicError("encode_2bit: non-ACGT base")
    padded[boff.astype(np.int64)[read_of_base] * 4 + within] = codes
    q = padded.reshape(-1, 4)
    packed = (q[:, 0] << 6) | (q[:, 1] << 4) | (q[:, 2] << 2) | q[:, 3]
    return packed.astype(np.uint8), boff


RECORD_DT = np.dtype((np.uint32, 6))


def decode_records(rec):
    """(n, 6) uint32 file words -> dict of decoded fields (ST/ovOverlap.H:93-124, ST/ovStoreFile.C:196-221)."""
    rec = np.asarray(rec, dtype=np.uint32).reshape(-1, 6)
    d0 = (rec[:, 2].astype(np.uint64) << np.uint64(32)) | rec[:, 3].astype(np.uint64)
    d1 = (rec[:, 4].astype(np.uint64) << np.uint64(32)) | rec[:, 5].astype(np.uint64)
    m21 = np.uint64((1 << 21) - 1)

    def f(x, sh, mask):
        return ((x >> np.uint64(sh)) & np.uint64(mask)).astype(np.uint32)
    return dict(a_iid=rec[:, 0].copy(), b_iid=rec[:, 1].copy(),
                ahg5=f(d0, 0, m21), ahg3=f(d0, 21, m21), evalue=f(d0, 42, 4095), flipped=f(d0, 54, 1),
                for_obt=f(d0, 55, 1), for_dup=f(d0, 56, 1), for_utg=f(d0, 57, 1),
                bhg5=f(d1, 0, m21), bhg3=f(d1, 21, m21), span=f(d1, 42, m21))


class Overlapper:
    """One context = one GPU.  Mirrors the life cycle of the reference process (OIC/overlapInCore.C:306-595)."""

    def __init__(self, params, device=0):
        self.lib = load_library()
        self.params = dict(params)
        self._p = Params(**params)
        self._h = C.c_void_p()
        rc = self.lib.oic_create(C.byref(self._p), device, C.byref(self._h))
        if rc != 0:
            raise OicError("oic_create failed (%d): %s" % (rc, self.lib.oic_last_error(None).decode()))
        self._keep = None

    def close(self):
        if self._h:
            self.lib.oic_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise OicError("%s failed (%d): %s" % (what, rc, self.lib.oic_last_error(self._h).decode()))

    # gkStore_loadReadData + tolower copy loops
    def load_reads(self, bases, offsets, lengths, first_id=1, hash_ok=None, ref_ok=None):
        bases = np.ascontiguousarray(bases, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        lengths = np.ascontiguousarray(lengths, dtype=np.uint32)
        n = len(lengths)
        ho = np.ascontiguousarray(hash_ok, dtype=np.uint8) if hash_ok is not None else None
        ro = np.ascontiguousarray(ref_ok, dtype=np.uint8) if ref_ok is not None else None
        self._check(self.lib.oic_load_reads(self._h, first_id, n, offsets.ctypes.data, lengths.ctypes.data,
                                            bases.ctypes.data, ho.ctypes.data if ho is not None else None,
                                            ro.ctypes.data if ro is not None else None), "oic_load_reads")
        self.n_reads = n
        self.first_id = first_id

    def load_reads_ptr(self, bases_ptr, offsets, lengths, first_id=1):
        """Same, with the bases given as a raw host pointer (e.g. a pinned torch tensor's data_ptr())."""
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        lengths = np.ascontiguousarray(lengths, dtype=np.uint32)
        self._check(self.lib.oic_load_reads(self._h, first_id, len(lengths), offsets.ctypes.data, lengths.ctypes.data,
                                            bases_ptr, None, None), "oic_load_reads")
        self.n_reads = len(lengths)
        self.first_id = first_id

    def load_reads_2bit(self, packed, byte_offsets, lengths, first_id=1, hash_ok=None, ref_ok=None):
        """Reads in the gkpStore's 2-bit encoding (see encode_2bit / include/oic_b200.h); `packed` may be a numpy array or a
        raw host pointer (e.g. a pinned torch tensor's data_ptr())."""
        if not isinstance(packed, int):
            packed = np.ascontiguousarray(packed, dtype=np.uint8)
            self._keep = packed
            packed = packed.ctypes.data
        offsets = np.ascontiguousarray(byte_offsets, dtype=np.uint64)
        lengths = np.ascontiguousarray(lengths, dtype=np.uint32)
        ho = np.ascontiguousarray(hash_ok, dtype=np.uint8) if hash_ok is not None else None
        ro = np.ascontiguousarray(ref_ok, dtype=np.uint8) if ref_ok is not None else None
        self._check(self.lib.oic_load_reads_2bit(self._h, first_id, len(lengths), offsets.ctypes.data, lengths.ctypes.data, packed,
                                                 ho.ctypes.data if ho is not None else None,
                                                 ro.ctypes.data if ro is not None else None), "oic_load_reads_2bit")
        self.n_reads = len(lengths)
        self.first_id = first_id

    def load_qualities(self, quals, offsets):
        """-w: one quality value per base of the loaded reads (read i at quals[offsets[i] ..])"""
        quals = np.ascontiguousarray(quals, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        self._check(self.lib.oic_load_qualities(self._h, offsets.ctypes.data, quals.ctypes.data), "oic_load_qualities")

    def set_job(self, bgn_hash_id, end_hash_id, bgn_ref_id, end_ref_id, max_hash_strings, max_hash_data_len):
        """New -h / -r / --hashstrings / --hashdatalen for the resident reads (one job of overlapInCorePartition)."""
        self._check(self.lib.oic_set_job(self._h, bgn_hash_id, end_hash_id, bgn_ref_id, end_ref_id, max_hash_strings,
                                         max_hash_data_len), "oic_set_job")

    def reset_stats(self):
        self.lib.oic_reset_stats(self._h)

    def set_skip_kmers(self, kmers):
        self._check(self.lib.oic_set_skip_kmers(self._h, len(kmers), "".join(kmers).encode()), "oic_set_skip_kmers")

    def count_kmers(self, K, min_count):
        """canonical K-mers of all loaded reads seen >= min_count times (meryl -B -C + meryl -Dt -n): (keys uint64 ascending, first
        base in the high bits; counts uint32; info dict with passes / device_ms)"""
        n, passes, ms = C.c_uint64(0), C.c_uint32(0), C.c_double(0.0)
        self._check(self.lib.oic_count_kmers(self._h, K, min_count, C.byref(n), C.byref(passes), C.byref(ms)), "oic_count_kmers")
        keys = np.zeros(n.value, dtype=np.uint64)
        counts = np.zeros(n.value, dtype=np.uint32)
        got = C.c_uint64(0)
        self._check(self.lib.oic_copy_frequent_kmers(self._h, keys.ctypes.data, counts.ctypes.data, n.value, C.byref(got)), "oic_copy_frequent_kmers")
        hl = C.c_uint64(0)
        self._check(self.lib.oic_copy_kmer_histogram(self._h, None, 0, C.byref(hl)), "oic_copy_kmer_histogram")
        hist = np.zeros(hl.value, dtype=np.uint64)
        self._check(self.lib.oic_copy_kmer_histogram(self._h, hist.ctypes.data, hl.value, C.byref(hl)), "oic_copy_kmer_histogram")
        return keys[:got.value], counts[:got.value], {"passes": passes.value, "device_ms": ms.value, "histogram": hist}

    def build_hash_index(self, bgn_id, end_id):
        last = C.c_uint32(0)
        self._check(self.lib.oic_build_hash_index(self._h, bgn_id, end_id, C.byref(last)), "oic_build_hash_index")
        return last.value

    def process_overlaps(self, bgn_ref_id, end_ref_id):
        self._check(self.lib.oic_process_overlaps(self._h, bgn_ref_id, end_ref_id), "oic_process_overlaps")

    def overlap_driver(self):
        self._check(self.lib.oic_overlap_driver(self._h), "oic_overlap_driver")

    def last_processed_ref(self, bgn, end, threads):
        return self.lib.oic_last_processed_ref(bgn, end, threads)

    def num_overlaps(self):
        return self.lib.oic_num_overlaps(self._h)

    def overlaps(self):
        """(n, 6) uint32: the 24-byte .ovb records."""
        n = self.num_overlaps()
        out = np.zeros((n, 6), dtype=np.uint32)
        got = C.c_uint64(0)
        self._check(self.lib.oic_copy_overlaps(self._h, out.ctypes.data, n, C.byref(got)), "oic_copy_overlaps")
        return out[:got.value]

    def clear_overlaps(self):
        self.lib.oic_clear_overlaps(self._h)

    def stats(self):
        s = Stats()
        self._check(self.lib.oic_get_stats(self._h, C.byref(s)), "oic_get_stats")
        return {n: getattr(s, n) for n in STATS_FIELDS + ["n_hash_blocks", "last_end_hash_id", "bad_short_window", "bad_long_window"]}

    def timings(self):
        t = Timings()
        self._check(self.lib.oic_get_timings(self._h, C.byref(t)), "oic_get_timings")
        return {n: getattr(t, n) for n, _ in Timings._fields_}

    def reset_timings(self):
        self.lib.oic_reset_timings(self._h)

    # ---- hash block replication (multi-GPU)
    def hash_block_meta(self):
        m = HashMeta()
        self._check(self.lib.oic_hash_block_meta(self._h, C.byref(m)), "oic_hash_block_meta")
        return m

    def hash_block_reserve(self, meta):
        self._check(self.lib.oic_hash_block_reserve(self._h, C.byref(meta)), "oic_hash_block_reserve")

    def hash_block_buffers(self):
        b = (DevBuffer * HASH_BLOCK_NBUF)()
        self._check(self.lib.oic_hash_block_buffers(self._h, b), "oic_hash_block_buffers")
        return [(x.ptr, x.bytes) for x in b]

    def hash_block_commit(self, meta):
        self._check(self.lib.oic_hash_block_commit(self._h, C.byref(meta)), "oic_hash_block_commit")

    # ---- the collective inside the library (NCCL)
    def comm_init_rank(self, unique_id, n_ranks, rank):
        """attach an NCCL communicator: `unique_id` = the 128 bytes rank 0 got from comm_unique_id(), shipped by the launcher"""
        buf = (C.c_uint8 * COMM_ID_BYTES).from_buffer_copy(bytes(unique_id))
        self._check(self.lib.oic_comm_init_rank(self._h, buf, n_ranks, rank), "oic_comm_init_rank")

    def comm_info(self):
        n, r = C.c_int(0), C.c_int(0)
        self._check(self.lib.oic_comm_info(self._h, C.byref(n), C.byref(r)), "oic_comm_info")
        return n.value, r.value

    def hash_block_replicate(self, root=0, comm=None):
        self._check(self.lib.oic_hash_block_replicate(self._h, comm, root), "oic_hash_block_replicate")

    def build_hash_index_sharded(self, bgn_id, end_id, comm=None):
        last = C.c_uint32(0)
        self._check(self.lib.oic_build_hash_index_sharded(self._h, comm, bgn_id, end_id, C.byref(last)), "oic_build_hash_index_sharded")
        return last.value

    # ---- overlap store construction (ovStoreBuild's filter + mirror + sort on the GPU)
    def ovlstore_sort(self, recs, read_len, skip_dup, library, max_evalue=4095):
        """(n, 6) uint32 .ovb records -> (m, 6) records of the store in ovOverlap::operator< order, filter tallies, device ms.
        read_len / skip_dup / library are indexed by read ID (entry 0 unused)."""
        recs = np.ascontiguousarray(recs, dtype=np.uint32).reshape(-1, 6)
        read_len = np.ascontiguousarray(read_len, dtype=np.uint32)
        skip_dup = np.ascontiguousarray(skip_dup, dtype=np.uint8)
        library = np.ascontiguousarray(library, dtype=np.uint8)
        assert len(read_len) == len(skip_dup) == len(library)
        out = np.zeros((2 * len(recs), 6), dtype=np.uint32)
        m, ms = C.c_uint64(0), C.c_double(0.0)
        counts = np.zeros(9, dtype=np.uint64)
        self._check(self.lib.oic_ovlstore_sort(self._h, recs.ctypes.data, len(recs), len(read_len), read_len.ctypes.data, skip_dup.ctypes.data,
                                               library.ctypes.data, max_evalue, out.ctypes.data, C.byref(m), counts.ctypes.data, C.byref(ms)),
                    "oic_ovlstore_sort")
        return out[:m.value], counts, ms.value

    # ---- stage dumps (tests)
    def set_debug(self, mask):
        self.lib.oic_set_debug(self._h, mask)

    def dump_index(self):
        nk, no = C.c_uint64(0), C.c_uint64(0)
        self._check(self.lib.oic_dump_index(self._h, C.byref(nk), C.byref(no), None, None, None, None), "oic_dump_index")
        keys = np.zeros(nk.value, dtype=np.uint64)
        start = np.zeros(nk.value + 1, dtype=np.uint64)
        occ = np.zeros(max(no.value, 1), dtype=np.uint64)
        flags = np.zeros(max(nk.value, 1), dtype=np.uint8)
        self._check(self.lib.oic_dump_index(self._h, C.byref(nk), C.byref(no), keys.ctypes.data, start.ctypes.data,
                                            occ.ctypes.data, flags.ctypes.data), "oic_dump_index")
        return dict(keys=keys, occ_start=start, occ=occ[:int(start[-1])], flags=flags[:nk.value])

    def dump_pairs(self):
        npair, nm = C.c_uint64(0), C.c_uint64(0)
        self._check(self.lib.oic_dump_pairs(self._h, C.byref(npair), C.byref(nm), None, None), "oic_dump_pairs")
        pairs = np.zeros((npair.value, 9), dtype=np.int64)
        matches = np.zeros((nm.value, 3), dtype=np.int32)
        self._check(self.lib.oic_dump_pairs(self._h, C.byref(npair), C.byref(nm), pairs.ctypes.data, matches.ctypes.data),
                    "oic_dump_pairs")
        return pairs, matches


def comm_unique_id():
    """ncclGetUniqueId through the library: 128 bytes for the launcher to ship to every rank"""
    lib = load_library()
    buf = (C.c_uint8 * COMM_ID_BYTES)()
    rc = lib.oic_comm_unique_id(buf)
    if rc != 0:
        raise OicError("oic_comm_unique_id failed (%d): %s" % (rc, lib.oic_last_error(None).decode()))
    return bytes(buf)


def comm_init_all(overlappers):
    """one process, one context per GPU: ncclCommInitAll over the contexts' devices"""
    lib = load_library()
    arr = (C.c_void_p * len(overlappers))(*[o._h for o in overlappers])
    rc = lib.oic_comm_init_all(arr, len(overlappers))
    if rc != 0:
        raise OicError("oic_comm_init_all failed (%d): %s" % (rc, lib.oic_last_error(overlappers[0]._h).decode()))


def measure_int32_peak(device=0):
    """(alu_gops, imad_gops): measured INT32 issue rates of the GPU (thread Gop/s); roofline denominator of the extension."""
    lib = load_library()
    a, m = C.c_double(0), C.c_double(0)
    lib.oic_measure_int32_peak.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    rc = lib.oic_measure_int32_peak(device, C.byref(a), C.byref(m))
    if rc != 0:
        raise OicError("oic_measure_int32_peak failed (%d)" % rc)
    return a.value, m.value
