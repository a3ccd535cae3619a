/*
 * include/oic_b200.h -- C ABI of the B200-native overlapInCore hot path (liboic_b200.so).
 *
 * The reference (sgblanch/canu, "Canu 1.6") has no plugin/FFI interface for this path: the
 * boundary it offers is the process `overlapInCore [options] gkpStore` (OIC/overlapInCore.C:306).
 * This ABI is what a reference-side maintainer (or our own drop-in executable
 * canu_b200/host/overlapInCore.cpp) binds instead of the in-process C++ calls; every entry point
 * names the reference function it replaces.  Paths: OIC/ = src/overlapInCore/,
 * LIBO/ = src/overlapInCore/liboverlap/, ST/ = src/stores/.
 *
 * Conventions: plain pointers and sizes only; every call returns 0 on success or a negative
 * OIC_ERR_* code (text via oic_last_error); the caller owns all host buffers; the library owns all
 * device memory; no exceptions cross the boundary; a context is bound to one CUDA device and is
 * not thread-safe (one context per host thread / per GPU).  There is NO CPU fallback: without a
 * CUDA device oic_create fails with OIC_ERR_CUDA.
 */
#ifndef OIC_B200_H
#define OIC_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define OIC_OK                0
#define OIC_ERR_CUDA        (-1)   /* CUDA runtime error / no device                          */
#define OIC_ERR_ARG         (-2)   /* bad argument                                            */
#define OIC_ERR_UNSUPPORTED (-3)   /* option outside this build's scope (K > 31 ...)          */
#define OIC_ERR_BASES       (-4)   /* a loaded read holds a letter other than ACGTN           */
#define OIC_ERR_CAPACITY    (-5)   /* an internal device buffer could not be sized            */
#define OIC_ERR_STATE       (-6)   /* call order violated (e.g. search before build)          */

/* Mirror of the reference's global `oicParameters G` (OIC/overlapInCore.H:418-530), restricted to
 * what the hot path reads.  Field meaning and defaults are the reference's. */
typedef struct {
  int32_t  kmer_len;             /* -k <n>            G.Kmer_Len                                     */
  double   max_erate;            /* --maxerate        G.maxErate; pass the strtof()-rounded value     */
  int32_t  min_olap_len;         /* --minlength       G.Min_Olap_Len                                  */
  int32_t  partial;              /* -G                G.Doing_Partial_Overlaps                        */
  int32_t  unique_olap;          /* -u(1)/-m(0)       G.Unique_Olap_Per_Pair                          */
  int32_t  use_hopeless_check;   /* !-z               G.Use_Hopeless_Check (forced 0 if erate > 0.06) */
  uint32_t hash_bits;            /* --hashbits        G.Hash_Mask_Bits (only feeds the load limit)    */
  double   hash_load;            /* --hashload        G.Max_Hash_Load                                 */
  uint32_t max_hash_strings;     /* --hashstrings     G.Max_Hash_Strings                              */
  uint64_t max_hash_data_len;    /* --hashdatalen     G.Max_Hash_Data_Len                             */
  uint64_t frag_olap_limit;      /* -l                G.Frag_Olap_Limit (UINT64_MAX = unlimited)      */
  uint64_t filter_by_kmer_count; /* --minkmers        G.Filter_By_Kmer_Count (0 = off)                */
  uint32_t num_threads;          /* -t                only feeds the scheduler quirk (see driver)     */
  uint32_t bgn_hash_id, end_hash_id;  /* -h a-b                                                       */
  uint32_t bgn_ref_id,  end_ref_id;   /* -r a-b                                                       */
  int32_t  use_window_filter;    /* -w                G.Use_Window_Filter (forced 0 if erate > 0.06); needs oic_load_qualities */
} oic_params_t;

/* The 8 counters of the -s stats file (OIC/overlapInCore.C:569-588) plus the work counters the
 * rooflines are computed from. */
typedef struct {
  uint64_t kmer_hits_without_olap, kmer_hits_with_olap, kmer_hits_skipped, multi_overlap;
  uint64_t total_overlaps, contained, dovetail;
  uint64_t n_hash_bases, n_inserts, n_distinct;               /* hash build                      */
  uint64_t n_query_bases, n_windows, n_found, n_refs, n_add_ref; /* seed scan                    */
  uint64_t n_runs, n_pairs;                                   /* exact-match runs / read pairs    */
  uint64_t ext_calls, ext_cells, ext_words;                   /* C and W of I_ext = 6C + 4W       */
  uint32_t n_hash_blocks, last_end_hash_id;
  uint64_t bad_short_window, bad_long_window;                 /* -w: "Rejected by short / long window"  */
} oic_stats_t;

/* Accumulated device time per stage, CUDA events on the library's stream (milliseconds) and launch counts. */
typedef struct {
  float    ms_pack, ms_build, ms_scan, ms_group, ms_extend, ms_total_search;
  uint64_t launches;             /* kernels launched by this context since the last reset */
  uint64_t h2d_bytes, d2h_bytes; /* bytes moved across PCIe by this context since reset   */
  /* inside ms_extend: first extension of every pair / speculative extension tasks (the dominant kernel) / wide reruns + commit */
  float    ms_ext_first, ms_ext_tasks, ms_ext_commit, ms_ext_retry;   /* retry = pairs rerun serially with worst-case scratch */
} oic_timings_t;

/* One overlap as written to the .ovb file: 6 x uint32 = a_iid, b_iid, dat0.hi, dat0.lo, dat1.hi, dat1.lo
 * (ST/ovOverlap.H:93-124 with AS_MAX_READLEN_BITS = 21, ST/ovStoreFile.C:196-221). */
typedef struct { uint32_t w[6]; } oic_ovb_record_t;

typedef struct oic_ctx oic_ctx_t;

/* main() set-up (OIC/overlapInCore.C:306-560): validates options, builds Edit_Match_Limit /
 * Error_Bound (LIBO/prefixEditDistance.C:35-97, LIBO/Binomial_Bound.C:119-194) and binds `device`. */
int  oic_create(const oic_params_t *params, int device, oic_ctx_t **out);
void oic_destroy(oic_ctx_t *ctx);
const char *oic_last_error(const oic_ctx_t *ctx);   /* ctx may be NULL: error of the last failed oic_create */

/* Replaces gkStore_loadReadData + the tolower copy loops (OIC/overlapInCore-Build_Hash_Index.C:563-586,
 * OIC/overlapInCore-Process_Overlaps.C:117-134): reads first_id .. first_id+n-1 (1-based IDs), ASCII bases
 * (any case) at bases[offsets[i] .. +lengths[i]); hash_ok / ref_ok are the -H / -R library filters per read
 * (NULL = all pass).  Bases are 2-bit packed on the device (both strands) and stay resident in HBM.
 * Replaces any previously loaded reads. */
int oic_load_reads(oic_ctx_t *ctx, uint32_t first_id, uint32_t n_reads, const uint64_t *offsets,
                   const uint32_t *lengths, const char *bases, const uint8_t *hash_ok, const uint8_t *ref_ok);

/* Same, taking the reads in the gkpStore's own 2-bit encoding instead of ASCII -- the `2SEQ` blob chunk as stored
 * (ST/gkStoreEncode.C:81-114, read by gkRead_loadData ST/gkStore.C:124-126): 4 bases per byte, first base in bits 7-6,
 * a=0 c=1 g=2 t=3, ceil(length/4) bytes per read at packed[byte_offsets[i] ..].  Replaces gkRead_decode2bit + the copy
 * loops: no decode on the host and a quarter of the PCIe bytes.  (Reads that the store keeps in a `USEQ` chunk contain
 * non-ACGT bases and go through oic_load_reads.) */
int oic_load_reads_2bit(oic_ctx_t *ctx, uint32_t first_id, uint32_t n_reads, const uint64_t *byte_offsets,
                        const uint32_t *lengths, const uint8_t *packed, const uint8_t *hash_ok, const uint8_t *ref_ok);

/* -w, the window filter (OIC/overlapInCore.C:398,417-419; OIC/overlapInCore-Process_String_Overlaps.C:365-388,560-620): the
 * quality values of the reads loaded by the last oic_load_reads / oic_load_reads_2bit -- what gkReadData_getQualities gives the
 * reference (ST/gkStore.C:117-145: the read's QVAL constant for every base, or its UQLT bytes), one byte per base, read i at
 * quals[offsets[i] .. + lengths[i]).  Needed before oic_process_overlaps when params.use_window_filter is set; loading new reads
 * drops them.  With the filter on, the pair stage runs its serial path (one warp per read pair for the whole Process_Matches
 * loop): every overlap about to be output is re-aligned and walked along its delta list, which the speculative path does not keep. */
int oic_load_qualities(oic_ctx_t *ctx, const uint64_t *offsets, const uint8_t *quals);

/* In-process sweep over the jobs of overlapInCorePartition (OIC/overlapInCorePartition.C, OverlapInCore.pm:184-233): the
 * pipeline starts one overlapInCore process per (hash range, query range) job, each re-opening the store and rebuilding its
 * hash block.  With the reads resident, a job is just a new set of ranges: oic_set_job replaces -h, -r, --hashstrings and
 * --hashdatalen of the context (every other option stays), after which oic_overlap_driver / oic_build_hash_index /
 * oic_process_overlaps behave exactly as a fresh process started with those options would.  A hash block that was just
 * built for the same (-h sub-range, --hashstrings, --hashdatalen) is reused, so the consecutive query blocks of one hash
 * block cost one build (only after oic_set_job: without it every oic_build_hash_index call builds).  oic_reset_stats
 * zeroes the -s counters between jobs. */
int  oic_set_job(oic_ctx_t *ctx, uint32_t bgn_hash_id, uint32_t end_hash_id, uint32_t bgn_ref_id, uint32_t end_ref_id,
                 uint32_t max_hash_strings, uint64_t max_hash_data_len);
void oic_reset_stats(oic_ctx_t *ctx);

/* Mark_Skip_Kmers input (OIC/overlapInCore-Build_Hash_Index.C:233-286): n k-mers of exactly kmer_len
 * characters, concatenated.  Applied at the end of every subsequent hash build. */
int oic_set_skip_kmers(oic_ctx_t *ctx, uint64_t n, const char *kmers);

/* Build_Hash_Index(gkp, bgnID, endID) (OIC/overlapInCore-Build_Hash_Index.C:443-675): builds the k-mer
 * index of reads bgn_id.. until --hashstrings / --hashdatalen / --hashload trips; *last_id = last read
 * actually loaded (the reference's return value). */
int oic_build_hash_index(oic_ctx_t *ctx, uint32_t bgn_id, uint32_t end_id, uint32_t *last_id);

/* Process_Overlaps over the query reads bgn_ref_id..end_ref_id (inclusive) against the current hash block
 * (OIC/overlapInCore-Process_Overlaps.C:78-181 -> Find_Overlaps -> Process_String_Olaps -> Extend_Alignment
 * -> Output_Overlap).  Every read in the range is processed: the scheduler quirk of the reference's
 * block hand-out (a final block that starts at endRefID is dropped) is applied by oic_overlap_driver,
 * or by the caller when it shards the range.  Overlaps are appended to the context's result buffer. */
int oic_process_overlaps(oic_ctx_t *ctx, uint32_t bgn_ref_id, uint32_t end_ref_id);

/* OverlapDriver() (OIC/overlapInCore.C:190-300): the whole hash-sub-block loop over params' -h / -r
 * ranges including the scheduler quirk for params.num_threads. */
int oic_overlap_driver(oic_ctx_t *ctx);

/* Last query read the reference would process for a -r range and -t (OIC/overlapInCore.C:249,
 * OIC/overlapInCore-Process_Overlaps.C:86); 0 if the range is empty.  Pure host arithmetic. */
uint32_t oic_last_processed_ref(uint32_t bgn_ref_id, uint32_t end_ref_id, uint32_t num_threads);

/* Results (what Output_Overlap pushes to ovFile::writeOverlap, OIC/overlapInCore-Output.C:75-313). */
uint64_t oic_num_overlaps(const oic_ctx_t *ctx);
int  oic_copy_overlaps(oic_ctx_t *ctx, oic_ovb_record_t *dst, uint64_t capacity, uint64_t *n_copied);
void oic_clear_overlaps(oic_ctx_t *ctx);
int  oic_get_stats(const oic_ctx_t *ctx, oic_stats_t *out);
int  oic_get_timings(oic_ctx_t *ctx, oic_timings_t *out);
void oic_reset_timings(oic_ctx_t *ctx);

/* ---- multi-GPU: the immutable product of a hash build, for replication by NCCL broadcast ------------
 * After oic_build_hash_index on the root rank, oic_hash_block_meta/oic_hash_block_buffers expose the
 * device buffers (table, occurrence array, string flags, Bloom filter); non-root ranks call oic_hash_block_reserve with
 * the root's meta to allocate identically sized buffers, the caller broadcasts each buffer in place
 * (ncclBroadcast / torch.distributed.broadcast over NVLink), then oic_hash_block_commit makes the block
 * current.  Reads must already be loaded identically on every rank. */
typedef struct {
  uint64_t table_slots;      /* n_slices << slice_bits                      */
  uint64_t n_occ;            /* occurrence entries                          */
  uint32_t n_strings;        /* string slots in the block                   */
  uint32_t bgn_id, last_id;  /* read IDs covered                            */
  uint32_t has_empty;        /* any screened (Empty) k-mer present          */
  uint64_t n_distinct;
  uint32_t n_slices;         /* table = n_slices independent slices ...     */
  uint32_t slice_bits;       /* ... of 2^slice_bits slots (probing wraps inside a slice) */
  uint64_t filter_words;     /* 32-bit words of the Bloom filter (power of two) */
} oic_hash_meta_t;
typedef struct { void *ptr; uint64_t bytes; } oic_dev_buffer_t;
#define OIC_HASH_BLOCK_NBUF 4   /* table, occurrence array, string flags, filter */
int oic_hash_block_meta(const oic_ctx_t *ctx, oic_hash_meta_t *meta);
int oic_hash_block_reserve(oic_ctx_t *ctx, const oic_hash_meta_t *meta);
int oic_hash_block_buffers(const oic_ctx_t *ctx, oic_dev_buffer_t bufs[OIC_HASH_BLOCK_NBUF]);
int oic_hash_block_commit(oic_ctx_t *ctx, const oic_hash_meta_t *meta);

/* ---- multi-GPU: the collective inside the library (NCCL over NVLink / NVSwitch) -----------------------
 * The reference parallelises with threads over ONE hash table in host memory (OIC/overlapInCore.C:271-273) and with
 * independent grid jobs (OIC/overlapInCorePartition.C:298-349); it has no collective to cite.  Here every GPU needs the hash
 * block in its own HBM.  A communicator is attached to a context either by the multi-process route (rank 0 calls
 * oic_comm_unique_id, the launcher ships the 128 bytes -- torch.distributed, MPI, a file -- and every rank calls
 * oic_comm_init_rank) or, inside ONE process with one context per GPU, by oic_comm_init_all.  NCCL is loaded with dlopen
 * ("libnccl.so.2", reusing a copy the process already holds) at the first of these calls; the library has no link-time
 * dependency on it.
 *
 * oic_hash_block_replicate: rank `root` has built the block (oic_build_hash_index); every rank of the communicator calls
 * this, the meta and the three buffers travel by ncclBroadcast on the context's own stream (no host synchronisation between
 * them), non-root ranks commit the block (occupancy bitmap rebuilt locally).  nccl_comm: a caller-owned ncclComm_t, or NULL
 * for the communicator attached to the context.
 *
 * oic_build_hash_index_sharded: the collective build.  Every rank calls it with the same arguments (the same reads must be
 * resident everywhere); rank r inserts only the k-mers whose home slot lies in slice r of the table (the table is cut into
 * n_ranks slices, linear probing wraps inside a slice, so slices are independent), builds the occurrence lists of its slice,
 * and the slices / list segments are exchanged with ncclAllGather / grouped ncclBroadcast.  Afterwards every rank holds the
 * identical, complete block: the serial build and the broadcast of a mostly empty table leave the critical path.  Falls
 * back to build-on-root + oic_hash_block_replicate when the rank count is not a power of two or the hash-load cut-off
 * trips (the cut needs the global first-occurrence histogram). */
#define OIC_COMM_ID_BYTES 128
int oic_comm_unique_id(uint8_t id[OIC_COMM_ID_BYTES]);
int oic_comm_init_rank(oic_ctx_t *ctx, const uint8_t id[OIC_COMM_ID_BYTES], int n_ranks, int rank);
int oic_comm_init_all(oic_ctx_t *const *ctxs, int n_ctx);
int oic_comm_info(const oic_ctx_t *ctx, int *n_ranks, int *rank);   /* n_ranks = 0: no communicator attached */
int oic_hash_block_replicate(oic_ctx_t *ctx, void *nccl_comm, int root);
int oic_build_hash_index_sharded(oic_ctx_t *ctx, void *nccl_comm, uint32_t bgn_id, uint32_t end_id, uint32_t *last_id);

/* ---- stage dumps for parity tests (not used by the product path) -----------------------------------
 * k-mer index of the current hash block as sorted CSR: keys ascending, occ (string<<32 | offset) DESCENDING
 * per key, flags bit0 = Empty.  Call with NULL pointers to get the sizes. */
int oic_dump_index(oic_ctx_t *ctx, uint64_t *n_keys, uint64_t *n_occ, uint64_t *keys, uint64_t *occ_start,
                   uint64_t *occ, uint8_t *flags);
/* per (query, strand, hash read) accumulators and their seed lists of the LAST oic_process_overlaps call,
 * when enabled with oic_set_debug(ctx, 1).  pair record: q_id, h_id, strand, consistent, diag_ct, diag_bgn,
 * diag_end, n_matches, first_match (9 x int64); match record: offset, len, start (3 x int32). */
void oic_set_debug(oic_ctx_t *ctx, uint32_t mask);
int oic_dump_pairs(oic_ctx_t *ctx, uint64_t *n_pairs, uint64_t *n_matches, int64_t *pairs, int32_t *matches);

/* Canonical K-mer counts of ALL loaded reads -- the step upstream of overlapInCore in the pipeline.
 * Replaces `meryl -B -C -m K -s <gkpStore>` (src/meryl/meryl-build.C: canonical = the lexicographically smaller of a K-mer and
 * its reverse complement; windows with a non-ACGT base are skipped) followed by `meryl -Dt -n min_count` (src/meryl/
 * meryl-dump.C, called at src/pipelines/canu/Meryl.pm:634 to write the `-k` skip list of overlapInCore).  The result stays in
 * the context, ascending by K-mer; n_passes = hash partitions the count was done in (1 unless the table did not fit),
 * device_ms = CUDA-event time of the count.  K 2..31. */
int oic_count_kmers(oic_ctx_t *ctx, uint32_t K, uint32_t min_count, uint64_t *n_found, uint32_t *n_passes, double *device_ms);

/* The K-mers found by oic_count_kmers, in the order `meryl -Dt` prints them.  keys[i] holds the K-mer with its FIRST base in the
 * HIGH bits (a=0 c=1 g=2 t=3, 2 bits per base); counts may be NULL. */
int oic_copy_frequent_kmers(oic_ctx_t *ctx, uint64_t *keys, uint32_t *counts, uint64_t capacity, uint64_t *n_copied);

/* Histogram of the counts of the last oic_count_kmers call: hist[i] = distinct canonical K-mers seen exactly i times, i <
 * *length = largest count + 1 (hist may be NULL to ask for the length).  What `meryl -Dh` prints (src/meryl/meryl-dump.C:120-152;
 * the pipeline feeds it to estimate-mer-threshold, src/pipelines/canu/Meryl.pm:449-470): distinct = sum hist[i], unique = hist[1],
 * total = sum i * hist[i]. */
int oic_copy_kmer_histogram(oic_ctx_t *ctx, uint64_t *hist, uint64_t capacity, uint64_t *length);

/* ---- overlap store construction: the step downstream of overlapInCore -------------------------------
 * Replaces the compute of ovStoreBuild's in-core path (ST/ovStoreBuild.C:500-645): every overlap of the input is filtered and
 * mirrored (ovStoreFilter::filterOverlap, ST/ovStoreFilter.C:118-260; ovOverlap::swapIDs, ST/ovOverlap.C:114-144), and the
 * records that still carry a forUTG / forOBT / forDUP flag are sorted by ovOverlap::operator< (ST/ovOverlap.H:327-339: a_iid,
 * b_iid, then the two 64-bit data words) on the GPU.  recs: n overlaps in .ovb word order (host memory); read_len / skip_dup /
 * library: per read ID, max_id = numReads + 1 entries (skip_dup[i] = 1 when the read's library does not ask for
 * deduplication, ST/ovStoreFilter.C:62-65); max_evalue = AS_OVS_encodeEvalue(-e).  sorted must hold 2 n records; *n_sorted
 * receives the number written.  counts (may be NULL) receives 9 tallies of the filter: records with an ID out of range (the
 * reference exits on the first one: the caller must treat a non-zero count as fatal), skipERATE, skipOBTbad, skipOBTshort,
 * skipDUP, skipDUPlib, saveUTG, saveOBT, saveDUP.  The store files themselves (info, index, data, evalueLen, overlapScores)
 * are host work: canu_b200/host/ovstore.cpp, tool canu_b200/bin/ovStoreBuild. */
int oic_ovlstore_sort(oic_ctx_t *ctx, const oic_ovb_record_t *recs, uint64_t n, uint32_t max_id, const uint32_t *read_len,
                      const uint8_t *skip_dup, const uint8_t *library, uint32_t max_evalue, oic_ovb_record_t *sorted,
                      uint64_t *n_sorted, uint64_t counts[9], double *device_ms);

/* ---- allocation-free variant of oic_ovlstore_sort ---------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same filter + mirror + sort, same kernels and
 * bit-identical output as oic_ovlstore_sort, but every intermediate buffer (mirrored records, two key and two index arrays of
 * the radix sort, digit histograms, counters) lives in ONE file-scope `static __device__` array of the library, and the call
 * makes no cudaMalloc / cudaFree and uses no allocating library routine.  It needs no context.
 *
 * Limit: at most oic_ovlstore_sort_noalloc_max() = 65 536 input overlaps per call (131 072 store records).  The array is
 * 6 357 504 bytes (6.06 MiB) and every GPU that loads the library holds it until the process exits.  A larger n is refused with
 * OIC_ERR_CAPACITY before anything runs (split the input by a_iid range, or use oic_ovlstore_sort).
 *
 * d_recs / d_read_len / d_skip_dup / d_library / d_sorted are DEVICE pointers on `device`, owned by the caller, with the
 * meaning and sizes of the host arrays of oic_ovlstore_sort (d_sorted holds 2 n records).  n_sorted and counts[9] (may be NULL)
 * are HOST pointers.  Work is issued on the legacy default stream of `device`; the call returns after the result is complete.
 * The first call on a device looks the array's address up (cudaGetSymbolAddress, once per device).
 *
 * Every call shares the array, which keeps the last call's contents (the counters are cleared at the start of each call;
 * nothing else is read before it is written).  Calls that overlap on the GPU -- from two threads, or two processes' worth of
 * streams on one device -- RACE on it: serialise them. */
uint64_t oic_ovlstore_sort_noalloc_max(void);
int oic_ovlstore_sort_noalloc(int device, const oic_ovb_record_t *d_recs, uint64_t n, uint32_t max_id, const uint32_t *d_read_len,
                              const uint8_t *d_skip_dup, const uint8_t *d_library, uint32_t max_evalue, oic_ovb_record_t *d_sorted,
                              uint64_t *n_sorted, uint64_t counts[9]);

/* Roofline denominator of the extension stage: INT32 issue rates of `device`, measured (thread Gop/s).  alu = chains of
 * LOP3 + IADD3 (the pipe the edit-distance rows live on; reference code measured against: LIBO/prefixEditDistance-forward.C
 * inner loop), imad = chains of IMAD.  Takes a few milliseconds. */
int oic_measure_int32_peak(int device, double *alu_gops, double *imad_gops);

#ifdef __cplusplus
}
#endif
#endif
