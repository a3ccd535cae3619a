/*
 * oracle/oic_oracle.h -- C ABI of the CPU restatement of Canu's overlapInCore.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg may load this.  The product (canu_b200/) never links or calls it.
 *
 * The restatement follows the reference's *behaviour* (file:line cited at each function
 * in oic_oracle.cpp) with its own data structures: a sorted k-mer index instead of the
 * 21-way bucket table, explicit band rows instead of the lazily grown edit array, and
 * vectors instead of arena linked lists.  It is pinned against the reference binary
 * built by oracle/build_ref.sh (tests/test_oracle_vs_ref.py) and emits stage-level
 * artefacts (k-mer index, per-pair seed lists, per-extension tuples) that the black-box
 * reference cannot, which is what the CUDA kernels are checked against.
 */
#ifndef OIC_ORACLE_H
#define OIC_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int32_t  kmer_len;           /* -k                                    */
  double   max_erate;          /* --maxerate, already strtof()-rounded   */
  int32_t  min_olap_len;       /* --minlength                            */
  int32_t  partial;            /* -G                                     */
  int32_t  unique_olap;        /* -u (1, default) / -m (0)               */
  int32_t  use_hopeless_check; /* !-z, forced off when erate > 0.06      */
  uint32_t hash_bits;          /* --hashbits                             */
  double   hash_load;          /* --hashload                             */
  uint32_t max_hash_strings;   /* --hashstrings                          */
  uint64_t max_hash_data_len;  /* --hashdatalen                          */
  uint64_t frag_olap_limit;    /* -l (UINT64_MAX = unlimited)            */
  uint64_t filter_by_kmer_count; /* --minkmers (0 = off)                 */
  uint32_t num_threads;        /* -t; only feeds the scheduler-quirk     */
  uint32_t bgn_hash_id, end_hash_id; /* -h (inclusive, 1-based)          */
  uint32_t bgn_ref_id, end_ref_id;   /* -r                               */
} oracle_params_t;

/* one overlap record, decoded fields (ST/ovOverlap.H:93-124) */
typedef struct {
  uint32_t a_iid, b_iid;
  uint32_t ahg5, ahg3, bhg5, bhg3, span, evalue;
  uint32_t flipped, for_obt, for_dup, for_utg;
} oracle_overlap_t;

/* one seed (exact-match run) of a pair, in final list order (head first) */
typedef struct { int32_t offset, len, start; } oracle_match_t;

/* per (query, strand, hash read) accumulator after the seed scan */
typedef struct {
  uint32_t q_id, h_id;   /* read IDs (1-based) */
  uint32_t strand;       /* 0 forward, 1 reverse */
  uint32_t consistent;
  int32_t  diag_ct, diag_bgn, diag_end;
  uint32_t n_matches;
  uint64_t first_match;  /* index into the matches dump */
} oracle_pair_t;

/* one Extend_Alignment call */
typedef struct {
  uint32_t q_id, h_id, strand;
  int32_t  m_offset, m_len, m_start;          /* the seed that was extended */
  int32_t  s_lo, s_hi, t_lo, t_hi, errors;    /* result */
  int32_t  kind;                              /* 0 NONE 1 LEFT_BRANCH 2 RIGHT_BRANCH 3 DOVETAIL */
  int32_t  delta_len;
  int32_t  delta_hash;                        /* FNV-style hash of the spliced delta list */
} oracle_ext_t;

typedef struct {
  uint64_t kmer_hits_without_olap, kmer_hits_with_olap, kmer_hits_skipped, multi_overlap;
  uint64_t total_overlaps, contained, dovetail;
  /* work counters for the rooflines */
  uint64_t n_hash_bases, n_inserts, n_distinct;
  uint64_t n_query_bases, n_windows, n_found, n_refs, n_add_ref;
  uint64_t ext_calls, ext_cells, ext_words; /* C and W of I_ext = 6C + 4W */
  uint32_t n_hash_blocks;
  uint32_t last_end_hash_id;
  uint64_t bad_short_window, bad_long_window;   /* -w: overlaps rejected by the 50- / 100-base window (appended: older fields keep their offsets) */
} oracle_stats_t;

typedef struct oracle_ctx oracle_ctx_t;

/* reads: 1-based IDs 1..n_reads; bases is ASCII (any case), read i at bases[offsets[i-1] .. +lengths[i-1]) .
 * hash_ok / ref_ok: optional per-read 0/1 library filters (NULL = all ok). */
oracle_ctx_t *oracle_create(const oracle_params_t *p, uint32_t n_reads, const uint64_t *offsets,
                            const uint32_t *lengths, const char *bases,
                            const uint8_t *hash_ok, const uint8_t *ref_ok);
/* skip k-mers: n strings of exactly kmer_len chars, concatenated (OIC/...Build_Hash_Index.C:233-286) */
void oracle_set_skip_kmers(oracle_ctx_t *c, uint64_t n, const char *kmers);
/* -w, the window filter (OIC/overlapInCore.C:398,417-419; Process_String_Overlaps.C:365-388,560-620): quals holds one quality value per
 * base (0..93, what gkReadData_getQualities returns: a QVAL constant or the UQLT bytes), read i at quals[offsets[i-1] ..] (the offsets
 * given to oracle_create).  Calling this turns the filter on (it stays off when --maxerate > 0.06, as in main()). */
void oracle_set_qualities(oracle_ctx_t *c, const uint64_t *offsets, const uint8_t *quals);
/* which stage dumps to keep (bit 0 pairs+matches, bit 1 extensions, bit 2 k-mer index of the LAST hash block) */
void oracle_set_dump(oracle_ctx_t *c, uint32_t mask);
/* run the whole OverlapDriver loop (OIC/overlapInCore.C:190-300); returns 0 */
int oracle_run(oracle_ctx_t *c);

uint64_t oracle_num_overlaps(const oracle_ctx_t *c);
const oracle_overlap_t *oracle_overlaps(const oracle_ctx_t *c);
const oracle_stats_t *oracle_stats(const oracle_ctx_t *c);
uint64_t oracle_num_pairs(const oracle_ctx_t *c);
const oracle_pair_t *oracle_pairs(const oracle_ctx_t *c);
uint64_t oracle_num_matches(const oracle_ctx_t *c);
const oracle_match_t *oracle_matches(const oracle_ctx_t *c);
uint64_t oracle_num_exts(const oracle_ctx_t *c);
const oracle_ext_t *oracle_exts(const oracle_ctx_t *c);
/* k-mer index of the last hash block: keys ascending; occ[occ_start[i] .. occ_start[i+1]) are
 * (string_num << 32 | offset) in DESCENDING order; flags bit0 = Empty (screened) */
uint64_t oracle_index_num_keys(const oracle_ctx_t *c);
const uint64_t *oracle_index_keys(const oracle_ctx_t *c);
const uint64_t *oracle_index_occ_start(const oracle_ctx_t *c);
const uint64_t *oracle_index_occ(const oracle_ctx_t *c);
const uint8_t *oracle_index_flags(const oracle_ctx_t *c);
void oracle_destroy(oracle_ctx_t *c);

/* host tables (LIBO/prefixEditDistance.C:35-97, LIBO/Binomial_Bound.C:47-206) */
int32_t oracle_max_errors(double erate);
void oracle_match_limit(double erate, int32_t max_errors, int32_t *out /* [max_errors+1] */);
int32_t oracle_error_bound(double erate, int32_t len);

/* stand-alone extension on two ASCII strings (for kernel unit tests):
 * returns kind; out[0..4] = s_lo,s_hi,t_lo,t_hi,errors; delta written to delta[] (cap entries), *delta_len set */
int32_t oracle_extend(double erate, int32_t partial, const char *S, int32_t s_len, const char *T, int32_t t_len,
                      int32_t m_start, int32_t m_offset, int32_t m_len, int32_t *out, int32_t *delta,
                      int32_t delta_cap, int32_t *delta_len, uint64_t *cells, uint64_t *words);

#ifdef __cplusplus
}
#endif
#endif
