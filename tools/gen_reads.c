/*
 * Seeded synthetic long/short-read generator (harness code, not on the product path).
 *
 * One generator serves tests, bench.py and the reference-binary parity runs so that the
 * CPU reference and the GPU path see byte-identical reads.
 *
 *   genome : splitmix64 stream, 2 bits per base (uniform ACGT); optional repeat families
 *   read i : own splitmix64 stream seeded from (seed, i) -> length uniform [lmin,lmax],
 *            start uniform, strand 50/50, per-base independent errors split
 *            sub:ins:del = 5:3:2 (or substitutions only)
 *
 * Built both as a shared library (ctypes, in-memory buffers) and as a CLI writing FASTA:
 *   gcc -O2 -fopenmp -shared -fPIC tools/gen_reads.c -o tools/libgenreads.so
 *   gcc -O2 -fopenmp -DGEN_MAIN   tools/gen_reads.c -o tools/gen_reads
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static inline uint64_t splitmix64(uint64_t *s) {
  uint64_t z = (*s += 0x9E3779B97F4A7C15ULL);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
static inline double u01(uint64_t *s) { return (double)(splitmix64(s) >> 11) * (1.0 / 9007199254740992.0); }

typedef struct {
  uint64_t genome_len;
  double   coverage;
  uint32_t len_min, len_max;
  double   err;        /* per-base error probability applied to each read */
  int      sub_only;   /* 1: substitutions only (Illumina-like) */
  uint64_t seed;
  /* repeat structure (config 5): fraction of genome overwritten by copies of repeat units */
  double   repeat_frac;     /* 0 = none */
  uint32_t repeat_families; /* number of families */
  uint32_t repeat_min, repeat_max; /* unit length range */
  double   repeat_div;      /* per-copy divergence */
} gen_params_t;

static const char ACGT[4] = {'A', 'C', 'G', 'T'};

/* genome as one byte per base, values 0..3 */
uint8_t *oic_gen_genome(const gen_params_t *p) {
  uint64_t G = p->genome_len;
  uint8_t *g = (uint8_t *)malloc(G + 64);
  if (!g) return NULL;
#pragma omp parallel for schedule(static)
  for (int64_t blk = 0; blk < (int64_t)((G + 31) / 32); blk++) {
    uint64_t s = p->seed * 0xD1342543DE82EF95ULL + (uint64_t)blk * 0x2545F4914F6CDD1DULL + 12345;
    uint64_t w = splitmix64(&s);
    for (int j = 0; j < 32; j++) {
      uint64_t pos = (uint64_t)blk * 32 + j;
      if (pos < G) g[pos] = (w >> (2 * j)) & 3;
    }
  }
  if (p->repeat_frac > 0 && p->repeat_families > 0) {
    /* serial, deterministic: paste diverged copies of each family until the budget is used */
    uint64_t s = p->seed ^ 0xABCDEF0123456789ULL;
    uint64_t budget = (uint64_t)(p->repeat_frac * (double)G);
    uint64_t used = 0;
    uint32_t F = p->repeat_families;
    uint8_t **unit = (uint8_t **)malloc(sizeof(uint8_t *) * F);
    uint32_t *ulen = (uint32_t *)malloc(sizeof(uint32_t) * F);
    for (uint32_t f = 0; f < F; f++) {
      ulen[f] = p->repeat_min + (uint32_t)(u01(&s) * (p->repeat_max - p->repeat_min + 1));
      unit[f] = (uint8_t *)malloc(ulen[f]);
      for (uint32_t i = 0; i < ulen[f]; i++) unit[f][i] = splitmix64(&s) & 3;
    }
    while (used < budget) {
      uint32_t f = (uint32_t)(u01(&s) * F);
      if (f >= F) f = F - 1;
      if (ulen[f] >= G) break;
      uint64_t at = (uint64_t)(u01(&s) * (double)(G - ulen[f]));
      for (uint32_t i = 0; i < ulen[f]; i++) {
        uint8_t b = unit[f][i];
        if (u01(&s) < p->repeat_div) b = (b + 1 + (splitmix64(&s) % 3)) & 3;
        g[at + i] = b;
      }
      used += ulen[f];
    }
    for (uint32_t f = 0; f < F; f++) free(unit[f]);
    free(unit); free(ulen);
  }
  return g;
}

uint64_t oic_gen_num_reads(const gen_params_t *p) {
  double mean = 0.5 * ((double)p->len_min + (double)p->len_max);
  return (uint64_t)((double)p->genome_len * p->coverage / mean + 0.5);
}

/* Generate read i. If out==NULL only the length is computed. Returns read length. */
static uint32_t gen_one(const gen_params_t *p, const uint8_t *g, uint64_t i, char *out) {
  uint64_t s = (p->seed + 0x632BE59BD9B4E019ULL) * 0x9E3779B97F4A7C15ULL + i * 0xD6E8FEB86659FD93ULL;
  splitmix64(&s);
  uint32_t len = p->len_min + (uint32_t)(u01(&s) * (double)(p->len_max - p->len_min + 1));
  if (len > p->len_max) len = p->len_max;
  if (len > p->genome_len) len = (uint32_t)p->genome_len;
  uint64_t start = (uint64_t)(u01(&s) * (double)(p->genome_len - len + 1));
  if (start + len > p->genome_len) start = p->genome_len - len;
  int rev = (splitmix64(&s) >> 17) & 1;
  uint32_t n = 0;
  for (uint32_t k = 0; k < len; k++) {
    uint8_t b = rev ? (uint8_t)(3 - g[start + len - 1 - k]) : g[start + k];
    double u = u01(&s);
    if (u < p->err) {
      double t = p->sub_only ? 0.0 : u / p->err; /* uniform in [0,1) given an error */
      if (t < 0.5) { /* substitution */
        b = (b + 1 + (splitmix64(&s) % 3)) & 3;
        if (out) out[n] = ACGT[b];
        n++;
      } else if (t < 0.8) { /* insertion: extra random base, then the true base */
        uint8_t ib = splitmix64(&s) & 3; /* drawn in both passes: the length pass must consume the same stream */
        if (out) out[n] = ACGT[ib];
        n++;
        if (out) out[n] = ACGT[b];
        n++;
      } else { /* deletion */
      }
    } else {
      if (out) out[n] = ACGT[b];
      n++;
    }
  }
  return n;
}

/* lengths[nreads]; returns total bases */
uint64_t oic_gen_lengths(const gen_params_t *p, const uint8_t *g, uint64_t nreads, uint32_t *lengths) {
  uint64_t total = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : total)
  for (int64_t i = 0; i < (int64_t)nreads; i++) {
    lengths[i] = gen_one(p, g, (uint64_t)i, NULL);
    total += lengths[i];
  }
  return total;
}

/* bases: ASCII, concatenated, offsets[i] = start of read i (no terminators) */
void oic_gen_fill(const gen_params_t *p, const uint8_t *g, uint64_t nreads, const uint64_t *offsets, char *bases) {
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t i = 0; i < (int64_t)nreads; i++) gen_one(p, g, (uint64_t)i, bases + offsets[i]);
}

void oic_gen_free(void *p) { free(p); }

/* K-mer counter of a read set -- CPU restatement (test infrastructure) of what the pipeline gets from the reference's meryl:
 * `meryl -B -C -m K -s store` counts CANONICAL K-mers (the lexicographically smaller of a K-mer and its reverse complement,
 * meryl-build.C) over every read, skipping windows with a non-ACGT base, and `meryl -Dt -n T` (meryl-dump.C, Meryl.pm:634) lists
 * those seen >= T times as ">count\nKMER\n" in ascending K-mer order.  Keys here are the K-mer with its FIRST base in the HIGH
 * bits (a=0 c=1 g=2 t=3), so ascending key order is that order.  Open-addressing table, OpenMP over reads.  Returns the number of
 * K-mers with count >= threshold (keys/counts written UNSORTED up to `cap`), or (uint64_t)-1 on allocation failure. */
static inline uint64_t fk_mix(uint64_t k) { k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33; return k; }
uint64_t oic_frequent_kmers(const char *bases, const uint64_t *offsets, const uint32_t *lengths, uint64_t nreads, int K, uint32_t threshold,
                            uint64_t *out_keys, uint32_t *out_counts, uint64_t cap) {
  uint64_t windows = 0;
  for (uint64_t i = 0; i < nreads; i++) if (lengths[i] >= (uint32_t)K) windows += lengths[i] - K + 1;
  uint64_t slots = 1024;
  while (slots < 2 * windows) slots <<= 1;
  const uint64_t mask = slots - 1, kmask = (K == 32) ? ~0ULL : ((1ULL << (2 * K)) - 1);
  uint64_t *keys = (uint64_t *)malloc(slots * sizeof(uint64_t));
  uint32_t *cnt = (uint32_t *)calloc(slots, sizeof(uint32_t));
  if (!keys || !cnt) { free(keys); free(cnt); return (uint64_t)-1; }
  memset(keys, 0xff, slots * sizeof(uint64_t));
#pragma omp parallel for schedule(dynamic, 64)
  for (uint64_t i = 0; i < nreads; i++) {
    const char *s = bases + offsets[i];
    const uint32_t len = lengths[i];
    uint64_t fw = 0, rv = 0;
    uint32_t good = 0;
    for (uint32_t j = 0; j < len; j++) {
      uint64_t c;
      switch (s[j]) {
        case 'a': case 'A': c = 0; break;
        case 'c': case 'C': c = 1; break;
        case 'g': case 'G': c = 2; break;
        case 't': case 'T': c = 3; break;
        default: c = 4; break;
      }
      if (c == 4) { good = 0; fw = rv = 0; continue; }
      fw = ((fw << 2) | c) & kmask;                    /* the window, first base in the high bits */
      rv = (rv >> 2) | ((3 - c) << (2 * (K - 1)));     /* its reverse complement, same convention */
      if (++good < (uint32_t)K) continue;
      const uint64_t key = fw < rv ? fw : rv;
      uint64_t h = fk_mix(key) & mask;
      for (;;) {
        uint64_t cur = __atomic_load_n(&keys[h], __ATOMIC_RELAXED);
        if (cur == ~0ULL) {
          uint64_t expect = ~0ULL;
          if (__atomic_compare_exchange_n(&keys[h], &expect, key, 0, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) cur = key; else cur = expect;
        }
        if (cur == key) { __atomic_fetch_add(&cnt[h], 1u, __ATOMIC_RELAXED); break; }
        h = (h + 1) & mask;
      }
    }
  }
  uint64_t n = 0;
  for (uint64_t h = 0; h < slots; h++)
    if (cnt[h] >= threshold && cnt[h] > 0) { if (n < cap) { out_keys[n] = keys[h]; if (out_counts) out_counts[n] = cnt[h]; } n++; }
  free(keys); free(cnt);
  return n;
}

/* ASCII reads -> the gkpStore 2-bit encoding (4 bases per byte, first base in bits 7-6, a=0 c=1 g=2 t=3, every read on a
 * byte boundary, ceil(len/4) bytes per read) -- what gatekeeperCreate writes into the `2SEQ` chunk of a read blob
 * (reference: src/stores/gkStoreEncode.C:45-78).  Returns the number of bytes written, or (uint64_t)-1 on a non-ACGT base. */
uint64_t oic_encode_2bit(const char *bases, const uint64_t *offsets, const uint32_t *lengths, uint64_t nreads, uint8_t *out,
                         uint64_t *byte_offsets) {
  uint64_t w = 0;
  int bad = 0;
  for (uint64_t i = 0; i < nreads; i++) { byte_offsets[i] = w; w += ((uint64_t)lengths[i] + 3) / 4; }
#pragma omp parallel for schedule(dynamic, 64) reduction(|:bad)
  for (uint64_t i = 0; i < nreads; i++) {
    const char *s = bases + offsets[i];
    uint8_t *o = out + byte_offsets[i];
    const uint32_t len = lengths[i];
    for (uint32_t j = 0; j < len; j += 4) {
      uint8_t b = 0;
      for (uint32_t k = 0; k < 4; k++) {
        uint8_t c = 0;
        if (j + k < len) {
          switch (s[j + k]) {
            case 'a': case 'A': c = 0; break;
            case 'c': case 'C': c = 1; break;
            case 'g': case 'G': c = 2; break;
            case 't': case 'T': c = 3; break;
            default: bad = 1; break;
          }
        }
        b = (uint8_t)((b << 2) | c);
      }
      o[j / 4] = b;
    }
  }
  return bad ? (uint64_t)-1 : w;
}

#ifdef GEN_MAIN
int main(int argc, char **argv) {
  gen_params_t p;
  memset(&p, 0, sizeof(p));
  p.genome_len = 100000; p.coverage = 10; p.len_min = 1000; p.len_max = 3000; p.err = 0.01; p.seed = 1;
  p.repeat_min = 1000; p.repeat_max = 6000; p.repeat_div = 0.02;
  const char *out = NULL;
  int fastq = 0;
  for (int a = 1; a < argc; a++) {
    if (!strcmp(argv[a], "--genome-len")) p.genome_len = strtoull(argv[++a], 0, 10);
    else if (!strcmp(argv[a], "--coverage")) p.coverage = atof(argv[++a]);
    else if (!strcmp(argv[a], "--len-min")) p.len_min = atoi(argv[++a]);
    else if (!strcmp(argv[a], "--len-max")) p.len_max = atoi(argv[++a]);
    else if (!strcmp(argv[a], "--err")) p.err = atof(argv[++a]);
    else if (!strcmp(argv[a], "--sub-only")) p.sub_only = 1;
    else if (!strcmp(argv[a], "--seed")) p.seed = strtoull(argv[++a], 0, 10);
    else if (!strcmp(argv[a], "--repeat-frac")) p.repeat_frac = atof(argv[++a]);
    else if (!strcmp(argv[a], "--repeat-families")) p.repeat_families = atoi(argv[++a]);
    else if (!strcmp(argv[a], "--repeat-div")) p.repeat_div = atof(argv[++a]);
    else if (!strcmp(argv[a], "--fastq")) fastq = 1;
    else if (!strcmp(argv[a], "--out")) out = argv[++a];
    else { fprintf(stderr, "unknown option %s\n", argv[a]); return 1; }
  }
  if (!out) { fprintf(stderr, "usage: gen_reads --genome-len G --coverage X --len-min a --len-max b --err e --seed s [--sub-only] [--fastq] --out reads.fasta\n"); return 1; }
  uint8_t *g = oic_gen_genome(&p);
  uint64_t n = oic_gen_num_reads(&p);
  uint32_t *len = (uint32_t *)malloc(sizeof(uint32_t) * n);
  uint64_t total = oic_gen_lengths(&p, g, n, len);
  uint64_t *off = (uint64_t *)malloc(sizeof(uint64_t) * (n + 1));
  off[0] = 0;
  for (uint64_t i = 0; i < n; i++) off[i + 1] = off[i] + len[i];
  char *bases = (char *)malloc(total + 1);
  oic_gen_fill(&p, g, n, off, bases);
  FILE *F = fopen(out, "w");
  if (!F) { perror(out); return 1; }
  char *q = NULL;
  if (fastq) { q = (char *)malloc(p.len_max * 2 + 16); memset(q, 'I', p.len_max * 2 + 16); }
  for (uint64_t i = 0; i < n; i++) {
    fprintf(F, "%cread%llu\n", fastq ? '@' : '>', (unsigned long long)(i + 1));
    fwrite(bases + off[i], 1, len[i], F);
    fputc('\n', F);
    if (fastq) { fputs("+\n", F); fwrite(q, 1, len[i], F); fputc('\n', F); }
  }
  fclose(F);
  fprintf(stderr, "gen_reads: %llu reads, %llu bases -> %s\n", (unsigned long long)n, (unsigned long long)total, out);
  return 0;
}
#endif
