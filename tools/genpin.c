/* genpin -- deterministic synthetic collections for the FULL-SIZE parity pins (test tooling).
 *
 * Same mutation model as tools/gen_synth.py / SURVEY.md 8(d) (reference = i.i.d. uniform ACGT; every
 * copy independently, per position u ~ U(0,1): u < d/3 substitute by another base, d/3 <= u < 2d/3
 * delete, 2d/3 <= u < d keep + insert one uniform base; config 5: `svs` block events of 1-50 kbp per
 * copy first -- 30 % deletion, 30 % tandem duplication, 20 % reversal, 20 % translocation), but with a
 * PRNG defined HERE (splitmix64 -> xoshiro256**, one stream per copy), so the same (ref_bp, copies,
 * div, seed, svs) gives the same bytes in this container (where the unmodified reference binary and
 * the GSA oracle produce the expected checksums, tests/golden/make_pins.py) and on the GPU box (where
 * tests/test_gpu_fullsize.py regenerates the input and checks the CUDA output against them).
 * Copies are independent, so they are generated on several threads.
 *
 * Library:  genpin_reference, genpin_copy, genpin_text          (ctypes, tests/genpin.py)
 * Program:  genpin <ref_bp> <copies> <div> <seed> <svs> <out_prefix>   -> .ref.fa .coll.fa .txt
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

typedef struct { uint64_t s[4]; } rng_t;
static uint64_t splitmix(uint64_t *x) {
  uint64_t z = (*x += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
static void rng_seed(rng_t *r, uint64_t seed, uint64_t stream) {
  uint64_t x = seed * 0xD1342543DE82EF95ull + stream * 0x2545F4914F6CDD1Dull + 0x1234567ull;
  for (int i = 0; i < 4; i++) r->s[i] = splitmix(&x);
}
static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static inline uint64_t rng_next(rng_t *r) {
  uint64_t *s = r->s, res = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
  s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
  return res;
}
static inline uint64_t rng_below(rng_t *r, uint64_t n) {          /* n > 0; bias < 2^-32 for n < 2^32 */
  return (uint64_t)(((__uint128_t)rng_next(r) * n) >> 64);
}
static const uint8_t ACGT[4] = {'A', 'C', 'G', 'T'};
static inline int code_of(uint8_t c) { return c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : 3; }

void genpin_reference(uint64_t ref_bp, uint64_t seed, uint8_t *out) {
  rng_t r;
  rng_seed(&r, seed, 0);
  uint64_t i = 0;
  while (i < ref_bp) {
    uint64_t w = rng_next(&r);
    for (int k = 0; k < 32 && i < ref_bp; k++, w >>= 2) out[i++] = ACGT[w & 3];
  }
}

/* block events on s (length *n, capacity cap); tmp holds at most 50001 bytes */
static void apply_svs(uint8_t *s, uint64_t *n, uint64_t cap, rng_t *r, int count) {
  uint8_t *tmp = (uint8_t *)malloc(50001 + 16);
  for (int e = 0; e < count; e++) {
    uint64_t len = *n, L = 1000 + rng_below(r, 49001);
    if (L >= len / 2) L = len / 4 ? len / 4 : 1;
    uint64_t a = rng_below(r, len - L);
    uint64_t u = rng_below(r, 100);
    if (u < 30) {                                        /* deletion */
      memmove(s + a, s + a + L, len - a - L);
      *n = len - L;
    } else if (u < 60) {                                 /* tandem duplication */
      if (len + L > cap) continue;
      memmove(s + a + 2 * L, s + a + L, len - a - L);
      memcpy(s + a + L, s + a, L);
      *n = len + L;
    } else if (u < 80) {                                 /* reversal */
      for (uint64_t i = 0, j = L - 1; i < j; i++, j--) { uint8_t t = s[a + i]; s[a + i] = s[a + j]; s[a + j] = t; }
    } else {                                             /* translocation */
      memcpy(tmp, s + a, L);
      memmove(s + a, s + a + L, len - a - L);
      uint64_t rest = len - L, b = rng_below(r, rest + 1);
      memmove(s + b + L, s + b, rest - b);
      memcpy(s + b, tmp, L);
    }
  }
  free(tmp);
}

/* copy k of the collection -> out (capacity cap); returns its length (0 on overflow) */
uint64_t genpin_copy(const uint8_t *ref, uint64_t ref_bp, double div, uint64_t seed, uint64_t k, int svs,
                     uint8_t *out, uint64_t cap) {
  rng_t r;
  rng_seed(&r, seed, k + 1);
  const uint8_t *base = ref;
  uint64_t n = ref_bp;
  uint8_t *sv = NULL;
  if (svs > 0) {
    uint64_t c2 = ref_bp + (uint64_t)svs * 50001 + 64;
    sv = (uint8_t *)malloc(c2);
    memcpy(sv, ref, ref_bp);
    apply_svs(sv, &n, c2, &r, svs);
    base = sv;
  }
  const double third = div / 3.0;
  const uint64_t t1 = (uint64_t)(third * 9007199254740992.0), t2 = (uint64_t)(2 * third * 9007199254740992.0),
                 t3 = (uint64_t)(div * 9007199254740992.0);
  uint64_t w = 0;
  for (uint64_t i = 0; i < n; i++) {
    if (w + 2 > cap) { free(sv); return 0; }
    const uint64_t u = rng_next(&r) >> 11;               /* 53 bits */
    const uint8_t c = base[i];
    if (u >= t3) { out[w++] = c; continue; }
    if (u < t1) out[w++] = ACGT[(code_of(c) + 1 + (int)rng_below(&r, 3)) & 3];
    else if (u < t2) { /* deleted */ }
    else { out[w++] = c; out[w++] = ACGT[rng_below(&r, 4)]; }
  }
  free(sv);
  return w;
}

/* ---- whole collection as the parsed text: 0x02, then every copy followed by 0x02 --------------- */
typedef struct {
  const uint8_t *ref; uint64_t ref_bp; double div; uint64_t seed; int svs;
  uint64_t first, count; uint8_t **bufs; uint64_t *lens; uint64_t cap; int tid, nth;
} job_t;
static void *worker(void *p) {
  job_t *j = (job_t *)p;
  for (uint64_t k = j->tid; k < j->count; k += j->nth) {
    j->bufs[k] = (uint8_t *)malloc(j->cap);
    j->lens[k] = j->bufs[k] ? genpin_copy(j->ref, j->ref_bp, j->div, j->seed, j->first + k, j->svs, j->bufs[k], j->cap) : 0;
  }
  return NULL;
}
static uint64_t copy_cap(uint64_t ref_bp, double div, int svs) {
  return (uint64_t)((double)(ref_bp + (uint64_t)svs * 50001) * (1.0 + div) + 4096.0 + 64.0 * 1024);
}
/* Generates copies [first, first + count) in waves of `threads` and hands each to sink in order. */
typedef void (*sink_fn)(void *ctx, uint64_t k, const uint8_t *s, uint64_t len);
static int generate(const uint8_t *ref, uint64_t ref_bp, uint64_t first, uint64_t count, double div, uint64_t seed,
                    int svs, int threads, sink_fn sink, void *ctx) {
  if (threads < 1) threads = 1;
  if (threads > 64) threads = 64;
  const uint64_t cap = copy_cap(ref_bp, div, svs);
  /* wave: enough copies to keep the threads busy without holding the whole collection twice */
  uint64_t wave = (uint64_t)threads * (ref_bp < (1u << 20) ? 64 : 1);
  uint8_t **bufs = (uint8_t **)calloc(wave, sizeof *bufs);
  uint64_t *lens = (uint64_t *)calloc(wave, sizeof *lens);
  int ok = 1;
  for (uint64_t k0 = 0; k0 < count && ok; k0 += wave) {
    uint64_t m = count - k0 < wave ? count - k0 : wave;
    pthread_t th[64];
    job_t jobs[64];
    int nt = threads < (int)m ? threads : (int)m;
    for (int t = 0; t < nt; t++) {
      jobs[t] = (job_t){ref, ref_bp, div, seed, svs, first + k0, m, bufs, lens, cap, t, nt};
      pthread_create(&th[t], NULL, worker, &jobs[t]);
    }
    for (int t = 0; t < nt; t++) pthread_join(th[t], NULL);
    for (uint64_t k = 0; k < m; k++) {
      if (!bufs[k] || (lens[k] == 0 && ref_bp > 0)) ok = 0;
      if (ok) sink(ctx, first + k0 + k, bufs[k], lens[k]);
      free(bufs[k]); bufs[k] = NULL;
    }
  }
  free(bufs); free(lens);
  return ok;
}

typedef struct { uint8_t *out; uint64_t w, cap; int overflow; } text_sink_t;
static void text_sink(void *c, uint64_t k, const uint8_t *s, uint64_t len) {
  text_sink_t *t = (text_sink_t *)c;
  (void)k;
  if (t->w + len + 1 > t->cap) { t->overflow = 1; return; }
  memcpy(t->out + t->w, s, len);
  t->w += len;
  t->out[t->w++] = 2;
}
/* text of copies [first, first + count): a leading 0x02 (the empty document a FASTA file starts with)
 * when first == 0.  Returns N (0 on overflow of cap); *ndocs = documents. */
uint64_t genpin_text(uint64_t ref_bp, uint64_t copies_first, uint64_t copies_count, double div, uint64_t seed,
                     int svs, int threads, uint8_t *ref_out, uint8_t *out, uint64_t cap, uint64_t *ndocs) {
  uint8_t *ref = ref_out ? ref_out : (uint8_t *)malloc(ref_bp + 1);
  genpin_reference(ref_bp, seed, ref);
  text_sink_t t = {out, 0, cap, 0};
  if (copies_first == 0 && cap > 0) out[t.w++] = 2;
  int ok = generate(ref, ref_bp, copies_first, copies_count, div, seed, svs, threads, text_sink, &t);
  if (!ref_out) free(ref);
  if (ndocs) *ndocs = copies_count + (copies_first == 0 ? 1 : 0);
  return ok && !t.overflow ? t.w : 0;
}
uint64_t genpin_text_cap(uint64_t ref_bp, uint64_t copies, double div, int svs) {
  return copies * (copy_cap(ref_bp, div, svs) - 64 * 1024 + 64) + 4096;
}

/* ---- program: FASTA files for the reference binary / the GSA oracle ------------------------------ */
static void fasta(FILE *f, const char *name, uint64_t k, const uint8_t *s, uint64_t len) {
  if (k == (uint64_t)-1) fprintf(f, ">%s\n", name); else fprintf(f, ">%s%llu\n", name, (unsigned long long)k);
  for (uint64_t a = 0; a < len; a += 60) {
    uint64_t m = len - a < 60 ? len - a : 60;
    fwrite(s + a, 1, m, f);
    fputc('\n', f);
  }
}
static void file_sink(void *c, uint64_t k, const uint8_t *s, uint64_t len) { fasta((FILE *)c, "s", k, s, len); }

#ifndef GENPIN_NO_MAIN
int main(int argc, char **argv) {
  if (argc < 7) { fprintf(stderr, "usage: genpin <ref_bp> <copies> <div> <seed> <svs> <out_prefix> [threads]\n"); return 2; }
  uint64_t ref_bp = strtoull(argv[1], 0, 10), copies = strtoull(argv[2], 0, 10), seed = strtoull(argv[4], 0, 10);
  double div = atof(argv[3]);
  int svs = atoi(argv[5]), threads = argc > 7 ? atoi(argv[7]) : 8;
  char path[4096];
  uint8_t *ref = (uint8_t *)malloc(ref_bp + 1);
  genpin_reference(ref_bp, seed, ref);
  snprintf(path, sizeof path, "%s.ref.fa", argv[6]);
  FILE *f = fopen(path, "wb");
  if (!f) return 1;
  fasta(f, "ref", (uint64_t)-1, ref, ref_bp);
  fclose(f);
  snprintf(path, sizeof path, "%s.coll.fa", argv[6]);
  f = fopen(path, "wb");
  if (!f) return 1;
  static char big[1 << 22];
  setvbuf(f, big, _IOFBF, sizeof big);
  int ok = generate(ref, ref_bp, 0, copies, div, seed, svs, threads, file_sink, f);
  fclose(f);
  snprintf(path, sizeof path, "%s.txt", argv[6]);
  f = fopen(path, "w");
  fprintf(f, "%s.ref.fa\n%s.coll.fa\n", argv[6], argv[6]);
  fclose(f);
  free(ref);
  return ok ? 0 : 1;
}
#endif
