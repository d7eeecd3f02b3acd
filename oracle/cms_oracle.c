/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the CMS-BWT hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this file's shared object.  The product (cms_bwt_b200/) never does; it fails loudly when
 * its CUDA library is missing.
 *
 * Every function cites the reference lines it restates (paths relative to /root/reference;
 * a bare :N means CMS-BWT-functions.cpp:N).  The restatement keeps the reference's algorithm
 * (SA-interval matching statistics with suffix-link contraction and the PLCP' skip loop, head
 * linking, per-position ordered keys, naming, suffix sort of the name string, next-rank
 * substitution, counterSmallerThanHead, bucket inversion, writers) but is plain C:
 *   - std::map<MatchInSet,ItemMatchInSet> per position  ->  one array of key groups sorted by
 *     (pos, MatchInSet order), with per-position [first,last) ranges;
 *   - libsais / libsais_int  ->  a small prefix-doubling suffix sorter (sa_doubling below): the
 *     suffix array of a string is unique, so any correct sorter gives the same result;
 *   - rmq_tree::psv / nsv  ->  plain scans with the same return convention;
 *   - nextHead is uint32 here (the reference stores j-i in a uint16, :1128, which wraps when
 *     more than 65535 heads share one match end);
 *   - the RLE writer has the remainder flush AFTER the sub-bucket loop, i.e. the maximal-run RLE
 *     of the .bwt (the as-shipped block at :1069-1078 / :1761-1770 is a
 *     defect, not a target).
 *
 * Parity pin: the reference has no tests or golden vectors.  This file is pinned
 * against (a) the unmodified reference binary built by oracle/Makefile into oracle/_ref/cmsbwt,
 * (b) the independent GSA oracle oracle/_ref/gsabwt, on the fuzz + config cases of
 * tests/test_oracle.py, and (c) the committed fixtures in tests/golden/ (generated from (a) by
 * tests/golden/make_golden.py).
 */
#define _GNU_SOURCE
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

typedef struct {
  uint32_t n;        /* augmented reference length (_n, CMS-BWT.h:45) */
  uint8_t *x;        /* augmented reference (_x) */
  int32_t *SA;       /* _SA */
  uint32_t *ISA;     /* _ISA  (:3-9) */
  int32_t *LCP;      /* _LCP, n+1 entries, LCP[n] = -1 (:260-261,271) */
  int32_t *PLCPp;    /* _PLCP after :272-274: max(LCP[ISA[i]], LCP[ISA[i]+1]) */
  uint8_t *BWT;      /* _BWT (:294-297) */
} cms_oracle;

/* ------------------------------------------------------------------------------------------ */
/* generic prefix-doubling suffix sorter over an int32 string (replaces libsais :247, and      */
/* libsais_int :648/:1456).  O(m log^2 m) worst case; inputs here are test-sized.              */

typedef struct { const int64_t *key; } cmp_ctx;
static int cmp_by_key(const void *a, const void *b, void *c) {
  const int64_t *k = ((cmp_ctx *)c)->key;
  int64_t ka = k[*(const int32_t *)a], kb = k[*(const int32_t *)b];
  return (ka > kb) - (ka < kb);
}

/* T[0..m) with values in [0,k); writes SA[0..m).  Suffix order: shorter prefix-equal suffix is
 * smaller (standard).  Returns 0. */
static int sa_doubling(const int32_t *T, int32_t *SA, int32_t m, int32_t k) {
  if (m == 0) return 0;
  int32_t *rank = (int32_t *)malloc(sizeof(int32_t) * (size_t)m);
  int32_t *tmp = (int32_t *)malloc(sizeof(int32_t) * (size_t)m);
  int64_t *key = (int64_t *)malloc(sizeof(int64_t) * (size_t)m);
  /* round 0: counting sort by first symbol */
  int64_t *cnt = (int64_t *)calloc((size_t)k + 1, sizeof(int64_t));
  for (int32_t i = 0; i < m; i++) cnt[T[i] + 1]++;
  for (int32_t c = 0; c < k; c++) cnt[c + 1] += cnt[c];
  for (int32_t i = 0; i < m; i++) SA[cnt[T[i]]++] = i;
  free(cnt);
  rank[SA[0]] = 0;
  for (int32_t i = 1; i < m; i++) rank[SA[i]] = rank[SA[i - 1]] + (T[SA[i]] != T[SA[i - 1]]);
  for (int64_t h = 1; rank[SA[m - 1]] != m - 1; h <<= 1) {
    for (int32_t i = 0; i < m; i++) {
      int64_t r2 = (i + h < m) ? (int64_t)rank[i + h] + 1 : 0;
      key[i] = ((int64_t)rank[i] << 32) | r2;
    }
    /* sort each group of equal rank by the second half of the key */
    cmp_ctx ctx = {key};
    int32_t s = 0;
    while (s < m) {
      int32_t e = s + 1;
      int32_t r = rank[SA[s]];
      while (e < m && rank[SA[e]] == r) e++;
      if (e - s > 1) qsort_r(SA + s, (size_t)(e - s), sizeof(int32_t), cmp_by_key, &ctx);
      s = e;
    }
    tmp[SA[0]] = 0;
    for (int32_t i = 1; i < m; i++) tmp[SA[i]] = tmp[SA[i - 1]] + (key[SA[i]] != key[SA[i - 1]]);
    memcpy(rank, tmp, sizeof(int32_t) * (size_t)m);
  }
  free(rank); free(tmp); free(key);
  return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* reference loader pieces                                                                     */

/* :208-213 (trim one trailing \n, \r or \0, then one '$') and :231-237 (append every byte
 * 3..127 that does not occur, then 0x01, 0x00).  out must hold len+127 bytes.  Returns the
 * augmented length. */
int64_t cms_oracle_augment(const uint8_t *ref, int64_t len, uint8_t *out) {
  if (len > 0 && (ref[len - 1] == '\n' || ref[len - 1] == '\r' || ref[len - 1] == 0)) len--;
  if (len > 0 && ref[len - 1] == '$') len--;
  uint8_t present[256];
  memset(present, 0, sizeof present);
  for (int64_t i = 0; i < len; i++) present[ref[i]] = 1;
  memcpy(out, ref, (size_t)len);
  int64_t n = len;
  /* `_x.find((char)i)` runs on the growing string, but appended symbols are the ones that were
   * absent, so membership never changes for later i. */
  for (int c = 3; c < 128; c++)
    if (!present[c]) out[n++] = (uint8_t)c;
  out[n++] = 1;
  out[n++] = 0;
  return n;
}

/* :171-189: FASTA reference = concatenation of all non-header, non-empty lines read with
 * getline while good() (an unterminated last line is dropped); :190-199 raw bytes otherwise.
 * out must hold `size` bytes.  Returns the sequence length (before trimming). */
int64_t cms_oracle_read_reference(const uint8_t *file, int64_t size, uint8_t *out) {
  if (size > 0 && file[0] == '>') {
    int64_t o = 0, p = 0;
    while (p < size) {
      const uint8_t *nl = (const uint8_t *)memchr(file + p, '\n', (size_t)(size - p));
      if (!nl) break;                      /* getline hits EOF -> !good() -> line dropped */
      int64_t e = nl - file;
      if (e > p && file[p] != '>') { memcpy(out + o, file + p, (size_t)(e - p)); o += e - p; }
      p = e + 1;
    }
    return o;
  }
  memcpy(out, file, (size_t)size);
  return size;
}

/* Collection reader, :1138-1140,1257-1266,1269-1275 (small path :344-346,464-482 is identical).
 * Produces the concatenated text: every document followed by the separator 0x02.
 *   - getline while good(): an unterminated last line is dropped;
 *   - an empty line or one starting with '>' closes the current document (so a file starting
 *     with '>' yields a leading empty document);
 *   - sn = min(file size, prefix); once charactersRead >= sn-1 the line is cut so that the
 *     total reaches sn-1 symbols and reading stops;
 *   - a non-empty remainder is one more document.
 * out_text must hold size+2 bytes; doc_start must hold size/1+2 entries at worst (callers pass
 * size+2).  Returns N (text length incl. terminators); *ndocs = number of documents. */
int64_t cms_oracle_parse_collection(const uint8_t *file, int64_t size, uint64_t prefix,
                                    uint8_t *out_text, int64_t *doc_start, int64_t *ndocs) {
  uint64_t sn = (uint64_t)size < prefix ? (uint64_t)size : prefix;
  uint64_t charactersRead = 0;
  int64_t o = 0, p = 0, nd = 0, cur0 = 0;
  int stop = 0;
  while (p < size && !stop) {
    const uint8_t *nl = (const uint8_t *)memchr(file + p, '\n', (size_t)(size - p));
    if (!nl) break;
    int64_t e = nl - file, L = e - p;
    if (L == 0 || file[p] == '>') {
      out_text[o++] = 2;
      charactersRead++;
      doc_start[nd++] = cur0;
      cur0 = o;
    } else {
      charactersRead += (uint64_t)L;
      if (charactersRead >= sn - 1) {
        /* line.substr(0, line.size() - (charactersRead - _sn) - 1), unsigned wrap-around */
        uint64_t keep = (uint64_t)L - (charactersRead - sn) - 1;
        if (keep > (uint64_t)L) keep = (uint64_t)L;     /* substr clamps */
        memcpy(out_text + o, file + p, (size_t)keep);
        o += (int64_t)keep;
        stop = 1;
      } else {
        memcpy(out_text + o, file + p, (size_t)L);
        o += L;
      }
    }
    p = e + 1;
  }
  if (o > cur0) {
    out_text[o++] = 2;
    doc_start[nd++] = cur0;
  }
  doc_start[nd] = o;
  *ndocs = nd;
  return o;
}

/* ------------------------------------------------------------------------------------------ */
/* index (initialize_reference :246-297)                                                        */

void cms_oracle_destroy(cms_oracle *o) {
  if (!o) return;
  free(o->x); free(o->SA); free(o->ISA); free(o->LCP); free(o->PLCPp); free(o->BWT); free(o);
}

/* x = augmented reference; sa may be NULL (then it is computed here). */
cms_oracle *cms_oracle_create(const uint8_t *x, uint32_t n, const int32_t *sa) {
  cms_oracle *o = (cms_oracle *)calloc(1, sizeof *o);
  o->n = n;
  o->x = (uint8_t *)malloc(n);
  memcpy(o->x, x, n);
  o->SA = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
  if (sa) memcpy(o->SA, sa, sizeof(int32_t) * (size_t)n);
  else {
    int32_t *T = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
    for (uint32_t i = 0; i < n; i++) T[i] = x[i];
    sa_doubling(T, o->SA, (int32_t)n, 256);         /* :247 libsais */
    free(T);
  }
  o->ISA = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)n);
  for (uint32_t i = 0; i < n; i++) o->ISA[o->SA[i]] = i;      /* constructISA :3-9 */
  /* :270-271 libsais_plcp + libsais_lcp: PLCP[i] = lcp(x[i..], x[phi(i)..]), LCP[r] = PLCP[SA[r]],
   * LCP[0] = 0; computed here with the Kasai recurrence PLCP[i] >= PLCP[i-1]-1. */
  int32_t *plcp = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
  o->LCP = (int32_t *)malloc(sizeof(int32_t) * ((size_t)n + 1));
  int32_t l = 0;
  for (uint32_t i = 0; i < n; i++) {
    uint32_t r = o->ISA[i];
    if (r == 0) { plcp[i] = 0; l = 0; continue; }
    uint32_t j = (uint32_t)o->SA[r - 1];
    while (i + l < n && j + l < n && x[i + l] == x[j + l]) l++;
    plcp[i] = l;
    if (l > 0) l--;
  }
  for (uint32_t r = 0; r < n; r++) o->LCP[r] = plcp[o->SA[r]];
  o->LCP[n] = -1;                                                  /* :261 */
  o->PLCPp = plcp;
  for (uint32_t i = 0; i < n; i++) {                                /* :272-274 */
    int32_t a = o->LCP[o->ISA[i]], b = o->LCP[o->ISA[i] + 1];
    o->PLCPp[i] = a > b ? a : b;
  }
  o->BWT = (uint8_t *)malloc(n);
  for (uint32_t i = 0; i < n; i++) o->BWT[i] = o->SA[i] > 0 ? x[o->SA[i] - 1] : 0;   /* :294-297 */
  return o;
}

uint32_t cms_oracle_n(const cms_oracle *o) { return o->n; }
const int32_t *cms_oracle_sa(const cms_oracle *o) { return o->SA; }
const uint32_t *cms_oracle_isa(const cms_oracle *o) { return o->ISA; }
const int32_t *cms_oracle_lcp(const cms_oracle *o) { return o->LCP; }
const int32_t *cms_oracle_plcp_prime(const cms_oracle *o) { return o->PLCPp; }
const uint8_t *cms_oracle_ref_bwt(const cms_oracle *o) { return o->BWT; }

/* rmq_tree.h:184-205: largest j <= i with LCP[j] < ub, -1 if none (plain scan). */
static int32_t psv(const cms_oracle *o, int32_t i, int32_t ub) {
  while (i >= 0 && o->LCP[i] >= ub) --i;
  return i >= 0 ? i : -1;
}
/* rmq_tree.h:208-225: smallest j >= i (j < n) with LCP[j] < ub, -1 if none. */
static int32_t nsv(const cms_oracle *o, int32_t i, int32_t ub) {
  int32_t n = (int32_t)o->n;
  while (i < n && o->LCP[i] >= ub) ++i;
  return i < n ? i : -1;
}
/* :11-25 */
static void adjustInterval(const cms_oracle *o, int32_t lo, int32_t hi, int32_t offset,
                           int32_t *L, int32_t *R) {
  int32_t p = psv(o, lo, offset);
  if (p == -1) p = 0;
  int32_t q = nsv(o, hi + 1, offset);
  if (q == -1) q = (int32_t)o->n - 1; else q--;
  *L = p; *R = q;
}
/* :34-43 */
static void contractLeft(const cms_oracle *o, int32_t lo, int32_t hi, int32_t offset,
                         int32_t *L, int32_t *R) {
  uint32_t suflo = (uint32_t)o->SA[lo], sufhi = (uint32_t)o->SA[hi];
  if (suflo == o->n - 1 || sufhi == o->n - 1) { *L = 0; *R = (int32_t)o->n - 1; return; }
  adjustInterval(o, (int32_t)o->ISA[suflo + 1], (int32_t)o->ISA[sufhi + 1], offset, L, R);
}
/* :94-120 */
static int32_t binarySearchLB(const cms_oracle *o, int32_t lo, int32_t hi, uint32_t offset, uint8_t c) {
  int32_t low = lo, high = hi;
  while (low <= high) {
    int32_t mid = (low + high) >> 1;
    uint8_t v = o->x[(uint32_t)o->SA[mid] + offset];
    if (v < c) low = mid + 1;
    else if (v > c) high = mid - 1;
    else {
      if (mid == lo) return mid;
      if (o->x[(uint32_t)o->SA[mid - 1] + offset] == v) high = mid - 1; else return mid;
    }
  }
  return -(low + 1);
}
/* :123-149 */
static int32_t binarySearchRB(const cms_oracle *o, int32_t lo, int32_t hi, uint32_t offset, uint8_t c) {
  int32_t low = lo, high = hi;
  while (low <= high) {
    int32_t mid = (low + high) >> 1;
    uint8_t v = o->x[(uint32_t)o->SA[mid] + offset];
    if (v < c) low = mid + 1;
    else if (v > c) high = mid - 1;
    else {
      if (mid == hi) return mid;
      if (o->x[(uint32_t)o->SA[mid + 1] + offset] == v) low = mid + 1; else return mid;
    }
  }
  return -(low + 1);
}
/* :45-88.  sx = one document incl. its 0x02 terminator, slen its length (the reference bounds
 * the loop by the global _sn; the terminator always stops it earlier because 0x02 never occurs
 * in the augmented reference). */
static void computeMSFactorAt(const cms_oracle *o, const uint8_t *sx, uint64_t slen, uint64_t i,
                              uint32_t *pos, uint32_t *len, int32_t *leftB, int32_t *rightB,
                              int *smaller) {
  uint32_t offset = *len;
  uint64_t j = i + offset;
  int32_t nlb = *leftB, nrb = *rightB, maxMatch;
  uint32_t match = (uint32_t)o->SA[nlb];
  while (j < slen) {
    if (nlb == nrb) {
      uint8_t xc = o->x[(uint32_t)o->SA[nlb] + offset];
      if (xc != sx[j]) { *smaller = xc > sx[j]; break; }
      *leftB = nlb; *rightB = nrb;
    } else {
      nlb = binarySearchLB(o, nlb, nrb, offset, sx[j]);
      if (nlb < 0) {
        maxMatch = -nlb - 1;
        *smaller = 1;
        if (maxMatch == nrb + 1) { maxMatch--; *smaller = 0; }
        match = (uint32_t)o->SA[maxMatch];
        break;
      }
      nrb = binarySearchRB(o, nlb, nrb, offset, sx[j]);
      *leftB = nlb; *rightB = nrb;
    }
    match = (uint32_t)o->SA[nlb];
    j++; offset++;
  }
  *pos = match;
  *len = offset;
}

/* ------------------------------------------------------------------------------------------ */
/* stage 1: per-document MS driver (:1138-1213 / :1269-1352; small path :344-419 / :476-559)   */

typedef struct {      /* MatchInVectorBigger, match.h:5-16 */
  uint32_t start, pos, len, toNext, idx;
  uint8_t smaller;
} head_t;

typedef struct {
  head_t *h; uint64_t nh, cap;
  uint8_t *chars;              /* listOfChars */
  int64_t *bucket;             /* bucketsForExpandedBWT */
  uint64_t ndocs;
} stage1_t;

static void push_head(stage1_t *s, uint32_t start, uint32_t pos, uint32_t len, int sm, uint8_t ch) {
  if (s->nh == s->cap) {
    s->cap = s->cap ? s->cap * 2 : 1024;
    s->h = (head_t *)realloc(s->h, sizeof(head_t) * s->cap);
    s->chars = (uint8_t *)realloc(s->chars, s->cap);
  }
  head_t *h = &s->h[s->nh];
  h->start = start; h->pos = pos; h->len = len; h->smaller = (uint8_t)sm; h->toNext = 0;
  h->idx = (uint32_t)s->nh;
  s->chars[s->nh] = ch;
  s->nh++;
}

/* One document `doc[0..dlen)` whose last byte is the 0x02 terminator.  ms_* (may be NULL) get
 * the per-position matching statistics (index = position in the document, terminator excluded;
 * SURVEY.md §8c O4). */
static void ms_document(const cms_oracle *o, stage1_t *s, const uint8_t *doc, uint64_t dlen,
                        uint32_t *ms_pos, uint32_t *ms_len, uint8_t *ms_sm, uint8_t *ms_head) {
  int64_t i = 0;
  int32_t leftB = 0, rightB = (int32_t)o->n - 1;
  int smaller = 0;
  uint64_t prevPos = (uint64_t)-2;
  uint32_t pos = o->n - 1, len = 0, iCurrentDoc = 0;
  int64_t *bucket = s->bucket;
  while (i < (int64_t)dlen - 1) {
    computeMSFactorAt(o, doc, dlen, (uint64_t)i, &pos, &len, &leftB, &rightB, &smaller);
    int is_head = (pos != prevPos + 1);
    if (ms_pos) { ms_pos[i] = pos; ms_len[i] = len; ms_sm[i] = (uint8_t)smaller; ms_head[i] = (uint8_t)is_head; }
    if (is_head) {                                                     /* :1154-1165 */
      push_head(s, iCurrentDoc, pos, len, smaller, i == 0 ? 2 : doc[i - 1]);
      if (bucket[pos] > 0) bucket[pos] = -bucket[pos] - 1; else bucket[pos]--;
    } else {                                                           /* :1166-1171 */
      if (bucket[pos] >= 0) bucket[pos]++; else bucket[pos]--;
    }
    iCurrentDoc++;
    len--;
    if (leftB == rightB) {                                             /* :1175-1189 */
      while ((int32_t)len > o->PLCPp[pos + 1]) {
        i++; iCurrentDoc++; len--; pos++;
        if (ms_pos) { ms_pos[i] = pos; ms_len[i] = len + 1; ms_sm[i] = (uint8_t)smaller; ms_head[i] = 0; }
        if (bucket[pos] >= 0) bucket[pos]++; else bucket[pos]--;
      }
      adjustInterval(o, (int32_t)o->ISA[pos + 1], (int32_t)o->ISA[pos + 1], (int32_t)len, &leftB, &rightB);
    } else {                                                           /* :1190-1194 */
      contractLeft(o, leftB, rightB, (int32_t)len, &leftB, &rightB);
    }
    i++;
    prevPos = pos;
  }
  /* document-end sentinel head :1198-1212 */
  push_head(s, iCurrentDoc, o->n - 1, 0, 0, dlen == 1 ? 2 : doc[dlen - 2]);
  bucket[o->n - 1]--;
  s->ndocs++;
}

/* Per-position MS of a concatenated text (documents each followed by 0x02).  Arrays are
 * indexed by text position; terminator positions get pos = n-1, len = 0, head = 1. */
int cms_oracle_ms(const cms_oracle *o, const uint8_t *text, int64_t N, uint32_t *ms_pos,
                  uint32_t *ms_len, uint8_t *ms_sm, uint8_t *ms_head) {
  stage1_t s; memset(&s, 0, sizeof s);
  s.bucket = (int64_t *)calloc(o->n, sizeof(int64_t));
  int64_t p = 0;
  while (p < N) {
    const uint8_t *t = (const uint8_t *)memchr(text + p, 2, (size_t)(N - p));
    if (!t) break;
    int64_t e = t - text;
    ms_document(o, &s, text + p, (uint64_t)(e - p + 1), ms_pos + p, ms_len + p, ms_sm + p, ms_head + p);
    ms_pos[e] = o->n - 1; ms_len[e] = 0; ms_sm[e] = 0; ms_head[e] = 1;
    p = e + 1;
  }
  free(s.h); free(s.chars); free(s.bucket);
  return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* stage 2 + 3                                                                                  */

/* MatchInSet::operator< (match.h:27-33) on (len, smaller, isaNext); isaNext lives in .start
 * after linking, exactly as in the reference (:1375-1378). */
static int key_less(const head_t *a, const head_t *b) {
  if (a->len != b->len)
    return a->smaller * (a->len < b->len) + !b->smaller * (a->len > b->len);
  return a->start < b->start;
}
/* match.h:23-25 */
static int key_eq(const head_t *a, const head_t *b) { return a->len == b->len && a->start == b->start; }

static int cmp_phrase(const void *pa, const void *pb) {
  const head_t *a = (const head_t *)pa, *b = (const head_t *)pb;
  if (a->pos != b->pos) return a->pos < b->pos ? -1 : 1;     /* :1387-1392 (pos, then map order) */
  if (key_less(a, b)) return -1;
  if (key_less(b, a)) return 1;
  return (a->idx > b->idx) - (a->idx < b->idx);               /* idxs appended in idx order */
}

typedef struct {          /* one (position, MatchInSet) map entry with its ItemMatchInSet */
  uint32_t pos, len, isaNext, untilNext;
  uint8_t smaller;
  uint64_t first, count;   /* range into the sorted phrase array / idxs */
} group_t;

static int cmp_u32(const void *a, const void *b) {
  uint32_t x = *(const uint32_t *)a, y = *(const uint32_t *)b;
  return (x > y) - (x < y);
}
typedef struct { const uint32_t *isa; const group_t *g; } isa_ctx;
static int cmp_pos_by_isa(const void *a, const void *b, void *c) {
  const uint32_t *isa = ((isa_ctx *)c)->isa;
  uint32_t x = isa[*(const uint32_t *)a], y = isa[*(const uint32_t *)b];
  return (x > y) - (x < y);
}

/* lower_bound of key (len, smaller, isaNext) among groups [lo,hi) of one position (std::map
 * order = key_less). */
static uint64_t group_lower_bound(const group_t *g, uint64_t lo, uint64_t hi, const head_t *k) {
  while (lo < hi) {
    uint64_t mid = (lo + hi) >> 1;
    head_t m; m.len = g[mid].len; m.smaller = g[mid].smaller; m.start = g[mid].isaNext;
    if (key_less(&m, k)) lo = mid + 1; else hi = mid;
  }
  return lo;
}

/* Whole path on a concatenated text (documents each followed by 0x02): writes the N BWT bytes.
 * Optional outputs (may be NULL): *n_heads. */
int cms_oracle_bwt(const cms_oracle *o, const uint8_t *text, int64_t N, uint8_t *out_bwt,
                   uint64_t *n_heads) {
  const uint32_t n = o->n;
  stage1_t s; memset(&s, 0, sizeof s);
  s.bucket = (int64_t *)calloc(n, sizeof(int64_t));
  /* ---- stage 1 ---- */
  int64_t p = 0;
  while (p < N) {
    const uint8_t *t = (const uint8_t *)memchr(text + p, 2, (size_t)(N - p));
    if (!t) break;
    int64_t e = t - text;
    ms_document(o, &s, text + p, (uint64_t)(e - p + 1), NULL, NULL, NULL, NULL);
    p = e + 1;
  }
  const uint64_t H = s.nh;
  const uint64_t D1 = s.ndocs;                       /* the reference's D-1 */
  if (n_heads) *n_heads = H;
  head_t *ph = s.h;
  if (H == 0) { free(s.bucket); return 0; }
  /* ---- head linking :1359-1385 ---- */
  uint32_t *nextHead = (uint32_t *)malloc(sizeof(uint32_t) * H);
  {
    uint64_t j = 0;
    for (uint64_t i = 0; i + 1 < H; i++) {
      while ((ph[j].start + ph[j].len <= ph[i].start + ph[i].len) & (ph[j].len != 0)) j++;
      ph[i].toNext = ph[i].len > 0 ? ph[i + 1].start - ph[i].start - 1 : 0;
      ph[i].start = o->ISA[ph[j].pos + (ph[i].start + ph[i].len - ph[j].start)];
      nextHead[i] = (uint32_t)(j - i);
      if (ph[i].len == 0) j++;
    }
    nextHead[H - 1] = 0;
    ph[H - 1].start = 0;
  }
  /* ---- phrase sort + keyed insertion :1387-1414 ---- */
  head_t *sp = (head_t *)malloc(sizeof(head_t) * H);
  memcpy(sp, ph, sizeof(head_t) * H);
  qsort(sp, H, sizeof(head_t), cmp_phrase);
  group_t *g = (group_t *)malloc(sizeof(group_t) * H);
  uint32_t *idxs = (uint32_t *)malloc(sizeof(uint32_t) * H);
  uint64_t G = 0;
  for (uint64_t i = 0; i < H; i++) {
    idxs[i] = sp[i].idx;
    if (i > 0 && sp[i].pos == sp[i - 1].pos && key_eq(&sp[i], &sp[i - 1])) { g[G - 1].count++; continue; }
    g[G].pos = sp[i].pos; g[G].len = sp[i].len; g[G].isaNext = sp[i].start;
    g[G].smaller = sp[i].smaller; g[G].untilNext = sp[i].toNext;
    g[G].first = i; g[G].count = 1; G++;
  }
  /* per-position group ranges (the map of each position), positions in text order */
  uint64_t *posFirst = (uint64_t *)malloc(sizeof(uint64_t) * ((size_t)n + 1));
  {
    uint64_t k = 0;
    for (uint32_t q = 0; q <= n; q++) {
      while (k < G && g[k].pos < q) k++;
      posFirst[q] = k;
    }
  }
  /* ---- naming :1432-1453: positions by ISA, keys in map order; doc-end heads one name each */
  uint32_t *keys = (uint32_t *)malloc(sizeof(uint32_t) * (G + 1));
  uint64_t nkeys = 0;
  for (uint64_t k = 0; k < G; k++) if (k == 0 || g[k].pos != g[k - 1].pos) keys[nkeys++] = g[k].pos;
  isa_ctx ictx = {o->ISA, g};
  qsort_r(keys, nkeys, sizeof(uint32_t), cmp_pos_by_isa, &ictx);
  int32_t *names = (int32_t *)malloc(sizeof(int32_t) * (H + 1));
  int32_t rank = 1;
  for (uint64_t a = 0; a < nkeys; a++) {
    uint32_t q = keys[a];
    for (uint64_t k = posFirst[q]; k < posFirst[q + 1]; k++) {
      for (uint64_t t = 0; t < g[k].count; t++) {
        names[idxs[g[k].first + t]] = rank;
        if (q == n - 1) rank++;
      }
      rank++;
    }
  }
  names[H] = 0;
  /* ---- suffix sort of the name string :1456 ---- */
  int32_t *headSA = (int32_t *)malloc(sizeof(int32_t) * (H + 1));
  sa_doubling(names, headSA, (int32_t)(H + 1), rank);
  /* ---- inverse + BWTheads :1468-1472 ---- */
  int32_t *rankOf = names;                            /* rankToHead reused as the inverse */
  uint8_t *BWTheads = (uint8_t *)malloc(H + 1);
  for (uint64_t i = 0; i < H; i++) {
    rankOf[headSA[i + 1]] = (int32_t)i;
    BWTheads[i] = s.chars[headSA[i + 1]];
  }
  free(headSA);
  /* ---- next-rank substitution :1480-1494 ---- */
  for (uint64_t k = 0; k < G; k++) {
    for (uint64_t t = 0; t < g[k].count; t++) {
      uint32_t id = idxs[g[k].first + t];
      idxs[g[k].first + t] = (uint32_t)rankOf[id + nextHead[id]];
    }
    qsort(idxs + g[k].first, g[k].count, sizeof(uint32_t), cmp_u32);
  }
  /* ---- prefixSumForPositions :1504-1514 == g[k].first (groups are stored in position order) */
  /* ---- counterSmallerThanHead :1525-1603 ---- */
  uint64_t *counter = (uint64_t *)calloc(H + 1, sizeof(uint64_t));
  const int64_t *bucket = s.bucket;
  for (uint64_t k = 0; k < G; k++) {
    const uint32_t i = g[k].pos;
    const uint32_t *pid = idxs + g[k].first;
    const uint64_t pc = g[k].count;
    for (uint32_t idx = 0; idx < g[k].untilNext; idx++) {
      const uint32_t b = i + 1 + idx;
      if (bucket[b] >= 0) continue;
      head_t toFind; toFind.len = g[k].len - 1 - idx; toFind.smaller = g[k].smaller; toFind.start = g[k].isaNext;
      uint64_t lo = posFirst[b], hi = posFirst[b + 1];
      uint64_t q = group_lower_bound(g, lo, hi, &toFind);
      if (q == hi) continue;
      head_t found; found.len = g[q].len; found.smaller = g[q].smaller; found.start = g[q].isaNext;
      if (key_eq(&found, &toFind)) {
        const uint32_t *qid = idxs + g[q].first;
        const uint64_t qc = g[q].count;
        if (pid[pc - 1] < qid[0]) { counter[g[q].first] += pc; continue; }
        if (pid[0] > qid[qc - 1]) { if (q + 1 < hi) counter[g[q + 1].first] += pc; continue; }
        uint64_t a = 0, c = 0;
        while (c < pc && a < qc) {
          if (pid[c] < qid[a]) { counter[g[q].first + a]++; c++; } else a++;
        }
        if (c < pc && q + 1 < hi) counter[g[q + 1].first] += pc - c;
      } else {
        counter[g[q].first] += pc;
      }
    }
  }
  /* ---- bucket inversion :1618-1626 + plain writer :1638-1688 ---- */
  uint64_t w = 0;
  memcpy(out_bwt, BWTheads, D1); w = D1;
  uint64_t headCounter = D1;
  for (uint32_t r = 1; r < n; r++) {
    const uint32_t q = (uint32_t)o->SA[r];
    const uint8_t c = o->BWT[r];
    int64_t b = bucket[q];
    if (b >= 0) { memset(out_bwt + w, c, (size_t)b); w += (uint64_t)b; continue; }
    for (uint64_t k = posFirst[q]; k < posFirst[q + 1]; k++)
      for (uint64_t t = g[k].first; t < g[k].first + g[k].count; t++) {
        memset(out_bwt + w, c, (size_t)counter[t]); w += counter[t]; b += (int64_t)counter[t];
        out_bwt[w++] = BWTheads[headCounter++]; b++;
      }
    memset(out_bwt + w, c, (size_t)(-b)); w += (uint64_t)(-b);
  }
  free(counter); free(BWTheads); free(names); free(keys); free(posFirst); free(idxs); free(g);
  free(sp); free(nextHead); free(s.h); free(s.chars); free(s.bucket);
  return w == (uint64_t)N ? 0 : -1;
}

/* Maximal-run RLE in the reference's record format (:1708-1709: uint64 LE run, uint8 symbol),
 * i.e. the RLE writer :1695-1777 with its remainder flush moved below the sub-bucket loop.
 * out may be NULL to just count.  Returns the number of records. */
int64_t cms_oracle_rle(const uint8_t *bwt, int64_t N, uint8_t *out) {
  int64_t R = 0, i = 0;
  while (i < N) {
    int64_t j = i + 1;
    while (j < N && bwt[j] == bwt[i]) j++;
    if (out) {
      uint64_t run = (uint64_t)(j - i);
      memcpy(out + 9 * R, &run, 8);
      out[9 * R + 8] = bwt[i];
    }
    R++;
    i = j;
  }
  return R;
}
