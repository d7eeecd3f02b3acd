// Fast matching-statistics path (2-bit data): one *lane* per text window.
//
// Replaces the reference's per-document MS driver (CMS-BWT-functions.cpp:1138-1213; small path
// :344-419) with computeMSFactorAt :45-88, binarySearchLB/RB :94-149 and the PLCP' skip loop
// :1176-1185.  MS is a pure function of the suffix, so every window of the
// text is walked independently by one lane.  A warp iteration has four phases:
//
//   A  search (lane-local):  S_POS reads 32 packed bases at i, looks the k-mer up in ktab and scans
//      the bucket [L,R) of {key = the 16 bases after the k-mer, sa} entries (S_SCAN; S_BIN first
//      when the bucket is large, S_EQ when an equal key sits at the edge of the scanned group).
//      Predecessor, successor and their LCE (< k+16) come from the keys alone.
//   B  extension (warp-cooperative): a lane whose key matched hands its single candidate to the
//      warp, which compares 1024 packed bases per round (32 lanes x 32 bases, coalesced).
//   C  head test + record (lane-local, one copy of the code); msf_generic_search(), an exact
//      binary search on the packed data, serves everything rare (separator within k+16 bases,
//      empty bucket, repeated key, suffixes within k+16 bases of a run of other symbols in the reference --
//      an N block, an IUPAC code, the augmentation; msf_room_at).
//   D  skip (warp-cooperative): after a long match, the first q in (pos, E) with G[q] >= E
//      (:1176-1185 as a search) from one coalesced read of the 32 entries below E.
//
// Symbols other than A, C, G, T in the TEXT (N, IUPAC codes, lower case: the reference program takes any
// byte 3..127 through its augmentation, :231-237) do not leave this path: the 2-bit stream holds an
// arbitrary code at such a position and the position is listed, with the separators, in the sorted
// BREAK list (position << 1 | escape).  A match that starts before a break is limited by it exactly as
// by a separator, the byte that decides the order at the limit is read from the 8-bit text, and the
// (rare) escape positions themselves -- and matches that run through an escape into the augmentation
// -- take an exact byte-wise search (msf_escape_search).  MS is still the sequential definition.
// Symbols other than A, C, G, T in the REFERENCE are runs between which the 2-bit comparisons stay
// (msf_room_at, msf_make_key); a match continues through a run byte-wise where the text has the same symbols.
//
// The neighbour rule is the one of :53-76 as restated in ms.cu: the successor wins ties
// (smaller = 1).  This header is plain C++ when compiled without nvcc so that the same lane code
// is run on the CPU by tests/harness (test infrastructure; the product only ever runs it inside
// k_ms_fast); the cooperative phases have a device form (shuffles) and a host form (loops over
// the lanes of an emulated warp) built from the same per-lane probes.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define MSF_HD __host__ __device__ __forceinline__
#define MSF_HDN static __host__ __device__ __forceinline__
#else
#define MSF_HD inline
#define MSF_HDN static inline
#endif

typedef uint8_t msf_u8;
typedef uint32_t msf_u32;
typedef uint64_t msf_u64;

#if defined(__CUDA_ARCH__)
#define MSF_LD64(p) ((msf_u64)__ldg((const unsigned long long *)(p)))
#define MSF_LD32(p) ((msf_u32)__ldg((const unsigned int *)(p)))
#define MSF_LD8(p) ((msf_u8)__ldg((const unsigned char *)(p)))
#define MSF_ATOMIC_INC32(p) atomicAdd((unsigned int *)(p), 1u)
#define MSF_ATOMIC_INC64(p) atomicAdd((unsigned long long *)(p), 1ull)
#define MSF_BREV64(x) ((msf_u64)__brevll((unsigned long long)(x)))
#define MSF_BREV32(x) ((msf_u32)__brev((unsigned int)(x)))
#define MSF_CLZ32(x) ((msf_u32)__clz((int)(x)))
#define MSF_CTZ64(x) ((msf_u32)(__ffsll((long long)(x)) - 1))
#else
#define MSF_LD64(p) (*(const msf_u64 *)(p))
#define MSF_LD32(p) (*(const msf_u32 *)(p))
#define MSF_LD8(p) (*(const msf_u8 *)(p))
#define MSF_ATOMIC_INC32(p) __sync_fetch_and_add((msf_u32 *)(p), 1u)
#define MSF_ATOMIC_INC64(p) __sync_fetch_and_add((unsigned long long *)(p), 1ull)
static inline msf_u64 msf_brev64_host(msf_u64 v) {
  v = ((v >> 1) & 0x5555555555555555ull) | ((v & 0x5555555555555555ull) << 1);
  v = ((v >> 2) & 0x3333333333333333ull) | ((v & 0x3333333333333333ull) << 2);
  v = ((v >> 4) & 0x0F0F0F0F0F0F0F0Full) | ((v & 0x0F0F0F0F0F0F0F0Full) << 4);
  return __builtin_bswap64(v);
}
#define MSF_BREV64(x) msf_brev64_host(x)
#define MSF_BREV32(x) ((msf_u32)(msf_brev64_host((msf_u64)(x)) >> 32))
#define MSF_CLZ32(x) ((msf_u32)__builtin_clz((unsigned)(x)))
#define MSF_CTZ64(x) ((msf_u32)__builtin_ctzll((unsigned long long)(x)))
#endif

#define MSF_RB 32                   // head records per pool batch (32 x 16 B; one atomic per batch)
#define MSF_STG 4                   // head records staged per lane before one 64-byte store to the pool
struct msf_rec4 { msf_u32 x, y, z, w; };
#define MSF_NOB 0xFFFFFFFFu
#define MSF_BADB 0xFFFFFFFEu
#define MSF_INVALID 0xFFFFFFFFu
#define MSF_PAD_WORDS 4             // zero words readable past the last packed word

enum { MSF_S_FETCH = 0, MSF_S_POS, MSF_S_BIN, MSF_S_SCAN, MSF_S_EQ, MSF_S_DONE };
// break list entries: position << 1 | (1 = escape: a symbol outside A,C,G,T; 0 = document separator)
#define MSF_BRK_POS(v) ((msf_u64)(v) >> 1)
#define MSF_BRK_ESC(v) ((msf_u32)((v) & 1ull))

// ---- 2-bit helpers -----------------------------------------------------------------------------
// Packed layout: base j lives in word j >> 5, bits [2 (j & 31), 2 (j & 31) + 2); A,C,G,T = 0,1,2,3.
MSF_HD msf_u32 msf_code_of(msf_u32 c) { return ((c >> 1) & 3u) ^ ((c >> 2) & 1u); }   // A C G T only
MSF_HD msf_u8 msf_sym_of(msf_u32 b) { return (msf_u8)(b == 0 ? 'A' : b == 1 ? 'C' : b == 2 ? 'G' : 'T'); }
// 32 bases starting at base j (little-endian: base j in the low bits)
MSF_HD msf_u64 msf_get64(const msf_u64 *p, msf_u64 j) {
  const msf_u64 *q = p + (j >> 5);
  msf_u32 sh = (msf_u32)(j & 31) * 2;
  msf_u64 lo = MSF_LD64(q);
  if (sh == 0) return lo;
  msf_u64 hi = MSF_LD64(q + 1);
  return (lo >> sh) | (hi << (64 - sh));
}
MSF_HD msf_u32 msf_base(const msf_u64 *p, msf_u64 j) {
  return (msf_u32)(MSF_LD64(p + (j >> 5)) >> ((j & 31) * 2)) & 3u;
}
// little-endian packed bases -> big-endian (first base in the top bits): integer order = lexicographic
MSF_HD msf_u64 msf_be64(msf_u64 w) {
  msf_u64 r = MSF_BREV64(w);
  return ((r & 0x5555555555555555ull) << 1) | ((r >> 1) & 0x5555555555555555ull);
}
// LCE of text bases [i, ...) and reference bases [s, ...), at most maxl
MSF_HD msf_u32 msf_lce(const msf_u64 *tp, msf_u64 i, const msf_u64 *xp, msf_u64 s, msf_u32 maxl) {
  msf_u32 m = 0;
  while (m < maxl) {
    msf_u64 d = msf_get64(tp, i + m) ^ msf_get64(xp, s + m);
    if (d) {
      m += MSF_CTZ64(d) >> 1;
      break;
    }
    m += 32;
  }
  return m < maxl ? m : maxl;
}

// ---- the reference's symbols outside A,C,G,T ---------------------------------------------------
// x = A/C/G/T segments between RUNS of other symbols (N blocks, IUPAC codes, lower case, and always the
// augmentation at the end, :231-237).  The 2-bit reference holds an arbitrary code inside a run; a match
// never enters one on 2-bit data: room(s) = number of A/C/G/T symbols from s up to the next run.
// xruns[j] = start | end << 32 (sorted, the last run ends at n), xbk[b] = first run that ends after
// position 64 b.  One run (the augmentation alone: every pure reference) needs no table at all.
MSF_HD msf_u32 msf_room_at(const msf_u64 *xruns, const msf_u32 *xbk, msf_u32 nruns, msf_u32 n_acgt, msf_u32 s) {
  if (nruns <= 1) return s < n_acgt ? n_acgt - s : 0u;
  msf_u32 j = MSF_LD32(xbk + (s >> 6));
  msf_u64 r = MSF_LD64(xruns + j);
  while ((msf_u32)(r >> 32) <= s) { j++; r = MSF_LD64(xruns + j); }
  const msf_u32 st = (msf_u32)r;
  return st > s ? st - s : 0u;
}

// ---- index construction helpers (shared by index.cu and the CPU harness) ----------------------
// key of the suffix s (room = msf_room_at(s)): the 16 bases after its first k, big-endian; suffixes with
// fewer than k pure bases get the largest key (they sort after every member of the k-mer bucket they
// fall into).  A suffix whose run starts inside the 16 bases gets, from there on, the bases that keep the
// keys of its bucket in suffix order: the largest of A,C,G,T below the run's first symbol and then T's
// (A's when the symbol sorts before 'A', as the augmentation's 0x03 does).  Such a key is never used
// for an LCE: every search that meets the suffix takes the exact path (room < k + 16).
MSF_HD msf_u32 msf_make_key(const msf_u64 *xp, const msf_u8 *x8, msf_u32 s, msf_u32 room, int k) {
  if (room < (msf_u32)k) return 0xFFFFFFFFu;
  msf_u32 valid = room - (msf_u32)k;
  msf_u64 be = msf_be64(msf_get64(xp, (msf_u64)s + (msf_u32)k));
  msf_u32 key = (msf_u32)(be >> 32);
  if (valid < 16) {
    const msf_u8 b = x8[s + room];
    const msf_u32 dl = (msf_u32)(b > 'A') + (b > 'C') + (b > 'G') + (b > 'T');
    key = valid == 0 ? 0u : (key & (0xFFFFFFFFu << (2 * (16 - valid))));
    if (dl) key |= ((dl - 1u) << (2 * (15 - valid))) | ((1u << (2 * (15 - valid))) - 1u);
  }
  return key;
}

// ---- k-mer slot table -----------------------------------------------------------------------------
// slots[code * 8 .. code * 8 + 8): the {key, sa} entries of the k-mer's bucket (sorted, as in sak),
// padded with MSF_SLOT_EMPTY; a bucket of more than 8 suffixes keeps its first 7 and ends with
// MSF_SLOT_OVF (the lane then takes the ktab + sak path).  One aligned 64-byte read -- one DRAM
// granule -- replaces the dependent ktab -> sak round trips of the common case.
#define MSF_SLOT_EMPTY 0xFFFFFFFFFFFFFFFFull
#define MSF_SLOT_OVF 0xFFFFFFFEFFFFFFFFull
MSF_HD msf_u64 msf_slot_entry(const msf_u64 *sak, msf_u32 lb, msf_u32 cnt, int j) {
  if (cnt > 8 && j == 7) return MSF_SLOT_OVF;
  return (msf_u32)j < cnt ? sak[lb + j] : MSF_SLOT_EMPTY;
}

// A k-mer that does not occur has an unused slot.  Its neighbours in suffix order are the ranks
// lb - 1 and lb (lb = ktab[code]); when both are pure suffixes of at least k bases the slot holds
// their {k-mer code, sa} in entries 1 and 2 and MSF_SLOT_NB in entry 3: the LCE of the text with a
// suffix of another bucket is the common prefix of the two k-mer codes, so the search of such a
// position (the novel k-mers behind a mutation) ends with the slot read instead of the exact
// bisection (variant bit 16; measured at config 3: k_ms_fast 6.25 -> 5.63 ms).  Entry 0 stays MSF_SLOT_EMPTY.
#define MSF_SLOT_NB 0xFFFFFFFDFFFFFFFFull
MSF_HD void msf_slot_fill(const msf_u64 *sak, const msf_u32 *ktab, const msf_u64 *xp, msf_u32 n,
                          const msf_u64 *xruns, const msf_u32 *xbk, msf_u32 nruns, msf_u32 n_acgt, int k,
                          msf_u64 code, msf_u64 *e) {
  const msf_u32 lb = ktab[code], cnt = ktab[code + 1] - lb;
#pragma unroll
  for (int j = 0; j < 8; j++) e[j] = msf_slot_entry(sak, lb, cnt, j);
  if (cnt == 0 && lb > 0 && lb < n && xp) {
    const msf_u32 ps = (msf_u32)(sak[lb - 1] >> 32), ss = (msf_u32)(sak[lb] >> 32);
    if (msf_room_at(xruns, xbk, nruns, n_acgt, ps) >= (msf_u32)k && msf_room_at(xruns, xbk, nruns, n_acgt, ss) >= (msf_u32)k) {
      const msf_u32 pc = (msf_u32)(msf_be64(msf_get64(xp, ps)) >> (64 - 2 * k));
      const msf_u32 sc = (msf_u32)(msf_be64(msf_get64(xp, ss)) >> (64 - 2 * k));
      e[1] = (msf_u64)pc | ((msf_u64)ps << 32);
      e[2] = (msf_u64)sc | ((msf_u64)ss << 32);
      e[3] = MSF_SLOT_NB;
    }
  }
}

struct MsfParams {
  // reference index
  const msf_u64 *slots;   // k-mer slot table (NULL: every search goes through ktab + sak)
  const msf_u64 *xp;      // 2-bit packed x (arbitrary codes inside the runs of other symbols), zero padded
  const msf_u64 *xruns;   // runs of symbols outside A,C,G,T in x (msf_room_at); unused when nruns <= 1
  const msf_u32 *xbk;
  msf_u32 nruns;
  const msf_u8 *x8;       // augmented reference bytes
  const msf_u64 *sak;     // per SA rank: key (low 32 bits) | sa (high 32 bits)
  const msf_u32 *ktab;    // 4^k + 1 entries: number of suffixes smaller than k-mer m
  const msf_u32 *G;       // G[q] = q + PLCP'[q]
  msf_u32 n, n_acgt;
  int k;
  // text
  const msf_u64 *tp;      // packed text (breaks hold arbitrary codes)
  const msf_u8 *t8;       // the text as parsed (8-bit): read only at breaks
  msf_u64 N;
  const msf_u64 *seps;    // sorted break list (separators and escapes, MSF_BRK_*), seps[nseps] = ~0
  msf_u64 nseps;
  const msf_u32 *wsep;    // per window: index of the first break >= window start
  msf_u32 WL;             // window length
  msf_u64 nwin;
  msf_u32 skip_min;       // try the G skip when len >= skip_min
  msf_u32 variant;        // bit 1: per-lane text word cache in shared memory; bit 3: L2 prefetch of the G
                          // line(s) a skip will probe, issued when the extension ends (both on by default)
  // output
  msf_u32 *pool;          // records of 4 x u32: rel, pos, len, cs (ch | sm << 8 | tentative << 9)
  msf_u32 *batch_next;
  msf_u32 nb_cap;
  msf_u32 *nb_used;
  msf_u32 *wfirst, *wcnt, *wfirstpos, *wlastpos;
  msf_u32 *flags;         // [0] pool overflow
  unsigned long long *win_counter;
  msf_u32 *dpos, *dlen;   // optional per-position dump (parity hook)
  msf_u8 *dsm, *dhead;
};

struct MsfLane {
  msf_u64 i, w1, sepnext;
  msf_u32 sepidx, w;
  msf_u32 sepesc, prevesc;   // the next break is an escape; the previous position was one
  msf_u32 state, prevpos, prevterm, first;
  msf_u32 curb, fill, cnt;
  msf_u32 limit;
  msf_u32 tkey, lo, hi, rb, prevkey, prevsa, haveprev;
  msf_u32 pos, len, sm;
  msf_u32 chb;            // code of the base before position i (4 = not loaded yet)
#if !defined(__CUDA_ARCH__)
  msf_rec4 hstage[MSF_STG];
  msf_u64 htc0, htc1;
  msf_u32 htcw;
#endif
};

// Head records are staged four at a time (64 bytes = one DRAM granule) before they go to the pool:
// a lane emits a record every ~20 us, long after L2 has dropped the sector of its previous one, so
// single 16-byte stores each cost a DRAM read-modify-write.  Device: shared memory, one column
// per thread; host harness: inside the lane.
#if defined(__CUDA_ARCH__)
extern __shared__ uint4 msf_stage_smem[];
#define MSF_STAGE(L, j) msf_stage_smem[(j) * blockDim.x + threadIdx.x]
// per-lane text cache behind the stage area: two consecutive packed words and the index of the first
#define MSF_TC0(L) (((msf_u64 *)(msf_stage_smem + MSF_STG * blockDim.x))[threadIdx.x])
#define MSF_TC1(L) (((msf_u64 *)(msf_stage_smem + MSF_STG * blockDim.x))[blockDim.x + threadIdx.x])
#define MSF_TCW(L) (((msf_u32 *)(msf_stage_smem + MSF_STG * blockDim.x))[4 * blockDim.x + threadIdx.x])
#else
#define MSF_STAGE(L, j) (L).hstage[j]
#define MSF_TC0(L) (L).htc0
#define MSF_TC1(L) (L).htc1
#define MSF_TCW(L) (L).htcw
#endif
#define MSF_SMEM_BYTES(threads) ((size_t)(threads) * (MSF_STG * 16 + 8 + 8 + 4))

// 32 text bases starting at base j through the lane's two-word cache: consecutive search positions
// (the cluster of short heads before a mutation) share their words, and the lane's line has long
// left L1 by the time it comes back (1024 lanes per SM each stream slots through it)
MSF_HD msf_u64 msf_text64_cached(const MsfParams &P, MsfLane &L, msf_u64 j) {
  const msf_u64 wi = j >> 5;
  const msf_u32 sh = (msf_u32)(j & 31) * 2, w32 = (msf_u32)wi;
  msf_u64 c0, c1;
  if (MSF_TCW(L) == w32) { c0 = MSF_TC0(L); c1 = MSF_TC1(L); }
  else {
    if (MSF_TCW(L) + 1u == w32 && w32 != 0) c0 = MSF_TC1(L); else c0 = MSF_LD64(P.tp + wi);
    c1 = MSF_LD64(P.tp + wi + 1);
    MSF_TC0(L) = c0; MSF_TC1(L) = c1; MSF_TCW(L) = w32;
  }
  return sh ? (c0 >> sh) | (c1 << (64 - sh)) : c0;
}

MSF_HD void msf_store_rec(const MsfParams &P, msf_u64 slot, msf_u32 a, msf_u32 b, msf_u32 c, msf_u32 d) {
#if defined(__CUDA_ARCH__)
  ((uint4 *)P.pool)[slot] = make_uint4(a, b, c, d);
#else
  msf_u32 *r = P.pool + 4 * slot;
  r[0] = a; r[1] = b; r[2] = c; r[3] = d;
#endif
}

// writes the last cnt staged records of the current batch (they end at record L.fill)
MSF_HD void msf_flush(const MsfParams &P, MsfLane &L, msf_u32 cnt) {
  if (L.curb == MSF_BADB || L.curb == MSF_NOB) return;
  const msf_u64 base = (msf_u64)L.curb * MSF_RB + (L.fill - cnt);
  for (msf_u32 j = 0; j < cnt; j++) {
#if defined(__CUDA_ARCH__)
    ((uint4 *)P.pool)[base + j] = MSF_STAGE(L, j);
#else
    msf_store_rec(P, base + j, MSF_STAGE(L, j).x, MSF_STAGE(L, j).y, MSF_STAGE(L, j).z, MSF_STAGE(L, j).w);
#endif
  }
}

MSF_HD void msf_emit(const MsfParams &P, MsfLane &L, msf_u32 rel, msf_u32 pos, msf_u32 len, msf_u32 cs) {
  if (L.curb == MSF_NOB || L.fill == MSF_RB) {
    msf_u32 nb = MSF_ATOMIC_INC32(P.nb_used);
    if (nb >= P.nb_cap) {
      P.flags[0] = 1;
      nb = MSF_BADB;
    } else if (L.curb == MSF_NOB) {
      P.wfirst[L.w] = nb;
    } else if (L.curb != MSF_BADB) {
      P.batch_next[L.curb] = nb;
    }
    L.curb = nb;
    L.fill = 0;
  }
  {
    const msf_u32 k = L.fill & (MSF_STG - 1);
#if defined(__CUDA_ARCH__)
    MSF_STAGE(L, k) = make_uint4(rel, pos, len, cs);
#else
    MSF_STAGE(L, k).x = rel; MSF_STAGE(L, k).y = pos; MSF_STAGE(L, k).z = len; MSF_STAGE(L, k).w = cs;
#endif
  }
  L.fill++;
  if ((L.fill & (MSF_STG - 1)) == 0) msf_flush(P, L, MSF_STG);
  L.cnt++;
}

// number of leading bytes on which a and b agree, counted from m on (they are known to agree before m).  The
// caller's streams always differ somewhere (0x02 occurs nowhere in x, x[n-1] = 0 is unique); eight bytes
// per step -- a match that runs through an N block of the reference is tens of thousands of bytes long
// -- both buffers are readable 64 bytes past their ends.
MSF_HD msf_u64 msf_ld64u(const msf_u8 *p) {
#if defined(__CUDA_ARCH__)
  const unsigned long long a = (unsigned long long)p;
  const msf_u32 sh = (msf_u32)(a & 7ull) * 8u;
  const msf_u64 *q = (const msf_u64 *)(a & ~7ull);
  const msf_u64 lo = MSF_LD64(q);
  if (sh == 0) return lo;
  return (lo >> sh) | (MSF_LD64(q + 1) << (64u - sh));
#else
  msf_u64 v;
  __builtin_memcpy(&v, p, 8);
  return v;
#endif
}
MSF_HD msf_u32 msf_match_bytes(const msf_u8 *a, const msf_u8 *b, msf_u32 m) {
  for (;;) {
    const msf_u64 d = msf_ld64u(a + m) ^ msf_ld64u(b + m);
    if (d) return m + (MSF_CTZ64(d) >> 3);
    m += 8;
  }
}

MSF_HD msf_u32 msf_room(const MsfParams &P, msf_u32 s) { return msf_room_at(P.xruns, P.xbk, P.nruns, P.n_acgt, s); }

// order of the reference suffix s against the text suffix at i (limit symbols up to its separator):
// m = LCE, returns true when the reference suffix is the smaller one.  Same semantics as probe()
// in ms.cu on the 8-bit data: past n_acgt the reference holds no A/C/G/T, and the separator 0x02
// is smaller than every reference symbol but x[n-2] = 1 and x[n-1] = 0.
MSF_HD bool msf_compare(const MsfParams &P, msf_u64 i, msf_u32 limit, msf_u32 s, msf_u32 &m, msf_u32 &xb,
                        msf_u32 &tb) {
  msf_u32 room = msf_room(P, s);
  msf_u32 maxl = limit < room ? limit : room;
  m = msf_lce(P.tp, i, P.xp, s, maxl);
  xb = MSF_LD8(P.x8 + s + m);
  tb = m == limit ? (msf_u32)MSF_LD8(P.t8 + i + m) : (msf_u32)msf_sym_of(msf_base(P.tp, i + m));
  if (m == limit && xb == tb) {        // an escape that the reference continues (a run of its own, its augmentation): bytes from here
    m = msf_match_bytes(P.x8 + s, P.t8 + i, m + 1);
    xb = MSF_LD8(P.x8 + s + m); tb = MSF_LD8(P.t8 + i + m);
  }
  return xb < tb;
}

// exact search of the text suffix at i on the 8-bit data (escape positions): same neighbour rule.
// The separator 0x02 occurs nowhere in x and x[n-1] = 0 is its unique smallest symbol, so the byte
// comparison stops by itself.
MSF_HDN void msf_escape_search(const MsfParams &P, msf_u64 i, msf_u32 &pos, msf_u32 &len, msf_u32 &sm) {
  msf_u32 lo = 0, hi = P.n;
  const msf_u8 *t = P.t8 + i;
  while (lo < hi) {
    const msf_u32 mid = lo + ((hi - lo) >> 1);
    const msf_u8 *x = P.x8 + (msf_u32)(MSF_LD64(P.sak + mid) >> 32);
    const msf_u32 m = msf_match_bytes(x, t, 0);
    if (MSF_LD8(x + m) < MSF_LD8(t + m)) lo = mid + 1; else hi = mid;
  }
  long long lp = -1, ls = -1;
  msf_u32 ppos = 0, spos = 0;
  if (lo > 0) {
    ppos = (msf_u32)(MSF_LD64(P.sak + lo - 1) >> 32);
    lp = msf_match_bytes(P.x8 + ppos, t, 0);
  }
  if (lo < P.n) {
    spos = (msf_u32)(MSF_LD64(P.sak + lo) >> 32);
    ls = msf_match_bytes(P.x8 + spos, t, 0);
  }
  if (ls >= lp) { pos = spos; len = (msf_u32)ls; sm = 1; }
  else { pos = ppos; len = (msf_u32)lp; sm = 0; }
}

// exact search on packed data (rare cases); same result as lane_search()/warp_search() in ms.cu
MSF_HDN void msf_generic_search(const MsfParams &P, msf_u64 i, msf_u32 limit, msf_u32 &pos, msf_u32 &len,
                                msf_u32 &sm) {
  msf_u32 lo = 0, hi = P.n;
  if (limit >= (msf_u32)P.k) {
    msf_u64 be = msf_be64(msf_get64(P.tp, i));
    msf_u32 code = (msf_u32)(be >> (64 - 2 * P.k));
    lo = MSF_LD32(P.ktab + code);
    hi = MSF_LD32(P.ktab + code + 1);
  }
  msf_u32 m, xb, tb;
  while (lo < hi) {
    msf_u32 mid = lo + ((hi - lo) >> 1);
    msf_u32 s = (msf_u32)(MSF_LD64(P.sak + mid) >> 32);
    if (msf_compare(P, i, limit, s, m, xb, tb)) lo = mid + 1; else hi = mid;
  }
  long long lp = -1, ls = -1;
  msf_u32 ppos = 0, spos = 0;
  if (lo > 0) {
    ppos = (msf_u32)(MSF_LD64(P.sak + lo - 1) >> 32);
    msf_compare(P, i, limit, ppos, m, xb, tb);
    lp = m;
  }
  if (lo < P.n) {
    spos = (msf_u32)(MSF_LD64(P.sak + lo) >> 32);
    msf_compare(P, i, limit, spos, m, xb, tb);
    ls = m;
  }
  if (ls >= lp) { pos = spos; len = (msf_u32)ls; sm = 1; }
  else { pos = ppos; len = (msf_u32)lp; sm = 0; }
}

// ---- the state machine ------------------------------------------------------------------------
MSF_HD void msf_dump(const MsfParams &P, msf_u64 p, msf_u32 pos, msf_u32 len, msf_u32 sm, msf_u32 head) {
  if (P.dpos) {
    P.dpos[p] = pos; P.dlen[p] = len; P.dsm[p] = (msf_u8)sm; P.dhead[p] = (msf_u8)head;
  }
}

// what the later phases of an iteration have to do for this lane
#define MSF_ACT_NONE 0u
#define MSF_ACT_FIN 1u        // (L.pos, L.len, L.sm) is the MS of position L.i
#define MSF_ACT_GENERIC 2u    // run the exact search first, then FIN
#define MSF_ACT_TERM 4u       // position L.i is a document terminator
#define MSF_ACT_EXT 8u        // extend the candidate L.pos from L.len matched bases (phase B)
#define MSF_ACT_SKIP 16u      // long match recorded: find the end of its non-head run (phase D)
#define MSF_ACT_ESC 32u       // position L.i holds an escape symbol: exact byte-wise search, then FIN

MSF_HD void msf_next_break(const MsfParams &P, MsfLane &L) {
  const msf_u64 v = MSF_LD64(P.seps + L.sepidx);
  L.sepnext = MSF_BRK_POS(v);
  L.sepesc = MSF_BRK_ESC(v);
}
// phase C: head test (:1154), record, step.  term: L.i holds a separator (:1198-1212).
// Returns MSF_ACT_SKIP when the non-head run after this head has to be looked up.
// esc: L.i holds an escape symbol (its MS comes from msf_escape_search; a long match is followed by its skip like any other)
MSF_HD msf_u32 msf_fin(const MsfParams &P, MsfLane &L, bool term, bool esc = false) {
  const bool tent = !term && L.first && !L.prevterm;   // window starts inside a document
  const bool head = term || L.prevterm || L.pos != L.prevpos + 1;
  if (L.first) P.wfirstpos[L.w] = (term || L.prevterm) ? MSF_INVALID : L.pos;
  if (head || tent) {
    msf_u32 ch = L.prevterm ? 2u : L.prevesc ? (msf_u32)MSF_LD8(P.t8 + L.i - 1)
                                             : (msf_u32)msf_sym_of(L.chb < 4 ? L.chb : msf_base(P.tp, L.i - 1));
    msf_emit(P, L, (msf_u32)(L.i - (msf_u64)L.w * P.WL), L.pos, L.len, ch | (L.sm << 8) | (tent ? 1u << 9 : 0u));
  }
  msf_dump(P, L.i, L.pos, L.len, L.sm, head || tent);
  L.first = 0;
  L.prevterm = term ? 1u : 0u;
  L.prevesc = esc ? 1u : 0u;
  L.state = MSF_S_POS;
  // a long match that starts AT an escape (a document or a window that begins inside an N block which the
  // reference has too) is followed by its non-head run like any other: without the skip every symbol of
  // the block would take its own exact search, each as long as the rest of the block.  msf_skip_to moves
  // the break cursor over the breaks of the run.
  if (esc && L.len >= P.skip_min && L.len > 1) return MSF_ACT_SKIP;
  if (term || esc) {
    L.sepidx++;
    msf_next_break(P, L);
    if (esc) L.prevpos = L.pos;
    L.i++;
    return MSF_ACT_NONE;
  }
  if (L.len >= P.skip_min && L.len > 1) return MSF_ACT_SKIP;
  L.prevpos = L.pos;
  L.i++;
  return MSF_ACT_NONE;
}

// the non-head run after a head ends at q = first position in (pos, E) with G[q] >= E (or E)
MSF_HD void msf_skip_to(const MsfParams &P, MsfLane &L, msf_u32 q) {
  msf_u32 skip = q - L.pos - 1;
  if (P.dpos)
    for (msf_u32 k2 = 1; k2 <= skip; k2++) msf_dump(P, L.i + k2, L.pos + k2, L.len - k2, L.sm, 0);
  L.i += 1 + (msf_u64)skip;
  L.prevpos = L.pos + skip;
  if (L.sepnext < L.i) {                 // the match ran through escapes: the run skipped over their breaks
    msf_u64 pb = 0;
    do { pb = MSF_LD64(P.seps + L.sepidx); L.sepidx++; msf_next_break(P, L); } while (L.sepnext < L.i);
    L.prevesc = (MSF_BRK_POS(pb) == L.i - 1 && MSF_BRK_ESC(pb)) ? 1u : 0u;
  }
}

// single candidate s whose key equals the text's: hand it to the warp (or to the exact search)
MSF_HD msf_u32 msf_start_ext(const MsfParams &P, MsfLane &L, msf_u32 s) {
  const msf_u32 KL = (msf_u32)P.k + 16;
  if (msf_room(P, s) < KL) return MSF_ACT_GENERIC;
  L.pos = s;
  L.len = KL;
  return MSF_ACT_EXT;
}
MSF_HD msf_u32 msf_ext_max(const MsfParams &P, msf_u32 limit, msf_u32 pos) {
  msf_u32 room = msf_room(P, pos);
  return limit < room ? limit : room;
}
// phase B probe: the 32 bases at offset off of the two streams.  Returns 0xFFFFFFFF when they are
// equal, else (offset of the first mismatch << 1) | (reference base > text base), so that a
// min-reduction over the lanes yields the match length and the smaller flag at once.
MSF_HD msf_u32 msf_ext_probe(const MsfParams &P, msf_u64 i, msf_u32 pos, msf_u32 off) {
  const msf_u64 tw = msf_get64(P.tp, i + off), xw = msf_get64(P.xp, (msf_u64)pos + off);
  const msf_u64 d = tw ^ xw;
  if (!d) return 0xFFFFFFFFu;
  const msf_u32 m = MSF_CTZ64(d) >> 1;
  const msf_u32 tc = (msf_u32)(tw >> (2 * m)) & 3u, xc = (msf_u32)(xw >> (2 * m)) & 3u;
  return ((off + m) << 1) | (xc > tc ? 1u : 0u);
}
// phase B result for the owner lane.  res = min over the probes (0xFFFFFFFF: no mismatch seen),
// bmax = min(limit, room in the pure reference).  A mismatch before bmax decides everything;
// otherwise the match ends at a separator or at the augmentation and the bytes decide
// (smaller = x[pos+len] > next text symbol, :53-58).
MSF_HD void msf_ext_done(const MsfParams &P, MsfLane &L, msf_u32 res, msf_u32 bmax) {
  if (res != 0xFFFFFFFFu && (res >> 1) < bmax) { L.len = res >> 1; L.sm = res & 1u; return; }
  L.len = bmax;
  msf_u32 xb = MSF_LD8(P.x8 + L.pos + bmax);
  msf_u32 tb = bmax == L.limit ? (msf_u32)MSF_LD8(P.t8 + L.i + bmax) : (msf_u32)msf_sym_of(msf_base(P.tp, L.i + bmax));
  if (bmax == L.limit && xb == tb) {   // the single candidate continues through an escape: bytes from here
    const msf_u32 m = msf_match_bytes(P.x8 + L.pos, P.t8 + L.i, bmax + 1);
    xb = MSF_LD8(P.x8 + L.pos + m); tb = MSF_LD8(P.t8 + L.i + m);
    L.len = m;
  }
  L.sm = xb > tb ? 1u : 0u;
}

MSF_HD void msf_load_slot(const msf_u64 *sp, msf_u64 *e) {
#if defined(__CUDA_ARCH__)
  const ulonglong2 *q = (const ulonglong2 *)sp;
  const ulonglong2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3);
  e[0] = a.x; e[1] = a.y; e[2] = b.x; e[3] = b.y; e[4] = c.x; e[5] = c.y; e[6] = d.x; e[7] = d.y;
#else
  for (int j = 0; j < 8; j++) e[j] = sp[j];
#endif
}

// predecessor (prevkey/prevsa, if any) and successor (skey/ssa, if found) are known by key only
MSF_HD msf_u32 msf_decide(const MsfParams &P, MsfLane &L, bool found, msf_u32 skey, msf_u32 ssa) {
  const msf_u32 KL = (msf_u32)P.k + 16;
  if ((found && msf_room(P, ssa) < KL) || (L.haveprev && msf_room(P, L.prevsa) < KL))
    return MSF_ACT_GENERIC;
  msf_u32 lp = 0, ls = 0;
  if (L.haveprev) lp = (msf_u32)P.k + (MSF_CLZ32(L.prevkey ^ L.tkey) >> 1);
  if (found) ls = (msf_u32)P.k + (MSF_CLZ32(skey ^ L.tkey) >> 1);
  if (found && (!L.haveprev || ls >= lp)) { L.pos = ssa; L.len = ls; L.sm = 1; }
  else { L.pos = L.prevsa; L.len = lp; L.sm = 0; }
  return MSF_ACT_FIN;
}

// phase A: one search step of this lane
MSF_HD msf_u32 msf_search_step(const MsfParams &P, MsfLane &L) {
  msf_u32 act = MSF_ACT_NONE;
  switch (L.state) {
    case MSF_S_FETCH: {
      msf_u64 w = MSF_ATOMIC_INC64(P.win_counter);
      if (w >= P.nwin) { L.state = MSF_S_DONE; break; }
      L.w = (msf_u32)w;
      L.i = w * P.WL;
      L.w1 = L.i + P.WL < P.N ? L.i + P.WL : P.N;
      L.sepidx = MSF_LD32(P.wsep + w);
      msf_next_break(P, L);
      {
        const msf_u64 pb = L.sepidx > 0 ? MSF_LD64(P.seps + L.sepidx - 1) : 0;
        const bool adj = L.sepidx > 0 && MSF_BRK_POS(pb) == L.i - 1;
        L.prevterm = (L.i == 0) || (adj && !MSF_BRK_ESC(pb));
        L.prevesc = (L.i != 0 && adj && MSF_BRK_ESC(pb)) ? 1u : 0u;
      }
      L.prevpos = MSF_INVALID;
      L.first = 1;
      MSF_TCW(L) = 0xFFFFFFFFu;
      L.curb = MSF_NOB;
      L.fill = 0;
      L.cnt = 0;
      L.state = MSF_S_POS;
    }
    // fall through
    case MSF_S_POS: {
      if (L.i >= L.w1) {                                  // close the window
        if (L.fill & (MSF_STG - 1)) msf_flush(P, L, L.fill & (MSF_STG - 1));
        P.wcnt[L.w] = L.cnt;
        P.wlastpos[L.w] = L.prevterm ? MSF_INVALID : L.prevpos - (msf_u32)(L.i - L.w1);
        L.state = MSF_S_FETCH;
        break;
      }
      L.chb = 4;
      if (L.i == L.sepnext) {
        if (L.sepesc) { act = MSF_ACT_ESC; break; }       // a symbol outside A,C,G,T
        L.pos = P.n - 1; L.len = 0; L.sm = 0;             // document terminator (:1198-1212)
        act = MSF_ACT_TERM;
        break;
      }
      const msf_u32 KL = (msf_u32)P.k + 16;
      msf_u64 dist = L.sepnext - L.i;
      L.limit = dist > 0xFFFFFFF0ull ? 0xFFFFFFF0u : (msf_u32)dist;
      if (L.limit < KL) { act = MSF_ACT_GENERIC; break; }
      // one read serves the k-mer, the key and the symbol before i (the head record needs it)
      msf_u64 tw = (P.variant & 2u) ? msf_text64_cached(P, L, L.i ? L.i - 1 : 0) : msf_get64(P.tp, L.i ? L.i - 1 : 0);
      if (L.i) { L.chb = (msf_u32)tw & 3u; tw >>= 2; }
      msf_u64 be = msf_be64(tw);
      msf_u32 code = (msf_u32)(be >> (64 - 2 * P.k));
      L.tkey = (msf_u32)((be << (2 * P.k)) >> 32);
      if (P.slots) {
        msf_u64 e[8];
        msf_load_slot(P.slots + (msf_u64)code * 8, e);
        if (e[0] == MSF_SLOT_EMPTY) {                                        // the k-mer does not occur
          if ((P.variant & 16u) && e[3] == MSF_SLOT_NB) {
            // neighbours from other buckets: LCE = common prefix of the k-mer codes (< k <= limit);
            // the successor wins ties, as in msf_generic_search
            const msf_u32 sh = 32u - 2u * (msf_u32)P.k;
            const msf_u32 lp = (MSF_CLZ32(((msf_u32)e[1] ^ code) << sh)) >> 1;
            const msf_u32 ls = (MSF_CLZ32(((msf_u32)e[2] ^ code) << sh)) >> 1;
            if (ls >= lp) { L.pos = (msf_u32)(e[2] >> 32); L.len = ls; L.sm = 1; }
            else { L.pos = (msf_u32)(e[1] >> 32); L.len = lp; L.sm = 0; }
            act = MSF_ACT_FIN;
            break;
          }
          act = MSF_ACT_GENERIC;
          break;
        }
        if (e[7] != MSF_SLOT_OVF) {
          // pred = last entry with a smaller key, succ = first entry with a key >= the text's, nxt =
          // the entry after succ (empty entries hold the largest key and sort last).  Written as
          // running selections over constant indices so that the slot stays in registers.
          msf_u64 pred = 0, succ = MSF_SLOT_EMPTY, nxt = MSF_SLOT_EMPTY;
#pragma unroll
          for (int j = 0; j < 8; j++)
            if ((msf_u32)e[j] < L.tkey) pred = e[j];
#pragma unroll
          for (int j = 7; j >= 0; j--)
            if (!((msf_u32)e[j] < L.tkey)) { nxt = succ; succ = e[j]; }
          L.haveprev = (msf_u32)e[0] < L.tkey;
          L.prevkey = (msf_u32)pred; L.prevsa = (msf_u32)(pred >> 32);
          if (succ == MSF_SLOT_EMPTY) act = msf_decide(P, L, false, 0, 0);   // every key is smaller
          else {
            const msf_u32 key = (msf_u32)succ, s = (msf_u32)(succ >> 32);
            if (key != L.tkey) act = msf_decide(P, L, true, key, s);
            else if (nxt != MSF_SLOT_EMPTY && (msf_u32)nxt == L.tkey) act = MSF_ACT_GENERIC;   // repeated key
            else act = msf_start_ext(P, L, s);
          }
          break;
        }
      }
      msf_u32 lb = MSF_LD32(P.ktab + code), rb = MSF_LD32(P.ktab + code + 1);
      if (lb == rb) { act = MSF_ACT_GENERIC; break; }
      L.lo = lb; L.hi = rb; L.rb = rb; L.haveprev = 0;
      if (rb - lb > 16) { L.state = MSF_S_BIN; break; }
      L.state = MSF_S_SCAN;
    }                                                     // small bucket: scan it in the same step
    // fall through
    case MSF_S_SCAN: {
      msf_u64 e[4];
#pragma unroll
      for (int j = 0; j < 4; j++) e[j] = L.lo + j < L.hi ? MSF_LD64(P.sak + L.lo + j) : ~0ull;
      // first entry (of these four) whose key is >= the text's; smaller ones become the predecessor
      int f = 4;
#pragma unroll
      for (int j = 3; j >= 0; j--)
        if (L.lo + j < L.hi && (msf_u32)e[j] >= L.tkey) f = j;
#pragma unroll
      for (int j = 0; j < 4; j++)
        if (j < f && L.lo + j < L.hi) { L.prevkey = (msf_u32)e[j]; L.prevsa = (msf_u32)(e[j] >> 32); L.haveprev = 1; }
      if (f < 4) {
        msf_u64 ef = f == 0 ? e[0] : f == 1 ? e[1] : f == 2 ? e[2] : e[3];
        msf_u64 en = f == 0 ? e[1] : f == 1 ? e[2] : e[3];             // entry after it (if loaded)
        msf_u32 key = (msf_u32)ef, s = (msf_u32)(ef >> 32), idx = L.lo + f;
        if (key != L.tkey) act = msf_decide(P, L, true, key, s);
        else if (idx + 1 >= L.rb) act = msf_start_ext(P, L, s);          // last of the bucket: unique
        else if (f < 3 && idx + 1 < L.hi) act = (msf_u32)en == L.tkey ? MSF_ACT_GENERIC : msf_start_ext(P, L, s);
        else { L.pos = s; L.lo = idx; L.state = MSF_S_EQ; }
      } else {
        L.lo += 4;
        if (L.lo >= L.hi) act = msf_decide(P, L, false, 0, 0);           // every key is smaller
      }
      break;
    }
    case MSF_S_EQ: {                                      // is the entry after the equal one equal too?
      msf_u64 e = MSF_LD64(P.sak + L.lo + 1);
      act = (msf_u32)e == L.tkey ? MSF_ACT_GENERIC : msf_start_ext(P, L, L.pos);
      break;
    }
    case MSF_S_BIN: {                                     // keys < tkey left of lo, >= tkey from hi on
      msf_u32 mid = L.lo + ((L.hi - L.lo) >> 1);
      msf_u64 e = MSF_LD64(P.sak + mid);
      if ((msf_u32)e < L.tkey) { L.lo = mid + 1; L.prevkey = (msf_u32)e; L.prevsa = (msf_u32)(e >> 32); L.haveprev = 1; }
      else L.hi = mid;
      if (L.hi - L.lo <= 3) {
        L.hi = L.hi + 1 < L.rb ? L.hi + 1 : L.rb;
        L.state = MSF_S_SCAN;
      }
      break;
    }
    default: break;
  }
  return act;
}

// phase D probe for one lane of the group looking at the 32 entries [E - 32 (w + 1), E - 32 w):
// returns 1 when entry idx is inside (pos, E) and fails G[idx] >= E, 2 when inside and it holds
MSF_HD msf_u32 msf_skip_probe(const MsfParams &P, msf_u32 pos, msf_u32 E, msf_u32 idx) {
  if (idx <= pos || idx >= E) return 0;
  return MSF_LD32(P.G + idx) >= E ? 2u : 1u;
}
// after MSF_SKIP_ROUNDS windows of 32 without a failing entry: bisection on (pos, hi)
#define MSF_SKIP_ROUNDS 4
MSF_HD msf_u32 msf_skip_bisect(const MsfParams &P, msf_u32 pos, msf_u32 E, msf_u32 hi) {
  msf_u32 lo = pos + 1;                                   // G[q] >= E for all q in [hi, E)
  while (lo < hi) {
    msf_u32 mid = lo + ((hi - lo) >> 1);
    if (MSF_LD32(P.G + mid) >= E) hi = mid; else lo = mid + 1;
  }
  return lo;
}

#if defined(__CUDACC__)
// ---- one warp iteration (device) ----------------------------------------------------------------
__device__ __forceinline__ void msf_warp_iter(const MsfParams &P, MsfLane &L) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  msf_u32 act = msf_search_step(P, L);
  // phase B: cooperative extension, one candidate at a time
  unsigned em = __ballot_sync(FULL, act == MSF_ACT_EXT);
  while (em) {
    const int src = __ffs(em) - 1;
    em &= em - 1;
    const msf_u64 bi = __shfl_sync(FULL, L.i, src);
    const msf_u32 bpos = __shfl_sync(FULL, L.pos, src);
    msf_u32 blen = __shfl_sync(FULL, L.len, src);
    const msf_u32 bmax = __shfl_sync(FULL, msf_ext_max(P, L.limit, L.pos), src);
    msf_u32 res;
    while (true) {
      const msf_u32 off = blen + 32u * lane;
      msf_u32 cand = 0xFFFFFFFFu;
      if (off < bmax) cand = msf_ext_probe(P, bi, bpos, off);
      res = __reduce_min_sync(FULL, cand);
      if (res != 0xFFFFFFFFu || blen + 1024u >= bmax) break;
      blen += 1024u;
    }
    if (lane == src) {
      msf_ext_done(P, L, res, bmax);
      act = MSF_ACT_FIN;
      // phase D will probe G[E - 32, E): have the line(s) on their way while phase C runs
      if ((P.variant & 8u) && L.len >= P.skip_min && L.len >= 32u) {
        const msf_u32 *gp = P.G + L.pos + L.len;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(gp - 32));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(gp - 1));
      }
    }
  }
  // phase C
  if (act & MSF_ACT_GENERIC) msf_generic_search(P, L.i, L.limit, L.pos, L.len, L.sm);
  if (act & MSF_ACT_ESC) msf_escape_search(P, L.i, L.pos, L.len, L.sm);
  if (act) act = msf_fin(P, L, (act & MSF_ACT_TERM) != 0, (act & MSF_ACT_ESC) != 0);
  // phase D: cooperative skip
  unsigned sk = __ballot_sync(FULL, act == MSF_ACT_SKIP);
  while (sk) {
    const int src = __ffs(sk) - 1;
    sk &= sk - 1;
    const msf_u32 bpos = __shfl_sync(FULL, L.pos, src);
    const msf_u32 E = bpos + __shfl_sync(FULL, L.len, src);
    msf_u32 q = E, round = 0;
    bool found = false;
    while (true) {
      const msf_u32 top = E - 32u * round;                // entries [top - 32, top); top > pos + 1
      const msf_u32 idx = top - 32u + lane;               // wraps below zero when top < 32: probe rejects it
      const msf_u32 r = msf_skip_probe(P, bpos, E, idx);
      const unsigned fail = __ballot_sync(FULL, r == 1u), hold = __ballot_sync(FULL, r == 2u);
      if (fail) {                                         // q = just above the highest failing entry
        q = top - 32u + (31 - __clz((int)fail)) + 1;
        found = true;
        break;
      }
      if (hold) q = top - 32u + (__ffs(hold) - 1);
      if (top <= bpos + 33u) { found = true; break; }     // the window reached pos + 1
      if (++round >= MSF_SKIP_ROUNDS) break;
    }
    if (lane == src) {
      if (!found) q = msf_skip_bisect(P, bpos, E, q);
      msf_skip_to(P, L, q);
    }
  }
}
#else
// ---- one warp iteration (host emulation for tests/harness: same phases, loops instead of shuffles)
// how often each action came out of a search step (harness statistics: [1] FIN, [2] GENERIC, [3] TERM,
// [4] EXT, [5] SKIP)
static unsigned long long msf_host_acts[8];
inline void msf_warp_iter_host(const MsfParams &P, MsfLane *W, int nl) {
  msf_u32 act[32];
  for (int l = 0; l < nl; l++) {
    act[l] = msf_search_step(P, W[l]);
    msf_host_acts[act[l] == MSF_ACT_FIN ? 1 : (act[l] & MSF_ACT_GENERIC) ? 2 : (act[l] & MSF_ACT_TERM) ? 3 :
                  act[l] == MSF_ACT_EXT ? 4 : 0]++;
  }
  for (int l = 0; l < nl; l++) {                          // phase B
    if (act[l] != MSF_ACT_EXT) continue;
    MsfLane &L = W[l];
    msf_u32 blen = L.len, res;
    const msf_u32 bmax = msf_ext_max(P, L.limit, L.pos);
    while (true) {
      res = 0xFFFFFFFFu;
      for (msf_u32 lane = 0; lane < 32; lane++) {
        const msf_u32 off = blen + 32u * lane;
        if (off < bmax) {
          msf_u32 cand = msf_ext_probe(P, L.i, L.pos, off);
          if (cand < res) res = cand;
        }
      }
      if (res != 0xFFFFFFFFu || blen + 1024u >= bmax) break;
      blen += 1024u;
    }
    msf_ext_done(P, L, res, bmax);
    act[l] = MSF_ACT_FIN;
  }
  for (int l = 0; l < nl; l++) {                          // phase C
    MsfLane &L = W[l];
    if (act[l] & MSF_ACT_GENERIC) msf_generic_search(P, L.i, L.limit, L.pos, L.len, L.sm);
    if (act[l] & MSF_ACT_ESC) msf_escape_search(P, L.i, L.pos, L.len, L.sm);
    if (act[l]) act[l] = msf_fin(P, L, (act[l] & MSF_ACT_TERM) != 0, (act[l] & MSF_ACT_ESC) != 0);
  }
  for (int l = 0; l < nl; l++) {                          // phase D
    if (act[l] != MSF_ACT_SKIP) continue;
    msf_host_acts[5]++;
    MsfLane &L = W[l];
    const msf_u32 bpos = L.pos, E = L.pos + L.len;
    msf_u32 q = E, round = 0;
    bool found = false;
    while (true) {
      const msf_u32 top = E - 32u * round;
      msf_u32 fail = 0, hold = 0;
      for (msf_u32 lane = 0; lane < 32; lane++) {
        msf_u32 r = msf_skip_probe(P, bpos, E, top - 32u + lane);
        if (r == 1u) fail |= 1u << lane;
        if (r == 2u) hold |= 1u << lane;
      }
      if (fail) { q = top - 32u + (31 - (msf_u32)__builtin_clz(fail)) + 1; found = true; break; }
      if (hold) q = top - 32u + (msf_u32)__builtin_ctz(hold);
      if (top <= bpos + 33u) { found = true; break; }
      if (++round >= MSF_SKIP_ROUNDS) break;
    }
    if (!found) q = msf_skip_bisect(P, bpos, E, q);
    msf_skip_to(P, L, q);
  }
}
#endif

// ---- window fix-up (head test across window borders) ------------------------------------------
// The first position of a window that starts inside a document is recorded tentatively; it is a
// head iff its pos differs from the previous position's pos + 1 (:1154), which the previous
// window reports in wlastpos.  Returns 1 when the tentative record has to be dropped.
MSF_HD msf_u32 msf_window_drop(const msf_u32 *wfirstpos, const msf_u32 *wlastpos, msf_u64 w) {
  if (w == 0) return 0;
  msf_u32 fp = wfirstpos[w];
  if (fp == MSF_INVALID) return 0;                 // window starts at a document start / terminator
  msf_u32 lp = wlastpos[w - 1];
  return (lp != MSF_INVALID && fp == lp + 1) ? 1u : 0u;
}

// number of k-mers <= the suffix s of the 8-bit augmented reference (non-decreasing in SA order);
// ktab[m] = first rank r with m < f(r)   (index.cu: k_ktab_mark + max-scan)
MSF_HD msf_u32 msf_kmers_le_suffix(const msf_u8 *x, msf_u32 s, int k) {
  msf_u32 m = 0;
  for (int j = 0; j < k; j++) {
    msf_u8 c = x[s + j];                // x[n-1] = 0 stops the walk before the padding
    int d = c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : c == 'T' ? 3 : -1;
    if (d < 0) {
      int dl = (c > 'A') + (c > 'C') + (c > 'G') + (c > 'T');
      return (m * 4 + dl) << (2 * (k - 1 - j));
    }
    m = m * 4 + d;
  }
  return m + 1;
}
