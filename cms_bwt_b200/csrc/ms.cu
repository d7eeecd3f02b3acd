// Stage 1: chunk-parallel matching statistics + phrase-head emission.
// Replaces the per-document MS driver (CMS-BWT-functions.cpp:1138-1213, small path :344-419):
// computeMSFactorAt :45-88, binarySearchLB/RB :94-149, contractLeft/adjustInterval :11-43 and the
// PLCP' skip loop :1176-1185.  MS is a pure function of the suffix, so every
// chunk is evaluated independently: "hard" positions are searched from scratch (k-mer interval
// table + exact binary search, one position per lane), long matches are extended by the whole
// warp and the non-head run that follows is skipped through G = q + PLCP'[q] (:1176-1185).
#include "common.cuh"
#include <cub/cub.cuh>

#define NOB 0xFFFFFFFFu      // no batch yet
#define BADB 0xFFFFFFFEu     // pool exhausted
#define INVALID_POS 0xFFFFFFFFu

struct MSParams {
  Index ix;
  const u8 *text;
  u64 N;
  u32 chunk_len;
  u64 nchunks;
  u64 *pg; u32 *pp; u32 *pl; u16 *pc;      // head pool, batches of 32 records
  u32 *batch_next;
  u32 nb_cap;
  u32 *nb_used;
  u32 *chunk_first, *chunk_cnt;
  u32 *flags;                              // [0] overflow, [1] bad symbol
  unsigned long long *chunk_counter;
  u32 *dpos, *dlen; u8 *dsm, *dhead;       // optional per-position dump
};

struct Hints { u32 s[4]; u32 m[4]; int n; };

__device__ __forceinline__ u32 lce_cap(const u8 *t, const u8 *xs) {
#pragma unroll
  for (int o = 0; o < CMS_CAP; o += 8) {
    u64 d = ld8u(t + o) ^ ld8u(xs + o);
    if (d) return o + ((__ffsll((long long)d) - 1) >> 3);
  }
  return CMS_CAP;
}

// full LCE of t[0..) and xs[0..) by the whole warp; result clamped to min(xmax, tmax)
__device__ __forceinline__ u32 warp_lce(const u8 *t, const u8 *xs, u32 xmax, u64 tmax) {
  const int lane = threadIdx.x & 31;
  u32 lim = (u64)xmax < tmax ? xmax : (u32)tmax;
  u32 o = 0;
  while (true) {
    u32 off = o + lane * 8;
    u64 d = ~0ull;
    if (off < lim) d = ld8u(t + off) ^ ld8u(xs + off);
    unsigned b = __ballot_sync(0xffffffffu, d != 0);
    if (b) {
      int l = __ffs(b) - 1;
      u64 dl = __shfl_sync(0xffffffffu, d, l);
      u32 m = o + l * 8 + ((__ffsll((long long)dl) - 1) >> 3);
      return m < lim ? m : lim;
    }
    o += 256;
  }
}

// insertion-point range from the k-mer table: [L, R] (whole SA when the k-mer is not pure ACGT)
__device__ __forceinline__ void kmer_range(const Index &ix, const u8 *t, u32 &L, u32 &R) {
  u64 w0 = ld8u(t), w1 = ld8u(t + 8);
  u32 code = 0;
  bool ok = true;
  for (int j = 0; j < ix.k; j++) {
    u32 c = (u32)((j < 8 ? (w0 >> (8 * j)) : (w1 >> (8 * (j - 8)))) & 0xFF);
    int d = c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : c == 'T' ? 3 : -1;
    if (d < 0) { ok = false; break; }
    code = code * 4 + d;
  }
  L = 0; R = ix.n;
  if (ok) { L = __ldg(ix.ktab + code); R = __ldg(ix.ktab + code + 1); }
}

// One probe: LCE of the text suffix with reference suffix s, and whether x-suffix < text suffix.
// Sets und (and s_und) when the per-lane budget is exhausted and no hint covers s.
__device__ __forceinline__ bool probe(const Index &ix, const u8 *t, u32 s, const Hints &h, u32 &m,
                                      bool &xless, u32 &s_und) {
  bool hit = false;
#pragma unroll
  for (int q = 0; q < 4; q++)
    if (q < h.n && h.s[q] == s) { m = h.m[q]; hit = true; }
  if (!hit) {
    m = lce_cap(t, ix.x + s);
    if (m >= CMS_CAP) { s_und = s; return false; }
  }
  u32 rem = ix.n - s;
  if (m >= rem) { m = rem; xless = true; }
  else xless = ix.x[s + m] < t[m];
  return true;
}

// Exact MS of the text suffix at t by one lane.  Returns false when a probe was undecided.
// Neighbour rule of :53-76: the successor wins ties (smaller = 1).
__device__ bool lane_search(const Index &ix, const u8 *t, const Hints &h, u32 &pos, u32 &len,
                            u32 &sm, u32 &s_und) {
  u32 lo, hi;
  kmer_range(ix, t, lo, hi);
  u32 m; bool xless;
  while (lo < hi) {
    u32 mid = lo + ((hi - lo) >> 1);
    u32 s = (u32)__ldg(ix.sa + mid);
    if (!probe(ix, t, s, h, m, xless, s_und)) return false;
    if (xless) lo = mid + 1; else hi = mid;
  }
  i64 lp = -1, ls = -1;
  u32 ppos = 0, spos = 0;
  if (lo > 0) {
    ppos = (u32)__ldg(ix.sa + lo - 1);
    if (!probe(ix, t, ppos, h, m, xless, s_und)) return false;
    lp = m;
  }
  if (lo < ix.n) {
    spos = (u32)__ldg(ix.sa + lo);
    if (!probe(ix, t, spos, h, m, xless, s_und)) return false;
    ls = m;
  }
  if (ls >= lp) { pos = spos; len = (u32)ls; sm = 1; }
  else { pos = ppos; len = (u32)lp; sm = 0; }
  return true;
}

// Same search with every comparison done by the whole warp (no budget); uniform across lanes.
__device__ void warp_search(const Index &ix, const u8 *t, u64 tmax, u32 &pos, u32 &len, u32 &sm) {
  u32 lo, hi;
  kmer_range(ix, t, lo, hi);
  while (lo < hi) {
    u32 mid = lo + ((hi - lo) >> 1);
    u32 s = (u32)__ldg(ix.sa + mid);
    u32 m = warp_lce(t, ix.x + s, ix.n - s, tmax);
    bool xless = (m >= ix.n - s) || ix.x[s + m] < t[m];
    if (xless) lo = mid + 1; else hi = mid;
  }
  i64 lp = -1, ls = -1;
  u32 ppos = 0, spos = 0;
  if (lo > 0) { ppos = (u32)__ldg(ix.sa + lo - 1); lp = warp_lce(t, ix.x + ppos, ix.n - ppos, tmax); }
  if (lo < ix.n) { spos = (u32)__ldg(ix.sa + lo); ls = warp_lce(t, ix.x + spos, ix.n - spos, tmax); }
  if (ls >= lp) { pos = spos; len = (u32)ls; sm = 1; }
  else { pos = ppos; len = (u32)lp; sm = 0; }
}

// first q in (pos, E) with G[q] >= E, else E  (G is non-decreasing); uniform across the warp
__device__ u32 first_G_ge(const u32 *G, u32 pos, u32 E) {
  const int lane = threadIdx.x & 31;
  u32 lo = pos + 1, hi = E;
  for (int it = 0; it < 2 && hi > lo; it++) {
    u32 base = hi - lo > 32 ? hi - 32 : lo;
    u32 q = base + lane;
    bool f = q < hi && __ldg(G + q) >= E;
    unsigned b = __ballot_sync(0xffffffffu, f);
    if (b == 0) return hi;
    int fl = __ffs(b) - 1;
    if (fl > 0 || base == lo) return base + fl;
    hi = base;
  }
  while (lo < hi) {
    u32 mid = lo + ((hi - lo) >> 1);
    if (__ldg(G + mid) >= E) hi = mid; else lo = mid + 1;
  }
  return lo;
}

__device__ __forceinline__ void warp_emit(const MSParams &P, bool has, u64 g, u32 pos, u32 len,
                                          u16 cs, u32 &curb, u32 &fill, u32 &total, u64 chunk) {
  const int lane = threadIdx.x & 31;
  unsigned m = __ballot_sync(0xffffffffu, has);
  if (!m) return;
  int cnt = __popc(m);
  int rk = __popc(m & ((1u << lane) - 1));
  int done = 0;
  while (done < cnt) {
    if (curb == NOB || fill == 32) {
      u32 nb = 0;
      if (lane == 0) nb = atomicAdd(P.nb_used, 1u);
      nb = __shfl_sync(0xffffffffu, nb, 0);
      if (nb >= P.nb_cap) {
        if (lane == 0) P.flags[0] = 1;
        nb = BADB;
      } else if (lane == 0) {
        if (curb == NOB) P.chunk_first[chunk] = nb;
        else if (curb != BADB) P.batch_next[curb] = nb;
      }
      curb = nb;
      fill = 0;
    }
    int take = min(32 - (int)fill, cnt - done);
    if (has && rk >= done && rk < done + take && curb != BADB) {
      u64 slot = (u64)curb * 32 + fill + (rk - done);
      P.pg[slot] = g; P.pp[slot] = pos; P.pl[slot] = len; P.pc[slot] = cs;
    }
    fill += take;
    done += take;
  }
  total += cnt;
}

__global__ void __launch_bounds__(128) k_ms(MSParams P) {
  const int lane = threadIdx.x & 31;
  const Index &ix = P.ix;
  const u32 n = ix.n;
  while (true) {
    unsigned long long chunk = 0;
    if (lane == 0) chunk = atomicAdd(P.chunk_counter, 1ull);
    chunk = __shfl_sync(0xffffffffu, chunk, 0);
    if (chunk >= P.nchunks) break;
    const u64 c0 = chunk * (u64)P.chunk_len;
    const u64 c1 = min(P.N, c0 + P.chunk_len);
    u64 i = c0;
    u32 prevpos = INVALID_POS;
    bool prevterm = true;          // previous position closes a document (or does not exist)
    bool skipFirst = false;
    if (c0 > 0 && P.text[c0 - 1] != 2) { i = c0 - 1; skipFirst = true; prevterm = false; }
    u32 curb = NOB, fill = 0, total = 0;
    while (i < c1) {
      const u64 p = i + lane;
      const bool act = p < c1;
      const u8 t0 = act ? P.text[p] : (u8)2;
      const bool term = act && t0 == 2;
      u32 pos = n - 1, len = 0, sm = 0, s_und = 0;
      bool und = false;
      Hints h; h.n = 0;
#pragma unroll
      for (int q = 0; q < 4; q++) { h.s[q] = 0; h.m[q] = 0; }
      if (act && !term) und = !lane_search(ix, P.text + p, h, pos, len, sm, s_und);
      unsigned um = __ballot_sync(0xffffffffu, und);
      int nvalid = (int)min((u64)32, c1 - i);
      int J = -1;
      if (um) {
        J = __ffs(um) - 1;
        nvalid = J + 1;
        const u64 pJ = i + J;
        const u8 *tJ = P.text + pJ;
        bool pending = true;
        while (pending) {
          u32 s = __shfl_sync(0xffffffffu, s_und, J);
          if (h.n == 4) {           // too many long candidates: fully cooperative search
            u32 a, b, c2;
            warp_search(ix, tJ, P.N - pJ, a, b, c2);
            if (lane == J) { pos = a; len = b; sm = c2; }
            break;
          }
          u32 m = warp_lce(tJ, ix.x + s, n - s, P.N - pJ);
#pragma unroll
          for (int q = 0; q < 4; q++) if (q == h.n) { h.s[q] = s; h.m[q] = m; }
          h.n++;
          bool ok = true;
          if (lane == J) ok = lane_search(ix, tJ, h, pos, len, sm, s_und);
          pending = !__shfl_sync(0xffffffffu, ok, J);
        }
      }
      // head flags
      u32 pv = __shfl_up_sync(0xffffffffu, pos, 1);
      bool pterm = __shfl_up_sync(0xffffffffu, term, 1);
      if (lane == 0) { pv = prevpos; pterm = prevterm; }
      const bool valid = lane < nvalid;
      const bool head = valid && !term && (pterm || pos != pv + 1);
      if (valid && !term && len == 0) P.flags[1] = 1;
      const bool rec = valid && (head || term) && !(skipFirst && lane == 0);
      u8 ch = 2;
      if (rec && !pterm) ch = P.text[p - 1];
      warp_emit(P, rec, p, pos, len, (u16)(ch | (sm << 8)), curb, fill, total, chunk);
      if (P.dpos && valid && !(skipFirst && lane == 0)) {
        P.dpos[p] = pos; P.dlen[p] = len; P.dsm[p] = (u8)sm; P.dhead[p] = (u8)(head || term);
      }
      skipFirst = false;
      if (J >= 0) {
        const u32 posJ = __shfl_sync(0xffffffffu, pos, J);
        const u32 lenJ = __shfl_sync(0xffffffffu, len, J);
        const u32 smJ = __shfl_sync(0xffffffffu, sm, J);
        const u64 pJ = i + J;
        u32 skip = 0;
        if (lenJ > 1) skip = first_G_ge(ix.G, posJ, posJ + lenJ) - posJ - 1;
        if (P.dpos)
          for (u32 k2 = 1 + lane; k2 <= skip; k2 += 32) {
            P.dpos[pJ + k2] = posJ + k2; P.dlen[pJ + k2] = lenJ - k2;
            P.dsm[pJ + k2] = (u8)smJ; P.dhead[pJ + k2] = 0;
          }
        i = pJ + 1 + skip;
        prevpos = posJ + skip;
        prevterm = false;
      } else {
        prevpos = __shfl_sync(0xffffffffu, pos, nvalid - 1);
        prevterm = __shfl_sync(0xffffffffu, term, nvalid - 1);
        i += nvalid;
      }
    }
    if (lane == 0) P.chunk_cnt[chunk] = total;
  }
}

// gather the linked batches of every chunk into dense text-order arrays
__global__ void k_compact(const u64 *pg, const u32 *pp, const u32 *pl, const u16 *pc,
                          const u32 *batch_next, const u32 *chunk_first, const u32 *chunk_cnt,
                          const u64 *chunk_off, u64 nchunks, u64 *g, u32 *pos, u32 *len, u16 *cs) {
  const int lane = threadIdx.x & 31;
  u64 c = ((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (c >= nchunks) return;
  u32 cnt = chunk_cnt[c];
  u64 off = chunk_off[c];
  u32 b = cnt ? chunk_first[c] : 0;
  for (u32 k = 0; k < cnt; k += 32) {
    if (k + lane < cnt) {
      u64 s = (u64)b * 32 + lane;
      g[off + k + lane] = pg[s]; pos[off + k + lane] = pp[s];
      len[off + k + lane] = pl[s]; cs[off + k + lane] = pc[s];
    }
    if (k + 32 < cnt) b = batch_next[b];
  }
}

void ms_stage(cmsbwt_ctx *c) {
  cudaStream_t st = c->st;
  c->stats.fast_path = 0;
  c->stats.n_escapes = 0;
  c->stats.pack_ms = 0;
  if (c->fast && c->pure && c->chunk_len <= 0x7FFFFFFF && ms_stage_fast(c)) return;
  const u64 N = c->N;
  const u32 CH = c->chunk_len ? (u32)c->chunk_len : 8192u;
  const u64 nchunks = (N + CH - 1) / CH;
  if (nchunks >= 0xFFFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "too many chunks"};
  DBuf cfirst, ccnt, coff, flags, counter, nbused, tmp;
  cfirst.alloc(nchunks * 4, st); ccnt.alloc(nchunks * 4, st); coff.alloc((nchunks + 1) * 8, st);
  flags.alloc(8, st); counter.alloc(8, st); nbused.alloc(4, st);
  if (c->dump) {
    c->dump_pos.alloc(N * 4, st); c->dump_len.alloc(N * 4, st);
    c->dump_sm.alloc(N, st); c->dump_head.alloc(N, st);
  }
  size_t tb = 0;
  EXSUM64(nullptr, tb, ccnt.as<u32>(), coff.as<u64>(), (int)(nchunks + 1), st);
  tmp.alloc(tb, st);
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->dev);
  u64 rec_cap = N / (u64)c->pool_div + 32 * nchunks + 1024;
  for (int attempt = 0; attempt < 2; attempt++) {
    u64 nb = (rec_cap + 31) / 32;
    if (nb >= 0xFFFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "head pool too large"};
    DBuf pg, pp, pl, pc, bnext;
    pg.alloc(nb * 32 * 8, st); pp.alloc(nb * 32 * 4, st); pl.alloc(nb * 32 * 4, st);
    pc.alloc(nb * 32 * 2, st); bnext.alloc(nb * 4, st);
    CK(cudaMemsetAsync(flags.p, 0, 8, st));
    CK(cudaMemsetAsync(counter.p, 0, 8, st));
    CK(cudaMemsetAsync(nbused.p, 0, 4, st));
    CK(cudaMemsetAsync(ccnt.p, 0, nchunks * 4, st));
    MSParams P;
    P.ix = c->ix; P.text = c->text; P.N = N; P.chunk_len = CH; P.nchunks = nchunks;
    P.pg = pg.as<u64>(); P.pp = pp.as<u32>(); P.pl = pl.as<u32>(); P.pc = pc.as<u16>();
    P.batch_next = bnext.as<u32>(); P.nb_cap = (u32)nb; P.nb_used = nbused.as<u32>();
    P.chunk_first = cfirst.as<u32>(); P.chunk_cnt = ccnt.as<u32>(); P.flags = flags.as<u32>();
    P.chunk_counter = counter.as<unsigned long long>();
    P.dpos = c->dump ? c->dump_pos.as<u32>() : nullptr; P.dlen = c->dump_len.as<u32>();
    P.dsm = c->dump_sm.as<u8>(); P.dhead = c->dump_head.as<u8>();
    u64 want_blocks = (nchunks + 3) / 4;
    u32 grid = (u32)min(want_blocks, (u64)sms * 12);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, st);
    { GroupSpan g(c, CMSBWT_G_MS);
    k_ms<<<grid, 128, 0, st>>>(P); }
    cudaEventRecord(e1, st);
    LAUNCH(c);
    CK(cudaGetLastError());
    // chunk_off = exclusive scan over nchunks+1 entries (ccnt has nchunks; the extra reads 0)
    GroupSpan ggather(c, CMSBWT_G_GATHER);
    size_t t = tb;
    CK(EXSUM64(tmp.p, t, ccnt.as<u32>(), coff.as<u64>(), (int)nchunks, st));
    LAUNCH(c);
    u32 hf[2]; u64 lastoff = 0; u32 lastcnt = 0;
    CK(cudaMemcpyAsync(hf, flags.p, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&lastoff, coff.as<u64>() + (nchunks - 1), 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&lastcnt, ccnt.as<u32>() + (nchunks - 1), 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    { float ms = 0; cudaEventElapsedTime(&ms, e0, e1); c->stats.ms_kernel_ms = ms; }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (hf[1]) throw CmsErr{CMSBWT_E_SYMBOL, "text holds a byte that does not occur in the augmented reference"};
    const u64 H = lastoff + lastcnt;
    if (hf[0]) {                 // pool exhausted: counts are exact, rerun with the exact size
      if (attempt == 1) throw CmsErr{CMSBWT_E_INTERNAL, "head pool overflow on exact-size rerun"};
      rec_cap = H + 32 * nchunks + 1024;
      continue;
    }
    if (H >= 0x7FFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "more than 2^31 heads on one device"};
    c->H = H;
    c->h_gstart.alloc((H + 1) * 8, st); c->h_pos.alloc((H + 1) * 4, st);
    c->h_len.alloc((H + 1) * 4, st); c->h_cs.alloc((H + 1) * 2, st);
    u64 threads = nchunks * 32;
    k_compact<<<(u32)((threads + 255) / 256), 256, 0, st>>>(
        pg.as<u64>(), pp.as<u32>(), pl.as<u32>(), pc.as<u16>(), bnext.as<u32>(), cfirst.as<u32>(),
        ccnt.as<u32>(), coff.as<u64>(), nchunks, c->h_gstart.as<u64>(), c->h_pos.as<u32>(),
        c->h_len.as<u32>(), c->h_cs.as<u16>());
    LAUNCH(c);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(st));
    break;
  }
  c->stats.n_heads = c->H;
}
