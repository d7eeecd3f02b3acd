// Stage 2/3, brute-force tile path: how many suffixes of its bucket precede every head
// (counterSmallerThanHead, CMS-BWT-functions.cpp:1517-1603 / small path :709-902), the bucket sizes
// (:368-390 as interval coverage) and the placement of the head symbols (:1640-1688).
//
// The reference walks every non-head of every phrase and does an ordered-map lower_bound per
// incidence.  Here the reference positions are cut into tiles of CMS_LT_W positions; every head is
// filed under the tile of its pos, every long phrase (toNext > 0) under each tile its non-head
// run [pos+1, pos+toNext] intersects.  One CTA per tile then
//   pass A (k_tile_cov):    bucket sizes cov[p] = heads at p + ranges covering p, coalesced
//   pass B (k_tile_place):  within(h) = #{heads at pos_h with smaller kappa} + #{ranges covering
//                           pos_h with smaller kappa}; the head symbol goes to
//                           out[base[ISA[pos_h]] + within(h)].
// No sort is involved; the pair tests are O(heads x depth) and run out of shared memory.  When a
// collection is so deep that this would cost more than `tile_budget` pair tests, or a tile does not
// fit shared memory, tiles_build returns false and the sort-based sweep of heads.cu is used.
#include "common.cuh"
#include <cub/cub.cuh>
#include <string.h>
#include <stdlib.h>

#define LT_SHIFT 8
#define LT_W (1u << LT_SHIFT)
#define LT_THREADS 256
#define LT_MAX_ENTRIES 11000          // heads + ranges of one tile that fit shared memory (16 B each)

struct __align__(16) TileEnt { u32 a, b; u64 kappa; };   // head: a = id, b = q | ch << 8; range: a = lo | hi << 16

// ---- histogram of entries per tile ---------------------------------------------------------------
__global__ void k_tile_hist(const u32 *__restrict__ pos, const u32 *__restrict__ len, const u32 *__restrict__ toNext, u64 H,
                            u32 *hcnt, u32 *rcnt) {
  const int lane = threadIdx.x & 31;
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  bool in = i < H && len[i] != 0;                  // sentinels (len 0) are placed by document index
  u32 p = in ? pos[i] : 0, tn = in ? toNext[i] : 0;
  if (in) atomicAdd(hcnt + (p >> LT_SHIFT), 1u);
  u32 ta = (p + 1) >> LT_SHIFT, tb = (p + tn) >> LT_SHIFT;
  u32 nt = (in && tn) ? tb - ta + 1 : 0;
  if (nt && nt <= 8)
    for (u32 t = ta; t <= tb; t++) atomicAdd(rcnt + t, 1u);
  unsigned big = __ballot_sync(0xffffffffu, nt > 8);
  while (big) {
    int src = __ffs(big) - 1;
    big &= big - 1;
    u32 a = __shfl_sync(0xffffffffu, ta, src), m = __shfl_sync(0xffffffffu, nt, src);
    for (u32 k = lane; k < m; k += 32) atomicAdd(rcnt + a + k, 1u);
  }
}

__device__ __forceinline__ void put_range(TileEnt *re, const u32 *roff, u32 *rcur, u32 t, u32 p, u32 tn,
                                          u64 kap) {
  u32 a = t << LT_SHIFT;
  u32 lo = max(p + 1, a) - a, hi = min(p + tn, a + LT_W - 1) - a;
  u32 slot = atomicAdd(rcur + t, 1u);            // cursors start at the tile's offset (tiles_scatter)
  TileEnt e; e.a = lo | (hi << 16); e.b = 0; e.kappa = kap;
  re[slot] = e;
}

// TS_ILP heads per thread: the slot atomics of all of them are in flight together (the kernel is
// bound by the latency of atomic -> dependent store, not by bandwidth)
template <int TS_ILP, int MINB>
__global__ void __launch_bounds__(256, MINB) k_tile_scatter(const u32 *__restrict__ pos, const u32 *__restrict__ len,
                                                      const u16 *__restrict__ cs, const u32 *__restrict__ toNext,
                                                      const u64 *__restrict__ kappa, u64 H, const u32 *hoff,
                                                      const u32 *roff, u32 *hcur, u32 *rcur, TileEnt *__restrict__ he,
                                                      TileEnt *__restrict__ re, const u32 *__restrict__ tail_incl,
                                                      const u32 *__restrict__ trank, const u32 *__restrict__ sexcl,
                                                      u8 *termsym) {
  // kappa == NULL: kappa = (cls, rank of the phrase's tail) is formed here (heads.cu k_kappa_count) instead of
  // being written by one kernel and read back by this one; tail_incl is monotone, so the rank reads are
  // nearly sequential
  const int lane = threadIdx.x & 31;
  const u64 i0 = (u64)blockIdx.x * (blockDim.x * TS_ILP) + threadIdx.x;
  u32 p[TS_ILP], tn[TS_ILP], c8[TS_ILP], slot[TS_ILP];
  u64 kap[TS_ILP];
  bool in[TS_ILP];
#pragma unroll
  for (int k = 0; k < TS_ILP; k++) {
    const u64 i = i0 + (u64)k * blockDim.x;
    const u32 ln = i < H ? len[i] : 0;
    in[k] = ln != 0;
    // a terminator (len 0): its symbol goes to the document's place in the bucket of SA rank 0 (:1640-1643)
    if (termsym && i < H && ln == 0) termsym[sexcl[i]] = (u8)(cs[i] & 0xFF);
    const u32 c16 = in[k] ? (u32)cs[i] : 0;
    p[k] = in[k] ? pos[i] : 0; tn[k] = in[k] ? toNext[i] : 0;
    c8[k] = c16 & 0xFF;
    if (kappa) kap[k] = in[k] ? kappa[i] : 0;
    else kap[k] = in[k] ? ((u64)cls_of(p[k], ln, c16 >> 8) << 32) | trank[tail_incl[i] - 1] : 0;
  }
#pragma unroll
  for (int k = 0; k < TS_ILP; k++)
    if (in[k]) { const u32 t = p[k] >> LT_SHIFT; slot[k] = atomicAdd(hcur + t, 1u); }
#pragma unroll
  for (int k = 0; k < TS_ILP; k++)
    if (in[k]) {
      TileEnt e; e.a = (u32)(i0 + (u64)k * blockDim.x); e.b = (p[k] & (LT_W - 1)) | (c8[k] << 8); e.kappa = kap[k];
      he[slot[k]] = e;
    }
#pragma unroll
  for (int k = 0; k < TS_ILP; k++) {
    const u32 ta = (p[k] + 1) >> LT_SHIFT, tb = (p[k] + tn[k]) >> LT_SHIFT;
    const u32 nt = (in[k] && tn[k]) ? tb - ta + 1 : 0;
    if (nt && nt <= 8) {
      // all slot reservations of this phrase first, then its stores: one round trip instead of one per tile
      u32 sl[8];
#pragma unroll
      for (int j = 0; j < 8; j++)
        if ((u32)j < nt) sl[j] = atomicAdd(rcur + ta + j, 1u);
#pragma unroll
      for (int j = 0; j < 8; j++)
        if ((u32)j < nt) {
          const u32 a = (ta + j) << LT_SHIFT;
          TileEnt e; e.a = (max(p[k] + 1, a) - a) | ((min(p[k] + tn[k], a + LT_W - 1) - a) << 16); e.b = 0; e.kappa = kap[k];
          re[sl[j]] = e;
        }
    }
    unsigned big = __ballot_sync(0xffffffffu, nt > 8);
    while (big) {
      int src = __ffs(big) - 1;
      big &= big - 1;
      u32 a = __shfl_sync(0xffffffffu, ta, src), m = __shfl_sync(0xffffffffu, nt, src);
      u32 bp = __shfl_sync(0xffffffffu, p[k], src), btn = __shfl_sync(0xffffffffu, tn[k], src);
      u64 bk = __shfl_sync(0xffffffffu, kap[k], src);
      for (u32 kk = lane; kk < m; kk += 32) put_range(re, roff, rcur, a + kk, bp, btn, bk);
    }
  }
}

// ---- the same two kernels for FEW tiles (short reference): every global counter would take the
// atomics of H / ntiles heads one after the other (measured: 11 + 12 ms for 58 M heads on 118 tiles),
// so a block first counts its slice of heads in shared memory and touches every global counter once.
#define TA_MAX_TILES 4096u
#define TA_SLICE 2048u
template <bool COUNT>
__device__ __forceinline__ void ta_ranges(u32 p, u32 tn, u32 *s_r, const u32 *s_rb, const u32 *roff, TileEnt *re,
                                          u64 kap) {
  const u32 ta = (p + 1) >> LT_SHIFT, tb = (p + tn) >> LT_SHIFT;
  for (u32 t = ta; t <= tb; t++) {
    const u32 k = atomicAdd(&s_r[t], 1u);
    if (!COUNT) {
      const u32 a = t << LT_SHIFT;
      TileEnt e; e.a = (max(p + 1, a) - a) | ((min(p + tn, a + LT_W - 1) - a) << 16); e.b = 0; e.kappa = kap;
      re[roff[t] + s_rb[t] + k] = e;
    }
  }
}
__global__ void __launch_bounds__(256) k_tile_hist_few(const u32 *pos, const u32 *len, const u32 *toNext, u64 H,
                                                       u32 ntiles, u32 *hcnt, u32 *rcnt) {
  extern __shared__ u32 s_ta[];
  u32 *s_h = s_ta, *s_r = s_ta + ntiles;
  for (u32 k = threadIdx.x; k < 2 * ntiles; k += blockDim.x) s_ta[k] = 0;
  __syncthreads();
  const u64 i0 = (u64)blockIdx.x * TA_SLICE, i1 = min(H, i0 + TA_SLICE);
  for (u64 i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
    if (len[i] == 0) continue;
    const u32 p = pos[i], tn = toNext[i];
    atomicAdd(&s_h[p >> LT_SHIFT], 1u);
    if (tn) ta_ranges<true>(p, tn, s_r, nullptr, nullptr, nullptr, 0);
  }
  __syncthreads();
  for (u32 k = threadIdx.x; k < ntiles; k += blockDim.x) {
    if (s_h[k]) atomicAdd(hcnt + k, s_h[k]);
    if (s_r[k]) atomicAdd(rcnt + k, s_r[k]);
  }
}
__global__ void __launch_bounds__(256) k_tile_scatter_few(const u32 *pos, const u32 *len, const u16 *cs,
                                                          const u32 *toNext, const u64 *kappa, u64 H, u32 ntiles,
                                                          const u32 *hoff, const u32 *roff, u32 *hcur, u32 *rcur,
                                                          TileEnt *he, TileEnt *re) {
  extern __shared__ u32 s_ta[];
  u32 *s_h = s_ta, *s_r = s_ta + ntiles, *s_hb = s_ta + 2 * ntiles, *s_rb = s_ta + 3 * ntiles;
  for (u32 k = threadIdx.x; k < 2 * ntiles; k += blockDim.x) s_ta[k] = 0;
  __syncthreads();
  const u64 i0 = (u64)blockIdx.x * TA_SLICE, i1 = min(H, i0 + TA_SLICE);
  for (u64 i = i0 + threadIdx.x; i < i1; i += blockDim.x) {          // pass 1: the slice's counts
    if (len[i] == 0) continue;
    const u32 p = pos[i], tn = toNext[i];
    atomicAdd(&s_h[p >> LT_SHIFT], 1u);
    if (tn) ta_ranges<true>(p, tn, s_r, nullptr, nullptr, nullptr, 0);
  }
  __syncthreads();
  for (u32 k = threadIdx.x; k < ntiles; k += blockDim.x) {           // one reservation per tile
    s_hb[k] = s_h[k] ? atomicAdd(hcur + k, s_h[k]) : 0u;
    s_rb[k] = s_r[k] ? atomicAdd(rcur + k, s_r[k]) : 0u;
    s_h[k] = 0; s_r[k] = 0;
  }
  __syncthreads();
  for (u64 i = i0 + threadIdx.x; i < i1; i += blockDim.x) {          // pass 2: the entries
    if (len[i] == 0) continue;
    const u32 p = pos[i], tn = toNext[i], t = p >> LT_SHIFT;
    const u64 kap = kappa[i];
    TileEnt e; e.a = (u32)i; e.b = (p & (LT_W - 1)) | ((u32)(cs[i] & 0xFF) << 8); e.kappa = kap;
    he[hoff[t] + s_hb[t] + atomicAdd(&s_h[t], 1u)] = e;
    if (tn) ta_ranges<false>(p, tn, s_r, s_rb, roff, re, kap);
  }
}

// ---- pass A: bucket sizes + cost of pass B -------------------------------------------------------
// pcov != NULL (single-GPU pull emit, emit.cu): pcov[p] = {bucket size without the terminators, index of
// the position's first head in the position-grouped head array hw[] that k_tile_place fills}, for p <= n.
__global__ void __launch_bounds__(LT_THREADS) k_tile_cov(const TileEnt *he, const TileEnt *re, const u32 *hoff,
                                                         const u32 *roff, u32 n, u32 ndocs, u32 t0, u32 *cov,
                                                         uint2 *pcov, unsigned long long *work, u32 *maxent) {
  __shared__ u32 s_h[LT_W];
  __shared__ int s_d[LT_W + 1];
  typedef cub::BlockScan<int, LT_THREADS> BS;
  typedef cub::BlockReduce<unsigned long long, LT_THREADS> BR;
  __shared__ typename BS::TempStorage ts;
  __shared__ typename BS::TempStorage ts2;
  __shared__ typename BR::TempStorage tr;
  // t0 = first tile of this device (multi-GPU: the entry arrays and cov are local to its tile range)
  const u32 t = blockIdx.x, tid = threadIdx.x;
  const u32 h0 = hoff[t], h1 = hoff[t + 1], r0 = roff[t], r1 = roff[t + 1];
  s_h[tid] = 0; s_d[tid] = 0;
  if (tid == 0) s_d[LT_W] = 0;
  __syncthreads();
  for (u32 j = h0 + tid; j < h1; j += LT_THREADS) atomicAdd(&s_h[he[j].b & (LT_W - 1)], 1u);
  for (u32 j = r0 + tid; j < r1; j += LT_THREADS) {
    u32 lh = re[j].a;
    atomicAdd(&s_d[lh & 0xFFFF], 1);
    atomicAdd(&s_d[(lh >> 16) + 1], -1);
  }
  __syncthreads();
  int d = s_d[tid], incl;
  BS(ts).InclusiveSum(d, incl);
  const u32 hq = s_h[tid];
  const u32 p = ((t + t0) << LT_SHIFT) + tid;
  if (cov && p < n) cov[(t << LT_SHIFT) + tid] = (u32)incl + hq + (p == n - 1 ? ndocs : 0u);
  if (pcov) {
    int hex;
    BS(ts2).ExclusiveSum((int)hq, hex);
    if (p <= n) pcov[p] = make_uint2((u32)incl + hq, h0 + (u32)hex);
  }
  // pair tests of pass B for this tile: every head against every range and its position mates
  unsigned long long w = (unsigned long long)hq * (hq + (r1 - r0));
  unsigned long long tot = BR(tr).Sum(w);
  if (tid == 0) {
    if (tot) atomicAdd(work, tot);
    atomicMax(maxent, (h1 - h0) + (r1 - r0));
  }
}

// ---- pass B: within + head symbols --------------------------------------------------------------
// A head needs the number of ranges that cover its position with a smaller kappa.  The tile's
// ranges are split into LT_NB kappa buckets by LT_NB-1 pivots sampled from the ranges themselves;
// C[b][q] = ranges of buckets < b covering position q comes from per-bucket difference arrays
// (two shared-memory atomics per range, one scan per bucket row, one prefix over the buckets), so
// a head only tests the ranges of its own bucket one by one (about rc / LT_NB of them).
#define LT_NB 16
__global__ void __launch_bounds__(LT_THREADS) k_tile_place(const TileEnt *he, const TileEnt *re, const u32 *hoff,
                                                           const u32 *roff, const u32 *isa, const u64 *base,
                                                           u32 n, u32 t0, u8 *out, u64 *plc, u32 *within_out,
                                                           u64 *off_out, int3 nb_t, u64 *hw) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ u32 s_cnt[LT_W], s_off[LT_W + 1], s_cur[LT_W];
  __shared__ u64 s_base[LT_W];
  __shared__ u64 s_praw[LT_NB], s_piv[LT_NB];
  __shared__ u32 s_bcnt[LT_NB + 1], s_boff[LT_NB + 2], s_bcur[LT_NB + 1];
  __shared__ int s_C[LT_NB + 1][LT_W + 1];
  typedef cub::BlockScan<u32, LT_THREADS> BS;
  __shared__ typename BS::TempStorage ts;
  const u32 t = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const u32 h0 = hoff[t], h1 = hoff[t + 1], r0 = roff[t], r1 = roff[t + 1];
  const u32 hc = h1 - h0, rc = r1 - r0;
  if (hc == 0) return;
  TileEnt *s_he = (TileEnt *)smem;              // hc heads, grouped by position
  u64 *s_rk = (u64 *)(s_he + hc);               // rc range kappas, grouped by bucket
  u32 *s_rl = (u32 *)(s_rk + rc);               // their lo | hi << 16
  // number of kappa buckets by depth: a shallow tile is cheaper by plain pair tests than by the
  // fixed cost of LT_NB difference rows
  const u32 NB = rc <= (u32)nb_t.x ? 1u : rc <= (u32)nb_t.y ? 4u : rc <= (u32)nb_t.z ? 8u : (u32)LT_NB;
  const u32 npiv = NB - 1;                      // rc >= npiv by construction
  s_cnt[tid] = 0; s_cur[tid] = 0;
  for (u32 k = tid; k < NB * (LT_W + 1); k += LT_THREADS) (&s_C[0][0])[k] = 0;
  if (tid <= LT_NB) { s_bcnt[tid] = 0; s_bcur[tid] = 0; }
  if (tid < npiv) s_praw[tid] = re[r0 + (u32)(((u64)(tid + 1) * rc) / (npiv + 1))].kappa;
  // the rank of position p is needed only if p has heads, but that is known a sync later: fetch
  // isa[p] now (coalesced) and do the random gather base[rank] once the head counts are in
  const u32 p_own = ((t + t0) << LT_SHIFT) + tid;
  const u32 rank_own = (!hw && p_own < n) ? isa[p_own] : 0;
  // the first range / head of every thread stays in registers between the two passes
  TileEnt re0, he0;
  re0.a = re0.b = 0; re0.kappa = 0; he0 = re0;
  if (tid < rc) re0 = re[r0 + tid];
  if (tid < hc) he0 = he[h0 + tid];
  __syncthreads();
  if (tid < npiv) {                             // sort the pivots (distinct kappas) by rank counting
    const u64 mine = s_praw[tid];
    u32 rk = 0;
    for (u32 k = 0; k < npiv; k++) rk += s_praw[k] < mine;
    s_piv[rk] = mine;
  }
  __syncthreads();
  // pass 1 over the ranges: bucket sizes and per-bucket coverage deltas; head counts per position
  for (u32 j = tid; j < rc; j += LT_THREADS) {
    const TileEnt e = j == tid ? re0 : re[r0 + j];
    u32 bk = 0;
    for (u32 k = 0; k < npiv; k++) bk += s_piv[k] <= e.kappa;
    atomicAdd(&s_bcnt[bk], 1u);
    if (NB > 1) {
      atomicAdd(&s_C[bk][e.a & 0xFFFF], 1);
      atomicAdd(&s_C[bk][(e.a >> 16) + 1], -1);
    }
  }
  for (u32 j = tid; j < hc; j += LT_THREADS) atomicAdd(&s_cnt[(j == tid ? he0 : he[h0 + j]).b & (LT_W - 1)], 1u);
  __syncthreads();
  s_base[tid] = (!hw && s_cnt[tid]) ? base[rank_own] : 0;      // 1/3 of the positions have no head: no gather
  u32 ex;
  BS(ts).ExclusiveSum(s_cnt[tid], ex);
  s_off[tid] = ex;
  if (tid == LT_THREADS - 1) s_off[LT_W] = ex + s_cnt[tid];
  if (tid == 0) {
    u32 run = 0;
    for (u32 k = 0; k < NB; k++) { s_boff[k] = run; run += s_bcnt[k]; }
    s_boff[NB] = run;
  }
  if (NB > 1) {
    // coverage per bucket: inclusive scan of every delta row (one warp per row, 8 entries per lane)
    for (u32 row = warp; row < NB; row += LT_THREADS / 32) {
      int v[8], sum = 0;
#pragma unroll
      for (int k = 0; k < 8; k++) { sum += s_C[row][lane * 8 + k]; v[k] = sum; }
      int incl = sum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { int o = __shfl_up_sync(0xffffffffu, incl, d); if ((int)lane >= d) incl += o; }
      const int pre = incl - sum;
#pragma unroll
      for (int k = 0; k < 8; k++) s_C[row][lane * 8 + k] = v[k] + pre;
    }
    __syncthreads();
    {                                           // exclusive prefix over the buckets, per position
      int run = 0;
      for (u32 bk = 0; bk < NB; bk++) { int c = s_C[bk][tid]; s_C[bk][tid] = run; run += c; }
    }
  } else {
    __syncthreads();
  }
  // pass 2: ranges grouped by bucket, heads grouped by position
  for (u32 j = tid; j < rc; j += LT_THREADS) {
    const TileEnt e = j == tid ? re0 : re[r0 + j];
    u32 bk = 0;
    for (u32 k = 0; k < npiv; k++) bk += s_piv[k] <= e.kappa;
    const u32 slot = s_boff[bk] + atomicAdd(&s_bcur[bk], 1u);
    s_rk[slot] = e.kappa;
    s_rl[slot] = e.a;
  }
  for (u32 j = tid; j < hc; j += LT_THREADS) {
    TileEnt e = j == tid ? he0 : he[h0 + j];
    u32 q = e.b & (LT_W - 1);
    s_he[s_off[q] + atomicAdd(&s_cur[q], 1u)] = e;
  }
  __syncthreads();
  for (u32 j = tid; j < hc; j += LT_THREADS) {
    const TileEnt e = s_he[j];
    const u32 q = e.b & (LT_W - 1);
    const u64 kap = e.kappa;
    u32 w = 0;
    for (u32 jj = s_off[q]; jj < s_off[q + 1]; jj++) w += s_he[jj].kappa < kap;
    const u32 mates = w;                        // kappas of one position are distinct: a permutation
    u32 bk = 0;
    for (u32 k = 0; k < npiv; k++) bk += s_piv[k] <= kap;
    if (NB > 1) w += (u32)s_C[bk][q];
    for (u32 r = s_boff[bk]; r < s_boff[bk + 1]; r++) {
      const u32 lh = s_rl[r];
      w += (s_rk[r] < kap) & ((lh & 0xFFFF) <= q) & (q <= (lh >> 16));
    }
    if (hw) {       // pull emit: (within << 8 | symbol), grouped by position, ascending within
      hw[h0 + s_off[q] + mates] = ((u64)w << 8) | ((e.b >> 8) & 0xFF);
      if (within_out) within_out[e.a] = w;
      continue;
    }
    const u64 off = s_base[q] + w;
    if (plc) plc[h0 + j] = (off << 8) | ((e.b >> 8) & 0xFF);   // multi-GPU: the byte belongs to another device
    else out[off] = (u8)(e.b >> 8);
    if (within_out) { within_out[e.a] = w; off_out[e.a] = off; }
  }
}

// ---- pass B of the pull emit, single GPU: bucket sizes AND within(h) in one kernel ----------------------
// within(h) = mates with a smaller kappa + ranges covering pos_h with a smaller kappa (:1517-1603).
// kappa = (cls, tail rank) with cls = smaller ? E : ~E (match.h:23-33): every "smaller" phrase precedes
// every other one, by ascending / descending match end E.  With X = end of the tile + TR_MARGIN, a range
// or head is NEAR when its E < X.  Then, exactly:
//   near head, smaller:      only near smaller ranges can have a lower kappa
//                            within = mates + #{near smaller ranges: kappa < kappa_h, covers q}
//   near head, not smaller:  only near not-smaller ranges can have a higher kappa
//                            within = mates + cov(q) - #{near not-smaller ranges: kappa > kappa_h, covers q}
//   far head:                every near smaller range covering q precedes it, no near not-smaller one does:
//                            within = mates + cov_near_smaller(q) + #{far ranges: kappa < kappa_h, covers q},
//                            the far ranges tested by a whole warp (32 per step)
// cov(q) = ranges covering q comes from the difference array the bucket sizes need anyway.  Nearly all
// heads are the short ones in front of a mutation (near); the near lists hold the few phrases that end
// around the tile, so a head tests ~15 ranges instead of a quarter of the tile (round 1: 4 kappa buckets
// by sampled pivots, per-bucket coverage rows; 37 % of the instructions in divergent pair tests), no
// sorting, no pivots, and lanes of a warp walk the same list.
#define TR_MARGIN 48u
#define TR_SHORT 8u                   // near ranges covering at most this many positions are filed per position
#define TR_MAX_ENTRIES 8192u          // heads + ranges of one tile (dynamic shared memory: 24 bytes per entry)
// Most near ranges are SHORT: the heads in front of a mutation that are followed by a few non-heads (ncu,
// config 3: ~190 ranges per tile, ~100 of them near, while a head has ~2 of them over its position).
// They are filed under every position they cover (at most TR_SHORT 16-bit indices each, room
// permitting: 40 instead of 24 bytes per entry), so that a head walks its position's incidences like
// its mates instead of the whole near list.
static size_t tile_rank_smem(u32 maxent, bool with_short) { return (size_t)maxent * (with_short ? 40 : 24) + 64; }
__device__ __forceinline__ u32 tr_covers(u32 lh, u32 q) { return ((lh & 0xFFFF) <= q) & (q <= (lh >> 16)); }
__global__ void __launch_bounds__(LT_THREADS) k_tile_rank(const TileEnt *__restrict__ he, const TileEnt *__restrict__ re,
                                                          const u32 *__restrict__ hoff, const u32 *__restrict__ roff,
                                                          u32 n, u32 margin, u32 smem_bytes, const u32 *__restrict__ isa,
                                                          u32 nd, uint4 *__restrict__ vrec, u64 *__restrict__ hw,
                                                          u32 *within_out, u32 t0, const u32 *__restrict__ slot) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ u32 s_cnt[LT_W], s_off[LT_W + 1], s_cur[LT_W], s_cov[LT_W];
  __shared__ u32 s_icnt[LT_W], s_ioff[LT_W + 1], s_cov1[LT_W];
  __shared__ int s_d[LT_W + 1], s_d1[LT_W + 1];
  __shared__ u32 s_c1, s_c0, s_cs, s_cf, s_nfar;
  typedef cub::BlockScan<u32, LT_THREADS> BS;
  typedef cub::BlockScan<int, LT_THREADS> BSI;
  __shared__ typename BS::TempStorage ts;
  __shared__ typename BS::TempStorage ts2;
  __shared__ typename BSI::TempStorage tsi;
  __shared__ typename BSI::TempStorage tsi2;
  const u32 t = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
  const u32 h0 = hoff[t], h1 = hoff[t + 1], r0 = roff[t], r1 = roff[t + 1];
  const u32 hc = h1 - h0, rc = r1 - r0;
  const bool shorts = (size_t)hc * 20 + (size_t)rc * 40 + 64 <= smem_bytes;   // room for the incidence lists
  TileEnt *s_he = (TileEnt *)smem;               // hc heads, grouped by position
  u64 *s_nk = (u64 *)(s_he + hc);                // near long ranges: smaller ones from the front, the others from the back
  u64 *s_fk = s_nk + rc;                         // short near ranges from the front, far ranges from the back
  u32 *s_nl = (u32 *)(s_fk + rc);                // lo | hi << 16 of the same
  u32 *s_fl = s_nl + rc;
  u32 *s_far = s_fl + rc;                        // far heads (indices into s_he)
  u16 *s_inc = (u16 *)(s_far + hc);              // per position: indices (into s_fk) of the short ranges covering it
  const u64 Xl = ((u64)(t + t0 + 1) << LT_SHIFT) + margin;      // t0: first tile of this device's range (multi-GPU)
  const u32 X = Xl > 0x7FFFFFFFull ? 0x7FFFFFFFu : (u32)Xl;
  s_cnt[tid] = 0; s_cur[tid] = 0; s_d[tid] = 0; s_d1[tid] = 0; s_icnt[tid] = 0;
  if (tid == 0) { s_d[LT_W] = 0; s_d1[LT_W] = 0; s_c1 = 0; s_c0 = 0; s_cs = 0; s_cf = 0; s_nfar = 0; }
  TileEnt he0;
  he0.a = he0.b = 0; he0.kappa = 0;
  if (tid < hc) he0 = he[h0 + tid];
  // SA rank of this thread's position (coalesced): its bucket record goes straight to virtual-rank order
  const u32 p_own = ((t + t0) << LT_SHIFT) + tid;
  const u32 rank_own = p_own < n ? isa[p_own] : 0u;
  // multi-GPU: the record goes to its slot of the send buffer (grouped by the owner of its SA rank)
  const u32 slot_own = slot ? slot[(t << LT_SHIFT) + tid] : 0u;
  __syncthreads();
  for (u32 j = tid; j < rc; j += LT_THREADS) {
    const TileEnt e = re[r0 + j];
    const u32 cls = (u32)(e.kappa >> 32), lo = e.a & 0xFFFF, hi = e.a >> 16;
    const bool near1 = cls < X, near0 = cls > 0xFFFFFFFFu - X;
    if ((near1 || near0) && shorts && hi - lo < TR_SHORT) {
      const u32 k = atomicAdd(&s_cs, 1u);
      s_fk[k] = e.kappa; s_fl[k] = e.a;
      for (u32 q = lo; q <= hi; q++) atomicAdd(&s_icnt[q], 1u);
    } else if (near1) { const u32 k = atomicAdd(&s_c1, 1u); s_nk[k] = e.kappa; s_nl[k] = e.a; }
    else if (near0) { const u32 k = rc - 1 - atomicAdd(&s_c0, 1u); s_nk[k] = e.kappa; s_nl[k] = e.a; }
    else { const u32 k = rc - 1 - atomicAdd(&s_cf, 1u); s_fk[k] = e.kappa; s_fl[k] = e.a; }
    atomicAdd(&s_d[lo], 1);
    atomicAdd(&s_d[hi + 1], -1);
    if (near1) { atomicAdd(&s_d1[lo], 1); atomicAdd(&s_d1[hi + 1], -1); }
  }
  for (u32 j = tid; j < hc; j += LT_THREADS) atomicAdd(&s_cnt[(j == tid ? he0 : he[h0 + j]).b & (LT_W - 1)], 1u);
  __syncthreads();
  {
    const u32 hq = s_cnt[tid];
    u32 ex, iex;
    int incl, incl1;
    BS(ts).ExclusiveSum(hq, ex);
    BS(ts2).ExclusiveSum(s_icnt[tid], iex);
    BSI(tsi).InclusiveSum(s_d[tid], incl);
    BSI(tsi2).InclusiveSum(s_d1[tid], incl1);
    s_off[tid] = ex;
    s_ioff[tid] = iex;
    s_cov[tid] = (u32)incl;
    s_cov1[tid] = (u32)incl1;
    if (tid == LT_THREADS - 1) { s_off[LT_W] = ex + hq; s_ioff[LT_W] = iex + s_icnt[tid]; }
  }
  __syncthreads();
  s_icnt[tid] = 0;                               // reused as the cursors of the incidence lists
  __syncthreads();
  const u32 c1 = s_c1, c0 = s_c0, cs = s_cs, cf = s_cf;
  const bool inl = slot == nullptr;              // multi-GPU records keep their SA rank in .w: no inline heads
  (void)cs;
  for (u32 k = tid; k < cs; k += LT_THREADS) {
    const u32 lh = s_fl[k], lo = lh & 0xFFFF, hi = lh >> 16;
    for (u32 q = lo; q <= hi; q++) s_inc[s_ioff[q] + atomicAdd(&s_icnt[q], 1u)] = (u16)k;
  }
  for (u32 j = tid; j < hc; j += LT_THREADS) {
    const TileEnt e = j == tid ? he0 : he[h0 + j];
    const u32 q = e.b & (LT_W - 1);
    s_he[s_off[q] + atomicAdd(&s_cur[q], 1u)] = e;
  }
  __syncthreads();
  for (u32 j = tid; j < hc; j += LT_THREADS) {
    const TileEnt e = s_he[j];
    const u32 q = e.b & (LT_W - 1);
    const u64 kap = e.kappa;
    const u32 cls = (u32)(kap >> 32);
    u32 mates = 0;
    for (u32 jj = s_off[q]; jj < s_off[q + 1]; jj++) mates += s_he[jj].kappa < kap;
    u32 c = 0, w;
    if (cls < X) {
      for (u32 r = 0; r < c1; r++) c += (s_nk[r] < kap) & tr_covers(s_nl[r], q);
      for (u32 r = s_ioff[q]; r < s_ioff[q + 1]; r++) c += s_fk[s_inc[r]] < kap;
      w = mates + c;
    } else if (cls > 0xFFFFFFFFu - X) {
      for (u32 r = rc - c0; r < rc; r++) c += (s_nk[r] > kap) & tr_covers(s_nl[r], q);
      for (u32 r = s_ioff[q]; r < s_ioff[q + 1]; r++) c += s_fk[s_inc[r]] > kap;
      w = mates + s_cov[q] - c;
    } else {
      s_far[atomicAdd(&s_nfar, 1u)] = j | (mates << 16);      // hc, mates <= 8192
      continue;
    }
    // at most two heads on the position: they travel inside the bucket record (s_d / s_d1 / s_cur are free by now)
    if (inl && s_cnt[q] <= 2) { (mates ? s_d1 : s_d)[q] = (int)w; ((u8 *)s_cur)[q * 4 + mates] = (u8)(e.b >> 8); }
    else hw[h0 + s_off[q] + mates] = ((u64)w << 8) | ((e.b >> 8) & 0xFF);     // grouped by position, ascending within
    if (within_out) within_out[e.a] = w;
  }
  __syncthreads();
  const u32 nfar = s_nfar;
  for (u32 f = tid >> 5; f < nfar; f += LT_THREADS / 32) {
    const u32 fj = s_far[f], j = fj & 0xFFFF, mates = fj >> 16;
    const TileEnt e = s_he[j];
    const u32 q = e.b & (LT_W - 1);
    const u64 kap = e.kappa;
    u32 c = 0;
    for (u32 r = rc - cf + lane; r < rc; r += 32) c += (s_fk[r] < kap) & tr_covers(s_fl[r], q);
    c = __reduce_add_sync(0xffffffffu, c);
    if (lane == 0) {
      const u32 w = mates + s_cov1[q] + c;
      if (inl && s_cnt[q] <= 2) { (mates ? s_d1 : s_d)[q] = (int)w; ((u8 *)s_cur)[q * 4 + mates] = (u8)(e.b >> 8); }
      else hw[h0 + s_off[q] + mates] = ((u64)w << 8) | ((e.b >> 8) & 0xFF);
      if (within_out) within_out[e.a] = w;
    }
  }
  __syncthreads();
  // {bucket size, heads}: one aligned 16-byte record per position (common.cuh), stored at its virtual rank (the
  // bucket inversion :1618-1628 as a scatter; the other CTAs of the SM hide it); SA rank 0 is the
  // terminators' bucket, which the virtual ranks below nd replace
  if (p_own < n && rank_own) {
    const u32 hq = s_cnt[tid], size = s_cov[tid] + hq;
    if (slot) vrec[slot_own] = make_uint4(size, h0 + s_off[tid], hq, rank_own - 1);
    else if (hq == 0 || hq > 2) vrec[nd + rank_own - 1] = make_uint4(size, h0 + s_off[tid], hq, 0u);
    else {
      const u32 sy = s_cur[tid];
      vrec[nd + rank_own - 1] = make_uint4(size, (u32)s_d[tid], hq | VREC_INLINE | ((sy & 0xFFu) << 16) |
                                           (hq > 1 ? ((sy >> 8) & 0xFFu) << 24 : 0u), hq > 1 ? (u32)s_d1[tid] : 0u);
    }
  }
}

// terminators: bucket of x[n-1..] (SA rank 0) holds one suffix per document, in document order
__global__ void k_place_sentinels(const u32 *len, const u16 *cs, const u32 *sexcl, u64 H, u8 *out,
                                  u32 *within_out, u64 *off_out) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= H || len[i] != 0) return;
  u32 d = sexcl[i];
  out[d] = (u8)(cs[i] & 0xFF);
  if (within_out) { within_out[i] = d; off_out[i] = d; }
}

// Builds the tile entries and cov[]; false = too heavy for the brute-force path (nothing kept).
// histogram + scan + scatter of this device's heads over ntiles tiles -> t_he, t_re, t_hoff, t_roff
__global__ void k_tile_deepest(const u32 *hcnt, const u32 *rcnt, u32 ntiles, u32 *mx) {
  u32 t = blockIdx.x * blockDim.x + threadIdx.x;
  u32 v = t < ntiles ? hcnt[t] + rcnt[t] : 0;
  v = __reduce_max_sync(0xffffffffu, v);
  if ((threadIdx.x & 31) == 0 && v) atomicMax(mx, v);
}
// depth_limit > 0: give up (false) as soon as a tile is known to hold more entries than that --
// before the histogram when even the average does, else right after it, before the scatter
static bool tiles_scatter(cmsbwt_ctx *c, const u64 *kappa, u32 ntiles, u32 *HE_out, u64 *R_out,
                          u32 depth_limit = 0) {
  cudaStream_t st = c->st;
  const u64 H = c->H;
  const int B = 256;
  if (depth_limit && H > (u64)ntiles * depth_limit) return false;
  const u32 gH = (u32)((H + B - 1) / B);
  DBuf hcnt, rcnt, tmp;
  hcnt.alloc(((size_t)ntiles + 1) * 4, st); rcnt.alloc(((size_t)ntiles + 1) * 4, st);
  c->t_hoff.alloc(((size_t)ntiles + 1) * 4, st); c->t_roff.alloc(((size_t)ntiles + 1) * 4, st);
  CK(cudaMemsetAsync(hcnt.p, 0, ((size_t)ntiles + 1) * 4, st));
  CK(cudaMemsetAsync(rcnt.p, 0, ((size_t)ntiles + 1) * 4, st));
  const bool few = ntiles <= TA_MAX_TILES;
  const u32 gS = (u32)((H + TA_SLICE - 1) / TA_SLICE);
  if (few)
    k_tile_hist_few<<<gS, B, (size_t)ntiles * 8, st>>>(c->h_pos.as<u32>(), c->h_len.as<u32>(), c->h_toNext.as<u32>(), H,
                                                       ntiles, hcnt.as<u32>(), rcnt.as<u32>());
  else
    k_tile_hist<<<gH, B, 0, st>>>(c->h_pos.as<u32>(), c->h_len.as<u32>(), c->h_toNext.as<u32>(), H,
                                  hcnt.as<u32>(), rcnt.as<u32>());
  size_t tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, hcnt.as<u32>(), c->t_hoff.as<u32>(), (int)(ntiles + 1), st);
  tmp.alloc(tb, st);
  size_t t1 = tb;
  CK(cub::DeviceScan::ExclusiveSum(tmp.p, t1, hcnt.as<u32>(), c->t_hoff.as<u32>(), (int)(ntiles + 1), st));
  t1 = tb;
  // the range total can exceed 2^32 only for collections far deeper than one device holds
  DBuf rtot; rtot.alloc(8, st);
  {
    size_t tr = 0;
    cub::DeviceReduce::Sum(nullptr, tr, rcnt.as<u32>(), rtot.as<unsigned long long>(), (int)ntiles, st);
    DBuf tmp2; tmp2.alloc(tr, st);
    CK(cub::DeviceReduce::Sum(tmp2.p, tr, rcnt.as<u32>(), rtot.as<unsigned long long>(), (int)ntiles, st));
  }
  unsigned long long R = 0;
  u32 deepest = 0;
  DBuf mx; mx.alloc(4, st);
  CK(cudaMemsetAsync(mx.p, 0, 4, st));
  k_tile_deepest<<<(ntiles + B - 1) / B, B, 0, st>>>(hcnt.as<u32>(), rcnt.as<u32>(), ntiles, mx.as<u32>());
  CK(cudaMemcpyAsync(&R, rtot.p, 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(&deepest, mx.p, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  c->launches += 5;
  if (R >= 0xFFFFFFF0ull) return false;
  c->tile_maxent = deepest;
  if (depth_limit && deepest > depth_limit) return false;
  CK(cub::DeviceScan::ExclusiveSum(tmp.p, t1, rcnt.as<u32>(), c->t_roff.as<u32>(), (int)(ntiles + 1), st));
  u32 HE = 0;
  CK(cudaMemcpyAsync(&HE, c->t_hoff.as<u32>() + ntiles, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  c->t_he.alloc(((size_t)HE + 1) * sizeof(TileEnt), st);
  c->t_re.alloc(((size_t)R + 1) * sizeof(TileEnt), st);
  // the counters become cursors: relative to the tile for the few-tiles variant, ABSOLUTE slots (starting at
  // the tile's offset) for k_tile_scatter -- one L2 request less per entry than offset + relative cursor
  if (few) {
    CK(cudaMemsetAsync(hcnt.p, 0, ((size_t)ntiles + 1) * 4, st));
    CK(cudaMemsetAsync(rcnt.p, 0, ((size_t)ntiles + 1) * 4, st));
  } else {
    CK(cudaMemcpyAsync(hcnt.p, c->t_hoff.p, ((size_t)ntiles + 1) * 4, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(rcnt.p, c->t_roff.p, ((size_t)ntiles + 1) * 4, cudaMemcpyDeviceToDevice, st));
  }
  if (few) {
    if (!kappa) { sort_kappa(c); kappa = c->s_kappa.as<u64>(); }
    if ((size_t)ntiles * 16 > 40 * 1024)
      CK(cudaFuncSetAttribute(k_tile_scatter_few, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(ntiles * 16)));
    k_tile_scatter_few<<<gS, B, (size_t)ntiles * 16, st>>>(c->h_pos.as<u32>(), c->h_len.as<u32>(), c->h_cs.as<u16>(),
                                                           c->h_toNext.as<u32>(), kappa, H, ntiles, c->t_hoff.as<u32>(),
                                                           c->t_roff.as<u32>(), hcnt.as<u32>(), rcnt.as<u32>(),
                                                           c->t_he.as<TileEnt>(), c->t_re.as<TileEnt>());
  } else {
    static int variant = -1;                      // measurement switch (A/B under profiles/)
    if (variant < 0) { const char *e = getenv("CMSBWT_SCATTER_VARIANT"); variant = e ? atoi(e) : 0; }
#define TS_LAUNCH(ILP, MINB)                                                                                            \
    k_tile_scatter<ILP, MINB><<<(u32)((H + B * ILP - 1) / (B * ILP)), B, 0, st>>>(                                      \
        c->h_pos.as<u32>(), c->h_len.as<u32>(), c->h_cs.as<u16>(), c->h_toNext.as<u32>(), kappa, H, c->t_hoff.as<u32>(), \
        c->t_roff.as<u32>(), hcnt.as<u32>(), rcnt.as<u32>(), c->t_he.as<TileEnt>(), c->t_re.as<TileEnt>(),              \
        c->s_tincl.as<u32>(), c->s_trank.as<u32>(), c->s_sexcl.as<u32>(), c->d_termsym.p ? c->d_termsym.as<u8>() : nullptr)
    // config 3, tiles group: 4 heads x 6 CTAs 3.74 ms | 4 x (5 by registers) 3.85 | 8 x 3 4.27 | 2 x 8 3.80 | 1 x 8 3.78
    if (variant == 1) TS_LAUNCH(4, 1);
    else if (variant == 2) TS_LAUNCH(8, 3);
    else if (variant == 3) TS_LAUNCH(2, 8);
    else if (variant == 4) TS_LAUNCH(1, 8);
    else TS_LAUNCH(4, 6);
#undef TS_LAUNCH
  }
  c->launches += 2;
  CK(cudaGetLastError());
  *HE_out = HE;
  *R_out = R;
  return true;
}

bool tiles_build(cmsbwt_ctx *c, const u64 *kappa, DBuf &sexcl) {
  const u32 n = c->ix.n;
  const u32 ntiles = (n >> LT_SHIFT) + 1;
  c->light = false;
  u32 HE = 0;
  u64 R = 0;
  // the scatter also files the terminators' symbols (document order) unless it is the few-tiles variant
  c->termsym_done = false;
  if (ntiles > TA_MAX_TILES) c->d_termsym.alloc((size_t)c->stats.n_docs + 16, c->st);
  // a tile deeper than shared memory holds sends the whole job to the sort-based sweep (heads.cu)
  if (!tiles_scatter(c, kappa, ntiles, &HE, &R, TR_MAX_ENTRIES)) {
    c->t_he.release(); c->t_re.release(); c->t_hoff.release(); c->t_roff.release(); c->d_termsym.release();
    return false;
  }
  c->termsym_done = ntiles > TA_MAX_TILES;
  c->stats.n_items = (u64)HE + R;
  c->stats.pair_tests = 0;
  c->n_hw = HE;
  c->h_sexcl.release();
  c->h_sexcl.swap(sexcl);
  c->light = true;
  return true;
}

// terminator symbols in document order (the bucket of SA rank 0, :1640-1643 "first D-1 head chars")
__global__ void k_term_syms(const u32 *len, const u16 *cs, const u32 *sexcl, u64 H, u8 *termsym, u32 *within_out) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= H || len[i] != 0) return;
  u32 d = sexcl[i];
  termsym[d] = (u8)(cs[i] & 0xFF);
  if (within_out) within_out[i] = d;
}

// pass B for the pull emit (emit.cu): within(h) of every head -> hw[] = (within << 8 | symbol), grouped
// by reference position in ascending within, and the terminators' symbols in document order.  Nothing
// is gathered from base[] and nothing is scattered into the output here: the fill kernel, which owns
// an output block in shared memory, pulls the heads of its buckets.
void tiles_count_heads(cmsbwt_ctx *c) {
  cudaStream_t st = c->st;
  const u64 H = c->H;
  const u32 n = c->ix.n;
  const u32 ntiles = (n >> LT_SHIFT) + 1;
  u32 *wo = nullptr;
  if (c->dump) { c->h_within.alloc(H * 4, st); wo = c->h_within.as<u32>(); }
  c->d_hw.alloc(((size_t)c->n_hw + 1) * 8, st);
  c->d_pcov.alloc(((size_t)c->stats.n_docs + n + 1) * 16, st);     // records in virtual-rank order
  if (!c->termsym_done) c->d_termsym.alloc((size_t)c->stats.n_docs + 16, st);
  // room for the per-position lists of short ranges where the deepest tile leaves it (else 24 bytes per entry)
  size_t sm = tile_rank_smem(c->tile_maxent, c->tile_short != 0);
  if (sm > 200 * 1024) sm = tile_rank_smem(c->tile_maxent, false);
  if (sm > 32 * 1024) CK(cudaFuncSetAttribute(k_tile_rank, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  k_tile_rank<<<ntiles, LT_THREADS, sm, st>>>(c->t_he.as<TileEnt>(), c->t_re.as<TileEnt>(), c->t_hoff.as<u32>(),
                                              c->t_roff.as<u32>(), n, (u32)c->tile_margin,
                                              c->tile_short ? (u32)sm : 0u, c->ix.isa, (u32)c->stats.n_docs, c->d_pcov.as<uint4>(),
                                              c->d_hw.as<u64>(), wo, 0u, nullptr);
  if (!c->termsym_done || wo) {
    k_term_syms<<<(u32)((H + 255) / 256), 256, 0, st>>>(c->h_len.as<u32>(), c->h_cs.as<u16>(), c->h_sexcl.as<u32>(), H,
                                                       c->d_termsym.as<u8>(), wo);
    LAUNCH(c);
  }
  LAUNCH(c);
  CK(cudaGetLastError());
}

// ================================================================================================
// Multi-GPU (SURVEY.md 8e "non-head counting in position-partitioned space"): the reference
// positions are cut into `world` ranges of tp tiles; every device files ITS heads under the global
// tile space, the entry arrays (already grouped by tile, so every destination's slab is one
// contiguous range) are exchanged all-to-all, and the owner merges the slabs tile by tile.
void shard_tiles_local(cmsbwt_ctx *c, u32 world) {
  const u32 n = c->ix.n;
  const u32 ntiles = (n >> LT_SHIFT) + 1;
  const u32 tp = (ntiles + world - 1) / world;
  c->sh_world = world;
  c->sh_tp = tp;
  u32 HE = 0;
  u64 R = 0;
  c->d_termsym.release();                        // the terminators of a shard travel separately (shard_term_syms)
  if (!tiles_scatter(c, c->s_kappa.p ? c->s_kappa.as<u64>() : nullptr, tp * world, &HE, &R))
    throw CmsErr{CMSBWT_E_LIMIT, "more than 2^32 tile ranges on one device"};
  c->sh_he = HE;
  c->sh_re = R;
  c->s_kappa.release(); c->s_nitems.release(); c->s_tincl.release(); c->s_trank.release();
  CK(cudaStreamSynchronize(c->st));
}

// per-tile entry counts summed over the W source slabs (offs: W rows of tp + 1 raw offsets)
__global__ void k_sh_tile_counts(const u32 *hoffs, const u32 *roffs, u32 W, u32 tp, u32 *hc, u32 *rc) {
  u32 t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t > tp) return;
  u32 a = 0, b = 0;
  if (t < tp)
    for (u32 s = 0; s < W; s++) {
      a += hoffs[(size_t)s * (tp + 1) + t + 1] - hoffs[(size_t)s * (tp + 1) + t];
      b += roffs[(size_t)s * (tp + 1) + t + 1] - roffs[(size_t)s * (tp + 1) + t];
    }
  hc[t] = a; rc[t] = b;
}
// one warp per tile copies the tile's segment of every slab behind one another
__global__ void k_sh_tile_merge(const TileEnt *src, const u32 *offs, const u64 *slab, u32 W, u32 tp, const u32 *moff,
                                TileEnt *dst) {
  const u32 t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (t >= tp) return;
  u32 d = moff[t];
  for (u32 s = 0; s < W; s++) {
    const u32 *o = offs + (size_t)s * (tp + 1);
    const u32 a = o[t] - o[0], cnt = o[t + 1] - o[t];
    const TileEnt *p = src + slab[s] + a;
    for (u32 k = lane; k < cnt; k += 32) dst[d + k] = p[k];
    d += cnt;
  }
}

// the same for deep tiles: block (t, s, z) copies every gridDim.z-th 256-entry piece of source s's segment
__global__ void __launch_bounds__(256) k_sh_tile_merge_deep(const TileEnt *src, const u32 *offs, const u64 *slab,
                                                            u32 W, u32 tp, const u32 *moff, TileEnt *dst) {
  const u32 t = blockIdx.x, s = blockIdx.y;
  u32 d = moff[t];
  for (u32 s2 = 0; s2 < s; s2++) { const u32 *o2 = offs + (size_t)s2 * (tp + 1); d += o2[t + 1] - o2[t]; }
  const u32 *o = offs + (size_t)s * (tp + 1);
  const u32 a = o[t] - o[0], cnt = o[t + 1] - o[t];
  const TileEnt *p = src + slab[s] + a;
  for (u32 k = blockIdx.z * 256 + threadIdx.x; k < cnt; k += gridDim.z * 256) dst[d + k] = p[k];
}

void shard_tiles_cov(cmsbwt_ctx *c, u64 ndocs_total);
// he/re: the received slabs (slab s starts at he_slab[s] / re_slab[s] entries); hoffs/roffs: the
// senders' offset slices for this device's tile range.  Leaves the merged entries in t_he/t_re
// (tile-local offsets t_hoff/t_roff) and the bucket sizes of the range in sh_cov (tp * 256 entries).
u32 shard_tiles_merge(cmsbwt_ctx *c, u32 world, u32 rank, const void *he, const void *re, const u32 *hoffs,
                      const u32 *roffs, const u64 *he_slab, const u64 *re_slab, u64 ndocs_total, bool with_cov) {
  cudaStream_t st = c->st;
  const u32 n = c->ix.n;
  const u32 tp = c->sh_tp;
  const int B = 256;
  c->sh_rank = rank;
  DBuf hc, rc, slab, tmp, acc, mh, mr;
  hc.alloc(((size_t)tp + 1) * 4, st); rc.alloc(((size_t)tp + 1) * 4, st);
  mh.alloc(((size_t)tp + 1) * 4, st); mr.alloc(((size_t)tp + 1) * 4, st);
  slab.alloc((size_t)world * 16, st); acc.alloc(16, st);
  CK(cudaMemsetAsync(acc.p, 0, 16, st));
  CK(cudaMemcpyAsync(slab.p, he_slab, (size_t)world * 8, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(slab.as<u64>() + world, re_slab, (size_t)world * 8, cudaMemcpyHostToDevice, st));
  k_sh_tile_counts<<<(tp + 1 + B - 1) / B, B, 0, st>>>(hoffs, roffs, world, tp, hc.as<u32>(), rc.as<u32>());
  size_t tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, hc.as<u32>(), mh.as<u32>(), (int)(tp + 1), st);
  tmp.alloc(tb, st);
  size_t t1 = tb;
  CK(cub::DeviceScan::ExclusiveSum(tmp.p, t1, hc.as<u32>(), mh.as<u32>(), (int)(tp + 1), st));
  t1 = tb;
  CK(cub::DeviceScan::ExclusiveSum(tmp.p, t1, rc.as<u32>(), mr.as<u32>(), (int)(tp + 1), st));
  u32 HE = 0, R = 0, deepest = 0;
  k_tile_deepest<<<(tp + B - 1) / B, B, 0, st>>>(hc.as<u32>(), rc.as<u32>(), tp, acc.as<u32>() + 3);
  CK(cudaMemcpyAsync(&HE, mh.as<u32>() + tp, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(&R, mr.as<u32>() + tp, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(&deepest, acc.as<u32>() + 3, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  DBuf mhe, mre;
  mhe.alloc(((size_t)HE + 1) * sizeof(TileEnt), st); mre.alloc(((size_t)R + 1) * sizeof(TileEnt), st);
  const u32 gW = (u32)(((u64)tp * 32 + B - 1) / B);
  if (((u64)HE + R) / ((u64)tp * world) >= 2048 && tp <= 65535) {     // deep tiles: blocks per (tile, source)
    const u64 per = ((u64)HE + R) / ((u64)tp * world);
    const dim3 g(tp, world, (u32)(per / 4096 < 1 ? 1 : per / 4096 > 32 ? 32 : per / 4096));
    k_sh_tile_merge_deep<<<g, 256, 0, st>>>((const TileEnt *)he, hoffs, slab.as<u64>(), world, tp, mh.as<u32>(),
                                            mhe.as<TileEnt>());
    k_sh_tile_merge_deep<<<g, 256, 0, st>>>((const TileEnt *)re, roffs, slab.as<u64>() + world, world, tp,
                                            mr.as<u32>(), mre.as<TileEnt>());
  } else {
    k_sh_tile_merge<<<gW, B, 0, st>>>((const TileEnt *)he, hoffs, slab.as<u64>(), world, tp, mh.as<u32>(),
                                      mhe.as<TileEnt>());
    k_sh_tile_merge<<<gW, B, 0, st>>>((const TileEnt *)re, roffs, slab.as<u64>() + world, world, tp, mr.as<u32>(),
                                      mre.as<TileEnt>());
  }
  c->t_he.swap(mhe);
  c->t_re.swap(mre);
  c->t_hoff.swap(mh);
  c->t_roff.swap(mr);
  c->sh_he = HE;
  c->sh_re = R;
  c->launches += 7;
  CK(cudaGetLastError());
  c->stats.n_items = (u64)HE + R;
  c->stats.pair_tests = 0;
  c->tile_maxent = deepest;
  if (with_cov) shard_tiles_cov(c, ndocs_total);
  return c->tile_maxent;
}

// bucket sizes of this device's tile range -> sh_cov (tp * 256 entries): the byte-range emit (shard_emit)
void shard_tiles_cov(cmsbwt_ctx *c, u64 ndocs_total) {
  cudaStream_t st = c->st;
  const u32 tp = c->sh_tp;
  DBuf acc;
  acc.alloc(16, st);
  CK(cudaMemsetAsync(acc.p, 0, 16, st));
  c->sh_cov.alloc((size_t)tp * LT_W * 4, st);
  CK(cudaMemsetAsync(c->sh_cov.p, 0, (size_t)tp * LT_W * 4, st));
  k_tile_cov<<<tp, LT_THREADS, 0, st>>>(c->t_he.as<TileEnt>(), c->t_re.as<TileEnt>(), c->t_hoff.as<u32>(),
                                        c->t_roff.as<u32>(), c->ix.n, (u32)ndocs_total, c->sh_rank * tp,
                                        c->sh_cov.as<u32>(), nullptr, acc.as<unsigned long long>(), acc.as<u32>() + 2);
  LAUNCH(c);
  CK(cudaGetLastError());
  unsigned long long hacc[2] = {0, 0};
  CK(cudaMemcpyAsync(hacc, acc.p, 16, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  c->stats.pair_tests = hacc[0];
  c->tile_maxent = (u32)hacc[1];
}

__global__ void k_sh_sentinel_plc(const u32 *len, const u16 *cs, const u32 *sexcl, u64 H, u64 doc_base, u64 *plc) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= H || len[i] != 0) return;
  u32 d = sexcl[i];
  plc[d] = ((doc_base + d) << 8) | (u64)(cs[i] & 0xFF);
}

// per-destination counts / scatter of placements (destination = offset / chunk), block-aggregated
#define PLC_MAXW 64
__global__ void __launch_bounds__(256) k_sh_plc_hist(const u64 *plc, u64 P, u64 chunk, u32 W,
                                                     unsigned long long *cnt) {
  __shared__ u32 s_c[PLC_MAXW];
  if (threadIdx.x < PLC_MAXW) s_c[threadIdx.x] = 0;
  __syncthreads();
  for (u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x; j < P; j += (u64)gridDim.x * blockDim.x)
    atomicAdd(&s_c[(u32)((plc[j] >> 8) / chunk)], 1u);
  __syncthreads();
  if (threadIdx.x < W && s_c[threadIdx.x]) atomicAdd(cnt + threadIdx.x, (unsigned long long)s_c[threadIdx.x]);
}
__global__ void __launch_bounds__(256) k_sh_plc_scatter(const u64 *plc, u64 P, u64 chunk, u32 W,
                                                        unsigned long long *cur, u64 *out) {
  __shared__ u32 s_c[PLC_MAXW];
  __shared__ unsigned long long s_b[PLC_MAXW];
  // every block handles one contiguous slice: counts, one reservation per destination, scatter
  const u64 per = (P + gridDim.x - 1) / gridDim.x;
  const u64 a = (u64)blockIdx.x * per, b = min(P, a + per);
  if (threadIdx.x < PLC_MAXW) s_c[threadIdx.x] = 0;
  __syncthreads();
  for (u64 j = a + threadIdx.x; j < b; j += blockDim.x) atomicAdd(&s_c[(u32)((plc[j] >> 8) / chunk)], 1u);
  __syncthreads();
  if (threadIdx.x < W) {
    s_b[threadIdx.x] = s_c[threadIdx.x] ? atomicAdd(cur + threadIdx.x, (unsigned long long)s_c[threadIdx.x]) : 0ull;
    s_c[threadIdx.x] = 0;
  }
  __syncthreads();
  for (u64 j = a + threadIdx.x; j < b; j += blockDim.x) {
    const u64 v = plc[j];
    const u32 d = (u32)((v >> 8) / chunk);
    out[s_b[d] + atomicAdd(&s_c[d], 1u)] = v;
  }
}

// ---- deep tile ranges (more entries per tile than shared memory holds, e.g. a short reference
// under very many documents): the owner counts with the sort-based sweep of heads.cu instead of
// k_tile_place.  Items live in 64-position sub-tiles (CMS_TILE_W) numbered inside the device's range.
#define SUBT (LT_W / CMS_TILE_W)
__global__ void __launch_bounds__(256) k_sh_deep_items(const TileEnt *he, const TileEnt *re, const u32 *hoff,
                                                       const u32 *roff, u32 HE, u32 pad_tile, int cshift,
                                                       u32 *it_tr, u64 *it_key, u64 *it_pay, u32 *epos) {
  const u64 cmask = (1ull << cshift) - 1;
  const u32 t = blockIdx.x;
  const u32 h0 = hoff[t], h1 = hoff[t + 1], r0 = roff[t], r1 = roff[t + 1];
  const u32 z0 = blockIdx.y * blockDim.x + threadIdx.x, zs = gridDim.y * blockDim.x;
  for (u32 j = h0 + z0; j < h1; j += zs) {
    const TileEnt e = he[j];
    const u32 q = e.b & (LT_W - 1), q6 = q & (CMS_TILE_W - 1);
    it_tr[j] = (u32)e.kappa;
    it_key[j] = ((u64)(t * SUBT + q / CMS_TILE_W) << cshift) | ((e.kappa >> 32) & cmask);
    it_pay[j] = ((u64)j << 32) | (q6 << 8) | q6;
    epos[j] = t * LT_W + q;
  }
  for (u32 j = r0 + z0; j < r1; j += zs) {
    const TileEnt e = re[j];
    const u32 lo = e.a & 0xFFFF, hi = e.a >> 16;
    for (u32 s = 0; s < SUBT; s++) {
      const u32 a = s * CMS_TILE_W, b = a + CMS_TILE_W - 1;
      const u64 o = (u64)HE + (u64)j * SUBT + s;
      const bool in = lo <= b && hi >= a;
      it_tr[o] = (u32)e.kappa;
      it_key[o] = ((u64)(in ? t * SUBT + s : pad_tile) << cshift) | ((e.kappa >> 32) & cmask);
      it_pay[o] = (0xFFFFFFFFull << 32) | ((in ? max(lo, a) - a : 1u) << 8) | (in ? min(hi, b) - a : 0u);
    }
  }
}
__global__ void __launch_bounds__(256) k_sh_deep_plc(const TileEnt *he, const u32 *hoff, const u32 *within,
                                                     const u32 *isa, const u64 *base, u32 n, u32 t0, u64 *plc) {
  const u32 t = blockIdx.x;
  for (u32 j = hoff[t] + blockIdx.y * blockDim.x + threadIdx.x; j < hoff[t + 1]; j += gridDim.y * blockDim.x) {
    const TileEnt e = he[j];
    const u32 p = ((t + t0) << LT_SHIFT) + (e.b & (LT_W - 1));
    const u64 off = (p < n ? base[isa[p]] : 0) + within[j];
    plc[j] = (off << 8) | ((e.b >> 8) & 0xFF);
  }
}
static void shard_place_deep(cmsbwt_ctx *c, const u64 *base, u64 *plc) {
  cudaStream_t st = c->st;
  const u32 tp = c->sh_tp, HE = c->sh_he;
  const u64 I = (u64)HE + (u64)c->sh_re * SUBT;
  const int cshift = sweep_cls_bits(c->ix.n);
  DBuf it_tr, it_key, it_pay, epos, within;
  it_tr.alloc((I + 1) * 4, st); it_key.alloc((I + 1) * 8, st); it_pay.alloc((I + 1) * 8, st);
  epos.alloc(((size_t)HE + 1) * 4, st); within.alloc(((size_t)HE + 1) * 4, st);
  // blocks per tile: the tiles of a deep range are few and hold 10^4..10^6 entries each
  u32 zy = (u32)((((u64)HE + c->sh_re) / tp) / 8192);
  zy = zy < 1 ? 1 : zy > 64 ? 64 : zy;
  k_sh_deep_items<<<dim3(tp, zy), 256, 0, st>>>(c->t_he.as<TileEnt>(), c->t_re.as<TileEnt>(), c->t_hoff.as<u32>(),
                                      c->t_roff.as<u32>(), HE, tp * SUBT, cshift, it_tr.as<u32>(), it_key.as<u64>(),
                                      it_pay.as<u64>(), epos.as<u32>());
  LAUNCH(c);
  sweep_count(c, I, it_tr, it_key, it_pay, epos.as<u32>(), tp * SUBT, 32, cshift, within.as<u32>());
  k_sh_deep_plc<<<dim3(tp, zy), 256, 0, st>>>(c->t_he.as<TileEnt>(), c->t_hoff.as<u32>(), within.as<u32>(), c->ix.isa, base,
                                    c->ix.n, c->sh_rank * tp, plc);
  LAUNCH(c);
  CK(cudaGetLastError());
}

// pass B for this device's tile range: (offset << 8 | symbol) per head entry, then this shard's
// document terminators (offset = global document index); grouped by destination in sh_plc.
void shard_place(cmsbwt_ctx *c, const u64 *base, u64 doc_base) {
  cudaStream_t st = c->st;
  const u32 n = c->ix.n;
  const u32 tp = c->sh_tp;
  const u64 H = c->H;
  const u64 nd = c->stats.n_docs;
  const u64 P = c->sh_he + nd;
  DBuf raw, tmp;
  raw.alloc((P + 1) * 8, st);
  c->sh_plc.alloc((P + 1) * 8, st);
  c->stats.light_tiles = c->tile_maxent <= LT_MAX_ENTRIES && c->tile_budget > 0;
  if (!c->stats.light_tiles) {
    shard_place_deep(c, base, raw.as<u64>());
  } else {
    size_t sm = (size_t)c->tile_maxent * sizeof(TileEnt) + 64;
    if (sm > 16 * 1024)
      CK(cudaFuncSetAttribute(k_tile_place, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    k_tile_place<<<tp, LT_THREADS, sm, st>>>(c->t_he.as<TileEnt>(), c->t_re.as<TileEnt>(), c->t_hoff.as<u32>(),
                                             c->t_roff.as<u32>(), c->ix.isa, base, n, c->sh_rank * tp, nullptr,
                                             raw.as<u64>(), nullptr, nullptr,
                                             make_int3((int)c->nb_t[0], (int)c->nb_t[1], (int)c->nb_t[2]), nullptr);
  }
  k_sh_sentinel_plc<<<(u32)((H + 255) / 256), 256, 0, st>>>(c->h_len.as<u32>(), c->h_cs.as<u16>(),
                                                           c->s_sexcl.as<u32>(), H, doc_base, raw.as<u64>() + c->sh_he);
  // group the placements by destination (= owner of the output byte range): histogram, offsets,
  // scatter -- two passes instead of a sort, the order inside a destination does not matter
  const u32 W = c->sh_world;
  DBuf cnt, cur;
  cnt.alloc((size_t)W * 8, st); cur.alloc((size_t)W * 8, st);
  CK(cudaMemsetAsync(cnt.p, 0, (size_t)W * 8, st));
  const u32 gP = (u32)min((u64)148 * 16, (P + 255) / 256);
  if (P) k_sh_plc_hist<<<gP, 256, 0, st>>>(raw.as<u64>(), P, c->sh_chunk, W, cnt.as<unsigned long long>());
  std::vector<unsigned long long> hc(W), ho(W);
  CK(cudaMemcpyAsync(hc.data(), cnt.p, (size_t)W * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  c->sh_plc_cnt.assign(W, 0);
  unsigned long long run = 0;
  for (u32 d = 0; d < W; d++) { ho[d] = run; run += hc[d]; c->sh_plc_cnt[d] = hc[d]; }
  CK(cudaMemcpyAsync(cur.p, ho.data(), (size_t)W * 8, cudaMemcpyHostToDevice, st));
  if (P) k_sh_plc_scatter<<<gP, 256, 0, st>>>(raw.as<u64>(), P, c->sh_chunk, W, cur.as<unsigned long long>(),
                                               c->sh_plc.as<u64>());
  c->launches += 4;
  c->sh_nplc = P;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));
}

// ================================================================================================
// Multi-GPU pull emit, sender side.  The output is owned by SA-RANK range (device D emits the buckets of
// the SA ranks [1 + D * rper, 1 + (D + 1) * rper), device 0 the terminators in front of them), so nothing
// n-sized is replicated: the owner of a reference-position range runs k_tile_rank over its merged tiles
// and sends every position's 16-byte bucket record and its head list to the owner of the position's SA
// rank.  Which record goes where depends on the index alone (ISA), so the slot of every position in
// the send buffer -- grouped by destination -- is a table built once per (index, world) and the
// records are stored into their slots by k_tile_rank itself (the bucket inversion :1618-1628 as a
// scatter across devices).
__global__ void k_sh_vdest(const u32 *__restrict__ isa, u32 n, u32 p0, u32 np, u32 *key, u32 *pl) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= np) return;
  const u32 p = p0 + i;
  const u32 r = p < n ? isa[p] : 0u;           // SA rank 0 (the sentinel suffix) is the terminators' bucket
  key[i] = r ? r - 1 : 0xFFFFFFFFu;            // sorted by SA rank: grouped by destination, and every destination
  pl[i] = i;                                   // receives its records in the order it stores them
}
// sorted by destination: slot of every local position, edges[d] = first slot of destination d (edges[W] = records)
__global__ void k_sh_vslot(const u32 *__restrict__ ks, const u32 *__restrict__ pls, u32 np, u32 rper, u32 W, u32 *slot,
                           u32 *edges) {
  const u32 s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= np) return;
  const u32 k = ks[s] == 0xFFFFFFFFu ? W : ks[s] / rper;
  slot[pls[s]] = k < W ? s : 0xFFFFFFFFu;
  if (s == 0) { for (u32 d = 0; d <= k && d <= W; d++) edges[d] = 0; }
  else {
    const u32 kp = ks[s - 1] == 0xFFFFFFFFu ? W : ks[s - 1] / rper;
    for (u32 d = kp + 1; d <= k && d <= W; d++) edges[d] = s;
  }
  if (s == np - 1) for (u32 d = k + 1; d <= W; d++) edges[d] = np;
}
__global__ void k_sh_rec_fields(const uint4 *__restrict__ rec, u32 ns, u32 *hz, u32 *sz) {
  const u32 s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s > ns) return;
  const uint4 a = s < ns ? rec[s] : make_uint4(0, 0, 0, 0);
  hz[s] = a.z; sz[s] = a.x;
}
// head lists into the send buffer in slot order; a record's "first head" becomes relative to its destination's slab
__global__ void __launch_bounds__(256) k_sh_hw_route(uint4 *rec, const u32 *__restrict__ hoff, const u32 *__restrict__ edges,
                                                     u32 W, u32 ns, const u64 *__restrict__ hw, u64 *hws) {
  __shared__ u32 s_e[PLC_MAXW + 1];
  if (threadIdx.x <= W) s_e[threadIdx.x] = edges[threadIdx.x];
  __syncthreads();
  const u32 s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= ns) return;
  uint4 a = rec[s];
  u32 d = 0;
  while (d + 1 < W && s >= s_e[d + 1]) d++;
  const u32 o = hoff[s];
  for (u32 j = 0; j < a.z; j++) hws[o + j] = hw[a.y + j];
  a.y = o - hoff[s_e[d]];
  rec[s] = a;
}
__global__ void k_sh_edge_counts(const u32 *hoff, const u64 *boff, const u32 *edges, u32 W, u64 *out) {
  const u32 d = threadIdx.x;
  if (d >= W) return;
  out[d] = hoff[edges[d + 1]] - hoff[edges[d]];
  out[W + d] = boff[edges[d + 1]] - boff[edges[d]];
}

// false: a tile of this device's range is deeper than k_tile_rank holds (the caller falls back to the
// byte-range emit with the sort-based counting).  Leaves sh_vsend (records by destination, edges in
// sh_slot_edges), sh_hwsend (head lists by destination, counts in sh_v_hwcnt) and the output bytes this
// device's records stand for, per destination (sh_v_bytes).
bool shard_rank_pull(cmsbwt_ctx *c, u32 W, u32 R) {
  cudaStream_t st = c->st;
  const u32 n = c->ix.n;
  const u32 tp = c->sh_tp;
  const int B = 256;
  if (W > PLC_MAXW) throw CmsErr{CMSBWT_E_LIMIT, "more than 64 devices"};
  if (c->tile_maxent > TR_MAX_ENTRIES || c->tile_budget <= 0 || n < 2) return false;
  const u32 rper = (n - 1 + W - 1) / W;          // SA ranks 1 .. n-1
  const u32 np = tp * LT_W;
  const u64 key[5] = {W, R, n, (u64)(uintptr_t)c->ix.isa, tp};
  if (memcmp(key, c->sh_slot_key, sizeof key) != 0 || !c->sh_slot.p) {
    DBuf k1, k2, v1, v2, ed;
    k1.alloc((size_t)np * 4, st); k2.alloc((size_t)np * 4, st); v1.alloc((size_t)np * 4, st); v2.alloc((size_t)np * 4, st);
    ed.alloc((size_t)(W + 1) * 4, st);
    c->sh_slot.alloc((size_t)np * 4, st);
    k_sh_vdest<<<(np + B - 1) / B, B, 0, st>>>(c->ix.isa, n, R * tp * LT_W, np, k1.as<u32>(), v1.as<u32>());
    rsort_pairs32(c, k1, k2, v1, v2, np, 0, 32);
    k_sh_vslot<<<(np + B - 1) / B, B, 0, st>>>(k1.as<u32>(), v1.as<u32>(), np, rper, W, c->sh_slot.as<u32>(), ed.as<u32>());
    std::vector<u32> he(W + 1);
    CK(cudaMemcpyAsync(he.data(), ed.p, (size_t)(W + 1) * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    c->sh_slot_edges.assign(he.begin(), he.end());
    memcpy(c->sh_slot_key, key, sizeof key);
    c->launches += 2;
  }
  c->sh_rper = rper;
  const u32 ns = (u32)c->sh_slot_edges[W];
  DBuf ed, hz, sz, hoff, boff, tmp, cnts;
  ed.alloc((size_t)(W + 1) * 4, st);
  {
    std::vector<u32> he(c->sh_slot_edges.begin(), c->sh_slot_edges.end());
    CK(cudaMemcpyAsync(ed.p, he.data(), (size_t)(W + 1) * 4, cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));             // he is a temporary
  }
  c->sh_vsend.alloc(((size_t)ns + 1) * 16, st);
  c->d_hw.alloc(((size_t)c->sh_he + 1) * 8, st);
  c->sh_hwsend.alloc(((size_t)c->sh_he + 1) * 8, st);
  size_t sm = tile_rank_smem(c->tile_maxent, c->tile_short != 0);
  if (sm > 200 * 1024) sm = tile_rank_smem(c->tile_maxent, false);
  if (sm > 32 * 1024) CK(cudaFuncSetAttribute(k_tile_rank, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  k_tile_rank<<<tp, LT_THREADS, sm, st>>>(c->t_he.as<TileEnt>(), c->t_re.as<TileEnt>(), c->t_hoff.as<u32>(),
                                          c->t_roff.as<u32>(), n, (u32)c->tile_margin, c->tile_short ? (u32)sm : 0u,
                                          c->ix.isa, 0u, c->sh_vsend.as<uint4>(), c->d_hw.as<u64>(), nullptr, R * tp,
                                          c->sh_slot.as<u32>());
  hz.alloc(((size_t)ns + 1) * 4, st); sz.alloc(((size_t)ns + 1) * 4, st);
  hoff.alloc(((size_t)ns + 1) * 4, st); boff.alloc(((size_t)ns + 1) * 8, st);
  cnts.alloc((size_t)2 * W * 8, st);
  k_sh_rec_fields<<<(ns + 1 + B - 1) / B, B, 0, st>>>(c->sh_vsend.as<uint4>(), ns, hz.as<u32>(), sz.as<u32>());
  size_t t1 = 0, t2 = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, t1, hz.as<u32>(), hoff.as<u32>(), (int)(ns + 1), st);
  EXSUM64(nullptr, t2, sz.as<u32>(), boff.as<u64>(), (int)(ns + 1), st);
  tmp.alloc(t1 > t2 ? t1 : t2, st);
  size_t tt = t1;
  CK(cub::DeviceScan::ExclusiveSum(tmp.p, tt, hz.as<u32>(), hoff.as<u32>(), (int)(ns + 1), st));
  tt = t2;
  CK(EXSUM64(tmp.p, tt, sz.as<u32>(), boff.as<u64>(), (int)(ns + 1), st));
  if (ns) k_sh_hw_route<<<(ns + B - 1) / B, B, 0, st>>>(c->sh_vsend.as<uint4>(), hoff.as<u32>(), ed.as<u32>(), W, ns,
                                                         c->d_hw.as<u64>(), c->sh_hwsend.as<u64>());
  k_sh_edge_counts<<<1, PLC_MAXW, 0, st>>>(hoff.as<u32>(), boff.as<u64>(), ed.as<u32>(), W, cnts.as<u64>());
  c->launches += 6;
  CK(cudaGetLastError());
  std::vector<u64> hc(2 * W);
  CK(cudaMemcpyAsync(hc.data(), cnts.p, (size_t)2 * W * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  c->sh_v_hwcnt.assign(hc.begin(), hc.begin() + W);
  c->sh_v_bytes.assign(hc.begin() + W, hc.end());
  c->d_hw.release();
  return true;
}

// terminator symbols of this device's documents, in document order (the receiver is device 0)
void shard_term_syms(cmsbwt_ctx *c, u8 *termsym) {
  const u64 H = c->H;
  if (!H) return;
  k_term_syms<<<(u32)((H + 255) / 256), 256, 0, c->st>>>(c->h_len.as<u32>(), c->h_cs.as<u16>(), c->s_sexcl.as<u32>(), H,
                                                        termsym, nullptr);
  LAUNCH(c);
  CK(cudaGetLastError());
}
