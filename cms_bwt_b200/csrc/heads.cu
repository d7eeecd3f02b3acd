// Stage 2: head linking, tail ranking, head keys and the count of non-head suffixes that precede
// every head inside its bucket.  Replaces CMS-BWT-functions.cpp:1359-1603 (head linking, phrase
// sort + keyed insertion, naming, libsais_int, next-rank substitution, prefix sums,
// counterSmallerThanHead).  Formulation (checked on the CPU by
// tests/model_pipeline.py):
//   tail(i)   = text suffix at g_end = gstart + len (shared by all heads with that match end)
//   label(t)  = (ISA[reference position of the suffix at g_end], cls of the phrase containing it)
//   tailrank  = rank of the label sequence along next-tail pointers (prefix doubling)
//   kappa(i)  = (cls_i, tailrank[tail(i)]),  cls = smaller ? E : ~E,  E = pos + len
//   within(h) = #{phrases p : pos_p <= pos_h <= pos_p + toNext_p, kappa_p < kappa_h}
#include "common.cuh"
#include <cub/cub.cuh>
#include <utility>

// ---- linking (toNext :1373, coverage deltas :368-390, tail flags) -------------------------
__global__ void k_link(const u64 *__restrict__ g, const u32 *__restrict__ pos, const u32 *__restrict__ len, u64 H,
                       u32 *__restrict__ toNext, i32 *diff, u32 *__restrict__ tflag, u32 *__restrict__ sflag) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= H) return;
  u32 l = len[i];
  u64 gi = g[i];
  u32 tn = l == 0 ? 0 : (u32)(g[i + 1] - gi - 1);     // a sentinel always follows a real head
  toNext[i] = tn;
  if (diff) {                     // sort-based path only: coverage deltas by global atomics
    u32 p = pos[i];
    atomicAdd(diff + p, 1);
    atomicAdd(diff + p + tn + 1, -1);
  }
  tflag[i] = (i == 0) || (gi + l != g[i - 1] + len[i - 1]);
  sflag[i] = l == 0;
}
// coverage deltas of the sort-based path (:368-390): +1 at pos, -1 past the phrase
__global__ void k_cov_diff(const u32 *pos, const u32 *toNext, u64 H, i32 *diff) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= H) return;
  u32 p = pos[i];
  atomicAdd(diff + p, 1);
  atomicAdd(diff + p + toNext[i] + 1, -1);
}
__global__ void k_tail_first(const u32 *tflag, const u32 *tail_incl, u64 H, u32 *tail_first) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= H) return;
  if (tflag[i]) tail_first[tail_incl[i] - 1] = (u32)i;
}
// label + next pointer of every tail
__global__ void k_tail_label(const u64 *__restrict__ g, const u32 *__restrict__ pos, const u32 *__restrict__ len,
                             const u16 *__restrict__ cs, const u32 *__restrict__ tail_incl,
                             const u32 *__restrict__ sent_excl, const u32 *__restrict__ tflag,
                             const u32 *__restrict__ isa, u64 H, u32 Tn, int qb, u32 doc_base, u64 *__restrict__ label,
                             u32 *__restrict__ next, u32 *__restrict__ tid, u32 *__restrict__ key32) {
  // one thread per HEAD; the first head of a tail (tflag) does the tail's work -- no list of first heads is
  // built (round 2 until q14: k_tail_first + a gather here), and g / len are read where the thread stands
  const u64 f = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= H || !tflag[f]) return;
  const u32 t = tail_incl[f] - 1;
  u64 e = g[f] + len[f];
  // j = last head with gstart <= e (gallop from f, then bisect)
  u64 lo = f, step = 1;
  while (lo + step < H && g[lo + step] <= e) { lo += step; step <<= 1; }
  u64 hi = min(H, lo + step);          // g[hi] > e or hi == H
  while (hi - lo > 1) {
    u64 mid = (lo + hi) >> 1;
    if (g[mid] <= e) lo = mid; else hi = mid;
  }
  u32 j = (u32)lo;
  tid[t] = t;
  // key32 = ISA rank, then qb bits of the cls order (weakly order preserving; tailrank.cu)
  const u32 Q = 1u << qb, half = Q >> 1;
  if (len[j] == 0) {                   // the suffix at g_end is a document terminator
    u32 d = doc_base + sent_excl[j];
    label[t] = (u64)d;
    next[t] = t;
    key32[t] = d < Q - 1 ? d : Q - 1;
  } else {
    u32 pe = pos[j] + (u32)(e - g[j]);
    u32 sm = cs[j] >> 8;
    u32 c = cls_of(pos[j], len[j], sm);
    u32 r = isa[pe];
    label[t] = ((u64)r << 32) | c;
    next[t] = tail_incl[j] - 1;
    u32 rest = (pos[j] + len[j] - pe) >> 5;            // bases left in the phrase, in bins of 32
    u32 f = half ? (rest < half - 1 ? rest : half - 1) : 0;
    key32[t] = (r << qb) | (sm ? f : Q - 1 - f);
  }
}
// dense ranks from sorted keys
__global__ void k_key_flags(const u64 *key, u32 n, u32 *flag) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) flag[j] = j > 0 && key[j] != key[j - 1];
}
__global__ void k_rank_scatter(const u32 *ids, const u32 *dense, u32 n, u32 *rank) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) rank[ids[j]] = dense[j];
}
__global__ void k_double(const u32 *rank, const u32 *next, u32 n, u64 *key, u32 *tid, u32 *next2) {
  u32 t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  u32 nx = next[t];
  key[t] = ((u64)rank[t] << 32) | rank[nx];
  tid[t] = t;
  next2[t] = next[nx];
}

// ---- head keys and tile items ---------------------------------------------------------------
__global__ void k_kappa_count(const u32 *pos, const u32 *len, const u16 *cs, const u32 *toNext,
                              const u32 *tail_incl, const u32 *trank, u64 H, u64 *kappa,
                              u32 *nitems) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= H) return;
  u32 p = pos[i];
  kappa[i] = ((u64)cls_of(p, len[i], cs[i] >> 8) << 32) | trank[tail_incl[i] - 1];
  nitems[i] = (p + toNext[i]) / CMS_TILE_W - p / CMS_TILE_W + 1;
}
// item = (tile, kappa) with payload (head id or ~0, lo, hi); one warp expands 32 heads
__global__ void k_items(const u32 *pos, const u32 *toNext, const u64 *kappa, const u32 *nitems,
                        const u64 *ioff, u64 H, int cshift, u32 *it_tr, u64 *it_key, u64 *it_pay) {
  const u64 cmask = (1ull << cshift) - 1;
  const int lane = threadIdx.x & 31;
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  bool in = i < H;
  u32 p = in ? pos[i] : 0, tn = in ? toNext[i] : 0, ni = in ? nitems[i] : 0;
  u64 kp = in ? kappa[i] : 0, off = in ? ioff[i] : 0;
  u32 t0 = p / CMS_TILE_W;
  if (in && ni <= 4) {
    for (u32 k = 0; k < ni; k++) {
      u32 t = t0 + k, a = t * CMS_TILE_W;
      u32 lo = max(p, a) - a, hi = min(p + tn, a + CMS_TILE_W - 1) - a;
      it_tr[off + k] = (u32)kp;
      it_key[off + k] = ((u64)t << cshift) | ((kp >> 32) & cmask);
      it_pay[off + k] = ((u64)(k == 0 ? (u32)i : 0xFFFFFFFFu) << 32) | (lo << 8) | hi;
    }
  }
  unsigned big = __ballot_sync(0xffffffffu, in && ni > 4);
  while (big) {
    int src = __ffs(big) - 1;
    big &= big - 1;
    u32 bp = __shfl_sync(0xffffffffu, p, src), btn = __shfl_sync(0xffffffffu, tn, src);
    u32 bni = __shfl_sync(0xffffffffu, ni, src), bt0 = __shfl_sync(0xffffffffu, t0, src);
    u64 bkp = __shfl_sync(0xffffffffu, kp, src), boff = __shfl_sync(0xffffffffu, off, src);
    u64 bi = i - lane + src;
    for (u32 k = lane; k < bni; k += 32) {
      u32 t = bt0 + k, a = t * CMS_TILE_W;
      u32 lo = max(bp, a) - a, hi = min(bp + btn, a + CMS_TILE_W - 1) - a;
      it_tr[boff + k] = (u32)bkp;
      it_key[boff + k] = ((u64)t << cshift) | ((bkp >> 32) & cmask);
      it_pay[boff + k] = ((u64)(k == 0 ? (u32)bi : 0xFFFFFFFFu) << 32) | (lo << 8) | hi;
    }
  }
}
__global__ void k_iota(u32 *a, u64 n) {
  u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) a[j] = (u32)j;
}
__global__ void k_gather_keys(const u32 *idx, const u64 *key, u64 n, u64 *out) {
  u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) out[j] = key[idx[j]];
}
__global__ void k_tile_bounds(const u64 *skey, u64 n, int cshift, u32 *tbeg, u32 *tend) {
  u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  u32 t = (u32)(skey[j] >> cshift);
  if (j == 0 || (u32)(skey[j - 1] >> cshift) != t) tbeg[t] = (u32)j;
  if (j + 1 == n || (u32)(skey[j + 1] >> cshift) != t) tend[t] = (u32)j + 1;
}
// one warp per tile: walk the tile's items in kappa order, 32 at a time; for every head item
// count the earlier items whose interval holds the head's position.  A tile is walked by ONE warp, so
// nothing else hides the three dependent reads of an item (sorted index -> payload -> position):
// they are issued TC_K blocks of 32 at a time, one group ahead of their use.
#define TC_K 8
#define TC_NONE 0xFFFFFFFF00000100ull       // no head, empty interval (lo 1 > hi 0)
__device__ __forceinline__ void tc_load(const u32 *__restrict__ sidx, const u64 *__restrict__ it_pay,
                                        const u32 *__restrict__ pos, u32 j0, u32 e, int lane, u64 *P, u32 *Q) {
  u32 s[TC_K];
#pragma unroll
  for (int k = 0; k < TC_K; k++) { const u32 j = j0 + 32u * k + lane; s[k] = j < e ? sidx[j] : 0xFFFFFFFFu; }
#pragma unroll
  for (int k = 0; k < TC_K; k++) P[k] = s[k] != 0xFFFFFFFFu ? it_pay[s[k]] : TC_NONE;
#pragma unroll
  for (int k = 0; k < TC_K; k++) { const u32 hid = (u32)(P[k] >> 32); Q[k] = hid != 0xFFFFFFFFu ? pos[hid] : 0u; }
}
// A tile of more than TC_CHUNK items is cut into chunks walked by different warps (a collection of
// 300,000 documents over 30 kbp has 470 tiles of 4x10^5 items each): work item w = (tile, chunk),
// woff[t] = first work item of tile t.  The coverage counters a chunk starts from are the column sums
// of the chunks before it (k_tc_cover + k_tc_prefix).
#define TC_CHUNK 2048u
__global__ void k_tc_nchunks(const u32 *tbeg, const u32 *tend, u32 ntiles, u32 *nch) {
  u32 t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t <= ntiles) nch[t] = t < ntiles ? (tend[t] - tbeg[t] + TC_CHUNK - 1) / TC_CHUNK : 0u;
}
__device__ __forceinline__ u32 tc_tile_of(const u32 *woff, u32 ntiles, u32 w) {
  u32 lo = 0, hi = ntiles;                    // last t with woff[t] <= w
  while (hi - lo > 1) { u32 mid = lo + ((hi - lo) >> 1); if (woff[mid] <= w) lo = mid; else hi = mid; }
  return lo;
}
// coverage of the 64 tile offsets by the items of one chunk (difference array in shared memory)
__global__ void __launch_bounds__(256) k_tc_cover(const u32 *__restrict__ sidx, const u64 *__restrict__ it_pay,
                                                  const u32 *tbeg, const u32 *tend, const u32 *woff, u32 ntiles,
                                                  u32 W, u32 *cov) {
  __shared__ int s_d[8][CMS_TILE_W + 2];
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  const u32 w = (u32)(((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  if (w >= W) return;
  const u32 t = tc_tile_of(woff, ntiles, w), c = w - woff[t];
  if (c + 1 == woff[t + 1] - woff[t]) return;             // nothing comes after the last chunk
  const u32 b = tbeg[t] + c * TC_CHUNK, e = min(tend[t], b + TC_CHUNK);
  s_d[wp][lane] = 0; s_d[wp][lane + 32] = 0;
  if (lane < 2) s_d[wp][64 + lane] = 0;
  __syncwarp();
  for (u32 j = b + lane; j < e; j += 32) {
    const u64 pay = it_pay[sidx[j]];
    const u32 lo = (u32)(pay >> 8) & 0xFF, hi = (u32)pay & 0xFF;
    if (lo <= hi) { atomicAdd(&s_d[wp][lo], 1); atomicAdd(&s_d[wp][hi + 1], -1); }
  }
  __syncwarp();
  int v0 = s_d[wp][2 * lane], v1 = s_d[wp][2 * lane + 1] + v0;     // offsets 2 lane, 2 lane + 1
  int incl = v1;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { int o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += o; }
  const int pre = incl - v1;
  cov[(size_t)w * CMS_TILE_W + 2 * lane] = (u32)(v0 + pre);
  cov[(size_t)w * CMS_TILE_W + 2 * lane + 1] = (u32)(v1 + pre);
}
// exclusive prefix of the chunk coverages inside every multi-chunk tile (one warp per tile)
__global__ void k_tc_prefix(const u32 *woff, u32 ntiles, u32 *cov) {
  const int lane = threadIdx.x & 31;
  const u32 t = (u32)(((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  if (t >= ntiles) return;
  const u32 w0 = woff[t], w1 = woff[t + 1];
  if (w1 - w0 < 2) return;
  u32 r0 = 0, r1 = 0;
  for (u32 w = w0; w < w1; w++) {
    u32 *p = cov + (size_t)w * CMS_TILE_W;
    const u32 a = p[lane], b = p[lane + 32];
    p[lane] = r0; p[lane + 32] = r1;
    r0 += a; r1 += b;
  }
}
__global__ void k_tile_count(const u32 *__restrict__ sidx, const u64 *__restrict__ it_pay,
                             const u32 *__restrict__ pos, const u32 *tbeg, const u32 *tend, const u32 *woff,
                             u32 ntiles, u32 W, const u32 *cov, u32 *__restrict__ within) {
  const int lane = threadIdx.x & 31;
  const u32 w = (u32)(((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  if (w >= W) return;
  const u32 t = tc_tile_of(woff, ntiles, w);
  const u32 b = tbeg[t] + (w - woff[t]) * TC_CHUNK, e = min(tend[t], b + TC_CHUNK);
  if (b >= e) return;
  // items so far covering tile offsets lane / lane + 32
  u32 run0 = cov[(size_t)w * CMS_TILE_W + lane], run1 = cov[(size_t)w * CMS_TILE_W + lane + 32];
  u64 pay[TC_K], npay[TC_K];
  u32 ps[TC_K], nps[TC_K];
  tc_load(sidx, it_pay, pos, b, e, lane, npay, nps);
  for (u32 g0 = b; g0 < e; g0 += 32u * TC_K) {
#pragma unroll
    for (int k = 0; k < TC_K; k++) { pay[k] = npay[k]; ps[k] = nps[k]; }
    if (g0 + 32u * TC_K < e) tc_load(sidx, it_pay, pos, g0 + 32u * TC_K, e, lane, npay, nps);
#pragma unroll
    for (int k = 0; k < TC_K; k++) {
      if (g0 + 32u * k >= e) break;                       // uniform over the warp
      const u32 hid = (u32)(pay[k] >> 32), lo = (u32)(pay[k] >> 8) & 0xFF, hi = (u32)pay[k] & 0xFF;
      const u32 q = hid != 0xFFFFFFFFu ? ps[k] - t * CMS_TILE_W : 0u;
      // the select must happen on the reading lane, so fetch both counters and pick here
      u32 r0 = __shfl_sync(0xffffffffu, run0, q & 31), r1 = __shfl_sync(0xffffffffu, run1, q & 31);
      u32 base = q < 32 ? r0 : r1;
      u32 cnt = 0, a0 = 0, a1 = 0;
      for (int jj = 0; jj < 32; jj++) {
        u32 l2 = __shfl_sync(0xffffffffu, lo, jj), h2 = __shfl_sync(0xffffffffu, hi, jj);
        cnt += (jj < lane) & (l2 <= q) & (q <= h2);
        a0 += (l2 <= (u32)lane) & ((u32)lane <= h2);
        a1 += (l2 <= (u32)lane + 32) & ((u32)lane + 32 <= h2);
      }
      if (hid != 0xFFFFFFFFu) within[hid] = base + cnt;
      run0 += a0; run1 += a1;
    }
  }
}

struct Tmp {                      // growable CUB scratch
  DBuf b;
  void need(size_t bytes, cudaStream_t st) { if (bytes > b.bytes) b.alloc(bytes, st); }
};

// Sort-based counting over I items (tailrank it_tr, key = tile << cshift | cls, payload = head id or ~0,
// lo, hi inside the 64-position tile): stable LSD by tailrank, then by (tile, cls); one warp per tile
// walks its items in kappa order.  pos[head id] = position in the same tile numbering.  Items of
// tile `ntiles` (one past the last) are padding and are never visited.  Used by the single-GPU
// fallback below and by the owners of deep tile ranges on the multi-GPU path (tiles.cu).
void sweep_count(cmsbwt_ctx *c, u64 I, DBuf &it_tr, DBuf &it_key, DBuf &it_pay, const u32 *pos, u32 ntiles,
                 int tbits, int cshift, u32 *within) {
  cudaStream_t st = c->st;
  const int B = 256;
  if (I == 0) return;
  if (I >= 0x7FFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "more than 2^31 tile items on one device"};
  Tmp tmp;
  size_t tb;
  DBuf it_tr2, it_key2, idx, idx2;
  it_tr2.alloc(I * 4, st); it_key2.alloc(I * 8, st); idx.alloc(I * 4, st); idx2.alloc(I * 4, st);
  const u32 gI = (u32)((I + B - 1) / B);
  k_iota<<<gI, B, 0, st>>>(idx2.as<u32>(), I);
  rsort_pairs32(c, it_tr, it_tr2, idx2, idx, I, 0, tbits);
  it_tr.swap(it_tr2); idx2.swap(idx);             // sorted by tail rank: it_tr2 / idx
  k_gather_keys<<<gI, B, 0, st>>>(idx.as<u32>(), it_key.as<u64>(), I, it_key2.as<u64>());
  int kb = cshift + 1; while ((1ull << (kb - cshift)) < (u64)ntiles + 1) kb++;
  rsort_pairs64(c, it_key2, it_key, idx, idx2, I, 0, kb);       // stable: the tail-rank order survives inside (tile, cls)
  it_key2.swap(it_key); idx.swap(idx2);           // sorted by (tile, cls): it_key / idx2
  DBuf tbeg, tend;
  tbeg.alloc(((size_t)ntiles + 1) * 4, st); tend.alloc(((size_t)ntiles + 1) * 4, st);
  CK(cudaMemsetAsync(tbeg.p, 0, ((size_t)ntiles + 1) * 4, st));
  CK(cudaMemsetAsync(tend.p, 0, ((size_t)ntiles + 1) * 4, st));
  k_tile_bounds<<<gI, B, 0, st>>>(it_key.as<u64>(), I, cshift, tbeg.as<u32>(), tend.as<u32>());
  DBuf nch, woff, cov;
  nch.alloc(((size_t)ntiles + 1) * 4, st); woff.alloc(((size_t)ntiles + 1) * 4, st);
  k_tc_nchunks<<<(ntiles + 1 + B - 1) / B, B, 0, st>>>(tbeg.as<u32>(), tend.as<u32>(), ntiles, nch.as<u32>());
  tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, nch.as<u32>(), woff.as<u32>(), (int)(ntiles + 1), st);
  tmp.need(tb, st);
  CK(cub::DeviceScan::ExclusiveSum(tmp.b.p, tb, nch.as<u32>(), woff.as<u32>(), (int)(ntiles + 1), st));
  u32 W = 0;
  CK(cudaMemcpyAsync(&W, woff.as<u32>() + ntiles, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (W) {
    cov.alloc((size_t)W * CMS_TILE_W * 4, st);
    CK(cudaMemsetAsync(cov.p, 0, (size_t)W * CMS_TILE_W * 4, st));
    const u32 gW = (u32)(((u64)W * 32 + B - 1) / B);
    if (I > TC_CHUNK) {                                    // else no tile can have a second chunk
      k_tc_cover<<<gW, B, 0, st>>>(idx2.as<u32>(), it_pay.as<u64>(), tbeg.as<u32>(), tend.as<u32>(), woff.as<u32>(),
                                   ntiles, W, cov.as<u32>());
      k_tc_prefix<<<(u32)(((u64)ntiles * 32 + B - 1) / B), B, 0, st>>>(woff.as<u32>(), ntiles, cov.as<u32>());
      c->launches += 2;
    }
    k_tile_count<<<gW, B, 0, st>>>(idx2.as<u32>(), it_pay.as<u64>(), pos, tbeg.as<u32>(), tend.as<u32>(),
                                   woff.as<u32>(), ntiles, W, cov.as<u32>(), within);
  }
  c->launches += 4;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));
}

// link + tail labels (everything here is local to a shard of whole documents: phrases, tails and
// next-tail pointers never leave a document).  doc_base = documents on lower shards.
void sort_prepare(cmsbwt_ctx *c, u64 doc_base) {
  cudaStream_t st = c->st;
  const u64 H = c->H;
  const u32 n = c->ix.n;
  const int B = 256;
  const u32 gH = (u32)((H + B - 1) / B);
  const u64 *g = c->h_gstart.as<u64>();
  const u32 *pos = c->h_pos.as<u32>(), *len = c->h_len.as<u32>();
  const u16 *cs = c->h_cs.as<u16>();
  Tmp tmp;
  size_t tb;
  c->h_toNext.alloc(H * 4, st);
  c->light = false;
  c->stats.light_tiles = 0;
  c->stats.pair_tests = 0;
  DBuf tflag, sflag, tfirst;
  tflag.alloc(H * 4, st); sflag.alloc(H * 4, st); c->s_tincl.alloc(H * 4, st); c->s_sexcl.alloc(H * 4, st);
  k_link<<<gH, B, 0, st>>>(g, pos, len, H, c->h_toNext.as<u32>(), nullptr, tflag.as<u32>(),
                           sflag.as<u32>());
  LAUNCH(c);
  tb = 0; cub::DeviceScan::InclusiveSum(nullptr, tb, tflag.as<u32>(), c->s_tincl.as<u32>(), (int)H, st);
  tmp.need(tb, st);
  CK(cub::DeviceScan::InclusiveSum(tmp.b.p, tb, tflag.as<u32>(), c->s_tincl.as<u32>(), (int)H, st));
  tb = 0; cub::DeviceScan::ExclusiveSum(nullptr, tb, sflag.as<u32>(), c->s_sexcl.as<u32>(), (int)H, st);
  tmp.need(tb, st);
  CK(cub::DeviceScan::ExclusiveSum(tmp.b.p, tb, sflag.as<u32>(), c->s_sexcl.as<u32>(), (int)H, st));
  c->launches += 2;
  u32 Tn = 0, lastS = 0, lastF = 0;
  CK(cudaMemcpyAsync(&Tn, c->s_tincl.as<u32>() + (H - 1), 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(&lastS, c->s_sexcl.as<u32>() + (H - 1), 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(&lastF, sflag.as<u32>() + (H - 1), 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  c->stats.n_docs = (u64)lastS + lastF;
  c->stats.n_tails = Tn;
  c->s_Tn = Tn;
  const u32 gT = (Tn + B - 1) / B;
  c->s_label.alloc((size_t)Tn * 8, st);
  c->s_next.alloc((size_t)Tn * 4, st); c->s_tid.alloc((size_t)Tn * 4, st);
  c->s_trank.alloc((size_t)Tn * 4, st);
  int isa_bits = 1;
  while (isa_bits < 32 && (1ull << isa_bits) < (u64)n) isa_bits++;
  const int qb = 32 - isa_bits < 6 ? 32 - isa_bits : 6;
  c->s_key_bits = isa_bits + qb;
  c->s_key32.alloc((size_t)Tn * 4, st);
  k_tail_label<<<gH, B, 0, st>>>(g, pos, len, cs, c->s_tincl.as<u32>(), c->s_sexcl.as<u32>(), tflag.as<u32>(),
                                 c->ix.isa, H, Tn, qb, (u32)doc_base, c->s_label.as<u64>(), c->s_next.as<u32>(),
                                 c->s_tid.as<u32>(), c->s_key32.as<u32>());
  c->launches += 1;
  CK(cudaGetLastError());
}

// kappa per head from the tail ranks (local)
void sort_kappa(cmsbwt_ctx *c) {
  cudaStream_t st = c->st;
  const u64 H = c->H;
  const int B = 256;
  const u32 gH = (u32)((H + B - 1) / B);
  c->s_kappa.alloc(H * 8, st); c->s_nitems.alloc((H + 1) * 4, st);
  CK(cudaMemsetAsync(c->s_nitems.p, 0, (H + 1) * 4, st));
  k_kappa_count<<<gH, B, 0, st>>>(c->h_pos.as<u32>(), c->h_len.as<u32>(), c->h_cs.as<u16>(), c->h_toNext.as<u32>(),
                                  c->s_tincl.as<u32>(), c->s_trank.as<u32>(), H, c->s_kappa.as<u64>(),
                                  c->s_nitems.as<u32>());
  LAUNCH(c);
  c->s_label.release(); c->s_next.release(); c->s_tid.release(); c->s_key32.release();
  CK(cudaGetLastError());
}

void sort_stage(cmsbwt_ctx *c) {
  cudaStream_t st = c->st;
  const u64 H = c->H;
  const u32 n = c->ix.n;
  const int B = 256;
  const u32 gH = (u32)((H + B - 1) / B);
  const u32 *pos = c->h_pos.as<u32>();
  Tmp tmp;
  size_t tb;
  u32 Tn;
  {
    GroupSpan g(c, CMSBWT_G_RANK);
    sort_prepare(c, 0);
    Tn = c->s_Tn;
    tail_rank(c, c->s_label.as<u64>(), c->s_key32, c->s_tid, c->s_key_bits, c->s_next.as<u32>(), Tn,
              c->s_trank.as<u32>());
    c->s_label.release(); c->s_next.release(); c->s_tid.release(); c->s_key32.release();
  }
  DBuf &kappa = c->s_kappa, &nitems = c->s_nitems, &sexcl = c->s_sexcl;
  // ---- brute-force tile path (tiles.cu) unless the collection is too deep for it; the tile scatter forms
  // kappa itself, the array is only built (sort_kappa) for the paths that need it more than once ----
  bool light;
  { GroupSpan g(c, CMSBWT_G_TILES); light = c->tile_budget > 0 && tiles_build(c, nullptr, sexcl); }
  if (!light && !c->s_kappa.p) { GroupSpan g(c, CMSBWT_G_RANK); sort_kappa(c); }
  if (light) {
    c->stats.light_tiles = 1;
    c->s_kappa.release(); c->s_nitems.release(); c->s_tincl.release(); c->s_trank.release();
    CK(cudaStreamSynchronize(st));
    return;
  }
  // ---- sort-based sweep ----
  GroupSpan gsweep(c, CMSBWT_G_PLACE);
  DBuf ioff;
  ioff.alloc((H + 1) * 8, st);
  tb = 0; EXSUM64(nullptr, tb, nitems.as<u32>(), ioff.as<u64>(), (int)(H + 1), st);
  tmp.need(tb, st);
  CK(EXSUM64(tmp.b.p, tb, nitems.as<u32>(), ioff.as<u64>(), (int)(H + 1), st));
  c->launches += 2;
  u64 I = 0;
  CK(cudaMemcpyAsync(&I, ioff.as<u64>() + H, 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (I >= 0x7FFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "more than 2^31 tile items on one device"};
  c->stats.n_items = I;
  c->d_cov.alloc(((size_t)n + 1) * 4, st);
  CK(cudaMemsetAsync(c->d_cov.p, 0, ((size_t)n + 1) * 4, st));
  k_cov_diff<<<gH, B, 0, st>>>(pos, c->h_toNext.as<u32>(), H, c->d_cov.as<i32>());
  LAUNCH(c);
  c->s_tincl.release(); c->s_trank.release();
  DBuf it_tr, it_key, it_pay;
  it_tr.alloc(I * 4, st); it_key.alloc(I * 8, st); it_pay.alloc(I * 8, st);
  const int cshift = sweep_cls_bits(n);
  k_items<<<gH, B, 0, st>>>(pos, c->h_toNext.as<u32>(), kappa.as<u64>(), nitems.as<u32>(),
                            ioff.as<u64>(), H, cshift, it_tr.as<u32>(), it_key.as<u64>(), it_pay.as<u64>());
  LAUNCH(c);
  {
    int tbits = 1; while ((1ull << tbits) < (u64)Tn) tbits++;
    c->h_within.alloc(H * 4, st);
    sweep_count(c, I, it_tr, it_key, it_pay, pos, n / CMS_TILE_W + 1, tbits, cshift, c->h_within.as<u32>());
  }
  c->s_kappa.release(); c->s_nitems.release(); c->s_sexcl.release();
}
