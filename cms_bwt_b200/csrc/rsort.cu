// Product-owned LSD radix sort (north-star subsystem 3: "a segmented LSD radix sort keyed on reference
// ISA rank").  Replaces the reference's std::sort of the phrases (CMS-BWT-functions.cpp:1387) and its
// libsais_int call on the name string (:1456) as the ordering primitive of stage 2, and every
// cub::DeviceRadixSort call the library used to make (tail keys, tie refinement, separators, sweep items,
// prefix doubling of the GPU suffix array).
//
// One pass per 8-bit digit, each pass ONE sweep over the data ("onesweep": Adinets & Merrill): the
// digit histograms of all passes come from one upfront read; a tile (4096 keys, one CTA of 256 threads)
// ranks its keys with warp-private digit counters (match_any: no shared-memory atomics), publishes its
// digit counts in a status word per (tile, digit) and obtains the counts of all earlier tiles by
// decoupled look-back; keys (and 32-bit values) are reordered in shared memory first so that every
// digit's run leaves the SM as one contiguous segment.  Tiles take their index from an atomic ticket,
// so a tile only ever waits for tiles that are already running.  The sort is stable.  Inputs of at most
// one tile are sorted by a single CTA entirely in shared memory (all passes, one launch).
#include "common.cuh"
#include <cub/block/block_scan.cuh>

#define RS_THREADS 256
#define RS_ITEMS 16
#define RS_TILE (RS_THREADS * RS_ITEMS)
#define RS_WARPS (RS_THREADS / 32)
#define RS_RADIX 256
#define RS_FLAG_AGG (1ull << 62)
#define RS_FLAG_INCL (1ull << 63)
#define RS_MASK ((1ull << 62) - 1)

template <class K>
__global__ void __launch_bounds__(RS_THREADS) k_rs_hist(const K *__restrict__ keys, u64 n, int begin_bit, int npass,
                                                        unsigned long long *ghist) {
  __shared__ u32 s_h[8][RS_RADIX];
  for (u32 k = threadIdx.x; k < 8 * RS_RADIX; k += RS_THREADS) (&s_h[0][0])[k] = 0;
  __syncthreads();
  // a block owns a contiguous slice small enough for 32-bit counters
  const u64 per = (n + gridDim.x - 1) / gridDim.x;
  const u64 a = (u64)blockIdx.x * per, b = min(n, a + per);
  for (u64 i = a + threadIdx.x; i < b; i += RS_THREADS) {
    const K key = keys[i];
#pragma unroll
    for (int p = 0; p < 8; p++)
      if (p < npass) atomicAdd(&s_h[p][(u32)(key >> (begin_bit + 8 * p)) & 255u], 1u);
  }
  __syncthreads();
  for (u32 k = threadIdx.x; k < (u32)npass * RS_RADIX; k += RS_THREADS) {
    const u32 v = (&s_h[0][0])[k];
    if (v) atomicAdd(ghist + k, (unsigned long long)v);
  }
}

// exclusive scan of every pass's 256 counts (one block per pass)
__global__ void __launch_bounds__(RS_RADIX) k_rs_scan(unsigned long long *ghist) {
  typedef cub::BlockScan<unsigned long long, RS_RADIX> BS;
  __shared__ typename BS::TempStorage ts;
  unsigned long long *h = ghist + (size_t)blockIdx.x * RS_RADIX;
  unsigned long long v = h[threadIdx.x], ex;
  BS(ts).ExclusiveSum(v, ex);
  h[threadIdx.x] = ex;
}

// ranks the RS_ITEMS keys of this thread among the tile's keys of the same digit (stable in tile order
// = warp-major, round, lane): returns in rank[] the position inside (warp, digit) and leaves the
// per-warp digit counts in s_cnt[warp][digit]
template <class K>
__device__ __forceinline__ void rs_rank(const K (&key)[RS_ITEMS], int shift, u32 (*s_cnt)[RS_RADIX], u32 (&rank)[RS_ITEMS]) {
  const u32 w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const u32 lt = (1u << l) - 1u;
#pragma unroll
  for (int r = 0; r < RS_ITEMS; r++) {
    const u32 d = (u32)(key[r] >> shift) & 255u;
    const unsigned m = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(m) - 1;
    // one shared-memory atomic per (round, digit): a warp's atomics on one address execute in program order,
    // so the rounds need no barrier between them and their latencies overlap (until q15 a read-modify-
    // write by the leader with __syncwarp per round: the 16 rounds were one serial chain; ncu: 15 of 25
    // stall cycles per issue on the short scoreboard)
    u32 old = 0;
    if ((int)l == leader) old = atomicAdd(&s_cnt[w][d], (u32)__popc(m));
    old = __shfl_sync(0xffffffffu, old, leader);
    rank[r] = old + __popc(m & lt);
  }
}

template <class K, bool VALS>
__global__ void __launch_bounds__(RS_THREADS, sizeof(K) == 4 ? 3 : 2) k_rs_pass(const K *__restrict__ kin, K *__restrict__ kout,
                                                        const u32 *__restrict__ vin, u32 *__restrict__ vout, u64 n,
                                                        int shift, const unsigned long long *__restrict__ gbase,
                                                        volatile unsigned long long *status, u32 *ticket) {
  extern __shared__ __align__(16) unsigned char rs_smem[];
  K *s_keys = (K *)rs_smem;
  u32 *s_vals = (u32 *)(s_keys + RS_TILE);
  __shared__ u32 s_cnt[RS_WARPS][RS_RADIX];
  __shared__ u32 s_excl[RS_RADIX];
  __shared__ long long s_gofs[RS_RADIX];
  __shared__ u32 s_tile;
  typedef cub::BlockScan<u32, RS_THREADS> BS;
  __shared__ typename BS::TempStorage ts;
  const u32 tid = threadIdx.x, w = tid >> 5, l = tid & 31;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);
  for (u32 k = tid; k < RS_WARPS * RS_RADIX; k += RS_THREADS) (&s_cnt[0][0])[k] = 0;
  __syncthreads();
  const u32 tile = s_tile;
  const u64 base = (u64)tile * RS_TILE;
  const u32 nvalid = (u32)min((u64)RS_TILE, n - base);
  K key[RS_ITEMS];
  u32 val[RS_ITEMS], rank[RS_ITEMS];
#pragma unroll
  for (int r = 0; r < RS_ITEMS; r++) {
    const u32 li = w * (32 * RS_ITEMS) + r * 32 + l;
    key[r] = li < nvalid ? kin[base + li] : (K)~(K)0;     // padding sorts behind every key of the last digit
  }
  rs_rank<K>(key, shift, s_cnt, rank);
  if (VALS) {                                             // values are not live while the keys are ranked
#pragma unroll
    for (int r = 0; r < RS_ITEMS; r++) {
      const u32 li = w * (32 * RS_ITEMS) + r * 32 + l;
      val[r] = li < nvalid ? vin[base + li] : 0u;
    }
  }
  __syncthreads();
  // exclusive prefix over the warps per digit; tile counts; their exclusive scan over the digits
  u32 tile_cnt = 0;
#pragma unroll
  for (int ww = 0; ww < RS_WARPS; ww++) { const u32 cw = s_cnt[ww][tid]; s_cnt[ww][tid] = tile_cnt; tile_cnt += cw; }
  // publish this tile's counts at once: later tiles can already add them
  if (tile == 0) status[tid] = RS_FLAG_INCL | (unsigned long long)tile_cnt;
  else status[(size_t)tile * RS_RADIX + tid] = RS_FLAG_AGG | (unsigned long long)tile_cnt;
  u32 ex;
  BS(ts).ExclusiveSum(tile_cnt, ex);
  s_excl[tid] = ex;
  __syncthreads();
#pragma unroll
  for (int r = 0; r < RS_ITEMS; r++) {
    const u32 d = (u32)(key[r] >> shift) & 255u;
    const u32 pos = s_excl[d] + s_cnt[w][d] + rank[r];
    s_keys[pos] = key[r];
    if (VALS) s_vals[pos] = val[r];
  }
  // decoupled look-back: thread d sums digit d over the earlier tiles
  {
    unsigned long long prev = 0;
    if (tile > 0) {
      size_t t = (size_t)tile - 1;
      for (;;) {
        const unsigned long long v = status[t * RS_RADIX + tid];
        if (v & RS_FLAG_INCL) { prev += v & RS_MASK; break; }
        if (v & RS_FLAG_AGG) { prev += v & RS_MASK; t--; }
      }
      status[(size_t)tile * RS_RADIX + tid] = RS_FLAG_INCL | (prev + tile_cnt);
    }
    s_gofs[tid] = (long long)(gbase[tid] + prev) - (long long)ex;   // global position = s_gofs[digit] + local index
  }
  __syncthreads();
  for (u32 i = tid; i < nvalid; i += RS_THREADS) {
    const K k = s_keys[i];
    const u32 d = (u32)(k >> shift) & 255u;
    const u64 g = (u64)(s_gofs[d] + (long long)i);
    kout[g] = k;
    if (VALS) vout[g] = s_vals[i];
  }
}

// n <= RS_TILE: every pass inside one CTA, ping-pong between two shared-memory buffers
template <class K, bool VALS>
__global__ void __launch_bounds__(RS_THREADS) k_rs_single(const K *__restrict__ kin, K *__restrict__ kout,
                                                          const u32 *__restrict__ vin, u32 *__restrict__ vout, u32 n,
                                                          int begin_bit, int npass) {
  extern __shared__ __align__(16) unsigned char rs_smem[];
  K *s_k[2] = {(K *)rs_smem, (K *)rs_smem + RS_TILE};
  u32 *s_v[2] = {(u32 *)((K *)rs_smem + 2 * RS_TILE), (u32 *)((K *)rs_smem + 2 * RS_TILE) + RS_TILE};
  __shared__ u32 s_cnt[RS_WARPS][RS_RADIX];
  __shared__ u32 s_excl[RS_RADIX];
  typedef cub::BlockScan<u32, RS_THREADS> BS;
  __shared__ typename BS::TempStorage ts;
  const u32 tid = threadIdx.x, w = tid >> 5, l = tid & 31;
  for (u32 i = tid; i < RS_TILE; i += RS_THREADS) {
    s_k[0][i] = i < n ? kin[i] : (K)~(K)0;
    if (VALS) s_v[0][i] = i < n ? vin[i] : 0u;
  }
  int cur = 0;
  for (int p = 0; p < npass; p++) {
    const int shift = begin_bit + 8 * p;
    for (u32 k = tid; k < RS_WARPS * RS_RADIX; k += RS_THREADS) (&s_cnt[0][0])[k] = 0;
    __syncthreads();
    K key[RS_ITEMS];
    u32 val[RS_ITEMS], rank[RS_ITEMS];
#pragma unroll
    for (int r = 0; r < RS_ITEMS; r++) {
      const u32 li = w * (32 * RS_ITEMS) + r * 32 + l;
      key[r] = s_k[cur][li];
      if (VALS) val[r] = s_v[cur][li];
    }
    rs_rank<K>(key, shift, s_cnt, rank);
    __syncthreads();
    u32 tile_cnt = 0;
#pragma unroll
    for (int ww = 0; ww < RS_WARPS; ww++) { const u32 cw = s_cnt[ww][tid]; s_cnt[ww][tid] = tile_cnt; tile_cnt += cw; }
    u32 ex;
    BS(ts).ExclusiveSum(tile_cnt, ex);
    s_excl[tid] = ex;
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RS_ITEMS; r++) {
      const u32 d = (u32)(key[r] >> shift) & 255u;
      const u32 pos = s_excl[d] + s_cnt[w][d] + rank[r];
      s_k[cur ^ 1][pos] = key[r];
      if (VALS) s_v[cur ^ 1][pos] = val[r];
    }
    cur ^= 1;
    __syncthreads();
  }
  // the all-ones padding has the largest digit in every pass and starts behind every real key, so (the
  // passes are stable) it stays behind them: the first n entries are the sorted input
  for (u32 i = tid; i < n; i += RS_THREADS) {
    kout[i] = s_k[cur][i];
    if (VALS) vout[i] = s_v[cur][i];
  }
}

template <class K>
static void rsort_run(cmsbwt_ctx *c, DBuf &k, DBuf &k2, DBuf *v, DBuf *v2, u64 n, int begin_bit, int end_bit) {
  cudaStream_t st = c->st;
  if (n == 0 || end_bit <= begin_bit) return;
  const bool vals = v != nullptr;
  int npass = (end_bit - begin_bit + 7) / 8;
  if (npass > 8) npass = 8;
  if (n <= RS_TILE) {
    const size_t sm = (size_t)RS_TILE * 2 * (sizeof(K) + 4);
    if (vals) {
      CK(cudaFuncSetAttribute(k_rs_single<K, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
      k_rs_single<K, true><<<1, RS_THREADS, sm, st>>>(k.as<K>(), k2.as<K>(), v->as<u32>(), v2->as<u32>(), (u32)n, begin_bit, npass);
    } else {
      CK(cudaFuncSetAttribute(k_rs_single<K, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
      k_rs_single<K, false><<<1, RS_THREADS, sm, st>>>(k.as<K>(), k2.as<K>(), nullptr, nullptr, (u32)n, begin_bit, npass);
    }
    LAUNCH(c);
    CK(cudaGetLastError());
    k.swap(k2);
    if (vals) v->swap(*v2);
    return;
  }
  const u64 ntiles = (n + RS_TILE - 1) / RS_TILE;
  if (ntiles >= 0xFFFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "radix sort: too many tiles"};
  DBuf ghist, status, ticket;
  ghist.alloc((size_t)npass * RS_RADIX * 8, st);
  status.alloc((size_t)ntiles * RS_RADIX * 8, st);
  ticket.alloc(4, st);
  CK(cudaMemsetAsync(ghist.p, 0, (size_t)npass * RS_RADIX * 8, st));
  u64 hb = (n + (1u << 12) - 1) >> 12;            // contiguous slices (n < 2^40: 32-bit block counters are safe)
  if (hb > 148 * 8) hb = 148 * 8;
  k_rs_hist<K><<<(u32)hb, RS_THREADS, 0, st>>>(k.as<K>(), n, begin_bit, npass, ghist.as<unsigned long long>());
  k_rs_scan<<<npass, RS_RADIX, 0, st>>>(ghist.as<unsigned long long>());
  c->launches += 2;
  const size_t sm = (size_t)RS_TILE * (sizeof(K) + (vals ? 4 : 0));
  if (vals) CK(cudaFuncSetAttribute(k_rs_pass<K, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  else CK(cudaFuncSetAttribute(k_rs_pass<K, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  for (int p = 0; p < npass; p++) {
    CK(cudaMemsetAsync(status.p, 0, (size_t)ntiles * RS_RADIX * 8, st));
    CK(cudaMemsetAsync(ticket.p, 0, 4, st));
    if (vals)
      k_rs_pass<K, true><<<(u32)ntiles, RS_THREADS, sm, st>>>(k.as<K>(), k2.as<K>(), v->as<u32>(), v2->as<u32>(), n,
                                                             begin_bit + 8 * p, ghist.as<unsigned long long>() + (size_t)p * RS_RADIX,
                                                             status.as<unsigned long long>(), ticket.as<u32>());
    else
      k_rs_pass<K, false><<<(u32)ntiles, RS_THREADS, sm, st>>>(k.as<K>(), k2.as<K>(), nullptr, nullptr, n, begin_bit + 8 * p,
                                                              ghist.as<unsigned long long>() + (size_t)p * RS_RADIX,
                                                              status.as<unsigned long long>(), ticket.as<u32>());
    LAUNCH(c);
    k.swap(k2);
    if (vals) v->swap(*v2);
  }
  CK(cudaGetLastError());
}

// Sorts n keys (and their 32-bit values) by the bits [begin_bit, end_bit) of the key, stable.  k / v hold
// the input and, on return, the sorted output; k2 / v2 are scratch of the same size (the buffers are
// swapped, not copied).  Whole 8-bit digits are used: bits of the key at or above end_bit rounded up to
// a byte take part in the order, so callers keep them zero (all of this library's keys do).
void rsort_pairs32(cmsbwt_ctx *c, DBuf &k, DBuf &k2, DBuf &v, DBuf &v2, u64 n, int begin_bit, int end_bit) {
  rsort_run<u32>(c, k, k2, &v, &v2, n, begin_bit, end_bit > 32 ? 32 : end_bit);
}
void rsort_pairs64(cmsbwt_ctx *c, DBuf &k, DBuf &k2, DBuf &v, DBuf &v2, u64 n, int begin_bit, int end_bit) {
  rsort_run<u64>(c, k, k2, &v, &v2, n, begin_bit, end_bit > 64 ? 64 : end_bit);
}
void rsort_keys64(cmsbwt_ctx *c, DBuf &k, DBuf &k2, u64 n, int begin_bit, int end_bit) {
  rsort_run<u64>(c, k, k2, nullptr, nullptr, n, begin_bit, end_bit > 64 ? 64 : end_bit);
}
