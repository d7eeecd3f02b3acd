// Stage 3: bucket inversion to SA order (CMS-BWT-functions.cpp:1618-1628) and the writers
// (:1631-1694 plain, :1695-1777 RLE).  Output = for SA rank r: cov[SA[r]] copies of refBWT[r],
// with the head symbols stored at offset base[ISA[pos]] + within.
#include "common.cuh"
#include <cub/cub.cuh>

__global__ void k_cnt_sa(const i32 *cov, const i32 *sa, u32 n, u32 *cnt) {
  u32 r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n) cnt[r] = (u32)cov[sa[r]];
}

#define FILL_THREADS 256
#define FILL_VEC_PER_THREAD 4
#define FILL_BYTES (FILL_THREADS * FILL_VEC_PER_THREAD * 16)
#define FILL_SMEM_RUNS 2048

// largest r in [lo, hi) with base[r] <= o   (base is non-decreasing, base[lo] <= o)
__device__ __forceinline__ u32 run_of(const u64 *base, u32 lo, u32 hi, u64 o) {
  while (hi - lo > 1) {
    u32 mid = lo + ((hi - lo) >> 1);
    if (base[mid] <= o) lo = mid; else hi = mid;
  }
  return lo;
}

// Run-length expansion: every thread writes 16-byte vectors of the output.
__global__ void __launch_bounds__(FILL_THREADS) k_fill(const u64 *base, const u8 *rbwt, u32 n, u64 N,
                                                       u8 *out) {
  __shared__ u64 sb[FILL_SMEM_RUNS + 1];
  __shared__ u32 s_r0, s_r1;
  const u64 o0 = (u64)blockIdx.x * FILL_BYTES;
  const u64 o1 = min(N, o0 + FILL_BYTES);
  if (threadIdx.x == 0) {
    s_r0 = run_of(base, 0, n + 1, o0);
    s_r1 = run_of(base, 0, n + 1, o1 - 1);
  }
  __syncthreads();
  const u32 r0 = s_r0, r1 = s_r1;
  const u32 nr = r1 - r0 + 1;                 // runs r0..r1 intersect this block
  const bool use_s = nr <= FILL_SMEM_RUNS;
  if (use_s)
    for (u32 k = threadIdx.x; k <= nr; k += FILL_THREADS) sb[k] = base[r0 + k];
  __syncthreads();
  const u64 *bb = use_s ? sb : base + r0;
  for (int v = 0; v < FILL_VEC_PER_THREAD; v++) {
    u64 o = o0 + ((u64)v * FILL_THREADS + threadIdx.x) * 16;
    if (o >= o1) break;
    u32 k = run_of(bb, 0, nr, o);             // bb[k] <= o < bb[k+1]
    u64 nextb = bb[k + 1];
    u32 c = rbwt[r0 + k];
    if (nextb >= o + 16 && o + 16 <= N) {
      u32 w = c * 0x01010101u;
      *(uint4 *)(out + o) = make_uint4(w, w, w, w);
    } else {
      u64 lim = min(N, o + 16);
      for (u64 b = o; b < lim; b++) {
        while (bb[k + 1] <= b) k++;
        out[b] = rbwt[r0 + k];
      }
    }
  }
}

__global__ void k_scatter_heads(const u32 *pos, const u16 *cs, const u32 *within, const u32 *isa,
                                const u64 *base, u64 H, u8 *out, u64 *off_out) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= H) return;
  u64 off = base[isa[pos[i]]] + within[i];
  out[off] = (u8)(cs[i] & 0xFF);
  off_out[i] = off;
}

// ---- RLE (9-byte records: LE u64 run, u8 symbol; maximal runs) -------------------------------
#define RLE_THREADS 256
#define RLE_PER_THREAD 64
#define RLE_BLOCK (RLE_THREADS * RLE_PER_THREAD)
__global__ void __launch_bounds__(RLE_THREADS) k_rle_count(const u8 *out, u64 N, u32 *bcnt) {
  typedef cub::BlockReduce<u32, RLE_THREADS> BR;
  __shared__ typename BR::TempStorage ts;
  u64 o0 = (u64)blockIdx.x * RLE_BLOCK + (u64)threadIdx.x * RLE_PER_THREAD;
  u32 c = 0;
  if (o0 < N) {
    u8 prev = o0 > 0 ? out[o0 - 1] : 0;
    u64 lim = min(N, o0 + RLE_PER_THREAD);
    for (u64 o = o0; o < lim; o++) {
      u8 v = out[o];
      c += (o == 0) || v != prev;
      prev = v;
    }
  }
  u32 s = BR(ts).Sum(c);
  if (threadIdx.x == 0) bcnt[blockIdx.x] = s;
}
__global__ void __launch_bounds__(RLE_THREADS) k_rle_starts(const u8 *out, u64 N, const u64 *boff,
                                                            u64 *starts) {
  typedef cub::BlockScan<u32, RLE_THREADS> BS;
  __shared__ typename BS::TempStorage ts;
  u64 o0 = (u64)blockIdx.x * RLE_BLOCK + (u64)threadIdx.x * RLE_PER_THREAD;
  u32 c = 0;
  u64 lim = min(N, o0 + RLE_PER_THREAD);
  if (o0 < N) {
    u8 prev = o0 > 0 ? out[o0 - 1] : 0;
    for (u64 o = o0; o < lim; o++) { u8 v = out[o]; c += (o == 0) || v != prev; prev = v; }
  }
  u32 ex;
  BS(ts).ExclusiveSum(c, ex);
  if (o0 < N) {
    u64 w = boff[blockIdx.x] + ex;
    u8 prev = o0 > 0 ? out[o0 - 1] : 0;
    for (u64 o = o0; o < lim; o++) {
      u8 v = out[o];
      if (o == 0 || v != prev) starts[w++] = o;
      prev = v;
    }
  }
}
__global__ void k_rle_records(const u8 *out, const u64 *starts, u64 R, u64 N, u8 *rec) {
  u64 k = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= R) return;
  u64 s = starts[k];
  u64 run = (k + 1 < R ? starts[k + 1] : N) - s;
  u8 *d = rec + 9 * k;
  for (int b = 0; b < 8; b++) d[b] = (u8)(run >> (8 * b));
  d[8] = out[s];
}

void emit_stage(cmsbwt_ctx *c, int mode) {
  cudaStream_t st = c->st;
  const u32 n = c->ix.n;
  const u64 N = c->N, H = c->H;
  const int B = 256;
  DBuf cnt, base, tmp;
  cnt.alloc(((size_t)n + 1) * 4, st);
  base.alloc(((size_t)n + 2) * 8, st);
  size_t tb = 0, tb2 = 0;
  // coverage = inclusive scan of the deltas (in place)
  cub::DeviceScan::InclusiveSum(nullptr, tb, c->d_cov.as<i32>(), c->d_cov.as<i32>(), (int)n, st);
  EXSUM64(nullptr, tb2, cnt.as<u32>(), base.as<u64>(), (int)(n + 1), st);
  if (tb2 > tb) tb = tb2;
  tmp.alloc(tb, st);
  size_t t = tb;
  CK(cub::DeviceScan::InclusiveSum(tmp.p, t, c->d_cov.as<i32>(), c->d_cov.as<i32>(), (int)n, st));
  CK(cudaMemsetAsync(cnt.p, 0, ((size_t)n + 1) * 4, st));
  k_cnt_sa<<<(n + B - 1) / B, B, 0, st>>>(c->d_cov.as<i32>(), c->ix.sa, n, cnt.as<u32>());
  t = tb;
  CK(EXSUM64(tmp.p, t, cnt.as<u32>(), base.as<u64>(), (int)(n + 1), st));
  c->launches += 3;
  u64 total = 0;
  CK(cudaMemcpyAsync(&total, base.as<u64>() + n, 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (total != N) {
    char buf[128];
    snprintf(buf, sizeof buf, "coverage total %llu != N %llu", (unsigned long long)total,
             (unsigned long long)N);
    throw CmsErr{CMSBWT_E_INTERNAL, buf};
  }
  c->d_out.alloc(N + CMS_PAD, st);
  c->h_off.alloc(H * 8, st);
  u8 *out = c->d_out.as<u8>();
  k_fill<<<(u32)((N + FILL_BYTES - 1) / FILL_BYTES), FILL_THREADS, 0, st>>>(base.as<u64>(), c->ix.rbwt,
                                                                          n, N, out);
  k_scatter_heads<<<(u32)((H + B - 1) / B), B, 0, st>>>(c->h_pos.as<u32>(), c->h_cs.as<u16>(),
                                                       c->h_within.as<u32>(), c->ix.isa,
                                                       base.as<u64>(), H, out, c->h_off.as<u64>());
  c->launches += 2;
  CK(cudaGetLastError());
  c->out_mode = mode;
  c->out_bytes = N;
  c->stats.n_runs = 0;
  if (mode == CMSBWT_MODE_RLE) {
    u64 nb = (N + RLE_BLOCK - 1) / RLE_BLOCK;
    DBuf bcnt, boff, starts;
    bcnt.alloc((nb + 1) * 4, st); boff.alloc((nb + 1) * 8, st);
    CK(cudaMemsetAsync(bcnt.p, 0, (nb + 1) * 4, st));
    k_rle_count<<<(u32)nb, RLE_THREADS, 0, st>>>(out, N, bcnt.as<u32>());
    size_t tb3 = 0;
    EXSUM64(nullptr, tb3, bcnt.as<u32>(), boff.as<u64>(), (int)(nb + 1), st);
    DBuf tmp3; tmp3.alloc(tb3, st);
    CK(EXSUM64(tmp3.p, tb3, bcnt.as<u32>(), boff.as<u64>(), (int)(nb + 1), st));
    u64 R = 0;
    CK(cudaMemcpyAsync(&R, boff.as<u64>() + nb, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    starts.alloc((R + 1) * 8, st);
    c->d_rle.alloc(R * 9 + CMS_PAD, st);
    k_rle_starts<<<(u32)nb, RLE_THREADS, 0, st>>>(out, N, boff.as<u64>(), starts.as<u64>());
    k_rle_records<<<(u32)((R + B - 1) / B), B, 0, st>>>(out, starts.as<u64>(), R, N, c->d_rle.as<u8>());
    c->launches += 4;
    CK(cudaGetLastError());
    c->out_bytes = R * 9;
    c->stats.n_runs = R;
  }
  CK(cudaStreamSynchronize(st));
}
