// Consumer of the path's output, fed from device memory: an FM-index over the BWT the
// emit stage left in HBM -- symbol counts C[], sampled occurrence counts (one checkpoint per FM_BLOCK
// symbols, exclusive prefix per symbol) -- and batched backward search (count queries) on it.  The
// reference stops at the .bwt / .rl_bwt files (CMS-BWT-functions.cpp:1631-1777) and leaves indexing
// to downstream tools; here the BWT never leaves the device between construction and first use.
// The BWT is the multi-string BWT with every terminator written as 0x02, so
// backward search counts the occurrences of any pattern without separators in the whole collection.
#include "common.cuh"
#include <cub/cub.cuh>
#include <string.h>

#define FM_BLOCK 256u            // symbols per checkpoint (one CTA pass of 256 threads, 16 uint4 per rank query)
#define FM_MAXSIG 16             // distinct symbols supported (DNA + separator + IUPAC leftovers)

__global__ void __launch_bounds__(256) k_fm_hist(const u8 *bwt, u64 N, unsigned long long *hist) {
  __shared__ u32 s_h[256];
  s_h[threadIdx.x] = 0;
  __syncthreads();
  const u64 per = (N + gridDim.x - 1) / gridDim.x;
  const u64 a = (u64)blockIdx.x * per, b = min(N, a + per);
  for (u64 i = a + threadIdx.x; i < b; i += 256) atomicAdd(&s_h[bwt[i]], 1u);
  __syncthreads();
  if (s_h[threadIdx.x]) atomicAdd(hist + threadIdx.x, (unsigned long long)s_h[threadIdx.x]);
}

// per checkpoint block: occurrences of every (dense) symbol inside the block
__global__ void __launch_bounds__(256) k_fm_block_counts(const u8 *bwt, u64 N, const u8 *dense, u32 sig, u64 nblk, u32 *cnt) {
  __shared__ u32 s_c[FM_MAXSIG];
  const u64 blk = blockIdx.x;
  if (threadIdx.x < FM_MAXSIG) s_c[threadIdx.x] = 0;
  __syncthreads();
  const u64 i = blk * FM_BLOCK + threadIdx.x;
  if (i < N) atomicAdd(&s_c[dense[bwt[i]]], 1u);
  __syncthreads();
  if (threadIdx.x < sig) cnt[(u64)threadIdx.x * nblk + blk] = s_c[threadIdx.x];      // symbol-major rows
}

// occurrences of dense symbol d in bwt[0, i)
__device__ __forceinline__ u64 fm_rank(const u8 *bwt, const u64 *occ, u64 nblk, u8 sym, u32 d, u64 i) {
  const u64 blk = i / FM_BLOCK;
  u64 r = occ[(u64)d * nblk + blk];
  const u64 a = blk * FM_BLOCK;
  const u32 rem = (u32)(i - a);
  const uint4 *q = (const uint4 *)(bwt + a);             // checkpoint blocks are 256-byte aligned
  const u32 pat = sym * 0x01010101u;
  u32 c = 0;
  for (u32 v = 0; v * 16 < rem; v++) {
    uint4 w = __ldg(q + v);
    u32 words[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const u32 left = rem - v * 16 - k * 4;             // bytes of this word inside [a, i)
      if ((int)left <= 0) break;
      u32 x = words[k] ^ pat;                            // zero bytes where equal
      u32 eq = ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & 0x80808080u;
      if (left < 4) eq &= (1u << (8 * left)) - 1u;
      c += __popc(eq);
    }
  }
  return r + c;
}

// one thread per pattern: backward search (LF on both interval ends)
__global__ void k_fm_count(const u8 *bwt, u64 N, const u64 *occ, const u64 *C, const u8 *dense, u64 nblk,
                           const u8 *pat, const u64 *poff, u32 npat, u64 *out) {
  const u32 k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= npat) return;
  u64 lo = 0, hi = N;
  for (u64 j = poff[k + 1]; j > poff[k] && lo < hi; j--) {
    const u8 c = pat[j - 1];
    const u32 d = dense[c];
    if (d == 0xFF) { lo = hi = 0; break; }               // symbol does not occur in the collection
    lo = C[d] + fm_rank(bwt, occ, nblk, c, d, lo);
    hi = C[d] + fm_rank(bwt, occ, nblk, c, d, hi);
  }
  out[k] = hi > lo ? hi - lo : 0;
}

struct FmIndex {
  DBuf occ, C, dense;
  u32 sig = 0;
  u64 N = 0, nblk = 0;
  const u8 *bwt = nullptr;        // the emit stage's result buffer (valid until the next ms_run)
};
void fm_release(cmsbwt_ctx *c) {
  if (c->fm) { delete (FmIndex *)c->fm; c->fm = nullptr; }
}

static int fm_fail(cmsbwt_ctx *c, int code, const char *msg) { c->err = msg; return code; }

extern "C" {

int cmsbwt_fm_build(cmsbwt_ctx *c, uint32_t *n_symbols) {
  if (!c) return CMSBWT_E_ARG;
  if (c->out_mode != CMSBWT_MODE_PLAIN || !c->d_out.p || c->out_bytes == 0)
    return fm_fail(c, CMSBWT_E_ARG, "cmsbwt_emit in plain mode first (the FM-index reads the BWT from device memory)");
  try {
    CK(cudaSetDevice(c->dev));
    current_cache() = &c->cache;
    cudaStream_t st = c->st;
    FmIndex *fm = (FmIndex *)c->fm;
    if (!fm) { fm = new FmIndex(); c->fm = fm; }
    const u64 N = c->out_bytes;
    const u8 *bwt = c->d_out.as<u8>();
    DBuf hist;
    hist.alloc(256 * 8, st);
    CK(cudaMemsetAsync(hist.p, 0, 256 * 8, st));
    k_fm_hist<<<148 * 8, 256, 0, st>>>(bwt, N, hist.as<unsigned long long>());
    unsigned long long hh[256];
    CK(cudaMemcpyAsync(hh, hist.p, sizeof hh, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    u8 dense[256];
    u64 Ch[FM_MAXSIG + 1];
    memset(dense, 0xFF, sizeof dense);
    u32 sig = 0;
    u64 run = 0;
    for (int v = 0; v < 256; v++)
      if (hh[v]) {
        if (sig == FM_MAXSIG) throw CmsErr{CMSBWT_E_LIMIT, "more than 16 distinct symbols in the BWT"};
        dense[v] = (u8)sig; Ch[sig] = run; run += hh[v]; sig++;
      }
    Ch[sig] = run;
    const u64 nblk = (N + FM_BLOCK - 1) / FM_BLOCK + 1;      // one extra block: rank(N) when N is a multiple of the block
    DBuf cnt;
    cnt.alloc(nblk * sig * 4, st);
    fm->occ.alloc(nblk * sig * 8, st); fm->C.alloc((FM_MAXSIG + 1) * 8, st); fm->dense.alloc(256, st);
    CK(cudaMemsetAsync(cnt.p, 0, nblk * sig * 4, st));
    CK(cudaMemcpyAsync(fm->dense.p, dense, 256, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(fm->C.p, Ch, (sig + 1) * 8, cudaMemcpyHostToDevice, st));
    k_fm_block_counts<<<(u32)(nblk - 1), 256, 0, st>>>(bwt, N, fm->dense.as<u8>(), sig, nblk, cnt.as<u32>());
    size_t tb = 0;
    EXSUM64(nullptr, tb, cnt.as<u32>(), fm->occ.as<u64>(), (int)nblk, st);
    DBuf tmp;
    tmp.alloc(tb, st);
    for (u32 d = 0; d < sig; d++) {       // exclusive prefix of every symbol's row of block counts
      size_t t = tb;
      CK(EXSUM64(tmp.p, t, cnt.as<u32>() + (size_t)d * nblk, fm->occ.as<u64>() + (size_t)d * nblk, (int)nblk, st));
    }
    c->launches += 2 + 2 * sig;
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(st));
    fm->sig = sig; fm->N = N; fm->nblk = nblk; fm->bwt = bwt;
    if (n_symbols) *n_symbols = sig;
  } catch (const CudaErr &e) { c->err = std::string(e.what) + ": " + cudaGetErrorString(e.e); cudaGetLastError(); return CMSBWT_E_CUDA; }
  catch (const CmsErr &e) { c->err = e.msg; return e.code; }
  return CMSBWT_OK;
}

int cmsbwt_fm_count(cmsbwt_ctx *c, const uint8_t *patterns, const uint64_t *offsets, uint32_t npat, uint64_t *counts) {
  if (!c || !patterns || !offsets || !counts) return CMSBWT_E_ARG;
  FmIndex *fm = (FmIndex *)c->fm;
  if (!fm || !fm->bwt || fm->bwt != c->d_out.as<u8>()) return fm_fail(c, CMSBWT_E_ARG, "cmsbwt_fm_build first");
  if (npat == 0) return CMSBWT_OK;
  try {
    CK(cudaSetDevice(c->dev));
    current_cache() = &c->cache;
    cudaStream_t st = c->st;
    const u64 total = offsets[npat];
    DBuf dp, d_off, dc;
    dp.alloc(total + 16, st); d_off.alloc(((size_t)npat + 1) * 8, st); dc.alloc((size_t)npat * 8, st);
    CK(cudaMemcpyAsync(dp.p, patterns, total, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_off.p, offsets, ((size_t)npat + 1) * 8, cudaMemcpyHostToDevice, st));
    k_fm_count<<<(npat + 127) / 128, 128, 0, st>>>(fm->bwt, fm->N, fm->occ.as<u64>(), fm->C.as<u64>(), fm->dense.as<u8>(),
                                                  fm->nblk, dp.as<u8>(), d_off.as<u64>(), npat, dc.as<u64>());
    LAUNCH(c);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(counts, dc.p, (size_t)npat * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
  } catch (const CudaErr &e) { c->err = std::string(e.what) + ": " + cudaGetErrorString(e.e); cudaGetLastError(); return CMSBWT_E_CUDA; }
  catch (const CmsErr &e) { c->err = e.msg; return e.code; }
  return CMSBWT_OK;
}

int cmsbwt_fm_free(cmsbwt_ctx *c) {
  if (!c) return CMSBWT_E_ARG;
  cudaSetDevice(c->dev);
  current_cache() = &c->cache;
  fm_release(c);
  return CMSBWT_OK;
}

}  // extern "C"
