// Run-length records of a BWT that already lies in device memory, WITHOUT any allocation: a variant of
// the encoder of emit.cu (k_rle_count / k_rle_starts / k_rle_records behind rle_encode) whose scratch -- per
// 4 KB block of the input: number of run starts, position of the last run start, and their scans -- lives
// in a file-scope __device__ array of fixed size instead of buffers from the context's allocator.  For
// callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
// it).  No counterpart in the reference.
//
// The array is CMSBWT_RLE_STATIC_SCRATCH bytes (4 MB + 64) on every GPU that loads the library, until
// exit.  It bounds the input at CMSBWT_RLE_STATIC_MAX = 2^30 symbols per call.  Every call on a device
// shares it: all the entries a call reads are written by that call first (nothing has to start at
// zero), but two calls that overlap on one GPU (different streams, or different host threads) race on
// it -- callers serialise them.
//
// Unlike the encoder of emit.cu there is no array of run starts (its size is the number of runs, which
// only the data bounds): the thread that finds the start of run k knows where run k-1 began (the last
// run start before it: in its own 16 bytes, else from a max-scan over the block, else from the scan
// over the blocks) and writes the LENGTH of run k-1 into record k-1 and the SYMBOL of run k into
// record k, so every byte of every record is written exactly once.
#include "common.cuh"
#include <cub/cub.cuh>
#include <mutex>

#define RS_T 256                       // threads per block
#define RS_V 16                        // input bytes per thread (one 16-byte load)
#define RS_BLK (RS_T * RS_V)           // 4096 input bytes per block
#define RS_MAXB (CMSBWT_RLE_STATIC_MAX / RS_BLK)      // 262144 blocks

struct RleStaticScratch {
  unsigned long long off[RS_MAXB];     // run starts per block, then (in place) their exclusive prefix sum
  unsigned long long last[RS_MAXB];    // 1 + position of the block's last run start (0: none), then (in place)
                                       // the same for everything BEFORE the block (exclusive max-scan)
  unsigned long long total;            // number of runs
  unsigned long long last_all;         // 1 + position of the last run start of the whole input
  unsigned long long pad[6];
};
static_assert(sizeof(RleStaticScratch) == CMSBWT_RLE_STATIC_SCRATCH, "header states the scratch size");
static __device__ __align__(256) RleStaticScratch g_rle_scratch;

// bit j of the result: a run starts at byte i0 + j (i0 + j < N); v[] receives the bytes
__device__ __forceinline__ u32 rs_flags(const u8 *__restrict__ in, u64 N, u64 i0, bool aligned, u8 v[RS_V]) {
  if (i0 >= N) return 0;
  const u32 cnt = (u32)min((u64)RS_V, N - i0);
  if (aligned && cnt == RS_V) {
    const uint4 w = __ldg((const uint4 *)(in + i0));
    const u32 ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int j = 0; j < RS_V; j++) v[j] = (u8)(ww[j >> 2] >> (8 * (j & 3)));
  } else {
#pragma unroll
    for (int j = 0; j < RS_V; j++) v[j] = (u32)j < cnt ? in[i0 + j] : 0;
  }
  u32 m = 0;
  u8 prev = i0 ? in[i0 - 1] : 0;
#pragma unroll
  for (int j = 0; j < RS_V; j++) {
    if ((u32)j < cnt && ((i0 + j == 0) || v[j] != prev)) m |= 1u << j;
    prev = v[j];
  }
  return m;
}

__global__ void __launch_bounds__(RS_T) k_rls_count(const u8 *__restrict__ in, u64 N, bool aligned, RleStaticScratch *S) {
  typedef cub::BlockReduce<u32, RS_T> BR;
  __shared__ typename BR::TempStorage t1;
  __shared__ typename BR::TempStorage t2;
  const u64 i0 = (u64)blockIdx.x * RS_BLK + (u64)threadIdx.x * RS_V;
  u8 v[RS_V];
  const u32 m = rs_flags(in, N, i0, aligned, v);
  const u32 c = __popc(m);
  const u32 l = m ? threadIdx.x * RS_V + (31 - __clz(m)) + 1 : 0;      // 1 + offset inside the block
  const u32 cs = BR(t1).Sum(c);
  const u32 lm = BR(t2).Reduce(l, cub::Max());
  if (threadIdx.x == 0) {
    S->off[blockIdx.x] = cs;
    S->last[blockIdx.x] = lm ? (u64)blockIdx.x * RS_BLK + lm : 0;
  }
}

// one CTA walks the blocks' entries 1024 at a time (at most 256 rounds)
__global__ void __launch_bounds__(1024) k_rls_scan(u32 nblk, RleStaticScratch *S, unsigned long long *n_out) {
  typedef cub::BlockScan<unsigned long long, 1024> BS;
  __shared__ typename BS::TempStorage t1;
  __shared__ typename BS::TempStorage t2;
  unsigned long long carry = 0, lcarry = 0;
  for (u32 b0 = 0; b0 < nblk; b0 += 1024) {
    const u32 b = b0 + threadIdx.x;
    const unsigned long long c = b < nblk ? S->off[b] : 0, l = b < nblk ? S->last[b] : 0;
    unsigned long long ex, agg, lex, lagg;
    BS(t1).ExclusiveSum(c, ex, agg);
    BS(t2).ExclusiveScan(l, lex, 0ull, cub::Max(), lagg);
    if (b < nblk) { S->off[b] = carry + ex; S->last[b] = max(lcarry, lex); }
    carry += agg;
    lcarry = max(lcarry, lagg);
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    S->total = carry;
    S->last_all = lcarry;
    if (n_out) *n_out = carry;
  }
}

__device__ __forceinline__ void rs_put_len(u8 *rec, u64 k, u64 run) {
  u8 *d = rec + 9 * k;
#pragma unroll
  for (int b = 0; b < 8; b++) d[b] = (u8)(run >> (8 * b));
}

__global__ void __launch_bounds__(RS_T) k_rls_write(const u8 *__restrict__ in, u64 N, bool aligned, const RleStaticScratch *S,
                                                    u8 *rec, u64 cap) {
  typedef cub::BlockScan<u32, RS_T> BS;
  __shared__ typename BS::TempStorage t1;
  __shared__ typename BS::TempStorage t2;
  const u64 blk0 = (u64)blockIdx.x * RS_BLK;
  const u64 i0 = blk0 + (u64)threadIdx.x * RS_V;
  u8 v[RS_V];
  const u32 m = rs_flags(in, N, i0, aligned, v);
  const u32 c = __popc(m);
  const u32 l = m ? threadIdx.x * RS_V + (31 - __clz(m)) + 1 : 0;
  u32 ex, lex;
  BS(t1).ExclusiveSum(c, ex);
  BS(t2).ExclusiveScan(l, lex, 0u, cub::Max());
  u64 k = S->off[blockIdx.x] + ex;                                 // rank of this thread's first run start
  u64 prev = lex ? blk0 + lex : S->last[blockIdx.x];               // 1 + start of the run before it (0: there is none)
#pragma unroll
  for (int j = 0; j < RS_V; j++) {
    if (!(m >> j & 1)) continue;
    const u64 i = i0 + j;
    if (k > 0 && k - 1 < cap) rs_put_len(rec, k - 1, i - (prev - 1));
    if (k < cap) rec[9 * k + 8] = v[j];
    prev = i + 1;
    k++;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {                       // the last run ends at N
    const u64 R = S->total;
    if (R > 0 && R - 1 < cap) rs_put_len(rec, R - 1, N - (S->last_all - 1));
  }
}

// device address of the scratch array, looked up once per device with that device current (in host
// code the array's name is only a stand-in for the symbol, not a device address)
static std::mutex g_rls_mu;
static RleStaticScratch *g_rls_ptr[64] = {};

static RleStaticScratch *rls_scratch(int dev) {
  if (dev < 0 || dev >= 64) throw CmsErr{CMSBWT_E_ARG, "device ordinal out of range"};
  std::lock_guard<std::mutex> lk(g_rls_mu);
  if (!g_rls_ptr[dev]) {
    RleStaticScratch *p = nullptr;
    CK(cudaGetSymbolAddress((void **)&p, g_rle_scratch));
    if (!p) throw CmsErr{CMSBWT_E_CUDA, "cudaGetSymbolAddress returned a null scratch address"};
    g_rls_ptr[dev] = p;
  }
  return g_rls_ptr[dev];
}

extern "C" int cmsbwt_rle_encode_static(cmsbwt_ctx *c, const void *d_bwt, uint64_t N, void *d_records,
                                        uint64_t cap_records, uint64_t *d_n_records, uint64_t *n_records,
                                        void *stream) {
  if (!c) return CMSBWT_E_ARG;
  if ((N && !d_bwt) || (cap_records && !d_records)) { c->err = "rle_encode_static: null buffer"; return CMSBWT_E_ARG; }
  if (N > CMSBWT_RLE_STATIC_MAX) {
    c->err = "rle_encode_static: N exceeds CMSBWT_RLE_STATIC_MAX (2^30 symbols per call: the static scratch array "
             "holds one entry per 4 KB of input); encode in pieces or use cmsbwt_emit(CMSBWT_MODE_RLE)";
    return CMSBWT_E_LIMIT;
  }
  try {
    CK(cudaSetDevice(c->dev));
    RleStaticScratch *S = rls_scratch(c->dev);
    cudaStream_t st = stream ? (cudaStream_t)stream : c->st;
    const u32 nblk = (u32)((N + RS_BLK - 1) / RS_BLK);
    const bool aligned = ((uintptr_t)d_bwt & 15) == 0;
    if (nblk) k_rls_count<<<nblk, RS_T, 0, st>>>((const u8 *)d_bwt, N, aligned, S);
    k_rls_scan<<<1, 1024, 0, st>>>(nblk, S, (unsigned long long *)d_n_records);
    if (nblk) k_rls_write<<<nblk, RS_T, 0, st>>>((const u8 *)d_bwt, N, aligned, S, (u8 *)d_records, cap_records);
    c->launches += nblk ? 3 : 1;
    CK(cudaGetLastError());
    if (n_records) {                       // optional: the count on the host (one 8-byte copy and a wait, no allocation)
      unsigned long long R = 0;
      CK(cudaMemcpyAsync(&R, &S->total, 8, cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
      *n_records = R;
      if (R > cap_records) {
        c->err = "rle_encode_static: the input has more runs than cap_records (only the first cap_records were written)";
        return CMSBWT_E_LIMIT;
      }
    }
  } catch (const CudaErr &e) { c->err = std::string(e.what) + ": " + cudaGetErrorString(e.e); cudaGetLastError(); return CMSBWT_E_CUDA; }
  catch (const CmsErr &e) { c->err = e.msg; return e.code; }
  return CMSBWT_OK;
}
