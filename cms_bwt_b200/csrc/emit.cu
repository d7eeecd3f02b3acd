// Stage 3: bucket inversion to SA order (CMS-BWT-functions.cpp:1618-1628) and the writers
// (:1631-1694 plain, :1695-1777 RLE).  Output = for SA rank r: cov[SA[r]] copies of refBWT[r],
// with the head symbols stored at offset base[ISA[pos]] + within.
#include "common.cuh"
#include <cub/cub.cuh>
#include <stdlib.h>
#include <algorithm>

__global__ void k_cnt_sa(const i32 *cov, const i32 *sa, u32 n, u32 *cnt) {
  u32 r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n) cnt[r] = (u32)cov[sa[r]];
}

#define FILL_THREADS 256
#define FILL_VEC_PER_THREAD 4
#define FILL_VECS (FILL_THREADS * FILL_VEC_PER_THREAD)
#define FILL_BYTES (FILL_VECS * 16)
#define FILL_SMEM_RUNS 2048

// largest r in [lo, hi) with base[r] <= o   (base is non-decreasing, base[lo] <= o)
__device__ __forceinline__ u32 run_of(const u64 *base, u32 lo, u32 hi, u64 o) {
  while (hi - lo > 1) {
    u32 mid = lo + ((hi - lo) >> 1);
    if (base[mid] <= o) lo = mid; else hi = mid;
  }
  return lo;
}

// merge-path style partition: the run that holds the first byte of every output block
// (blk0: first output block of this device; bstart is indexed from it)
__global__ void k_fill_part(const u64 *base, u32 n, u64 N, u64 nblocks, u64 blk0, u32 *bstart) {
  u64 b = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (b > nblocks) return;
  u64 o = (b + blk0) * FILL_BYTES;
  bstart[b] = o < N ? run_of(base, 0, n, o) : n - 1;
}

// Run-length expansion (:1640-1688 without the head symbols): cov[SA[r]] copies of refBWT[r].
// A block owns FILL_BYTES of output; the run covering every 16-byte vector comes from a max-scan
// over "last run starting in this vector" marks, vectors inside one run are splatted.
__global__ void __launch_bounds__(FILL_THREADS) k_fill(const u64 *base, const u8 *rbwt, const u32 *bstart,
                                                       u32 n, u64 N, u64 blk0, u8 *out) {
  __shared__ u64 sb[FILL_SMEM_RUNS + 2];
  __shared__ u8 ssym[FILL_SMEM_RUNS + 2];
  __shared__ u32 s_m[FILL_VECS];
  typedef cub::BlockScan<u32, FILL_THREADS> BS;
  __shared__ typename BS::TempStorage ts;
  const u32 tid = threadIdx.x;
  const u64 o0 = ((u64)blockIdx.x + blk0) * FILL_BYTES;
  const u64 o1 = min(N, o0 + FILL_BYTES);
  out -= blk0 * FILL_BYTES;                   // out points at this device's first byte
  const u32 r0 = bstart[blockIdx.x], r1 = bstart[blockIdx.x + 1];
  const u32 nr = r1 - r0 + 1;                 // runs r0..r1 intersect this block
  const bool use_s = nr <= FILL_SMEM_RUNS;
  if (use_s) {
    for (u32 k = tid; k <= nr; k += FILL_THREADS) { sb[k] = base[r0 + k]; ssym[k] = k < nr ? rbwt[r0 + k] : 0; }
    if (tid == 0) sb[nr + 1] = ~0ull;          // the last block runs one entry past base[n] = N
    for (u32 v = tid; v < FILL_VECS; v += FILL_THREADS) s_m[v] = 0;
    __syncthreads();
    for (u32 k = 1 + tid; k < nr; k += FILL_THREADS) {
      u64 s = sb[k];
      if (s >= o0 && s < o1) atomicMax(&s_m[(u32)(s - o0) >> 4], k);
    }
    __syncthreads();
    u32 it[FILL_VEC_PER_THREAD];
#pragma unroll
    for (int v = 0; v < FILL_VEC_PER_THREAD; v++) it[v] = s_m[tid * FILL_VEC_PER_THREAD + v];
    BS(ts).InclusiveScan(it, it, cub::Max());
    __syncthreads();
#pragma unroll
    for (int v = 0; v < FILL_VEC_PER_THREAD; v++) s_m[tid * FILL_VEC_PER_THREAD + v] = it[v];
    __syncthreads();
  }
#pragma unroll
  for (int v = 0; v < FILL_VEC_PER_THREAD; v++) {
    const u32 vi = (u32)v * FILL_THREADS + tid;
    const u64 o = o0 + (u64)vi * 16;
    if (o >= o1) break;
    u32 k;
    u64 lo, hi;
    if (use_s) {
      k = vi ? s_m[vi - 1] : 0;
      while (sb[k + 1] <= o) k++;              // runs that start exactly here, empty runs
      lo = hi = (u64)ssym[k] * 0x0101010101010101ull;
      u64 nb = sb[k + 1];
      while (nb < o + 16) {                    // a run starts inside this vector: overwrite its bytes
        k++;
        const u32 d = (u32)(nb - o);
        const u64 sp = (u64)ssym[k] * 0x0101010101010101ull;
        const u64 mlo = d < 8 ? ~0ull << (8 * d) : 0ull;
        const u64 mhi = d < 8 ? ~0ull : ~0ull << (8 * (d - 8));
        lo = (lo & ~mlo) | (sp & mlo);
        hi = (hi & ~mhi) | (sp & mhi);
        nb = sb[k + 1];
      }
    } else {
      const u64 *bb = base + r0;
      k = run_of(bb, 0, nr, o);
      lo = hi = (u64)rbwt[r0 + k] * 0x0101010101010101ull;
      u64 nb = bb[k + 1];
      while (nb < o + 16) {
        k++;
        const u32 d = (u32)(nb - o);
        const u64 sp = (u64)rbwt[r0 + k] * 0x0101010101010101ull;
        const u64 mlo = d < 8 ? ~0ull << (8 * d) : 0ull;
        const u64 mhi = d < 8 ? ~0ull : ~0ull << (8 * (d - 8));
        lo = (lo & ~mlo) | (sp & mlo);
        hi = (hi & ~mhi) | (sp & mhi);
        nb = bb[k + 1];
      }
    }
    // out has CMS_PAD bytes of slack, so the last vector may be written whole
    *(uint4 *)(out + o) = make_uint4((u32)lo, (u32)(lo >> 32), (u32)hi, (u32)(hi >> 32));
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


// maximal-run records (9 bytes: LE u64 run, u8 symbol) of out[0, N) -> d_rle (:1695-1777)
void rle_encode(cmsbwt_ctx *c, const u8 *out, u64 N) {
  cudaStream_t st = c->st;
  const int B = 256;
  u64 nb = (N + RLE_BLOCK - 1) / RLE_BLOCK;
  DBuf bcnt, boff, starts;
  bcnt.alloc((nb + 1) * 4, st); boff.alloc((nb + 1) * 8, st);
  CK(cudaMemsetAsync(bcnt.p, 0, (nb + 1) * 4, st));
  if (nb) k_rle_count<<<(u32)nb, RLE_THREADS, 0, st>>>(out, N, bcnt.as<u32>());
  size_t tb3 = 0;
  EXSUM64(nullptr, tb3, bcnt.as<u32>(), boff.as<u64>(), (int)(nb + 1), st);
  DBuf tmp3; tmp3.alloc(tb3, st);
  CK(EXSUM64(tmp3.p, tb3, bcnt.as<u32>(), boff.as<u64>(), (int)(nb + 1), st));
  u64 R = 0;
  CK(cudaMemcpyAsync(&R, boff.as<u64>() + nb, 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  starts.alloc((R + 1) * 8, st);
  c->d_rle.alloc(R * 9 + CMS_PAD, st);
  if (nb) k_rle_starts<<<(u32)nb, RLE_THREADS, 0, st>>>(out, N, boff.as<u64>(), starts.as<u64>());
  if (R) k_rle_records<<<(u32)((R + B - 1) / B), B, 0, st>>>(out, starts.as<u64>(), R, N, c->d_rle.as<u8>());
  c->launches += 4;
  CK(cudaGetLastError());
  c->out_mode = CMSBWT_MODE_RLE;
  c->out_bytes = R * 9;
  c->rle_off = 0;
  c->stats.n_runs = R;
  CK(cudaStreamSynchronize(st));
}

// ================================================================================================
// Pull emit (single GPU, tile path).  The output is walked in "virtual rank" order: v < nd is the
// terminator of document v (the bucket of SA rank 0 holds exactly one suffix per document, in document
// order: the writers' "first D-1 head chars", :1640-1643), v >= nd is SA rank r = v - nd + 1.  Per
// virtual rank: cnt[v] = bucket size (:1618-1628 bucket inversion: k_tile_rank stores every position's
// record at its virtual rank, the scatter overlapping that kernel's arithmetic), vsym[v] = reference
// BWT symbol, vrec[v] = the bucket's record: its heads, inside the record or as a list in hw[] (common.cuh).  The fill
// kernel expands the runs of its 16 KB output block in SHARED memory, pulls the head symbols of the
// buckets it intersects into it and stores the finished block with one bulk copy (TMA,
// cp.async.bulk.global.shared::cta) -- every output sector is written exactly once and nothing is
// scattered into the cold output (round 1: one sector read-modify-write per head symbol).
// ndl = terminators in front of this device's SA ranks (multi-GPU: all of them on device 0, none elsewhere),
// r0 = SA ranks before this device's range: entry i >= ndl is SA rank 1 + r0 + (i - ndl)
__global__ void k_vrank(uint4 *__restrict__ vrec, const u8 *__restrict__ rbwt, const u8 *__restrict__ termsym,
                        u32 ndl, u32 r0, u32 nv, u32 *__restrict__ cnt, u8 *__restrict__ vsym) {
  const u32 v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nv) return;
  if (v < ndl) { cnt[v] = 1; vrec[v] = make_uint4(1u, 0u, 0u, 0u); vsym[v] = termsym[v]; return; }
  cnt[v] = vrec[v].x;                            // bucket records (common.cuh), written by k_tile_rank
  vsym[v] = rbwt[r0 + (v - ndl) + 1];
}

#define PULL_SERIAL 24          // heads of one bucket a single thread pulls; longer lists go to a warp
#define PULL_QCAP 128

__device__ __forceinline__ void bulk_store_block(u8 *gdst, const u8 *s_src, u32 bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               :: "l"(gdst), "r"((u32)__cvta_generic_to_shared(s_src)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");    // smem may be reused / the CTA may exit
}

template <int STORE>      // 1 = bulk copy (TMA), 0 = 16-byte stores from shared memory
__global__ void __launch_bounds__(FILL_THREADS) k_fill_pull(const u64 *__restrict__ base, const u8 *__restrict__ vsym,
                                                            const uint4 *__restrict__ vrec, const u64 *__restrict__ hw,
                                                            const u32 *__restrict__ bstart, u64 N, u8 *out) {
  // run starts RELATIVE to the block's first byte (a block is 16 KB: 32 bits do; starts before the block
  // clamp to 0 -- only the first run can --, starts past it to 0xFFFFFFFE)
  __shared__ u32 sb[FILL_SMEM_RUNS + 2];
  __shared__ u8 ssym[FILL_SMEM_RUNS + 2];
  __shared__ u32 s_m[FILL_VECS];
  __shared__ __align__(128) u8 s_out[FILL_BYTES];
  __shared__ u32 s_q[PULL_QCAP], s_qn;
  typedef cub::BlockScan<u32, FILL_THREADS> BS;
  __shared__ typename BS::TempStorage ts;
  const u32 tid = threadIdx.x;
  const u64 o0 = (u64)blockIdx.x * FILL_BYTES;
  const u64 o1 = min(N, o0 + FILL_BYTES);
  const u32 olen = (u32)(o1 - o0);
  const u32 r0 = bstart[blockIdx.x], r1 = bstart[blockIdx.x + 1];
  const u32 nr = r1 - r0 + 1;                 // runs r0..r1 intersect this block
  const bool use_s = nr <= FILL_SMEM_RUNS;
  // the head list of this thread's first run is requested before anything else: its two dependent
  // reads (record -> hw) are in flight while the block expands the runs
  uint4 h_pre = make_uint4(0u, 0u, 0u, 0u);
  u64 e_pre0 = 0, e_pre1 = 0;
  if (tid < nr) h_pre = __ldg(vrec + r0 + tid);
  const u64 base_r0 = __ldg(base + r0);
  if (tid == 0) s_qn = 0;
  if (use_s) {
    for (u32 k = tid; k <= nr; k += FILL_THREADS) {
      const u64 bk = base[r0 + k];
      sb[k] = bk <= o0 ? 0u : (bk - o0 >= 0xFFFFFFFEull ? 0xFFFFFFFEu : (u32)(bk - o0));
      ssym[k] = k < nr ? vsym[r0 + k] : 0;
    }
    if (tid == 0) sb[nr + 1] = 0xFFFFFFFFu;    // the last block runs one entry past base[nv] = N
    for (u32 v = tid; v < FILL_VECS; v += FILL_THREADS) s_m[v] = 0;
  }
  if (vrec_heads(h_pre)) {                       // heads inside the record cost nothing here
    e_pre0 = vrec_head(h_pre, hw, 0);
    if (vrec_heads(h_pre) > 1) e_pre1 = vrec_head(h_pre, hw, 1);
  }
  if (use_s) {
    __syncthreads();
    for (u32 k = 1 + tid; k < nr; k += FILL_THREADS) {
      const u32 s = sb[k];
      if (s < olen) atomicMax(&s_m[s >> 4], k);
    }
    __syncthreads();
    u32 it[FILL_VEC_PER_THREAD];
#pragma unroll
    for (int v = 0; v < FILL_VEC_PER_THREAD; v++) it[v] = s_m[tid * FILL_VEC_PER_THREAD + v];
    BS(ts).InclusiveScan(it, it, cub::Max());
    __syncthreads();
#pragma unroll
    for (int v = 0; v < FILL_VEC_PER_THREAD; v++) s_m[tid * FILL_VEC_PER_THREAD + v] = it[v];
    __syncthreads();
  }
#pragma unroll
  for (int v = 0; v < FILL_VEC_PER_THREAD; v++) {
    const u32 vi = (u32)v * FILL_THREADS + tid;
    const u32 orel = vi * 16;
    if (orel >= olen) break;
    u32 k;
    u64 lo, hi;
    if (use_s) {
      k = vi ? s_m[vi - 1] : 0;
      while (sb[k + 1] <= orel) k++;           // runs that start exactly here, empty runs
      lo = hi = (u64)ssym[k] * 0x0101010101010101ull;
      u32 nb = sb[k + 1];
      while (nb < orel + 16) {                 // a run starts inside this vector: overwrite its bytes
        k++;
        const u32 d = nb - orel;
        const u64 sp = (u64)ssym[k] * 0x0101010101010101ull;
        const u64 mlo = d < 8 ? ~0ull << (8 * d) : 0ull;
        const u64 mhi = d < 8 ? ~0ull : ~0ull << (8 * (d - 8));
        lo = (lo & ~mlo) | (sp & mlo);
        hi = (hi & ~mhi) | (sp & mhi);
        nb = sb[k + 1];
      }
    } else {
      const u64 *bb = base + r0;
      const u64 o = o0 + orel;
      k = run_of(bb, 0, nr, o);
      lo = hi = (u64)vsym[r0 + k] * 0x0101010101010101ull;
      u64 nb = bb[k + 1];
      while (nb < o + 16) {
        k++;
        const u32 d = (u32)(nb - o);
        const u64 sp = (u64)vsym[r0 + k] * 0x0101010101010101ull;
        const u64 mlo = d < 8 ? ~0ull << (8 * d) : 0ull;
        const u64 mhi = d < 8 ? ~0ull : ~0ull << (8 * (d - 8));
        lo = (lo & ~mlo) | (sp & mlo);
        hi = (hi & ~mhi) | (sp & mhi);
        nb = bb[k + 1];
      }
    }
    *(uint4 *)(s_out + orel) = make_uint4((u32)lo, (u32)(lo >> 32), (u32)hi, (u32)(hi >> 32));
  }
  __syncthreads();
  // ---- pull the head symbols of the buckets that intersect this block (:1653-1676 as a gather) ----
  // offset of a head inside the block = (start of its run - o0) + within; only the first run can start
  // before the block (signed arithmetic on its true start)
  for (u32 k = tid; k < nr; k += FILL_THREADS) {
    const uint4 h = k == tid ? h_pre : __ldg(vrec + r0 + k);
    const u32 nh = vrec_heads(h);
    if (!nh) continue;
    if (nh > PULL_SERIAL) {
      const u32 q = atomicAdd(&s_qn, 1u);
      if (q < PULL_QCAP) { s_q[q] = k; continue; }
    }
    const i64 b0 = k == 0 ? (i64)(base_r0 - o0) : (i64)(use_s ? (u64)sb[k] : base[r0 + k] - o0);
    for (u32 j = 0; j < nh; j++) {
      const u64 x = (k == tid && j < 2) ? (j ? e_pre1 : e_pre0) : vrec_head(h, hw, j);
      const i64 off = b0 + (i64)(x >> 8);
      if (off >= (i64)olen) break;             // ascending within
      if (off >= 0) s_out[off] = (u8)x;
    }
  }
  __syncthreads();
  {
    const u32 qn = min(s_qn, (u32)PULL_QCAP), lane = tid & 31;
    for (u32 q = tid >> 5; q < qn; q += FILL_THREADS / 32) {
      const u32 k = s_q[q];
      const uint4 h = __ldg(vrec + r0 + k);
      const u32 nh = vrec_heads(h);                // > PULL_SERIAL: never an inline record
      const i64 b0 = k == 0 ? (i64)(base_r0 - o0) : (i64)(use_s ? (u64)sb[k] : base[r0 + k] - o0);
      const u64 *e = hw + h.y;
      for (u32 j = lane; j < nh; j += 32) {
        const u64 x = __ldg(e + j);
        const i64 off = b0 + (i64)(x >> 8);
        if (off >= 0 && off < (i64)olen) s_out[off] = (u8)x;
      }
    }
  }
  // generic-proxy writes to shared memory must be visible to the async proxy before the bulk copy:
  // every writer fences, then the barrier, then one thread issues the copy
  if (STORE == 1) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  const u32 bytes = (olen + 15) & ~15u;                        // out has CMS_PAD bytes of slack
  if (STORE == 1) {
    if (tid == 0) bulk_store_block(out + o0, s_out, bytes);
  } else {
    for (u32 vi = tid; vi * 16 < bytes; vi += FILL_THREADS) *(uint4 *)(out + o0 + (u64)vi * 16) = *(const uint4 *)(s_out + vi * 16);
  }
}

// text-order offsets for the parity hook (cmsbwt_heads_download): off = base[virtual rank of pos] + within
__global__ void k_head_offsets(const u32 *pos, const u32 *len, const u32 *within, const u32 *isa, const u64 *base,
                               u32 nd, u64 H, u64 *off) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= H) return;
  off[i] = len[i] == 0 ? (u64)within[i] : base[nd + isa[pos[i]] - 1] + within[i];
}

// ---- run-length records without the plain BWT (:1695-1777 with the remainder
// flush after the sub-bucket loop): the symbol sequence of bucket v is vsym[v] everywhere except at
// the heads' offsets, so its runs follow from O(1 + heads) segments; runs are merged across buckets
// through the last symbol before each bucket (a scan), and a record's length is the distance to the
// next run start.  Nothing here is O(N).
#define RLE_NONE 0xFFu
__global__ void k_rled_last(const u32 *cnt, const u8 *vsym, const uint4 *vrec, const u64 *hw, u32 nv, u8 *lastsym) {
  const u32 v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nv) return;
  const u32 c = cnt[v];
  u8 s = RLE_NONE;
  if (c) {
    s = vsym[v];
    const uint4 h = vrec[v];
    const u32 nh = vrec_heads(h);
    if (nh) { const u64 x = vrec_head(h, hw, nh - 1); if ((u32)(x >> 8) == c - 1) s = (u8)x; }
  }
  lastsym[v] = s;
}
struct LastValid {
  __device__ __forceinline__ u8 operator()(const u8 &a, const u8 &b) const { return b != RLE_NONE ? b : a; }
};
// walks the segments of bucket v; F(offset in the bucket, symbol) is called at every run start
template <class F>
__device__ __forceinline__ void rled_walk(u32 c, u8 vs, const uint4 h, const u64 *hw, u8 prev, F f) {
  if (!c) return;
  const u32 nh = vrec_heads(h);
  u32 at = 0;
  for (u32 j = 0; j < nh; j++) {
    const u64 x = vrec_head(h, hw, j);
    const u32 w = (u32)(x >> 8);
    const u8 s = (u8)x;
    if (w > at) { if (vs != prev) f(at, vs); prev = vs; }
    if (s != prev) f(w, s);
    prev = s;
    at = w + 1;
  }
  if (at < c && vs != prev) f(at, vs);
}
__global__ void k_rled_count(const u32 *cnt, const u8 *vsym, const uint4 *vrec, const u64 *hw, const u8 *prevsym, u32 nv,
                             u32 *nstart) {
  const u32 v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nv) return;
  u32 k = 0;
  rled_walk(cnt[v], vsym[v], vrec[v], hw, prevsym[v], [&](u32, u8) { k++; });
  nstart[v] = k;
}
__global__ void k_rled_write(const u32 *cnt, const u8 *vsym, const uint4 *vrec, const u64 *hw, const u8 *prevsym,
                             const u64 *base, const u64 *soff, u32 nv, u64 *starts, u8 *syms) {
  const u32 v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nv) return;
  u64 w = soff[v];
  const u64 b0 = base[v];
  rled_walk(cnt[v], vsym[v], vrec[v], hw, prevsym[v], [&](u32 at, u8 s) { starts[w] = b0 + at; syms[w] = s; w++; });
}
__global__ void k_rled_records(const u64 *starts, const u8 *syms, u64 R, u64 N, u8 *rec) {
  u64 k = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= R) return;
  const u64 s = starts[k];
  const u64 run = (k + 1 < R ? starts[k + 1] : N) - s;
  u8 *d = rec + 9 * k;
  for (int b = 0; b < 8; b++) d[b] = (u8)(run >> (8 * b));
  d[8] = syms[k];
}

static int fill_store_variant() {     // measurement switch: CMSBWT_FILL_STORE=0 selects 16-byte stores instead of the bulk copy
  static int v = -1;
  if (v < 0) { const char *e = getenv("CMSBWT_FILL_STORE"); v = e ? atoi(e) : 1; }
  return v;
}

// Buckets of nv virtual ranks (records vrec[], head lists hw[], terminators termsym[]) -> d_out (plain) or
// d_rle (records) of N bytes; throws unless the bucket sizes add up to N.  keep_base: the caller wants
// base[] (parity hook).
static void emit_from_vrec(cmsbwt_ctx *c, int mode, uint4 *vrec, const u64 *hw, const u8 *termsym, u32 ndl,
                           u32 r0, u32 nv, u64 N, DBuf *keep_base) {
  cudaStream_t st = c->st;
  const int B = 256;
  DBuf cnt, base, vsym, tmp, bstart;
  cnt.alloc(((size_t)nv + 1) * 4, st);
  base.alloc(((size_t)nv + 2) * 8, st);
  vsym.alloc((size_t)nv + 16, st);
  GroupSpan *goff = new GroupSpan(c, CMSBWT_G_OFFSETS);
  if (nv) k_vrank<<<(nv + B - 1) / B, B, 0, st>>>(vrec, c->ix.rbwt, termsym, ndl, r0, nv, cnt.as<u32>(), vsym.as<u8>());
  CK(cudaMemsetAsync(cnt.as<u32>() + nv, 0, 4, st));
  size_t tb = 0;
  EXSUM64(nullptr, tb, cnt.as<u32>(), base.as<u64>(), (int)(nv + 1), st);
  tmp.alloc(tb, st);
  CK(EXSUM64(tmp.p, tb, cnt.as<u32>(), base.as<u64>(), (int)(nv + 1), st));
  CK(cudaMemsetAsync(base.as<u64>() + (nv + 1), 0xFF, 8, st));     // sentinel past base[nv] = N
  c->launches += 3;
  const u64 nblocks = (N + FILL_BYTES - 1) / FILL_BYTES;
  if (mode == CMSBWT_MODE_PLAIN) {
    bstart.alloc((nblocks + 1) * 4, st);
    k_fill_part<<<(u32)((nblocks + 1 + B - 1) / B), B, 0, st>>>(base.as<u64>(), nv, N, nblocks, 0, bstart.as<u32>());
    LAUNCH(c);
  }
  delete goff;
  u64 total = 0;
  CK(cudaMemcpyAsync(&total, base.as<u64>() + nv, 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (total != N) {
    char buf[128];
    snprintf(buf, sizeof buf, "coverage total %llu != N %llu", (unsigned long long)total, (unsigned long long)N);
    throw CmsErr{CMSBWT_E_INTERNAL, buf};
  }
  c->out_mode = mode;
  c->stats.n_runs = 0;
  if (mode == CMSBWT_MODE_PLAIN) {
    GroupSpan g(c, CMSBWT_G_FILL);
    c->d_out.alloc(N + CMS_PAD, st);
    u8 *out = c->d_out.as<u8>();
    if (nblocks) {
      if (fill_store_variant() == 1)
        k_fill_pull<1><<<(u32)nblocks, FILL_THREADS, 0, st>>>(base.as<u64>(), vsym.as<u8>(), vrec, hw,
                                                              bstart.as<u32>(), N, out);
      else
        k_fill_pull<0><<<(u32)nblocks, FILL_THREADS, 0, st>>>(base.as<u64>(), vsym.as<u8>(), vrec, hw,
                                                              bstart.as<u32>(), N, out);
      LAUNCH(c);
    }
    c->out_bytes = N;
  } else {
    GroupSpan g(c, CMSBWT_G_RLE);
    DBuf lastsym, prevsym, nstart, soff, starts, syms;
    lastsym.alloc((size_t)nv + 16, st); prevsym.alloc((size_t)nv + 16, st);
    nstart.alloc(((size_t)nv + 1) * 4, st); soff.alloc(((size_t)nv + 2) * 8, st);
    const u32 gV = (nv + B - 1) / B;
    u64 R = 0;
    if (nv) {
      k_rled_last<<<gV, B, 0, st>>>(cnt.as<u32>(), vsym.as<u8>(), vrec, hw, nv, lastsym.as<u8>());
      size_t t1 = 0, t2 = 0;
      cub::DeviceScan::ExclusiveScan(nullptr, t1, lastsym.as<u8>(), prevsym.as<u8>(), LastValid(), (u8)RLE_NONE, (int)nv, st);
      EXSUM64(nullptr, t2, nstart.as<u32>(), soff.as<u64>(), (int)(nv + 1), st);
      DBuf tmp2; tmp2.alloc(t1 > t2 ? t1 : t2, st);
      size_t tt = t1;
      CK(cub::DeviceScan::ExclusiveScan(tmp2.p, tt, lastsym.as<u8>(), prevsym.as<u8>(), LastValid(), (u8)RLE_NONE, (int)nv, st));
      k_rled_count<<<gV, B, 0, st>>>(cnt.as<u32>(), vsym.as<u8>(), vrec, hw, prevsym.as<u8>(), nv, nstart.as<u32>());
      CK(cudaMemsetAsync(nstart.as<u32>() + nv, 0, 4, st));
      tt = t2;
      CK(EXSUM64(tmp2.p, tt, nstart.as<u32>(), soff.as<u64>(), (int)(nv + 1), st));
      CK(cudaMemcpyAsync(&R, soff.as<u64>() + nv, 8, cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
    }
    starts.alloc((R + 1) * 8, st); syms.alloc(R + 16, st);
    c->d_rle.alloc(R * 9 + CMS_PAD, st);
    if (nv) k_rled_write<<<gV, B, 0, st>>>(cnt.as<u32>(), vsym.as<u8>(), vrec, hw, prevsym.as<u8>(),
                                           base.as<u64>(), soff.as<u64>(), nv, starts.as<u64>(), syms.as<u8>());
    if (R) k_rled_records<<<(u32)((R + B - 1) / B), B, 0, st>>>(starts.as<u64>(), syms.as<u8>(), R, N, c->d_rle.as<u8>());
    c->launches += 7;
    c->out_bytes = R * 9;
    c->rle_off = 0;
    c->stats.n_runs = R;
    c->d_out.release();
  }
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));
  if (keep_base) keep_base->swap(base);
}

static void emit_pull(cmsbwt_ctx *c, int mode) {
  cudaStream_t st = c->st;
  const u32 n = c->ix.n;
  const u64 N = c->N, H = c->H;
  const int B = 256;
  const u64 nd64 = c->stats.n_docs;
  if (nd64 + n >= 0xFFFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "documents + reference length exceed 2^32"};
  const u32 nd = (u32)nd64, nv = nd + n - 1;
  {
    GroupSpan g(c, CMSBWT_G_PLACE);
    tiles_count_heads(c);
  }
  DBuf base;
  emit_from_vrec(c, mode, c->d_pcov.as<uint4>(), c->d_hw.as<u64>(), c->d_termsym.as<u8>(), nd, 0, nv, N,
                 c->dump ? &base : nullptr);
  if (c->dump) {
    c->h_off.alloc(H * 8, st);
    k_head_offsets<<<(u32)((H + B - 1) / B), B, 0, st>>>(c->h_pos.as<u32>(), c->h_len.as<u32>(), c->h_within.as<u32>(),
                                                        c->ix.isa, base.as<u64>(), nd, H, c->h_off.as<u64>());
    LAUNCH(c);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(st));
  }
  c->d_hw.release(); c->d_pcov.release(); c->d_termsym.release();
  c->t_he.release(); c->t_re.release(); c->t_hoff.release(); c->t_roff.release();
}

void emit_stage(cmsbwt_ctx *c, int mode) {
  if (c->light) { emit_pull(c, mode); return; }
  cudaStream_t st = c->st;
  const u32 n = c->ix.n;
  const u64 N = c->N, H = c->H;
  const int B = 256;
  DBuf cnt, base, tmp, bstart;
  cnt.alloc(((size_t)n + 1) * 4, st);
  base.alloc(((size_t)n + 2) * 8, st);
  GroupSpan *goff = new GroupSpan(c, CMSBWT_G_OFFSETS);
  size_t tb = 0, tb2 = 0;
  cub::DeviceScan::InclusiveSum(nullptr, tb, c->d_cov.as<i32>(), c->d_cov.as<i32>(), (int)n, st);
  EXSUM64(nullptr, tb2, cnt.as<u32>(), base.as<u64>(), (int)(n + 1), st);
  if (tb2 > tb) tb = tb2;
  tmp.alloc(tb, st);
  size_t t = tb;
  // sort-based path: coverage = inclusive scan of the deltas (in place)
  CK(cub::DeviceScan::InclusiveSum(tmp.p, t, c->d_cov.as<i32>(), c->d_cov.as<i32>(), (int)n, st));
  LAUNCH(c);
  CK(cudaMemsetAsync(cnt.as<u32>() + n, 0, 4, st));
  k_cnt_sa<<<(n + B - 1) / B, B, 0, st>>>(c->d_cov.as<i32>(), c->ix.sa, n, cnt.as<u32>());
  t = tb;
  CK(EXSUM64(tmp.p, t, cnt.as<u32>(), base.as<u64>(), (int)(n + 1), st));
  CK(cudaMemsetAsync(base.as<u64>() + (n + 1), 0xFF, 8, st));     // sentinel past base[n] = N
  c->launches += 2;
  const u64 nblocks = (N + FILL_BYTES - 1) / FILL_BYTES;
  bstart.alloc((nblocks + 1) * 4, st);
  k_fill_part<<<(u32)((nblocks + 1 + B - 1) / B), B, 0, st>>>(base.as<u64>(), n, N, nblocks, 0, bstart.as<u32>());
  LAUNCH(c);
  delete goff;
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
  u8 *out = c->d_out.as<u8>();
  { GroupSpan g(c, CMSBWT_G_FILL);
  k_fill<<<(u32)nblocks, FILL_THREADS, 0, st>>>(base.as<u64>(), c->ix.rbwt, bstart.as<u32>(), n, N, 0, out); }
  LAUNCH(c);
  {
    GroupSpan gplace(c, CMSBWT_G_PLACE);
    c->h_off.alloc(H * 8, st);
    k_scatter_heads<<<(u32)((H + B - 1) / B), B, 0, st>>>(c->h_pos.as<u32>(), c->h_cs.as<u16>(),
                                                         c->h_within.as<u32>(), c->ix.isa,
                                                         base.as<u64>(), H, out, c->h_off.as<u64>());
    LAUNCH(c);
  }
  CK(cudaGetLastError());
  c->out_mode = mode;
  c->out_bytes = N;
  c->stats.n_runs = 0;
  if (mode == CMSBWT_MODE_RLE) { GroupSpan g(c, CMSBWT_G_RLE); rle_encode(c, out, N); }
  CK(cudaStreamSynchronize(st));
}

// ================================================================================================
// Multi-GPU emit: every device gets the all-gathered bucket sizes, computes the
// (replicated, n-sized) offsets, expands the reference symbols of ITS contiguous output range and
// produces the head placements of ITS tile range sorted by output offset (exchanged by the host
// layer, then stored with shard_apply).
void shard_place(cmsbwt_ctx *c, const u64 *base, u64 doc_base);

void shard_emit(cmsbwt_ctx *c, const u32 *cov_full, u64 N_total, u32 world, u32 rank, u64 doc_base) {
  cudaStream_t st = c->st;
  const u32 n = c->ix.n;
  const int B = 256;
  DBuf cnt, tmp, bstart;
  cnt.alloc(((size_t)n + 1) * 4, st);
  c->sh_base.alloc(((size_t)n + 2) * 8, st);
  u64 *base = c->sh_base.as<u64>();
  size_t tb = 0;
  EXSUM64(nullptr, tb, cnt.as<u32>(), base, (int)(n + 1), st);
  tmp.alloc(tb, st);
  CK(cudaMemsetAsync(cnt.as<u32>() + n, 0, 4, st));
  k_cnt_sa<<<(n + B - 1) / B, B, 0, st>>>((const i32 *)cov_full, c->ix.sa, n, cnt.as<u32>());
  CK(EXSUM64(tmp.p, tb, cnt.as<u32>(), base, (int)(n + 1), st));
  CK(cudaMemsetAsync(base + (n + 1), 0xFF, 8, st));
  c->launches += 2;
  u64 total = 0;
  CK(cudaMemcpyAsync(&total, base + n, 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (total != N_total) {
    char buf[128];
    snprintf(buf, sizeof buf, "coverage total %llu != N %llu", (unsigned long long)total,
             (unsigned long long)N_total);
    throw CmsErr{CMSBWT_E_INTERNAL, buf};
  }
  c->sh_Ntotal = N_total;
  const u64 nblocks = (N_total + FILL_BYTES - 1) / FILL_BYTES;
  const u64 bp = (nblocks + world - 1) / world;
  const u64 b0 = min(nblocks, (u64)rank * bp), b1 = min(nblocks, b0 + bp);
  c->sh_chunk = bp * FILL_BYTES;
  c->sh_obeg = b0 * FILL_BYTES;
  c->sh_olen = min(N_total, b1 * FILL_BYTES) - min(N_total, c->sh_obeg);
  const u64 nb = b1 - b0;
  c->d_out.alloc(c->sh_olen + CMS_PAD, st);
  if (nb) {
    bstart.alloc((nb + 1) * 4, st);
    k_fill_part<<<(u32)((nb + 1 + B - 1) / B), B, 0, st>>>(base, n, N_total, nb, b0, bstart.as<u32>());
    k_fill<<<(u32)nb, FILL_THREADS, 0, st>>>(base, c->ix.rbwt, bstart.as<u32>(), n, N_total, b0, c->d_out.as<u8>());
    c->launches += 2;
  }
  CK(cudaGetLastError());
  shard_place(c, base, doc_base);
  c->out_mode = CMSBWT_MODE_PLAIN;
  c->out_bytes = c->sh_olen;
}

__global__ void k_sh_apply(const u64 *plc, u64 cnt, u64 obeg, u64 olen, u8 *out, u32 *bad) {
  u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= cnt) return;
  u64 p = plc[j], off = (p >> 8) - obeg;
  if (off < olen) out[off] = (u8)(p & 0xFF); else *bad = 1;
}

void shard_apply(cmsbwt_ctx *c, const u64 *plc, u64 cnt) {
  cudaStream_t st = c->st;
  DBuf bad;
  bad.alloc(4, st);
  CK(cudaMemsetAsync(bad.p, 0, 4, st));
  if (cnt) {
    k_sh_apply<<<(u32)((cnt + 255) / 256), 256, 0, st>>>(plc, cnt, c->sh_obeg, c->sh_olen, c->d_out.as<u8>(),
                                                        bad.as<u32>());
    LAUNCH(c);
  }
  u32 hb = 0;
  CK(cudaMemcpyAsync(&hb, bad.p, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (hb) throw CmsErr{CMSBWT_E_INTERNAL, "a head placement fell outside this device's output range"};
  c->sh_base.release(); c->sh_plc.release(); c->t_he.release(); c->t_re.release();
}

// ================================================================================================
// Multi-GPU pull emit, receiver side (sender: shard_rank_pull in tiles.cu).  This device owns the SA
// ranks [1 + R * rper, 1 + (R + 1) * rper) (device 0 also the terminators in front of them); the records
// of its buckets arrived grouped by source, their head lists likewise.  Scatter the records to their
// SA rank -- a region of 16 * n / world bytes, L2-sized from 8 devices on -- and run the single-GPU
// emit on the local buckets: offsets, fill and run-length records are all local, the output slices
// of the devices follow each other in rank order.
__global__ void __launch_bounds__(256) k_sh_vscatter(const uint4 *__restrict__ rec, u64 nrec, const u64 *__restrict__ rec_base,
                                                     const u64 *__restrict__ hw_base, u32 W, u32 ndl, u32 r0, u32 nvl,
                                                     uint4 *vrec, u32 *bad) {
  __shared__ u64 s_rb[65], s_hb[65];
  if (threadIdx.x <= W) s_rb[threadIdx.x] = rec_base[threadIdx.x];
  if (threadIdx.x < W) s_hb[threadIdx.x] = hw_base[threadIdx.x];
  __syncthreads();
  const u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nrec) return;
  u32 s = 0;
  while (s + 1 < W && j >= s_rb[s + 1]) s++;
  uint4 a = rec[j];
  a.y += (u32)s_hb[s];
  const u32 idx = ndl + (a.w - r0);
  if (a.w < r0 || idx >= nvl) { *bad = 1; return; }
  a.w = 0;
  vrec[idx] = a;
}

void shard_emit_pull(cmsbwt_ctx *c, int mode, u32 W, u32 R, u32 nd_total, const void *recs, u64 nrecs,
                     const u64 *rec_base, const u64 *hw_base, const void *hw, const u8 *termsym, u64 expect_bytes) {
  cudaStream_t st = c->st;
  const u32 n = c->ix.n;
  const u32 rper = c->sh_rper;
  const u64 r0 = (u64)R * rper;
  const u32 nranks = r0 >= (u64)n - 1 ? 0u : (u32)std::min<u64>(rper, (u64)n - 1 - r0);
  const u32 ndl = R == 0 ? nd_total : 0u;
  const u32 nvl = ndl + nranks;
  if (nrecs != nranks) throw CmsErr{CMSBWT_E_INTERNAL, "bucket records received != SA ranks owned"};
  DBuf vrec, rb, hb, bad;
  vrec.alloc(((size_t)nvl + 1) * 16, st);
  rb.alloc((size_t)(W + 1) * 8, st); hb.alloc((size_t)W * 8, st); bad.alloc(4, st);
  CK(cudaMemsetAsync(bad.p, 0, 4, st));
  CK(cudaMemcpyAsync(rb.p, rec_base, (size_t)(W + 1) * 8, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(hb.p, hw_base, (size_t)W * 8, cudaMemcpyHostToDevice, st));
  {
    GroupSpan g(c, CMSBWT_G_OFFSETS);
    if (nrecs) {
      k_sh_vscatter<<<(u32)((nrecs + 255) / 256), 256, 0, st>>>((const uint4 *)recs, nrecs, rb.as<u64>(), hb.as<u64>(), W, ndl,
                                                               (u32)r0, nvl, vrec.as<uint4>(), bad.as<u32>());
      LAUNCH(c);
    }
  }
  u32 hbad = 0;
  CK(cudaMemcpyAsync(&hbad, bad.p, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (hbad) throw CmsErr{CMSBWT_E_INTERNAL, "a bucket record arrived at a device that does not own its SA rank"};
  emit_from_vrec(c, mode, vrec.as<uint4>(), (const u64 *)hw, termsym, ndl, (u32)r0, nvl, expect_bytes, nullptr);
  c->sh_olen = expect_bytes;
}
