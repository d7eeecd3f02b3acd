// Reference index on the device: suffix array (prefix doubling), ISA, PLCP' -> G, reference BWT,
// k-mer interval table.  Replaces initialize_reference (CMS-BWT-functions.cpp:246-297).
#include "common.cuh"
#include "ms_fast.cuh"
#include <cub/cub.cuh>

// ------------------------------------------------------------------------------------------
// Suffix array by prefix doubling (replaces libsais() at :247 when the caller passes no SA).
// x[n-1] = 0 is the unique smallest symbol, so no comparison ever runs past the end.

__global__ void k_sa_init_keys(const u8 *x, u32 n, u64 *key, u32 *idx) {
  u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  u64 k = 0;
  for (int j = 0; j < 8; j++) {
    u32 p = i + j;
    u64 c = p < n ? x[p] : 0;
    k = (k << 8) | c;
  }
  key[i] = k;
  idx[i] = i;
}

// flag[j] = j if key starts a new group else 0  (group start index, for a max-scan)
__global__ void k_group_start(const u64 *key, u32 n, u32 *gs) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  gs[j] = (j == 0 || key[j] != key[j - 1]) ? j : 0;
}
__global__ void k_count_groups(const u64 *key, u32 n, u32 *cnt) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  bool f = j < n && (j == 0 || key[j] != key[j - 1]);
  unsigned b = __ballot_sync(0xffffffffu, f);
  if ((threadIdx.x & 31) == 0 && b) atomicAdd(cnt, __popc(b));
}
// rank[idx[j]] = group start + 1  (0 is reserved for "past the end")
__global__ void k_scatter_rank(const u32 *idx, const u32 *gs, u32 n, u32 *rank) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  rank[idx[j]] = gs[j] + 1;
}
__global__ void k_double_keys(const u32 *rank, u32 n, u32 h, u64 *key, u32 *idx) {
  u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  u64 r2 = (u64)i + h < n ? rank[i + h] : 0;
  key[i] = ((u64)rank[i] << 32) | r2;
  idx[i] = i;
}
__global__ void k_copy_u32_i32(const u32 *a, u32 n, i32 *b) {
  u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) b[i] = (i32)a[i];
}

void sa_build_device(cmsbwt_ctx *c, const u8 *d_x, u32 n, i32 *d_sa) {
  cudaStream_t st = c->st;
  DBuf key, key2, idx, idx2, rank, gs, cnt, tmp;
  key.alloc((size_t)n * 8, st); key2.alloc((size_t)n * 8, st);
  idx.alloc((size_t)n * 4, st); idx2.alloc((size_t)n * 4, st);
  rank.alloc((size_t)n * 4, st); gs.alloc((size_t)n * 4, st);
  cnt.alloc(4, st);
  size_t tb = 0, tb2 = 0;
  cub::DeviceScan::InclusiveScan(nullptr, tb2, gs.as<u32>(), gs.as<u32>(), cub::Max(), (int)n, st);
  if (tb2 > tb) tb = tb2;
  tmp.alloc(tb, st);
  const int B = 256;
  const u32 g = (n + B - 1) / B;
  k_sa_init_keys<<<g, B, 0, st>>>(d_x, n, key.as<u64>(), idx.as<u32>());
  u32 h = 8;
  for (int round = 0; round < 40; round++) {
    size_t t = tb;
    rsort_pairs64(c, key, key2, idx, idx2, n, 0, 64);
    key.swap(key2); idx.swap(idx2);               // sorted: key2 / idx2
    CK(cudaMemsetAsync(cnt.p, 0, 4, st));
    k_count_groups<<<g, B, 0, st>>>(key2.as<u64>(), n, cnt.as<u32>());
    u32 groups = 0;
    CK(cudaMemcpyAsync(&groups, cnt.p, 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (groups == n || (u64)h >= n) break;
    k_group_start<<<g, B, 0, st>>>(key2.as<u64>(), n, gs.as<u32>());
    t = tb;
    CK(cub::DeviceScan::InclusiveScan(tmp.p, t, gs.as<u32>(), gs.as<u32>(), cub::Max(), (int)n, st));
    k_scatter_rank<<<g, B, 0, st>>>(idx2.as<u32>(), gs.as<u32>(), n, rank.as<u32>());
    k_double_keys<<<g, B, 0, st>>>(rank.as<u32>(), n, h, key.as<u64>(), idx.as<u32>());
    h *= 2;
  }
  k_copy_u32_i32<<<g, B, 0, st>>>(idx2.as<u32>(), n, d_sa);
  CK(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------
// ISA (:3-9), Phi, PLCP (libsais_plcp :270), PLCP' (:272-274) as G, reference BWT (:294-297)

__global__ void k_isa_phi_bwt(const i32 *sa, const u8 *x, u32 n, u32 *isa, i32 *phi, u8 *rbwt) {
  u32 r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  u32 s = (u32)sa[r];
  isa[s] = r;
  phi[s] = r > 0 ? sa[r - 1] : -1;
  rbwt[r] = s > 0 ? x[s - 1] : 0;
}

// Kasai recurrence over runs of PB consecutive text positions per thread (PLCP[i] >= PLCP[i-1]-1)
#define PLCP_PB 32
__global__ void k_plcp(const u8 *x, const i32 *phi, u32 n, u32 *plcp) {
  u32 t = blockIdx.x * blockDim.x + threadIdx.x;
  u64 i0 = (u64)t * PLCP_PB;
  if (i0 >= n) return;
  u32 i1 = (u32)min((u64)n, i0 + PLCP_PB);
  u32 l = 0;
  for (u32 i = (u32)i0; i < i1; i++) {
    i32 j = phi[i];
    if (j < 0) { plcp[i] = 0; l = 0; continue; }
    u32 lim = n - max(i, (u32)j);     // the two suffixes differ before either ends
    while (l < lim) {
      u64 d = ld8u(x + i + l) ^ ld8u(x + (u32)j + l);
      if (d) { l += (__ffsll((long long)d) - 1) >> 3; break; }
      l += 8;
    }
    if (l > lim) l = lim;
    plcp[i] = l;
    if (l > 0) l--;
  }
}
__global__ void k_G(const i32 *sa, const u32 *isa, const u32 *plcp, u32 n, u32 *G) {
  u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  u32 a = plcp[i];
  u32 r = isa[i];
  if (r + 1 < n) { u32 b = plcp[(u32)sa[r + 1]]; if (b > a) a = b; }
  G[i] = i + a;
}

// ------------------------------------------------------------------------------------------
// k-mer table: ktab[m] = number of suffixes smaller than the ACGT k-mer m (m in [0,4^k]).
// f(r) = number of k-mers <= suffix r (non-decreasing in r); ktab[m] = r for f(r-1) <= m < f(r).
// (kmers_le_suffix lives in ms_fast.cuh as msf_kmers_le_suffix, shared with the CPU harness)
__global__ void k_ktab_mark(const u8 *x, const i32 *sa, u32 n, int k, u32 *ktab) {
  u32 r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  u32 f = msf_kmers_le_suffix(x, (u32)sa[r], k);
  u32 fp = r > 0 ? msf_kmers_le_suffix(x, (u32)sa[r - 1], k) : 0;
  if (f > fp) ktab[fp] = r;
  if (r == n - 1) { u32 K = 1u << (2 * k); if (f <= K) ktab[f] = n; }
}

// fast path: per SA rank the 16 bases after the k-mer (big-endian) and the suffix itself
__global__ void k_sak(const i32 *sa, const u64 *xp, const u8 *x8, u32 n, const u64 *xruns, const u32 *xbk, u32 nruns,
                      u32 n_acgt, int k, u64 *sak) {
  u32 r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  u32 s = (u32)sa[r];
  sak[r] = (u64)msf_make_key(xp, x8, s, msf_room_at(xruns, xbk, nruns, n_acgt, s), k) | ((u64)s << 32);
}

// k-mer slot table (ms_fast.cuh): one thread per k-mer writes its 64-byte slot
__global__ void k_slots(const u64 *sak, const u32 *ktab, const u64 *xp, u32 n, const u64 *xruns, const u32 *xbk,
                        u32 nruns, u32 n_acgt, int k, u64 K, ulonglong2 *slots) {
  u64 code = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (code >= K) return;
  u64 e[8];
  msf_slot_fill(sak, ktab, xp, n, xruns, xbk, nruns, n_acgt, k, code, e);
#pragma unroll
  for (int j = 0; j < 4; j++) slots[code * 4 + j] = make_ulonglong2(e[2 * j], e[2 * j + 1]);
}

// largest k with 4^k <= n: a k-mer then has n / 4^k < 4 suffixes on average, so nearly every
// search ends inside its 8-entry slot (k = 12 at 59 Mbp, 13 at 248 Mbp: measured 80 -> 25 ms for
// the MS kernel of the config-5 shape against k = 12, where most slots overflow)
static int pick_k(u32 n, i64 forced) {
  if (forced > 0) return (int)(forced > 14 ? 14 : forced);
  int k = 1;
  while (k < 14 && (1ull << (2 * (k + 1))) <= (u64)n) k++;
  return k;
}

void index_build(cmsbwt_ctx *c) {
  cudaStream_t st = c->st;
  const u32 n = c->ix.n;
  const int B = 256;
  const u32 g = (n + B - 1) / B;
  c->d_isa.alloc((size_t)n * 4, st);
  c->d_G.alloc((size_t)n * 4, st);
  c->d_rbwt.alloc((size_t)n + CMS_PAD, st);
  DBuf phi, plcp;
  phi.alloc((size_t)n * 4, st);
  plcp.alloc((size_t)n * 4, st);
  const u8 *x = c->d_x.as<u8>();
  const i32 *sa = c->d_sa.as<i32>();
  k_isa_phi_bwt<<<g, B, 0, st>>>(sa, x, n, c->d_isa.as<u32>(), phi.as<i32>(), c->d_rbwt.as<u8>());
  u32 nt = (n + PLCP_PB - 1) / PLCP_PB;
  k_plcp<<<(nt + B - 1) / B, B, 0, st>>>(x, phi.as<i32>(), n, plcp.as<u32>());
  k_G<<<g, B, 0, st>>>(sa, c->d_isa.as<u32>(), plcp.as<u32>(), n, c->d_G.as<u32>());
  int k = pick_k(n, c->kmer_k);
  size_t K = (size_t)1 << (2 * k);
  // k-mer table and (fast path) the packed reference share one allocation: together they are the
  // hot, L2-sized part of the index, pinned in L2 while the MS kernel runs (ms_fast.cu)
  const size_t ktab_bytes = (((K + 2) * 4) + 255) & ~(size_t)255;
  const u64 xp_words = c->pure ? ((u64)n + 31) / 32 + MSF_PAD_WORDS : 0;
  c->d_ktab.alloc(ktab_bytes + xp_words * 8, st);
  c->xp = c->pure ? (u64 *)((u8 *)c->d_ktab.p + ktab_bytes) : nullptr;
  c->hot_bytes = ktab_bytes + xp_words * 8;
  CK(cudaMemsetAsync(c->d_ktab.p, 0, (K + 2) * 4, st));
  k_ktab_mark<<<g, B, 0, st>>>(x, sa, n, k, c->d_ktab.as<u32>());
  size_t tb = 0;
  cub::DeviceScan::InclusiveScan(nullptr, tb, c->d_ktab.as<u32>(), c->d_ktab.as<u32>(), cub::Max(),
                                 (int)(K + 1), st);
  DBuf tmp;
  tmp.alloc(tb, st);
  CK(cub::DeviceScan::InclusiveScan(tmp.p, tb, c->d_ktab.as<u32>(), c->d_ktab.as<u32>(), cub::Max(),
                                    (int)(K + 1), st));
  CK(cudaGetLastError());
  c->ix.k = k;
  c->ix.x = x;
  c->ix.sa = sa;
  c->ix.isa = c->d_isa.as<u32>();
  c->ix.G = c->d_G.as<u32>();
  c->ix.rbwt = c->d_rbwt.as<u8>();
  c->ix.ktab = c->d_ktab.as<u32>();
  c->stats.kmer_k = (u32)k;
  if (c->pure) {
    const u64 nwords = xp_words;
    c->d_sak.alloc((size_t)n * 8, st);
    DBuf fl;
    fl.alloc(8, st);
    CK(cudaMemsetAsync(fl.p, 0, 8, st));
    // the whole of x is packed: inside the runs of other symbols the codes are arbitrary and never compared
    pack_2bit(c, x, n, c->xp, nwords, nullptr, 0, nullptr, fl.as<u32>());
    k_sak<<<g, B, 0, st>>>(sa, c->xp, x, n, c->d_xruns.as<u64>(), c->d_xbk.as<u32>(), c->nruns, c->n_acgt, k,
                           c->d_sak.as<u64>());
    if (c->use_slots) {
      c->d_slots.alloc(K * 64, st);
      k_slots<<<(u32)((K + B - 1) / B), B, 0, st>>>(c->d_sak.as<u64>(), c->d_ktab.as<u32>(), c->xp, n,
                                                    c->d_xruns.as<u64>(), c->d_xbk.as<u32>(), c->nruns, c->n_acgt, k,
                                                    (u64)K, (ulonglong2 *)c->d_slots.p);
    } else {
      c->d_slots.release();
    }
    CK(cudaGetLastError());
  } else {
    c->d_sak.release();
    c->d_slots.release();
  }
  CK(cudaStreamSynchronize(st));
}
