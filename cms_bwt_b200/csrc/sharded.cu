// Multi-GPU host layer in C++ over NCCL (north-star: "C++ host code ... NCCL all-to-all
// over NVLink").  One context per GPU, one rank per context (one process per GPU under torchrun/mpirun,
// or one host thread per GPU inside the cms_bwt command line).  The reference is single threaded and
// has no counterpart of this file; the order of the steps is the order of its stages
// (CMS-BWT-functions.cpp:1138-1213 MS per document, :1359-1603 head order, :1618-1777 emit) with a
// collective wherever data changes owner:
//
//   stage 1   matching statistics of this rank's documents                 (no communication)
//   stage 2a  tails -> owners by 32-bit key range          all-to-all      (fixed ranges of the key space)
//             ranks + tie flags back                       all-to-all
//             refinement of the still tied tails           all-gather per round (small)
//   stage 2b  tile entries -> owners by reference-position range           all-to-all
//   stage 3   bucket records + head lists -> owners by SA-rank range       all-to-all (pull emit; by output
//             byte range with head placements for collections whose tiles are too deep for k_tile_rank)
//
// All collectives are enqueued on the context's own stream, so they are ordered with the kernels
// either side of them without host synchronisation; the host only waits where it needs a count.
// NCCL is loaded at run time (dlopen of libnccl.so.2, or CMSBWT_NCCL=<path>): the library has no
// link-time dependency on it and single-GPU users never touch it.  cms_bwt_b200/multi.py keeps the
// same orchestration in Python for the gloo model tests (CPU) only.
#include "common.cuh"
#include <nccl.h>
#include <dlfcn.h>
#include <algorithm>
#include <chrono>
#include <mutex>
#include <string.h>
#include <stdlib.h>

struct NcclApi {
  void *h = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;
static std::once_flag g_nccl_once;

static NcclApi &nccl() {
  std::call_once(g_nccl_once, [] {
    const char *cands[] = {getenv("CMSBWT_NCCL"), "libnccl.so.2", "libnccl.so"};
    for (const char *p : cands) {
      if (!p) continue;
      g_nccl.h = dlopen(p, RTLD_NOW | RTLD_GLOBAL);
      if (g_nccl.h) break;
    }
    if (!g_nccl.h) return;
#define NS(field, name) g_nccl.field = (decltype(g_nccl.field))dlsym(g_nccl.h, name)
    NS(GetUniqueId, "ncclGetUniqueId"); NS(CommInitRank, "ncclCommInitRank"); NS(CommDestroy, "ncclCommDestroy");
    NS(AllGather, "ncclAllGather"); NS(Send, "ncclSend"); NS(Recv, "ncclRecv"); NS(GroupStart, "ncclGroupStart");
    NS(GroupEnd, "ncclGroupEnd"); NS(GetErrorString, "ncclGetErrorString");
#undef NS
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommDestroy || !g_nccl.AllGather || !g_nccl.Send ||
        !g_nccl.Recv || !g_nccl.GroupStart || !g_nccl.GroupEnd)
      g_nccl.h = nullptr;
  });
  if (!g_nccl.h) throw CmsErr{CMSBWT_E_CUDA, "NCCL (libnccl.so.2) could not be loaded; set CMSBWT_NCCL=<path>"};
  return g_nccl;
}
#define NK(call)                                                                                     \
  do {                                                                                               \
    ncclResult_t _r = (call);                                                                        \
    if (_r != ncclSuccess)                                                                           \
      throw CmsErr{CMSBWT_E_CUDA, std::string("NCCL: ") + #call + ": " +                             \
                                      (nccl().GetErrorString ? nccl().GetErrorString(_r) : "error")}; \
  } while (0)

// ---- small kernels of the host layer ------------------------------------------------------------
__global__ void k_shx_sample(const u32 *ks, u32 Tn, u32 S, u32 *out) {
  u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= S) return;
  out[i] = Tn ? ks[(u32)(((u64)i * (Tn - 1)) / (S - 1))] : 0xFFFFFFFFu;
}
__global__ void k_shx_lb(const u32 *a, u32 n, const u32 *v, u32 cnt, u32 *out) {
  u32 k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= cnt) return;
  u32 lo = 0, hi = n, key = v[k];
  while (lo < hi) { u32 mid = lo + ((hi - lo) >> 1); if (a[mid] < key) lo = mid + 1; else hi = mid; }
  out[k] = lo;
}
__global__ void k_shx_pack2(const u32 *hi, const u32 *lo, u32 n, u64 *out) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) out[j] = ((u64)hi[j] << 32) | lo[j];
}
__global__ void k_shx_unpack2(const u64 *in, u32 n, u32 *hi, u32 *lo) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) { u64 v = in[j]; hi[j] = (u32)(v >> 32); lo[j] = (u32)v; }
}
__global__ void k_shx_pack_rt(const u32 *rk, const u8 *tie, u32 n, u64 *out) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) out[j] = ((u64)tie[j] << 32) | rk[j];
}
__global__ void k_shx_unpack_rt(const u64 *in, u32 n, u32 *rk, u8 *tie) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) { u64 v = in[j]; rk[j] = (u32)v; tie[j] = (u8)(v >> 32); }
}
// ranks below 2^31: rank | tie << 31 in one word (halves the bytes of the exchange back to the tails' origins)
__global__ void k_shx_pack_rt32(const u32 *rk, const u8 *tie, u32 n, u32 *out) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) out[j] = rk[j] | ((u32)(tie[j] != 0) << 31);
}
__global__ void k_shx_unpack_rt32(const u32 *in, u32 n, u32 *rk, u8 *tie) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) { u32 v = in[j]; rk[j] = v & 0x7FFFFFFFu; tie[j] = (u8)(v >> 31); }
}
__global__ void k_shx_edges(const u32 *hoff, const u32 *roff, u32 tp, u32 W, u32 *out) {
  u32 d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d > W) return;
  out[d] = hoff[(size_t)d * tp];
  out[W + 1 + d] = roff[(size_t)d * tp];
}

// tails -> destination = owner of the key's range (fixed ranges of the key space: keys are SA ranks of
// random-looking reference positions, so equal ranges hold about equal numbers of tails, and equal keys
// always meet on one owner).  Counts, then one reservation per (block, destination).
#define SHX_MAXW 64
__global__ void __launch_bounds__(256) k_shx_part_hist(const u32 *__restrict__ key, u32 Tn, u32 kper, u32 W,
                                                       unsigned long long *cnt) {
  __shared__ u32 s_c[SHX_MAXW];
  if (threadIdx.x < SHX_MAXW) s_c[threadIdx.x] = 0;
  __syncthreads();
  for (u32 j = blockIdx.x * blockDim.x + threadIdx.x; j < Tn; j += gridDim.x * blockDim.x)
    atomicAdd(&s_c[min(W - 1, key[j] / kper)], 1u);
  __syncthreads();
  if (threadIdx.x < W && s_c[threadIdx.x]) atomicAdd(cnt + threadIdx.x, (unsigned long long)s_c[threadIdx.x]);
}
__global__ void k_shx_part_offsets(const unsigned long long *cnt, u32 W, unsigned long long *cur) {
  if (threadIdx.x == 0) { unsigned long long run = 0; for (u32 d = 0; d < W; d++) { cur[d] = run; run += cnt[d]; } }
}
__global__ void __launch_bounds__(256) k_shx_part_scatter(const u32 *__restrict__ key, const u64 *__restrict__ label, u32 Tn,
                                                          u32 kper, u32 W, unsigned long long *cur, u64 *kc, u32 *perm) {
  __shared__ u32 s_c[SHX_MAXW];
  __shared__ unsigned long long s_b[SHX_MAXW];
  const u32 per = (Tn + gridDim.x - 1) / gridDim.x;
  const u32 a = blockIdx.x * per, b = min(Tn, a + per);
  if (threadIdx.x < SHX_MAXW) s_c[threadIdx.x] = 0;
  __syncthreads();
  for (u32 j = a + threadIdx.x; j < b; j += blockDim.x) atomicAdd(&s_c[min(W - 1, key[j] / kper)], 1u);
  __syncthreads();
  if (threadIdx.x < W) {
    s_b[threadIdx.x] = s_c[threadIdx.x] ? atomicAdd(cur + threadIdx.x, (unsigned long long)s_c[threadIdx.x]) : 0ull;
    s_c[threadIdx.x] = 0;
  }
  __syncthreads();
  for (u32 j = a + threadIdx.x; j < b; j += blockDim.x) {
    const u32 k = key[j];
    const u32 d = min(W - 1, k / kper);
    const u64 o = s_b[d] + atomicAdd(&s_c[d], 1u);
    kc[o] = ((u64)k << 32) | (u32)label[j];          // (key, cls): what the owner ranks by
    perm[o] = j;
  }
}

namespace {
struct Phase {
  cmsbwt_ctx *c;
  std::chrono::high_resolution_clock::time_point t;
  explicit Phase(cmsbwt_ctx *c_) : c(c_), t(std::chrono::high_resolution_clock::now()) {}
  void mark(int id) {        // every phase ends on a host wait, so wall time is the phase's time
    auto now = std::chrono::high_resolution_clock::now();
    c->sh_phase_ms[id] += std::chrono::duration<double, std::milli>(now - t).count();
    t = now;
  }
};

struct Comm {
  cmsbwt_ctx *c;
  ncclComm_t comm;
  u32 W, R;
  cudaStream_t st;
  DBuf stage_in, stage_out;     // staging of small host vectors
  explicit Comm(cmsbwt_ctx *c_) : c(c_), comm((ncclComm_t)c_->comm), W(c_->comm_world), R(c_->comm_rank), st(c_->st) {}

  // all-gather of `per` u64 values per rank through the device -> host matrix [W][per]
  std::vector<u64> gather_u64(const std::vector<u64> &mine) {
    const size_t per = mine.size();
    stage_in.alloc(per * 8, st); stage_out.alloc(per * 8 * W, st);
    CK(cudaMemcpyAsync(stage_in.p, mine.data(), per * 8, cudaMemcpyHostToDevice, st));
    NK(nccl().AllGather(stage_in.p, stage_out.p, per, ncclUint64, comm, st));
    std::vector<u64> all(per * W);
    CK(cudaMemcpyAsync(all.data(), stage_out.p, per * 8 * W, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    c->sh_collectives++;
    return all;
  }
  void all_gather(const void *in, void *out, size_t bytes_per_rank) {
    NK(nccl().AllGather(in, out, bytes_per_rank, ncclUint8, comm, st));
    c->sh_collectives++;
    c->sh_comm_bytes += bytes_per_rank * W;
  }
  // slabs in rank order; counts in elements of `elt` bytes.  May be called between group_begin/end.
  void a2a(const void *send, const std::vector<u64> &scnt, void *recv, const std::vector<u64> &rcnt, size_t elt) {
    u64 so = 0, ro = 0;
    for (u32 p = 0; p < W; p++) {
      if (scnt[p]) NK(nccl().Send((const u8 *)send + so * elt, scnt[p] * elt, ncclUint8, (int)p, comm, st));
      if (rcnt[p]) NK(nccl().Recv((u8 *)recv + ro * elt, rcnt[p] * elt, ncclUint8, (int)p, comm, st));
      so += scnt[p]; ro += rcnt[p];
      c->sh_comm_bytes += (scnt[p] + rcnt[p]) * elt;
    }
  }
  // the same `cnt` elements to every rank; rcnt[p] elements from rank p (an all-gather of ragged slabs)
  void a2a_same(const void *send, u64 cnt, void *recv, const std::vector<u64> &rcnt, size_t elt) {
    u64 ro = 0;
    for (u32 p = 0; p < W; p++) {
      if (cnt) NK(nccl().Send(send, cnt * elt, ncclUint8, (int)p, comm, st));
      if (rcnt[p]) NK(nccl().Recv((u8 *)recv + ro * elt, rcnt[p] * elt, ncclUint8, (int)p, comm, st));
      ro += rcnt[p];
      c->sh_comm_bytes += (cnt + rcnt[p]) * elt;
    }
  }
  void group_begin() { NK(nccl().GroupStart()); }
  void group_end() { NK(nccl().GroupEnd()); c->sh_collectives++; }
};

u64 sum(const std::vector<u64> &v) { u64 s = 0; for (u64 x : v) s += x; return s; }
int isa_bits_of(u32 n) { int b = 1; while (b < 32 && (1ull << b) < (u64)n) b++; return b; }
}  // namespace

enum { PH_SIZES = 0, PH_TAILS, PH_X_TAILS, PH_RANK_OWNER, PH_X_RANKS, PH_REFINE, PH_TILES_LOCAL, PH_X_TILES,
       PH_TILES_MERGE, PH_X_COV, PH_EMIT, PH_X_PLC, PH_APPLY, PH_RLE, PH_PLACE, PH_X_VREC, PH_COUNT };
static_assert(PH_COUNT <= 16, "cmsbwt_ctx::sh_phase_ms holds 16 phases");

void shard_place(cmsbwt_ctx *c, const u64 *base, u64 doc_base);
void shard_term_syms(cmsbwt_ctx *c, u8 *termsym);

static void run_sharded_impl(cmsbwt_ctx *c, u64 N_local, u64 ndocs_local, int mode) {
  cudaStream_t st = c->st;
  Comm cm(c);
  const u32 W = cm.W, R = cm.R;
  const int B = 256;
  Phase ph(c);
  for (double &p : c->sh_phase_ms) p = 0;
  c->sh_collectives = 0;
  c->sh_comm_bytes = 0;

  // ---- the shard plan: symbols and documents of every rank
  std::vector<u64> plan = cm.gather_u64({N_local, ndocs_local});
  u64 N_total = 0, ndocs_total = 0, doc_base = 0;
  for (u32 r = 0; r < W; r++) {
    N_total += plan[2 * r]; ndocs_total += plan[2 * r + 1];
    if (r < R) doc_base += plan[2 * r + 1];
  }
  if (ndocs_total >= 0xFFFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "more than 2^32 documents"};
  ph.mark(PH_SIZES);

  // ---- stage 2a: global order of the tails (:1432-1494)
  {
    Timer t(st);
    sort_prepare(c, doc_base);
    c->stats.sort_ms = t.stop();
  }
  const u32 Tn = c->s_Tn;
  std::vector<u64> send(W, 0), recv(W, 0);
  std::vector<u64> cmat;
  DBuf kc;
  kc.alloc((size_t)Tn * 8 + 16, st);
  c->sh_perm.alloc((size_t)Tn * 4 + 16, st);
  {
    if (W > SHX_MAXW) throw CmsErr{CMSBWT_E_LIMIT, "more than 64 devices"};
    const int kb = c->s_key_bits > 32 ? 32 : c->s_key_bits;
    // keys are (SA rank << qb | quantised cls order): they span [0, n << qb), not the whole 2^kb
    const u64 kspan = std::min<u64>(1ull << kb, ((u64)c->ix.n + 1) << (kb - isa_bits_of(c->ix.n)));
    const u32 kper = (u32)std::max<u64>(1, (kspan + W - 1) / W);
    DBuf cnt, cur;
    cnt.alloc((size_t)W * 8, st); cur.alloc((size_t)W * 8, st);
    CK(cudaMemsetAsync(cnt.p, 0, (size_t)W * 8, st));
    const u32 gP = (u32)std::min<u64>(148 * 8, ((u64)Tn + 255) / 256);
    if (Tn) {
      k_shx_part_hist<<<gP, 256, 0, st>>>(c->s_key32.as<u32>(), Tn, kper, W, cnt.as<unsigned long long>());
      k_shx_part_offsets<<<1, 32, 0, st>>>(cnt.as<unsigned long long>(), W, cur.as<unsigned long long>());
      k_shx_part_scatter<<<gP, 256, 0, st>>>(c->s_key32.as<u32>(), c->s_label.as<u64>(), Tn, kper, W,
                                             cur.as<unsigned long long>(), kc.as<u64>(), c->sh_perm.as<u32>());
      c->launches += 3;
    }
    CK(cudaMemcpyAsync(send.data(), cnt.p, (size_t)W * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    ph.mark(PH_TAILS);
    cmat = cm.gather_u64(send);                     // count matrix cmat[src * W + dst]
  }
  u64 M64 = 0, rank_base = 0, tails_total = 0;
  for (u32 s = 0; s < W; s++) { recv[s] = cmat[(size_t)s * W + R]; M64 += recv[s]; }
  for (u32 d = 0; d < W; d++)
    for (u32 s = 0; s < W; s++) { tails_total += cmat[(size_t)s * W + d]; if (d < R) rank_base += cmat[(size_t)s * W + d]; }
  if (tails_total >= 0xFFFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "more than 2^32 tails in the collection"};
  const u32 M = (u32)M64;
  DBuf got, rk, rc, r_out, t_out, back, backr, rs, ts;
  got.alloc((size_t)M * 8 + 16, st);
  rk.alloc((size_t)M * 4 + 16, st); rc.alloc((size_t)M * 4 + 16, st);
  cm.group_begin();
  cm.a2a(kc.p, send, got.p, recv, 8);
  cm.group_end();
  if (M) k_shx_unpack2<<<(M + B - 1) / B, B, 0, st>>>(got.as<u64>(), M, rk.as<u32>(), rc.as<u32>());
  c->launches += 2;
  CK(cudaStreamSynchronize(st));
  ph.mark(PH_X_TAILS);
  r_out.alloc((size_t)M * 4 + 16, st); t_out.alloc((size_t)M + 16, st);
  shard_rank_owner(c, rk.as<u32>(), rc.as<u32>(), M, (u32)rank_base, r_out.as<u32>(), t_out.as<u8>());
  ph.mark(PH_RANK_OWNER);
  back.alloc((size_t)M * 8 + 16, st); backr.alloc((size_t)Tn * 8 + 16, st);
  rs.alloc((size_t)Tn * 4 + 16, st); ts.alloc((size_t)Tn + 16, st);
  const bool rt32 = tails_total < 0x7FFFFFF0ull;
  if (M) {
    if (rt32) k_shx_pack_rt32<<<(M + B - 1) / B, B, 0, st>>>(r_out.as<u32>(), t_out.as<u8>(), M, back.as<u32>());
    else k_shx_pack_rt<<<(M + B - 1) / B, B, 0, st>>>(r_out.as<u32>(), t_out.as<u8>(), M, back.as<u64>());
  }
  cm.group_begin();
  cm.a2a(back.p, recv, backr.p, send, rt32 ? 4 : 8);
  cm.group_end();
  if (Tn) {
    if (rt32) k_shx_unpack_rt32<<<(Tn + B - 1) / B, B, 0, st>>>(backr.as<u32>(), Tn, rs.as<u32>(), ts.as<u8>());
    else k_shx_unpack_rt<<<(Tn + B - 1) / B, B, 0, st>>>(backr.as<u64>(), Tn, rs.as<u32>(), ts.as<u8>());
  }
  c->launches += 2;
  CK(cudaStreamSynchronize(st));
  ph.mark(PH_X_RANKS);
  kc.release(); got.release(); rk.release(); rc.release(); r_out.release(); t_out.release(); back.release(); backr.release();
  u32 m = shard_rank_apply(c, rs.as<u32>(), ts.as<u8>());
  rs.release(); ts.release();
  // refinement rounds on the all-gathered (rank, rank[next^(2^h)]) keys of the tied subset (:1480-1494)
  u32 rounds = 0;
  std::vector<u64> ms = cm.gather_u64({(u64)m});     // later rounds: every rank derives the counts itself
  for (;;) {
    const u64 mt = sum(ms);
    if (mt == 0) break;
    if (mt >= 0xFFFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "more than 2^32 tied tails"};
    if (m) shard_tie_keys(c);
    DBuf allk;
    allk.alloc((size_t)mt * 8 + 16, st);
    cm.group_begin();
    cm.a2a_same(c->sh_tiekeys.p, m, allk.p, ms, 8);
    cm.group_end();
    CK(cudaStreamSynchronize(st));
    std::vector<u64> off(W + 1, 0), nxt(W, 0);
    for (u32 s = 0; s < W; s++) off[s + 1] = off[s] + ms[s];
    m = shard_tie_refine(c, allk.as<u64>(), (u32)mt, (u32)off[R], off.data(), W, nxt.data());
    if (nxt[R] != m) throw CmsErr{CMSBWT_E_INTERNAL, "tie refinement: slab count disagrees with the local subset"};
    ms = nxt;
    rounds++;
  }
  c->stats.rank_rounds = rounds;
  ph.mark(PH_REFINE);

  // ---- stage 2b: tile entries to the owners of the reference-position ranges (:1387-1416, :368-390)
  {
    Timer t(st);
    c->s_label.release(); c->s_next.release(); c->s_tid.release(); c->s_key32.release();
    shard_tiles_local(c, W);                      // the tile scatter forms kappa from the tail ranks itself
    c->stats.sort_ms += t.stop();
  }
  const u32 tp = c->sh_tp;
  ph.mark(PH_TILES_LOCAL);
  std::vector<u64> hsend(W), rsend(W), hrecv(W), rrecv(W), eq(W, (u64)tp + 1);
  {
    DBuf ed;
    ed.alloc((size_t)(2 * (W + 1)) * 4, st);
    k_shx_edges<<<(W + 1 + 63) / 64, 64, 0, st>>>(c->t_hoff.as<u32>(), c->t_roff.as<u32>(), tp, W, ed.as<u32>());
    LAUNCH(c);
    std::vector<u32> he(2 * (W + 1));
    CK(cudaMemcpyAsync(he.data(), ed.p, he.size() * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    std::vector<u64> mine(2 * W);
    for (u32 d = 0; d < W; d++) { hsend[d] = he[d + 1] - he[d]; rsend[d] = he[W + 1 + d + 1] - he[W + 1 + d]; mine[d] = hsend[d]; mine[W + d] = rsend[d]; }
    std::vector<u64> all = cm.gather_u64(mine);
    for (u32 s = 0; s < W; s++) { hrecv[s] = all[(size_t)s * 2 * W + R]; rrecv[s] = all[(size_t)s * 2 * W + W + R]; }
  }
  DBuf he, re, hoffs, roffs;
  he.alloc((sum(hrecv) + 1) * 16, st); re.alloc((sum(rrecv) + 1) * 16, st);
  hoffs.alloc((size_t)W * (tp + 1) * 4, st); roffs.alloc((size_t)W * (tp + 1) * 4, st);
  cm.group_begin();
  cm.a2a(c->t_he.p, hsend, he.p, hrecv, 16);
  cm.a2a(c->t_re.p, rsend, re.p, rrecv, 16);
  // per destination its slice of tp + 1 offsets: slices overlap by one entry, so they are sent from
  // where they lie (t_hoff + d * tp) instead of as slabs
  for (u32 p = 0; p < W; p++) {
    NK(nccl().Send(c->t_hoff.as<u32>() + (size_t)p * tp, (size_t)(tp + 1) * 4, ncclUint8, (int)p, cm.comm, st));
    NK(nccl().Recv(hoffs.as<u32>() + (size_t)p * (tp + 1), (size_t)(tp + 1) * 4, ncclUint8, (int)p, cm.comm, st));
    NK(nccl().Send(c->t_roff.as<u32>() + (size_t)p * tp, (size_t)(tp + 1) * 4, ncclUint8, (int)p, cm.comm, st));
    NK(nccl().Recv(roffs.as<u32>() + (size_t)p * (tp + 1), (size_t)(tp + 1) * 4, ncclUint8, (int)p, cm.comm, st));
  }
  cm.group_end();
  CK(cudaStreamSynchronize(st));
  ph.mark(PH_X_TILES);
  std::vector<u64> he_slab(W), re_slab(W);
  { u64 a = 0, b = 0; for (u32 s = 0; s < W; s++) { he_slab[s] = a; re_slab[s] = b; a += hrecv[s]; b += rrecv[s]; } }
  {
    Timer t(st);
    shard_tiles_merge(c, W, R, he.p, re.p, hoffs.as<u32>(), roffs.as<u32>(), he_slab.data(), re_slab.data(), ndocs_total,
                      false);
    c->stats.sort_ms += t.stop();
  }
  he.release(); re.release(); hoffs.release(); roffs.release();
  ph.mark(PH_TILES_MERGE);

  // ---- stage 3, pull emit by SA-rank range (:1618-1694): within(h) and the bucket records where the tiles
  // are, records + head lists to the owners of the SA ranks, offsets / fill / run-length records local.
  // One count exchange decides the path for everybody: a device whose tiles are deeper than
  // k_tile_rank holds sends everyone to the byte-range emit with the sort-based counting.
  bool pulled = false;
  {
    const int want = getenv("CMSBWT_SHARD_EMIT") ? atoi(getenv("CMSBWT_SHARD_EMIT")) : 1;   // 0: byte-range emit (A/B)
    bool mine = false;
    {
      Timer t(st);
      if (want && ndocs_total + c->ix.n < 0xFFFFFFF0ull) mine = shard_rank_pull(c, W, R);
      c->stats.sort_ms += t.stop();
    }
    std::vector<u64> my(1 + 3 * (size_t)W, 0);
    my[0] = mine ? 1 : 0;
    if (mine)
      for (u32 d = 0; d < W; d++) {
        my[1 + d] = c->sh_slot_edges[d + 1] - c->sh_slot_edges[d];
        my[1 + W + d] = c->sh_v_hwcnt[d];
        my[1 + 2 * W + d] = c->sh_v_bytes[d];
      }
    const size_t per = my.size();
    std::vector<u64> all = cm.gather_u64(my);
    pulled = true;
    for (u32 s = 0; s < W; s++) pulled = pulled && all[s * per] != 0;
    ph.mark(PH_PLACE);
    if (pulled) {
      std::vector<u64> qsend(W), qrecv(W), hsend2(W), hrecv2(W), rbase(W + 1, 0), hbase(W + 1, 0), olen(W, 0);
      for (u32 d = 0; d < W; d++) {
        qsend[d] = my[1 + d]; hsend2[d] = my[1 + W + d];
        for (u32 s2 = 0; s2 < W; s2++) olen[d] += all[s2 * per + 1 + 2 * W + d];
      }
      olen[0] += ndocs_total;                          // the terminators (one byte per document) lead the output
      for (u32 s2 = 0; s2 < W; s2++) {
        qrecv[s2] = all[s2 * per + 1 + R]; hrecv2[s2] = all[s2 * per + 1 + W + R];
        rbase[s2 + 1] = rbase[s2] + qrecv[s2]; hbase[s2 + 1] = hbase[s2] + hrecv2[s2];
      }
      if (hbase[W] >= 0xFFFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "more than 2^32 heads in one device's SA-rank range"};
      u64 tot = 0, obeg = 0;
      for (u32 d = 0; d < W; d++) { if (d < R) obeg += olen[d]; tot += olen[d]; }
      if (tot != N_total) {
        char buf[128];
        snprintf(buf, sizeof buf, "coverage total %llu != N %llu", (unsigned long long)tot, (unsigned long long)N_total);
        throw CmsErr{CMSBWT_E_INTERNAL, buf};
      }
      c->t_he.release(); c->t_re.release(); c->t_hoff.release(); c->t_roff.release();
      // terminator symbols of this rank's documents (document order) -> rank 0
      DBuf term_mine, term_all, vrecv, hwrecv;
      term_mine.alloc(ndocs_local + 16, st);
      shard_term_syms(c, term_mine.as<u8>());
      if (R == 0) term_all.alloc(ndocs_total + 16, st);
      vrecv.alloc((rbase[W] + 1) * 16, st); hwrecv.alloc((hbase[W] + 1) * 8, st);
      cm.group_begin();
      cm.a2a(c->sh_vsend.p, qsend, vrecv.p, qrecv, 16);
      cm.a2a(c->sh_hwsend.p, hsend2, hwrecv.p, hrecv2, 8);
      if (ndocs_local) NK(nccl().Send(term_mine.p, ndocs_local, ncclUint8, 0, cm.comm, st));
      if (R == 0) {
        u64 off = 0;
        for (u32 s2 = 0; s2 < W; s2++) {
          const u64 nds = plan[2 * s2 + 1];
          if (nds) NK(nccl().Recv(term_all.as<u8>() + off, nds, ncclUint8, (int)s2, cm.comm, st));
          off += nds;
        }
      }
      cm.group_end();
      CK(cudaStreamSynchronize(st));
      c->sh_vsend.release(); c->sh_hwsend.release();
      ph.mark(PH_X_VREC);
      {
        Timer t(st);
        shard_emit_pull(c, mode, W, R, (u32)ndocs_total, vrecv.p, rbase[W], rbase.data(), hbase.data(), hwrecv.p,
                        term_all.as<u8>(), olen[R]);
        c->stats.emit_ms = t.stop();
      }
      c->sh_obeg = obeg;
      c->sh_Ntotal = N_total;
      ph.mark(PH_EMIT);
    } else {
      c->sh_vsend.release(); c->sh_hwsend.release();
    }
  }
  if (!pulled) {
    shard_tiles_cov(c, ndocs_total);
    DBuf cov_full;
    cov_full.alloc((size_t)tp * 256 * 4 * W, st);
    cm.all_gather(c->sh_cov.p, cov_full.p, (size_t)tp * 256 * 4);
    CK(cudaStreamSynchronize(st));
    ph.mark(PH_X_COV);

    // ---- stage 3: own output range + placements to the owners of the output ranges (:1618-1694)
    {
      Timer t(st);
      shard_emit(c, cov_full.as<u32>(), N_total, W, R, doc_base);
      c->stats.emit_ms = t.stop();
    }
    cov_full.release();
    ph.mark(PH_EMIT);
    std::vector<u64> psend(W), precv(W);
    for (u32 d = 0; d < W; d++) psend[d] = c->sh_plc_cnt[d];
    {
      std::vector<u64> all = cm.gather_u64(psend);
      for (u32 s = 0; s < W; s++) precv[s] = all[(size_t)s * W + R];
    }
    DBuf plc;
    plc.alloc((sum(precv) + 1) * 8, st);
    cm.group_begin();
    cm.a2a(c->sh_plc.p, psend, plc.p, precv, 8);
    cm.group_end();
    CK(cudaStreamSynchronize(st));
    ph.mark(PH_X_PLC);
    {
      Timer t(st);
      shard_apply(c, plc.as<u64>(), sum(precv));
      c->stats.emit_ms += t.stop();
    }
    ph.mark(PH_APPLY);
  }
  c->sh_rle_first = 0;
  c->sh_rle_skip = 0;
  if (mode == CMSBWT_MODE_RLE) {
    // -r: every rank encodes its slice; a run that straddles rank borders belongs to the leftmost rank
    // it touches (the concatenated records are then the maximal-run RLE of the whole BWT, :1695-1777)
    if (!pulled) rle_encode(c, c->d_out.as<u8>(), c->sh_olen);       // the pull emit wrote this slice's records itself
    const u64 nrec = c->stats.n_runs;
    u8 edge[18] = {0};
    if (nrec) {
      CK(cudaMemcpyAsync(edge, c->d_rle.p, 9, cudaMemcpyDeviceToHost, st));
      CK(cudaMemcpyAsync(edge + 9, c->d_rle.as<u8>() + 9 * (nrec - 1), 9, cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
    }
    u64 frun = 0, lrun = 0;
    memcpy(&frun, edge, 8); memcpy(&lrun, edge + 9, 8);
    std::vector<u64> info = cm.gather_u64({nrec, frun, (u64)edge[8], lrun, (u64)edge[17]});
    std::vector<char> drop(W, 0);
    std::vector<u64> add(W, 0);
    long owner = -1; u64 osym = 0;
    for (u32 r = 0; r < W; r++) {
      const u64 Rr = info[5 * r], fr = info[5 * r + 1], fs = info[5 * r + 2], ls = info[5 * r + 4];
      if (Rr == 0) continue;
      if (owner >= 0 && fs == osym) {
        drop[r] = 1; add[owner] += fr;
        if (Rr == 1) continue;                 // the whole slice went into the open run; it stays open
      }
      owner = r; osym = ls;
    }
    u64 first = 0;
    for (u32 r = 0; r < R; r++) first += info[5 * r] - (drop[r] ? 1 : 0);
    c->sh_rle_first = first;
    c->sh_rle_skip = drop[R] ? 1 : 0;
    c->rle_off = 9 * c->sh_rle_skip;
    c->out_bytes = 9 * (nrec - c->sh_rle_skip);
    if (add[R]) {
      const u64 nl = lrun + add[R];
      CK(cudaMemcpyAsync(c->d_rle.as<u8>() + 9 * (nrec - 1), &nl, 8, cudaMemcpyHostToDevice, st));
      CK(cudaStreamSynchronize(st));
    }
    ph.mark(PH_RLE);
  }
  c->out_mode = mode;
  if (mode == CMSBWT_MODE_PLAIN) c->out_bytes = c->sh_olen;
  group_collect(c);
}

extern "C" {

int cmsbwt_nccl_unique_id(void *out128) {
  if (!out128) return CMSBWT_E_ARG;
  try {
    ncclUniqueId id;
    if (nccl().GetUniqueId(&id) != ncclSuccess) return CMSBWT_E_CUDA;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    memcpy(out128, &id, 128);
  } catch (...) { return CMSBWT_E_CUDA; }
  return CMSBWT_OK;
}

int cmsbwt_comm_init(cmsbwt_ctx *c, const void *id128, uint32_t rank, uint32_t world) {
  if (!c || !id128 || world == 0 || rank >= world || world > 64) return CMSBWT_E_ARG;
  try {
    CK(cudaSetDevice(c->dev));
    if (c->comm) { nccl().CommDestroy((ncclComm_t)c->comm); c->comm = nullptr; }
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    ncclComm_t comm;
    NK(nccl().CommInitRank(&comm, (int)world, id, (int)rank));
    c->comm = comm;
    c->comm_rank = rank;
    c->comm_world = world;
  } catch (const CudaErr &e) { c->err = std::string(e.what) + ": " + cudaGetErrorString(e.e); cudaGetLastError(); return CMSBWT_E_CUDA; }
  catch (const CmsErr &e) { c->err = e.msg; return e.code; }
  return CMSBWT_OK;
}

int cmsbwt_comm_destroy(cmsbwt_ctx *c) {
  if (!c) return CMSBWT_E_ARG;
  if (c->comm) {
    cudaSetDevice(c->dev);
    cudaStreamSynchronize(c->st);
    try { nccl().CommDestroy((ncclComm_t)c->comm); } catch (...) {}
    c->comm = nullptr;
  }
  return CMSBWT_OK;
}

int cmsbwt_sharded_after_ms(cmsbwt_ctx *c, uint64_t ndocs_local, int mode, uint64_t *out_begin, uint64_t *out_len,
                            uint64_t *total_bytes) {
  if (!c || !c->comm) { if (c) c->err = "cmsbwt_comm_init first"; return CMSBWT_E_ARG; }
  if (!c->have_ms) { c->err = "ms_run first"; return CMSBWT_E_ARG; }
  if (mode != CMSBWT_MODE_PLAIN && mode != CMSBWT_MODE_RLE) { c->err = "bad mode"; return CMSBWT_E_ARG; }
  try {
    CK(cudaSetDevice(c->dev));
    current_cache() = &c->cache;
    run_sharded_impl(c, c->N, ndocs_local, mode);
    if (out_begin) *out_begin = c->sh_obeg;
    if (out_len) *out_len = c->sh_olen;
    if (total_bytes) *total_bytes = c->sh_Ntotal;
  } catch (const CudaErr &e) { c->err = std::string(e.what) + ": " + cudaGetErrorString(e.e); cudaGetLastError(); return CMSBWT_E_CUDA; }
  catch (const CmsErr &e) { c->err = e.msg; return e.code; }
  catch (const std::exception &e) { c->err = e.what(); return CMSBWT_E_NOMEM; }
  return CMSBWT_OK;
}

int cmsbwt_run_sharded(cmsbwt_ctx *c, const void *d_text, uint64_t N_local, uint64_t ndocs_local, int mode,
                       uint64_t *out_begin, uint64_t *out_len, uint64_t *total_bytes) {
  if (!c || !c->comm) { if (c) c->err = "cmsbwt_comm_init first"; return CMSBWT_E_ARG; }
  int rc = cmsbwt_ms_run_device(c, d_text, N_local);
  if (rc) return rc;
  return cmsbwt_sharded_after_ms(c, ndocs_local, mode, out_begin, out_len, total_bytes);
}

int cmsbwt_sharded_rle(cmsbwt_ctx *c, const void **d_records, uint64_t *n_records, uint64_t *first_record) {
  if (!c || !d_records || !n_records) return CMSBWT_E_ARG;
  if (c->out_mode != CMSBWT_MODE_RLE || !c->d_rle.p) { c->err = "run_sharded in RLE mode first"; return CMSBWT_E_ARG; }
  const u64 skip = c->sh_rle_skip;
  *d_records = c->d_rle.as<u8>() + 9 * skip;
  *n_records = c->stats.n_runs - skip;
  if (first_record) *first_record = c->sh_rle_first;
  return CMSBWT_OK;
}

int cmsbwt_sharded_stats(const cmsbwt_ctx *c, uint64_t *collectives, uint64_t *exchange_bytes, double *phase_ms,
                         uint32_t phase_cap) {
  if (!c) return CMSBWT_E_ARG;
  if (collectives) *collectives = c->sh_collectives;
  if (exchange_bytes) *exchange_bytes = c->sh_comm_bytes;
  if (phase_ms)
    for (uint32_t k = 0; k < phase_cap && k < (uint32_t)PH_COUNT; k++) phase_ms[k] = c->sh_phase_ms[k];
  return PH_COUNT;
}

}  // extern "C"
