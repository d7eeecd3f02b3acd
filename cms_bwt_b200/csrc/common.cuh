// Shared declarations of the cmsbwt_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>
#include <utility>
#include "../../include/cmsbwt_b200.h"

typedef uint8_t u8;
typedef uint16_t u16;
typedef uint32_t u32;
typedef uint64_t u64;
typedef int32_t i32;
typedef int64_t i64;

#define CMS_PAD 64          // readable bytes past the end of x / text
#define CMS_TILE_W 64       // reference positions per counting tile (one 64-bit mask)
#define CMS_CAP 32          // per-lane LCE budget (bytes) before a probe is handed to the warp

struct CudaErr { cudaError_t e; const char *what; };

#define CK(call)                                                                      \
  do {                                                                                \
    cudaError_t _e = (call);                                                          \
    if (_e != cudaSuccess) throw CudaErr{_e, #call};                                  \
  } while (0)

struct CmsErr { int code; std::string msg; };

// Caching device allocator, one per context: blocks are recycled by size class, so a repeated job
// of the same shape performs no cudaMalloc / cudaFree at all.  All work of a context runs on one
// stream, so handing a released block to a later kernel is ordered correctly; every API entry
// makes its context's cache the calling thread's current one (contexts may be driven from
// different host threads).
struct DevCache {
  std::vector<std::pair<size_t, void *>> free_blocks;
  size_t held = 0;
  static size_t round(size_t b) {
    if (b < 512) return 512;
    if (b < (1u << 20)) return (b + 511) & ~(size_t)511;
    return (b + (1u << 20) - 1) & ~(size_t)((1u << 20) - 1);
  }
  void *get(size_t b) {
    b = round(b);
    int best = -1;
    for (size_t i = 0; i < free_blocks.size(); i++)
      if (free_blocks[i].first >= b && free_blocks[i].first <= b + b / 8 &&
          (best < 0 || free_blocks[i].first < free_blocks[best].first)) best = (int)i;
    if (best >= 0) {
      void *p = free_blocks[best].second;
      free_blocks.erase(free_blocks.begin() + best);
      return p;
    }
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, b);
    if (e != cudaSuccess) {          // give cached blocks back to the driver and retry once
      cudaGetLastError();
      trim();
      e = cudaMalloc(&p, b);
      if (e != cudaSuccess) throw CudaErr{e, "cudaMalloc"};
    }
    held += b;
    return p;
  }
  void put(void *p, size_t b) { free_blocks.push_back({round(b), p}); }
  void trim() {
    if (free_blocks.empty()) return;
    cudaDeviceSynchronize();
    for (auto &fb : free_blocks) { cudaFree(fb.second); held -= fb.first; }
    free_blocks.clear();
  }
  ~DevCache() { trim(); }
};
DevCache *&current_cache();        // thread-local

struct DBuf {
  void *p = nullptr;
  size_t bytes = 0;
  DevCache *owner = nullptr;
  void alloc(size_t b, cudaStream_t) {
    release();
    bytes = b;
    owner = current_cache();
    p = owner->get(b ? b : 16);
  }
  void release() {
    if (p) owner->put(p, bytes ? bytes : 16);
    p = nullptr;
    bytes = 0;
  }
  void swap(DBuf &o) { std::swap(p, o.p); std::swap(bytes, o.bytes); std::swap(owner, o.owner); }
  template <class T> T *as() const { return (T *)p; }
  ~DBuf() { release(); }
  DBuf() {}
  DBuf(const DBuf &) = delete;
  DBuf &operator=(const DBuf &) = delete;
};

struct Index {             // device-resident reference index
  u32 n = 0;
  int k = 0;               // k-mer table order
  const u8 *x = nullptr;   // augmented reference, CMS_PAD bytes of 0xFF after it
  const i32 *sa = nullptr;
  const u32 *isa = nullptr;
  const u32 *G = nullptr;  // G[q] = q + PLCP'[q]
  const u8 *rbwt = nullptr;
  const u32 *ktab = nullptr;   // 4^k + 1 entries: #suffixes < k-mer m
};

struct cmsbwt_ctx {
  DevCache cache;            // first member: destroyed after every DBuf below has been returned to it
  int dev = 0;
  cudaStream_t st = 0;
  std::string err;
  cmsbwt_stats stats{};
  u64 launches = 0;
  // index
  Index ix;
  DBuf d_x, d_sa, d_isa, d_G, d_rbwt, d_ktab;
  bool have_sa = false, have_index = false;
  // fast path (pure-ACGT reference): packed reference, {key, sa} per rank
  bool pure = false;         // x = A/C/G/T segments between a bounded number of runs of other symbols (the last run
                             // is the augmentation); n_acgt = start of the last run (all there is for a pure reference)
  u32 n_acgt = 0, nruns = 0;
  DBuf d_xruns, d_xbk;       // the runs (start | end << 32) and, per 64 positions, the first run ending after them
  DBuf d_sak, d_slots;       // {key, sa} per rank; k-mer slot table (64 B per k-mer)
  u64 *xp = nullptr;         // packed reference, inside d_ktab's allocation
  size_t hot_bytes = 0;      // k-mer table + packed reference (L2 persistence window)
  i64 l2_pin = 0;
  i64 l2_persist_max = 0, l2_window_max = 0;   // device limits (bytes), read at ctx_create
  DBuf d_tp, d_seps;         // packed text, sorted separator positions (+ sentinel)
  u64 sep_hint = 0, head_hint = 0;
  // text
  const u8 *text = nullptr;
  u64 N = 0;
  DBuf d_text;
  // streamed ingest / egress (cmsbwt_text_*, cmsbwt_result_read): a second stream for the copies,
  // one event per host staging slot
  cudaStream_t st_copy = 0;
  cudaEvent_t ev_slot[CMSBWT_COPY_SLOTS] = {};
  cudaEvent_t ev_in0 = 0, ev_in1 = 0;
  u64 staged = 0, staged_cap = 0;
  bool staging = false;
  // stage 1 products (text-order heads)
  u64 H = 0;
  DBuf h_gstart, h_pos, h_len, h_cs;     // u64, u32, u32, u16 (ch | sm << 8)
  bool dump = false;
  DBuf dump_pos, dump_len, dump_sm, dump_head;
  // stage 2 products
  DBuf h_toNext, h_within, h_off, d_cov;  // u32, u32, u64, i32[n+1]
  bool have_ms = false, have_sort = false;
  // stage 2, brute-force tile path (tiles.cu)
  bool light = false;
  DBuf t_hoff, t_roff, t_he, t_re, h_sexcl;
  u32 tile_maxent = 0;
  // pull emit (single GPU, light path): per position {bucket size, first head slot}; heads grouped by
  // position as (within << 8 | symbol); terminator symbols in document order
  DBuf d_pcov, d_hw, d_termsym;
  bool termsym_done = false;  // d_termsym was filled by the tile scatter
  u64 n_hw = 0;
  // stage 2 working set (kept in the context so that the multi-GPU shard calls can interleave
  // collectives between the phases)
  DBuf s_tincl, s_sexcl, s_label, s_next, s_tid, s_key32, s_trank, s_kappa, s_nitems;
  u32 s_Tn = 0;
  int s_key_bits = 32;
  // shard state (shard.cu / tailrank.cu): locally sorted tail keys, tied subset, merged tiles
  DBuf sh_ks, sh_cs, sh_perm, sh_sub, sh_tiekeys, sh_cov, sh_plc, sh_base;
  // pull emit across devices: static routing table of this device's reference positions (slot in the send
  // buffer, grouped by the owner of the position's SA rank), the records / head lists to send
  DBuf sh_slot, sh_vsend, sh_hwsend;
  u64 sh_slot_key[5] = {0, 0, 0, 0, 0};
  std::vector<u64> sh_slot_edges, sh_v_hwcnt, sh_v_bytes;
  u32 sh_rper = 0;                               // SA ranks per device
  u32 sh_m = 0, sh_round = 0;
  std::vector<u64> sh_plc_cnt;                   // placements per destination rank
  u32 sh_world = 1, sh_rank = 0, sh_tp = 0;      // tiles per rank
  u64 sh_he = 0, sh_re = 0, sh_nplc = 0, sh_obeg = 0, sh_olen = 0, sh_chunk = 0, sh_Ntotal = 0;
  // multi-GPU host layer (sharded.cu): NCCL communicator of this rank, phase wall times, counters
  void *comm = nullptr;
  u32 comm_rank = 0, comm_world = 1;
  double sh_phase_ms[16] = {};
  u64 sh_collectives = 0, sh_comm_bytes = 0, sh_rle_first = 0, sh_rle_skip = 0;
  void *fm = nullptr;        // FM-index built from the last plain result (fmindex.cu)
  // stage 3
  DBuf d_out, d_rle;
  u64 out_bytes = 0, rle_off = 0;   // rle_off: bytes of d_rle that belong to the rank on the left (sharded -r)
  int out_mode = 0;
  // kernel-group timers: pairs of events per span, collected after the stage
  std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> spans;
  std::vector<cudaEvent_t> ev_pool;
  // params
  i64 chunk_len = 0;         // text window per lane; 0 = by text length (ms_fast.cu), 8192 in k_ms
  i64 pool_div = 6;          // initial head-pool capacity = N / pool_div records
  i64 kmer_k = 0;            // 0 = auto
  i64 skip_min = 20;         // fast path: try the G skip for matches at least this long
  i64 fast = 1;              // 0 = always use the general 8-bit kernel
  i64 ms_occ = 8;            // fast path: resident CTAs per SM the MS kernel is compiled for (8, 10, 12)
  i64 ms_variant = 26;       // MsfParams::variant: 2 text word cache, 8 G prefetch, 16 empty-bucket slots (k_ms_fast)
  i64 use_slots = 1;         // fast path: k-mer slot table (one 64-byte read per search) on / off
  i64 nb_t[3] = {48, 160, 512};   // tile placement: most ranges per tile for 1 / 4 / 8 kappa buckets
  i64 scatter_mb = 64;       // random 4-byte scatters run in passes over target regions of this many MB (L2)
  i64 tie_local = 1;          // tail ranking: small tied groups ordered by cls inside the group (0: by the 64-bit sort)
  i64 tile_short = 1;         // k_tile_rank: short near ranges filed per position (0: walked with the near lists)
  i64 tile_margin = 48;       // k_tile_rank: heads / ranges with a match end below tile end + margin are "near"
  i64 tile_budget = 1ll << 36; // stage 2: brute-force tile counting up to this many pair tests per job
};

#define LAUNCH(ctx) ((ctx)->launches++)

struct Timer {
  cudaEvent_t a, b;
  cudaStream_t st;
  Timer(cudaStream_t s) : st(s) { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, s); }
  double stop() {
    cudaEventRecord(b, st);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    return ms;
  }
  ~Timer() { cudaEventDestroy(a); cudaEventDestroy(b); }
};


// Kernel-group timers (stats.group_ms): events on the context's stream around a group of launches;
// read back by group_collect() once the stage has synchronised.  A group may be opened several
// times per job (its spans add up).
struct GroupSpan {
  cmsbwt_ctx *c; int id;
  GroupSpan(cmsbwt_ctx *c_, int id_);
  ~GroupSpan();
};
void group_reset(cmsbwt_ctx *c);
void group_collect(cmsbwt_ctx *c);

// stage entry points (each throws CudaErr / CmsErr)
void sa_build_device(cmsbwt_ctx *c, const u8 *d_x, u32 n, i32 *d_sa);
void index_build(cmsbwt_ctx *c);
void ms_stage(cmsbwt_ctx *c);
bool ms_stage_fast(cmsbwt_ctx *c);
void pack_2bit(cmsbwt_ctx *c, const u8 *src, u64 len, u64 *dst, u64 nwords, u64 *seps, u64 sepcap,
               unsigned long long *nsep, u32 *flags);
void sort_stage(cmsbwt_ctx *c);
void sort_prepare(cmsbwt_ctx *c, u64 doc_base);
void sort_kappa(cmsbwt_ctx *c);
void sweep_count(cmsbwt_ctx *c, u64 I, DBuf &it_tr, DBuf &it_key, DBuf &it_pay, const u32 *pos, u32 ntiles,
                 int tbits, int cshift, u32 *within);
// bits that hold cls = smaller ? E : ~E order-preservingly (E = pos + len <= n): the low bits(n) + 1
static inline int sweep_cls_bits(u32 n) { int b = 1; while (b < 32 && (1ull << b) <= (u64)n) b++; return b + 1 > 32 ? 32 : b + 1; }
void tail_rank(cmsbwt_ctx *c, const u64 *label, DBuf &key32, DBuf &id, int key_bits, u32 *nexth, u32 Tn,
               u32 *trank);
void tail_rank_ex(cmsbwt_ctx *c, const u64 *label, const u32 *cls32, DBuf &key32, DBuf &id, int key_bits,
                  u32 *nexth, u32 Tn, u32 *trank, u32 rank_base, bool cls_only, u8 *tie_by_id);
bool tiles_build(cmsbwt_ctx *c, const u64 *kappa, DBuf &sexcl);
// multi-GPU shard steps (tailrank.cu, tiles.cu, emit.cu); the collectives between them are the
// host layer's (cms_bwt_b200/multi.py)
void shard_tails_sort(cmsbwt_ctx *c);
void shard_rank_owner(cmsbwt_ctx *c, const u32 *keys, const u32 *cls, u32 M, u32 rank_base, u32 *rank_out,
                      u8 *tie_out);
u32 shard_rank_apply(cmsbwt_ctx *c, const u32 *rk_sorted, const u8 *tie_sorted);
void shard_tie_keys(cmsbwt_ctx *c);
u32 shard_tie_refine(cmsbwt_ctx *c, const u64 *allkeys, u32 m_total, u32 my_off, const u64 *slab_off = nullptr,
                     u32 world = 0, u64 *next_counts = nullptr);
void shard_tiles_local(cmsbwt_ctx *c, u32 world);
u32 shard_tiles_merge(cmsbwt_ctx *c, u32 world, u32 rank, const void *he, const void *re, const u32 *hoffs,
                      const u32 *roffs, const u64 *he_slab, const u64 *re_slab, u64 ndocs_total, bool with_cov);
void shard_tiles_cov(cmsbwt_ctx *c, u64 ndocs_total);
// multi-GPU pull emit (tiles.cu / emit.cu): bucket records routed to the owners of the SA-rank ranges
bool shard_rank_pull(cmsbwt_ctx *c, u32 world, u32 rank);
void shard_emit_pull(cmsbwt_ctx *c, int mode, u32 world, u32 rank, u32 nd_total, const void *recs, u64 nrecs,
                     const u64 *rec_base, const u64 *hw_base, const void *hw, const u8 *termsym, u64 expect_bytes);
void shard_emit(cmsbwt_ctx *c, const u32 *cov_full, u64 N_total, u32 world, u32 rank, u64 doc_base);
void shard_apply(cmsbwt_ctx *c, const u64 *plc, u64 cnt);
void rle_encode(cmsbwt_ctx *c, const u8 *out, u64 N);
void tiles_count_heads(cmsbwt_ctx *c);
void fm_release(cmsbwt_ctx *c);
// product-owned LSD radix sort (rsort.cu): k / v = input and sorted output, k2 / v2 = scratch (swapped)
void rsort_pairs32(cmsbwt_ctx *c, DBuf &k, DBuf &k2, DBuf &v, DBuf &v2, u64 n, int begin_bit, int end_bit);
void rsort_pairs64(cmsbwt_ctx *c, DBuf &k, DBuf &k2, DBuf &v, DBuf &v2, u64 n, int begin_bit, int end_bit);
void rsort_keys64(cmsbwt_ctx *c, DBuf &k, DBuf &k2, u64 n, int begin_bit, int end_bit);
void emit_stage(cmsbwt_ctx *c, int mode);

// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__
// unaligned little-endian 8-byte load; the buffer base must be 8-byte aligned and padded.
__device__ __forceinline__ u64 ld8u(const u8 *p) {
  uintptr_t a = (uintptr_t)p;
  unsigned sh = (unsigned)(a & 7) * 8;
  const u64 *q = (const u64 *)(a & ~(uintptr_t)7);
  u64 lo = __ldg(q);
  if (sh == 0) return lo;
  u64 hi = __ldg(q + 1);
  return (lo >> sh) | (hi << (64 - sh));
}
// exclusive prefix sum of u32 counts into u64 offsets (64-bit accumulator)
#define EXSUM64(tmp_ptr, tmp_bytes, in, out, num, stream) \
  cub::DeviceScan::ExclusiveScan(tmp_ptr, tmp_bytes, in, out, cub::Sum(), (u64)0, num, stream)
// Bucket record of the pull emit, one uint4 per virtual rank (written by k_tile_rank, read by the fill and
// run-length kernels): x = bucket size; z = number of heads (15 bits) | VREC_INLINE | sym0 << 16 | sym1 << 24.
// A bucket with at most two heads carries them in the record (y = within of the first, w = within of
// the second): the fill kernel then never touches hw[] for it -- buckets are walked in SA order, their
// head lists lie in reference-position order, so each list costs a random sector (ncu r2 final: 5.2 GB
// read by k_fill_pull against 1.5 GB of sequential inputs).  Else y = first slot of the list in hw[].
#define VREC_INLINE 0x8000u
__device__ __forceinline__ u32 vrec_heads(const uint4 &a) { return a.z & 0x7FFFu; }
// head j of the bucket as (within << 8 | symbol)
__device__ __forceinline__ u64 vrec_head(const uint4 &a, const u64 *__restrict__ hw, u32 j) {
  if (a.z & VREC_INLINE) return j ? (((u64)a.w << 8) | (a.z >> 24)) : (((u64)a.y << 8) | ((a.z >> 16) & 0xFFu));
  return __ldg(hw + a.y + j);
}
__device__ __forceinline__ u32 cls_of(u32 pos, u32 len, u32 sm) {
  u32 E = pos + len;
  return sm ? E : 0xFFFFFFFFu - E;
}
#endif
