// Stage 2: total order of the tails (text suffixes at the match ends), the tie-breaker of the
// bucket-internal order.  Replaces naming + libsais_int on the name string + next-rank
// substitution (CMS-BWT-functions.cpp:1432-1494; small path :627-687).
//
// A tail t carries label(t) = (ISA[reference position it is assigned to], cls of its phrase) and
// next(t) = the tail of that phrase; tails compare as the sequences label(t), label(next t), ...
// (heads.cu).  Segmented LSD radix sort keyed on the reference ISA rank, then refinement of the
// tied groups only (Larsson-Sadakane style prefix doubling on the subset):
//   1. one radix sort of all tails on a 32-bit key = ISA rank, then as many bits of the cls order
//      as fit (a weakly order-preserving quantisation of the phrase's remaining length; it
//      separates most tails that merely land on the same reference position);  rank = first
//      sorted index of the tail's group, so ranks of already-unique tails never change again;
//   2. the tails whose key is shared are sorted by (rank, cls);
//   3. the tails still tied are sorted by (rank, rank[next^(2^h) t]) for h = 0, 1, ... until no
//      group of two is left -- only they are touched, so a collection with a handful of shared
//      variants costs a handful of small sorts instead of log(chain) full ones.
#include "common.cuh"
#include <cub/cub.cuh>

struct Scratch {
  DBuf b;
  void need(size_t bytes, cudaStream_t st) { if (bytes > b.bytes) b.alloc(bytes, st); }
};

__global__ void k_tr_split(const u64 *label, u32 n, u32 *key32, u32 *id) {
  u32 t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  key32[t] = (u32)(label[t] >> 32);
  id[t] = t;
}
// group start index (for a max-scan) of sorted 32-bit keys
__global__ void k_tr_gs32(const u32 *key, u32 n, u32 *gs) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  gs[j] = (j == 0 || key[j] != key[j - 1]) ? j : 0;
}
// rank[id] = group start; tie flag = group has more than one member
// The scatter rank[id[j]] is a random 4-byte write per tail; done over the whole array it costs a
// DRAM read-modify-write of a 64-byte granule each.  It is therefore run in passes over id ranges
// [lo, hi) small enough for the written region to stay in L2 (the inputs are streamed past it).
__global__ void k_tr_rank0(const u32 *key, const u32 *id, const u32 *gs, u32 n, u32 base, u32 *rank, u8 *tie,
                           u32 lo, u32 hi, int first) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const u32 t = __ldcs(id + j), g = __ldcs(gs + j);
  if (t >= lo && t < hi) rank[t] = base + g;
  if (first) tie[j] = (g != j) || (j + 1 < n && __ldcs(key + j + 1) == __ldcs(key + j));
}
// The (rank, cls) round for SMALL groups, inside the group: tails that share a key sit next to each other
// after the radix sort (2.9 M of 48.8 M at config 3, in groups of 2-3), so the thread at the first
// position of a group orders its members by cls itself instead of sending them through an 8-pass 64-bit
// sort.  rank = group start + start of the member's equal-cls run; tie = the run has another member.
// Groups of more than TR_GRP_MAX members are left to the sort-based round below.
#define TR_GRP_MAX 16
__global__ void k_tr_group_cls(const u32 *__restrict__ gs, const u32 *__restrict__ ids, const u64 *__restrict__ label,
                               const u32 *__restrict__ cls32, u32 n, u32 base, u32 *rank, u8 *tie) {
  const u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n || !tie[j] || gs[j] != j) return;          // first position of a tied group only
  u32 g = 1;
  while (g <= TR_GRP_MAX && j + g < n && gs[j + g] == j) g++;
  if (g > TR_GRP_MAX) return;
  u32 cl[TR_GRP_MAX], at[TR_GRP_MAX];
  for (u32 s = 0; s < g; s++) {
    const u32 t = ids[j + s];
    const u32 v = cls32 ? cls32[t] : (u32)label[t];
    u32 k = s;                                          // insertion sort by cls, carrying the member's place
    while (k > 0 && cl[k - 1] > v) { cl[k] = cl[k - 1]; at[k] = at[k - 1]; k--; }
    cl[k] = v; at[k] = s;
  }
  u32 run = 0;
  for (u32 s = 0; s < g; s++) {
    if (s > 0 && cl[s] != cl[s - 1]) run = s;
    const bool shared = (s > 0 && cl[s - 1] == cl[s]) || (s + 1 < g && cl[s + 1] == cl[s]);
    rank[ids[j + at[s]]] = base + j + run;
    tie[j + at[s]] = shared ? 1 : 0;
  }
}
// subset keys: (current rank, secondary) with secondary = cls (round 0) or rank of the far tail
__global__ void k_tr_key_cls(const u32 *sub, u32 m, const u32 *rank, const u64 *label, const u32 *cls32,
                             u64 *key) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  u32 t = sub[j];
  key[j] = ((u64)rank[t] << 32) | (cls32 ? cls32[t] : (u32)label[t]);
}
__global__ void k_tr_mark(const u32 *sub, u32 m, u8 *flag) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < m) flag[sub[j]] = 1;
}
__global__ void k_tr_key_next(const u32 *sub, u32 m, const u32 *rank, const u32 *nexth, u64 *key) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  u32 t = sub[j];
  key[j] = ((u64)rank[t] << 32) | rank[nexth[t]];
}
__global__ void k_tr_gs64(const u64 *key, u32 m, u32 *grp, u32 *sg) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  bool g = j == 0 || (key[j] >> 32) != (key[j - 1] >> 32);
  bool s = j == 0 || key[j] != key[j - 1];
  grp[j] = g ? j : 0;
  sg[j] = s ? j : 0;
}
__global__ void k_tr_refine(const u64 *key, const u32 *id, const u32 *grp, const u32 *sg, u32 m, u32 *rank,
                            u8 *tie) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  rank[id[j]] = (u32)(key[j] >> 32) + (sg[j] - grp[j]);
  tie[j] = (sg[j] != j) || (j + 1 < m && key[j + 1] == key[j]);
}
// pointer doubling for the subset: tmp = nexth[nexth[t]], then written back
__global__ void k_tr_jump(const u32 *sub, u32 m, const u32 *nexth, u32 *tmp) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < m) tmp[j] = nexth[nexth[sub[j]]];
}
__global__ void k_tr_jump_store(const u32 *sub, u32 m, const u32 *tmp, u32 *nexth) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < m) nexth[sub[j]] = tmp[j];
}

static void max_scan(Scratch &tmp, u32 *a, u32 n, cudaStream_t st) {
  size_t tb = 0;
  cub::DeviceScan::InclusiveScan(nullptr, tb, a, a, cub::Max(), (int)n, st);
  tmp.need(tb, st);
  CK(cub::DeviceScan::InclusiveScan(tmp.b.p, tb, a, a, cub::Max(), (int)n, st));
}

// ids[j] with tie[j] != 0 -> out; returns the count (synchronises)
static u32 select_tied(Scratch &tmp, const u32 *ids, const u8 *tie, u32 n, u32 *out, DBuf &cnt, cudaStream_t st) {
  size_t tb = 0;
  cub::DeviceSelect::Flagged(nullptr, tb, ids, tie, out, cnt.as<u32>(), (int)n, st);
  tmp.need(tb, st);
  CK(cub::DeviceSelect::Flagged(tmp.b.p, tb, ids, tie, out, cnt.as<u32>(), (int)n, st));
  u32 m = 0;
  CK(cudaMemcpyAsync(&m, cnt.p, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return m;
}

// label[Tn], key32[Tn] (ISA rank + quantised cls order, consumed), id[Tn] = iota (consumed),
// nexth[Tn] (overwritten) -> trank[Tn]: distinct, order-preserving ranks
void tail_rank(cmsbwt_ctx *c, const u64 *label, DBuf &key32, DBuf &id, int key_bits, u32 *nexth, u32 Tn,
               u32 *trank) {
  tail_rank_ex(c, label, nullptr, key32, id, key_bits, nexth, Tn, trank, 0, false, nullptr);
}

// General form.  cls32 != NULL: the cls words come from there instead of the labels.  rank_base is
// added to every rank.  cls_only: stop after the (rank, cls) round (the owner side of the multi-GPU
// ranking); tie_by_id[t] (zeroed here) is then set for the tails that are still tied.
void tail_rank_ex(cmsbwt_ctx *c, const u64 *label, const u32 *cls32, DBuf &key32, DBuf &id, int key_bits,
                  u32 *nexth, u32 Tn, u32 *trank, u32 rank_base, bool cls_only, u8 *tie_by_id) {
  cudaStream_t st = c->st;
  const int B = 256;
  const u32 gT = (Tn + B - 1) / B;
  Scratch tmp;
  DBuf key32s, ids, gs, tie, cnt;
  key32s.alloc((size_t)Tn * 4, st);
  ids.alloc((size_t)Tn * 4, st); gs.alloc((size_t)Tn * 4, st); tie.alloc((size_t)Tn, st); cnt.alloc(4, st);
  const int bits = key_bits > 32 ? 32 : key_bits;
  size_t tb = 0;
  if (tie_by_id) CK(cudaMemsetAsync(tie_by_id, 0, Tn, st));
  rsort_pairs32(c, key32, key32s, id, ids, Tn, 0, bits);
  key32.swap(key32s); id.swap(ids);             // sorted: key32s / ids; key32 / id are scratch from here on
  k_tr_gs32<<<gT, B, 0, st>>>(key32s.as<u32>(), Tn, gs.as<u32>());
  max_scan(tmp, gs.as<u32>(), Tn, st);
  {
    // entries per pass: an L2-sized slice of the rank array, but never more than 8 passes over the
    // sorted keys (measured at 417 M tails: 26 passes 89 ms, 7 passes 81 ms for the whole stage)
    u64 region = (u64)(c->scatter_mb > 0 ? c->scatter_mb : 1 << 20) * (1u << 20) / 4;
    if (region < (Tn + 7) / 8) region = (Tn + 7) / 8;
    int first = 1;
    for (u64 lo = 0; lo < Tn; lo += region) {
      const u64 hi = lo + region < Tn ? lo + region : Tn;
      k_tr_rank0<<<gT, B, 0, st>>>(key32s.as<u32>(), ids.as<u32>(), gs.as<u32>(), Tn, rank_base, trank, tie.as<u8>(),
                                   (u32)lo, (u32)hi, first);
      first = 0;
      LAUNCH(c);
    }
  }
  c->launches += 8;
  if (c->tie_local) {
    k_tr_group_cls<<<gT, B, 0, st>>>(gs.as<u32>(), ids.as<u32>(), label, cls32, Tn, rank_base, trank, tie.as<u8>());
    LAUNCH(c);
  }
  // subset of tied tails (reuses id as the compacted list)
  u32 m = select_tied(tmp, ids.as<u32>(), tie.as<u8>(), Tn, id.as<u32>(), cnt, st);
  key32.release(); key32s.release(); gs.release();
  u32 rounds = 0;
  c->stats.rank_tied = m;
  if (m) {
    DBuf key, keys, sub2, grp, sg, jt;
    key.alloc((size_t)m * 8, st); keys.alloc((size_t)m * 8, st); sub2.alloc((size_t)m * 4, st);
    grp.alloc((size_t)m * 4, st); sg.alloc((size_t)m * 4, st); jt.alloc((size_t)m * 4, st);
    u32 *sub = id.as<u32>(), *subs = sub2.as<u32>();
    int tbits = 1;
    while (tbits < 32 && (1ull << tbits) < (u64)Tn) tbits++;
    bool first = true;
    while (m) {
      const u32 gM = (m + B - 1) / B;
      if (first) k_tr_key_cls<<<gM, B, 0, st>>>(sub, m, trank, label, cls32, key.as<u64>());
      else k_tr_key_next<<<gM, B, 0, st>>>(sub, m, trank, nexth, key.as<u64>());
      rsort_pairs64(c, key, keys, id, sub2, m, 0, 32 + tbits);
      key.swap(keys); id.swap(sub2);              // sorted: keys / sub2
      sub = id.as<u32>(); subs = sub2.as<u32>();
      k_tr_gs64<<<gM, B, 0, st>>>(keys.as<u64>(), m, grp.as<u32>(), sg.as<u32>());
      max_scan(tmp, grp.as<u32>(), m, st);
      max_scan(tmp, sg.as<u32>(), m, st);
      k_tr_refine<<<gM, B, 0, st>>>(keys.as<u64>(), subs, grp.as<u32>(), sg.as<u32>(), m, trank, tie.as<u8>());
      c->launches += 12;
      u32 m2 = select_tied(tmp, subs, tie.as<u8>(), m, sub, cnt, st);
      if (cls_only) {
        if (tie_by_id && m2) { k_tr_mark<<<(m2 + B - 1) / B, B, 0, st>>>(sub, m2, tie_by_id); LAUNCH(c); }
        break;
      }
      if (m2 && !first) {             // next round looks twice as far
        const u32 g2 = (m2 + B - 1) / B;
        k_tr_jump<<<g2, B, 0, st>>>(sub, m2, nexth, jt.as<u32>());
        k_tr_jump_store<<<g2, B, 0, st>>>(sub, m2, jt.as<u32>(), nexth);
        c->launches += 2;
      }
      if (!first) rounds++;
      first = false;
      m = m2;
      if (rounds > 64) throw CmsErr{CMSBWT_E_INTERNAL, "tail ranking did not converge"};
    }
  }
  c->stats.rank_rounds = rounds;
  CK(cudaGetLastError());
}

// ================================================================================================
// Multi-GPU ranking (SURVEY.md 8e "global head order").  Tails are range-partitioned by their
// 32-bit key; the owner of a key range ranks what it receives (radix sort + the (rank, cls) round)
// and sends ranks and tie flags back; the tails that are still tied are refined by every rank on the
// all-gathered (rank, rank[next^(2^h)]) keys of the tied subset -- the same rounds as above, with
// the collectives (driven by the host layer, cms_bwt_b200/multi.py) between the calls.
__global__ void k_sh_gather_cls(const u32 *perm, const u64 *label, u32 n, u32 *cs) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) cs[j] = (u32)label[perm[j]];
}
__global__ void k_sh_iota(u32 *a, u32 n) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) a[j] = j;
}
__global__ void k_sh_unperm(const u32 *perm, const u32 *rk, u32 n, u32 *trank) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) trank[perm[j]] = rk[j];
}
__global__ void k_sh_refine_scatter(const u64 *key, const u32 *id, const u32 *grp, const u32 *sg, u32 m,
                                    u32 *nr, u8 *nt) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  u32 o = id[j];
  nr[o] = (u32)(key[j] >> 32) + (sg[j] - grp[j]);
  nt[o] = (sg[j] != j) || (j + 1 < m && key[j + 1] == key[j]);
}
__global__ void k_sh_apply_own(const u32 *sub, u32 m, const u32 *nr, u32 *trank) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < m) trank[sub[j]] = nr[j];
}

// local sort of the shard's tails by key: sh_ks (sorted keys), sh_perm, sh_cs (cls in that order)
void shard_tails_sort(cmsbwt_ctx *c) {
  cudaStream_t st = c->st;
  const u32 Tn = c->s_Tn;
  const int B = 256;
  const u32 gT = (Tn + B - 1) / B;
  Scratch tmp;
  c->sh_ks.alloc((size_t)Tn * 4 + 16, st); c->sh_perm.alloc((size_t)Tn * 4 + 16, st);
  c->sh_cs.alloc((size_t)Tn * 4 + 16, st);
  const int bits = c->s_key_bits > 32 ? 32 : c->s_key_bits;
  size_t tb = 0;
  rsort_pairs32(c, c->s_key32, c->sh_ks, c->s_tid, c->sh_perm, Tn, 0, bits);
  c->s_key32.swap(c->sh_ks); c->s_tid.swap(c->sh_perm);      // sorted: sh_ks / sh_perm
  k_sh_gather_cls<<<gT, B, 0, st>>>(c->sh_perm.as<u32>(), c->s_label.as<u64>(), Tn, c->sh_cs.as<u32>());
  c->launches += 6;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));
}

// owner side: M received (key, cls) pairs -> rank_out[M] (rank_base + order), tie_out[M]
void shard_rank_owner(cmsbwt_ctx *c, const u32 *keys, const u32 *cls, u32 M, u32 rank_base, u32 *rank_out,
                      u8 *tie_out) {
  cudaStream_t st = c->st;
  if (M == 0) return;
  DBuf key32, id;
  key32.alloc((size_t)M * 4, st); id.alloc((size_t)M * 4, st);
  CK(cudaMemcpyAsync(key32.p, keys, (size_t)M * 4, cudaMemcpyDeviceToDevice, st));
  k_sh_iota<<<(M + 255) / 256, 256, 0, st>>>(id.as<u32>(), M);
  LAUNCH(c);
  tail_rank_ex(c, nullptr, cls, key32, id, c->s_key_bits, nullptr, M, rank_out, rank_base, true, tie_out);
  CK(cudaStreamSynchronize(st));
}

// origin side: ranks and tie flags came back in sh_ks order -> trank, tied subset; returns its size
u32 shard_rank_apply(cmsbwt_ctx *c, const u32 *rk_sorted, const u8 *tie_sorted) {
  cudaStream_t st = c->st;
  const u32 Tn = c->s_Tn;
  Scratch tmp;
  DBuf cnt;
  cnt.alloc(4, st);
  k_sh_unperm<<<(Tn + 255) / 256, 256, 0, st>>>(c->sh_perm.as<u32>(), rk_sorted, Tn, c->s_trank.as<u32>());
  LAUNCH(c);
  c->sh_sub.alloc((size_t)Tn * 4 + 16, st);
  c->sh_m = select_tied(tmp, c->sh_perm.as<u32>(), tie_sorted, Tn, c->sh_sub.as<u32>(), cnt, st);
  c->launches += 2;
  c->sh_round = 0;
  c->stats.rank_tied = c->sh_m;
  c->sh_ks.release(); c->sh_cs.release(); c->sh_perm.release();
  return c->sh_m;
}

// keys (rank, rank[next^(2^h)]) of the local tied subset -> sh_tiekeys
void shard_tie_keys(cmsbwt_ctx *c) {
  cudaStream_t st = c->st;
  const u32 m = c->sh_m;
  c->sh_tiekeys.alloc((size_t)m * 8 + 16, st);
  if (m) {
    k_tr_key_next<<<(m + 255) / 256, 256, 0, st>>>(c->sh_sub.as<u32>(), m, c->s_trank.as<u32>(), c->s_next.as<u32>(),
                                                   c->sh_tiekeys.as<u64>());
    LAUNCH(c);
  }
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));
}

// still-tied entries per source slab (every rank refines all the keys, so every rank knows the next
// round's counts of every other rank without asking)
__global__ void __launch_bounds__(256) k_sh_slab_ties(const u8 *__restrict__ nt, const u64 *__restrict__ off, u64 *out) {
  __shared__ u32 s_c;
  if (threadIdx.x == 0) s_c = 0;
  __syncthreads();
  u32 k = 0;
  for (u64 j = off[blockIdx.x] + threadIdx.x; j < off[blockIdx.x + 1]; j += blockDim.x) k += nt[j] != 0;
  k = __reduce_add_sync(0xffffffffu, k);
  if ((threadIdx.x & 31) == 0 && k) atomicAdd(&s_c, k);
  __syncthreads();
  if (threadIdx.x == 0) out[blockIdx.x] = s_c;
}

// every rank refines the all-gathered keys; this shard's entries are [my_off, my_off + sh_m).
// slab_off (world + 1 offsets of the sources' slabs) != NULL: next_counts[s] = entries of source s still tied.
u32 shard_tie_refine(cmsbwt_ctx *c, const u64 *allkeys, u32 m_total, u32 my_off, const u64 *slab_off, u32 world,
                     u64 *next_counts) {
  cudaStream_t st = c->st;
  const int B = 256;
  const u32 m = c->sh_m;
  if (m_total == 0) return 0;
  Scratch tmp;
  DBuf keys, ids0, ids, grp, sg, nr, nt, cnt, sub2, jt;
  keys.alloc((size_t)m_total * 8, st); ids0.alloc((size_t)m_total * 4, st); ids.alloc((size_t)m_total * 4, st);
  grp.alloc((size_t)m_total * 4, st); sg.alloc((size_t)m_total * 4, st); nr.alloc((size_t)m_total * 4, st);
  nt.alloc((size_t)m_total, st); cnt.alloc(4, st);
  const u32 gM = (m_total + B - 1) / B;
  k_sh_iota<<<gM, B, 0, st>>>(ids0.as<u32>(), m_total);
  size_t tb = 0;
  {
    DBuf kin;
    kin.alloc((size_t)m_total * 8, st);
    CK(cudaMemcpyAsync(kin.p, allkeys, (size_t)m_total * 8, cudaMemcpyDeviceToDevice, st));
    rsort_pairs64(c, kin, keys, ids0, ids, m_total, 0, 64);
    kin.swap(keys); ids0.swap(ids);               // sorted: keys / ids
  }
  k_tr_gs64<<<gM, B, 0, st>>>(keys.as<u64>(), m_total, grp.as<u32>(), sg.as<u32>());
  max_scan(tmp, grp.as<u32>(), m_total, st);
  max_scan(tmp, sg.as<u32>(), m_total, st);
  k_sh_refine_scatter<<<gM, B, 0, st>>>(keys.as<u64>(), ids.as<u32>(), grp.as<u32>(), sg.as<u32>(), m_total,
                                        nr.as<u32>(), nt.as<u8>());
  c->launches += 14;
  DBuf so, sc;
  if (slab_off && world) {
    so.alloc((size_t)(world + 1) * 8, st); sc.alloc((size_t)world * 8, st);
    CK(cudaMemcpyAsync(so.p, slab_off, (size_t)(world + 1) * 8, cudaMemcpyHostToDevice, st));
    k_sh_slab_ties<<<world, 256, 0, st>>>(nt.as<u8>(), so.as<u64>(), sc.as<u64>());
    CK(cudaMemcpyAsync(next_counts, sc.p, (size_t)world * 8, cudaMemcpyDeviceToHost, st));
    LAUNCH(c);
  }
  u32 m2 = 0;
  if (m) {
    k_sh_apply_own<<<(m + B - 1) / B, B, 0, st>>>(c->sh_sub.as<u32>(), m, nr.as<u32>() + my_off, c->s_trank.as<u32>());
    sub2.alloc((size_t)m * 4 + 16, st);
    m2 = select_tied(tmp, c->sh_sub.as<u32>(), nt.as<u8>() + my_off, m, sub2.as<u32>(), cnt, st);
    c->launches += 3;
    if (m2) {                           // the next round looks twice as far
      jt.alloc((size_t)m2 * 4, st);
      const u32 g2 = (m2 + B - 1) / B;
      k_tr_jump<<<g2, B, 0, st>>>(sub2.as<u32>(), m2, c->s_next.as<u32>(), jt.as<u32>());
      k_tr_jump_store<<<g2, B, 0, st>>>(sub2.as<u32>(), m2, jt.as<u32>(), c->s_next.as<u32>());
      c->launches += 2;
    }
    c->sh_sub.swap(sub2);
  }
  c->sh_m = m2;
  c->sh_round++;
  c->stats.rank_rounds = c->sh_round;
  if (c->sh_round > 64) throw CmsErr{CMSBWT_E_INTERNAL, "tail ranking did not converge"};
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));
  return m2;
}
