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
__global__ void k_tr_rank0(const u32 *key, const u32 *id, const u32 *gs, u32 n, u32 *rank, u8 *tie) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  rank[id[j]] = gs[j];
  tie[j] = (gs[j] != j) || (j + 1 < n && key[j + 1] == key[j]);
}
// subset keys: (current rank, secondary) with secondary = cls (round 0) or rank of the far tail
__global__ void k_tr_key_cls(const u32 *sub, u32 m, const u32 *rank, const u64 *label, u64 *key) {
  u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  u32 t = sub[j];
  key[j] = ((u64)rank[t] << 32) | (u32)label[t];
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
  cudaStream_t st = c->st;
  const int B = 256;
  const u32 gT = (Tn + B - 1) / B;
  Scratch tmp;
  DBuf key32s, ids, gs, tie, cnt;
  key32s.alloc((size_t)Tn * 4, st);
  ids.alloc((size_t)Tn * 4, st); gs.alloc((size_t)Tn * 4, st); tie.alloc((size_t)Tn, st); cnt.alloc(4, st);
  const int bits = key_bits > 32 ? 32 : key_bits;
  size_t tb = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tb, key32.as<u32>(), key32s.as<u32>(), id.as<u32>(), ids.as<u32>(),
                                  (int)Tn, 0, bits, st);
  tmp.need(tb, st);
  CK(cub::DeviceRadixSort::SortPairs(tmp.b.p, tb, key32.as<u32>(), key32s.as<u32>(), id.as<u32>(),
                                     ids.as<u32>(), (int)Tn, 0, bits, st));
  k_tr_gs32<<<gT, B, 0, st>>>(key32s.as<u32>(), Tn, gs.as<u32>());
  max_scan(tmp, gs.as<u32>(), Tn, st);
  k_tr_rank0<<<gT, B, 0, st>>>(key32s.as<u32>(), ids.as<u32>(), gs.as<u32>(), Tn, trank, tie.as<u8>());
  c->launches += 8;
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
      if (first) k_tr_key_cls<<<gM, B, 0, st>>>(sub, m, trank, label, key.as<u64>());
      else k_tr_key_next<<<gM, B, 0, st>>>(sub, m, trank, nexth, key.as<u64>());
      tb = 0;
      cub::DeviceRadixSort::SortPairs(nullptr, tb, key.as<u64>(), keys.as<u64>(), sub, subs, (int)m, 0,
                                      32 + tbits, st);
      tmp.need(tb, st);
      CK(cub::DeviceRadixSort::SortPairs(tmp.b.p, tb, key.as<u64>(), keys.as<u64>(), sub, subs, (int)m, 0,
                                         32 + tbits, st));
      k_tr_gs64<<<gM, B, 0, st>>>(keys.as<u64>(), m, grp.as<u32>(), sg.as<u32>());
      max_scan(tmp, grp.as<u32>(), m, st);
      max_scan(tmp, sg.as<u32>(), m, st);
      k_tr_refine<<<gM, B, 0, st>>>(keys.as<u64>(), subs, grp.as<u32>(), sg.as<u32>(), m, trank, tie.as<u8>());
      c->launches += 12;
      u32 m2 = select_tied(tmp, subs, tie.as<u8>(), m, sub, cnt, st);
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
