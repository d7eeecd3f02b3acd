// Stage 1, fast path (2-bit data; other symbols through the text's break list and the reference's run list): 2-bit packing, the lane-per-window matching
// statistics kernel (state machine in ms_fast.cuh) and the compaction of its head records into the
// text-order arrays the rest of the pipeline reads.  Replaces CMS-BWT-functions.cpp:1138-1213
// (small path :344-419) for such inputs; anything else goes through k_ms (ms.cu).
#include "common.cuh"
#include "ms_fast.cuh"
#include <cub/cub.cuh>
#include <string.h>
#include <stdlib.h>

// ---- 8-bit -> 2-bit ---------------------------------------------------------------------------
// One thread packs 32 symbols into one word.  Symbols at positions >= len pack as 0.  Separators
// (0x02) are appended to `seps` (unsorted); any other byte outside A,C,G,T raises flags[0].
__device__ __forceinline__ u32 pack4(u32 w, bool &clean) {
  // 2-bit codes of four symbols (A,C,G,T -> 0,1,2,3), one per byte, and back through a PRMT table:
  // the round trip equals the input exactly when all four bytes are A, C, G or T
  u32 c = ((w >> 1) & 0x03030303u) ^ ((w >> 2) & 0x01010101u);
  u32 t = c | (c >> 4);                         // byte 0: c0 | c1 << 4, byte 2: c2 | c3 << 4
  u32 sel = __byte_perm(t, 0, 0x4420);          // selector nibbles c0, c1, c2, c3
  clean = __byte_perm(0x54474341u, 0, sel) == w;
  u32 p = c | (c >> 6);                         // byte 0: c0 | c1 << 2, byte 2: c2 | c3 << 2
  return (p & 0xFu) | ((p >> 12) & 0xF0u);
}

// one 64-bit word = 32 symbols (a, b = their 32 bytes); breaks are appended, illegal bytes raise *bad
__device__ __forceinline__ u64 pack_word(uint4 a, uint4 b, u64 p0, u64 len, u64 *seps, u64 sepcap,
                                         unsigned long long *nsep, bool &bad) {
  u32 v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  u64 word = 0;
#pragma unroll
  for (int t = 0; t < 8; t++) {
    u32 w = v[t];
    bool clean;
    u32 c8 = pack4(w, clean);
    if (!clean || p0 + 4 * t + 4 > len) {     // rare: separator, foreign symbol, end of the text
      const u64 pb = p0 + 4 * t;
      c8 = 0;
#pragma unroll
      for (int bb = 0; bb < 4; bb++) {
        u32 ch = (w >> (8 * bb)) & 0xFF;
        if (pb + bb >= len) continue;
        if (ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T') c8 |= msf_code_of(ch) << (2 * bb);
        else if (seps && (ch == 2 || (ch >= 3 && ch <= 127))) {     // separator, or an escape (any other legal symbol)
          unsigned long long k = atomicAdd(nsep, 1ull);
          if (k < sepcap) seps[k] = ((pb + bb) << 1) | (ch != 2 ? 1ull : 0ull);
          if (ch != 2) atomicAdd(nsep + 1, 1ull);
        } else bad = true;
      }
    }
    word |= (u64)c8 << (8 * t);
  }
  return word;
}

__global__ void __launch_bounds__(256) k_pack(const u8 *src, u64 len, u64 *dst, u64 nwords, u64 *seps,
                                              u64 sepcap, unsigned long long *nsep, u32 *flags) {
  for (u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x; j < nwords; j += (u64)gridDim.x * blockDim.x) {
    const u64 p0 = j * 32;
    u64 word = 0;
    if (p0 < len) {
      const uint4 *q = (const uint4 *)(src + p0);
      bool bad = false;
      word = pack_word(__ldg(q), __ldg(q + 1), p0, len, seps, sepcap, nsep, bad);
      if (bad) flags[0] = 1;
    }
    dst[j] = word;
  }
}

// The same packing with the text staged into shared memory by the TMA engine (north-star subsystem 1:
// "2-bit-packed DNA text is staged into shared memory with TMA"): persistent CTAs, two 8 KB stages,
// cp.async.bulk global -> shared completing on an mbarrier; a stage is one packed word per thread.
// The text must be readable up to the next multiple of 16 bytes (CMS_PAD guarantees it).
#define PK_CHUNK 8192u
__device__ __forceinline__ void pk_mbar_init(u64 *bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((u32)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void pk_load(u8 *sdst, const u8 *gsrc, u32 bytes, u64 *bar) {
  const u32 b = (u32)__cvta_generic_to_shared(bar);
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"((u32)__cvta_generic_to_shared(sdst)), "l"(gsrc), "r"(bytes), "r"(b) : "memory");
}
__device__ __forceinline__ void pk_wait(u64 *bar, u32 phase) {
  const u32 b = (u32)__cvta_generic_to_shared(bar);
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(b), "r"(phase) : "memory");
}
__global__ void __launch_bounds__(256) k_pack_tma(const u8 *src, u64 len, u64 *dst, u64 nwords, u64 *seps,
                                                  u64 sepcap, unsigned long long *nsep, u32 *flags) {
  __shared__ __align__(128) u8 s_buf[2][PK_CHUNK];
  __shared__ __align__(8) u64 s_bar[2];
  const u32 tid = threadIdx.x;
  const u64 nchunks = (nwords * 32 + PK_CHUNK - 1) / PK_CHUNK;
  const u64 padded = (len + 15) & ~15ull;                  // bytes that may be read
  if (tid == 0) {
    pk_mbar_init(&s_bar[0]); pk_mbar_init(&s_bar[1]);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  auto chunk_bytes = [&](u64 ch) -> u32 {
    const u64 a = ch * PK_CHUNK;
    return a >= padded ? 0u : (u32)min((u64)PK_CHUNK, padded - a);
  };
  u64 ch = blockIdx.x;
  u32 stage = 0, phase[2] = {0, 0};
  bool bad = false;
  if (tid == 0 && ch < nchunks && chunk_bytes(ch)) pk_load(s_buf[0], src + ch * PK_CHUNK, chunk_bytes(ch), &s_bar[0]);
  for (; ch < nchunks; ch += gridDim.x) {
    const u64 nx = ch + gridDim.x;
    // the other stage was consumed in the previous iteration (barrier at its end): refill it
    if (tid == 0 && nx < nchunks && chunk_bytes(nx)) pk_load(s_buf[stage ^ 1], src + nx * PK_CHUNK, chunk_bytes(nx), &s_bar[stage ^ 1]);
    const u32 nb = chunk_bytes(ch);
    const u64 j = ch * (PK_CHUNK / 32) + tid, p0 = j * 32;
    u64 word = 0;
    if (nb) {
      pk_wait(&s_bar[stage], phase[stage]);
      phase[stage] ^= 1;
      if (p0 < len) {
        const uint4 *q = (const uint4 *)(s_buf[stage] + tid * 32);
        uint4 a = q[0], b = make_uint4(0, 0, 0, 0);
        if (tid * 32 + 16 < nb) b = q[1];                  // bytes past the copied ones are never symbols (>= len)
        word = pack_word(a, b, p0, len, seps, sepcap, nsep, bad);
      }
    }
    if (j < nwords) dst[j] = word;
    __syncthreads();
    stage ^= 1;
  }
  if (bad) flags[0] = 1;
}

// measurement switch: CMSBWT_PACK_TMA=1 selects the TMA-staged kernel.  Measured at
// config 3 (profiles/r2_q5_bench_packtma*.json): 1.89 ms against 1.59 ms for the plain 16-byte loads --
// the packing is bound by the PRMT/validation arithmetic and the separator test, not by load issue, and
// the bulk copy adds the shared-memory round trip; the plain kernel stays the default.
static int pack_tma() {
  static int v = -1;
  if (v < 0) { const char *e = getenv("CMSBWT_PACK_TMA"); v = e ? atoi(e) : 0; }
  return v;
}

void pack_2bit(cmsbwt_ctx *c, const u8 *src, u64 len, u64 *dst, u64 nwords, u64 *seps, u64 sepcap,
               unsigned long long *nsep, u32 *flags) {
  u64 blocks = (nwords + 255) / 256;
  if (blocks > 148 * 64) blocks = 148 * 64;
  if (pack_tma() && ((uintptr_t)src & 15) == 0) {
    u64 nch = (nwords * 32 + PK_CHUNK - 1) / PK_CHUNK;
    u32 grid = (u32)(nch < 148 * 8 ? nch : 148 * 8);       // persistent: 8 CTAs per SM, 16 KB of stages each
    k_pack_tma<<<grid ? grid : 1, 256, 0, c->st>>>(src, len, dst, nwords, seps, sepcap, nsep, flags);
  } else
  k_pack<<<(u32)blocks, 256, 0, c->st>>>(src, len, dst, nwords, seps, sepcap, nsep, flags);
  LAUNCH(c);
  CK(cudaGetLastError());
}

// index of the first break (separator or escape, MSF_BRK_*) at or after the window start
__global__ void k_wsep(const u64 *seps, u64 nseps, u32 WL, u64 nwin, u32 *wsep) {
  u64 w = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= nwin) return;
  u64 key = w * WL, lo = 0, hi = nseps;
  while (lo < hi) {
    u64 mid = (lo + hi) >> 1;
    if (MSF_BRK_POS(seps[mid]) < key) lo = mid + 1; else hi = mid;
  }
  wsep[w] = (u32)lo;
}

// ---- the kernel ---------------------------------------------------------------------------------
template <int OCC>
__global__ void __launch_bounds__(128, OCC) k_ms_fast(MsfParams P) {
  MsfLane L;
  L.i = L.w1 = L.sepnext = 0;
  L.sepidx = L.w = 0;
  L.prevpos = MSF_INVALID; L.prevterm = 0; L.first = 0;
  L.curb = MSF_NOB; L.fill = 0; L.cnt = 0; L.limit = 0;
  L.tkey = L.lo = L.hi = L.rb = L.prevkey = L.prevsa = L.haveprev = 0;
  L.pos = L.len = L.sm = 0; L.chb = 4;
  L.state = MSF_S_FETCH;
  while (!__all_sync(0xffffffffu, L.state == MSF_S_DONE)) msf_warp_iter(P, L);
}

// ---- compaction -------------------------------------------------------------------------------
__global__ void k_msf_fix(const u32 *wcnt, const u32 *wfirstpos, const u32 *wlastpos, u64 nwin, u32 WL,
                          u32 *wadj, u8 *dhead) {
  u64 w = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= nwin) return;
  u32 drop = msf_window_drop(wfirstpos, wlastpos, w);
  wadj[w] = wcnt[w] - drop;
  if (drop && dhead) dhead[w * WL] = 0;
}

// GL threads per window walk its batch list and write the text-order arrays (a batch is 32 records = 512
// bytes: with 32 lanes a warp moves one batch per step and the steps of a window are independent loads)
template <int GL>
__global__ void __launch_bounds__(256) k_msf_gather(const uint4 *__restrict__ pool, const u32 *__restrict__ batch_next,
                                                    const u32 *__restrict__ wfirst, const u32 *__restrict__ wcnt,
                                                    const u32 *__restrict__ wadj, const u64 *__restrict__ woff, u64 nwin,
                                                    u32 WL, u64 *g, u32 *pos, u32 *len, u16 *cs) {
  static_assert(MSF_RB % GL == 0, "a step never straddles two batches");
  u64 gid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  u64 w = gid / GL;
  u32 l = (u32)(gid % GL);
  if (w >= nwin) return;
  u32 cnt = wcnt[w];
  if (cnt == 0) return;
  u32 drop = cnt - wadj[w];
  u64 off = woff[w];
  u32 b = wfirst[w];
  for (u32 j0 = 0; j0 < cnt; j0 += GL) {
    if (j0 && (j0 % MSF_RB) == 0) b = batch_next[b];
    u32 j = j0 + l;
    if (j < cnt && !(j == 0 && drop)) {
      uint4 r = pool[(u64)b * MSF_RB + (j % MSF_RB)];
      u64 o = off + j - drop;
      g[o] = w * WL + r.x; pos[o] = r.y; len[o] = r.z; cs[o] = (u16)(r.w & 0x1FF);
    }
  }
}

// Returns false (nothing changed) when the text holds an illegal byte or is mostly symbols outside A,C,G,T.
bool ms_stage_fast(cmsbwt_ctx *c) {
  cudaStream_t st = c->st;
  const u64 N = c->N;
  int sms = 148, occ = 4;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->dev);
  void (*kern)(MsfParams) = c->ms_occ >= 12 ? k_ms_fast<12> : c->ms_occ >= 10 ? k_ms_fast<10> : c->ms_occ >= 8 ? k_ms_fast<8> : k_ms_fast<6>;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 128, MSF_SMEM_BYTES(128));
  if (occ < 1) occ = 1;
  // window length: a lane walks its window serially, so the kernel ends with its slowest window; at
  // least 4 windows per resident lane keep that tail short (measured: 8192 is best at 5.9 Gbp,
  // 1024-2048 at <= 1 Gbp or when whole windows are novel sequence), 0 = pick by the text length
  u32 WL = (u32)c->chunk_len;
  if (c->chunk_len == 0) {
    u64 per = N / ((u64)sms * occ * 128 * 4);
    WL = per >= 8192 ? 8192u : per <= 1024 ? 1024u : (u32)(per & ~63ull);
  }
  const u64 nwin = (N + WL - 1) / WL;
  if (nwin >= 0xFFFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "too many windows"};
  // ---- pack the text, collect separators ----
  const u64 nwords = (N + 31) / 32 + MSF_PAD_WORDS;
  c->d_tp.alloc(nwords * 8, st);
  DBuf flags, nsep, tmp;
  flags.alloc(8, st); nsep.alloc(16, st);
  u64 sepcap = c->sep_hint ? c->sep_hint + 16 : (N / 1024 + (1u << 16));
  unsigned long long hsep2[2] = {0, 0};
  unsigned long long &hsep = hsep2[0];
  u32 hf[2] = {0, 0};
  struct Ev3 {                                   // destroyed on every way out, the throwing ones included
    cudaEvent_t a = nullptr, b = nullptr, c = nullptr;
    Ev3() { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventCreate(&c); }
    ~Ev3() { cudaEventDestroy(a); cudaEventDestroy(b); cudaEventDestroy(c); }
  } ev;
  cudaEvent_t e0 = ev.a, e1 = ev.b, e2 = ev.c;
  cudaEventRecord(e0, st);
  GroupSpan *gpack = new GroupSpan(c, CMSBWT_G_PACK);
  for (int attempt = 0; attempt < 2; attempt++) {
    c->d_seps.alloc((sepcap + 1) * 8, st);
    CK(cudaMemsetAsync(flags.p, 0, 8, st));
    CK(cudaMemsetAsync(nsep.p, 0, 16, st));
    pack_2bit(c, c->text, N, c->d_tp.as<u64>(), nwords, c->d_seps.as<u64>(), sepcap,
              nsep.as<unsigned long long>(), flags.as<u32>());
    CK(cudaMemcpyAsync(hsep2, nsep.p, 16, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(hf, flags.p, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    // a byte outside 2..127, or so many escapes (lower-case genomes: every other symbol) that the break
    // list would outgrow the text: the general 8-bit kernel takes the job
    if (hf[0] || hsep2[1] > N / 32 + 4096) { delete gpack; return false; }
    if (hsep <= sepcap) break;
    if (attempt == 1) throw CmsErr{CMSBWT_E_INTERNAL, "separator list overflow on exact-size rerun"};
    sepcap = hsep;
  }
  c->sep_hint = hsep;
  c->stats.n_escapes = hsep2[1];
  const u64 nseps = hsep;
  if (nseps >= 0xFFFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "more than 2^32 documents"};
  {                                   // sort the separator positions, add the sentinel
    DBuf alt;
    alt.alloc((nseps + 1) * 8, st);
    int bits = 2; while (bits < 64 && (1ull << bits) < 2 * N) bits++;       // position << 1 | escape
    rsort_keys64(c, c->d_seps, alt, nseps, 0, bits);
    CK(cudaMemsetAsync(c->d_seps.as<u64>() + nseps, 0xFF, 8, st));
  }
  DBuf wsep, wfirst, wcnt, wfp, wlp, wadj, woff, counter, nbused;
  wsep.alloc(nwin * 4, st); wfirst.alloc(nwin * 4, st); wcnt.alloc(nwin * 4, st);
  wfp.alloc(nwin * 4, st); wlp.alloc(nwin * 4, st); wadj.alloc((nwin + 1) * 4, st);
  woff.alloc((nwin + 1) * 8, st); counter.alloc(8, st); nbused.alloc(4, st);
  k_wsep<<<(u32)((nwin + 255) / 256), 256, 0, st>>>(c->d_seps.as<u64>(), nseps, WL, nwin, wsep.as<u32>());
  LAUNCH(c);
  delete gpack;
  cudaEventRecord(e1, st);
  if (c->dump) {
    c->dump_pos.alloc(N * 4, st); c->dump_len.alloc(N * 4, st);
    c->dump_sm.alloc(N, st); c->dump_head.alloc(N, st);
  }
  size_t tb = 0;
  EXSUM64(nullptr, tb, wadj.as<u32>(), woff.as<u64>(), (int)(nwin + 1), st);
  tmp.alloc(tb, st);
  // pool: batches of MSF_RB records; every window wastes at most one partly filled batch
  u64 rec_est = c->head_hint ? c->head_hint + c->head_hint / 8 : N / 16;
  u64 nb = rec_est / MSF_RB + nwin + 1024;
  for (int attempt = 0; attempt < 2; attempt++) {
    if (nb >= 0xFFFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "head pool too large"};
    DBuf pool, bnext;
    pool.alloc(nb * MSF_RB * 16, st); bnext.alloc(nb * 4, st);
    CK(cudaMemsetAsync(flags.p, 0, 8, st));
    CK(cudaMemsetAsync(counter.p, 0, 8, st));
    CK(cudaMemsetAsync(nbused.p, 0, 4, st));
    MsfParams P;
    P.slots = c->d_slots.as<u64>();
    P.xp = c->xp; P.x8 = c->ix.x; P.sak = c->d_sak.as<u64>(); P.ktab = c->ix.ktab; P.G = c->ix.G;
    P.n = c->ix.n; P.n_acgt = c->n_acgt; P.k = c->ix.k;
    P.xruns = c->d_xruns.as<u64>(); P.xbk = c->d_xbk.as<u32>(); P.nruns = c->nruns;
    P.tp = c->d_tp.as<u64>(); P.t8 = c->text; P.N = N; P.seps = c->d_seps.as<u64>(); P.nseps = nseps; P.wsep = wsep.as<u32>();
    P.WL = WL; P.nwin = nwin; P.skip_min = (u32)c->skip_min; P.variant = (u32)c->ms_variant;
    P.pool = pool.as<u32>(); P.batch_next = bnext.as<u32>(); P.nb_cap = (u32)nb; P.nb_used = nbused.as<u32>();
    P.wfirst = wfirst.as<u32>(); P.wcnt = wcnt.as<u32>(); P.wfirstpos = wfp.as<u32>(); P.wlastpos = wlp.as<u32>();
    P.flags = flags.as<u32>(); P.win_counter = counter.as<unsigned long long>();
    P.dpos = c->dump ? c->dump_pos.as<u32>() : nullptr; P.dlen = c->dump_len.as<u32>();
    P.dsm = c->dump_sm.as<u8>(); P.dhead = c->dump_head.as<u8>();
    u64 want = (nwin + 127) / 128;
    u32 grid = (u32)min(want, (u64)sms * occ);
    cudaEventRecord(e1, st);
    // pin the k-mer table + packed reference in L2 for this kernel: every search starts there,
    // while the {key, sa} array, G and the text stream through
    cudaStreamAttrValue apw;
    memset(&apw, 0, sizeof apw);
    if (c->l2_pin && c->l2_persist_max) {
      apw.accessPolicyWindow.base_ptr = c->d_ktab.p;
      apw.accessPolicyWindow.num_bytes = c->hot_bytes < (size_t)c->l2_window_max ? c->hot_bytes : (size_t)c->l2_window_max;
      double r = (double)c->l2_persist_max / (double)apw.accessPolicyWindow.num_bytes;
      apw.accessPolicyWindow.hitRatio = r > 1.0 ? 1.0f : (float)r;
      apw.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
      apw.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
      cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &apw);
    }
    { GroupSpan g(c, CMSBWT_G_MS);
    kern<<<grid, 128, MSF_SMEM_BYTES(128), st>>>(P); }
    if (c->l2_pin && c->l2_persist_max) {
      apw.accessPolicyWindow.num_bytes = 0;
      cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &apw);
    }
    cudaEventRecord(e2, st);
    LAUNCH(c);
    CK(cudaGetLastError());
    GroupSpan ggather(c, CMSBWT_G_GATHER);
    k_msf_fix<<<(u32)((nwin + 255) / 256), 256, 0, st>>>(wcnt.as<u32>(), wfp.as<u32>(), wlp.as<u32>(), nwin, WL,
                                                        wadj.as<u32>(), c->dump ? c->dump_head.as<u8>() : nullptr);
    size_t t = tb;
    CK(EXSUM64(tmp.p, t, wadj.as<u32>(), woff.as<u64>(), (int)nwin, st));
    c->launches += 2;
    u64 lastoff = 0; u32 lastcnt = 0;
    CK(cudaMemcpyAsync(hf, flags.p, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&lastoff, woff.as<u64>() + (nwin - 1), 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&lastcnt, wadj.as<u32>() + (nwin - 1), 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    { float ms = 0; cudaEventElapsedTime(&ms, e1, e2); c->stats.ms_kernel_ms = ms;
      cudaEventElapsedTime(&ms, e0, e1); c->stats.pack_ms = ms; }
    const u64 H = lastoff + lastcnt;
    if (hf[0]) {                    // pool exhausted: counts are exact, rerun with the exact size
      if (attempt == 1) throw CmsErr{CMSBWT_E_INTERNAL, "head pool overflow on exact-size rerun"};
      nb = (H + nwin) / MSF_RB + 2 * nwin + 1024;
      continue;
    }
    if (H >= 0x7FFFFFF0ull) throw CmsErr{CMSBWT_E_LIMIT, "more than 2^31 heads on one device"};
    c->H = H;
    c->head_hint = H;
    c->h_gstart.alloc((H + 1) * 8, st); c->h_pos.alloc((H + 1) * 4, st);
    c->h_len.alloc((H + 1) * 4, st); c->h_cs.alloc((H + 1) * 2, st);
    static int gl = -1;                           // measurement switch: lanes per window (8, 16, 32)
    if (gl < 0) { const char *e = getenv("CMSBWT_GATHER_LANES"); gl = e ? atoi(e) : 16; }   // gather group at config 3: 8 lanes 0.71, 16 0.69, 32 0.71 ms
#define MSF_GATHER(GL)                                                                                              \
    k_msf_gather<GL><<<(u32)((nwin * GL + 255) / 256), 256, 0, st>>>(                                               \
        pool.as<uint4>(), bnext.as<u32>(), wfirst.as<u32>(), wcnt.as<u32>(), wadj.as<u32>(), woff.as<u64>(), nwin, WL, \
        c->h_gstart.as<u64>(), c->h_pos.as<u32>(), c->h_len.as<u32>(), c->h_cs.as<u16>())
    if (gl == 8) MSF_GATHER(8); else if (gl == 32) MSF_GATHER(32); else MSF_GATHER(16);
#undef MSF_GATHER
    LAUNCH(c);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(st));
    break;
  }
  c->stats.n_heads = c->H;
  c->stats.fast_path = 1;
  return true;
}
