// TEST INFRASTRUCTURE ONLY: runs the lane state machine of cms_bwt_b200/csrc/ms_fast.cuh on the CPU,
// lane by lane, so that its logic can be checked against the oracle without a GPU.  The product
// never loads this; k_ms_fast (ms_fast.cu) is the only shipped user of that header.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>
#include "../../cms_bwt_b200/csrc/ms_fast.cuh"

typedef uint8_t u8;
typedef uint32_t u32;
typedef uint64_t u64;

static void pack(const u8 *s, u64 len, std::vector<u64> &out) {
  out.assign((len + 31) / 32 + MSF_PAD_WORDS, 0);
  for (u64 j = 0; j < len; j++) {
    u8 c = s[j];
    if (c == 'A' || c == 'C' || c == 'G' || c == 'T') out[j >> 5] |= (u64)msf_code_of(c) << ((j & 31) * 2);
  }
}

extern "C" {
// step counts per state of the last runs (used to model the kernel's cost per mutation)
unsigned long long g_steps[16];
void msf_host_steps(unsigned long long *out) {
  for (int i = 0; i < 16; i++) { out[i] = g_steps[i]; g_steps[i] = 0; }
  for (int i = 1; i < 6; i++) { out[8 + i] = msf_host_acts[i]; msf_host_acts[i] = 0; }
}

// returns the number of heads (text order) or -1 when the input is not eligible for the fast path
// (reference or text not pure ACGT), -2 on pool overflow.
// outputs: per-position dump arrays (N entries) and the compacted head list (capacity N).
long long msf_host_run(const u8 *x, u32 n, const int32_t *sa, const int32_t *plcp_prime, const u8 *text,
                       u64 N, u32 WL, int k, u32 skip_min, u32 nlanes, int use_slots, u32 *dpos, u32 *dlen, u8 *dsm,
                       u8 *dhead, u64 *hg, u32 *hpos, u32 *hlen, uint16_t *hcs) {
  // ---- eligibility (capi.cu: classify_x): A/C/G/T segments between a bounded number of runs of other symbols ----
  auto acgt = [](u8 b) { return b == 'A' || b == 'C' || b == 'G' || b == 'T'; };
  std::vector<u64> runs;
  std::vector<u32> xbk;
  u64 bases = 0;
  for (u32 j = 0; j < n;) {
    if (acgt(x[j])) { j++; bases++; continue; }
    u32 e = j;
    while (e < n && !acgt(x[e])) e++;
    runs.push_back((u64)j | ((u64)e << 32));
    j = e;
  }
  if (runs.size() > (size_t)n / 8 + 16 || bases == 0 || runs.empty() || (u32)(runs.back() >> 32) != n) return -1;
  const u32 nruns = (u32)runs.size(), n_acgt = (u32)runs.back();
  if (nruns > 1) {
    xbk.resize(((size_t)n >> 6) + 2);
    u32 j = 0;
    for (size_t b = 0; b < xbk.size(); b++) {
      while (j + 1 < runs.size() && (u32)(runs[j] >> 32) <= (b << 6)) j++;
      xbk[b] = j;
    }
  }
  runs.push_back(~0ull);
  for (u64 j = 0; j < N; j++) {          // any legal symbol: the others are escapes in the break list
    u8 c = text[j];
    if (c < 2 || c > 127) return -1;
  }
  // ---- index ----
  std::vector<u64> xp, tp;
  pack(x, n, xp);
  xp.resize((n + 31) / 32 + MSF_PAD_WORDS, 0);
  std::vector<u8> x8(x, x + n);
  x8.resize(n + 64, 0xFF);
  const u64 K = 1ull << (2 * k);
  std::vector<u32> ktab(K + 2, 0);
  {
    u32 fp = 0;
    for (u32 r = 0; r < n; r++) {
      u32 f = msf_kmers_le_suffix(x8.data(), (u32)sa[r], k);
      if (f > fp) ktab[fp] = r;
      if (r == n - 1 && f <= K) ktab[f] = n;
      fp = f;
    }
    u32 mx = 0;
    for (u64 m = 0; m <= K; m++) { mx = std::max(mx, ktab[m]); ktab[m] = mx; }
  }
  std::vector<u64> sak(n);
  for (u32 r = 0; r < n; r++)
    sak[r] = (u64)msf_make_key(xp.data(), x8.data(), (u32)sa[r],
                               msf_room_at(runs.data(), xbk.data(), nruns, n_acgt, (u32)sa[r]), k) | ((u64)(u32)sa[r] << 32);
  std::vector<u64> slots(K * 8);
  for (u64 code = 0; code < K; code++)
    msf_slot_fill(sak.data(), ktab.data(), xp.data(), n, runs.data(), xbk.data(), nruns, n_acgt, k, code, &slots[code * 8]);
  std::vector<u32> G(n + 8);
  for (u32 i = 0; i < n; i++) G[i] = i + (u32)plcp_prime[i];
  // ---- text ----
  pack(text, N, tp);
  std::vector<u64> seps;                 // break list: position << 1 | escape (ms_fast.cu: k_pack + sort)
  for (u64 j = 0; j < N; j++) {
    const u8 c = text[j];
    if (c == 2) seps.push_back(j << 1);
    else if (!(c == 'A' || c == 'C' || c == 'G' || c == 'T')) seps.push_back((j << 1) | 1);
  }
  const u64 nseps = seps.size();
  seps.push_back(~0ull);
  const u64 nwin = (N + WL - 1) / WL;
  std::vector<u32> wsep(nwin);
  for (u64 w = 0; w < nwin; w++)
    wsep[w] = (u32)(std::lower_bound(seps.begin(), seps.begin() + nseps, (w * WL) << 1) - seps.begin());
  std::vector<u8> t8(text, text + N);
  t8.resize(N + 64, 2);
  // ---- run ----
  const u64 nb_cap = N / MSF_RB + 2 * nwin + 16;
  std::vector<u32> pool(nb_cap * MSF_RB * 4), bnext(nb_cap), wfirst(nwin), wcnt(nwin, 0), wfp(nwin), wlp(nwin);
  u32 nb_used = 0, flags[2] = {0, 0};
  unsigned long long counter = 0;
  MsfParams P;
  P.slots = use_slots ? slots.data() : nullptr;
  P.xp = xp.data(); P.x8 = x8.data(); P.sak = sak.data(); P.ktab = ktab.data(); P.G = G.data();
  P.n = n; P.n_acgt = n_acgt; P.k = k;
  P.xruns = runs.data(); P.xbk = xbk.data(); P.nruns = nruns;
  P.tp = tp.data(); P.t8 = t8.data(); P.N = N; P.seps = seps.data(); P.nseps = nseps; P.wsep = wsep.data();
  P.WL = WL; P.nwin = nwin; P.skip_min = skip_min; P.variant = use_slots == 2 ? 2u | 16u : 2u;   // use_slots 2: empty-bucket slots too
  P.pool = pool.data(); P.batch_next = bnext.data(); P.nb_cap = (u32)nb_cap; P.nb_used = &nb_used;
  P.wfirst = wfirst.data(); P.wcnt = wcnt.data(); P.wfirstpos = wfp.data(); P.wlastpos = wlp.data();
  P.flags = flags; P.win_counter = &counter;
  P.dpos = dpos; P.dlen = dlen; P.dsm = dsm; P.dhead = dhead;
  if (nlanes < 1) nlanes = 1;
  if (nlanes > 32) nlanes = 32;
  std::vector<MsfLane> lanes(nlanes);
  for (auto &L : lanes) { memset(&L, 0, sizeof L); L.state = MSF_S_FETCH; }
  bool any = true;
  while (any) {                       // one emulated warp of nlanes lanes
    any = false;
    for (auto &L : lanes)
      if (L.state != MSF_S_DONE) { g_steps[L.state]++; any = true; }
    if (any) { g_steps[15]++; msf_warp_iter_host(P, lanes.data(), (int)nlanes); }
  }
  if (flags[0]) return -2;
  // ---- compaction (ms_fast.cu: k_msf_fix + scan + k_msf_gather) ----
  u64 H = 0;
  for (u64 w = 0; w < nwin; w++) {
    u32 drop = msf_window_drop(wfp.data(), wlp.data(), w);
    if (dhead && drop) dhead[w * WL] = 0;
    u32 b = wcnt[w] ? wfirst[w] : 0;
    for (u32 j = 0; j < wcnt[w]; j++) {
      if (j && j % MSF_RB == 0) b = bnext[b];
      const u32 *r = &pool[((u64)b * MSF_RB + j % MSF_RB) * 4];
      if (j == 0 && drop) continue;
      hg[H] = w * WL + r[0]; hpos[H] = r[1]; hlen[H] = r[2]; hcs[H] = (uint16_t)(r[3] & 0x1FF);
      H++;
    }
  }
  return (long long)H;
}

}  // extern "C"
