// Host-side parsing of the collection for streamed ingest (no device code): the reference's
// document rules (:1138-1140, :1257-1275; small path :344-346, :464-482) applied to pieces of the
// file on several threads.  A line is classified on its own (header / blank line -> one 0x02,
// sequence line -> its bytes), so pieces cut at line starts parse independently; what crosses
// pieces is only whether the last line left a document open.
#include "common.cuh"
#include <string.h>
#include <thread>

extern "C" {

int64_t cmsbwt_line_start(const uint8_t *file, int64_t size, int64_t pos) {
  if (!file || size < 0) return CMSBWT_E_ARG;
  if (pos <= 0) return 0;
  if (pos >= size) return size;
  const uint8_t *e = (const uint8_t *)memchr(file + pos - 1, '\n', (size_t)(size - pos + 1));
  return e ? (e - file) + 1 : size;
}

int64_t cmsbwt_parse_piece(const uint8_t *file, int64_t begin, int64_t end, uint8_t *out, int64_t *ndocs,
                           int32_t *open) {
  if (!file || begin < 0 || end < begin) return CMSBWT_E_ARG;
  int64_t a = begin, w = 0, docs = 0;
  int32_t op = -1;
  while (a < end) {
    const uint8_t *e = (const uint8_t *)memchr(file + a, '\n', (size_t)(end - a));
    if (!e) break;
    const int64_t L = (e - file) - a;
    if (L == 0 || file[a] == '>') {
      if (out) out[w] = 2;
      w++;
      docs++;
      op = 0;
    } else {
      if (out) memcpy(out + w, file + a, (size_t)L);
      w += L;
      op = 1;
    }
    a += L + 1;
  }
  if (ndocs) *ndocs += docs;
  if (open && op >= 0) *open = op;
  return w;
}

int64_t cmsbwt_parse_collection_mt(const uint8_t *file, int64_t size, uint64_t prefix, uint8_t *out,
                                   int64_t *ndocs, int threads) {
  if (!file || !out || size < 0) return CMSBWT_E_ARG;
  if (prefix < (uint64_t)size) return cmsbwt_parse_collection(file, size, prefix, out, ndocs);
  int T = threads;
  if (T <= 0) { T = (int)std::thread::hardware_concurrency(); if (T < 1) T = 1; if (T > 32) T = 32; }
  if ((int64_t)T > size / 4096 + 1) T = (int)(size / 4096 + 1);
  std::vector<int64_t> cut((size_t)T + 1), w((size_t)T, 0), d((size_t)T, 0), off((size_t)T + 1, 0);
  std::vector<int32_t> op((size_t)T, -1);
  for (int t = 0; t <= T; t++) cut[t] = t == T ? size : cmsbwt_line_start(file, size, size / T * t);
  auto each = [&](auto fn) {
    std::vector<std::thread> th;
    for (int t = 1; t < T; t++) th.emplace_back(fn, t);
    fn(0);
    for (auto &x : th) x.join();
  };
  each([&](int t) { w[t] = cmsbwt_parse_piece(file, cut[t], cut[t + 1], nullptr, &d[t], &op[t]); });
  for (int t = 0; t < T; t++) off[t + 1] = off[t] + w[t];
  each([&](int t) { cmsbwt_parse_piece(file, cut[t], cut[t + 1], out + off[t], nullptr, nullptr); });
  int64_t N = off[T], docs = 0;
  int32_t last = 0;
  for (int t = 0; t < T; t++) { docs += d[t]; if (op[t] >= 0) last = op[t]; }
  if (last == 1) { out[N++] = 2; docs++; }
  if (ndocs) *ndocs = docs;
  return N;
}

}  // extern "C"
