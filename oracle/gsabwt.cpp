// TEST INFRASTRUCTURE ONLY (see oracle/README.md) -- never linked into the product.
//
// Independent oracle O2: the multi-string BWT of a document list with
// document-ordered terminators, computed straight from a generalised suffix array
// (libsais_int over an int32 text, one distinct terminator per document).  It shares no
// logic with CMS-BWT: no matching statistics, no heads, no buckets.
//
// Document parsing follows the reference's collection reader
// (/root/reference/CMS-BWT-functions.cpp:1138-1140,1257-1266,1269): getline while good();
// an empty line or a line starting with '>' closes the current document; a non-empty
// remainder after EOF is one more document.
//
// usage: gsabwt <collection file> <out.bwt>
#include <libsais.h>
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <string>
#include <vector>

int main(int argc, char **argv) {
  if (argc != 3) { fprintf(stderr, "usage: %s <collection> <out.bwt>\n", argv[0]); return 2; }
  std::ifstream in(argv[1]);
  if (!in) { fprintf(stderr, "cannot open %s\n", argv[1]); return 1; }
  std::vector<std::string> docs;
  std::string line, cur;
  while (std::getline(in, line).good()) {
    if (line.empty() || line[0] == '>') { docs.push_back(cur); cur.clear(); }
    else cur += line;
  }
  if (!cur.empty()) docs.push_back(cur);
  const int32_t D = (int32_t)docs.size();
  size_t N = 0;
  for (auto &d : docs) N += d.size() + 1;
  if (N >= (size_t)INT32_MAX - 6000) { fprintf(stderr, "too large for the int32 oracle\n"); return 1; }
  std::vector<int32_t> T(N), SA(N + 6000);
  std::vector<uint8_t> is_start(N, 0);
  size_t p = 0;
  for (int32_t k = 0; k < D; k++) {
    is_start[p] = 1;
    for (unsigned char c : docs[k]) T[p++] = D + c;
    T[p++] = k;                       // terminator of document k sorts by k
  }
  if (libsais_int(T.data(), SA.data(), (int32_t)N, D + 256, 6000) != 0) {
    fprintf(stderr, "libsais_int failed\n"); return 1;
  }
  std::vector<uint8_t> out(N);
  for (size_t i = 0; i < N; i++) {
    int32_t s = SA[i];
    out[i] = is_start[s] ? 2 : (uint8_t)(T[s - 1] - D);
  }
  FILE *f = fopen(argv[2], "wb");
  if (!f) { fprintf(stderr, "cannot write %s\n", argv[2]); return 1; }
  fwrite(out.data(), 1, N, f);
  fclose(f);
  fprintf(stderr, "docs=%d N=%zu\n", D, N);
  return 0;
}
