// cms_bwt -- host driver of the B200-native CMS-BWT path.
//
// Same process contract as the reference's main.cpp (/root/reference/main.cpp:14-132):
//   cms_bwt [-p <chars>] [-b <GB>] [-r] [-o <basename>] [-h] <list file>
// list file: line 1 = reference path (verbatim), line 2 = collection path (trailing whitespace
// trimmed) (main.cpp:97-115); outputs <out>.bwt or <out>.rl_bwt and always <out>.log; the stdout
// banner lines and exit codes (usage -> -1, unreadable/empty inputs -> 1) follow main.cpp:31-34,
// :55-59,:74-80,:91-115,:121-126 and CMS-BWT-functions.cpp:156-160,:1789-1791.
//
// Where the reference calls computeBWT (CMS-BWT.h:69) this driver calls the C ABI of
// include/cmsbwt_b200.h: cmsbwt_index_upload, then cmsbwt_run.  There is no CPU fallback: without
// a CUDA device the program exits 1.  Extra, non-reference options are long options only
// (--device N, --no-host-sa, --no-stream, --threads T) so the reference's getopt string "p:b:o:rh"
// keeps its meaning.
//
// Ingest and egress are streamed (--no-stream keeps the plain read / parse / copy /
// write sequence): the collection file is mapped and parsed in pieces by several host threads into
// pinned staging slots (cmsbwt_parse_piece), each piece goes to the device while the next one is
// parsed (cmsbwt_text_append), the host suffix array (libsais, one thread) is built meanwhile on
// its own thread, and the result comes back range by range into pinned slots that writer threads
// pwrite() into the output file while the next range is copied (cmsbwt_result_read).
//
// Reference SA: north-star keeps it on the host with libsais (the reference's call at
// CMS-BWT-functions.cpp:247).  libsais is a third-party dependency of the product: third_party/Makefile
// compiles it into third_party/_build/libsais.a and this program links it statically
// (CMSBWT_HAVE_LIBSAIS); a build without the archive, or --no-host-sa, builds the SA on the GPU with
// cmsbwt_build_sa.  CMSBWT_LIBSAIS=<path of a libsais shared object> overrides the linked copy.  The
// route taken is printed ("Computed SA in ... (libsais ...)" / "(GPU prefix doubling)").  No path
// outside the files named on the command line (and CMSBWT_LIBSAIS) is ever opened by this program.
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <fstream>
#include <getopt.h>
#include <iostream>
#include <string>
#include <vector>
#include <unistd.h>
#include <csignal>
#include <fcntl.h>
#include <future>
#include <sys/mman.h>
#include <sys/stat.h>
#include <thread>
#include "cmsbwt_b200.h"

struct Args {                       // utils.h:8-14
  std::string filename, outname;
  bool format = false;              // 0 full BWT | 1 run-length encoded BWT
  unsigned buffer = 2;              // accepted for compatibility; device staging is sized from N
  unsigned long long prefixLength = UINT64_MAX;
  int device = 0;
  bool host_sa = true;
  bool stream = true;               // streamed ingest / egress
  int threads = 0;                  // host threads of the parser (0 = hardware threads, at most 16)
  int gpus = 0;                     // > 0: the multi-GPU path on devices 0..gpus-1 (one host thread per GPU, NCCL)
};

static void print_help(char **argv) {
  std::cout << "Usage: " << argv[0] << " [options] <input filename>" << std::endl;
  std::cout << "<input filename> is the name of the file containing paths to the reference sequence "
               "(in the first line) and to the collection file (in the second line)." << std::endl;
  std::cout << "  Options: " << std::endl
            << "\t-p \tread only a prefix of the file expressed in number of characters, def. whole file" << std::endl
            << "\t-b \tsize for the additional memory buffer in GB, def. 2 " << std::endl
            << "\t-r \toutputs the run-length encoded BWT, def. false " << std::endl
            << "\t-o \tbasename for the output files, def. <input filename>" << std::endl
            << "\t-h \tprints this help" << std::endl
            << "\t--device N \tCUDA device ordinal, def. 0" << std::endl
            << "\t--no-host-sa \tbuild the reference suffix array on the GPU even if libsais is available" << std::endl
            << "\t--no-stream \tread, parse, copy and write one after the other (no overlap)" << std::endl
            << "\t--threads T \thost threads of the collection parser, def. all (at most 16)" << std::endl
            << "\t--gpus N \tshard the collection's documents over GPUs 0..N-1 (NCCL), def. one GPU" << std::endl;
  exit(-1);
}

static void parse_args(int argc, char **argv, Args &arg) {
  puts("==== Command line:");
  for (int i = 0; i < argc; i++) printf(" %s", argv[i]);
  puts("");
  static const option longopts[] = {{"device", required_argument, nullptr, 1000},
                                    {"no-host-sa", no_argument, nullptr, 1001},
                                    {"no-stream", no_argument, nullptr, 1002},
                                    {"threads", required_argument, nullptr, 1003},
                                    {"gpus", required_argument, nullptr, 1004},
                                    {nullptr, 0, nullptr, 0}};
  int c;
  while ((c = getopt_long(argc, argv, "p:b:o:rh", longopts, nullptr)) != -1) {
    switch (c) {
      case 'p': arg.prefixLength = (unsigned long long)atoll(optarg); break;
      case 'b': arg.buffer = (unsigned)atoi(optarg); break;
      case 'r': arg.format = true; break;
      case 'o': arg.outname = optarg; break;
      case 1000: arg.device = atoi(optarg); break;
      case 1001: arg.host_sa = false; break;
      case 1002: arg.stream = false; break;
      case 1003: arg.threads = atoi(optarg); break;
      case 1004: arg.gpus = atoi(optarg); break;
      case 'h': print_help(argv); break;
      default:
        std::cout << "Unknown option. Use -h for help." << std::endl;
        exit(-1);
    }
  }
  if (argc == optind + 1) arg.filename = argv[optind];
  else {
    std::cout << "Invalid number of arguments" << std::endl;
    print_help(argv);
  }
  if (arg.outname.empty()) arg.outname = arg.filename;
  std::cout << "==== Parameters:" << std::endl;
  std::cout << "Input file: " << arg.filename << std::endl;
  std::cout << "Output basename: " << arg.outname << std::endl;
  std::cout << "Prefix length: " << arg.prefixLength << std::endl;
  std::cout << "Buffer size: " << arg.buffer << " GB" << std::endl;
  std::cout << "Output format: " << (arg.format ? "RLE" : "FULL") << std::endl;
  std::cout << std::endl;
}

static bool read_file(const std::string &path, std::vector<uint8_t> &out) {
  FILE *f = fopen(path.c_str(), "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  long sz = ftell(f);
  fseek(f, 0, SEEK_SET);
  if (sz < 0) { fclose(f); return false; }
  out.resize((size_t)sz);
  size_t got = sz ? fread(out.data(), 1, (size_t)sz, f) : 0;
  fclose(f);
  return got == (size_t)sz;
}

typedef int32_t (*libsais_fn)(const uint8_t *, int32_t *, int32_t, int32_t, int32_t *);
#ifdef CMSBWT_HAVE_LIBSAIS
extern "C" int32_t libsais(const uint8_t *T, int32_t *SA, int32_t n, int32_t fs, int32_t *freq);
#endif
static libsais_fn find_libsais(const char **route) {
  if (const char *e = getenv("CMSBWT_LIBSAIS")) {          // explicit override, never a default path
    void *h = dlopen(e, RTLD_NOW | RTLD_LOCAL);
    if (h)
      if (void *s = dlsym(h, "libsais")) { *route = "libsais, CMSBWT_LIBSAIS"; return (libsais_fn)s; }
    fprintf(stderr, "Error: CMSBWT_LIBSAIS=%s is not a libsais shared object\n", e);
    exit(1);
  }
#ifdef CMSBWT_HAVE_LIBSAIS
  *route = "libsais, linked";
  return &libsais;
#else
  *route = "GPU prefix doubling";
  return nullptr;
#endif
}

static double ms_since(std::chrono::high_resolution_clock::time_point t) {
  return std::chrono::duration<double, std::milli>(std::chrono::high_resolution_clock::now() - t).count();
}

// ---- streamed ingest / egress ---------------------------------------------------
struct Mapped {                     // read-only view of a file: mmap, or a heap copy where mapping fails
  const uint8_t *p = nullptr;
  int64_t size = 0;
  bool mapped = false;
  std::vector<uint8_t> heap;
  bool open(const std::string &path) {
    int fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) return false;
    struct stat sb;
    if (fstat(fd, &sb) == 0 && S_ISREG(sb.st_mode) && sb.st_size > 0) {
      void *m = mmap(nullptr, (size_t)sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
      if (m != MAP_FAILED) {
        madvise(m, (size_t)sb.st_size, MADV_SEQUENTIAL);
        p = (const uint8_t *)m;
        size = (int64_t)sb.st_size;
        mapped = true;
        ::close(fd);
        return true;
      }
    }
    ::close(fd);
    if (!read_file(path, heap)) return false;
    p = heap.data();
    size = (int64_t)heap.size();
    return true;
  }
  void close() {
    if (mapped) munmap((void *)p, (size_t)size);
    std::vector<uint8_t>().swap(heap);
    p = nullptr; size = 0; mapped = false;
  }
};

static int64_t env_mb(const char *name, int64_t def_mb) {
  const char *e = getenv(name);
  int64_t v = e ? atoll(e) : 0;          // CMSBWT_PIECE_KB / CMSBWT_OUT_KB: sizes for tests
  return v > 0 ? v * 1024 : def_mb * 1024 * 1024;
}

// Parses file[0, size) piece by piece on T threads into two pinned slots and appends every piece to
// the device text while the next one is parsed.  Returns N (0: no document) or a negative code.
static int64_t ingest_streamed(cmsbwt_ctx *ctx, const uint8_t *file, int64_t size, int T, int64_t *ndocs) {
  const int64_t piece = env_mb("CMSBWT_PIECE_KB", 64);   // 2 pinned slots of this size (pinning costs ~0.4 ms/MB)
  int rc = cmsbwt_text_begin(ctx, (uint64_t)size + 1);
  if (rc) return rc;
  void *slot[2] = {nullptr, nullptr};
  int64_t cap = 0;
  int64_t N = 0, docs = 0;
  int32_t open = 0;
  std::vector<int64_t> cut((size_t)T + 1), w((size_t)T), d((size_t)T);
  std::vector<int32_t> op((size_t)T);
  int k = 0;
  for (int64_t a = 0; a < size && rc == 0; k++) {
    const int64_t b = cmsbwt_line_start(file, size, a + piece < size ? a + piece : size);
    const int s = k & 1;
    if (b - a + 16 > cap) {                 // first piece, or a line longer than a piece
      for (int q = 0; q < 2; q++) { if ((rc = cmsbwt_copy_wait(ctx, q))) break; cmsbwt_host_free(slot[q]); slot[q] = nullptr; }
      cap = b - a + 16;
      if (cap < piece + (1 << 20) && size > piece) cap = piece + (1 << 20);   // pieces end at line starts: leave slack so that the slots are pinned once
      for (int q = 0; q < 2 && rc == 0; q++) rc = cmsbwt_host_alloc(&slot[q], (uint64_t)cap);
      if (rc) break;
    }
    if ((rc = cmsbwt_copy_wait(ctx, s))) break;      // the copy that last used this slot is done
    uint8_t *dst = (uint8_t *)slot[s];
    for (int t = 0; t <= T; t++) cut[t] = t == T ? b : cmsbwt_line_start(file, b, a + (b - a) / T * t);
    auto work = [&](int t) {
      d[t] = 0; op[t] = -1;
      w[t] = cmsbwt_parse_piece(file, cut[t], cut[t + 1], dst + (cut[t] - a), &d[t], &op[t]);
    };
    std::vector<std::thread> th;
    for (int t = 1; t < T; t++) th.emplace_back(work, t);
    work(0);
    for (auto &x : th) x.join();
    for (int t = 0; t < T && rc == 0; t++) {
      if (w[t] < 0) { rc = (int)w[t]; break; }
      rc = cmsbwt_text_append(ctx, dst + (cut[t] - a), (uint64_t)w[t]);
      N += w[t]; docs += d[t];
      if (op[t] >= 0) open = op[t];
    }
    if (rc == 0) rc = cmsbwt_copy_fence(ctx, s);
    a = b;
  }
  if (rc == 0 && open == 1) {               // a non-empty remainder is one more document (:1269-1275)
    const int s = k & 1;
    if (!slot[s]) rc = cmsbwt_host_alloc(&slot[s], 16);
    if (rc == 0) rc = cmsbwt_copy_wait(ctx, s);
    if (rc == 0) {
      *(uint8_t *)slot[s] = 2;
      rc = cmsbwt_text_append(ctx, (const uint8_t *)slot[s], 1);
      N++; docs++;
    }
    if (rc == 0) rc = cmsbwt_copy_fence(ctx, s);
  }
  for (int q = 0; q < 2; q++) { cmsbwt_copy_wait(ctx, q); cmsbwt_host_free(slot[q]); }
  if (ndocs) *ndocs = docs;
  return rc ? rc : N;
}

static bool write_all(int fd, const uint8_t *p, uint64_t bytes, uint64_t off) {
  while (bytes) {
    ssize_t k = pwrite(fd, p, (size_t)bytes, (off_t)off);
    if (k <= 0) return false;
    p += k; off += (uint64_t)k; bytes -= (uint64_t)k;
  }
  return true;
}

// Reads the result back range by range into pinned slots; writer threads store a range into the
// output file while the next one is copied.  Buffered write()s to one file serialise on its inode
// lock (measured: 2.4 GB/s with 5 writers in flight), so the file is sized first and written through
// a shared mapping, where the page faults of different threads run in parallel; pwrite() remains
// where the mapping fails (CMSBWT_OUT_MMAP=0 forces it).  A full disk then shows up as SIGBUS
// instead of a short write: the handler reports it the way the reference reports I/O errors.
// Returns 0, a negative library code, or 1 for a write error.
static const char *g_outpath = "";
static void on_sigbus(int) {
  const char *m = "Error writing ";
  if (write(2, m, strlen(m)) < 0 || write(2, g_outpath, strlen(g_outpath)) < 0 || write(2, "\n", 1) < 0) {}
  _exit(1);
}

static int egress_streamed(cmsbwt_ctx *ctx, int fd, uint64_t out_bytes) {
  enum { NS = 6 };
  const uint64_t ch = (uint64_t)env_mb("CMSBWT_OUT_KB", 16);
  void *slot[NS] = {};
  std::future<bool> pend[NS];
  int rc = 0;
  bool ok = true;
  uint8_t *map = nullptr;
  const char *e = getenv("CMSBWT_OUT_MMAP");
  if (out_bytes && !(e && e[0] == '0') && ftruncate(fd, (off_t)out_bytes) == 0) {
    void *m = mmap(nullptr, (size_t)out_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (m != MAP_FAILED) { map = (uint8_t *)m; signal(SIGBUS, on_sigbus); }
  }
  for (int q = 0; q < NS && rc == 0 && (uint64_t)q * ch < out_bytes; q++) rc = cmsbwt_host_alloc(&slot[q], ch);
  uint64_t j = 0;
  for (uint64_t off = 0; off < out_bytes && rc == 0 && ok; off += ch, j++) {
    const int s = (int)(j % NS);
    if (pend[s].valid()) ok = pend[s].get() && ok;
    const uint64_t m = out_bytes - off < ch ? out_bytes - off : ch;
    if ((rc = cmsbwt_result_read(ctx, slot[s], off, m))) break;
    if ((rc = cmsbwt_copy_fence(ctx, 2 + s))) break;
    if ((rc = cmsbwt_copy_wait(ctx, 2 + s))) break;
    const uint8_t *src = (const uint8_t *)slot[s];
    pend[s] = std::async(std::launch::async, [=] {
      if (map) { memcpy(map + off, src, (size_t)m); return true; }
      return write_all(fd, src, m, off);
    });
  }
  for (int q = 0; q < NS; q++) if (pend[q].valid()) ok = pend[q].get() && ok;
  for (int q = 0; q < NS; q++) cmsbwt_host_free(slot[q]);
  if (map) { if (munmap(map, (size_t)out_bytes) != 0) ok = false; signal(SIGBUS, SIG_DFL); }
  return rc ? rc : (ok ? 0 : 1);
}

// ---- multi-GPU: whole documents sharded over GPUs 0..W-1 by cumulative length, one host
// thread per GPU driving its context; the exchanges are the library's (cmsbwt_sharded_after_ms over
// NCCL).  Every rank writes its slice of <out>.bwt / <out>.rl_bwt at its own offset.
static int run_multi_gpu(const Args &arg, const std::vector<uint8_t> &x, int64_t n, const int32_t *sa,
                         const std::vector<uint8_t> &text, int64_t N, std::ofstream &log) {
  const int W = arg.gpus;
  // contiguous ranges of whole documents with balanced symbol counts
  std::vector<int64_t> ends;
  for (int64_t i = 0; i < N; i++) if (text[(size_t)i] == 2) ends.push_back(i);
  const int64_t D = (int64_t)ends.size();
  if (D < W) { fprintf(stderr, "Error: more GPUs (%d) than documents (%lld)\n", W, (long long)D); return 1; }
  std::vector<int64_t> bound(W + 1, 0);
  {
    int64_t d = 0;
    for (int r = 1; r < W; r++) {
      const int64_t target = N * r / W;
      while (d < D && (d == 0 ? 0 : ends[(size_t)d - 1] + 1) + (ends[(size_t)d] - (d == 0 ? -1 : ends[(size_t)d - 1])) / 2 < target) d++;
      if (d <= bound[(size_t)r - 1]) d = bound[(size_t)r - 1] + 1;   // every rank gets at least one document
      if (d > D - (W - r)) d = D - (W - r);
      bound[r] = d;
    }
    bound[W] = D;
  }
  uint8_t id[128];
  if (cmsbwt_nccl_unique_id(id) != CMSBWT_OK) { fprintf(stderr, "Error: NCCL is not available (libnccl.so.2 / CMSBWT_NCCL)\n"); return 1; }
  const std::string outpath = arg.outname + (arg.format ? ".rl_bwt" : ".bwt");
  int fd = ::open(outpath.c_str(), O_RDWR | O_CREAT | O_TRUNC, 0644);
  if (fd < 0) { fprintf(stderr, "Error writing %s\n", outpath.c_str()); return 1; }
  std::vector<int> rcs((size_t)W, 0);
  std::vector<std::string> errs((size_t)W);
  std::vector<cmsbwt_stats> sts((size_t)W);
  std::vector<std::vector<double>> phases((size_t)W, std::vector<double>(16, 0.0));
  std::vector<uint64_t> colls((size_t)W, 0), xbytes((size_t)W, 0);
  const int mode = arg.format ? CMSBWT_MODE_RLE : CMSBWT_MODE_PLAIN;
  auto work = [&](int r) {
    cmsbwt_ctx *ctx = nullptr;
    auto failed = [&](const char *what, int rc) {
      errs[(size_t)r] = std::string(what) + ": " + (ctx ? cmsbwt_last_error(ctx) : "no context");
      rcs[(size_t)r] = rc ? rc : 1;
      // a rank that stops here would leave the others waiting inside a collective: end the process
      fprintf(stderr, "Error in %s (GPU %d): %s\n", what, r, ctx ? cmsbwt_last_error(ctx) : "no usable CUDA device");
      _exit(1);
    };
    int rc = cmsbwt_ctx_create(r, &ctx);
    if (rc) failed("ctx_create", rc);
    if ((rc = cmsbwt_index_upload(ctx, x.data(), (uint32_t)n, sa))) failed("index_upload", rc);
    if ((rc = cmsbwt_comm_init(ctx, id, (uint32_t)r, (uint32_t)W))) failed("comm_init", rc);
    const int64_t d0 = bound[(size_t)r], d1 = bound[(size_t)r + 1];
    const int64_t lo = d0 == 0 ? 0 : ends[(size_t)d0 - 1] + 1, hi = ends[(size_t)d1 - 1] + 1;
    if ((rc = cmsbwt_ms_run(ctx, text.data() + lo, (uint64_t)(hi - lo)))) failed("ms_run", rc);
    uint64_t ob = 0, ol = 0, tot = 0;
    if ((rc = cmsbwt_sharded_after_ms(ctx, (uint64_t)(d1 - d0), mode, &ob, &ol, &tot))) failed("sharded_after_ms", rc);
    uint64_t bytes = ol, off = ob;
    if (mode == CMSBWT_MODE_RLE) {
      const void *rec = nullptr; uint64_t nrec = 0, first = 0;
      if ((rc = cmsbwt_sharded_rle(ctx, &rec, &nrec, &first))) failed("sharded_rle", rc);
      bytes = 9 * nrec; off = 9 * first;
    }
    std::vector<uint8_t> out((size_t)bytes + 16);
    if (bytes && (rc = cmsbwt_result_download(ctx, out.data(), bytes))) failed("result_download", rc);
    if (bytes && !write_all(fd, out.data(), bytes, off)) { fprintf(stderr, "Error writing %s\n", outpath.c_str()); _exit(1); }
    cmsbwt_get_stats(ctx, &sts[(size_t)r]);
    cmsbwt_sharded_stats(ctx, &colls[(size_t)r], &xbytes[(size_t)r], phases[(size_t)r].data(), 16);
    cmsbwt_ctx_destroy(ctx);
  };
  std::vector<std::thread> th;
  for (int r = 1; r < W; r++) th.emplace_back(work, r);
  work(0);
  for (auto &t : th) t.join();
  if (::close(fd) != 0) { fprintf(stderr, "Error writing %s\n", outpath.c_str()); return 1; }
  static const char *PH[] = {"sizes", "tails", "x_tails", "rank_owner", "x_ranks", "refine", "tiles_local", "x_tiles",
                             "tiles_merge", "x_cov", "emit", "x_placements", "apply", "rle"};
  uint64_t heads = 0;
  for (int r = 0; r < W; r++) heads += sts[(size_t)r].n_heads;
  log << "Multi-GPU path on " << W << " GPUs (documents sharded; NCCL): headNumber: " << heads << ", collectives per job: "
      << colls[0] << ", exchanged bytes (rank 0): " << xbytes[0] << std::endl;
  for (int r = 0; r < W; r++) {
    log << "  GPU " << r << ": documents " << bound[(size_t)r] << ".." << bound[(size_t)r + 1] << ", MS " << sts[(size_t)r].ms_ms
        << " ms, phases";
    for (int k = 0; k < 14; k++) log << " " << PH[k] << "=" << phases[(size_t)r][(size_t)k];
    log << std::endl;
  }
  return 0;
}

int main(int argc, char **argv) {
  Args arg;
  parse_args(argc, argv, arg);

  std::ifstream infile(arg.filename);
  if (!infile) { fprintf(stderr, "Error opening file %s\n", arg.filename.c_str()); exit(1); }
  std::string line;
  std::getline(infile, line);
  if (line.empty()) { fprintf(stderr, "Error: first line of file %s is empty\n", arg.filename.c_str()); exit(1); }
  std::string refFileName = line;
  std::getline(infile, line);
  if (line.empty()) { fprintf(stderr, "Error: second line of file %s is empty\n", arg.filename.c_str()); exit(1); }
  std::string collFileName = line;
  collFileName.erase(collFileName.find_last_not_of(" \n\r\t") + 1);
  if (collFileName.empty()) { fprintf(stderr, "Error: second line of file %s is empty\n", arg.filename.c_str()); exit(1); }

  auto t1 = std::chrono::high_resolution_clock::now();
  std::ofstream log(arg.outname + ".log");
  std::cout << "==== CMS-BWT" << std::endl;
  std::cout << "==== For more information about the execution, please check the log file: "
            << arg.outname + ".log" << std::endl;
  std::cout << "==== Processing reference sequence" << std::endl;        // CMS-BWT-functions.cpp:1789

  // ---- device: context creation (0.6-2.9 s measured) starts now, beside the host work below ----
  cmsbwt_ctx *ctx = nullptr;
  int ctx_rc = 0;
  double ctx_ms = 0;
  std::thread ctx_thread, sa_thread;
  auto make_ctx = [&] {
    auto tc = std::chrono::high_resolution_clock::now();
    ctx_rc = cmsbwt_ctx_create(arg.device, &ctx);
    ctx_ms = ms_since(tc);
  };
  if (arg.gpus > 0) arg.stream = false;     // the multi-GPU path parses the whole collection first
  if (arg.stream) ctx_thread = std::thread(make_ctx);
  auto bail = [&](int code) {          // no thread may be left running when exit() tears the process down
    if (ctx_thread.joinable()) ctx_thread.join();
    if (sa_thread.joinable()) sa_thread.join();
    exit(code);
  };

  // ---- reference: read, trim, augment (:155-237) ----
  std::vector<uint8_t> reffile;
  if (!read_file(refFileName, reffile)) {
    fprintf(stderr, "Error opening file %s\n", refFileName.c_str());
    bail(1);
  }
  std::vector<uint8_t> ref(reffile.size() + 1);
  int64_t rlen = cmsbwt_read_reference(reffile.data(), (int64_t)reffile.size(), ref.data());
  if (rlen <= 0) { fprintf(stderr, "Error: reference %s is empty\n", refFileName.c_str()); bail(1); }
  log << "Reference (size = " << rlen << ") loaded" << std::endl;
  std::vector<uint8_t> x((size_t)rlen + 128);
  int64_t n = cmsbwt_augment(ref.data(), rlen, x.data());
  if (n < 0) { fprintf(stderr, "Error: reference holds a reserved byte (0, 1, 2 or > 127)\n"); bail(1); }
  if (n >= 0x7FFFFF00ll) { fprintf(stderr, "Error: reference longer than 2^31 symbols\n"); bail(1); }
  std::vector<uint8_t>().swap(reffile);
  std::vector<uint8_t>().swap(ref);

  // ---- collection: the reference's document list (:1138-1140,:1257-1275) ----
  Mapped coll;
  if (!coll.open(collFileName)) {
    fprintf(stderr, "Error opening file %s\n", collFileName.c_str());
    bail(1);
  }
  if (coll.size == 0) { fprintf(stderr, "Error: collection %s is empty\n", collFileName.c_str()); bail(1); }
  int T = arg.threads > 0 ? arg.threads : (int)std::thread::hardware_concurrency();
  if (T < 1) T = 1;
  if (T > 16 && arg.threads <= 0) T = 16;
  const bool streamed = arg.stream && arg.prefixLength >= (unsigned long long)coll.size;

  // ---- SA (:247) on its own host thread while the context comes up and the collection is parsed ----
  std::vector<int32_t> sa;
  const char *sa_route = "GPU prefix doubling";
  libsais_fn sais = arg.host_sa ? find_libsais(&sa_route) : nullptr;
  auto ts = std::chrono::high_resolution_clock::now();
  double sa_ms = 0;
  int sa_rc = 0;
  auto build_sa = [&] {
    sa_rc = sais(x.data(), sa.data(), (int32_t)n, 0, nullptr);
    sa_ms = ms_since(ts);
  };
  if (sais && arg.stream) {
    sa.resize((size_t)n);
    sa_thread = std::thread(build_sa);
  }

  if (arg.gpus > 0) {
    // ---- multi-GPU: SA on the host (libsais) or on GPU 0, whole parse, then one thread per GPU ----
    std::vector<int32_t> sa_m((size_t)n);
    if (sais) {
      ts = std::chrono::high_resolution_clock::now();
      if (sais(x.data(), sa_m.data(), (int32_t)n, 0, nullptr) != 0) { fprintf(stderr, "Error: libsais failed\n"); exit(1); }
      log << "Computed SA in " << ms_since(ts) << " ms (" << sa_route << ", host)" << std::endl;
    } else {
      cmsbwt_ctx *c0 = nullptr;
      if (cmsbwt_ctx_create(0, &c0) != CMSBWT_OK || cmsbwt_build_sa(c0, x.data(), (uint32_t)n, sa_m.data()) != CMSBWT_OK) {
        fprintf(stderr, "Error: no usable CUDA device 0 (cms_bwt has no CPU fallback)\n");
        exit(1);
      }
      cmsbwt_stats s0{};
      cmsbwt_get_stats(c0, &s0);
      log << "Computed SA in " << s0.sa_build_ms << " ms (GPU prefix doubling)" << std::endl;
      cmsbwt_ctx_destroy(c0);
    }
    std::vector<uint8_t> text_m((size_t)coll.size + 2);
    int64_t nd_m = 0;
    const int64_t N_m = cmsbwt_parse_collection_mt(coll.p, coll.size, arg.prefixLength, text_m.data(), &nd_m, T);
    coll.close();
    if (N_m <= 0) { fprintf(stderr, "Error: collection %s holds no document\n", collFileName.c_str()); exit(1); }
    log << "Collection parsed: " << N_m << " symbols, " << nd_m << " documents" << std::endl;
    std::cout << "==== Processing collection" << std::endl;
    const int rc_m = run_multi_gpu(arg, x, n, sa_m.data(), text_m, N_m, log);
    if (rc_m) exit(1);
    std::cout << "==== Time elapsed: "
              << std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::high_resolution_clock::now() - t1).count()
              << " ms" << std::endl;
    log.close();
    return 0;
  }

  // ---- device ----
  if (ctx_thread.joinable()) ctx_thread.join();
  else make_ctx();
  if (ctx_rc != CMSBWT_OK) {
    fprintf(stderr, "Error: no usable CUDA device %d (cms_bwt has no CPU fallback)\n", arg.device);
    bail(1);
  }
  log << "Device context created in " << ctx_ms << " ms" << (arg.stream ? " (beside the reference and the SA)" : "") << std::endl;
  auto die = [&](const char *what) {
    fprintf(stderr, "Error in %s: %s\n", what, cmsbwt_last_error(ctx));
    log << "Error in " << what << ": " << cmsbwt_last_error(ctx) << std::endl;
    if (sa_thread.joinable()) sa_thread.join();
    cmsbwt_ctx_destroy(ctx);
    exit(1);
  };
  if (sais && !arg.stream) {
    sa.resize((size_t)n);
    ts = std::chrono::high_resolution_clock::now();
    build_sa();
  }

  // ---- collection -> device text ----
  std::vector<uint8_t> text;
  int64_t ndocs = 0, N = 0;
  auto tp = std::chrono::high_resolution_clock::now();
  if (streamed) {
    N = ingest_streamed(ctx, coll.p, coll.size, T, &ndocs);
    if (N < 0) { if (sa_thread.joinable()) sa_thread.join(); die("ingest"); }
  } else {
    text.resize((size_t)coll.size + 2);
    N = cmsbwt_parse_collection_mt(coll.p, coll.size, arg.prefixLength, text.data(), &ndocs, arg.stream ? T : 1);
  }
  coll.close();
  if (N <= 0) {
    if (sa_thread.joinable()) sa_thread.join();
    fprintf(stderr, "Error: collection %s holds no document\n", collFileName.c_str());
    exit(1);
  }
  log << "Collection parsed" << (streamed ? " and copied to the device" : "") << " in " << ms_since(tp)
      << " ms: " << N << " symbols, " << ndocs << " documents (" << (arg.stream ? T : 1) << " threads)" << std::endl;

  // ---- device index (:263-297) ----
  if (sa_thread.joinable()) sa_thread.join();
  if (sais) {
    if (sa_rc != 0) { fprintf(stderr, "Error: libsais failed\n"); exit(1); }
    log << "Computed SA in " << sa_ms << " ms (" << sa_route << ", host" << (arg.stream ? ", beside the context creation and the ingest" : "") << ")" << std::endl;
  }
  auto ti = std::chrono::high_resolution_clock::now();
  if (cmsbwt_index_upload(ctx, x.data(), (uint32_t)n, sais ? sa.data() : nullptr) != CMSBWT_OK) die("index_upload");
  cmsbwt_stats s0{};
  cmsbwt_get_stats(ctx, &s0);
  if (!sais) log << "Computed SA in " << s0.sa_build_ms << " ms (GPU prefix doubling)" << std::endl;
  log << "Computed ISA, LCP, PLCP', reference BWT and k-mer table (k=" << s0.kmer_k << ") in "
      << s0.index_build_ms << " ms on the device" << std::endl;
  log << "Preprocessing done in " << ms_since(ts) << " ms (incl. host-device copies " << ms_since(ti)
      << " ms)" << std::endl;
  std::vector<int32_t>().swap(sa);

  std::cout << "==== Processing collection" << std::endl;                // :1791
  // ---- the path: MS -> head sort -> emit ----
  const int mode = arg.format ? CMSBWT_MODE_RLE : CMSBWT_MODE_PLAIN;
  auto tr = std::chrono::high_resolution_clock::now();
  if (streamed) { if (cmsbwt_ms_run_staged(ctx) != CMSBWT_OK) die("ms_run_staged"); }
  else if (cmsbwt_ms_run(ctx, text.data(), (uint64_t)N) != CMSBWT_OK) die("ms_run");
  std::vector<uint8_t>().swap(text);
  if (cmsbwt_heads_sort(ctx) != CMSBWT_OK) die("heads_sort");
  uint64_t out_bytes = 0;
  if (cmsbwt_emit(ctx, mode, &out_bytes) != CMSBWT_OK) die("emit");
  std::string outpath = arg.outname + (arg.format ? ".rl_bwt" : ".bwt");
  auto tw = std::chrono::high_resolution_clock::now();
  if (arg.stream) {
    g_outpath = outpath.c_str();
    int fd = ::open(outpath.c_str(), O_RDWR | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) { fprintf(stderr, "Error writing %s\n", outpath.c_str()); exit(1); }
    int rc = egress_streamed(ctx, fd, out_bytes);
    if (::close(fd) != 0 && rc == 0) rc = 1;
    if (rc < 0) die("result_read");
    if (rc > 0) { fprintf(stderr, "Error writing %s\n", outpath.c_str()); exit(1); }
  } else {
    std::vector<uint8_t> out((size_t)out_bytes);
    if (cmsbwt_result_download(ctx, out.data(), out_bytes) != CMSBWT_OK) die("result_download");
    FILE *fo = fopen(outpath.c_str(), "wb");
    if (!fo || (out_bytes && fwrite(out.data(), 1, (size_t)out_bytes, fo) != (size_t)out_bytes)) {
      fprintf(stderr, "Error writing %s\n", outpath.c_str());
      exit(1);
    }
    fclose(fo);
  }
  double write_ms = ms_since(tw);
  double path_ms = ms_since(tr);
  cmsbwt_stats st{};
  cmsbwt_get_stats(ctx, &st);
  log << "Time to compute matching statistics: " << st.ms_ms << " ms (kernel " << st.ms_kernel_ms
      << " ms, H2D " << st.h2d_ms << " ms)" << std::endl;
  log << "headNumber: " << st.n_heads << std::endl;
  log << "ndoc: " << st.n_docs << std::endl;
  log << "tails: " << st.n_tails << ", ranking rounds: " << st.rank_rounds << std::endl;
  log << "Done computing ranks and counterSmallerThanHead in " << st.sort_ms << " ms" << std::endl;
  log << "Done emitting in " << st.emit_ms << " ms (D2H " << st.d2h_ms << " ms)";
  if (mode == CMSBWT_MODE_RLE) log << ", runs: " << st.n_runs;
  log << std::endl;
  log << "Path (MS + head sort + emit + copies) wall: " << path_ms << " ms, kernel launches: "
      << st.kernel_launches << std::endl;

  log << "Result copied back and written in " << write_ms << " ms" << (arg.stream ? " (streamed)" : "") << std::endl;
  cmsbwt_ctx_destroy(ctx);

  std::cout << "==== Time elapsed: "
            << std::chrono::duration_cast<std::chrono::milliseconds>(
                   std::chrono::high_resolution_clock::now() - t1).count()
            << " ms" << std::endl;
  log.close();
  return 0;
}
