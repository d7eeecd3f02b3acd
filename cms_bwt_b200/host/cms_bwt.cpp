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
// (--device N, --no-host-sa) so the reference's getopt string "p:b:o:rh" keeps its meaning.
//
// Reference SA: north-star keeps it on the host with libsais.  libsais is a third-party library
// that is not vendored in this repository; when a libsais shared object can be dlopen()ed
// (CMSBWT_LIBSAIS=<path>, or libsais.so on the loader path, or oracle/_ref/libsais.so next to the
// package) the SA is built with it (timed separately, "Computed SA in"); otherwise the SA is built
// on the GPU by cmsbwt_build_sa.
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
#include "cmsbwt_b200.h"

struct Args {                       // utils.h:8-14
  std::string filename, outname;
  bool format = false;              // 0 full BWT | 1 run-length encoded BWT
  unsigned buffer = 2;              // accepted for compatibility; device staging is sized from N
  unsigned long long prefixLength = UINT64_MAX;
  int device = 0;
  bool host_sa = true;
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
            << "\t--no-host-sa \tbuild the reference suffix array on the GPU even if libsais is available" << std::endl;
  exit(-1);
}

static void parse_args(int argc, char **argv, Args &arg) {
  puts("==== Command line:");
  for (int i = 0; i < argc; i++) printf(" %s", argv[i]);
  puts("");
  static const option longopts[] = {{"device", required_argument, nullptr, 1000},
                                    {"no-host-sa", no_argument, nullptr, 1001},
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
static libsais_fn find_libsais(const char *argv0) {
  std::vector<std::string> cands;
  if (const char *e = getenv("CMSBWT_LIBSAIS")) cands.push_back(e);
  cands.push_back("libsais.so");
  std::string self(argv0);
  size_t sl = self.find_last_of('/');
  std::string dir = sl == std::string::npos ? "." : self.substr(0, sl);
  cands.push_back(dir + "/../oracle/_ref/libsais.so");
  for (auto &p : cands) {
    void *h = dlopen(p.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!h) continue;
    if (void *s = dlsym(h, "libsais")) return (libsais_fn)s;
  }
  return nullptr;
}

static double ms_since(std::chrono::high_resolution_clock::time_point t) {
  return std::chrono::duration<double, std::milli>(std::chrono::high_resolution_clock::now() - t).count();
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

  // ---- reference: read, trim, augment (:155-237) ----
  std::vector<uint8_t> reffile, collfile;
  if (!read_file(refFileName, reffile)) {
    fprintf(stderr, "Error opening file %s\n", refFileName.c_str());
    exit(1);
  }
  std::vector<uint8_t> ref(reffile.size() + 1);
  int64_t rlen = cmsbwt_read_reference(reffile.data(), (int64_t)reffile.size(), ref.data());
  if (rlen <= 0) { fprintf(stderr, "Error: reference %s is empty\n", refFileName.c_str()); exit(1); }
  log << "Reference (size = " << rlen << ") loaded" << std::endl;
  std::vector<uint8_t> x((size_t)rlen + 128);
  int64_t n = cmsbwt_augment(ref.data(), rlen, x.data());
  if (n < 0) { fprintf(stderr, "Error: reference holds a reserved byte (0, 1, 2 or > 127)\n"); exit(1); }
  if (n >= 0x7FFFFF00ll) { fprintf(stderr, "Error: reference longer than 2^31 symbols\n"); exit(1); }
  std::vector<uint8_t>().swap(reffile);
  std::vector<uint8_t>().swap(ref);

  // ---- collection: the reference's document list (:1138-1140,:1257-1275) ----
  if (!read_file(collFileName, collfile)) {
    fprintf(stderr, "Error opening file %s\n", collFileName.c_str());
    exit(1);
  }
  if (collfile.empty()) { fprintf(stderr, "Error: collection %s is empty\n", collFileName.c_str()); exit(1); }
  std::vector<uint8_t> text(collfile.size() + 2);
  int64_t ndocs = 0;
  auto tp = std::chrono::high_resolution_clock::now();
  int64_t N = cmsbwt_parse_collection(collfile.data(), (int64_t)collfile.size(), arg.prefixLength,
                                      text.data(), &ndocs);
  std::vector<uint8_t>().swap(collfile);
  if (N <= 0) { fprintf(stderr, "Error: collection %s holds no document\n", collFileName.c_str()); exit(1); }
  log << "Collection parsed in " << ms_since(tp) << " ms: " << N << " symbols, " << ndocs
      << " documents" << std::endl;

  // ---- device ----
  cmsbwt_ctx *ctx = nullptr;
  if (cmsbwt_ctx_create(arg.device, &ctx) != CMSBWT_OK) {
    fprintf(stderr, "Error: no usable CUDA device %d (cms_bwt has no CPU fallback)\n", arg.device);
    exit(1);
  }
  auto die = [&](const char *what) {
    fprintf(stderr, "Error in %s: %s\n", what, cmsbwt_last_error(ctx));
    log << "Error in " << what << ": " << cmsbwt_last_error(ctx) << std::endl;
    cmsbwt_ctx_destroy(ctx);
    exit(1);
  };

  // ---- SA (:247) + device index (:263-297) ----
  std::vector<int32_t> sa;
  libsais_fn sais = arg.host_sa ? find_libsais(argv[0]) : nullptr;
  auto ts = std::chrono::high_resolution_clock::now();
  if (sais) {
    sa.resize((size_t)n);
    if (sais(x.data(), sa.data(), (int32_t)n, 0, nullptr) != 0) { fprintf(stderr, "Error: libsais failed\n"); exit(1); }
    log << "Computed SA in " << ms_since(ts) << " ms (libsais, host)" << std::endl;
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
  if (cmsbwt_ms_run(ctx, text.data(), (uint64_t)N) != CMSBWT_OK) die("ms_run");
  std::vector<uint8_t>().swap(text);
  if (cmsbwt_heads_sort(ctx) != CMSBWT_OK) die("heads_sort");
  uint64_t out_bytes = 0;
  if (cmsbwt_emit(ctx, mode, &out_bytes) != CMSBWT_OK) die("emit");
  std::vector<uint8_t> out((size_t)out_bytes);
  if (cmsbwt_result_download(ctx, out.data(), out_bytes) != CMSBWT_OK) die("result_download");
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

  std::string outpath = arg.outname + (arg.format ? ".rl_bwt" : ".bwt");
  FILE *fo = fopen(outpath.c_str(), "wb");
  if (!fo || (out_bytes && fwrite(out.data(), 1, (size_t)out_bytes, fo) != (size_t)out_bytes)) {
    fprintf(stderr, "Error writing %s\n", outpath.c_str());
    exit(1);
  }
  fclose(fo);
  cmsbwt_ctx_destroy(ctx);

  std::cout << "==== Time elapsed: "
            << std::chrono::duration_cast<std::chrono::milliseconds>(
                   std::chrono::high_resolution_clock::now() - t1).count()
            << " ms" << std::endl;
  log.close();
  return 0;
}
