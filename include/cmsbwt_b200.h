/*
 * cmsbwt_b200 -- C ABI of the B200-native CMS-BWT hot path.
 *
 * The reference (fmasillo/CMS-BWT) exposes no FFI; its hot path sits behind one C++ call,
 *     void computeBWT(Args arg, std::string& refFile, std::string& collFile)      CMS-BWT.h:69
 * which runs initialize_reference (CMS-BWT-functions.cpp:151-310) and then
 * process_collection_{small,large}_reference (:312-1091 / :1093-1785) on file-scope globals.
 * This header is what a binding of that path would call instead: the three stages (matching
 * statistics, head sort, emit) plus the index upload that precedes them, with plain pointers
 * and sizes.  INTEGRATION.md shows the reference-side stub.
 *
 * Conventions
 *   - every function returns 0 on success or a negative CMSBWT_E_* code; nothing throws, nothing
 *     calls exit() (the reference exit()s on errors, main.cpp:91-115, :156-160,:648-652);
 *   - `x` is the AUGMENTED reference (reference :208-213 trim + :231-237 augmentation: every byte
 *     3..127 that does not occur, then 0x01, 0x00); n = its length (the reference's _n), n < 2^31;
 *   - `text` is the parsed collection: every document followed by the separator 0x02
 *     (reference :1138-1213: documents are closed by header / blank lines; _sx += 0x02);
 *     bytes must be in [3,127] or 0x02; N = its length (the reference's _sn worth of symbols);
 *   - host pointers unless the name says _device; device buffers passed in must be 16-byte
 *     aligned and have at least 64 readable bytes after their last element;
 *   - one context per GPU; calls on one context are serialised by the caller (the reference is
 *     single threaded); contexts on different GPUs may be driven from different host threads;
 *   - STREAM ORDERING: a context runs its kernels on its own non-blocking stream, which is ordered
 *     against no stream of the caller (not even the legacy default stream).  Every function that
 *     takes a device pointer produced by the caller (cmsbwt_ms_run_device and the cmsbwt_shard_*
 *     calls with d_* arguments) first synchronises the whole device, so all work launched on the
 *     device before the call -- on any stream -- has completed before the library reads the buffer;
 *     cmsbwt_ms_run_device_ordered replaces that host synchronisation by an event wait on the
 *     producer's stream.  Every function returns with the context's stream synchronised, so results
 *     (cmsbwt_result_device, cmsbwt_shard_buffer) may be read from any stream afterwards.
 */
#ifndef CMSBWT_B200_H
#define CMSBWT_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cmsbwt_ctx cmsbwt_ctx;

enum {
  CMSBWT_OK = 0,
  CMSBWT_E_CUDA = -1,        /* a CUDA call failed; see cmsbwt_last_error() */
  CMSBWT_E_ARG = -2,         /* bad argument / call order */
  CMSBWT_E_LIMIT = -3,       /* input exceeds an index-width limit of this build */
  CMSBWT_E_SYMBOL = -4,      /* text or reference holds a byte outside the allowed set */
  CMSBWT_E_INTERNAL = -5,    /* self-check failed (coverage total != N etc.) */
  CMSBWT_E_NOMEM = -6
};

enum { CMSBWT_MODE_PLAIN = 0, CMSBWT_MODE_RLE = 1 };

/* Stage timings (CUDA events on the context's stream), milliseconds, and counters.
 * Mirrors what the reference prints to <out>.log (e.g. :248-252 "Computed SA in", :618-622
 * "Time to compute matching statistics", :907-917). */
typedef struct {
  double sa_build_ms;       /* GPU suffix array (only when cmsbwt_index_upload got sa == NULL) */
  double index_build_ms;    /* ISA, PLCP', G, reference BWT, k-mer table         (:263-297) */
  double ms_ms;             /* stage 1: matching statistics + head compaction    (:1138-1213) */
  double sort_ms;           /* stage 2: linking, tail ranking, head keys, counting (:1359-1603) */
  double emit_ms;           /* stage 3: coverage->offsets, fill, head scatter, RLE (:1618-1777) */
  double h2d_ms, d2h_ms;
  double ms_kernel_ms;      /* the matching-statistics kernel alone (k_ms), last attempt */
  uint64_t n_heads;         /* the reference's headNumber (incl. one sentinel per document) */
  uint64_t n_docs;          /* the reference's D-1 */
  uint64_t n_tails;         /* distinct match-end text positions */
  uint64_t n_items;         /* (phrase, tile) incidences used by the counting step */
  uint64_t n_runs;          /* records written in RLE mode */
  uint32_t rank_rounds;     /* doubling rounds of the tail ranking */
  uint32_t kmer_k;
  uint64_t kernel_launches; /* launches issued by ms_run + heads_sort + emit of the last job */
  double pack_ms;           /* fast path: 2-bit packing of the text + separator list (inside ms_ms) */
  uint32_t fast_path;       /* 1 = the last ms_run used the 2-bit lane-per-window kernel */
  uint32_t light_tiles;     /* 1 = the last heads_sort counted every tile by brute force */
  uint64_t pair_tests;      /* head x (range | position mate) tests of that brute-force pass */
  uint64_t rank_tied;       /* tails whose reference ISA rank is shared (refined after the radix sort) */
  /* CUDA-event time of the kernel groups of the last job (single-GPU path), milliseconds; the sum is
   * ms_ms + sort_ms + emit_ms minus host gaps.  Indices: CMSBWT_G_* below. */
  double group_ms[12];
  uint64_t n_escapes;       /* text symbols outside A,C,G,T that the 2-bit kernel took through its break list */
} cmsbwt_stats;

enum {
  CMSBWT_G_PACK = 0,      /* 8-bit -> 2-bit text + separator list (k_pack, sort of separators)          */
  CMSBWT_G_MS = 1,        /* the matching-statistics kernel (k_ms_fast / k_ms)                           */
  CMSBWT_G_GATHER = 2,    /* head records -> text-order arrays (k_msf_fix, scan, k_msf_gather)           */
  CMSBWT_G_RANK = 3,      /* linking, tail labels, radix sort + tie refinement, kappa (:1359-1494)       */
  CMSBWT_G_TILES = 4,     /* tile filing + bucket sizes: histogram, scan, scatter, k_tile_cov (:1387-1416) */
  CMSBWT_G_OFFSETS = 5,   /* bucket sizes in SA order + offsets (k_vrank / k_cnt_sa, scan, k_fill_part; :1618-1628) */
  CMSBWT_G_FILL = 6,      /* run expansion of the reference symbols (k_fill; :1640-1688)                 */
  CMSBWT_G_PLACE = 7,     /* counterSmallerThanHead + head symbols (k_tile_place / sweep; :1517-1603)    */
  CMSBWT_G_RLE = 8,       /* 9-byte run records (:1695-1777)                                             */
  CMSBWT_G_COUNT = 9
};

/* replaces: process start-up of the reference (no equivalent; it has no device). */
int cmsbwt_ctx_create(int device_ordinal, cmsbwt_ctx **out);
void cmsbwt_ctx_destroy(cmsbwt_ctx *ctx);
const char *cmsbwt_last_error(const cmsbwt_ctx *ctx);
int cmsbwt_get_stats(const cmsbwt_ctx *ctx, cmsbwt_stats *out);

/* Host-side loaders with the reference's exact semantics (helpers for bindings and tests).
 * cmsbwt_augment: :208-213 + :231-237.  `out` must hold len + 127 bytes; returns n (or <0). */
int64_t cmsbwt_augment(const uint8_t *ref, int64_t len, uint8_t *out);
/* cmsbwt_read_reference: :171-199 (FASTA records concatenated / raw bytes). out: size bytes. */
int64_t cmsbwt_read_reference(const uint8_t *file, int64_t size, uint8_t *out);
/* cmsbwt_parse_collection: :1138-1140,:1257-1275 incl. the -p rule.  out_text: size+2 bytes.
 * Returns N; *ndocs = number of documents. */
int64_t cmsbwt_parse_collection(const uint8_t *file, int64_t size, uint64_t prefix,
                                uint8_t *out_text, int64_t *ndocs);

/* Parallel form of cmsbwt_parse_collection (same bytes, same *ndocs) on `threads` host threads
 * (<= 0: one per hardware thread, at most 32).  A -p prefix shorter than the file is parsed by the
 * sequential function (the cut depends on everything before it). */
int64_t cmsbwt_parse_collection_mt(const uint8_t *file, int64_t size, uint64_t prefix,
                                   uint8_t *out_text, int64_t *ndocs, int threads);
/* One piece of the same parse, for streamed ingest: the complete lines of file[begin, end) --
 * `begin` must be the start of a line; a last line without '\n' before `end` is not consumed
 * (the reference's getline().good() loop drops it at the end of the file, :1138-1140).  Writes the
 * text bytes to `out` (may be NULL to count only; needs at most end - begin bytes), adds the
 * documents closed inside the piece to *ndocs and sets *open to 1 / 0 when the piece's last line
 * was a sequence / a header or blank line (left untouched when the piece holds no complete line).
 * The caller appends one 0x02 (and one document) after the last piece when *open is 1 (:1269-1275).
 * No -p handling.  Returns the bytes written. */
int64_t cmsbwt_parse_piece(const uint8_t *file, int64_t begin, int64_t end, uint8_t *out,
                           int64_t *ndocs, int32_t *open);
/* first line start at or after `pos` (= pos when file[pos-1] is '\n' or pos is 0; size if none) */
int64_t cmsbwt_line_start(const uint8_t *file, int64_t size, int64_t pos);

/* replaces: libsais(x, sa, n, 0, NULL) at :247 when the caller has no host SA builder.
 * GPU prefix doubling; sa_out (host, n entries) may be NULL to keep it on the device only. */
int cmsbwt_build_sa(cmsbwt_ctx *ctx, const uint8_t *x, uint32_t n, int32_t *sa_out);

/* replaces: initialize_reference :246-297 after the SA: constructISA (:3-9, :263),
 * libsais_plcp/lcp (:270-271), PLCP' (:272-274), reference BWT (:294-297); the rmq_tree
 * (:284) is superseded by a k-mer interval table + exact search.
 * sa == NULL: the SA is built on the GPU (cmsbwt_build_sa). */
int cmsbwt_index_upload(cmsbwt_ctx *ctx, const uint8_t *x, uint32_t n, const int32_t *sa);
/* copies back index arrays for parity tests (any pointer may be NULL).
 * isa[n], plcp_prime[n] (= G[i]-i), ref_bwt[n], sa[n] */
int cmsbwt_index_download(cmsbwt_ctx *ctx, int32_t *sa, uint32_t *isa, uint32_t *plcp_prime,
                          uint8_t *ref_bwt);

/* replaces: stage 1, the per-document MS driver :1138-1213 (small path :344-419) with
 * computeMSFactorAt :45-88, binarySearchLB/RB :94-149, the PLCP' skip loop :1176-1185 and the
 * head records of match.h:5-16.  Host text is copied to the device first. */
int cmsbwt_ms_run(cmsbwt_ctx *ctx, const uint8_t *text, uint64_t N);
/* same, text already resident in device memory (not copied, must stay valid until emit).  The
 * whole device is synchronised before d_text is read (see STREAM ORDERING above). */
int cmsbwt_ms_run_device(cmsbwt_ctx *ctx, const void *d_text, uint64_t N);
/* same without the device-wide synchronisation: the library's stream waits for everything enqueued
 * so far on `producer_stream` (a cudaStream_t passed as void*; NULL = the legacy default stream). */
int cmsbwt_ms_run_device_ordered(cmsbwt_ctx *ctx, const void *d_text, uint64_t N, void *producer_stream);
/* parity hook: per-position matching statistics of the last ms_run, which
 * must have been preceded by cmsbwt_set_ms_dump(ctx, 1).  Arrays of N entries (may be NULL). */
int cmsbwt_set_ms_dump(cmsbwt_ctx *ctx, int enable);
int cmsbwt_ms_download(cmsbwt_ctx *ctx, uint32_t *pos, uint32_t *len, uint8_t *smaller,
                       uint8_t *is_head);

/* replaces: stage 2, :1359-1603 (head linking, phrase sort / keyed insertion, naming,
 * libsais_int :1456, next-rank substitution, prefix sums, counterSmallerThanHead). */
int cmsbwt_heads_sort(cmsbwt_ctx *ctx);
/* intermediate arrays for parity tests (text-order heads; any pointer may be NULL):
 * gstart[H], pos[H], len[H], within[H] (suffixes of the same bucket that precede the head),
 * offset[H] (final position in the BWT). */
int cmsbwt_heads_download(cmsbwt_ctx *ctx, uint64_t *gstart, uint32_t *pos, uint32_t *len,
                          uint64_t *within, uint64_t *offset);

/* replaces: stage 3, bucket inversion :1618-1628 and the writers :1631-1694 (plain) /
 * :1695-1777 (RLE: 9-byte records, LE uint64 run + uint8 symbol, maximal runs -- i.e. the
 * writer with its remainder flush after the sub-bucket loop).
 * Leaves the result on the device; *out_bytes = N (plain) or 9*R (RLE). */
int cmsbwt_emit(cmsbwt_ctx *ctx, int mode, uint64_t *out_bytes);
/* device pointer of the last emit's result (valid until the next ms_run / destroy). */
const void *cmsbwt_result_device(const cmsbwt_ctx *ctx);
/* copy the last emit's result to host memory (cap >= out_bytes). */
int cmsbwt_result_download(cmsbwt_ctx *ctx, void *host_out, uint64_t cap);

/* Streamed ingest / egress (replaces the getline loop :1138-1140 / :1257-1275 in
 * front of stage 1 and the 1 MiB write buffer of the writers :1640-1688 behind stage 3).  Copies
 * run on a second stream of the context, so the host can parse the next piece (or write the
 * previous range to the file) while a copy is in flight.  Host buffers handed to text_append /
 * result_read should come from cmsbwt_host_alloc (pinned) and must not be reused before
 * cmsbwt_copy_wait(slot) returned for a cmsbwt_copy_fence(slot) issued after the copy.
 *   text_begin(capacity)  device text buffer for up to `capacity` bytes
 *   text_append           async H2D of the next bytes of the text (in order)
 *   ms_run_staged         stage 1 on what was appended (= cmsbwt_ms_run on the same bytes)
 *   result_read           async D2H of result[offset, offset + bytes) after cmsbwt_emit */
enum { CMSBWT_COPY_SLOTS = 8 };
int cmsbwt_host_alloc(void **out, uint64_t bytes);
void cmsbwt_host_free(void *p);
int cmsbwt_text_begin(cmsbwt_ctx *ctx, uint64_t capacity);
int cmsbwt_text_append(cmsbwt_ctx *ctx, const uint8_t *chunk, uint64_t bytes);
int cmsbwt_ms_run_staged(cmsbwt_ctx *ctx);
int cmsbwt_result_read(cmsbwt_ctx *ctx, void *host_out, uint64_t offset, uint64_t bytes);
int cmsbwt_copy_fence(cmsbwt_ctx *ctx, int slot);
int cmsbwt_copy_wait(cmsbwt_ctx *ctx, int slot);

/* Whole path on host buffers: ms_run + heads_sort + emit + result_download.  This is the call
 * that stands where computeBWT's process_collection_* call stands (:1792-1796). */
int cmsbwt_run(cmsbwt_ctx *ctx, const uint8_t *text, uint64_t N, int mode, void *host_out,
               uint64_t cap, uint64_t *out_bytes);

/* ---- multi-GPU shard API (no reference counterpart: the reference is one thread) ----
 * One context per GPU, one process per GPU.  Documents are sharded whole (the MS state is reset per
 * document, reference :349-355 / :1143-1149), so cmsbwt_ms_run[_device] is called on the shard's
 * text as is.  Stages 2-3 then alternate these device steps with collectives that the host layer
 * owns (cms_bwt_b200/multi.py: torch.distributed over NCCL; the library itself links no NCCL):
 *
 *   shard_tails            link + tail labels (:1359-1385) + local sort by the 32-bit tail key
 *   [all-to-all of keys/cls by key range]
 *   shard_rank_owner       rank what was received (replaces naming + libsais_int :1432-1472 for the range)
 *   [all-to-all back]      shard_rank_apply
 *   loop: shard_tie_keys   [all-gather]  shard_tie_refine     (next-rank substitution :1480-1494)
 *   shard_tiles_local      kappa + tile entries of this shard's heads over the global tile space
 *   [all-to-all of entries by reference-position range]
 *   shard_tiles_merge      merged tiles of this device's position range + bucket sizes (:368-390)
 *   [all-gather of bucket sizes]
 *   shard_emit             offsets (:1618-1628), reference symbols of this device's output range
 *                          (:1640-1688), head placements of its position range
 *   [all-to-all of placements by output range]   shard_apply
 *
 * All pointers named d_* are device pointers on the context's GPU (16-byte aligned).  Every call
 * synchronises the context's stream before it returns. */
int cmsbwt_shard_tails(cmsbwt_ctx *ctx, uint64_t doc_base, uint32_t *n_tails);
/* named device arrays of the current shard state: "tail_keys" u32 (sorted), "tail_cls" u32,
 * "tie_keys" u64, "tile_heads"/"tile_ranges" 16-byte entries grouped by tile, "tile_hoff"/
 * "tile_roff" u32[world * tiles_per_rank + 1], "cov_slice" u32[tiles_per_rank * 256],
 * "placements" u64 (offset << 8 | symbol, grouped by destination rank), "result" u8,
 * "rle" u8 (9-byte records) */
int cmsbwt_shard_buffer(cmsbwt_ctx *ctx, const char *name, const void **dptr, uint64_t *count);
/* first index whose element is >= values[k], for the sorted buffer "tail_keys" */
int cmsbwt_shard_lower_bound(cmsbwt_ctx *ctx, const char *name, const uint64_t *values, uint32_t count,
                             uint64_t *out_idx);
int cmsbwt_shard_rank_owner(cmsbwt_ctx *ctx, const void *d_keys, const void *d_cls, uint32_t M,
                            uint32_t rank_base, void *d_rank_out, void *d_tie_out);
int cmsbwt_shard_rank_apply(cmsbwt_ctx *ctx, const void *d_rank_sorted, const void *d_tie_sorted,
                            uint32_t *m_tied);
int cmsbwt_shard_tie_keys(cmsbwt_ctx *ctx, uint32_t *m);
int cmsbwt_shard_tie_refine(cmsbwt_ctx *ctx, const void *d_allkeys, uint32_t m_total, uint32_t my_off,
                            uint32_t *m_tied);
int cmsbwt_shard_tiles_local(cmsbwt_ctx *ctx, uint32_t world, uint32_t *tiles_per_rank);
int cmsbwt_shard_tiles_merge(cmsbwt_ctx *ctx, uint32_t world, uint32_t rank, const void *d_he,
                             const void *d_re, const void *d_hoffs, const void *d_roffs,
                             const uint64_t *he_slab, const uint64_t *re_slab, uint64_t ndocs_total,
                             uint32_t *max_entries);
int cmsbwt_shard_emit(cmsbwt_ctx *ctx, const void *d_cov_full, uint64_t N_total, uint32_t world,
                      uint32_t rank, uint64_t doc_base, uint64_t *out_begin, uint64_t *out_len,
                      uint64_t *plc_counts /* [world]: placements per destination rank */);
int cmsbwt_shard_apply(cmsbwt_ctx *ctx, const void *d_plc, uint64_t count);
/* -r for a shard: maximal-run records (:1695-1777) of this device's slice of the BWT -> buffer "rle";
 * a run that straddles a device border is merged by the host layer (multi.rle_sharded). */
int cmsbwt_shard_rle(cmsbwt_ctx *ctx, uint64_t *n_records);

/* ---- the multi-GPU path behind ONE call (C++ host layer over NCCL, cms_bwt_b200/csrc/sharded.cu) ----
 * One rank per context.  The shard steps above, with every collective (all-to-all = grouped
 * ncclSend/ncclRecv, all-gather) issued by the library on the context's stream; NCCL is loaded at run
 * time (libnccl.so.2, or CMSBWT_NCCL=<path>).  No reference counterpart (the reference is one thread);
 * the stage order is the reference's (:1138-1213, :1359-1603, :1618-1777).
 *   cmsbwt_nccl_unique_id   128-byte ncclUniqueId made on one rank; the caller hands it to every rank
 *                           (torch.distributed / MPI broadcast, or shared memory between host threads)
 *   cmsbwt_comm_init        ncclCommInitRank on the context's device (collective: every rank calls it)
 *   cmsbwt_run_sharded      stage 1 on this rank's whole documents (d_text as in cmsbwt_ms_run_device,
 *                           ndocs_local = its separators), then stages 2-3 with the exchanges.  PLAIN:
 *                           this rank's slice of the BWT = bytes [*out_begin, *out_begin + *out_len) of
 *                           the whole output (*total_bytes long), in cmsbwt_shard_buffer("result"); the
 *                           slices follow each other in rank order and are NOT of equal length (a rank
 *                           emits the buckets of its range of SA ranks).
 *                           RLE: additionally the slice's records with border-straddling runs merged,
 *                           see cmsbwt_sharded_rle.  Collective: every rank must call it.
 *   cmsbwt_sharded_after_ms the same for a text already run through cmsbwt_ms_run / _staged
 *   cmsbwt_sharded_rle      device pointer / count of this rank's 9-byte records and the index of its
 *                           first record in the concatenated .rl_bwt
 *   cmsbwt_sharded_stats    collectives and exchanged bytes of the last job, wall time per phase (ms):
 *                           sizes, tails, x_tails, rank_owner, x_ranks, refine, tiles_local, x_tiles,
 *                           tiles_merge, x_cov, emit, x_placements, apply, rle, place, x_records (x_cov,
 *                           x_placements and apply are 0 unless the byte-range emit ran: deep tiles);
 *                           returns the phase count */
int cmsbwt_nccl_unique_id(void *out128);
int cmsbwt_comm_init(cmsbwt_ctx *ctx, const void *id128, uint32_t rank, uint32_t world);
int cmsbwt_comm_destroy(cmsbwt_ctx *ctx);
int cmsbwt_run_sharded(cmsbwt_ctx *ctx, const void *d_text, uint64_t N_local, uint64_t ndocs_local, int mode,
                       uint64_t *out_begin, uint64_t *out_len, uint64_t *total_bytes);
int cmsbwt_sharded_after_ms(cmsbwt_ctx *ctx, uint64_t ndocs_local, int mode, uint64_t *out_begin,
                            uint64_t *out_len, uint64_t *total_bytes);
int cmsbwt_sharded_rle(cmsbwt_ctx *ctx, const void **d_records, uint64_t *n_records, uint64_t *first_record);
int cmsbwt_sharded_stats(const cmsbwt_ctx *ctx, uint64_t *collectives, uint64_t *exchange_bytes, double *phase_ms,
                         uint32_t phase_cap);

/* ---- consumer of the output, fed from device memory (the reference stops at the files it
 * writes, :1631-1777) ----
 * cmsbwt_fm_build: FM-index over the plain BWT the last cmsbwt_emit left on the device (symbol counts +
 * one occurrence checkpoint per 256 symbols); *n_symbols = distinct symbols (at most 16).
 * cmsbwt_fm_count: batched backward search: counts[k] = occurrences of pattern k (bytes
 * patterns[offsets[k] .. offsets[k+1]), host arrays) in the whole collection; a pattern must not hold
 * the separator.  The index is valid until the next cmsbwt_ms_run* on the context. */
int cmsbwt_fm_build(cmsbwt_ctx *ctx, uint32_t *n_symbols);
int cmsbwt_fm_count(cmsbwt_ctx *ctx, const uint8_t *patterns, const uint64_t *offsets, uint32_t npat,
                    uint64_t *counts);
int cmsbwt_fm_free(cmsbwt_ctx *ctx);

/* ---- run-length records of a device-resident BWT without any allocation ----
 * No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Encodes d_bwt[0, N) (device memory) into maximal-run records (9 bytes each: little-endian u64 run
 * length, u8 symbol -- the .rl_bwt record of cmsbwt_emit(CMSBWT_MODE_RLE)) at d_records (device memory
 * of the caller, room for cap_records records; records past the capacity are dropped, the count is the
 * true one).  The number of records goes to *d_n_records (device, may be NULL) and, if n_records is
 * not NULL, to *n_records on the host -- that alone makes the call wait for the stream; with
 * n_records == NULL the call only enqueues three kernels.  stream: a cudaStream_t, NULL = the
 * context's own stream.  The call performs no cudaMalloc / cudaFree and calls no allocating library
 * routine: its scratch is a file-scope __device__ array of CMSBWT_RLE_STATIC_SCRATCH bytes that every
 * GPU which loads the library holds until exit, which bounds N at CMSBWT_RLE_STATIC_MAX symbols per
 * call (larger N: CMSBWT_E_LIMIT, nothing is launched).  All calls on one GPU share that array: a call
 * writes every entry it reads, so nothing has to be cleared in between, but calls that OVERLAP on
 * the GPU (different streams or host threads) race on it -- serialise them. */
#define CMSBWT_RLE_STATIC_MAX (1ull << 30)
#define CMSBWT_RLE_STATIC_SCRATCH (2 * (CMSBWT_RLE_STATIC_MAX / 4096) * 8 + 64)
int cmsbwt_rle_encode_static(cmsbwt_ctx *ctx, const void *d_bwt, uint64_t N, void *d_records,
                             uint64_t cap_records, uint64_t *d_n_records, uint64_t *n_records, void *stream);

/* tuning knobs: "chunk_len" (text window of an MS lane; 0 = by text length, 1024..8192), "kmer_k"
 * (order of the k-mer table, 0 = largest k <= 14 with 4^k <= n; takes effect at the next
 * index_upload), "fast" (0: 8-bit warp kernel for every input), "tile_budget" (0: sort-based sweep
 * instead of the tile kernels), "slots", "skip_min", "ms_occ", "scatter_mb", "nb_t1/4/8",
 * "ms_variant" (measurement switches; ms_variant bits: 2 text word cache, 8 G
 * prefetch, 16 empty k-mer buckets resolved from their slot; default 26).  Results never depend on
 * them. */
int cmsbwt_set_param(cmsbwt_ctx *ctx, const char *name, int64_t value);

#ifdef __cplusplus
}
#endif
#endif
