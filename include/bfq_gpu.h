/* bfq_gpu.h -- C ABI of the B200-native BFQzip data path (libbfqgpu_M<M>_B<B>.so).
 *
 * The reference has no in-process API for this path: BFQzip_ext.py runs two tools as
 * subprocesses and exchanges flat files (BFQzip_ext.py:173-177, 207-215, 310-318).  The
 * entry points below are what the two drop-in tools (tools/step1_main.cpp, tools/bfq_ext_main.cpp)
 * bind, and what a maintainer would bind from the reference's own C/C++ tools:
 *
 *   bfq_build_ebwt      replaces  gsufsort's load + cat_all + gsacak + store_to_disk
 *                       (external/gsufsort/main.c:280-311, 476-486, 848-895;
 *                        external/gsufsort/lib/utils.c:193-249; external/gsacak.c:2503-2518)
 *                       == eGap's three phases (external/egap/eGap:255-365) up to the EOS byte.
 *   bfq_smooth_rebuild  replaces  bfq_ext main phases 1-5 (src_ext_mem/bfq_ext.cpp:1067-1106):
 *                       buildFreq, splitIntoPartial, computeVectorUnbuildBCR, buildBitVectors,
 *                       run/process_cluster, BCRInverse (src_ext_mem/decode.cpp:238-317).
 *   bfq_pipeline        both, fused on the device (no file / host round trip in between).
 *
 * Compile-time modes: the reference selects the smoothing rule with -DM=0..3 and Illumina
 * binning with -DB=0|1 (src_ext_mem/makefile:10-22, bfq_ext.cpp:643-654, parameters.h:41-45),
 * one binary per mode.  Likewise there is one shared object per (M,B); bfq_mode_M/B report it.
 *
 * Conventions: plain C types only; every function returns 0 on success or a negative
 * BFQ_E* code, with a message available from bfq_last_error() (thread-local).  There is no
 * CPU fallback: without a CUDA device bfq_ctx_create fails with BFQ_ENODEVICE.
 * A bfq_ctx owns one device, one stream and a device arena; it is not thread-safe -- use one
 * ctx per host thread / per GPU (BFQzip_parallel.py:102-119 runs one tool process per block).
 */
#ifndef BFQ_GPU_H
#define BFQ_GPU_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BFQ_OK          0
#define BFQ_ENODEVICE  -1   /* no CUDA device / CUDA runtime error at start-up        */
#define BFQ_ECUDA      -2   /* CUDA error during a call                                */
#define BFQ_EINVAL     -3   /* bad argument (null pointer, size, offsets)              */
#define BFQ_EALPHABET  -4   /* more than 5 distinct non-EOS symbols, or a base byte <= EOS */
#define BFQ_ETOOLARGE  -5   /* collection does not fit this build's 32-bit row indices */
#define BFQ_EFREQ3     -6   /* >2 frequent symbols in a cluster: the reference asserts (bfq_ext.cpp:683) */
#define BFQ_ENOMEM     -7   /* device or host allocation failed                        */
#define BFQ_ECAPACITY  -8   /* caller-provided output capacity too small               */

typedef struct bfq_ctx bfq_ctx;

/* Run-time flags of bfq_ext (bfq_ext.cpp:105-129, 969-1031). */
typedef struct {
    int32_t K;          /* -k  minimum LCP inside a cluster            (default 16)   */
    int32_t m;          /* -m  minimum cluster length                  (default 2; the driver passes 5) */
    int32_t qthr;       /* -t  quality threshold for trusted bases     (default 20)   */
    int32_t max_len;    /* -l  maximum read length, <= 255             (driver passes 250) */
    float   fthr;       /* -f  percentage threshold for frequent bases (default 40)   */
    uint8_t eos_byte;   /* -s  end-marker byte: 35 (gsufsort) or 0 (eGap; what the driver passes) */
    uint8_t default_q;  /* -v  replacement quality, ASCII, used when M=2 (default '>') */
} bfq_params;

/* The statistics block bfq_ext prints (bfq_ext.cpp:1113-1128). */
typedef struct {
    uint64_t clusters;      /* "is the total number of clusters"      */
    uint64_t discarded;     /* "Discarded"                             */
    uint64_t ambiguous;     /* "Ambiguous discarded"                   */
    uint64_t processed;     /* "Processed"                             */
    uint64_t bases_inside;  /* "bases fall inside clusters"            */
    uint64_t qs_smoothed;   /* "qualities have been modified"          */
    uint64_t modified;      /* "bases have been modified"              */
    uint64_t nreads;        /* numSeq                                  */
    uint64_t n;             /* BWT length                              */
} bfq_stats;

int         bfq_mode_M(void);
int         bfq_mode_B(void);
const char *bfq_last_error(void);
void        bfq_params_default(bfq_params *p);          /* reference defaults, eos 35, max_len 250 */

int  bfq_ctx_create(bfq_ctx **out, int device);
void bfq_ctx_destroy(bfq_ctx *ctx);
/* Tunables (tests use them to force rarely-taken paths).  Unknown key -> BFQ_EINVAL.
 *   "initial_k"   symbols in the initial radix key (1..21, 0 = auto)
 *   "group_cap"   suffix groups with more members go to the next MSD level even when they fit a tile
 *   "batch_min_groups" a tile of the sorted suffix list with more groups than this (default 64: low coverage, a few rows per
 *                 group) is ordered in batches of whole groups, one warp per batch, instead of one warp per group; 0 = always
 *   "rebuild"     part C of the fused pipeline: 0 auto, 1 streaming un-permute through the suffix array, 2 inverse-LF walk
 *   "flat_min_rows" auto mode switches to the un-permute from this many BWT rows on
 *   "flat_min_rows_dev" the same threshold for device-resident calls (bfq_pipeline_dev), default 2^26: with nothing to copy out
 *                 while part C runs, the un-permute wins as soon as the walk lines outgrow L2
 *   "long_cluster" clusters longer than this many rows go to the block-per-cluster kernel
 *   "profile"     1 = record CUDA events around every kernel class (see bfq_prof_get)
 *   "host_slices" host-buffer calls (bfq_pipeline) launch the group kernel in this many slices and
 *                 copy every finished slice of .bwt/.bwt.qs/.1.lcp rows to the host while the next
 *                 one runs (default 16, 1 = one launch and one copy); the rebuilt reads leave the
 *                 same way, one GPU-full of reads at a time
 *   "trace"       1 = print a timeline of the stage boundaries of every bfq_pipeline call to stderr */
int  bfq_ctx_set_option(bfq_ctx *ctx, const char *key, int64_t value);
/* Run on a caller-owned CUDA stream (a cudaStream_t cast to void*) instead of the context's own,
 * so that the caller's events bracket the library's kernels.  NULL restores the own stream. */
int  bfq_ctx_set_stream(bfq_ctx *ctx, void *cuda_stream);

/* ---- host-buffer entry points (what the tools and the e2e benchmark call) ------------- */

/* Part A.  bases/quals: the reads concatenated without separators, off[nreads+1] their
 * boundaries.  Writes N = off[nreads] + nreads bytes to each of bwt, qs, lcp. */
int bfq_build_ebwt(bfq_ctx *ctx, const uint8_t *bases, const uint8_t *quals, const uint64_t *off,
                   uint64_t nreads, uint8_t eos_byte, uint8_t *bwt, uint8_t *qs, uint8_t *lcp);

/* Parts B+C.  out_bases/out_quals need n bytes of capacity, out_off off_capacity entries
 * (>= nreads+1, nreads = number of EOS bytes in bwt).  Reads come back in input order. */
int bfq_smooth_rebuild(bfq_ctx *ctx, const uint8_t *bwt, const uint8_t *qs, const uint8_t *lcp,
                       uint64_t n, const bfq_params *params,
                       uint8_t *out_bases, uint8_t *out_quals, uint64_t *out_off,
                       uint64_t off_capacity, bfq_stats *stats);

/* Parts A+B+C fused.  bwt/qs/lcp may be NULL (then they are not copied back).
 * out_bases/out_quals: off[nreads] bytes (read lengths are unchanged; offsets == off).
 * A read longer than params->max_len is refused with BFQ_EINVAL by every fused call (bfq_pipeline, bfq_pipeline_dev,
 * bfq_pipeline_fastq): the reference truncates such a read (decode.cpp:605), which would move every later offset;
 * bfq_build_ebwt + bfq_smooth_rebuild serve that case and return the new offsets. */
int bfq_pipeline(bfq_ctx *ctx, const uint8_t *bases, const uint8_t *quals, const uint64_t *off,
                 uint64_t nreads, const bfq_params *params,
                 uint8_t *bwt, uint8_t *qs, uint8_t *lcp,
                 uint8_t *out_bases, uint8_t *out_quals, bfq_stats *stats);

/* ---- device-buffer entry point (inputs already resident in HBM; what `value` times) --- */
/* All pointers are device pointers on ctx's device, except stats (host).  d_bwt/d_qs/d_lcp
 * may be NULL.  total_bases == off[nreads].  Runs on ctx's stream and synchronises it. */
int bfq_pipeline_dev(bfq_ctx *ctx, const uint8_t *d_bases, const uint8_t *d_quals,
                     const uint64_t *d_off, uint64_t nreads, uint64_t total_bases,
                     const bfq_params *params, uint8_t *d_bwt, uint8_t *d_qs, uint8_t *d_lcp,
                     uint8_t *d_out_bases, uint8_t *d_out_quals, bfq_stats *stats);

/* ---- FASTQ text in, FASTQ text out ----------------------------
 * The same three calls with the FASTQ file contents on either side instead of flat arrays: the
 * newline scan and field split of the loaders (external/gsufsort/lib/file.c:182-275 keeps line 2
 * and line 4 of every 4-line record; empty reads are dropped, lib/utils.c:229-236; a trailing
 * partial record is ignored) and the record writer of step 3
 * (src_ext_mem/decode.cpp:1080-1127: "@\nBASES\n+\nQUALS\n", header "@" alone -- use the flat calls
 * for -H header files) run on the device.  text: host pointer (pageable or pinned) to `size` bytes.
 * REPO-DEFINED, not the reference's behaviour: a '\r' before '\n' is stripped and a last line without '\n' is kept
 * whole.  The reference's loader overwrites the LAST byte of every line it reads (lib/file.c:210-212), so it keeps '\r'
 * as a symbol of the read and cuts the final symbol off an unterminated last line; parity is claimed on LF-terminated
 * files only.
 * bwt/qs/lcp hold `capacity` bytes each; size / 2 always suffices.  The FASTQ written for a
 * collection of R reads and T bases takes 2 T + 6 R bytes; `size` always suffices when it is the
 * text the reads came from.  Errors: BFQ_EINVAL for a record whose two payload lines differ in
 * length, BFQ_ECAPACITY when an output buffer is too small (the needed size is in the message). */
int bfq_build_ebwt_fastq(bfq_ctx *ctx, const uint8_t *text, uint64_t size, uint8_t eos_byte,
                         uint8_t *bwt, uint8_t *qs, uint8_t *lcp, uint64_t capacity,
                         uint64_t *nreads, uint64_t *n);
int bfq_smooth_rebuild_fastq(bfq_ctx *ctx, const uint8_t *bwt, const uint8_t *qs, const uint8_t *lcp,
                             uint64_t n, const bfq_params *params, uint8_t *fastq_out,
                             uint64_t capacity, uint64_t *fastq_size, uint8_t *dna_out,
                             uint8_t *qs_out, bfq_stats *stats);
/* dna_out / qs_out (both or NULL, n bytes each; section 8(f) row 2): the component streams step 4
 * makes with `sed -n 2~4p` / `sed -n 4~4p` (BFQzip_ext.py:19-21, 224-244) -- line 2 / line 4 of every
 * record of the FASTQ just written, one per line -- straight from the device buffers.  They take
 * at most n bytes (bases + one newline per read; fewer when -l truncates reads). */
/* bwt/qs/lcp may be NULL (then they are not produced on the host). */
int bfq_pipeline_fastq(bfq_ctx *ctx, const uint8_t *text, uint64_t size, const bfq_params *params,
                       uint8_t *bwt, uint8_t *qs, uint8_t *lcp, uint64_t capacity,
                       uint8_t *fastq_out, uint64_t fastq_capacity, uint64_t *fastq_size,
                       uint64_t *nreads, uint64_t *n, bfq_stats *stats);

/* ---- one collection over several GPUs (second row) ----------------
 * Every rank holds the whole collection (device pointers as in bfq_pipeline_dev) and builds the
 * slice [row_base, row_base + n_local) of the global eBWT / QS / LCP: the suffixes whose first four
 * symbols fall in the rank's range (ranges are derived identically on every rank, no exchange).
 * The slice is smoothed locally (exact for -k >= 4) and leaves as (text position, base | quality << 8)
 * pairs sorted by position; send_counts[d] of them belong to the rank that rebuilds text positions
 * [text_bounds[d], text_bounds[d + 1]).  After the all-to-all (the caller's: NCCL, MPI ...)
 * bfq_dist_place writes reads [read_lo, read_hi) -- text positions [text_lo, text_lo + count) -- in
 * input order.  Together they replace eGap + bfq_ext on one collection (BFQzip_ext.py:173-196).
 * d_bwt/d_qs/d_lcp/d_keys/d_vals hold `capacity` entries; BFQ_ECAPACITY if the slice is larger.
 * text_bounds, n_local, row_base, send_counts (world entries) and stats are host pointers;
 * stats counters are this rank's share (sum them over ranks). */
int bfq_dist_build(bfq_ctx *ctx, const uint8_t *d_bases, const uint8_t *d_quals, const uint64_t *d_off,
                   uint64_t nreads, uint64_t total_bases, const bfq_params *params, int rank, int world,
                   uint64_t capacity, const uint64_t *text_bounds, uint8_t *d_bwt, uint8_t *d_qs,
                   uint8_t *d_lcp, uint32_t *d_keys, uint16_t *d_vals, uint64_t *n_local,
                   uint64_t *row_base, uint64_t *send_counts, bfq_stats *stats);
int bfq_dist_place(bfq_ctx *ctx, uint32_t *d_keys, uint16_t *d_vals, uint64_t count, uint64_t text_lo,
                   const uint64_t *d_off, uint64_t read_lo, uint64_t read_hi, uint8_t *d_out_bases,
                   uint8_t *d_out_quals);

/* ---- introspection ---------------------------------------------------------------------- */
/* After a smooth_rebuild/pipeline call: the per-row state the reference keeps in
 * tmp/bwt_qs_<p>.aux and BWT_MOD/rbv (bfq_ext.cpp:555-556, 810-811), in global row order.
 * qs_out = stored qualities (after smoothing; after binning too when B=1);
 * obwt_out = output bases (bwt with noise-reduced rows replaced).  Host buffers of n bytes. */
int bfq_export_rows(bfq_ctx *ctx, uint8_t *qs_out, uint8_t *obwt_out);

/* Profile counters, filled when option "profile"=1.  Classes: "radix_pass", "radix_hist",
 * "pack", "keys", "group", "defer", "hist", "rank", "cluster", "cluster_long", "rule11",
 * "walk", "assemble", "unpermute" (the sort passes of the flat-mode rebuild).  Returns accumulated device ms, launch count and algorithmic bytes
 * since bfq_prof_reset.  Unknown class -> BFQ_EINVAL. */
int  bfq_prof_get(bfq_ctx *ctx, const char *klass, double *ms, uint64_t *launches, uint64_t *bytes);
void bfq_prof_reset(bfq_ctx *ctx);
/* number of kernel launches issued by this ctx since creation (bench.py's gpu_launches) */
uint64_t bfq_launch_count(bfq_ctx *ctx);

/* ---- quality-stream pre-transform for the host entropy coders --------------------
 * The stream of step 4 (BFQzip_ext.py:224-244: `sed -n 4~4p OUT.fq`, one line of qualities per read; what `bfq_ext -C`
 * writes as OUT.fq.qs) as two byte streams of equal length: run values and run lengths.  After smoothing the stream is
 * dominated by runs of the clusters' new quality; the reference hands it to 7z PPMd / bsc unchanged
 * (BFQzip_ext.py:248-272) -- this transform is optional, invertible and changes no file contract.
 * Defined on newline-separated lines of at most 255 symbols: a run never exceeds 255 bytes (BFQ_EINVAL otherwise).
 *   encode: qs[n] -> vals[*nruns], lens[*nruns]   (capacity = bytes available in each output; n always suffices)
 *   decode: the inverse; qs holds `capacity` bytes, *n = bytes written.   Host buffers; n and nruns below 2^32. */
int bfq_qs_rle_encode(bfq_ctx *ctx, const uint8_t *qs, uint64_t n, uint8_t *vals, uint8_t *lens, uint64_t capacity,
                      uint64_t *nruns);
int bfq_qs_rle_decode(bfq_ctx *ctx, const uint8_t *vals, const uint8_t *lens, uint64_t nruns, uint8_t *qs,
                      uint64_t capacity, uint64_t *n);

/* ---- allocation-free entry point ------------------------------------------------------- */
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * Stable LSD radix sort of (u64 key, u32 value) pairs on the low `bits` key bits -- the onesweep kernels of
 * bfq_debug_sort_pairs -- on DEVICE buffers of the CURRENT device, enqueued on `cuda_stream` (NULL = default
 * stream).  It needs no context, allocates nothing (no cudaMalloc / cudaFree / allocating library call) and does
 * not synchronise: the result is valid once the work enqueued on the stream has finished.
 *   d_keys_a / d_vals_a  input (overwritten); d_keys_b / d_vals_b  second ping-pong buffer, n entries each;
 *   *result_in_b = 0: the sorted pairs are in (a), 1: in (b)   (= number of 8-bit passes, ceil(bits/8), modulo 2).
 * LIMIT: n <= BFQ_NOALLOC_MAX_PAIRS.  The scratch of the passes (per-tile status words, tile counter, digit
 * histograms: 1 MiB + 8.25 KiB) is a fixed device array inside the library, held by every GPU that loads it until the
 * process exits.  Larger n -> BFQ_ECAPACITY (use the context-based calls).  ALL calls on a device share that array:
 * calls that overlap on the GPU (different streams, or several host threads) race on it -- serialise them. */
#define BFQ_NOALLOC_MAX_PAIRS (1u << 22)
int bfq_sort_pairs_noalloc(uint64_t *d_keys_a, uint64_t *d_keys_b, uint32_t *d_vals_a, uint32_t *d_vals_b,
                           uint64_t n, int bits, void *cuda_stream, int *result_in_b);

/* ---- test hooks (exercise one kernel family through the same ABI) ---------------------- */
/* Stable LSD radix sort of (key, value) pairs on the low `bits` bits; host buffers, in place. */
int bfq_debug_sort_pairs(bfq_ctx *ctx, uint64_t *keys, uint32_t *vals, uint64_t n, int bits);
/* Suffix array (text positions, terminators included) of the collection; sa has N entries. */
int bfq_debug_suffix_array(bfq_ctx *ctx, const uint8_t *bases, const uint64_t *off,
                           uint64_t nreads, uint32_t *sa);

#ifdef __cplusplus
}
#endif
#endif /* BFQ_GPU_H */
