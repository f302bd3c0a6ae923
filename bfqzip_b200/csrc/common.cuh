// common.cuh -- context, device arena, error plumbing and launch accounting shared by all
// translation units of libbfqgpu.  sm_100a only; no CPU fallback anywhere in this library.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include <map>
#include "../../include/bfq_gpu.h"

#ifndef BFQ_M
#define BFQ_M 2
#endif
#ifndef BFQ_B
#define BFQ_B 0
#endif

namespace bfq {

void set_error(const char *fmt, ...);

struct Error { int code; };   // thrown inside the library, caught at the C ABI

#define BFQ_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) {                                                        \
            bfq::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call,                 \
                           cudaGetErrorString(e__));                                     \
            throw bfq::Error{e__ == cudaErrorMemoryAllocation ? BFQ_ENOMEM : BFQ_ECUDA}; \
        }                                                                                \
    } while (0)

#define BFQ_FAIL(code, ...)          \
    do {                             \
        bfq::set_error(__VA_ARGS__); \
        throw bfq::Error{code};      \
    } while (0)

// ---------------------------------------------------------------- profiling classes
enum ProfClass {
    PC_RADIX_PASS = 0, PC_RADIX_HIST, PC_PACK, PC_KEYS, PC_GROUP, PC_DEFER, PC_HIST, PC_RANK,
    PC_CLUSTER, PC_CLUSTER_LONG, PC_RULE11, PC_WALK, PC_ASSEMBLE, PC_UNPERMUTE, PC_COUNT
};
static const char *const kProfNames[PC_COUNT] = {
    "radix_pass", "radix_hist", "pack", "keys", "group", "defer", "hist", "rank",
    "cluster", "cluster_long", "rule11", "walk", "assemble", "unpermute"};   // unpermute: the passes of the flat-mode rebuild sort

struct ProfEntry {
    double ms = 0; uint64_t launches = 0; uint64_t bytes = 0;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending;
};

// ---------------------------------------------------------------- device arena
// One growing cudaMalloc'd slab list with stack discipline: mark()/release(mark).
struct Arena {
    struct Slab { char *base; size_t size; size_t used; };
    std::vector<Slab> slabs;
    size_t min_slab = (size_t)256 << 20;
    void *alloc(size_t bytes);
    struct Mark { size_t slab; size_t used; };
    Mark mark() const;
    void release(Mark m);
    void destroy();
    size_t total() const { size_t t = 0; for (auto &s : slabs) t += s.size; return t; }
};

// ---- packed text (3-bit codes, 64 symbols per three u64 words, first symbol in the most significant bits)
#ifdef __CUDACC__
// ---- walk sectors: what one inverse-LF step reads.  32 bytes per 16 BWT rows -- ONE DRAM sector per step:
//   u16 rel[5]   occurrences of code 1..5 in the rows of this super-block before the sector
//   u16 plane[3] bit-sliced codes of the 16 rows
//   u8  q[16]    stored quality (bit 7 = "base modified", the base is then in obase[row])
// plus a table of u32[8] per super-block of 8192 rows (absolute occurrences before the block + pile start; cache resident).
constexpr int WS_SUPER_SHIFT = 13;
struct WalkSector { uint32_t rel[5]; uint32_t plane[3]; };
__device__ __forceinline__ WalkSector ws_load(const uint8_t *__restrict__ lines, uint64_t row)
{
    const uint4 a = *reinterpret_cast<const uint4 *>(lines + (row >> 4) * 32);
    WalkSector s;
    s.rel[0] = a.x & 0xFFFFu; s.rel[1] = a.x >> 16; s.rel[2] = a.y & 0xFFFFu; s.rel[3] = a.y >> 16; s.rel[4] = a.z & 0xFFFFu;
    s.plane[0] = a.z >> 16; s.plane[1] = a.w & 0xFFFFu; s.plane[2] = a.w >> 16;
    return s;
}
__device__ __forceinline__ uint32_t ws_code(const WalkSector &s, uint32_t r)          // r = row & 15
{
    return ((s.plane[0] >> r) & 1u) | (((s.plane[1] >> r) & 1u) << 1) | (((s.plane[2] >> r) & 1u) << 2);
}
__device__ __forceinline__ uint64_t ws_lf(const WalkSector &s, const uint32_t *__restrict__ super, uint64_t row, uint32_t c)   // c in 1..5
{
    const uint32_t r = (uint32_t)(row & 15);
    const uint32_t mm = ((c & 1) ? s.plane[0] : ~s.plane[0]) & ((c & 2) ? s.plane[1] : ~s.plane[1]) &
                        ((c & 4) ? s.plane[2] : ~s.plane[2]);
    return (uint64_t)super[(row >> WS_SUPER_SHIFT) * 8 + (c - 1)] + s.rel[c - 1] + __popc(mm & ((1u << r) - 1u));
}
__device__ __forceinline__ uint8_t *ws_q(uint8_t *lines, uint64_t row) { return lines + (row >> 4) * 32 + 16 + (row & 15); }
__device__ __forceinline__ const uint8_t *ws_q(const uint8_t *lines, uint64_t row) { return lines + (row >> 4) * 32 + 16 + (row & 15); }

// 21 symbols starting at text position p, in bits 63..1 (bit 0 cleared)
__device__ __forceinline__ uint64_t text_window(const uint64_t *__restrict__ T3, uint64_t p)
{
    const uint64_t bit = 3 * p;
    const uint64_t wi = bit >> 6;
    const int s = (int)(bit & 63);
    const uint64_t w0 = T3[wi], w1 = T3[wi + 1];
    uint64_t r = s ? ((w0 << s) | (w1 >> (64 - s))) : w0;
    return r & ~1ull;
}

// x / 3 for 0 <= x <= 127 (bit counts of a 64-bit word) as one multiply and one shift; the compiler's signed form costs six
__device__ __forceinline__ uint32_t div3_small(uint32_t x) { return (x * 86u) >> 8; }

// number of leading non-terminator symbols in a window (0..21)
__device__ __forceinline__ int leading_nonzero_fields(uint64_t w)
{
    const uint64_t M4 = 0x9249249249249248ull;   // top bit of each of the 21 fields (bits 63,60,...,3)
    const uint64_t M3 = 0x6DB6DB6DB6DB6DB6ull;   // low two bits of each field
    uint64_t t = ((w & M3) + M3) | w;            // field top bit set <=> field != 0
    uint64_t zeros = ~t & M4;
    const int lz = zeros ? __clzll((long long)zeros) : 63;
    return (int)div3_small((uint32_t)lz);
}

// the window with every symbol after its first terminator cleared (what `f = leading_nonzero_fields(w); if (f < 21)
// w &= f ? ~0ull << (64 - 3 * f) : 0` computes, without the division: the count of leading zeros IS 3 f)
__device__ __forceinline__ uint64_t cut_after_terminator(uint64_t w)
{
    const uint64_t M4 = 0x9249249249249248ull, M3 = 0x6DB6DB6DB6DB6DB6ull;
    const uint64_t zeros = ~(((w & M3) + M3) | w) & M4;
    if (!zeros) return w;
    const int lz = __clzll((long long)zeros);
    return lz ? (w & (~0ull << (64 - lz))) : 0ull;
}

__device__ __forceinline__ uint32_t sym_code(const uint64_t *__restrict__ T3, uint64_t p)
{
    const uint64_t bit = 3 * p;
    const uint64_t wi = bit >> 6;
    const int s = (int)(bit & 63);
    const uint64_t w0 = T3[wi];
    if (s <= 61) return (uint32_t)(w0 >> (61 - s)) & 7u;              // the symbol lies inside one word
    return (uint32_t)(((w0 << s) | (T3[wi + 1] >> (64 - s))) >> 61);
}
#endif

} // namespace bfq

struct bfq_ctx {
    int device = 0;
    int num_sms = 148;
    size_t l2_persist_max = 0;        // bytes of L2 that may be set aside for persisting lines
    size_t l2_window_max = 0;         // largest access-policy window
    cudaStream_t stream = nullptr;
    cudaStream_t own_stream = nullptr;
    // Host destinations of the current host-buffer call (bfq_pipeline): stages copy finished row / read ranges out on
    // the side stream while later stages still run, and say so in the *_copied flags.
    struct HostSink {
        uint8_t *bwt = nullptr, *qs = nullptr, *lcp = nullptr, *out_b = nullptr, *out_q = nullptr;
        const uint64_t *off = nullptr;       // host copy of the read offsets
        bool files_copied = false, out_copied = false;
    } *sink = nullptr;
    // option "trace": timed events at stage boundaries of a host-buffer call, printed to stderr when it ends
    int opt_trace = 0;
    int opt_wide_positions = 0;       // option "wide_positions": 64-bit text positions in single-collection mode even below 2^32 rows (tests)
    int opt_host_slices = 16;         // option "host_slices": slices of the group kernel in a host-buffer call (1 = off)
    std::vector<std::pair<const char *, cudaEvent_t>> trace_events;
    void trace(const char *name, cudaStream_t s)
    {
        if (!opt_trace) return;
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        cudaEventRecord(e, s);
        trace_events.emplace_back(name, e);
    }
    void trace_dump()
    {
        if (trace_events.empty()) return;
        cudaDeviceSynchronize();
        for (auto &p : trace_events) {
            float ms = 0;
            cudaEventElapsedTime(&ms, trace_events[0].second, p.second);
            fprintf(stderr, "[bfq trace] %8.3f ms  %s\n", ms, p.first);
        }
        for (auto &p : trace_events) cudaEventDestroy(p.second);
        trace_events.clear();
    }
    cudaStream_t copy_stream = nullptr;   // D2H of the part-A files overlaps parts B and C
    cudaEvent_t ev_a = nullptr, ev_copy = nullptr;
    bfq::Arena arena;
    // host pinned scratch for small read-backs
    void *h_small_dev = nullptr;      // device view of h_small
    uint64_t *h_small = nullptr;      // 4 KiB pinned (mapped: fetch_small writes it from a kernel)
    uint64_t *d_small = nullptr;      // 4 KiB device
    // options
    int opt_initial_k = 0;
    int64_t opt_long_cluster = 1024;
    int64_t opt_group_cap = 1 << 20;
    int64_t opt_batch_min_groups = 64;    // group_tile_kernel: tiles with more groups than this order them in batches
    int opt_rebuild = 0;                  // fused pipeline, part C: 0 auto, 1 streaming un-permute (flat), 2 inverse-LF walk
    int64_t opt_flat_min_rows = 1 << 28;  // auto: flat mode from this many rows on
    int64_t opt_flat_min_rows_dev = 1 << 26;   // the same for device-resident calls (bfq_pipeline_dev)
    int opt_profile = 0;
    int opt_bulk_tile = 1;                // group_tile_kernel phase 0 through cp.async.bulk (see sufsort.cu); 0 = per-thread loads
    int64_t opt_r11_cap = 0;              // test hook: first-attempt capacity of the two-base cluster list (0 = n / 8 + 1024)
    // accounting
    uint64_t launches = 0;
    bfq::ProfEntry prof[bfq::PC_COUNT];
    std::vector<cudaEvent_t> event_pool;
    // state kept after smooth/pipeline for bfq_export_rows
    const uint8_t *last_lines = nullptr;   // walk lines (device)
    const uint8_t *last_obase = nullptr;   // modified-base side array (device)
    const uint8_t *last_flat_bwt = nullptr; // flat mode: the BWT bytes the stored qualities belong to
    uint64_t last_n = 0;
    uint64_t base_hist[256] = {0};   // byte histogram of the reads of the last part-A call
    uint64_t qual_hist[256] = {0};   // ... and of their qualities (has_qual_hist), so that the fused pipeline does not histogram bwt / qs again
    bool has_qual_hist = false;
    uint8_t last_code2byte[8] = {0};
    bfq::Arena::Mark last_mark{0, 0};
    bool has_last = false;
};

namespace bfq {

// RAII profile scope: records two events around a kernel class when profiling is on,
// always counts the launch.
struct ProfScope {
    bfq_ctx *c; ProfClass k; cudaEvent_t e0 = nullptr, e1 = nullptr;
    ProfScope(bfq_ctx *ctx, ProfClass klass, uint64_t bytes, uint64_t nlaunch = 1);
    ~ProfScope();
};
void prof_collect(bfq_ctx *c);   // resolves pending event pairs (after a stream sync)

// Keeps [ptr, ptr + bytes) in the persisting part of L2 for the kernels launched on the context's stream while the
// object lives (as much of it as the set-aside takes: the hit ratio is scaled), everything else streams through.
struct L2Pin {
    cudaStream_t st = nullptr; bool on = false;
    L2Pin(bfq_ctx *c, const void *ptr, size_t bytes);
    ~L2Pin();
    L2Pin(const L2Pin &) = delete;
    L2Pin &operator=(const L2Pin &) = delete;
};

template <typename T> inline T *arena_alloc(bfq_ctx *c, size_t count)
{
    return (T *)c->arena.alloc(count * sizeof(T));
}

inline unsigned div_up(uint64_t a, uint64_t b) { return (unsigned)((a + b - 1) / b); }

// ---- pipeline stages (implemented across the .cu files) -----------------------------
struct Alphabet {
    int nsym = 0;                 // non-EOS symbols, <= 5
    uint8_t code2byte[8] = {0};   // code 0 = EOS byte, 1..nsym = symbols in byte order
    uint8_t byte2code[256];       // 0xFF = not in alphabet
    int code_N = -1;              // code of 'N' or -1
    uint8_t acgt_mask = 0;        // bit c set when code c is one of A,C,G,T
};

struct EbwtDev {                 // device results of part A
    uint8_t *bwt = nullptr, *qs = nullptr, *lcp = nullptr;
    uint64_t n = 0;
};

// ctx.cu
// Small results (counters, histograms) come back through a kernel that stores into the context's mapped pinned block
// instead of a device-to-host copy: a copy would queue on the copy engine BEHIND the bulk copies of the side stream
// and serialise the pipeline with them.  host_dst must lie inside c->h_small; bytes is a multiple of 4.
void fetch_small(bfq_ctx *c, void *host_dst, const void *dev_src, size_t bytes);

// sufsort.cu
struct KeyRange {                // single-collection mode: this rank builds only its slice of the global eBWT
    int rank = 0, world = 1;
    uint64_t capacity = 0;        // rows the output arrays can take
    uint64_t n_local = 0;         // out: rows of this rank
    uint64_t row_base = 0;        // out: global index of its first row
};
void build_ebwt_dev(bfq_ctx *c, const uint8_t *d_bases, const uint8_t *d_quals, const uint64_t *d_off,
                    uint64_t nreads, uint64_t total_bases, uint64_t max_len, uint8_t eos_byte,
                    uint8_t *d_bwt, uint8_t *d_qs, uint8_t *d_lcp, uint32_t *d_sa_out, uint64_t *d_t3_keep = nullptr,
                    KeyRange *range = nullptr);
void build_ebwt_dev_wide(bfq_ctx *c, const uint8_t *d_bases, const uint8_t *d_quals, const uint64_t *d_off,
                         uint64_t nreads, uint64_t total_bases, uint64_t max_len, uint8_t eos_byte,
                         uint8_t *d_bwt, uint8_t *d_qs, uint8_t *d_lcp, uint64_t *d_sa_out, uint64_t *d_t3_keep,
                         KeyRange *range);       // 64-bit text positions; key ranges only
size_t packed_text_words(uint64_t N);     // u64 words build_ebwt_dev needs for the packed text of N rows
// fastq.cu
struct DeviceReads {             // reads parsed on the device (arena memory)
    uint8_t *bases = nullptr, *quals = nullptr;
    uint64_t *off = nullptr;      // nreads + 1
    uint64_t nreads = 0, total = 0;
};
// FASTQ text (device) -> flat reads: line 2 / line 4 of every 4-line record, empty reads dropped, trailing partial record ignored
void fastq_ingest_dev(bfq_ctx *c, const uint8_t *d_text, uint64_t size, DeviceReads *out);
// flat reads -> FASTQ text with "@" headers; d_text holds 2 * total + 6 * nreads bytes
void fastq_emit_dev(bfq_ctx *c, const uint8_t *d_bases, const uint8_t *d_quals, const uint64_t *d_off, uint64_t nreads,
                    uint64_t total, uint8_t *d_text, uint8_t *d_dna = nullptr, uint8_t *d_qs = nullptr);
// d_dna / d_qs (both or none, total + nreads bytes each): line 2 / line 4 of every record, one per line
// rebuild.cu: exclusive scan u32 -> u64, out[n] = total
void scan_u32_u64_dev(bfq_ctx *c, const uint32_t *in, uint64_t n, unsigned long long *out);
// smooth.cu
struct SmoothResult {
    uint8_t *lines = nullptr;     // walk sectors: 32 B per 16 rows (see ws_load)
    const uint32_t *super = nullptr;   // u32[8] per 8192 rows: occurrences before the super-block + pile start
    uint8_t *obase = nullptr;     // n bytes, valid where q&0x80
    bool flat = false;            // lines is a plain per-row array of stored qualities (no walk lines were built)
    uint64_t n = 0, nreads = 0;
    Alphabet alpha;
};
// d_sa / d_t3 (both or none): suffix array of the rows and the packed text from part A -> flat mode
void smooth_dev(bfq_ctx *c, const uint8_t *d_bwt, const uint8_t *d_qs, const uint8_t *d_lcp,
                uint64_t n, const bfq_params *P, SmoothResult *out, bfq_stats *stats,
                const void *d_sa = nullptr, const uint64_t *d_t3 = nullptr, const uint64_t *alpha_hist = nullptr,
                bool sa_wide = false,       // sa_wide: d_sa holds u64 positions (else u32)
                const uint64_t *known_bwt_hist = nullptr, const uint64_t *known_qs_hist = nullptr);
// known_*_hist (fused pipeline): the byte histograms of d_bwt and d_qs, which part A knows -- the BWT holds the reads' bases plus one
// end-marker per read, the permuted qualities are the reads' qualities plus one end-marker byte per read -- so they are not recomputed.
// alpha_hist: byte histogram that fixes the alphabet (single-collection mode: the rows are a slice of the global eBWT
// and may miss a symbol); NULL = histogram of d_bwt itself, as buildFreq does.
// rebuild.cu
// d_len_known: per-read lengths offsets (nreads+1, u64) or NULL.  Writes reads in input order.
void rebuild_dev(bfq_ctx *c, const SmoothResult &S, const bfq_params *P,
                 const uint64_t *d_off_known, uint8_t *d_out_bases, uint8_t *d_out_quals,
                 uint64_t *d_out_off, uint64_t *total_out);
// Flat-mode rebuild: rows carry their text position (SA), so the smoothed (base, quality) of every row is put
// back into text order by a streaming sort on SA - 1 instead of an inverse-LF walk.  off = read offsets.
void unpermute_dev(bfq_ctx *c, const SmoothResult &S, const uint8_t *d_bwt, const uint32_t *d_sa,
                   const uint64_t *d_off, uint64_t nreads, uint8_t *d_out_bases, uint8_t *d_out_quals);
// Single-collection mode, sending side: (text position - text_bounds[d], base | quality << 8) of every local row, grouped by
// destination d; send_counts[d] = pairs whose position lies in [text_bounds[d], text_bounds[d + 1]).  n_total = rows of the global eBWT.
void emit_sorted_pairs_dev(bfq_ctx *c, const SmoothResult &S, const uint8_t *d_bwt, const void *d_sa, bool sa_wide, uint64_t n_total,
                           const uint64_t *text_bounds, int world, uint32_t *d_keys, uint16_t *d_vals, uint64_t *send_counts);
// receiving side: pairs keyed by (text position - t_lo), one per position of [t_lo, t_lo + cnt), in any order -> reads
// [read_lo, read_hi) in input order
void place_pairs_dev(bfq_ctx *c, uint32_t *d_keys, uint16_t *d_vals, uint64_t cnt, uint64_t t_lo, const uint64_t *d_off,
                     uint64_t read_lo, uint64_t read_hi, uint8_t *d_out_bases, uint8_t *d_out_quals);
// radix.cu
// Single-collection mode: the same, restricted to the suffixes whose key falls in the range of `rank` (ranges cut on
// the first four symbols from a histogram all ranks compute identically).  n_local = suffixes of this rank,
// row_base = rows of the global eBWT before them, prev_key / has_prev = key of the row just before the range.
template <typename V>        // V = position type, u32 or u64 (collections of 2^32 rows and more)
int radix_sort_suffix_keys_range(bfq_ctx *c, const uint64_t *T3, const uint16_t *QL, uint64_t n, int k, int rank, int world,
                                 uint64_t *keys_a, uint64_t *keys_b, V *pos_a, V *pos_b, uint64_t capacity,
                                 uint64_t *n_local, uint64_t *row_base, uint64_t *prev_key, bool *has_prev);
// Sorts (key, 16-bit payload) pairs whose keys are exactly the integers 0..n-1 in some order (applying a
// permutation with streaming passes instead of a random scatter); the digit histograms are computed on the host.
// Only the VALUES come out in key order (vals_a when the returned flag is 0, vals_b when 1): the passes cover the high key bits
// and a placement pass in shared memory finishes the low 15 (a bucket of a permutation sits exactly where it belongs).
// src (optional, only when radix_permutation_generates(n)): the first pass reads the smoothed rows instead of keys_a / vals_a.
struct PermSource { const uint32_t *SA; const uint8_t *bwt, *qflat, *obase; uint32_t n_total; };
bool radix_permutation_generates(uint64_t n);
// verify: the keys come from outside the library (bfq_dist_place) -- check that they are a permutation of 0..n-1 (BFQ_EINVAL)
int radix_sort_permutation(bfq_ctx *c, uint32_t *keys_a, uint32_t *keys_b, uint16_t *vals_a, uint16_t *vals_b, uint64_t n,
                           const PermSource *src = nullptr, bool verify = false);
// Level-0 keys of the suffix sort: k symbols of every text position (cut at the terminator, QL[p] in the
// unsorted bits 63..48 when given) with payload = position, built and histogrammed in one kernel, then sorted
// on the 3k symbol bits.  Same result convention as radix_sort_pairs64.
int radix_sort_suffix_keys(bfq_ctx *c, const uint64_t *T3, const uint16_t *QL, uint64_t n, int k, uint64_t *keys_a,
                           uint64_t *keys_b, uint32_t *pos_a, uint32_t *pos_b);
// Stable LSD sort of (key,val) on key bits [0,bits).  Result ends in keys_a/vals_a when the
// returned flag is 0, in keys_b/vals_b when 1 (ping-pong buffers, both n long).
int radix_sort_pairs64(bfq_ctx *c, uint64_t *keys_a, uint64_t *keys_b, uint32_t *vals_a,
                       uint32_t *vals_b, uint64_t n, int bits);
int radix_sort_pairs64(bfq_ctx *c, uint64_t *keys_a, uint64_t *keys_b, uint64_t *vals_a,
                       uint64_t *vals_b, uint64_t n, int bits);
int radix_sort_pairs32(bfq_ctx *c, uint32_t *keys_a, uint32_t *keys_b, uint32_t *vals_a,
                       uint32_t *vals_b, uint64_t n, int bits);
int radix_sort_pairs32_16(bfq_ctx *c, uint32_t *keys_a, uint32_t *keys_b, uint16_t *vals_a,
                          uint16_t *vals_b, uint64_t n, int bits);

} // namespace bfq
