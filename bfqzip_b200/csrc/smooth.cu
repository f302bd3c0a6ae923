// smooth.cu -- part B on the device: rank/occ "walk lines", positional-cluster detection from the
// LCP array, quality smoothing and noise reduction.
//
// Replaces: src_ext_mem/bfq_ext.cpp buildFreq :132-187,
// splitIntoPartial/tableOcc :190-348, buildBitVectors :350-412, run :837-948,
// process_cluster :581-824, modBasesSmoothQS :546-573, buildLF_Cluster :415-476,
// max_qs/mean_error/avg_qs :479-543; src_ext_mem/decode.cpp computeVectorUnbuildBCR :87-235;
// external/rankbv (the modified-row bitvector becomes bit 7 of the stored quality byte).
//
// The compile-time modes of the reference (-DM=0..3, -DB=0|1, src_ext_mem/makefile:10-22) are
// compile-time here too: BFQ_M / BFQ_B.
//
// HBM layout (see ws_load in common.cuh).  One 32-byte WALK SECTOR per 16 BWT rows:
//     u16 rel[5]   occurrences of codes 1..5 in the rows of this super-block (8192 rows) before the sector
//     u16 plane[3] bit b of the 3-bit code of each of the 16 rows (row r = bit r)
//     u8  q[16]    stored quality of each row (smoothed, binned when B=1); bit 7 set when the
//                  output base differs from the BWT symbol (then obase[row] holds it)
// plus u32[8] per super-block (absolute occurrences before it + pile start C[code]; cache resident), so one
// inverse-LF step (part C) touches exactly one aligned 32-byte DRAM sector.  Flat mode (fused pipeline with the
// suffix array at hand) builds no sectors at all: the stored qualities are a plain per-row array.
#include "common.cuh"
#include <math.h>
#include <algorithm>
#include <mutex>

namespace bfq {

void byte_histogram(bfq_ctx *c, const uint8_t *d_data, uint64_t n, uint64_t hist_out[256]);
Alphabet alphabet_from_hist(const uint64_t hist[256], uint8_t eos, bool eos_may_appear);

// ------------------------------------------------------------------ constants for a run
struct SmoothConst {
    uint8_t byte2code[256];
    uint8_t code2byte[8];
    int32_t K, m, qthr;
    float fthr;
    int32_t fT;          // ceil(fthr): (uchar)(100 * f / n) >= fthr  <=>  100 * f >= fT * n  (the left side is an integer)
    uint8_t default_q;
    int8_t code_N;
    uint8_t acgt_mask;
    uint8_t nsym;
    uint32_t long_len;
};
// SmoothConst travels as a kernel argument (__grid_constant__): it depends on the call (flags, alphabet of THIS
// collection), and several contexts may run on one device.  The two M=1 tables are pure functions of nothing: they
// are module constants, written once per device.
__constant__ double g_errtab[256];     // pow(10, -(q-33)/10), host glibc (bfq_ext.cpp:532)
__constant__ double g_m1thr[128];      // g_m1thr[v] = largest x with round(-10*log10(x)) >= v
#define SC_ARG const __grid_constant__ SmoothConst g_sc

// decode.cpp:72-84 on an ASCII quality (the caller's `foundQual-33` wraps in uchar like the reference)
__device__ __forceinline__ uint8_t bin8(uint8_t ascii)
{
    uint8_t q = (uint8_t)(ascii - 33);
    if (q >= 40) q = 40;
    else if (q >= 35) q = 37;
    else if (q >= 30) q = 33;
    else if (q >= 25) q = 27;
    else if (q >= 20) q = 22;
    else if (q >= 10) q = 15;
    else if (q >= 2) q = 6;
    return (uint8_t)(q + 33);
}
// quality byte as stored in a walk line: bit 7 = "base modified", low 7 bits = emitted quality
__device__ __forceinline__ uint8_t store_q(uint8_t q_with_flag)
{
#if BFQ_B
    return (uint8_t)(bin8(q_with_flag & 0x7F) | (q_with_flag & 0x80));
#else
    return q_with_flag;
#endif
}

// ------------------------------------------------------------------ walk lines: counts
constexpr int RK_THREADS = 256;
constexpr int RK_ROWS = RK_THREADS * 32;     // rows per block (one 32-row group per thread)

__device__ __forceinline__ void load_group_codes(const uint8_t *__restrict__ bwt, uint64_t row0,
                                                 uint64_t n, const uint8_t *s_lut, uint32_t plane[3],
                                                 uint32_t cnt[5])
{
    plane[0] = plane[1] = plane[2] = 0;
#pragma unroll
    for (int k = 0; k < 5; k++) cnt[k] = 0;
    if (row0 >= n) return;
    uint8_t b[32];
    if (row0 + 32 <= n && ((reinterpret_cast<uintptr_t>(bwt) + row0) & 15) == 0) {
        const uint4 *v = reinterpret_cast<const uint4 *>(bwt + row0);
        uint4 x0 = v[0], x1 = v[1];
        uint32_t w[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
#pragma unroll
        for (int r = 0; r < 32; r++) b[r] = (uint8_t)(w[r >> 2] >> (8 * (r & 3)));
    } else {
#pragma unroll
        for (int r = 0; r < 32; r++) b[r] = (row0 + r < n) ? bwt[row0 + r] : 0;
    }
#pragma unroll
    for (int r = 0; r < 32; r++) {
        uint32_t code = (row0 + r < n) ? s_lut[b[r]] : 0u;
        plane[0] |= (code & 1u) << r;
        plane[1] |= ((code >> 1) & 1u) << r;
        plane[2] |= ((code >> 2) & 1u) << r;
    }
#pragma unroll
    for (int k = 0; k < 5; k++) {
        uint32_t c = k + 1;
        uint32_t mm = ((c & 1) ? plane[0] : ~plane[0]) & ((c & 2) ? plane[1] : ~plane[1]) &
                      ((c & 4) ? plane[2] : ~plane[2]);
        cnt[k] = __popc(mm);
    }
}

__global__ void __launch_bounds__(RK_THREADS)
rank_count_kernel(SC_ARG, const uint8_t *__restrict__ bwt, uint64_t n, uint32_t *__restrict__ blk_counts)
{
    __shared__ uint8_t s_lut[256];
    __shared__ uint32_t s_cnt[5];
    s_lut[threadIdx.x] = g_sc.byte2code[threadIdx.x];
    if (threadIdx.x < 5) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    uint32_t plane[3], cnt[5];
    load_group_codes(bwt, (uint64_t)blockIdx.x * RK_ROWS + (uint64_t)threadIdx.x * 32, n, s_lut, plane, cnt);
#pragma unroll
    for (int k = 0; k < 5; k++) {
        uint32_t v = cnt[k];
#pragma unroll
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(&s_cnt[k], v);
    }
    __syncthreads();
    if (threadIdx.x < 5) blk_counts[(size_t)blockIdx.x * 8 + threadIdx.x] = s_cnt[threadIdx.x];
}

// single block: exclusive scan over blocks for each code, plus the pile start C[code] (the five codes side by side)
__global__ void __launch_bounds__(1024)
rank_scan_kernel(uint32_t *__restrict__ blk_counts, uint32_t nblk, const uint32_t *__restrict__ pile_start)
{
    __shared__ uint32_t s_warp[32][5];
    __shared__ uint32_t s_carry[5];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid < 5) s_carry[tid] = pile_start[tid];
    __syncthreads();
    for (uint32_t base = 0; base < nblk; base += 1024) {
        const uint32_t i = base + tid;
        uint32_t v[5], x[5];
#pragma unroll
        for (int k = 0; k < 5; k++) {
            v[k] = i < nblk ? blk_counts[(size_t)i * 8 + k] : 0u;
            x[k] = v[k];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x[k], o); if (lane >= o) x[k] += y; }
            if (lane == 31) s_warp[w][k] = x[k];
        }
        __syncthreads();
        uint32_t tot[5];
#pragma unroll
        for (int k = 0; k < 5; k++) {
            uint32_t add = 0;
            for (int j = 0; j < w; j++) add += s_warp[j][k];
            tot[k] = s_carry[k] + add + x[k];
            if (i < nblk) blk_counts[(size_t)i * 8 + k] = tot[k] - v[k];
        }
        __syncthreads();
        if (tid == 1023) {
#pragma unroll
            for (int k = 0; k < 5; k++) s_carry[k] = tot[k];
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(RK_THREADS)
rank_build_kernel(SC_ARG, const uint8_t *__restrict__ bwt, uint64_t n, uint8_t *__restrict__ lines)
{
    __shared__ uint8_t s_lut[256];
    __shared__ uint32_t s_warp[RK_THREADS / 32][5];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    s_lut[tid] = g_sc.byte2code[tid];
    __syncthreads();
    const uint64_t row0 = (uint64_t)blockIdx.x * RK_ROWS + (uint64_t)tid * 32;
    uint32_t plane[3], cnt[5], excl[5];
    load_group_codes(bwt, row0, n, s_lut, plane, cnt);
#pragma unroll
    for (int k = 0; k < 5; k++) {
        uint32_t x = cnt[k];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) s_warp[w][k] = x;
        excl[k] = x - cnt[k];
    }
    __syncthreads();
    if (row0 >= n) return;
    uint32_t lo[5], hi[5];                      // occurrences inside this block before each of the thread's two sectors
#pragma unroll
    for (int k = 0; k < 5; k++) {
        uint32_t add = 0;
        for (int j = 0; j < w; j++) add += s_warp[j][k];
        const uint32_t c = k + 1;
        const uint32_t mm = ((c & 1) ? plane[0] : ~plane[0]) & ((c & 2) ? plane[1] : ~plane[1]) & ((c & 4) ? plane[2] : ~plane[2]);
        lo[k] = excl[k] + add;
        hi[k] = lo[k] + __popc(mm & 0xFFFFu);
    }
    uint8_t *S = lines + (row0 >> 4) * 32;
    *reinterpret_cast<uint4 *>(S) = make_uint4(lo[0] | (lo[1] << 16), lo[2] | (lo[3] << 16), lo[4] | ((plane[0] & 0xFFFFu) << 16),
                                               (plane[1] & 0xFFFFu) | ((plane[2] & 0xFFFFu) << 16));
    *reinterpret_cast<uint4 *>(S + 32) = make_uint4(hi[0] | (hi[1] << 16), hi[2] | (hi[3] << 16), hi[4] | ((plane[0] >> 16) << 16),
                                                    (plane[1] >> 16) | ((plane[2] >> 16) << 16));
}

// ------------------------------------------------------------------ line access helpers
// where the stored quality of a row lives: inside its walk line, or (FLAT, fused pipeline with the suffix
// array at hand: no walk lines are built) in a plain per-row array
template <bool FLAT> __device__ __forceinline__ uint8_t *q_slot(uint8_t *lines, uint64_t row)
{
    return FLAT ? lines + row : ws_q(lines, row);
}

// ------------------------------------------------------------------ cluster rules (shared)
// Device counters: 0 clusters 1 discarded 2 ambiguous 3 processed 4 bases_inside
//                  5 qs_smoothed 6 modified 7 error(>2 frequent symbols)
enum { ST_CLUSTERS = 0, ST_DISCARDED, ST_AMBIGUOUS, ST_PROCESSED, ST_BASES, ST_SMOOTHED, ST_MODIFIED, ST_ERR, ST_COUNT };

struct ClAcc {                     // per-cluster accumulators (rule 8, bfq_ext.cpp:612-654)
    uint32_t freq[6];              // by code; [0] unused
    uint32_t hiq;                  // bit c: some row with code c has q >= qthr+33 ("lowQS" in the reference)
    uint32_t base_num;
    uint32_t mx;                   // M=0
    int32_t isum;                  // M=3 (int, wraps like the reference's)
    double dsum;                   // M=1, added in row order
};

enum { ACT_NONE = 0, ACT_RULE10 = 1, ACT_RULE11 = 2 };
struct Decision { int act; uint32_t sym, f0, f1; uint8_t newqs; };

__device__ __forceinline__ uint8_t newqs_from(const SmoothConst &g_sc, const ClAcc &a)
{
#if BFQ_M == 0
    return (uint8_t)a.mx;                                            // bfq_ext.cpp:502-518
#elif BFQ_M == 1
    // bfq_ext.cpp:537-541: (uchar)round(-10*log10(sum/num)) + 33, through host-built thresholds
    double avg = __ddiv_rn(a.dsum, (double)a.base_num);
    int lo = 0, hi = 127;                                            // largest v with avg <= thr[v]
    while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (avg <= g_m1thr[mid]) lo = mid; else hi = mid - 1; }
    return (uint8_t)(lo + 33);
#elif BFQ_M == 2
    return g_sc.default_q;                                           // bfq_ext.cpp:648
#else
    if (a.isum == 0) return 0;                                       // bfq_ext.cpp:495-496
    float f = __fdiv_rn((float)a.isum, (float)(unsigned long long)a.base_num);
    return (uint8_t)(int)floor((double)f + 0.5);
#endif
}

// rules 7-9 (bfq_ext.cpp:584, 633, 665-713).  st[] are this thread's counters.
__device__ __forceinline__ Decision decide(const SmoothConst &g_sc, uint32_t len, const ClAcc &a, uint32_t *st)
{
    Decision d{ACT_NONE, 0, 0, 0, 0};
    if ((int)len < g_sc.m) return d;
    if (a.base_num == 0) return d;
    st[ST_CLUSTERS]++;
    st[ST_BASES] += a.base_num;
    d.newqs = newqs_from(g_sc, a);
    uint32_t F[3]; int nF = 0;
#pragma unroll
    for (uint32_t c = 1; c <= 5; c++) {
        if (a.freq[c] > 0) {
            // bfq_ext.cpp:665-677: perc = (uchar)((100 * freq) / base_num), frequent when perc >= fthr.  perc is an integer
            // in 0..100, so perc >= fthr <=> perc >= ceil(fthr) <=> 100 * freq >= ceil(fthr) * base_num: no division
            if (100ll * (long long)a.freq[c] >= (long long)g_sc.fT * (long long)a.base_num) { if (nF < 3) F[nF] = c; nF++; }
        }
    }
    if (nF >= 3) { st[ST_ERR]++; return d; }
    const uint32_t cN = (uint32_t)(int)g_sc.code_N;
    if (nF == 0) st[ST_DISCARDED]++;
    else if (nF == 1) {
        if (F[0] == cN) st[ST_DISCARDED]++;
        else { d.act = ACT_RULE10; d.sym = F[0]; st[ST_PROCESSED]++; }
    } else if (a.base_num < (uint32_t)g_sc.m) st[ST_DISCARDED]++;
    else if (F[0] == cN) { d.act = ACT_RULE10; d.sym = F[1]; st[ST_PROCESSED]++; }
    else if (F[1] == cN) { d.act = ACT_RULE10; d.sym = F[0]; st[ST_PROCESSED]++; }
    else { d.act = ACT_RULE11; d.f0 = F[0]; d.f1 = F[1]; }
    return d;
}

// rule 10 on one row (bfq_ext.cpp:546-573).  Returns the stored byte (bit 7 = modified).
__device__ __forceinline__ uint8_t rule10_row(uint32_t c, uint8_t q, uint32_t sym, uint32_t hiq,
                                              uint8_t newqs, uint32_t *st)
{
    if (c == 0) return q;
    if (c != sym && !((hiq >> c) & 1u)) { st[ST_MODIFIED]++; return (uint8_t)(q | 0x80); }
    if (c == sym) { st[ST_SMOOTHED]++; return newqs; }
    if (newqs < q) { st[ST_SMOOTHED]++; return newqs; }
    return q;
}

__device__ __forceinline__ void acc_row(const SmoothConst &g_sc, ClAcc &a, uint32_t c, uint8_t q)
{
    if (c == 0) return;
    a.freq[c]++;
    a.base_num++;
    if ((int)q >= g_sc.qthr + 33) a.hiq |= 1u << c;
#if BFQ_M == 0
    a.mx = max(a.mx, (uint32_t)q);
#elif BFQ_M == 1
    a.dsum = __dadd_rn(a.dsum, g_errtab[q]);
#elif BFQ_M == 3
    a.isum += (int32_t)q;
#endif
}

// rule 5-6 flag of row i (bfq_ext.cpp:375-392, 880): lcp[i] >= K and row i is not a local minimum
__device__ __forceinline__ bool f_flag(const SmoothConst &g_sc, int lm1, int l0, int lp1, bool is_first, bool is_last)
{
    if (is_first) return false;
    bool thr = l0 >= g_sc.K;
    bool mn = !is_last && (lm1 > l0) && (l0 <= lp1);
    return thr && !mn;
}

struct R11Desc { uint32_t start, len, packed /* f0 | f1<<8 | newqs<<16 | hiq<<24 */, pad; };

struct SmoothLists {
    uint32_t *long_starts;   uint32_t *long_count;   uint32_t long_cap;
    R11Desc *r11;            uint32_t *r11_count;    uint32_t r11_cap;
    unsigned long long *stats;                        // ST_COUNT counters
};

// ------------------------------------------------------------------ tile kernel (rules 5-10)
// Row-parallel: one block stages CL_S = CL_TILE + CL_EXT rows, 16 consecutive rows per thread, and keeps them in
// registers from the loads to the stores.
//   1. 16-byte loads of bwt / qs / lcp; codes through the LUT; the rule 5-6 flag of every row from the 3-point LCP
//      stencil (the LCP itself is not kept); the thread's rows as bit planes (code bits, "quality >= qthr").
//   2. A cluster is an unflagged row followed by a maximal run of flagged rows.  Cluster starts are the bit pattern
//      ~f & (f >> 1); their block-wide exclusive count (warp scan + warp totals) numbers the clusters of the window, so
//      every row knows its cluster's ORDINAL without searching.
//   3. Segmented reduction of the rule-8 statistics: a thread cuts its 16 rows into segments at the cluster ends,
//      gets the per-symbol row counts and "has a high quality" counts of a segment from popcounts on the bit planes,
//      and adds them to its cluster's accumulator in shared memory (a cluster spans 2-4 threads at 40x coverage).
//   4. One thread per cluster: length from the flag bitmap, rule 9 decision, statistics, hand-over of clusters that
//      leave the window / exceed long_len (cluster_long_kernel) and of two-base clusters (rule11_kernel).  M=1 sums
//      the error probabilities here, serially in row order (double addition is not associative, bfq_ext.cpp:522-543).
//   5. Rule 10 row by row in registers with the decision of the row's cluster; the stored qualities leave as one
//      16-byte store per thread (= the quality half of one walk sector).
// A tile owns the rows [t0, t0 + CL_TILE) except its leading run of flagged rows (they belong to a cluster that
// started in an earlier tile, whose block writes them from its staged extension) plus the rows of its own last
// cluster that lie past the tile.  More than CL_CAP clusters in a window (tiny -k) are processed in rounds.
constexpr int CL_THREADS = 256;             // cluster_long_kernel
constexpr int CLT_THREADS = 288;            // tile kernel: one 16-row chunk per thread
constexpr int CL_TILE = 4096;
constexpr int CL_EXT = 512;                 // rows staged past the tile: a cluster that starts in the tile and ends inside
constexpr int CL_S = CL_TILE + CL_EXT;      // the staged window is finished here, longer ones go to cluster_long_kernel
constexpr int CL_CHUNKS = CL_S / 16;
constexpr int CL_CAP = 1024;                // cluster accumulators per round
constexpr int CL_ACCW = 5;                  // words per accumulator: 11 counters of 10 bits + the M=0 / M=3 word
constexpr uint32_t CL_LONG_MAX = 1023;      // 10-bit counters: longer clusters always go to cluster_long_kernel
static_assert(CLT_THREADS == CL_CHUNKS, "one chunk per thread");

__device__ __forceinline__ void flush_stats(const uint32_t *st, uint32_t *s_stats, unsigned long long *g_stats)
{
    // caller zeroed s_stats and synchronised before; all threads of the block call this
#pragma unroll
    for (int k = 0; k < ST_COUNT; k++) {
        const uint32_t v = __reduce_add_sync(0xffffffffu, st[k]);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(&s_stats[k], v);
    }
    __syncthreads();
    if (threadIdx.x < ST_COUNT && s_stats[threadIdx.x])
        atomicAdd(&g_stats[threadIdx.x], (unsigned long long)s_stats[threadIdx.x]);
}

// bit 0 of each of the four bytes of x -> bits 0..3
__device__ __forceinline__ uint32_t gather_bit0(uint32_t x) { return ((x & 0x01010101u) * 0x01020408u) >> 24; }
// ---- four rows per instruction: packed-byte helpers (flags live in bit 7 of every byte)
// a >= b per unsigned byte
__device__ __forceinline__ uint32_t ge4(uint32_t a, uint32_t b)
{
    const uint32_t t = (a | 0x80808080u) - (b & 0x7F7F7F7Fu);         // bit 7: low 7 bits of a >= low 7 bits of b (no borrow crosses a byte)
    return ((a & ~b) | (~(a ^ b) & t)) & 0x80808080u;
}
// bit 7 of each of the four bytes -> bits 0..3
__device__ __forceinline__ uint32_t gather_bit7(uint32_t x) { return (((x >> 7) & 0x01010101u) * 0x01020408u) >> 24; }
// bits 0..3 -> 0xFF in the bytes whose bit is set
__device__ __forceinline__ uint32_t expand4(uint32_t m4) { return (((m4 & 15u) * 0x00204081u) & 0x01010101u) * 0xFFu; }
// rows (bits) whose 3-bit code, given as bit planes, equals c
__device__ __forceinline__ uint32_t code_rows(uint32_t p0, uint32_t p1, uint32_t p2, uint32_t c)
{
    return ((c & 1u) ? p0 : ~p0) & ((c & 2u) ? p1 : ~p1) & ((c & 4u) ? p2 : ~p2);
}

template <bool FLAT>
__global__ void __launch_bounds__(CLT_THREADS, 4)
cluster_tile_kernel(SC_ARG, const uint8_t *__restrict__ bwt, const uint8_t *__restrict__ qs,
                    const uint8_t *__restrict__ lcp, uint32_t n, uint8_t *__restrict__ lines,
                    uint8_t *__restrict__ obase, SmoothLists L)
{
    __shared__ uint32_t s_acc[CL_CAP * CL_ACCW];
    __shared__ uint32_t s_dec[CL_CAP];          // bit 0 rule 10, bit 1 "closed here: its rows past the tile are mine", sym << 4, hiq << 8, newqs << 16
    __shared__ uint16_t s_start[CL_S / 2 + 2];  // window row of every cluster start, by ordinal
    __shared__ uint32_t s_fb[CL_CHUNKS + 1];    // flags of chunk k in bits 0..15; [CL_CHUNKS] bit 0 = flag of row CL_S
    __shared__ uint32_t s_wsum[CLT_THREADS / 32];
    __shared__ uint32_t s_stats[ST_COUNT];
    __shared__ uint8_t s_lut[256];
#if BFQ_M == 1
    // M=1 adds the error probabilities of a cluster serially, in row order (double addition is not associative).  Per row
    // that loop reads ONE index byte -- the quality, or 128 for an end-marker row -- and one table entry from shared memory
    // (entry 128 = +0.0, whose addition is exact); the table in constant memory would serialise its divergent lookups.
    __shared__ __align__(16) uint8_t s_qi[CL_S];
    __shared__ double s_err[129];
#endif
    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const int64_t t0 = (int64_t)blockIdx.x * CL_TILE;
    const int64_t nn = (int64_t)n;
    for (int i = tid; i < 256; i += CLT_THREADS) s_lut[i] = g_sc.byte2code[i];
    if (tid < ST_COUNT) s_stats[tid] = 0;
#if BFQ_M == 1
    for (int i = tid; i < 129; i += CLT_THREADS) s_err[i] = i < 128 ? g_errtab[i] : 0.0;
#endif
    __syncthreads();

    // ---- 1. this thread's 16 rows
    const int r0 = 16 * tid;
    const int64_t i0 = t0 + r0;
    uint32_t wl[4] = {0, 0, 0, 0}, cw[4] = {0, 0, 0, 0}, qw[4] = {0, 0, 0, 0};     // 16 packed bytes each
    const bool vec = ((reinterpret_cast<uintptr_t>(bwt) | reinterpret_cast<uintptr_t>(qs) | reinterpret_cast<uintptr_t>(lcp)) & 15) == 0;
    const bool full = i0 + 16 <= nn;
    if (vec && full) {
        const uint4 vl = *reinterpret_cast<const uint4 *>(lcp + i0);
        const uint4 vb = *reinterpret_cast<const uint4 *>(bwt + i0);
        const uint4 vq = *reinterpret_cast<const uint4 *>(qs + i0);
        wl[0] = vl.x; wl[1] = vl.y; wl[2] = vl.z; wl[3] = vl.w;
        qw[0] = vq.x; qw[1] = vq.y; qw[2] = vq.z; qw[3] = vq.w;
        const uint32_t wb[4] = {vb.x, vb.y, vb.z, vb.w};
#pragma unroll
        for (int j = 0; j < 16; j++) cw[j >> 2] |= (uint32_t)s_lut[(wb[j >> 2] >> (8 * (j & 3))) & 255u] << (8 * (j & 3));
    } else {
#pragma unroll
        for (int j = 0; j < 16; j++) {
            const int64_t i = i0 + j;
            if (i < nn) {
                wl[j >> 2] |= (uint32_t)lcp[i] << (8 * (j & 3));
                cw[j >> 2] |= (uint32_t)s_lut[bwt[i]] << (8 * (j & 3));
                qw[j >> 2] |= (uint32_t)qs[i] << (8 * (j & 3));
            }
        }
    }
    const uint32_t valid = full ? 0xFFFFu : (i0 < nn ? (1u << (nn - i0)) - 1u : 0u);          // rows of the chunk inside the BWT
    uint32_t f16 = 0;
    if (vec && i0 > 0 && i0 + 17 <= nn && g_sc.K >= 1 && g_sc.K <= 255) {
        // interior chunk (no row is the first or the last of the BWT): the 3-point stencil on four rows at a time.
        // f = (l0 >= K) && !(lm > l0 && l0 <= lp)  =  (l0 >= K) && (l0 >= lm || !(lp >= l0))
        const uint32_t K4 = (uint32_t)g_sc.K * 0x01010101u;
        const uint32_t prevw = (uint32_t)lcp[i0 - 1] << 24, nextw = lcp[i0 + 16];
#pragma unroll
        for (int w = 0; w < 4; w++) {
            const uint32_t l0 = wl[w];
            const uint32_t lm = __funnelshift_l(w ? wl[w - 1] : prevw, l0, 8);          // the LCP of the row before, per byte
            const uint32_t lp = __funnelshift_r(l0, w < 3 ? wl[w + 1] : nextw, 8);       // ... of the row after
            const uint32_t f = ge4(l0, K4) & (ge4(l0, lm) | ~ge4(lp, l0));
            f16 |= gather_bit7(f) << (4 * w);
        }
    } else {
        const uint32_t lprev = (i0 > 0 && i0 - 1 < nn) ? lcp[i0 - 1] : 0u, lnext = (i0 + 16 < nn) ? lcp[i0 + 16] : 0u;
#pragma unroll
        for (int j = 0; j < 16; j++) {
            const int64_t i = i0 + j;
            const uint32_t lm = j == 0 ? lprev : (wl[(j - 1) >> 2] >> (8 * ((j - 1) & 3))) & 255u;
            const uint32_t l0 = (wl[j >> 2] >> (8 * (j & 3))) & 255u;
            const uint32_t lp = j == 15 ? lnext : (wl[(j + 1) >> 2] >> (8 * ((j + 1) & 3))) & 255u;
            const bool f = (i > 0 && i < nn) ? f_flag(g_sc, (int)lm, (int)l0, (int)lp, false, i == nn - 1) : false;
            f16 |= (uint32_t)f << j;
        }
    }
    s_fb[tid] = f16;
    if (tid == CLT_THREADS - 1) {               // flag of the row just past the window (is the last cluster still open there?)
        const int64_t i = t0 + CL_S;
        bool f = false;
        if (i > 0 && i < nn) f = f_flag(g_sc, (int)lcp[i - 1], (int)lcp[i], i + 1 < nn ? (int)lcp[i + 1] : 0, false, i == nn - 1);
        s_fb[CL_CHUNKS] = f ? 1u : 0u;
    }
#if BFQ_M == 1
    {
        uint32_t qi[4];
#pragma unroll
        for (int w = 0; w < 4; w++) {
            const uint32_t nz = (((cw[w] + 0x7F7F7F7Fu) & 0x80808080u) >> 7) * 0xFFu;       // 0xFF where the code is not the end-marker (codes are <= 5)
            qi[w] = (qw[w] & 0x7F7F7F7Fu & nz) | (0x80808080u & ~nz);
        }
        *reinterpret_cast<uint4 *>(s_qi + r0) = make_uint4(qi[0], qi[1], qi[2], qi[3]);
    }
#endif
    // bit planes of the 16 codes and the "quality >= qthr + 33" bits (qualities are < 128: smooth_dev checks)
    uint32_t p0 = 0, p1 = 0, p2 = 0, hq = 0;
    {
        const uint32_t thr4 = (uint32_t)min(max(g_sc.qthr + 33, 0), 128) * 0x01010101u;
#pragma unroll
        for (int w = 0; w < 4; w++) {
            p0 |= gather_bit0(cw[w]) << (4 * w);
            p1 |= gather_bit0(cw[w] >> 1) << (4 * w);
            p2 |= gather_bit0(cw[w] >> 2) << (4 * w);
            hq |= gather_bit0((((qw[w] | 0x80808080u) - thr4) & 0x80808080u) >> 7) << (4 * w);
        }
    }
    __syncthreads();

    // ---- 2. cluster starts, ordinals
    const uint32_t f17 = f16 | ((s_fb[tid + 1] & 1u) << 16);
    const uint32_t startm = ~f16 & (f17 >> 1) & 0xFFFFu;      // an unflagged row followed by a flagged one opens a cluster
    const uint32_t inm = f16 | startm;                         // rows that belong to some cluster
    const uint32_t endm = inm & ~(f17 >> 1);                   // last row of its cluster
    uint32_t ord0, total;
    {
        const uint32_t cnt = __popc(startm);
        uint32_t x = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) s_wsum[wrp] = x;
        __syncthreads();
        uint32_t add = 0, tot = 0;
#pragma unroll
        for (int i = 0; i < CLT_THREADS / 32; i++) { const uint32_t v = s_wsum[i]; if (i < wrp) add += v; tot += v; }
        ord0 = add + x - cnt;                                  // clusters that start before this chunk
        total = tot;
    }
    {
        uint32_t m = startm, k = ord0;
        while (m) { s_start[k++] = (uint16_t)(r0 + __ffs(m) - 1); m &= m - 1; }
    }
    // rows of this chunk that continue a cluster of an earlier tile (ordinal -1): not mine
    uint32_t leadm = 0;
    if (ord0 == 0 && (inm & 1u) && !(startm & 1u)) {
        const uint32_t sm = startm ? ((1u << (__ffs(startm) - 1)) - 1u) : 0xFFFFu;      // rows before the first start of the window
        const uint32_t em = endm & sm;
        leadm = em ? ((2u << (__ffs(em) - 1)) - 1u) : (inm & sm);                          // up to the run's end, or all of them
        leadm &= inm;
    }

    uint32_t st[ST_COUNT];
#pragma unroll
    for (int k = 0; k < ST_COUNT; k++) st[k] = 0;
    uint32_t extm = 0;                                          // rows past the tile that this block writes
    const uint32_t long_len = min(g_sc.long_len, CL_LONG_MAX);

    for (uint32_t base = 0; base < total; base += CL_CAP) {    // (block-uniform; one round unless -k is tiny)
        const uint32_t cnt = min((uint32_t)CL_CAP, total - base);
        for (uint32_t i = tid; i < cnt * CL_ACCW; i += CLT_THREADS) s_acc[i] = 0;
        __syncthreads();
        // ---- 3. segmented reduction into the cluster accumulators
        {
            uint32_t m = inm & ~leadm;
            while (m) {
                const int a = __ffs(m) - 1;
                const uint32_t em = endm >> a;
                const int slen = em ? __ffs(em) : 16 - a;
                const uint32_t seg = ((1u << slen) - 1u) << a;
                m &= ~seg;
                const uint32_t o = ord0 + __popc(startm & ((2u << a) - 1u)) - 1u - base;
                if (o < cnt) {
                    const uint32_t n0 = ~p0 & seg, n1 = ~p1 & seg, n2 = ~p2 & seg;
                    const uint32_t m0 = n0 & n1 & n2, m1 = p0 & n1 & n2, m2 = n0 & p1 & n2, m3 = p0 & p1 & n2, m4 = n0 & n1 & p2, m5 = p0 & n1 & p2;
                    uint32_t *acc = s_acc + o * CL_ACCW;
                    atomicAdd(acc + 0, __popc(m0) | (__popc(m1) << 10) | (__popc(m2) << 20));
                    atomicAdd(acc + 1, __popc(m3) | (__popc(m4) << 10) | (__popc(m5) << 20));
                    atomicAdd(acc + 2, __popc(m1 & hq) | (__popc(m2 & hq) << 10) | (__popc(m3 & hq) << 20));
                    atomicAdd(acc + 3, __popc(m4 & hq) | (__popc(m5 & hq) << 10));
#if BFQ_M == 0 || BFQ_M == 3
                    // largest quality (M=0) / sum of the qualities (M=3) of the segment's non-EOS rows, four rows at a time
                    const uint32_t rows = seg & ~m0;
                    uint32_t X = 0;
#pragma unroll
                    for (int w = 0; w < 4; w++) {
                        const uint32_t qm = qw[w] & expand4(rows >> (4 * w));
#if BFQ_M == 0
                        X = __vmaxu4(X, qm);
#else
                        X = __dp4a(qm, 0x01010101u, X);
#endif
                    }
#if BFQ_M == 0
                    X = __vmaxu4(X, X >> 16);
                    X = max(X & 255u, (X >> 8) & 255u);
                    if (X) atomicMax(acc + 4, X);
#else
                    if (X) atomicAdd(acc + 4, X);
#endif
#endif
                }
            }
        }
        __syncthreads();
        // ---- 4. one thread per cluster: rules 7-9
        for (uint32_t k = tid; k < cnt; k += CLT_THREADS) {
            const int a = s_start[base + k];
            uint32_t dec = 0;
            if (a < CL_TILE) {                                  // (clusters that start in the extension belong to the next tile)
                int ch = (a + 1) >> 4;
                uint32_t bits = ~s_fb[ch] & 0xFFFFu & (0xFFFFu << ((a + 1) & 15));
                while (!bits) bits = ~s_fb[++ch] & 0xFFFFu;    // s_fb[CL_CHUNKS] has zero bits: terminates
                const int e = 16 * ch + __ffs(bits) - 1;       // first unflagged row after the start
                const uint32_t len = (uint32_t)(e - a);
                if (e > CL_S || len > long_len) {               // open at the end of the window, or too long: hand over
                    const uint32_t kk = atomicAdd(L.long_count, 1u);
                    if (kk < L.long_cap) L.long_starts[kk] = (uint32_t)(t0 + a);
                } else {
                    dec = 2u;
                    const uint32_t *acc = s_acc + k * CL_ACCW;
                    const uint32_t w0 = acc[0], w1 = acc[1], w2 = acc[2], w3 = acc[3];
                    ClAcc A;
                    A.freq[0] = 0;
                    A.freq[1] = (w0 >> 10) & 1023u; A.freq[2] = w0 >> 20;
                    A.freq[3] = w1 & 1023u; A.freq[4] = (w1 >> 10) & 1023u; A.freq[5] = w1 >> 20;
                    A.hiq = ((w2 & 1023u) ? 2u : 0u) | (((w2 >> 10) & 1023u) ? 4u : 0u) | ((w2 >> 20) ? 8u : 0u) |
                            ((w3 & 1023u) ? 16u : 0u) | ((w3 >> 10) ? 32u : 0u);
                    A.base_num = A.freq[1] + A.freq[2] + A.freq[3] + A.freq[4] + A.freq[5];
                    A.mx = acc[4]; A.isum = (int32_t)acc[4]; A.dsum = 0.0;
#if BFQ_M == 1
                    for (int r = a; r < e; r++) A.dsum = __dadd_rn(A.dsum, s_err[s_qi[r]]);
#endif
                    const Decision d = decide(g_sc, len, A, st);
                    if (d.act == ACT_RULE11) {
                        const uint32_t kk = atomicAdd(L.r11_count, 1u);
                        if (kk < L.r11_cap)
                            L.r11[kk] = R11Desc{(uint32_t)(t0 + a), len, d.f0 | (d.f1 << 8) | ((uint32_t)d.newqs << 16) | (A.hiq << 24), 0};
                    } else if (d.act == ACT_RULE10)
                        dec = 3u | (d.sym << 4) | (A.hiq << 8) | ((uint32_t)d.newqs << 16);
                }
            }
            s_dec[k] = dec;
        }
        __syncthreads();
        // ---- 5. rule 10 (bfq_ext.cpp:546-573) on the thread's 16 rows in registers.  Per segment (= the rows of one cluster
        // inside this chunk) the decision gives three row sets as bit masks: rows whose base is replaced (bit 7 of the stored
        // quality), rows that take the new quality, rows that take min(quality, new quality).  They are applied to the packed
        // qualities four rows at a time; only the (rare) replaced bases are written row by row.
        if (inm & ~leadm) {
            const uint32_t nz = p0 | p1 | p2;
            uint32_t MOD = 0, EQ = 0, MINC = 0;                  // rows of the chunk, over all its segments
            uint32_t nq[4] = {0, 0, 0, 0};                       // the new quality of the row's cluster, per byte
            uint32_t m = inm & ~leadm;
            while (m) {
                const int a = __ffs(m) - 1;
                const uint32_t em = endm >> a;
                const int slen = em ? __ffs(em) : 16 - a;
                const uint32_t seg = ((1u << slen) - 1u) << a;
                m &= ~seg;
                const uint32_t o = ord0 + __popc(startm & ((2u << a) - 1u)) - 1u - base;
                const uint32_t dec = o < cnt ? s_dec[o] : 0u;
                if (r0 >= CL_TILE && (dec & 2u)) extm |= seg;   // rows past the tile of a cluster this block closed
                if (dec & 1u) {
                    const uint32_t sym = (dec >> 4) & 15u, hiq = (dec >> 8) & 255u;
                    const uint32_t eq = code_rows(p0, p1, p2, sym) & seg;
                    uint32_t hi = 0;                             // rows whose symbol has a high-quality occurrence in the cluster
#pragma unroll
                    for (uint32_t c = 1; c <= 5; c++) if ((hiq >> c) & 1u) hi |= code_rows(p0, p1, p2, c);
                    const uint32_t other = nz & seg & ~eq;
                    const uint32_t mod = other & ~hi;
                    MOD |= mod; EQ |= eq; MINC |= other & hi;
                    const uint32_t nq4 = ((dec >> 16) & 255u) * 0x01010101u;
#pragma unroll
                    for (int w = 0; w < 4; w++) nq[w] |= nq4 & expand4(seg >> (4 * w));
                    if (mod) {
                        const uint8_t symb = g_sc.code2byte[sym];
                        uint32_t mm = mod;
                        while (mm) { obase[i0 + __ffs(mm) - 1] = symb; mm &= mm - 1; }
                    }
                }
            }
            if (MOD | EQ | MINC) {
                st[ST_MODIFIED] += __popc(MOD);
                st[ST_SMOOTHED] += __popc(EQ);
#pragma unroll
                for (int w = 0; w < 4; w++) {
                    const uint32_t q = qw[w];
                    // rows of the min set whose quality is above the new one (both are below 128: bit 7 of the difference).
                    // Bit 7 of q is masked out: a row that an EARLIER round of this loop marked as replaced carries it, and
                    // its borrow would flip the compare of the next byte when that row's quality equals the new one.
                    const uint32_t lt = ~((nq[w] | 0x80808080u) - (q & 0x7F7F7F7Fu)) & 0x80808080u & expand4(MINC >> (4 * w));
                    st[ST_SMOOTHED] += __popc(lt);
                    const uint32_t take = expand4(EQ >> (4 * w)) | ((lt >> 7) * 0xFFu);
                    qw[w] = (q & ~take) | (nq[w] & take) | (expand4(MOD >> (4 * w)) & 0x80808080u);
                }
            }
        }
        __syncthreads();
    }

    // ---- stored qualities of the rows this block owns
    const uint32_t own = (r0 < CL_TILE ? (valid & ~leadm) : extm) & valid;
    if (own) {
        uint8_t *dst = q_slot<FLAT>(lines, (uint64_t)i0);         // 16 rows = the quality half of one walk sector
#if BFQ_B
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t x = qw[j];
            qw[j] = (uint32_t)store_q((uint8_t)x) | ((uint32_t)store_q((uint8_t)(x >> 8)) << 8) |
                    ((uint32_t)store_q((uint8_t)(x >> 16)) << 16) | ((uint32_t)store_q((uint8_t)(x >> 24)) << 24);
        }
#endif
        if (own == 0xFFFFu && (reinterpret_cast<uintptr_t>(dst) & 15) == 0)
            *reinterpret_cast<uint4 *>(dst) = make_uint4(qw[0], qw[1], qw[2], qw[3]);
        else {
#pragma unroll
            for (int j = 0; j < 16; j++)
                if ((own >> j) & 1u) *q_slot<FLAT>(lines, (uint64_t)(i0 + j)) = (uint8_t)(qw[j >> 2] >> (8 * (j & 3)));
        }
    }
    flush_stats(st, s_stats, L.stats);
}

// ------------------------------------------------------------------ long clusters: one block each
template <bool FLAT>
__global__ void __launch_bounds__(CL_THREADS)
cluster_long_kernel(SC_ARG, const uint8_t *__restrict__ bwt, const uint8_t *__restrict__ qs,
                    const uint8_t *__restrict__ lcp, uint32_t n, uint8_t *__restrict__ lines,
                    uint8_t *__restrict__ obase, SmoothLists L)
{
    __shared__ uint8_t s_lut[256];
    __shared__ uint32_t s_stats[ST_COUNT];
    __shared__ uint32_t s_end;
    __shared__ uint32_t s_freq[6], s_hiq, s_base, s_mx, s_isum;
    __shared__ double s_dsum;
    __shared__ Decision s_dec;
    __shared__ uint8_t s_c[1024], s_qq[1024];
    const int tid = threadIdx.x;
    const int64_t nn = (int64_t)n;
    s_lut[tid] = g_sc.byte2code[tid];
    if (tid < ST_COUNT) s_stats[tid] = 0;
    __syncthreads();
    uint32_t st[ST_COUNT];
#pragma unroll
    for (int k = 0; k < ST_COUNT; k++) st[k] = 0;

    auto lcp_at = [&](int64_t i) -> int { return (i < 0 || i >= nn) ? 0 : (int)lcp[i]; };
    auto F = [&](int64_t i) -> bool {
        if (i <= 0 || i >= nn) return false;
        return f_flag(g_sc, lcp_at(i - 1), lcp_at(i), lcp_at(i + 1), false, i == nn - 1);
    };
    const uint32_t count = min(*L.long_count, L.long_cap);
    for (uint32_t ci = blockIdx.x; ci < count; ci += gridDim.x) {
        const int64_t a = L.long_starts[ci];
        // ---- end of the run: first e > a with !F(e)
        if (tid == 0) s_end = 0xFFFFFFFFu;
        __syncthreads();
        for (int64_t base = a + 1;; base += CL_THREADS) {
            int64_t i = base + tid;
            if (i <= nn && !F(i)) atomicMin(&s_end, (uint32_t)i);     // F(nn) is false
            __syncthreads();
            if (s_end != 0xFFFFFFFFu) break;
            __syncthreads();
        }
        const int64_t e = s_end;
        const uint32_t len = (uint32_t)(e - a);
        // ---- statistics
        if (tid < 6) s_freq[tid] = 0;
        if (tid == 0) { s_hiq = 0; s_base = 0; s_mx = 0; s_isum = 0; s_dsum = 0.0; }
        __syncthreads();
        {
            uint32_t f[6] = {0, 0, 0, 0, 0, 0}, hiq = 0, mx = 0, isum = 0;
            for (int64_t i = a + tid; i < e; i += CL_THREADS) {
                uint32_t c = s_lut[bwt[i]]; uint8_t q = qs[i];
                if (!c) continue;
#pragma unroll
                for (uint32_t k = 1; k < 6; k++) f[k] += (c == k);
                if ((int)q >= g_sc.qthr + 33) hiq |= 1u << c;
                mx = max(mx, (uint32_t)q);
                isum += q;
            }
#pragma unroll
            for (int k = 1; k < 6; k++) if (f[k]) { atomicAdd(&s_freq[k], f[k]); atomicAdd(&s_base, f[k]); }
            if (hiq) atomicOr(&s_hiq, hiq);
            atomicMax(&s_mx, mx);
            atomicAdd(&s_isum, isum);          // wraps mod 2^32 exactly like the reference's int
        }
#if BFQ_M == 1
        // the reference adds the error probabilities in row order, in double: keep that order
        for (int64_t base = a; base < e; base += 1024) {
            __syncthreads();
            for (int r = tid; r < 1024; r += CL_THREADS) {
                int64_t i = base + r;
                if (i < e) { s_c[r] = s_lut[bwt[i]]; s_qq[r] = qs[i]; }
            }
            __syncthreads();
            if (tid == 0) {
                double s = s_dsum;
                int lim = (int)min((int64_t)1024, e - base);
                for (int r = 0; r < lim; r++) if (s_c[r]) s = __dadd_rn(s, g_errtab[s_qq[r]]);
                s_dsum = s;
            }
        }
#endif
        __syncthreads();
        if (tid == 0) {
            ClAcc acc;
#pragma unroll
            for (int k = 0; k < 6; k++) acc.freq[k] = s_freq[k];
            acc.hiq = s_hiq; acc.base_num = s_base; acc.mx = s_mx; acc.isum = (int32_t)s_isum; acc.dsum = s_dsum;
            Decision d = decide(g_sc, len, acc, st);
            if (d.act == ACT_RULE11) {
                uint32_t k = atomicAdd(L.r11_count, 1u);
                if (k < L.r11_cap)
                    L.r11[k] = R11Desc{(uint32_t)a, len, d.f0 | (d.f1 << 8) | ((uint32_t)d.newqs << 16) | (acc.hiq << 24), 0};
            }
            s_dec = d;
        }
        __syncthreads();
        const Decision d = s_dec;
        const uint32_t hiq = s_hiq;
        const uint8_t symb = g_sc.code2byte[d.sym];
        // ---- every row of a long cluster is written here (later tiles skipped them)
        for (int64_t i = a + tid; i < e; i += CL_THREADS) {
            uint32_t c = s_lut[bwt[i]]; uint8_t q = qs[i];
            uint8_t o = q;
            if (d.act == ACT_RULE10) o = rule10_row(c, q, d.sym, hiq, d.newqs, st);
            if (o & 0x80) obase[i] = symb;
            *q_slot<FLAT>(lines, (uint64_t)i) = store_q(o);
        }
        __syncthreads();
    }
    flush_stats(st, s_stats, L.stats);
}

// ------------------------------------------------------------------ rule 11: two frequent bases
// bfq_ext.cpp:713-802 with buildLF_Cluster :415-476.  One warp per cluster; the symbol that
// precedes row j in its read is bwt[LF(j)], two walk-line fetches.
// FLAT: the row's symbol comes from the BWT bytes and the symbol before it in its read straight from the
// packed text at SA[row] - 2 (code 0 there = the previous read's terminator = "no preceding base").
template <bool FLAT>
__global__ void __launch_bounds__(256)
rule11_kernel(SC_ARG, const uint8_t *__restrict__ qs, uint8_t *__restrict__ lines, uint8_t *__restrict__ obase,
              SmoothLists L, const uint8_t *__restrict__ bwt, const void *__restrict__ SA, int sa_wide,
              const uint64_t *__restrict__ T3, const uint32_t *__restrict__ super)
{
    __shared__ uint32_t s_stats[ST_COUNT];
    __shared__ uint8_t s_lut[256];
    const int tid = threadIdx.x, lane = tid & 31;
    s_lut[tid] = g_sc.byte2code[tid];
    if (tid < ST_COUNT) s_stats[tid] = 0;
    __syncthreads();
    auto row_code = [&](uint64_t i) -> uint32_t {
        if (FLAT) return s_lut[bwt[i]];
        return ws_code(ws_load(lines, i), (uint32_t)(i & 15));
    };
    auto prec_code = [&](uint64_t i, uint32_t c) -> uint32_t {      // symbol before row i's symbol c in its read
        if (FLAT) {
            const uint64_t p = sa_wide ? ((const uint64_t *)SA)[i] : (uint64_t)((const uint32_t *)SA)[i];
            return p >= 2 ? sym_code(T3, p - 2) : 0u;
        }
        const uint64_t r = ws_lf(ws_load(lines, i), super, i, c);
        return ws_code(ws_load(lines, r), (uint32_t)(r & 15));
    };
    uint32_t st[ST_COUNT];
#pragma unroll
    for (int k = 0; k < ST_COUNT; k++) st[k] = 0;
    const uint32_t count = min(*L.r11_count, L.r11_cap);
    const uint32_t warp0 = (blockIdx.x * 256u + tid) >> 5, nwarps = (gridDim.x * 256u) >> 5;
    const uint32_t acgt = g_sc.acgt_mask;
    for (uint32_t di = warp0; di < count; di += nwarps) {
        const R11Desc D = L.r11[di];
        const uint32_t f0 = D.packed & 255u, f1 = (D.packed >> 8) & 255u, hiq = D.packed >> 24;
        const uint8_t newqs = (uint8_t)(D.packed >> 16);
        uint32_t P0 = 0, P1 = 0;
        for (uint32_t j = lane; j < D.len; j += 32) {
            const uint64_t i = (uint64_t)D.start + j;
            const uint32_t c = row_code(i);
            if (c == f0 || c == f1) {
                const uint32_t pc = prec_code(i, c);
                if ((acgt >> pc) & 1u) { if (c == f0) P0 |= 1u << pc; else P1 |= 1u << pc; }
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) { P0 |= __shfl_xor_sync(0xffffffffu, P0, o); P1 |= __shfl_xor_sync(0xffffffffu, P1, o); }
        const bool ok = __popc(P0) == 1 && __popc(P1) == 1 && P0 != P1;
        if (!ok) { if (lane == 0) st[ST_AMBIGUOUS]++; continue; }
        if (lane == 0) st[ST_PROCESSED]++;
        const uint32_t prec0 = __ffs(P0) - 1, prec1 = __ffs(P1) - 1;
        for (uint32_t j = lane; j < D.len; j += 32) {
            const uint64_t i = (uint64_t)D.start + j;
            const uint32_t c = row_code(i);
            if (c == 0) continue;
            const uint8_t q = qs[i];
            if (c != f0 && c != f1 && !((hiq >> c) & 1u)) {
                const uint32_t pc = prec_code(i, c);
                uint32_t to = pc == prec0 ? f0 : (pc == prec1 ? f1 : 0u);
                if (to) {
                    obase[i] = g_sc.code2byte[to];
                    *q_slot<FLAT>(lines, i) = store_q((uint8_t)(q | 0x80));
                    st[ST_MODIFIED]++;
                }
            } else if (c == f0 || c == f1) {
                *q_slot<FLAT>(lines, i) = store_q(newqs);
                st[ST_SMOOTHED]++;
            } else if (newqs < q) {
                *q_slot<FLAT>(lines, i) = store_q(newqs);
                st[ST_SMOOTHED]++;
            }
        }
    }
    flush_stats(st, s_stats, L.stats);
}

// ------------------------------------------------------------------ host side
// M=1 tables, computed with the host libm the reference itself links (bfq_ext.cpp:532, 539).
static void build_m1_tables(double errtab[256], double thr[128])
{
    for (int q = 0; q < 256; q++) errtab[q] = pow(10, -((double)q - 33.0) / 10.0);
    auto pred = [](double x, int v) { return round(-10 * log10(x)) >= (double)v; };
    thr[0] = INFINITY;
    for (int v = 1; v < 128; v++) {
        // largest double x with round(-10*log10(x)) >= v; the predicate is monotone in x
        double lo = 1e-300, hi = 1e10;                 // pred(lo) true, pred(hi) false
        uint64_t a, b;
        memcpy(&a, &lo, 8); memcpy(&b, &hi, 8);
        while (b - a > 1) {
            uint64_t mid = a + (b - a) / 2; double x;
            memcpy(&x, &mid, 8);
            if (pred(x, v)) a = mid; else b = mid;
        }
        memcpy(&thr[v], &a, 8);
    }
}

static void launch_check(bfq_ctx *c) { BFQ_CUDA(cudaGetLastError()); (void)c; }

void smooth_dev(bfq_ctx *c, const uint8_t *d_bwt, const uint8_t *d_qs, const uint8_t *d_lcp,
                uint64_t n, const bfq_params *P, SmoothResult *out, bfq_stats *stats,
                const void *d_sa, const uint64_t *d_t3, const uint64_t *alpha_hist, bool sa_wide,
                const uint64_t *known_bwt_hist, const uint64_t *known_qs_hist)
{
    const bool flat = d_sa != nullptr && d_t3 != nullptr;
    out->flat = flat;
    cudaStream_t st = c->stream;
    if (n >= ((uint64_t)1 << 32) - 64) BFQ_FAIL(BFQ_ETOOLARGE, "BWT has %llu rows; this build indexes rows with 32 bits", (unsigned long long)n);
    if (P->max_len <= 0 || P->max_len > 255) BFQ_FAIL(BFQ_EINVAL, "max_len must be 1..255 (dataTypelenSeq is uchar, parameters.h:68)");
    if (BFQ_M == 2 && P->default_q >= 128) BFQ_FAIL(BFQ_EINVAL, "default_q (-v) must be an ASCII quality below 128 (bit 7 of a stored quality marks a replaced base)");

    // ---- buildFreq (bfq_ext.cpp:132-187): symbol counts, alphabet in byte order, pile starts
    uint64_t hist[256], qhist[256];
    if (known_bwt_hist && known_qs_hist) { memcpy(hist, known_bwt_hist, sizeof hist); memcpy(qhist, known_qs_hist, sizeof qhist); }
    else { byte_histogram(c, d_bwt, n, hist); byte_histogram(c, d_qs, n, qhist); }
    for (int b = 128; b < 256; b++)
        if (qhist[b]) BFQ_FAIL(BFQ_EINVAL, "quality byte %d >= 128 is not a FASTQ quality", b);
    Alphabet A = alphabet_from_hist(alpha_hist ? alpha_hist : hist, P->eos_byte, /*eos_may_appear=*/true);
    out->alpha = A;
    out->n = n;
    out->nreads = hist[P->eos_byte];

    SmoothConst sc;
    memset(&sc, 0, sizeof sc);
    for (int b = 0; b < 256; b++) sc.byte2code[b] = A.byte2code[b] == 0xFF ? 0 : A.byte2code[b];
    memcpy(sc.code2byte, A.code2byte, 8);
    sc.K = P->K; sc.m = P->m; sc.qthr = P->qthr; sc.fthr = P->fthr;
    sc.fT = (int32_t)std::min(1000.0f, std::max(-1000.0f, ceilf(P->fthr)));
    sc.default_q = P->default_q; sc.code_N = (int8_t)A.code_N; sc.acgt_mask = A.acgt_mask;
    sc.nsym = (uint8_t)A.nsym;
    sc.long_len = (uint32_t)std::min<int64_t>(c->opt_long_cluster, 4000);
#if BFQ_M == 1
    {   // the M=1 tables are the same for every call: written once per device (several contexts may share one)
        static std::once_flag once[64];
        std::call_once(once[c->device & 63], [&] {
            static double errtab[256], thr[128];
            build_m1_tables(errtab, thr);
            BFQ_CUDA(cudaMemcpyToSymbol(g_errtab, errtab, sizeof errtab));
            BFQ_CUDA(cudaMemcpyToSymbol(g_m1thr, thr, sizeof thr));
        });
    }
#endif
    uint32_t *d_pile = (uint32_t *)((char *)c->d_small + 2048);
    if (!flat) {                          // (flat mode builds no walk lines: the pile starts are not needed)
        uint32_t pile[8] = {0};
        uint64_t acc = hist[P->eos_byte];
        for (int k = 1; k <= 5; k++) { pile[k - 1] = (uint32_t)acc; if (k <= A.nsym) acc += hist[A.code2byte[k]]; }
        uint32_t *h = (uint32_t *)((char *)c->h_small + 2048);
        memcpy(h, pile, sizeof pile);
        BFQ_CUDA(cudaMemcpyAsync(d_pile, h, sizeof pile, cudaMemcpyHostToDevice, st));
        BFQ_CUDA(cudaStreamSynchronize(st));     // the pinned staging above is reused by later calls
    }

    // ---- walk lines (computeVectorUnbuildBCR, decode.cpp:87-235, + tableOcc)
    static_assert(RK_ROWS == (1 << WS_SUPER_SHIFT), "a rank block is one super-block of the walk sectors");
    const uint64_t nlines = (n + 31) / 32;                 // two 32-byte sectors each
    out->lines = flat ? arena_alloc<uint8_t>(c, n + 64) : arena_alloc<uint8_t>(c, nlines * 64 + 64);
    out->obase = arena_alloc<uint8_t>(c, n + 64);
    const uint32_t nblk = div_up(n, RK_ROWS);
    uint32_t *blk = flat ? nullptr : arena_alloc<uint32_t>(c, (size_t)nblk * 8);       // stays: the super-block table
    out->super = blk;
    Arena::Mark mk = c->arena.mark();
    if (!flat) {
        ProfScope ps(c, PC_RANK, 2 * n + nlines * 32, 3);
        rank_count_kernel<<<nblk, RK_THREADS, 0, st>>>(sc, d_bwt, n, blk);
        rank_scan_kernel<<<1, 1024, 0, st>>>(blk, nblk, d_pile);
        rank_build_kernel<<<nblk, RK_THREADS, 0, st>>>(sc, d_bwt, n, out->lines);
        launch_check(c);
    }

    // ---- clusters
    SmoothLists L;
    L.long_cap = (uint32_t)(n / std::max<uint32_t>(std::min<uint32_t>(sc.long_len, CL_LONG_MAX), 1) + 1024);
    L.r11_cap = c->opt_r11_cap > 0 ? (uint32_t)c->opt_r11_cap : (uint32_t)(n / 8 + 1024);
    const uint64_t *h = c->h_small;
    const uint32_t *cnts = (const uint32_t *)(h + ST_COUNT);
    for (int attempt = 0;; attempt++) {
        Arena::Mark mk2 = c->arena.mark();
        L.long_starts = arena_alloc<uint32_t>(c, L.long_cap);
        L.r11 = arena_alloc<R11Desc>(c, L.r11_cap);
        unsigned long long *d_stats = arena_alloc<unsigned long long>(c, ST_COUNT + 2);
        L.stats = d_stats;
        L.long_count = (uint32_t *)(d_stats + ST_COUNT);
        L.r11_count = L.long_count + 1;
        BFQ_CUDA(cudaMemsetAsync(d_stats, 0, (ST_COUNT + 2) * 8, st));
        {
            ProfScope ps(c, PC_CLUSTER, 4 * n);
            if (flat) cluster_tile_kernel<true><<<div_up(n, CL_TILE), CLT_THREADS, 0, st>>>(sc, d_bwt, d_qs, d_lcp, (uint32_t)n, out->lines, out->obase, L);
            else cluster_tile_kernel<false><<<div_up(n, CL_TILE), CLT_THREADS, 0, st>>>(sc, d_bwt, d_qs, d_lcp, (uint32_t)n, out->lines, out->obase, L);
            launch_check(c);
        }
        {
            ProfScope ps(c, PC_CLUSTER_LONG, 0);
            if (flat) cluster_long_kernel<true><<<c->num_sms * 2, CL_THREADS, 0, st>>>(sc, d_bwt, d_qs, d_lcp, (uint32_t)n, out->lines, out->obase, L);
            else cluster_long_kernel<false><<<c->num_sms * 2, CL_THREADS, 0, st>>>(sc, d_bwt, d_qs, d_lcp, (uint32_t)n, out->lines, out->obase, L);
            launch_check(c);
        }
        {
            ProfScope ps(c, PC_RULE11, 0);
            if (flat) rule11_kernel<true><<<c->num_sms * 4, 256, 0, st>>>(sc, d_qs, out->lines, out->obase, L, d_bwt, d_sa, sa_wide ? 1 : 0, d_t3, nullptr);
            else rule11_kernel<false><<<c->num_sms * 4, 256, 0, st>>>(sc, d_qs, out->lines, out->obase, L, d_bwt, d_sa, 0, d_t3, out->super);
            launch_check(c);
        }
        fetch_small(c, c->h_small, d_stats, (ST_COUNT + 2) * 8);
        BFQ_CUDA(cudaStreamSynchronize(st));
        c->arena.release(mk2);
        if (cnts[0] <= L.long_cap && cnts[1] <= L.r11_cap) break;
        // A list overflowed (a two-base cluster needs only two rows, so valid input can hold up to n/2 of them): the kernels
        // counted what they would have stored, and every pass rewrites all rows from the unchanged inputs, so run again with
        // lists of the counted size.
        if (attempt) BFQ_FAIL(BFQ_ECAPACITY, "internal cluster list overflow (long %u/%u, two-base %u/%u)", cnts[0], L.long_cap, cnts[1], L.r11_cap);
        L.long_cap = std::max(L.long_cap, cnts[0]);
        L.r11_cap = std::max(L.r11_cap, cnts[1]);
    }
    c->arena.release(mk);
    if (h[ST_ERR])
        BFQ_FAIL(BFQ_EFREQ3, "a cluster has more than two frequent symbols (the reference asserts, bfq_ext.cpp:683); raise -f");
    if (stats) {
        stats->clusters = h[ST_CLUSTERS]; stats->discarded = h[ST_DISCARDED]; stats->ambiguous = h[ST_AMBIGUOUS];
        stats->processed = h[ST_PROCESSED]; stats->bases_inside = h[ST_BASES]; stats->qs_smoothed = h[ST_SMOOTHED];
        stats->modified = h[ST_MODIFIED]; stats->nreads = out->nreads; stats->n = n;
    }
}

} // namespace bfq
