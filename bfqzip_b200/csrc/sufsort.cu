// sufsort.cu -- part A on the device: generalized suffix sort of a read collection by MSD
// radix (k symbols, stable LSD) + in-shared-memory string sort, with the eBWT / permuted-quality /
// LCP rows produced by the same kernel.
//
// Replaces: external/gsufsort/lib/utils.c:193-249 (cat_all),
// external/gsufsort/external/gsacak.c:1648-1930, 2503-2518 (gSACA-K + LCP) and
// external/gsufsort/main.c:848-867, 891-895 (bwt / bwt.qs / lcp writers).
//
// Text model: the reads are laid out back to back, each followed by ONE terminator (code 0);
// symbol codes 1..5 follow byte order (A<C<G<N<T), exactly gsufsort's "+1" remap
// (utils.c:229-246).  Every terminator is a distinct symbol ordered by read index
// (gsacak.c:113-133), so after sorting on the first k symbols (cut at the terminator) a key
// that contains a terminator is unique and is ranked by text position -- a stable sort from
// position order gives that for free.  Groups of equal terminator-free keys are finished by
// sorting the rest of their suffixes inside shared memory (group_tile_kernel).
//
// Layouts in HBM (N = sum(len+1) rows, N < 2^32):
//   T3   3-bit codes, big-endian bit stream, 64 symbols per three u64 words (fits L2 for cfg2)
//   QL   u16[N], quality of position p-1 | (symbols from p to its terminator) << 8
//   key  u64[N] x2, pos u32[N] x2   sort ping-pong: (k-symbol key, text position)
//   level >= 1 lists (members of groups too large for one tile) reuse the same ping-pong pair
#include "common.cuh"
#include <algorithm>

namespace bfq {

// ============================================================================ byte histogram
__global__ void __launch_bounds__(256)
hist256_kernel(const uint8_t *__restrict__ data, uint64_t n, unsigned long long *__restrict__ hist)
{
    __shared__ uint32_t sh[256];
    sh[threadIdx.x] = 0;
    __syncthreads();
    const uint64_t nvec = ((reinterpret_cast<uintptr_t>(data) & 15) == 0) ? n / 16 : 0;
    const uint4 *v = reinterpret_cast<const uint4 *>(data);
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        uint4 x = v[i];
        uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
        for (int a = 0; a < 4; a++)
#pragma unroll
            for (int b = 0; b < 4; b++) atomicAdd(&sh[(w[a] >> (8 * b)) & 255], 1u);
    }
    for (uint64_t i = nvec * 16 + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        atomicAdd(&sh[data[i]], 1u);
    __syncthreads();
    if (sh[threadIdx.x]) atomicAdd(&hist[threadIdx.x], (unsigned long long)sh[threadIdx.x]);
}

// two arrays under one synchronisation (the reads' bases and qualities in part A)
void byte_histogram2(bfq_ctx *c, const uint8_t *d_a, const uint8_t *d_b, uint64_t n, uint64_t hist_a[256], uint64_t hist_b[256])
{
    unsigned long long *d_hist = (unsigned long long *)c->d_small;   // the whole 4 KiB block: 2 x 256 x 8 bytes
    BFQ_CUDA(cudaMemsetAsync(d_hist, 0, 2 * 256 * 8, c->stream));
    if (n) {
        ProfScope ps(c, PC_HIST, 2 * n, 2);
        int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 8, div_up(n, 256 * 16) + 1);
        hist256_kernel<<<grid, 256, 0, c->stream>>>(d_a, n, d_hist);
        hist256_kernel<<<grid, 256, 0, c->stream>>>(d_b, n, d_hist + 256);
    }
    fetch_small(c, c->h_small, d_hist, 2 * 256 * 8);
    BFQ_CUDA(cudaStreamSynchronize(c->stream));
    for (int i = 0; i < 256; i++) { hist_a[i] = c->h_small[i]; hist_b[i] = c->h_small[256 + i]; }
}

void byte_histogram(bfq_ctx *c, const uint8_t *d_data, uint64_t n, uint64_t hist_out[256])
{
    unsigned long long *d_hist = (unsigned long long *)c->d_small;   // 256*8 = 2 KiB of the 4 KiB
    BFQ_CUDA(cudaMemsetAsync(d_hist, 0, 256 * 8, c->stream));
    if (n) {
        ProfScope ps(c, PC_HIST, n);
        int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 8, div_up(n, 256 * 16) + 1);
        hist256_kernel<<<grid, 256, 0, c->stream>>>(d_data, n, d_hist);
    }
    fetch_small(c, c->h_small, d_hist, 256 * 8);
    BFQ_CUDA(cudaStreamSynchronize(c->stream));
    for (int i = 0; i < 256; i++) hist_out[i] = c->h_small[i];
}

// ============================================================================ text packing
struct Lut256 { uint8_t v[256]; };

// one warp per read: T8[p] = code (terminator code 0 after the read);
// QL[p] = quality of position p-1 | (symbols from p to the read's terminator) << 8
__global__ void __launch_bounds__(256)
pack_reads_kernel(const uint8_t *__restrict__ bases, const uint8_t *__restrict__ quals,
                  const uint64_t *__restrict__ off, uint64_t nreads, Lut256 lut, uint8_t eos,
                  uint8_t *__restrict__ T8, uint16_t *__restrict__ QL)
{
    __shared__ uint8_t s_lut[256];
    s_lut[threadIdx.x] = lut.v[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const uint64_t warp0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t d = warp0; d < nreads; d += nwarps) {
        const uint64_t s = off[d], e = off[d + 1], len = e - s, dst = s + d;
        for (uint64_t j = lane; j < len; j += 32) {
            T8[dst + j] = s_lut[bases[s + j]];
            QL[dst + j] = (uint16_t)((j ? quals[s + j - 1] : eos) | ((len - j) << 8));
        }
        if (lane == 0) { T8[dst + len] = 0; QL[dst + len] = quals[e - 1]; }
    }
}

// 64 symbols -> three u64 words, first symbol in the most significant bits
__global__ void __launch_bounds__(256)
pack3_kernel(const uint8_t *__restrict__ T8, uint64_t n, uint64_t ngroups, uint64_t *__restrict__ T3)
{
    const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= ngroups) return;
    const uint64_t base = g * 64;
    // 8 codes (one 64-bit load, first code in the low byte) -> 24 bits, first code in the top 3 bits
    auto pack8 = [](uint64_t x) -> uint64_t {
        uint64_t r = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) r |= ((x >> (8 * j)) & 7ull) << (21 - 3 * j);
        return r;
    };
    uint64_t f[8];                                        // 8 fields of 24 bits = 192 bits
    if (base + 64 <= n) {
        const uint4 *src = reinterpret_cast<const uint4 *>(T8 + base);      // T8 is 256-byte aligned, base a multiple of 64
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint4 v = __ldcs(src + q);
            f[2 * q] = pack8((uint64_t)v.x | ((uint64_t)v.y << 32));
            f[2 * q + 1] = pack8((uint64_t)v.z | ((uint64_t)v.w << 32));
        }
    } else {
#pragma unroll
        for (int q = 0; q < 8; q++) {
            uint64_t x = 0;
            for (int j = 0; j < 8; j++) {
                const uint64_t i = base + 8 * q + j;
                if (i < n) x |= (uint64_t)T8[i] << (8 * j);
            }
            f[q] = pack8(x);
        }
    }
    T3[g * 3 + 0] = (f[0] << 40) | (f[1] << 16) | (f[2] >> 8);
    T3[g * 3 + 1] = (f[2] << 56) | (f[3] << 32) | (f[4] << 8) | (f[5] >> 16);
    T3[g * 3 + 2] = (f[5] << 48) | (f[6] << 24) | f[7];
}


// ============================================================================ group tile kernel
// After a stable radix sort on a key of `ksym` symbols (cut at the terminator) the list is a
// sequence of GROUPS (runs of equal keys).  A key that holds a terminator is unique (every
// terminator is its own symbol, gsacak.c:113-133) and already in read order because the sort is
// stable from position order, so such rows are final.  Every other group shares depth = depth0 +
// ksym symbols and is finished here, inside shared memory: the rest of every member's suffix
// (<= W words of 21 symbols, cut at its terminator) is staged from the packed text, one warp
// sorts the group with a bitonic network over 16-bit member indices (ties = identical strings ->
// text position, i.e. read index), and the rows leave the SM complete: SA, bwt, qs and lcp.
// The LCP of a sorted row is depth + the common prefix of two staged strings; the LCP of a
// group head comes from the two keys.  This fuses what used to be the prefix-doubling rounds and
// the eBWT/LCP gather (main.c:848-895) into one pass over the sorted list.
//
// A block owns the groups that START in its nominal range of GT_T list slots and stages
// GT_S = GT_T + GT_E slots; a group that does not end inside the staged window is DEFERRED: its
// start is appended to a list and the host runs one more level on those groups with the next
// symbols as key (MSD recursion; at most ceil(max_len / 14) levels, only repeats get there).
#ifndef GT_THREADS_N
#define GT_THREADS_N 256
#endif
constexpr int GT_THREADS = GT_THREADS_N;
#ifndef GT_T_ROWS
#define GT_T_ROWS 1024
#endif
#ifndef GT_E_ROWS
#define GT_E_ROWS 256
#endif
constexpr int GT_T = GT_T_ROWS;
constexpr int GT_E = GT_E_ROWS;
constexpr int GT_S = GT_T + GT_E;
constexpr int GT_HW = (GT_S + 1 + 31) / 32 + 1;       // head bitmap words (bit GT_S is the sentinel)
constexpr int GT_GMAX = GT_S / 2;
#ifndef GT_BATCH_ROWS
#define GT_BATCH_ROWS 48
#endif
constexpr int GT_BATCH = GT_BATCH_ROWS;                          // rows of a tile whose groups form one batch (see phase 1b)
constexpr int GT_RW = 13;                             // words of 21 symbols that hold the longest suffix tail (255 symbols)

template <typename P>
struct GroupTileArgsT {
    const P *pos;               // list payload: text position (u32, or u64 for collections of 2^32 rows and more)
    uint32_t m;                 // list length
    int level0;                 // 1: slot == list index, key has no group id
    int ksym;                   // symbols in this level's key (low 3*ksym bits)
    int depth0;                 // symbols known equal before this key (0 at level 0)
    int W;                      // words of 21 symbols to stage per member
    const uint64_t *T3;
    const uint16_t *QL;         // per text position p: quality of p-1 | (symbols from p to its terminator) << 8
    const uint32_t *g_slot;     // level >= 1: SA slot of the first member of list group g
    const uint32_t *g_off;      // level >= 1: list offset of the first member of list group g
    P *SA;                      // optional
    uint8_t *bwt, *qs, *lcp;    // optional (all three or none)
    uint64_t code2byte;
    uint8_t eos;
    uint32_t *defer_count;      // [0] groups deferred, [1] members (filled by defer_len_kernel)
    uint32_t *defer_start;      // list index of the head of every deferred group
    uint32_t defer_cap;
    uint32_t tile0;             // first tile of this launch (level 0 may be launched in slices)
    uint64_t prev_key;          // level 0 of a key range: key of the global row just before this list (has_prev)
    int has_prev;
    uint64_t kmask;             // key bits that take part in group equality (level 0: the 3*ksym symbol bits)
    int embed;                  // level 0, k <= 16: key bits 63..48 hold QL[pos]
    uint32_t group_cap;         // groups with more members are deferred even when they fit (test hook)
    uint32_t batch_min_groups;  // tiles with more groups than this are ordered batch-wise
    int bulk_tile;              // phase 0: the window comes in with two cp.async.bulk copies (mbarrier completion) instead of per-thread loads
};
using GroupTileArgs = GroupTileArgsT<uint32_t>;

__device__ __forceinline__ uint64_t cut_at_terminator(uint64_t w, bool &terminated)
{
    if (terminated) return 0ull;
    const int f = leading_nonzero_fields(w);
    if (f < 21) { terminated = true; return f ? (w & (~0ull << (64 - 3 * f))) : 0ull; }
    return w;
}

// LCP of the head of a group with the row before it, from the two keys (symbol parts)
template <typename K>
__device__ __forceinline__ uint32_t head_lcp(K kprev, K kcur, int ksym)
{
    // left-align the ksym fields so that field 0 sits in bits 63..61
    const uint64_t a = (uint64_t)kprev << (64 - 3 * ksym), b = (uint64_t)kcur << (64 - 3 * ksym);
    const uint64_t x = a ^ b;
    const int fm = x ? (int)div3_small((uint32_t)__clzll((long long)x)) : ksym;
    // symbols before the terminator (fields below ksym are zero padding -> count as terminator)
    const int z = leading_nonzero_fields(b);
    return (uint32_t)min(min(fm, z), ksym);
}

// ---- in-group ordering -------------------------------------------------------------------
// Rows of a group start in position order as ONE unresolved run at depth 0.  Per iteration every
// member of an unresolved run is compared once with the run's REFERENCE (its longest member):
// l = common prefix, c / r = the member's / the reference's symbol at l (0 = terminator).  With
//   type L (c < r): before the reference, ordered by (l ascending, c)
//   type E (c = r = 0): identical to the reference
//   type G (c > r): after it, ordered by (l descending, c)
// the key (run, type, l', c, position index) sorts the run; members with equal (run, type, l, c != 0)
// share l + 1 symbols and form a new run for the next iteration, everything else is final.  Two
// neighbours with different key classes have LCP = min(l, l') -- no string compare.  Typical read
// data: most members are prefixes of the longest one and a single iteration resolves the group.
struct RefCmp { uint32_t l, c, r; };

template <typename P>
__device__ __forceinline__ RefCmp ref_compare(const uint64_t *__restrict__ T3, const P *s_pos, const uint8_t *s_len,
                                              int depth, uint32_t mem, uint32_t ref, uint32_t D)
{
    // The strings are read straight from the packed text (L1 / pinned L2): a member is compared once per
    // iteration, so staging it in shared memory would cost more than it saves (and occupancy with it).
    const uint32_t lm = s_len[mem];                       // the reference is its run's longest member: lm <= its length
    const uint64_t pm = (uint64_t)s_pos[mem] + depth, pr = (uint64_t)s_pos[ref] + depth;
    RefCmp o;
    o.l = lm; o.c = 0u;
    if (mem != ref) {
        const uint32_t jend = (lm + 20u) / 21u;           // a difference cannot lie past the member's terminator
        for (uint32_t j = D / 21u; j < jend; j++) {
            const uint64_t wm = text_window(T3, pm + 21u * j), wr = text_window(T3, pr + 21u * j);
            const uint64_t x = wm ^ wr;
            if (x) {
                const uint32_t f = div3_small((uint32_t)__clzll((long long)x));
                if (21u * j + f < lm) {                   // both symbols at the difference come from the two windows
                    o.l = 21u * j + f;
                    o.c = (uint32_t)((wm << (3u * f)) >> 61);
                    o.r = (uint32_t)((wr << (3u * f)) >> 61);
                    return o;
                }
                break;
            }
        }
    }
    // The member is a prefix of the reference (or the reference itself): c = its terminator.  What the reference holds
    // there does not matter -- a terminator sorts below every symbol and equal strings keep their position order --
    // so the row is classed "below the reference at l = its length" without reading it.
    o.r = 1u;
    return o;
}

// First iteration of a group: every member is compared with the same string V, whose words the warp keeps in shared
// memory.  ANY fixed string orders the members correctly (the (type, l, c) classes and the min(l, l') LCP rule only need
// "the same string for everybody"), so V is the group's longest member a corrected by a per-symbol majority vote with
// the second and third longest members b and c wherever all three reach: one sequencing error in a -- more than half of
// the longest reads of a 100 bp collection have one -- would otherwise make every longer member tie at that symbol and
// cost the whole group another iteration.
__device__ __forceinline__ RefCmp v_compare(const uint64_t *__restrict__ T3, const uint64_t *s_v, uint64_t pm, uint32_t lm)
{
    RefCmp o;
    o.l = lm; o.c = 0u;
    const uint32_t jend = (lm + 20u) / 21u;
    for (uint32_t j = 0; j < jend; j++) {
        const uint64_t wm = text_window(T3, pm + 21u * j), wr = s_v[j];
        const uint64_t x = wm ^ wr;
        if (x) {
            const uint32_t f = div3_small((uint32_t)__clzll((long long)x));
            if (21u * j + f < lm) {
                o.l = 21u * j + f;
                o.c = (uint32_t)((wm << (3u * f)) >> 61);
                o.r = (uint32_t)((wr << (3u * f)) >> 61);
                return o;
            }
            break;
        }
    }
    o.r = 1u;                                             // a prefix of V: below it at l = its length (see ref_compare)
    return o;
}

// Key fields above the position index (IB bits): c [2:0], l' [10:3], type [12:11], run [..:13]
__device__ __forceinline__ uint32_t class_of(const RefCmp &o)
{
    const uint32_t type = o.c < o.r ? 0u : (o.c == o.r ? 1u : 2u);
    const uint32_t lkey = type == 2u ? 255u - o.l : o.l;
    return (type << 11) | (lkey << 3) | o.c;
}
__device__ __forceinline__ uint32_t class_l(uint32_t cls)
{
    const uint32_t lk = (cls >> 3) & 255u;
    return ((cls >> 11) & 3u) == 2u ? 255u - lk : lk;
}

// groups of at most 32 * R rows: one row per (slot, lane), keys in registers, shuffles only.
// MULTI: the rows [gs, gs + len) are a BATCH of whole small groups and single rows (low coverage leaves a few rows per
// group; a warp per group would idle most lanes): every group starts as its own run -- the same state a single group
// is in after its first iteration -- and all of them are ordered by the same sorting network.
template <int R, bool MULTI, typename P>
__device__ __forceinline__ void order_group_regs(uint32_t gs, uint32_t len, int lane, int depth, const uint64_t *__restrict__ T3,
                                                 const P *s_pos, const uint8_t *s_len, uint32_t *s_best, uint16_t *s_sorted,
                                                 uint8_t *s_lcpout, uint64_t *s_refw, const uint16_t *s_gs)
{
    constexpr uint32_t IB = 6, IM = 63u;                  // position index bits
    uint32_t key[R], runa[R], rdepth[R];
#pragma unroll
    for (int s = 0; s < R; s++) {
        const uint32_t t = s * 32 + lane;
        key[s] = t < len ? t : 0xFFFFFFFFu;
        runa[s] = 0;
        rdepth[s] = t < len ? 0u : 0xFFu;
    }
    bool first = !MULTI;
    if constexpr (MULTI) {
        bool work = false;
#pragma unroll
        for (int s = 0; s < R; s++) {
            const uint32_t t = s * 32 + lane;
            if (t < len) {
                const uint32_t row = gs + t, g0 = s_gs[row];      // (0xFFFF: a group handed to the next level, its rows stay put)
                const bool head = g0 == row;
                const bool single = g0 == 0xFFFFu || (head && (t + 1 >= len || s_gs[row + 1] != g0));
                runa[s] = single ? t : g0 - gs;
                if (single) rdepth[s] = 0xFFu;
                else { work = true; if (head) s_best[row] = 0; }
            }
        }
        if (!__any_sync(0xffffffffu, work)) return;       // single rows only: s_sorted is the identity already
        __syncwarp();
    } else {   // V: the three longest members (valid bit | length | row), then lane j builds word j
        uint32_t v[R], a = 0, b = 0, cc = 0;
#pragma unroll
        for (int s = 0; s < R; s++) {
            const uint32_t t = s * 32 + lane;
            v[s] = t < len ? (0x80000000u | ((uint32_t)s_len[gs + t] << 16) | t) : 0u;
            a = max(a, v[s]);
        }
        a = __reduce_max_sync(0xffffffffu, a);
#pragma unroll
        for (int s = 0; s < R; s++) if (v[s] != a) b = max(b, v[s]);
        b = __reduce_max_sync(0xffffffffu, b);
#pragma unroll
        for (int s = 0; s < R; s++) if (v[s] != a && v[s] != b) cc = max(cc, v[s]);
        cc = __reduce_max_sync(0xffffffffu, cc);
        const uint32_t la = (a >> 16) & 255u, lc = (cc >> 16) & 255u;
        const uint32_t nw = la / 21u + 1u;                // the word that holds a's terminator included
        if ((uint32_t)lane < nw) {
            const uint32_t j = lane;
            uint64_t w = text_window(T3, (uint64_t)s_pos[gs + (a & 0xFFFFu)] + depth + 21u * j);
            if (cc && lc > 21u * j) {                     // symbols of this word that all three members reach
                const uint32_t cnt = min(21u, lc - 21u * j);
                const uint64_t vote = ~0ull << (64 - 3 * cnt);
                const uint64_t wb = text_window(T3, (uint64_t)s_pos[gs + (b & 0xFFFFu)] + depth + 21u * j);
                const uint64_t wc = text_window(T3, (uint64_t)s_pos[gs + (cc & 0xFFFFu)] + depth + 21u * j);
                const uint64_t maj = (w & wb) | (w & wc) | (wb & wc);          // per symbol: the value two of them share
                w = (maj & vote) | (w & ~vote);
            }
            s_refw[j] = w;
        }
        __syncwarp();
    }
    while (true) {
        if (!first) {
#pragma unroll
            for (int s = 0; s < R; s++)
                if (rdepth[s] != 0xFFu) {
                    const uint32_t mem = gs + (key[s] & IM);
                    atomicMax(&s_best[gs + runa[s]], ((uint32_t)s_len[mem] << 16) | mem);
                }
            __syncwarp();
        }
#pragma unroll
        for (int s = 0; s < R; s++) {
            const uint32_t t = s * 32 + lane;
            if (t < len) {
                const uint32_t idx = key[s] & IM;
                if (rdepth[s] == 0xFFu) key[s] = (t << (13 + IB)) | idx;
                else {
                    RefCmp o;
                    if (!MULTI && first) o = v_compare(T3, s_refw, (uint64_t)s_pos[gs + idx] + depth, s_len[gs + idx]);
                    else o = ref_compare(T3, s_pos, s_len, depth, gs + idx, s_best[gs + runa[s]] & 0xFFFFu, rdepth[s]);
                    key[s] = (runa[s] << (13 + IB)) | (class_of(o) << IB) | idx;
                }
            }
        }
        first = false;
        // bitonic network, ascending comparators only (rows >= len hold +inf and never move)
#pragma unroll
        for (uint32_t k = 2; k <= 32u * R; k <<= 1) {
            if ((k >> 1) >= len) break;
            const uint32_t half = k >> 1;
            if (k <= 32) {
#pragma unroll
                for (int s = 0; s < R; s++) {
                    const uint32_t o = __shfl_xor_sync(0xffffffffu, key[s], k - 1);
                    key[s] = (lane & half) ? max(key[s], o) : min(key[s], o);
                }
            } else {                                      // k == 64: partner = other slot, lane ^ 31
                const uint32_t o0 = __shfl_xor_sync(0xffffffffu, key[R - 1], 31);
                const uint32_t o1 = __shfl_xor_sync(0xffffffffu, key[0], 31);
                key[0] = min(key[0], o0);
                key[R - 1] = max(key[R - 1], o1);
            }
#pragma unroll
            for (uint32_t jj = 16; jj > 0; jj >>= 1) {
                if (jj < half) {
#pragma unroll
                    for (int s = 0; s < R; s++) {
                        const uint32_t o = __shfl_xor_sync(0xffffffffu, key[s], jj);
                        key[s] = (lane & jj) ? max(key[s], o) : min(key[s], o);
                    }
                }
            }
        }
        // new runs and final LCPs
        bool any = false;
        uint32_t startbits[R];
        uint32_t cls[R];
        bool st[R], un[R];
#pragma unroll
        for (int s = 0; s < R; s++) {
            const uint32_t t = s * 32 + lane;
            uint32_t prevk = __shfl_up_sync(0xffffffffu, key[s], 1);
            uint32_t nextk = __shfl_down_sync(0xffffffffu, key[s], 1);
            if (s > 0) { const uint32_t x = __shfl_sync(0xffffffffu, key[s - 1], 31); if (lane == 0) prevk = x; }
            if (s + 1 < R) { const uint32_t x = __shfl_sync(0xffffffffu, key[s + 1], 0); if (lane == 31) nextk = x; }
            else if (lane == 31) nextk = 0xFFFFFFFFu;
            st[s] = true; un[s] = false; cls[s] = 0;
            if (t < len && rdepth[s] != 0xFFu) {
                const uint32_t kt = key[s] >> IB;
                cls[s] = kt & 0x1FFFu;
                const bool open = (kt & 7u) != 0;
                bool tiedprev = false;
                if (t > runa[s]) {
                    const uint32_t kp = prevk >> IB;
                    tiedprev = kp == kt && open;
                    if (!tiedprev) s_lcpout[gs + t] = (uint8_t)min(class_l(kt), class_l(kp));
                }
                const bool tiednext = (nextk >> IB) == kt && open;
                st[s] = !tiedprev;
                un[s] = tiedprev || tiednext;
            }
            startbits[s] = __ballot_sync(0xffffffffu, t < len && st[s]);
            any |= un[s];
        }
        if (!__any_sync(0xffffffffu, any)) break;         // everything resolved: no run bookkeeping needed
#pragma unroll
        for (int s = 0; s < R; s++) {
            const uint32_t t = s * 32 + lane;
            if (t < len) {
                const uint32_t below = startbits[s] & (0xffffffffu >> (31 - lane));
                uint32_t a = below ? s * 32 + 31 - __clz(below) : 0u;
                if (!below && s > 0) a = 31 - __clz(startbits[0]);          // R <= 2: the run began in slot 0
                runa[s] = a;
                rdepth[s] = un[s] ? class_l(cls[s]) + 1 : 0xFFu;
                if (st[s] && un[s]) s_best[gs + t] = 0;
            }
        }
        __syncwarp();
    }
#pragma unroll
    for (int s = 0; s < R; s++) {
        const uint32_t t = s * 32 + lane;
        if (t < len) s_sorted[gs + t] = (uint16_t)(gs + (key[s] & IM));
    }
}

// groups of any size that fits the window: keys in shared memory, lanes stride over the rows
template <typename P>
__device__ __forceinline__ void order_group_smem(uint32_t gs, uint32_t len, int lane, int depth, const uint64_t *__restrict__ T3,
                                                 const P *s_pos, const uint8_t *s_len, uint32_t *s_best, uint16_t *s_sorted,
                                                 uint8_t *s_lcpout, uint64_t *s_key, uint16_t *s_runa, uint8_t *s_rdepth)
{
    uint32_t n2 = 2;
    while (n2 < len) n2 <<= 1;
    for (uint32_t t = lane; t < len; t += 32) { s_runa[gs + t] = (uint16_t)gs; s_rdepth[gs + t] = 0; s_key[gs + t] = t; }
    if (lane == 0) s_best[gs] = 0;
    __syncwarp();
    uint64_t *key = s_key + gs;
    while (true) {
        for (uint32_t t = lane; t < len; t += 32) {
            const uint32_t row = gs + t;
            if (s_rdepth[row] != 0xFF) {
                const uint32_t mem = gs + (uint32_t)(key[t] & 0x7FFu);
                atomicMax(&s_best[s_runa[row]], ((uint32_t)s_len[mem] << 16) | mem);
            }
        }
        __syncwarp();
        for (uint32_t t = lane; t < len; t += 32) {
            const uint32_t row = gs + t;
            const uint32_t idx = (uint32_t)(key[t] & 0x7FFu);
            const uint32_t D = s_rdepth[row];
            if (D == 0xFF) key[t] = ((uint64_t)t << 24) | idx;
            else {
                const uint32_t a = s_runa[row];
                const RefCmp o = ref_compare(T3, s_pos, s_len, depth, gs + idx, s_best[a] & 0xFFFFu, D);
                key[t] = ((uint64_t)(a - gs) << 24) | ((uint64_t)class_of(o) << 11) | idx;
            }
        }
        __syncwarp();
        for (uint32_t k = 2; k <= n2; k <<= 1) {
            const uint32_t half = k >> 1;
            for (uint32_t t = lane; t < (n2 >> 1); t += 32) {
                const uint32_t i = ((t & ~(half - 1)) << 1) | (t & (half - 1));
                const uint32_t l = i ^ (k - 1);
                if (l < len) { const uint64_t x = key[i], y = key[l]; if (y < x) { key[i] = y; key[l] = x; } }
            }
            __syncwarp();
            for (uint32_t jj = half >> 1; jj > 0; jj >>= 1) {
                for (uint32_t t = lane; t < (n2 >> 1); t += 32) {
                    const uint32_t i = ((t & ~(jj - 1)) << 1) | (t & (jj - 1));
                    const uint32_t l = i | jj;
                    if (l < len) { const uint64_t x = key[i], y = key[l]; if (y < x) { key[i] = y; key[l] = x; } }
                }
                __syncwarp();
            }
        }
        bool any = false;
        uint32_t carry = gs;
        for (uint32_t base = 0; base < len; base += 32) {
            const uint32_t t = base + lane, row = gs + t;
            const bool valid = t < len;
            bool start = true, unres = false;
            uint32_t newdepth = 0xFF;
            if (valid && s_rdepth[row] != 0xFF) {
                const uint32_t kt = (uint32_t)(key[t] >> 11);
                const bool open = (kt & 7u) != 0;
                bool tiedprev = false;
                if (row > s_runa[row]) {
                    const uint32_t kp = (uint32_t)(key[t - 1] >> 11);
                    tiedprev = kp == kt && open;
                    if (!tiedprev) s_lcpout[row] = (uint8_t)min(class_l(kt), class_l(kp));
                }
                const bool tiednext = t + 1 < len && (uint32_t)(key[t + 1] >> 11) == kt && open;
                start = !tiedprev;
                unres = tiedprev || tiednext;
                newdepth = class_l(kt) + 1;
            }
            const uint32_t bits = __ballot_sync(0xffffffffu, valid && start);
            const uint32_t below = bits & (0xffffffffu >> (31 - lane));
            const uint32_t anew = below ? gs + base + 31 - __clz(below) : carry;
            __syncwarp();
            if (valid) {
                s_runa[row] = (uint16_t)anew;
                s_rdepth[row] = unres ? (uint8_t)newdepth : (uint8_t)0xFF;
                if (start && unres) s_best[row] = 0;
            }
            if (bits) carry = gs + base + 31 - __clz(bits);
            any |= unres;
        }
        __syncwarp();
        if (!__any_sync(0xffffffffu, any)) break;
    }
    for (uint32_t t = lane; t < len; t += 32) s_sorted[gs + t] = (uint16_t)(gs + (uint32_t)(key[t] & 0x7FFu));
}

template <typename K, typename P>
__global__ void __launch_bounds__(GT_THREADS, 1280 / GT_THREADS)
group_tile_kernel(const K *__restrict__ keys, GroupTileArgsT<P> A)
{
    extern __shared__ __align__(16) unsigned char gt_smem[];
    uint64_t *s_key = reinterpret_cast<uint64_t *>(gt_smem) + 1;                    // [GT_S + 2]; &s_key[1] is 16-byte aligned (bulk copy target)
    P *s_pos = reinterpret_cast<P *>(s_key + GT_S + 3);                                 // [GT_S] text position of the row's suffix (16-byte aligned)
    uint32_t *s_best = reinterpret_cast<uint32_t *>(s_pos + GT_S);                                                // [GT_S] (length << 16 | row) of a run's reference
    uint32_t *s_head = s_best + GT_S;                                               // [GT_HW]
    uint16_t *s_prev = reinterpret_cast<uint16_t *>(s_head + GT_HW);                // [GT_S] previous symbol code | its quality << 8
    uint16_t *s_gs = s_prev + GT_S;                                                 // [GT_S] group start or 0xFFFF (row not owned)
    uint16_t *s_sorted = s_gs + GT_S;                                               // [GT_S] member at sorted row
    uint16_t *s_runa = s_sorted + GT_S;                                             // [GT_S] first row of the run a row is in
    uint16_t *s_gstart = s_runa + GT_S;                                             // [GT_GMAX]
    uint16_t *s_glen = s_gstart + GT_GMAX;                                          // [GT_GMAX]
    uint8_t *s_len = reinterpret_cast<uint8_t *>(s_glen + GT_GMAX);                 // [GT_S] staged symbols before the terminator
    uint8_t *s_rdepth = s_len + GT_S;                                               // [GT_S] depth of the row's run, 0xFF = resolved
    uint8_t *s_lcpout = s_rdepth + GT_S;                                            // [GT_S] LCP with the previous sorted row (staged part)
    uint64_t *s_refw = reinterpret_cast<uint64_t *>((reinterpret_cast<uintptr_t>(s_lcpout + GT_S) + 7) & ~(uintptr_t)7) +
                       (threadIdx.x >> 5) * GT_RW;                                  // [warps][GT_RW] words of the group's reference
    __shared__ uint32_t s_ngroups, s_next, s_nsmall;     // s_ngroups: work items listed from the front, s_nsmall: small groups listed from the back

    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t t0 = (A.tile0 + blockIdx.x) * (uint32_t)GT_T;
    const uint32_t m = A.m;
    const int ksym = A.ksym;
    const int depth = A.depth0 + ksym;
    const int kshift = 3 * ksym;                          // group id sits above the symbols (level >= 1)
    if (tid == 0) { s_ngroups = 0; s_next = 0; s_nsmall = 0; }

    // ---- phase 0: the window's keys and positions come in with one batch of independent loads per thread (slot r of
    // the window sits at s_key[r + 1]; [0] is the key before the window, [GT_S + 1] the one after it), then the head
    // bitmap: bit r <=> list slot t0 + r starts a group
    __shared__ __align__(8) uint64_t s_bar;
    const bool bulk = A.bulk_tile && sizeof(K) == 8 && (uint64_t)t0 + GT_S <= m;      // a full window: sizes are multiples of 16 bytes
    if (bulk) {
        // Blackwell/Hopper bulk asynchronous copy (SASS UBLKCP): ONE thread moves the window's keys (10 KB) and positions
        // straight into shared memory; the copy engine signals the mbarrier with the byte count.  No thread stages data
        // through registers (the per-thread path below holds 15 registers and issues 10 loads + 10 shared stores).
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
        if (tid == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (tid == 0) {
            constexpr uint32_t kbytes = GT_S * sizeof(K), pbytes = GT_S * sizeof(P);
            uint64_t pol;
            asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));      // the list streams through once
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kbytes + pbytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                         ::"r"((uint32_t)__cvta_generic_to_shared(s_key + 1)), "l"(keys + t0), "r"(kbytes), "r"(bar), "l"(pol) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                         ::"r"((uint32_t)__cvta_generic_to_shared(s_pos)), "l"(A.pos + t0), "r"(pbytes), "r"(bar), "l"(pol) : "memory");
            s_key[0] = t0 > 0 ? keys[t0 - 1] : (K)0;
        }
        if (tid == 32) s_key[GT_S + 1] = (uint64_t)t0 + GT_S < m ? keys[(uint64_t)t0 + GT_S] : (K)0;
        asm volatile("{\n\t.reg .pred p;\n\tBULK_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra BULK_DONE;\n\tbra BULK_WAIT;\n\tBULK_DONE:\n\t}"
                     ::"r"(bar) : "memory");
    } else {
        constexpr int NB = (GT_S + GT_THREADS - 1) / GT_THREADS;
        K kk[NB]; P pp[NB];
#pragma unroll
        for (int i = 0; i < NB; i++) {
            const uint64_t j = (uint64_t)t0 + i * GT_THREADS + tid;
            const bool in = i * GT_THREADS + tid < GT_S && j < m;
            kk[i] = in ? __ldcs(&keys[j]) : (K)0;
            pp[i] = in ? __ldcs(&A.pos[j]) : (P)0;
        }
        if (tid == 0) s_key[0] = t0 > 0 ? keys[t0 - 1] : (K)0;
        if (tid == 32) s_key[GT_S + 1] = (uint64_t)t0 + GT_S < m ? keys[(uint64_t)t0 + GT_S] : (K)0;
#pragma unroll
        for (int i = 0; i < NB; i++)
            if (i * GT_THREADS + tid < GT_S) { s_key[i * GT_THREADS + tid + 1] = kk[i]; s_pos[i * GT_THREADS + tid] = pp[i]; }
    }
    __syncthreads();
    for (int base = 0; base < GT_HW * 32; base += GT_THREADS) {
        const int r = base + tid;
        const uint64_t j = (uint64_t)t0 + r;
        bool h = true;                                    // slots past the list end count as heads
        if (r <= GT_S && j < m && j > 0) {
            const K kc = s_key[r + 1] & (K)A.kmask, kp = s_key[r] & (K)A.kmask;
            h = (kc != kp) || ((kc & (K)7) == 0);
        }
        const uint32_t bits = __ballot_sync(0xffffffffu, h);
        if (lane == 0 && (r >> 5) < GT_HW) s_head[r >> 5] = bits;
    }
    __syncthreads();

    auto next_head = [&](int r) -> int {                  // first head > r (exists: sentinel region)
        int w = (r + 1) >> 5;
        uint32_t bits = s_head[w] & (0xffffffffu << ((r + 1) & 31));
        while (!bits) bits = s_head[++w];
        return (w << 5) + __ffs(bits) - 1;
    };
    auto prev_head = [&](int r) -> int {                  // last head <= r, or -1
        int w = r >> 5;
        uint32_t bits = s_head[w] & (0xffffffffu >> (31 - (r & 31)));
        while (!bits) { if (--w < 0) return -1; bits = s_head[w]; }
        return (w << 5) + 31 - __clz(bits);
    };
    auto slot_of = [&](uint32_t j) -> uint32_t {
        if (A.level0) return j;
        const uint32_t g = (uint32_t)((uint64_t)keys[j] >> kshift);
        return A.g_slot[g] + (j - A.g_off[g]);
    };

    // ---- phase 1: ownership, group list, staging of previous symbol / suffix words
    for (int r = tid; r < GT_S; r += GT_THREADS) {
        const uint64_t j = (uint64_t)t0 + r;
        uint16_t gs16 = 0xFFFF;
        if (j < m) {
            const int gs = prev_head(r);
            if (gs >= 0 && gs < GT_T) {
                const int ge = next_head(r);
                const int size = ge - gs;
                const bool fits = ge <= GT_S && (uint32_t)size <= A.group_cap;
                if (r == gs) {
                    // head row: its LCP comes from the keys and is final at this level
                    if (A.lcp) {
                        bool write = true; uint32_t l = 0;
                        if (j == 0) {
                            write = A.level0;
                            if (A.level0 && A.has_prev) l = head_lcp<K>((K)A.prev_key, s_key[1] & (K)A.kmask, ksym);
                        }
                        else {
                            const K kc = s_key[r + 1] & (K)A.kmask, kp = s_key[r] & (K)A.kmask;
                            if (!A.level0 && ((uint64_t)kc >> kshift) != ((uint64_t)kp >> kshift)) write = false;   // written one level up
                            else l = A.depth0 + head_lcp<K>(kp, kc, ksym);
                        }
                        if (write) A.lcp[slot_of((uint32_t)j)] = (uint8_t)l;
                    }
                    if (size > 1) {
                        if (fits) {
                            // groups of more than 32 rows (the two-slot network, twice the work) are listed from the front and
                            // handed out first, the others from the back: the warps finish closer together (longest first)
                            const uint32_t g = size > 32 ? atomicAdd(&s_ngroups, 1u) : (uint32_t)GT_GMAX - 1u - atomicAdd(&s_nsmall, 1u);
                            s_gstart[g] = (uint16_t)gs; s_glen[g] = (uint16_t)size;
                        } else {
                            const uint32_t d = atomicAdd(A.defer_count, 1u);
                            if (d < A.defer_cap) A.defer_start[d] = (uint32_t)j;
                        }
                    }
                }
                if (fits) {
                    gs16 = (uint16_t)gs;
                    const P p = s_pos[r];
                    s_sorted[r] = (uint16_t)r;
                    const uint32_t ql = A.embed ? (uint32_t)((uint64_t)s_key[r + 1] >> 48) : (uint32_t)A.QL[p];
                    if (A.bwt) {
                        const uint32_t c = p > 0 ? sym_code(A.T3, (uint64_t)p - 1) : 0u;
                        s_prev[r] = (uint16_t)(c | ((c ? (ql & 255u) : (uint32_t)A.eos) << 8));
                    }
                    s_len[r] = (uint8_t)max(0, (int)(ql >> 8) - depth);   // symbols of this suffix past the shared prefix
                }
            }
        }
        s_gs[r] = gs16;
    }
    __syncthreads();

    // ---- phase 1b: a tile of many small groups (low coverage) is cut into BATCHES instead -- batch i = the groups that
    // start in rows [i * GT_BATCH, (i + 1) * GT_BATCH) -- and a warp orders a whole batch at once.  Only the last group of a
    // batch can be longer than GT_BATCH rows; it becomes a work item of its own when the batch would exceed 64 rows.
    if (s_ngroups + s_nsmall > A.batch_min_groups) {     // (block-uniform)
        __syncthreads();
        if (tid == 0) { s_ngroups = 0; s_nsmall = 0; }
        __syncthreads();
        constexpr int NBATCH = (GT_T + GT_BATCH - 1) / GT_BATCH;
        if (tid < NBATCH) {
            const int lo = tid * GT_BATCH, hi = min(lo + GT_BATCH, GT_T);
            const int last = (int)min((uint64_t)GT_S, (uint64_t)m - t0);          // rows of the window inside the list
            const int a = next_head(lo - 1);
            int b = min(next_head(hi - 1), last);
            if (a < hi && a < b) {
                if (s_gs[b - 1] == 0xFFFFu) b = prev_head(b - 1);               // the last group does not fit the window: deferred
                if (b - a > 64) {
                    const int ls = prev_head(b - 1);
                    const uint32_t g = atomicAdd(&s_ngroups, 1u);
                    s_gstart[g] = (uint16_t)ls; s_glen[g] = (uint16_t)(b - ls);
                    b = ls;
                }
                if (b - a > 1) {
                    const uint32_t g = atomicAdd(&s_ngroups, 1u);
                    s_gstart[g] = (uint16_t)a; s_glen[g] = (uint16_t)((b - a) | 0x8000);
                }
            }
        }
        __syncthreads();
    }

    // ---- phase 2: one warp orders one group (or one batch of small groups)
    const uint32_t nbig = s_ngroups, ngroups = nbig + s_nsmall;
    while (true) {
        uint32_t g = 0;
        if (lane == 0) g = atomicAdd(&s_next, 1u);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= ngroups) break;
        if (g >= nbig) g = (uint32_t)GT_GMAX - 1u - (g - nbig);
        const uint32_t gs = s_gstart[g], len = s_glen[g] & 0x7FFFu;
        if (s_glen[g] & 0x8000u) {
            if (len <= 32) order_group_regs<1, true, P>(gs, len, lane, depth, A.T3, s_pos, s_len, s_best, s_sorted, s_lcpout, s_refw, s_gs);
            else order_group_regs<2, true, P>(gs, len, lane, depth, A.T3, s_pos, s_len, s_best, s_sorted, s_lcpout, s_refw, s_gs);
        }
        else if (len <= 32) order_group_regs<1, false, P>(gs, len, lane, depth, A.T3, s_pos, s_len, s_best, s_sorted, s_lcpout, s_refw, s_gs);
        else if (len <= 64) order_group_regs<2, false, P>(gs, len, lane, depth, A.T3, s_pos, s_len, s_best, s_sorted, s_lcpout, s_refw, s_gs);
        else order_group_smem<P>(gs, len, lane, depth, A.T3, s_pos, s_len, s_best, s_sorted, s_lcpout, s_key, s_runa, s_rdepth);
        __syncwarp();
    }
    __syncthreads();

    // ---- phase 3: finished rows leave the SM in slot order
    for (int r = tid; r < GT_S; r += GT_THREADS) {
        const uint32_t gs = s_gs[r];
        if (gs == 0xFFFFu) continue;
        const uint32_t j = t0 + r;
        const uint32_t e = s_sorted[r];
        const uint32_t slot = slot_of(j);
        if (A.SA) A.SA[slot] = s_pos[e];
        if (A.bwt) {
            const uint32_t pv = s_prev[e];
            const uint32_t c = pv & 7u;
            __stcs(&A.bwt[slot], c ? (uint8_t)(A.code2byte >> (8 * c)) : A.eos);
            __stcs(&A.qs[slot], (uint8_t)(pv >> 8));
            if ((uint32_t)r != gs) __stcs(&A.lcp[slot], (uint8_t)(depth + s_lcpout[r]));
        }
    }
}

static size_t group_tile_smem(size_t pos_bytes)
{
    return (size_t)(GT_S + 4) * 8 + (size_t)GT_S * (4 + pos_bytes) + (size_t)GT_HW * 4 + (size_t)GT_S * 2 * 4 + (size_t)GT_GMAX * 2 * 2 +
           (size_t)GT_S * 3 + (size_t)(GT_THREADS / 32) * GT_RW * 8 + 8 + 16;
}

// ============================================================================ deferred groups
// key of `s` symbols starting at text position p, cut at the terminator, right-aligned
__device__ __forceinline__ uint64_t cut_key(const uint64_t *__restrict__ T3, uint64_t p, int s)
{
    uint64_t w = text_window(T3, p);
    const int f = leading_nonzero_fields(w);
    if (f < 21) w &= f ? (~0ull << (64 - 3 * f)) : 0ull;
    return w >> (64 - 3 * s);
}

// one warp per deferred group: its length (distance to the next head) and the SA slot of its head
template <typename K>
__global__ void __launch_bounds__(256)
defer_len_kernel(const K *__restrict__ keys, K kmask, uint32_t m, int level0, int kshift,
                 const uint32_t *__restrict__ g_slot, const uint32_t *__restrict__ g_off,
                 const uint32_t *__restrict__ dstart, uint32_t G, uint32_t *__restrict__ dlen,
                 uint32_t *__restrict__ dslot)
{
    const int lane = threadIdx.x & 31;
    const uint32_t d = (blockIdx.x * 256u + threadIdx.x) >> 5;
    if (d >= G) return;
    const uint32_t start = dstart[d];
    uint32_t end = m;
    for (uint64_t base = (uint64_t)start + 1; base < m; base += 32) {
        const uint64_t j = base + lane;
        bool h = false;
        if (j < m) { const K kc = keys[j] & kmask, kp = keys[j - 1] & kmask; h = (kc != kp) || ((kc & (K)7) == 0); }
        const uint32_t b = __ballot_sync(0xffffffffu, h);
        if (b) { end = (uint32_t)base + __ffs(b) - 1; break; }
    }
    if (lane == 0) {
        dlen[d] = end - start;
        uint32_t slot = start;
        if (!level0) { const uint32_t g = (uint32_t)((uint64_t)keys[start] >> kshift); slot = g_slot[g] + (start - g_off[g]); }
        dslot[d] = slot;
    }
}

// single block: exclusive scan of n counters, total to *total
__global__ void __launch_bounds__(1024)
scan_excl_u32_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, uint32_t n, uint32_t *__restrict__ total)
{
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n; base += 1024) {
        const uint32_t i = base + tid;
        const uint32_t v = i < n ? in[i] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) s_warp[w] = x;
        __syncthreads();
        uint32_t add = 0;
        for (int k = 0; k < w; k++) add += s_warp[k];
        const uint32_t carry = s_carry;
        if (i < n) out[i] = carry + add + x - v;
        __syncthreads();
        if (tid == 1023) s_carry = carry + add + x;
        __syncthreads();
    }
    if (tid == 0) *total = s_carry;
}

// next-level list: members of deferred group d get key (d << 3s) | next s symbols
template <typename P>
__global__ void __launch_bounds__(256)
defer_fill_kernel(const P *__restrict__ pos, const uint32_t *__restrict__ dstart,
                  const uint32_t *__restrict__ dlen, const uint32_t *__restrict__ doff, uint32_t G,
                  const uint64_t *__restrict__ T3, int depth, int s, uint64_t *__restrict__ keys_out,
                  P *__restrict__ pos_out)
{
    for (uint32_t d = blockIdx.x; d < G; d += gridDim.x) {
        const uint32_t start = dstart[d], len = dlen[d], off = doff[d];
        for (uint32_t i = threadIdx.x; i < len; i += 256) {
            const P p = pos[start + i];
            keys_out[off + i] = ((uint64_t)d << (3 * s)) | cut_key(T3, (uint64_t)p + depth, s);
            pos_out[off + i] = p;
        }
    }
}

// ============================================================================ host driver
static int ceil_log2(uint64_t x) { int b = 0; while (((uint64_t)1 << b) < x) b++; return b; }

Alphabet alphabet_from_hist(const uint64_t hist[256], uint8_t eos, bool eos_may_appear)
{
    Alphabet A;
    memset(A.byte2code, 0xFF, sizeof A.byte2code);
    A.code2byte[0] = eos;
    A.byte2code[eos] = 0;
    for (int b = 0; b < 256; b++) {
        if (!hist[b] || b == eos) continue;
        if (b < eos) BFQ_FAIL(BFQ_EALPHABET, "symbol byte %d sorts below the end-marker byte %d", b, (int)eos);
        if (A.nsym == 5) BFQ_FAIL(BFQ_EALPHABET, "more than 5 distinct symbols besides the end-marker (DNA alphabets only)");
        A.nsym++;
        A.code2byte[A.nsym] = (uint8_t)b;
        A.byte2code[b] = (uint8_t)A.nsym;
        if (b == 'N') A.code_N = A.nsym;
        if (b == 'A' || b == 'C' || b == 'G' || b == 'T') A.acgt_mask |= (uint8_t)(1u << A.nsym);
    }
    if (!eos_may_appear && hist[eos]) BFQ_FAIL(BFQ_EALPHABET, "the end-marker byte %d occurs inside the reads", (int)eos);
    return A;
}

template <typename K, typename P>
static void launch_group_tile(bfq_ctx *c, const K *keys, const GroupTileArgsT<P> &A)
{
    const size_t smem = group_tile_smem(sizeof(P));
    BFQ_CUDA(cudaFuncSetAttribute(group_tile_kernel<K, P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const uint32_t ntiles = div_up(A.m, GT_T);
    bfq_ctx::HostSink *sink = c->sink;
    // host-buffer call: the finished rows leave for the host slice by slice while the next slice is ordered
    const uint32_t nslices = (sink && A.level0 && A.bwt && (sink->bwt || sink->qs || sink->lcp) && ntiles >= 4096) ? (uint32_t)c->opt_host_slices : 1;
    const uint32_t per = div_up(ntiles, nslices);
    for (uint32_t s = 0; s < nslices; s++) {
        GroupTileArgsT<P> B = A;
        B.tile0 = s * per;
        const uint32_t cnt = std::min(per, ntiles - B.tile0);
        {
            // DRAM bytes per row: key + position in, the three file bytes out, the suffix array when it is kept.  The packed text
            // the members are compared through is served by L2 (pinned) and is NOT counted: this kernel is issue-bound.
            ProfScope ps(c, PC_GROUP, (uint64_t)cnt * GT_T * (sizeof(K) + sizeof(P) + 3 + (A.SA ? sizeof(P) : 0)));
            group_tile_kernel<K, P><<<cnt, GT_THREADS, smem, c->stream>>>(keys, B);
            BFQ_CUDA(cudaGetLastError());
        }
        if (nslices > 1) {
            // rows below the slice end are final unless a group was deferred (then everything is copied again later)
            const uint64_t lo = (uint64_t)B.tile0 * GT_T, hi = std::min<uint64_t>(A.m, (uint64_t)(B.tile0 + cnt) * GT_T);
            BFQ_CUDA(cudaEventRecord(c->ev_a, c->stream));
            BFQ_CUDA(cudaStreamWaitEvent(c->copy_stream, c->ev_a, 0));
            if (sink->bwt) BFQ_CUDA(cudaMemcpyAsync(sink->bwt + lo, A.bwt + lo, hi - lo, cudaMemcpyDeviceToHost, c->copy_stream));
            if (sink->qs) BFQ_CUDA(cudaMemcpyAsync(sink->qs + lo, A.qs + lo, hi - lo, cudaMemcpyDeviceToHost, c->copy_stream));
            if (sink->lcp) BFQ_CUDA(cudaMemcpyAsync(sink->lcp + lo, A.lcp + lo, hi - lo, cudaMemcpyDeviceToHost, c->copy_stream));
            sink->files_copied = true;
        }
    }
}

size_t packed_text_words(uint64_t N) { return ((N + 320 + 63) / 64) * 3 + 2; }

// P = type of a text position: u32, or u64 for a collection of 2^32 rows and more (only as a key range: the rows of ONE
// GPU are always indexed with 32 bits)
template <typename P>
static void build_ebwt_impl(bfq_ctx *c, const uint8_t *d_bases, const uint8_t *d_quals, const uint64_t *d_off,
                            uint64_t nreads, uint64_t total_bases, uint64_t max_len, uint8_t eos, uint8_t *d_bwt,
                            uint8_t *d_qs, uint8_t *d_lcp, P *d_sa_out, uint64_t *d_t3_keep, KeyRange *range)
{
    const uint64_t N = total_bases + nreads;
    if (N == 0) return;
    if (sizeof(P) == 4 && N >= ((uint64_t)1 << 32) - 64)
        BFQ_FAIL(BFQ_ETOOLARGE, "collection has %llu rows; this build indexes rows with 32 bits", (unsigned long long)N);
    if (sizeof(P) == 8 && !range) BFQ_FAIL(BFQ_ETOOLARGE, "64-bit text positions are only built for key ranges (single-collection mode)");
    if (max_len > 255)
        BFQ_FAIL(BFQ_EINVAL, "a read has %llu symbols; reads are limited to 255 (1-byte LCP entries, dataTypelenSeq is uchar, parameters.h:68)",
                 (unsigned long long)max_len);
    cudaStream_t st = c->stream;

    // ---- alphabet from the data (codes follow byte order, like gsufsort's +1 remap)
    uint64_t hist[256];
    if (d_quals && d_quals != d_bases) { byte_histogram2(c, d_bases, d_quals, total_bases, hist, c->qual_hist); c->has_qual_hist = true; }
    else { byte_histogram(c, d_bases, total_bases, hist); c->has_qual_hist = false; }
    c->trace("part A: byte histograms", st);
    memcpy(c->base_hist, hist, sizeof hist);
    Alphabet A = alphabet_from_hist(hist, eos, /*eos_may_appear=*/false);
    // gsufsort drops bytes <= 1 and 255 (utils.c:229-236); refuse them instead of silently differing
    if (hist[0] || hist[1] || hist[255]) BFQ_FAIL(BFQ_EALPHABET, "bytes 0, 1 or 255 inside reads are not supported");
    Lut256 lut;
    for (int b = 0; b < 256; b++) lut.v[b] = A.byte2code[b] == 0xFF ? 0 : A.byte2code[b];
    uint64_t c2b = 0;
    for (int k = 0; k < 8; k++) c2b |= (uint64_t)A.code2byte[k] << (8 * k);

    // ---- buffers
    const uint64_t ngroups = (N + 320 + 63) / 64;           // padded so windows past the end read zeros
    uint64_t *T3 = d_t3_keep ? d_t3_keep : arena_alloc<uint64_t>(c, ngroups * 3 + 2);
    uint16_t *QL = arena_alloc<uint16_t>(c, N + 16);
    // a key range sorts about N / world suffixes; keyB also hosts the byte text (N + 1 bytes) while packing
    const uint64_t cap = range ? std::max<uint64_t>(range->capacity, N / 8 + 64) : N;
    uint64_t *keyA = arena_alloc<uint64_t>(c, cap), *keyB = arena_alloc<uint64_t>(c, cap);
    P *posA = arena_alloc<P>(c, cap), *posB = arena_alloc<P>(c, cap);
    const uint32_t dcap = (uint32_t)(N / std::min<uint64_t>(GT_E, (uint64_t)c->opt_group_cap + 1) + 64);
    uint32_t *dtab[2][4];                                   // per level parity: start, len, slot, off
    for (int a = 0; a < 2; a++)
        for (int b = 0; b < 4; b++) dtab[a][b] = arena_alloc<uint32_t>(c, dcap + 1);
    uint32_t *dcount = arena_alloc<uint32_t>(c, 4);

    {   // ---- pack: T8/Q, then 3-bit stream (T8 lives in keyB until the first sort needs it)
        uint8_t *T8 = reinterpret_cast<uint8_t *>(keyB);
        ProfScope ps(c, PC_PACK, 2 * total_bases + 2 * N + N + ngroups * 24, 2);
        BFQ_CUDA(cudaMemsetAsync(T3 + ngroups * 3, 0, 16, st));
        int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 16, div_up(nreads * 32, 256));
        pack_reads_kernel<<<std::max(grid, 1), 256, 0, st>>>(d_bases, d_quals, d_off, nreads, lut, eos, T8, QL);
        pack3_kernel<<<div_up(ngroups, 256), 256, 0, st>>>(T8, N, ngroups, T3);
        BFQ_CUDA(cudaGetLastError());
    }

    c->trace("text packed", st);
    // ---- level 0: global sort on k symbols, groups finished in shared memory
    int k = c->opt_initial_k;
    if (k <= 0) k = 16;
    k = std::min(k, 21);
    int flip;
    uint64_t n_sort = N, prev_key = 0;
    bool has_prev = false;
    if (range) {
        flip = radix_sort_suffix_keys_range<P>(c, T3, k <= 16 ? QL : nullptr, N, k, range->rank, range->world, keyA, keyB, posA, posB,
                                               range->capacity, &range->n_local, &range->row_base, &prev_key, &has_prev);
        n_sort = range->n_local;
        if (n_sort == 0) return;
    } else if constexpr (sizeof(P) == 4)
        flip = radix_sort_suffix_keys(c, T3, k <= 16 ? QL : nullptr, N, k, keyA, keyB, posA, posB);
    else
        flip = 0;
    uint64_t *keys = flip ? keyB : keyA, *keys_alt = flip ? keyA : keyB;
    P *pos = flip ? posB : posA, *pos_alt = flip ? posA : posB;

    c->trace("suffix keys sorted", st);
    // The packed text is what every group gathers from at random: pin it in L2 (persisting access
    // window, set aside only for the group kernels) while the sorted list streams through; T3 of config 2 is 38 MB of the 126 MB L2.
    L2Pin pin_text(c, T3, (ngroups * 3 + 2) * 8);

    auto words_for = [&](int depth) { return (int)std::max<int64_t>(0, ((int64_t)max_len - depth + 20) / 21); };
    GroupTileArgsT<P> G;
    memset(&G, 0, sizeof G);
    G.T3 = T3; G.QL = QL; G.SA = d_sa_out; G.bwt = d_bwt; G.qs = d_qs; G.lcp = d_lcp;
    G.code2byte = c2b; G.eos = eos; G.defer_count = dcount; G.defer_cap = dcap;
    G.group_cap = (uint32_t)c->opt_group_cap;
    G.batch_min_groups = (uint32_t)c->opt_batch_min_groups;
    G.bulk_tile = c->opt_bulk_tile;
    G.prev_key = prev_key; G.has_prev = has_prev;

    int parity = 0, depth = 0, ksym = k, level = 0;
    uint32_t m = (uint32_t)n_sort;
    while (true) {
        BFQ_CUDA(cudaMemsetAsync(dcount, 0, 16, st));
        G.pos = pos; G.m = m; G.level0 = level == 0; G.ksym = ksym; G.depth0 = depth;
        G.W = words_for(depth + ksym);
        G.kmask = level == 0 ? ((1ull << (3 * ksym)) - 1) : ~0ull;
        G.embed = level == 0 && ksym <= 16;
        G.g_slot = level ? dtab[parity ^ 1][2] : nullptr;
        G.g_off = level ? dtab[parity ^ 1][3] : nullptr;
        G.defer_start = dtab[parity][0];
        launch_group_tile<uint64_t, P>(c, keys, G);
        fetch_small(c, c->h_small, dcount, 4);
        BFQ_CUDA(cudaStreamSynchronize(st));
        c->trace("part A: groups of a level ordered", st);
        const uint32_t ndef = ((uint32_t *)c->h_small)[0];
        if (ndef == 0) break;
        if (c->sink) c->sink->files_copied = false;      // deferred groups rewrite rows that already left
        if (ndef > dcap) BFQ_FAIL(BFQ_ECUDA, "internal error: %u deferred groups exceed the table of %u", ndef, dcap);
        if (++level > 64) BFQ_FAIL(BFQ_ECUDA, "suffix sort did not converge (internal error)");
        // ---- next level: the members of the deferred groups, keyed by (group, next s symbols)
        uint32_t *dstart = dtab[parity][0], *dlen = dtab[parity][1], *dslot = dtab[parity][2], *doff = dtab[parity][3];
        {
            ProfScope ps(c, PC_DEFER, 0, 2);
            defer_len_kernel<uint64_t><<<div_up((uint64_t)ndef * 32, 256), 256, 0, st>>>(
                keys, G.kmask, m, G.level0, 3 * ksym, G.g_slot, G.g_off, dstart, ndef, dlen, dslot);
            scan_excl_u32_kernel<<<1, 1024, 0, st>>>(dlen, doff, ndef, dcount + 1);
            BFQ_CUDA(cudaGetLastError());
        }
        fetch_small(c, c->h_small, dcount + 1, 4);
        BFQ_CUDA(cudaStreamSynchronize(st));
        const uint32_t m2 = ((uint32_t *)c->h_small)[0];
        depth += ksym;
        const int gbits = std::max(1, ceil_log2(ndef));
        const int s = std::min(21, (64 - gbits) / 3);
        {
            ProfScope ps(c, PC_DEFER, (uint64_t)m2 * (4 + 12 + 16));
            int grid = (int)std::min<uint64_t>(ndef, (uint64_t)c->num_sms * 8);
            defer_fill_kernel<P><<<grid, 256, 0, st>>>(pos, dstart, dlen, doff, ndef, T3, depth, s, keys_alt, pos_alt);
            BFQ_CUDA(cudaGetLastError());
        }
        flip = radix_sort_pairs64(c, keys_alt, keys, pos_alt, pos, m2, gbits + 3 * s);     // overloaded on the payload type
        if (!flip) { std::swap(keys, keys_alt); std::swap(pos, pos_alt); }
        m = m2;
        ksym = s;
        parity ^= 1;
    }
}

void build_ebwt_dev(bfq_ctx *c, const uint8_t *d_bases, const uint8_t *d_quals, const uint64_t *d_off,
                    uint64_t nreads, uint64_t total_bases, uint64_t max_len, uint8_t eos, uint8_t *d_bwt,
                    uint8_t *d_qs, uint8_t *d_lcp, uint32_t *d_sa_out, uint64_t *d_t3_keep, KeyRange *range)
{
    build_ebwt_impl<uint32_t>(c, d_bases, d_quals, d_off, nreads, total_bases, max_len, eos, d_bwt, d_qs, d_lcp, d_sa_out, d_t3_keep, range);
}

void build_ebwt_dev_wide(bfq_ctx *c, const uint8_t *d_bases, const uint8_t *d_quals, const uint64_t *d_off,
                         uint64_t nreads, uint64_t total_bases, uint64_t max_len, uint8_t eos, uint8_t *d_bwt,
                         uint8_t *d_qs, uint8_t *d_lcp, uint64_t *d_sa_out, uint64_t *d_t3_keep, KeyRange *range)
{
    build_ebwt_impl<uint64_t>(c, d_bases, d_quals, d_off, nreads, total_bases, max_len, eos, d_bwt, d_qs, d_lcp, d_sa_out, d_t3_keep, range);
}

} // namespace bfq
