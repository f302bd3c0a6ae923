// sufsort.cu -- part A on the device: generalized suffix sort of a read collection by prefix
// doubling, then the eBWT / permuted-quality gather and the LCP array.
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
// position order gives that for free.  Unresolved groups never contain a terminator in
// their first h symbols, hence ISA[p+h] is always a position of the same read.
//
// Layouts in HBM (N = sum(len+1) rows, N < 2^32):
//   T3   3-bit codes, big-endian bit stream, 64 symbols per three u64 words (fits L2 for cfg2)
//   Q    N bytes, quality of position p (EOS byte at terminators)
//   SA   u32[N] final suffix array (text positions); ISA u32[N] group-head rank of each position
//   active list (sorted by group): pos u32, slot u32 (global SA index of the list slot), gid u32
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

void byte_histogram(bfq_ctx *c, const uint8_t *d_data, uint64_t n, uint64_t hist_out[256])
{
    unsigned long long *d_hist = (unsigned long long *)c->d_small;   // 256*8 = 2 KiB of the 4 KiB
    BFQ_CUDA(cudaMemsetAsync(d_hist, 0, 256 * 8, c->stream));
    if (n) {
        ProfScope ps(c, PC_HIST, n);
        int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 8, div_up(n, 256 * 16) + 1);
        hist256_kernel<<<grid, 256, 0, c->stream>>>(d_data, n, d_hist);
    }
    BFQ_CUDA(cudaMemcpyAsync(c->h_small, d_hist, 256 * 8, cudaMemcpyDeviceToHost, c->stream));
    BFQ_CUDA(cudaStreamSynchronize(c->stream));
    for (int i = 0; i < 256; i++) hist_out[i] = c->h_small[i];
}

// ============================================================================ text packing
struct Lut256 { uint8_t v[256]; };

// one warp per read: T8[p] = code, Q[p] = quality; terminator code 0 / EOS byte after the read
__global__ void __launch_bounds__(256)
pack_reads_kernel(const uint8_t *__restrict__ bases, const uint8_t *__restrict__ quals,
                  const uint64_t *__restrict__ off, uint64_t nreads, Lut256 lut, uint8_t eos,
                  uint8_t *__restrict__ T8, uint8_t *__restrict__ Q)
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
            Q[dst + j] = quals[s + j];
        }
        if (lane == 0) { T8[dst + len] = 0; Q[dst + len] = eos; }
    }
}

// 64 symbols -> three u64 words, first symbol in the most significant bits
__global__ void __launch_bounds__(256)
pack3_kernel(const uint8_t *__restrict__ T8, uint64_t n, uint64_t ngroups, uint64_t *__restrict__ T3)
{
    const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= ngroups) return;
    uint64_t w[4] = {0, 0, 0, 0};
    const uint64_t base = g * 64;
#pragma unroll 8
    for (int j = 0; j < 64; j++) {
        uint64_t code = (base + j < n) ? (uint64_t)T8[base + j] : 0ull;
        int bit = 3 * j, wi = bit >> 6, sh = 61 - (bit & 63);
        if (sh >= 0) w[wi] |= code << sh;
        else { w[wi] |= code >> (-sh); w[wi + 1] |= code << (64 + sh); }
    }
    T3[g * 3 + 0] = w[0];
    T3[g * 3 + 1] = w[1];
    T3[g * 3 + 2] = w[2];
}

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

// number of leading non-terminator symbols in a window (0..21)
__device__ __forceinline__ int leading_nonzero_fields(uint64_t w)
{
    const uint64_t M4 = 0x9249249249249248ull;   // top bit of each of the 21 fields (bits 63,60,...,3)
    const uint64_t M3 = 0x6DB6DB6DB6DB6DB6ull;   // low two bits of each field
    uint64_t t = ((w & M3) + M3) | w;            // field top bit set <=> field != 0
    uint64_t zeros = ~t & M4;
    int lz = zeros ? __clzll((long long)zeros) : 63;
    return lz / 3;
}

__device__ __forceinline__ uint32_t sym_code(const uint64_t *__restrict__ T3, uint64_t p)
{
    return (uint32_t)(text_window(T3, p) >> 61);
}

// ============================================================================ initial keys
template <typename K>
__global__ void __launch_bounds__(256)
init_keys_kernel(const uint64_t *__restrict__ T3, uint32_t n, int k, K *__restrict__ keys,
                 uint32_t *__restrict__ vals)
{
    const uint32_t p = blockIdx.x * 256u + threadIdx.x;
    if (p >= n) return;
    uint64_t w = text_window(T3, p);
    int f = leading_nonzero_fields(w);          // symbols before the first terminator
    if (f < 21) w &= f ? (~0ull << (64 - 3 * f)) : 0ull;
    keys[p] = (K)(w >> (64 - 3 * k));
    vals[p] = p;
}

// ============================================================================ regroup
// After a sort: find the new groups, rank them, retire singletons, compact the rest.
// One chained scan (three look-back words per tile: max, kept count, kept-head count).
constexpr int RG_THREADS = 256;
constexpr int RG_ITEMS = 4;
constexpr int RG_TILE = RG_THREADS * RG_ITEMS;

struct RgState { uint32_t mx, keep, heads; };

__device__ __forceinline__ RgState rg_combine(RgState a, RgState b)
{
    return RgState{max(a.mx, b.mx), a.keep + b.keep, a.heads + b.heads};
}

template <typename K, bool INITIAL>
__global__ void __launch_bounds__(RG_THREADS)
regroup_kernel(const K *__restrict__ keys, const uint32_t *__restrict__ pos,
               const uint32_t *__restrict__ slot_in, uint32_t m,
               uint32_t *__restrict__ ISA, uint32_t *__restrict__ SA,
               uint32_t *__restrict__ pos_out, uint32_t *__restrict__ slot_out,
               uint32_t *__restrict__ gid_out,
               unsigned long long *__restrict__ status /*[tiles][3]*/,
               uint32_t *__restrict__ tile_counter, uint32_t *__restrict__ totals /*[2]: kept, groups*/)
{
    __shared__ uint32_t s_tile;
    __shared__ RgState s_warp[RG_THREADS / 32];
    __shared__ RgState s_prefix;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t j0 = tile * RG_TILE + tid * RG_ITEMS;

    K key[RG_ITEMS + 2];            // key[0] = predecessor, key[RG_ITEMS+1] = successor
    uint32_t ps[RG_ITEMS], sl[RG_ITEMS];
    bool head[RG_ITEMS + 1];
#pragma unroll
    for (int i = 0; i < RG_ITEMS + 2; i++) {
        int64_t j = (int64_t)j0 + i - 1;
        key[i] = (j >= 0 && j < (int64_t)m) ? keys[j] : (K)0;
    }
#pragma unroll
    for (int i = 0; i < RG_ITEMS; i++) {
        uint32_t j = j0 + i;
        ps[i] = j < m ? pos[j] : 0u;
        sl[i] = INITIAL ? j : (j < m ? slot_in[j] : 0u);
    }
#pragma unroll
    for (int i = 0; i <= RG_ITEMS; i++) {
        uint32_t j = j0 + i;
        bool h = (j == 0) || (j >= m) || (key[i + 1] != key[i]);
        if (INITIAL) h = h || ((key[i + 1] & 7) == 0);     // key holds a terminator: unique
        head[i] = h;
    }
    // local aggregate
    RgState loc{0, 0, 0};
    bool keep[RG_ITEMS];
#pragma unroll
    for (int i = 0; i < RG_ITEMS; i++) {
        uint32_t j = j0 + i;
        bool valid = j < m;
        keep[i] = valid && !(head[i] && head[i + 1]);
        if (valid && head[i]) loc.mx = max(loc.mx, sl[i]);
        loc.keep += keep[i] ? 1u : 0u;
        loc.heads += (keep[i] && head[i]) ? 1u : 0u;
    }
    // block exclusive scan of RgState
    RgState inc = loc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        RgState y{__shfl_up_sync(0xffffffffu, inc.mx, o), __shfl_up_sync(0xffffffffu, inc.keep, o),
                  __shfl_up_sync(0xffffffffu, inc.heads, o)};
        if (lane >= o) inc = rg_combine(y, inc);
    }
    if (lane == 31) s_warp[w] = inc;
    __syncthreads();
    RgState wpre{0, 0, 0};
    for (int i = 0; i < w; i++) wpre = rg_combine(wpre, s_warp[i]);
    // exclusive for this thread = wpre + (inc of previous lane)
    RgState prev{__shfl_up_sync(0xffffffffu, inc.mx, 1), __shfl_up_sync(0xffffffffu, inc.keep, 1),
                 __shfl_up_sync(0xffffffffu, inc.heads, 1)};
    RgState excl = (lane == 0) ? wpre : rg_combine(wpre, prev);

    // tile aggregate + chained scan (lanes 0..2 of warp 0 each own one word)
    if (w == 0) {
        RgState agg{0, 0, 0};
        for (int i = 0; i < RG_THREADS / 32; i++) agg = rg_combine(agg, s_warp[i]);
        if (lane < 3) {
            const unsigned long long FL = 62;
            unsigned long long mine = lane == 0 ? agg.mx : (lane == 1 ? agg.keep : agg.heads);
            volatile unsigned long long *st = status;
            st[(size_t)tile * 3 + lane] = ((unsigned long long)(tile == 0 ? 2 : 1) << FL) | mine;
            unsigned long long ex = 0;
            if (tile > 0) {
                int64_t t = (int64_t)tile - 1;
                while (true) {
                    unsigned long long v = st[(size_t)t * 3 + lane];
                    unsigned f = (unsigned)(v >> FL);
                    if (f == 0) continue;
                    unsigned long long val = v & ((1ull << FL) - 1);
                    ex = (lane == 0) ? (ex > val ? ex : val) : ex + val;
                    if (f == 2) break;
                    t--;
                }
                unsigned long long incl = (lane == 0) ? (ex > mine ? ex : mine) : ex + mine;
                st[(size_t)tile * 3 + lane] = (2ull << FL) | incl;
            }
            if (lane == 0) s_prefix.mx = (uint32_t)ex;
            if (lane == 1) s_prefix.keep = (uint32_t)ex;
            if (lane == 2) s_prefix.heads = (uint32_t)ex;
            // the last tile publishes the totals
            if ((uint64_t)(tile + 1) * RG_TILE >= m) {
                if (lane == 1) totals[0] = (uint32_t)(ex + mine);
                if (lane == 2) totals[1] = (uint32_t)(ex + mine);
            }
        }
    }
    __syncthreads();
    RgState run = rg_combine(s_prefix, excl);
#pragma unroll
    for (int i = 0; i < RG_ITEMS; i++) {
        uint32_t j = j0 + i;
        if (j >= m) break;
        if (head[i]) run.mx = max(run.mx, sl[i]);
        // group-head rank of this suffix (global SA index of the first member of its group)
        ISA[ps[i]] = run.mx;
        if (head[i] && head[i + 1]) {
            SA[sl[i]] = ps[i];                    // singleton: final
        } else {
            if (head[i]) run.heads++;
            pos_out[run.keep] = ps[i];
            slot_out[run.keep] = sl[i];
            gid_out[run.keep] = run.heads - 1;
            run.keep++;
        }
    }
}

// key = (compact group id << nbits) | rank of the suffix h symbols further on
__global__ void __launch_bounds__(256)
doubling_keys_kernel(const uint32_t *__restrict__ pos, const uint32_t *__restrict__ gid,
                     const uint32_t *__restrict__ ISA, uint32_t m, uint32_t h, int nbits,
                     uint64_t *__restrict__ keys)
{
    const uint32_t j = blockIdx.x * 256u + threadIdx.x;
    if (j >= m) return;
    const uint32_t p = pos[j];
    keys[j] = ((uint64_t)gid[j] << nbits) | (uint64_t)ISA[p + h];
}

// ============================================================================ final gather
// row i: p = SA[i]; bwt/qs from position p-1 (EOS at a read start); lcp by comparing the
// packed text of SA[i-1] and SA[i] 21 symbols at a time, stopping at either terminator.
__global__ void __launch_bounds__(256)
gather_kernel(const uint32_t *__restrict__ SA, const uint64_t *__restrict__ T3,
              const uint8_t *__restrict__ Q, uint32_t n, uint8_t eos, uint64_t code2byte_packed,
              uint8_t *__restrict__ bwt, uint8_t *__restrict__ qs, uint8_t *__restrict__ lcp)
{
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i >= n) return;
    const uint32_t p = SA[i];
    uint8_t b = eos, q = eos;
    if (p > 0) {
        uint32_t c = sym_code(T3, (uint64_t)p - 1);
        if (c) { b = (uint8_t)(code2byte_packed >> (8 * c)); q = Q[p - 1]; }
    }
    bwt[i] = b;
    qs[i] = q;
    uint32_t l = 0;
    if (i > 0) {
        uint64_t a = SA[i - 1], bb = p;
        while (true) {
            uint64_t wa = text_window(T3, a + l), wb = text_window(T3, bb + l);
            int za = leading_nonzero_fields(wa), zb = leading_nonzero_fields(wb);
            uint64_t x = wa ^ wb;
            int fm = x ? (__clzll((long long)x) / 3) : 21;
            int step = min(fm, min(za, zb));
            l += step;
            if (step < 21) break;
        }
    }
    lcp[i] = (uint8_t)l;      // low byte, as main.c:891-895 writes it
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

void build_ebwt_dev(bfq_ctx *c, const uint8_t *d_bases, const uint8_t *d_quals, const uint64_t *d_off,
                    uint64_t nreads, uint64_t total_bases, uint8_t eos, uint8_t *d_bwt, uint8_t *d_qs,
                    uint8_t *d_lcp, uint32_t *d_sa_out)
{
    const uint64_t N = total_bases + nreads;
    if (N == 0) return;
    if (N >= ((uint64_t)1 << 32) - 64) BFQ_FAIL(BFQ_ETOOLARGE, "collection has %llu rows; this build indexes rows with 32 bits", (unsigned long long)N);
    cudaStream_t st = c->stream;

    // ---- alphabet from the data (codes follow byte order, like gsufsort's +1 remap)
    uint64_t hist[256];
    byte_histogram(c, d_bases, total_bases, hist);
    Alphabet A = alphabet_from_hist(hist, eos, /*eos_may_appear=*/false);
    // gsufsort drops bytes <= 1 and 255 (utils.c:229-236); refuse them instead of silently differing
    if (hist[0] || hist[1] || hist[255]) BFQ_FAIL(BFQ_EALPHABET, "bytes 0, 1 or 255 inside reads are not supported");
    Lut256 lut;
    for (int b = 0; b < 256; b++) lut.v[b] = A.byte2code[b] == 0xFF ? 0 : A.byte2code[b];
    uint64_t c2b = 0;
    for (int k = 0; k < 8; k++) c2b |= (uint64_t)A.code2byte[k] << (8 * k);

    // ---- buffers
    const uint64_t ngroups = (N + 64 + 63) / 64;            // padded so windows past the end read zeros
    uint64_t *T3 = arena_alloc<uint64_t>(c, ngroups * 3 + 2);
    uint8_t *Q = arena_alloc<uint8_t>(c, N + 16);
    uint32_t *SA = d_sa_out ? d_sa_out : arena_alloc<uint32_t>(c, N);
    uint32_t *ISA = arena_alloc<uint32_t>(c, N + 256);
    uint64_t *keyA = arena_alloc<uint64_t>(c, N), *keyB = arena_alloc<uint64_t>(c, N);
    uint32_t *posA = arena_alloc<uint32_t>(c, N), *posB = arena_alloc<uint32_t>(c, N);
    uint32_t *slotA = arena_alloc<uint32_t>(c, N), *slotB = arena_alloc<uint32_t>(c, N);
    uint32_t *gid = arena_alloc<uint32_t>(c, N);
    const uint32_t rg_tiles_max = div_up(N, RG_TILE);
    unsigned long long *rg_status = arena_alloc<unsigned long long>(c, (size_t)rg_tiles_max * 3 + 4);
    uint32_t *rg_counter = (uint32_t *)(rg_status + (size_t)rg_tiles_max * 3);   // [0]=tile ctr [1..2]=totals

    {   // ---- pack: T8/Q, then 3-bit stream (T8 lives in keyB until the first sort needs it)
        uint8_t *T8 = reinterpret_cast<uint8_t *>(keyB);
        ProfScope ps(c, PC_PACK, 2 * total_bases + 2 * N + N + ngroups * 24, 2);
        BFQ_CUDA(cudaMemsetAsync(T3 + ngroups * 3, 0, 16, st));
        int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 16, div_up(nreads * 32, 256));
        pack_reads_kernel<<<std::max(grid, 1), 256, 0, st>>>(d_bases, d_quals, d_off, nreads, lut, eos, T8, Q);
        pack3_kernel<<<div_up(ngroups, 256), 256, 0, st>>>(T8, N, ngroups, T3);
        BFQ_CUDA(cudaGetLastError());
    }

    // ---- initial sort on k symbols
    int k = c->opt_initial_k;
    if (k <= 0) k = 16;
    k = std::min(k, 21);
    {
        ProfScope ps(c, PC_KEYS, N * (8 + 4) + ngroups * 24);
        init_keys_kernel<uint64_t><<<div_up(N, 256), 256, 0, st>>>(T3, (uint32_t)N, k, keyA, posA);
    }
    int flip = radix_sort_pairs64(c, keyA, keyB, posA, posB, N, 3 * k);
    uint64_t *keys = flip ? keyB : keyA, *keys_alt = flip ? keyA : keyB;
    uint32_t *pos = flip ? posB : posA, *pos_alt = flip ? posA : posB;
    uint32_t *slot = slotA, *slot_alt = slotB;

    uint32_t m = (uint32_t)N;
    uint32_t h = (uint32_t)k;
    const int nbits = ceil_log2(N);
    bool initial = true;
    int round = 0;
    while (m > 0) {
        // regroup the sorted list: ranks, retire singletons, compact the rest
        const uint32_t tiles = div_up(m, RG_TILE);
        BFQ_CUDA(cudaMemsetAsync(rg_status, 0, ((size_t)tiles * 3 + 4) * 8, st));
        uint32_t *ctr = (uint32_t *)(rg_status + (size_t)tiles * 3);
        {
            ProfScope ps(c, PC_REGROUP, (uint64_t)m * (8 + 4 + 4 + 4 + 12));
            if (initial)
                regroup_kernel<uint64_t, true><<<tiles, RG_THREADS, 0, st>>>(
                    keys, pos, nullptr, m, ISA, SA, pos_alt, slot_alt, gid, rg_status, ctr, ctr + 1);
            else
                regroup_kernel<uint64_t, false><<<tiles, RG_THREADS, 0, st>>>(
                    keys, pos, slot, m, ISA, SA, pos_alt, slot_alt, gid, rg_status, ctr, ctr + 1);
        }
        BFQ_CUDA(cudaGetLastError());
        BFQ_CUDA(cudaMemcpyAsync(c->h_small, ctr + 1, 8, cudaMemcpyDeviceToHost, st));
        BFQ_CUDA(cudaStreamSynchronize(st));
        const uint32_t kept = ((uint32_t *)c->h_small)[0], groups = ((uint32_t *)c->h_small)[1];
        std::swap(pos, pos_alt);
        std::swap(slot, slot_alt);
        initial = false;
        m = kept;
        if (m == 0) break;
        if (++round > 40) BFQ_FAIL(BFQ_ECUDA, "suffix sort did not converge (internal error)");
        // next round: sort the survivors by (group, rank of the suffix h further on)
        {
            ProfScope ps(c, PC_KEYS, (uint64_t)m * (4 + 4 + 4 + 8));
            doubling_keys_kernel<<<div_up(m, 256), 256, 0, st>>>(pos, gid, ISA, m, h, nbits, keys);
        }
        const int gbits = ceil_log2(groups);
        flip = radix_sort_pairs64(c, keys, keys_alt, pos, pos_alt, m, nbits + gbits);
        if (flip) { std::swap(keys, keys_alt); std::swap(pos, pos_alt); }
        h *= 2;
    }
    (void)rg_counter;

    // ---- eBWT, permuted qualities, LCP
    if (d_bwt) {
        ProfScope ps(c, PC_GATHER, N * (4 + 3 + 2 + 16));
        gather_kernel<<<div_up(N, 256), 256, 0, st>>>(SA, T3, Q, (uint32_t)N, eos, c2b, d_bwt, d_qs, d_lcp);
        BFQ_CUDA(cudaGetLastError());
    }
}

} // namespace bfq
