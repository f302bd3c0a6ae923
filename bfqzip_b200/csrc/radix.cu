// radix.cu -- stable LSD radix sort of (key, 32-bit payload) pairs, one-sweep style.
//
// This is the hot kernel of part A (SURVEY.md 8(a) row a3 replaces the sequential induced
// sorting of external/gsufsort/external/gsacak.c:1648-1930 with prefix doubling over radix
// passes).  Per pass every key and payload is read once and written once:
//   * rs_hist_kernel    one read of the keys builds the digit histograms of ALL passes
//                       (shared-memory histograms, warp-aggregated with __match_any_sync);
//   * rs_onesweep_kernel per tile: warp-striped coalesced loads, match-based stable ranking,
//                       chained-scan ("decoupled look-back") of the per-tile digit counts so no
//                       second read is needed, shared-memory staging so that every digit's run
//                       leaves the SM as contiguous, coalesced stores.
// HBM-bound integer work: no tensor cores.  Algorithmic bytes per pass = 2*(sizeof(K)+4)*n.
#include "common.cuh"
#include <algorithm>
#include <mutex>

namespace bfq {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_RADIX = 256;
constexpr int RS_MAX_PASSES = 8;
#define RS_BLOCKS_PER_SM(items) ((items) >= 22 ? 3 : (items) >= 16 ? 4 : (items) >= 12 ? 5 : 6)
#ifndef RS_ITEMS_MAIN
#define RS_ITEMS_MAIN 16
#endif

struct RsPasses {
    int num;
    int shift[RS_MAX_PASSES];
    int bits[RS_MAX_PASSES];
};

static RsPasses plan_passes(int bits)
{
    RsPasses p{};
    p.num = (bits + 7) / 8;
    int base = p.num ? bits / p.num : 0, rem = p.num ? bits % p.num : 0, sh = 0;
    for (int i = 0; i < p.num; i++) {
        p.bits[i] = base + (i < rem ? 1 : 0);
        p.shift[i] = sh;
        sh += p.bits[i];
    }
    return p;
}

template <typename K> __device__ __forceinline__ uint32_t digit_of(K k, int shift, uint32_t mask)
{
    return (uint32_t)(k >> shift) & mask;
}

// ------------------------------------------------------------------------------ histogram
template <typename K>
__global__ void __launch_bounds__(RS_THREADS)
rs_hist_kernel(const K *__restrict__ keys, uint64_t n, RsPasses P, uint32_t *__restrict__ hist)
{
    // Two shared copies of the per-pass histograms (even / odd warps) halve same-address collisions;
    // a warp whose 32 digits are all equal (constant high bits) adds once instead of 32 times.
    __shared__ uint32_t sh[2][RS_MAX_PASSES][RS_RADIX];
    for (int i = threadIdx.x; i < 2 * RS_MAX_PASSES * RS_RADIX; i += RS_THREADS) (&sh[0][0][0])[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, copy = (threadIdx.x >> 5) & 1;
    const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
    for (uint64_t base = (uint64_t)blockIdx.x * RS_THREADS; base < n; base += stride) {
        const uint64_t i = base + threadIdx.x;          // block-uniform trip count
        const bool ok = i < n;
        const K k = ok ? keys[i] : (K)0;
#pragma unroll
        for (int p = 0; p < RS_MAX_PASSES; p++) {
            if (p < P.num) {
                const uint32_t d = digit_of(k, P.shift[p], (1u << P.bits[p]) - 1);
                const uint32_t d0 = __shfl_sync(0xffffffffu, d, 0);
                const uint32_t same = __ballot_sync(0xffffffffu, ok && d == d0);
                const uint32_t act = __ballot_sync(0xffffffffu, ok);
                if (same == act) { if (lane == 0 && act) atomicAdd(&sh[copy][p][d0], __popc(act)); }
                else if (ok) atomicAdd(&sh[copy][p][d], 1u);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < P.num * RS_RADIX; i += RS_THREADS) {
        const uint32_t v = (&sh[0][0][0])[i] + (&sh[1][0][0])[i];
        if (v) atomicAdd(&hist[i], v);
    }
}

// level-0 suffix keys + their digit histograms in one pass over the packed text (see common.cuh).
// A thread takes FOUR consecutive text positions per step: their windows lie in three consecutive words of the packed
// text (loaded once), their QL entries are one 8-byte load, and the four keys / positions leave as 16-byte stores.
// SIX8: the plan is six 8-bit digits (k = 16, the default), so the digits are the key's low six bytes -- cut at compile-time
// shifts instead of the variable 64-bit shifts and the per-pass branches of the general form.
template <bool SIX8>
__global__ void __launch_bounds__(RS_THREADS)
rs_suffix_keys_kernel(const uint64_t *__restrict__ T3, const uint16_t *__restrict__ QL, uint32_t n, int k, RsPasses P,
                      uint64_t *__restrict__ keys, uint32_t *__restrict__ vals, uint32_t *__restrict__ hist)
{
    __shared__ uint32_t sh[2][RS_MAX_PASSES][RS_RADIX];
    for (int i = threadIdx.x; i < 2 * RS_MAX_PASSES * RS_RADIX; i += RS_THREADS) (&sh[0][0][0])[i] = 0;
    __syncthreads();
    const int copy = (threadIdx.x >> 5) & 1;
    const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS * 4;
    const bool vec = ((reinterpret_cast<uintptr_t>(keys) | reinterpret_cast<uintptr_t>(vals)) & 15) == 0 &&
                     (!QL || (reinterpret_cast<uintptr_t>(QL) & 7) == 0);
    for (uint64_t p0 = ((uint64_t)blockIdx.x * RS_THREADS + threadIdx.x) * 4; p0 < n; p0 += stride) {
        const uint64_t bit = 3 * p0;
        const uint64_t wi = bit >> 6;
        const int s0 = (int)(bit & 63);
        const uint64_t w0 = T3[wi], w1 = T3[wi + 1], w2 = T3[wi + 2];     // (the text is padded: windows past the end read zeros)
        const bool full = vec && p0 + 4 <= n;
        const int valid = p0 + 4 <= n ? 4 : (int)(n - p0);
        uint32_t ql[4] = {0, 0, 0, 0};
        if (QL) {
            if (full) {
                const uint2 v = *reinterpret_cast<const uint2 *>(QL + p0);
                ql[0] = v.x & 0xFFFFu; ql[1] = v.x >> 16; ql[2] = v.y & 0xFFFFu; ql[3] = v.y >> 16;
            } else {
#pragma unroll
                for (int i = 0; i < 4; i++) if (p0 + i < n) ql[i] = QL[p0 + i];
            }
        }
        // the 128 bits that start at the first position's window; the other three windows are cut from them at fixed shifts
        const uint64_t a0 = s0 ? ((w0 << s0) | (w1 >> (64 - s0))) : w0, a1 = s0 ? ((w1 << s0) | (w2 >> (64 - s0))) : w1;
        uint64_t key[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            uint64_t w = i ? ((a0 << (3 * i)) | (a1 >> (64 - 3 * i))) : a0;
            w = cut_after_terminator(w & ~1ull);              // symbols after the first terminator do not belong to the suffix
            const uint64_t kk = SIX8 ? (w >> 16) : (w >> (64 - 3 * k));
            if (i < valid) {
                if constexpr (SIX8) {
                    const uint32_t lo = (uint32_t)kk, hi = (uint32_t)(kk >> 32);
                    atomicAdd(&sh[copy][0][lo & 255u], 1u);
                    atomicAdd(&sh[copy][1][(lo >> 8) & 255u], 1u);
                    atomicAdd(&sh[copy][2][(lo >> 16) & 255u], 1u);
                    atomicAdd(&sh[copy][3][lo >> 24], 1u);
                    atomicAdd(&sh[copy][4][hi & 255u], 1u);
                    atomicAdd(&sh[copy][5][(hi >> 8) & 255u], 1u);
                } else {
#pragma unroll
                    for (int q = 0; q < RS_MAX_PASSES; q++)
                        if (q < P.num) atomicAdd(&sh[copy][q][digit_of(kk, P.shift[q], (1u << P.bits[q]) - 1)], 1u);
                }
            }
            key[i] = kk | ((uint64_t)ql[i] << 48);            // QL rides in bits the sort ignores (k <= 16); 0 without QL
        }
        if (full) {
            *reinterpret_cast<ulonglong2 *>(keys + p0) = make_ulonglong2(key[0], key[1]);
            *reinterpret_cast<ulonglong2 *>(keys + p0 + 2) = make_ulonglong2(key[2], key[3]);
            *reinterpret_cast<uint4 *>(vals + p0) = make_uint4((uint32_t)p0, (uint32_t)p0 + 1, (uint32_t)p0 + 2, (uint32_t)p0 + 3);
        } else {
#pragma unroll
            for (int i = 0; i < 4; i++) if (p0 + i < n) { keys[p0 + i] = key[i]; vals[p0 + i] = (uint32_t)(p0 + i); }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < P.num * RS_RADIX; i += RS_THREADS) {
        const uint32_t v = (&sh[0][0][0])[i] + (&sh[1][0][0])[i];
        if (v) atomicAdd(&hist[i], v);
    }
}

// exclusive scan of each pass's 256 counters, in place (one block per pass)
__global__ void rs_scan_hist_kernel(uint32_t *hist)
{
    __shared__ uint32_t warp_tot[RS_WARPS];
    uint32_t *h = hist + blockIdx.x * RS_RADIX;
    int t = threadIdx.x, lane = t & 31, w = t >> 5;
    uint32_t v = h[t], x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) warp_tot[w] = x;
    __syncthreads();
    uint32_t add = 0;
    for (int i = 0; i < w; i++) add += warp_tot[i];
    h[t] = x - v + add;
}

// ------------------------------------------------------------------------------ onesweep
// status word: [flag:2][value].  flag 1 = tile aggregate, 2 = inclusive prefix.
template <typename S> struct StatusTraits;
template <> struct StatusTraits<uint32_t> { static constexpr int FS = 30; };
template <> struct StatusTraits<uint64_t> { static constexpr int FS = 62; };

template <typename S> __device__ __forceinline__ S ld_status(const S *p)
{
    return *((const volatile S *)p);
}
template <typename S> __device__ __forceinline__ void st_status(S *p, S v)
{
    *((volatile S *)p) = v;
}

// m &= lanes whose digit has the same bit B.  Four instructions (test with predicate output, vote, conditional
// complement, and); the C form compiles to six or seven per bit.
template <int B> __device__ __forceinline__ void match_bit(uint32_t d, uint32_t &m)
{
    asm volatile("{\n\t.reg .pred p, q;\n\t.reg .b32 t, v;\n\t"
                 "setp.ne.u32 q, 0, 0;\n\t"
                 "lop3.or.b32 t|p, %1, %2, 0, 0xC0, q;\n\t"
                 "vote.sync.ballot.b32 v, p, 0xffffffff;\n\t"
                 "@!p not.b32 v, v;\n\tand.b32 %0, %0, v;\n\t}"
                 : "+r"(m) : "r"(d), "n"(1u << B));
}

// EMIT = true: the first pass of the flat-mode rebuild makes its input on the fly from the smoothed rows -- element i is
// (text position SA[i] - 1, base | quality << 8), see PermSource in common.cuh -- so the pair list is never written
// unsorted and read back (the loads are as many bytes as the pairs would be).
// MB: digit bits the in-warp match looks at (8, or 6 for the passes of the permutation sort, whose digits have at most 6 bits)
template <typename K, typename V, typename S, int ITEMS, bool EMIT = false, int MB = 8>
__global__ void __launch_bounds__(RS_THREADS, RS_BLOCKS_PER_SM(ITEMS))
rs_onesweep_kernel(const K *__restrict__ keys_in, K *__restrict__ keys_out,
                   const V *__restrict__ vals_in, V *__restrict__ vals_out,
                   uint32_t n, int shift, int bits, const uint32_t *__restrict__ digit_base,
                   S *__restrict__ status, uint32_t *__restrict__ tile_counter, PermSource src = PermSource{})
{
    constexpr int TILE = RS_THREADS * ITEMS;
    static_assert(sizeof(V) <= sizeof(K), "the payload is staged in the key buffer");
    constexpr int FS = StatusTraits<S>::FS;
    constexpr S VMASK = (((S)1) << FS) - 1;

    __shared__ uint32_t s_tile;
    __shared__ uint32_t warp_cnt[RS_WARPS][RS_RADIX];
    __shared__ uint32_t digit_start[RS_RADIX];
    __shared__ uint32_t digit_gbase[RS_RADIX];
    __shared__ uint32_t warp_tot[RS_WARPS];
    extern __shared__ __align__(16) unsigned char rs_stage[];     // TILE keys (dynamic: more than 48 KiB for the larger tiles)
    K *stage = reinterpret_cast<K *>(rs_stage);

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t mask = (1u << bits) - 1;
    const uint32_t lt = (1u << lane) - 1;

    if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
    for (int i = tid; i < RS_WARPS * RS_RADIX; i += RS_THREADS) (&warp_cnt[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t tile_base = tile * (uint32_t)TILE;
    const uint32_t valid = min((uint32_t)TILE, n - tile_base);
    // EMIT: the tile's stored-quality and BWT bytes are contiguous -- the block fetches them with 16-byte loads (one per thread
    // and array for a 4096-row tile) into shared memory instead of 2 * ITEMS one-byte loads per thread
    __shared__ __align__(16) uint8_t s_emit[EMIT ? 2 * TILE : 16];
    bool emit_staged = false;
    if constexpr (EMIT) {
        emit_staged = valid == (uint32_t)TILE && ((reinterpret_cast<uintptr_t>(src.qflat) | reinterpret_cast<uintptr_t>(src.bwt)) & 15) == 0;
        if (emit_staged) {
            for (int i = tid; i < TILE / 16; i += RS_THREADS) {
                reinterpret_cast<uint4 *>(s_emit)[i] = __ldcs(reinterpret_cast<const uint4 *>(src.qflat + tile_base) + i);
                reinterpret_cast<uint4 *>(s_emit + TILE)[i] = __ldcs(reinterpret_cast<const uint4 *>(src.bwt + tile_base) + i);
            }
        }
    }

    // ---- load keys (warp-striped: item i of lane l sits at warp_base + i*32 + l)
    K key[ITEMS];
    const uint32_t wbase = w * (ITEMS * 32) + lane;
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        uint32_t loc = wbase + i * 32;
        if constexpr (EMIT) {
            key[i] = (K)~(K)0;
            if (loc < valid) { const uint32_t p = src.SA[tile_base + loc]; key[i] = (K)(p ? p - 1 : src.n_total - 1); }
        } else
            key[i] = (loc < valid) ? keys_in[tile_base + loc] : (K)~(K)0;
    }
    // ---- stable in-warp ranking
    uint16_t rank[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        uint32_t d = digit_of(key[i], shift, mask);
        // lanes with the same digit: one ballot per digit bit (match.any is far slower with ~30 distinct values per warp)
        // (always all 8 digit bits: a pass with fewer bits has zeros above them, which match everywhere)
        uint32_t m = 0xffffffffu;
        match_bit<0>(d, m); match_bit<1>(d, m); match_bit<2>(d, m); match_bit<3>(d, m);
        match_bit<4>(d, m); match_bit<5>(d, m);
        if constexpr (MB > 6) { match_bit<6>(d, m); match_bit<7>(d, m); }
        uint32_t prev = warp_cnt[w][d];
        __syncwarp();
        if ((__ffs(m) - 1) == lane) warp_cnt[w][d] = prev + __popc(m);
        __syncwarp();
        rank[i] = (uint16_t)(prev + __popc(m & lt));
    }
    __syncthreads();

    // ---- per-digit totals over warps (thread d owns digit d), publish tile aggregate
    uint32_t count;
    {
        uint32_t acc = 0;
#pragma unroll
        for (int ww = 0; ww < RS_WARPS; ww++) {
            uint32_t t = warp_cnt[ww][tid];
            warp_cnt[ww][tid] = acc;
            acc += t;
        }
        count = acc;
        // padding keys are all-ones: they were ranked into the top digit; do not count them
        if ((uint32_t)tid == mask) count -= ((uint32_t)TILE - valid);
        // (padding has digit == mask because the key is all ones)
    }
    S *my_status = status + (size_t)tile * RS_RADIX + tid;
    st_status<S>(my_status, ((S)(tile == 0 ? 2 : 1) << FS) | (S)count);

    // ---- exclusive scan of the 256 counts -> local start of each digit inside the tile
    {
        uint32_t x = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) warp_tot[w] = x;
        __syncthreads();
        uint32_t add = 0;
#pragma unroll
        for (int i = 0; i < RS_WARPS; i++) if (i < w) add += warp_tot[i];
        digit_start[tid] = x - count + add;
    }
    __syncthreads();

    // ---- scatter keys into the staging buffer at their tile-local sorted position
    uint16_t lpos[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        uint32_t d = digit_of(key[i], shift, mask);
        uint32_t p = digit_start[d] + warp_cnt[w][d] + rank[i];
        lpos[i] = (uint16_t)p;
        stage[p] = key[i];
    }

    // the payload loads are issued here: they are in flight during the chained scan and the key output
    V val[ITEMS];
    if constexpr (EMIT) {
        // both bytes are loaded unconditionally (independent loads, all 2 * ITEMS of them in flight); only the rare rows whose
        // base was replaced (bit 7) touch obase.  Two loops, so that the staged case carries no global-load alternative.
        if (emit_staged) {                                                   // (a full tile: every loc is valid)
#pragma unroll
            for (int i = 0; i < ITEMS; i++) {
                const uint32_t loc = wbase + i * 32;
                const uint32_t q = s_emit[loc];                              // (staged before the first barrier)
                uint32_t base = s_emit[TILE + loc];
                if (q & 0x80u) base = src.obase[tile_base + loc];
                val[i] = (V)(base | ((q & 0x7Fu) << 8));
            }
        } else {
#pragma unroll
            for (int i = 0; i < ITEMS; i++) {
                const uint32_t loc = wbase + i * 32;
                val[i] = (V)0;
                if (loc < valid) {
                    const uint32_t r = tile_base + loc;
                    const uint8_t q = __ldcs(src.qflat + r);
                    uint8_t base = __ldcs(src.bwt + r);
                    if (q & 0x80) base = src.obase[r];
                    val[i] = (V)(base | ((uint32_t)(q & 0x7F) << 8));
                }
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < ITEMS; i++) {
            const uint32_t loc = wbase + i * 32;
            val[i] = (loc < valid) ? vals_in[tile_base + loc] : (V)0;
        }
    }

    // ---- chained scan: exclusive prefix of this digit over all earlier tiles
    {
        uint32_t excl = 0;
        if (tile > 0) {
            int64_t t = (int64_t)tile - 1;
            while (true) {
                S v = ld_status<S>(status + (size_t)t * RS_RADIX + tid);
                uint32_t f = (uint32_t)(v >> FS);
                if (f == 0) continue;
                excl += (uint32_t)(v & VMASK);
                if (f == 2) break;
                t--;
            }
            st_status<S>(my_status, ((S)2 << FS) | (S)((S)excl + (S)count));
        }
        digit_gbase[tid] = digit_base[tid] + excl - digit_start[tid];
    }
    __syncthreads();

    // ---- coalesced output of keys; the destination of each slot stays in registers for the payload pass
    uint32_t dst[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        uint32_t slot = i * RS_THREADS + tid;
        K k = stage[slot];
        dst[i] = digit_gbase[digit_of(k, shift, mask)] + slot;
        if (slot < valid) keys_out[dst[i]] = k;
    }
    __syncthreads();

    // ---- payload through the same permutation
    V *stage32 = reinterpret_cast<V *>(stage);
#pragma unroll
    for (int i = 0; i < ITEMS; i++) stage32[lpos[i]] = val[i];
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        uint32_t slot = i * RS_THREADS + tid;
        if (slot < valid) vals_out[dst[i]] = stage32[slot];
    }
}

// ------------------------------------------------------------------------------ host driver
// the passes, given the (unscanned) digit histograms of all of them
template <typename K, typename V, int ITEMS>
static int radix_passes(bfq_ctx *c, K *keys_a, K *keys_b, V *vals_a, V *vals_b, uint64_t n,
                        const RsPasses &P, uint32_t *hist, const PermSource *src = nullptr, ProfClass cls = PC_RADIX_PASS)
{
    constexpr int TILE = RS_THREADS * ITEMS;
    const uint32_t tiles = div_up(n, TILE);
    const bool wide = n >= ((uint64_t)1 << 30);
    const size_t status_bytes = (size_t)tiles * RS_RADIX * (wide ? 8 : 4) + 256;
    char *status = (char *)c->arena.alloc(status_bytes);
    uint32_t *tile_counter = (uint32_t *)(status + (size_t)tiles * RS_RADIX * (wide ? 8 : 4));
    {
        ProfScope ps(c, PC_RADIX_HIST, 0, 1);
        rs_scan_hist_kernel<<<P.num, RS_RADIX, 0, c->stream>>>(hist);
    }
    K *kin = keys_a, *kout = keys_b;
    V *vin = vals_a, *vout = vals_b;
    if ((size_t)TILE * sizeof(K) > 32 * 1024) {
        BFQ_CUDA(cudaFuncSetAttribute(rs_onesweep_kernel<K, V, uint64_t, ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(TILE * sizeof(K))));
        BFQ_CUDA(cudaFuncSetAttribute(rs_onesweep_kernel<K, V, uint32_t, ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(TILE * sizeof(K))));
    }
    for (int p = 0; p < P.num; p++) {
        BFQ_CUDA(cudaMemsetAsync(status, 0, status_bytes, c->stream));
        {
            ProfScope ps(c, cls, 2 * n * (sizeof(K) + sizeof(V)));
            const size_t dyn = (size_t)TILE * sizeof(K);
            if constexpr (sizeof(K) == 4 && sizeof(V) == 2) {
                if (!(src && p == 0) && P.bits[p] <= 6 && !wide) {          // a plain pass of the permutation sort: 6-bit match
                    rs_onesweep_kernel<K, V, uint32_t, ITEMS, false, 6><<<tiles, RS_THREADS, dyn, c->stream>>>(
                        kin, kout, vin, vout, (uint32_t)n, P.shift[p], P.bits[p], hist + p * RS_RADIX, (uint32_t *)status, tile_counter);
                    BFQ_CUDA(cudaGetLastError());
                    std::swap(kin, kout);
                    std::swap(vin, vout);
                    continue;
                }
                if (src && p == 0) {
                    if (wide)
                        rs_onesweep_kernel<K, V, uint64_t, ITEMS, true><<<tiles, RS_THREADS, dyn, c->stream>>>(
                            kin, kout, vin, vout, (uint32_t)n, P.shift[p], P.bits[p], hist + p * RS_RADIX,
                            (uint64_t *)status, tile_counter, *src);
                    else if (P.bits[p] <= 6)
                        rs_onesweep_kernel<K, V, uint32_t, ITEMS, true, 6><<<tiles, RS_THREADS, dyn, c->stream>>>(
                            kin, kout, vin, vout, (uint32_t)n, P.shift[p], P.bits[p], hist + p * RS_RADIX,
                            (uint32_t *)status, tile_counter, *src);
                    else
                        rs_onesweep_kernel<K, V, uint32_t, ITEMS, true><<<tiles, RS_THREADS, dyn, c->stream>>>(
                            kin, kout, vin, vout, (uint32_t)n, P.shift[p], P.bits[p], hist + p * RS_RADIX,
                            (uint32_t *)status, tile_counter, *src);
                    BFQ_CUDA(cudaGetLastError());
                    std::swap(kin, kout);
                    std::swap(vin, vout);
                    continue;
                }
            }
            if (wide)
                rs_onesweep_kernel<K, V, uint64_t, ITEMS><<<tiles, RS_THREADS, dyn, c->stream>>>(
                    kin, kout, vin, vout, (uint32_t)n, P.shift[p], P.bits[p], hist + p * RS_RADIX,
                    (uint64_t *)status, tile_counter);
            else
                rs_onesweep_kernel<K, V, uint32_t, ITEMS><<<tiles, RS_THREADS, dyn, c->stream>>>(
                    kin, kout, vin, vout, (uint32_t)n, P.shift[p], P.bits[p], hist + p * RS_RADIX,
                    (uint32_t *)status, tile_counter);
        }
        BFQ_CUDA(cudaGetLastError());
        std::swap(kin, kout);
        std::swap(vin, vout);
    }
    return (P.num & 1);
}

template <typename K, int ITEMS, typename V = uint32_t>
static int radix_sort_pairs(bfq_ctx *c, K *keys_a, K *keys_b, V *vals_a, V *vals_b,
                            uint64_t n, int bits)
{
    if (n == 0 || bits <= 0) return 0;
    if (n >= ((uint64_t)1 << 32)) BFQ_FAIL(BFQ_ETOOLARGE, "radix sort: n=%llu needs 64-bit indices", (unsigned long long)n);
    RsPasses P = plan_passes(bits);
    Arena::Mark mk = c->arena.mark();
    uint32_t *hist = arena_alloc<uint32_t>(c, RS_MAX_PASSES * RS_RADIX);
    BFQ_CUDA(cudaMemsetAsync(hist, 0, RS_MAX_PASSES * RS_RADIX * sizeof(uint32_t), c->stream));
    {
        ProfScope ps(c, PC_RADIX_HIST, n * sizeof(K), 1);
        int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 8, div_up(n, RS_THREADS));
        rs_hist_kernel<K><<<grid, RS_THREADS, 0, c->stream>>>(keys_a, n, P, hist);
    }
    const int flip = radix_passes<K, V, ITEMS>(c, keys_a, keys_b, vals_a, vals_b, n, P, hist);
    // The arena memory (hist/status) is reused by later allocations only after stream order
    // guarantees completion: everything runs on c->stream.
    c->arena.release(mk);
    return flip;
}

// ------------------------------------------------------------------------------ key-range selection
// Single-collection mode over several GPUs: every rank holds the whole packed text and sorts only the suffixes
// whose key lies in its range.  The ranges are cut on the first four symbols (top 12 of the 48 key bits) from a
// histogram every rank computes identically, so no splitter exchange is needed and equal keys never straddle ranks.
// the 21-symbol window that starts `s` bits (s < 128) into three consecutive words
__device__ __forceinline__ uint64_t window3(uint64_t w0, uint64_t w1, uint64_t w2, int s)
{
    uint64_t w;
    if (s < 64) w = s ? ((w0 << s) | (w1 >> (64 - s))) : w0;
    else { const int t = s - 64; w = t ? ((w1 << t) | (w2 >> (64 - t))) : w1; }
    return w & ~1ull;
}
// its first four symbols, cut at the terminator: the top RS_TOP_BITS bits of the suffix key, which is all the range test needs
__device__ __forceinline__ uint32_t window_top12(uint64_t w)
{
    const uint32_t t = (uint32_t)(w >> 52);
    if (!(t & 0xE00u)) return 0u;
    if (!(t & 0x1C0u)) return t & 0xE00u;
    if (!(t & 0x038u)) return t & 0xFC0u;
    return t;
}
__device__ __forceinline__ uint64_t window_key(uint64_t w, int k)
{
    return cut_after_terminator(w) >> (64 - 3 * k);
}

constexpr int RS_TOP_BITS = 12;
constexpr int RS_SEL_ITEMS = 16;                      // consecutive positions per thread in the selection kernels
constexpr int RS_SEL_CHUNK = RS_THREADS * RS_SEL_ITEMS;
static_assert(3 * (RS_SEL_ITEMS - 1) + 63 + 63 < 192, "the windows of a thread's run of positions lie in three words");

__global__ void __launch_bounds__(RS_THREADS)
rs_top_hist_kernel(const uint64_t *__restrict__ T3, uint64_t n, int k, uint32_t *__restrict__ hist)
{
    __shared__ uint32_t sh[1 << RS_TOP_BITS];
    for (int i = threadIdx.x; i < (1 << RS_TOP_BITS); i += RS_THREADS) sh[i] = 0;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * RS_SEL_CHUNK;
    for (uint64_t p0 = ((uint64_t)blockIdx.x * RS_THREADS + threadIdx.x) * RS_SEL_ITEMS; p0 < n; p0 += stride) {
        const uint64_t bit = 3 * p0, wi = bit >> 6;
        const int s0 = (int)(bit & 63);
        const uint64_t w0 = T3[wi], w1 = T3[wi + 1], w2 = T3[wi + 2];      // 16 windows: 3 * 15 + 63 + s0 < 192 bits
#pragma unroll
        for (int i = 0; i < RS_SEL_ITEMS; i++)
            if (p0 + i < n) atomicAdd(&sh[window_top12(window3(w0, w1, w2, s0 + 3 * i))], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < (1 << RS_TOP_BITS); i += RS_THREADS)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// WRITE = false: per-chunk count of selected positions and the largest key below the range;
// WRITE = true : ordered compaction (position order is kept: the sort relies on it for ties) + digit histograms
template <bool WRITE, typename V>
__global__ void __launch_bounds__(RS_THREADS)
rs_select_kernel(const uint64_t *__restrict__ T3, const uint16_t *__restrict__ QL, uint64_t n, int k, uint32_t b0,
                 uint32_t b1, int pb, RsPasses P, uint32_t *__restrict__ chunk_count, const uint32_t *__restrict__ chunk_base,
                 unsigned long long *__restrict__ prev_key, uint64_t *__restrict__ keys, V *__restrict__ vals,
                 uint32_t *__restrict__ hist)
{
    __shared__ uint32_t sh[WRITE ? RS_MAX_PASSES * RS_RADIX : 1];
    __shared__ uint32_t s_warp[RS_WARPS];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (WRITE) { for (int i = tid; i < RS_MAX_PASSES * RS_RADIX; i += RS_THREADS) sh[i] = 0; __syncthreads(); }
    const uint64_t p0 = (uint64_t)blockIdx.x * RS_SEL_CHUNK + (uint64_t)tid * RS_SEL_ITEMS;
    uint64_t key[RS_SEL_ITEMS];
    uint32_t mine = 0;
    unsigned long long below = 0; bool any_below = false;
    uint64_t w0 = 0, w1 = 0, w2 = 0;
    const int s0 = (int)((3 * p0) & 63);
    if (p0 < n) { const uint64_t wi = (3 * p0) >> 6; w0 = T3[wi]; w1 = T3[wi + 1]; w2 = T3[wi + 2]; }
#pragma unroll
    for (int i = 0; i < RS_SEL_ITEMS; i++) {
        const uint64_t p = p0 + i;
        key[i] = ~0ull;
        if (p < n) {
            // The ranges are cut on the first four symbols, so those decide membership; the full key is built only for the
            // positions of this rank (1 / world of them) and, in the counting pass, for the positions of the last non-empty
            // bin below the range (pb, known from the histogram): the largest key below the range lies there.
            const uint64_t w = window3(w0, w1, w2, s0 + 3 * i);
            const uint32_t t = window_top12(w);
            if (t >= b0 && t < b1) { key[i] = window_key(w, k); mine++; }
            else if (!WRITE && (int)t == pb) { below = max(below, (unsigned long long)window_key(w, k)); any_below = true; }
        }
    }
    uint32_t x = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) s_warp[w] = x;
    __syncthreads();
    uint32_t add = 0, total = 0;
    for (int i = 0; i < RS_WARPS; i++) { if (i < w) add += s_warp[i]; total += s_warp[i]; }
    if (!WRITE) {
        if (tid == 0) chunk_count[blockIdx.x] = total;
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            below = max(below, (unsigned long long)__shfl_xor_sync(0xffffffffu, below, o));
            any_below |= (bool)__shfl_xor_sync(0xffffffffu, (int)any_below, o);
        }
        if (lane == 0 && any_below) atomicMax(prev_key, below | (1ull << 63));      // bit 63 = "a key below the range exists"
        return;
    }
    uint32_t dst = chunk_base[blockIdx.x] + add + x - mine;
#pragma unroll
    for (int i = 0; i < RS_SEL_ITEMS; i++) {
        if (key[i] != ~0ull) {
            const uint64_t p = p0 + i;
#pragma unroll
            for (int q = 0; q < RS_MAX_PASSES; q++)
                if (q < P.num) atomicAdd(&sh[q * RS_RADIX + digit_of(key[i], P.shift[q], (1u << P.bits[q]) - 1)], 1u);
            keys[dst] = QL ? (key[i] | ((uint64_t)QL[p] << 48)) : key[i];
            vals[dst] = (V)p;
            dst++;
        }
    }
    __syncthreads();
    for (int i = tid; i < P.num * RS_RADIX; i += RS_THREADS)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

__global__ void __launch_bounds__(1024)
rs_scan_chunks_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, uint32_t n, uint32_t *__restrict__ total)
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
        for (int j = 0; j < w; j++) add += s_warp[j];
        const uint32_t carry = s_carry;
        if (i < n) out[i] = carry + add + x - v;
        __syncthreads();
        if (tid == 1023) s_carry = carry + add + x;
        __syncthreads();
    }
    if (tid == 0) *total = s_carry;
}

template <typename V>
int radix_sort_suffix_keys_range(bfq_ctx *c, const uint64_t *T3, const uint16_t *QL, uint64_t n, int k, int rank, int world,
                                 uint64_t *keys_a, uint64_t *keys_b, V *pos_a, V *pos_b, uint64_t capacity,
                                 uint64_t *n_local, uint64_t *row_base, uint64_t *prev_key, bool *has_prev)
{
    if (sizeof(V) == 4 && n >= ((uint64_t)1 << 32)) BFQ_FAIL(BFQ_ETOOLARGE, "radix sort: n=%llu needs 64-bit indices", (unsigned long long)n);
    if (n >= ((uint64_t)1 << 43)) BFQ_FAIL(BFQ_ETOOLARGE, "collection of %llu rows", (unsigned long long)n);
    if (3 * k < RS_TOP_BITS) BFQ_FAIL(BFQ_EINVAL, "the key-range split needs an initial key of at least 4 symbols");
    cudaStream_t st = c->stream;
    RsPasses P = plan_passes(3 * k);
    Arena::Mark mk = c->arena.mark();
    const uint32_t nchunks = div_up(n, RS_SEL_CHUNK);
    uint32_t *hist = arena_alloc<uint32_t>(c, RS_MAX_PASSES * RS_RADIX);
    uint32_t *top = arena_alloc<uint32_t>(c, 1 << RS_TOP_BITS);
    uint32_t *cc = arena_alloc<uint32_t>(c, nchunks + 1), *cb = arena_alloc<uint32_t>(c, nchunks + 1);
    unsigned long long *d_prev = arena_alloc<unsigned long long>(c, 2);
    BFQ_CUDA(cudaMemsetAsync(hist, 0, RS_MAX_PASSES * RS_RADIX * sizeof(uint32_t), st));
    BFQ_CUDA(cudaMemsetAsync(top, 0, sizeof(uint32_t) << RS_TOP_BITS, st));
    BFQ_CUDA(cudaMemsetAsync(d_prev, 0, 16, st));
    std::vector<uint32_t> h(1 << RS_TOP_BITS);
    {
        ProfScope ps(c, PC_KEYS, n * 3 / 8);
        int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 8, div_up(n, RS_THREADS));
        rs_top_hist_kernel<<<grid, RS_THREADS, 0, st>>>(T3, n, k, top);
        BFQ_CUDA(cudaGetLastError());
    }
    BFQ_CUDA(cudaMemcpyAsync(h.data(), top, h.size() * 4, cudaMemcpyDeviceToHost, st));
    BFQ_CUDA(cudaStreamSynchronize(st));
    // rank r takes the bins [b_r, b_{r+1}) with b_r = first bin whose cumulative count reaches r * n / world
    auto bin_of = [&](int r) -> uint32_t {
        if (r <= 0) return 0;
        if (r >= world) return 1u << RS_TOP_BITS;
        const uint64_t want = n / (uint64_t)world * (uint64_t)r;
        uint64_t cum = 0;
        for (uint32_t b = 0; b < (1u << RS_TOP_BITS); b++) { if (cum >= want) return b; cum += h[b]; }
        return 1u << RS_TOP_BITS;
    };
    const uint32_t b0 = bin_of(rank), b1 = bin_of(rank + 1);
    uint64_t below = 0, mine = 0;
    for (uint32_t b = 0; b < b0; b++) below += h[b];
    for (uint32_t b = b0; b < b1; b++) mine += h[b];
    // (the 32-bit bin counters wrap for a bin of 2^32 suffixes; such a collection fails the capacity check below)
    if (mine >= ((uint64_t)1 << 32) - 64)
        BFQ_FAIL(BFQ_ETOOLARGE, "rank %d of %d would sort %llu suffixes; one GPU indexes its rows with 32 bits", rank, world, (unsigned long long)mine);
    int pb = -1;                                           // last non-empty bin below the range: it holds the key before the rank's first row
    for (uint32_t b = 0; b < b0; b++) if (h[b]) pb = (int)b;
    {
        ProfScope ps(c, PC_KEYS, n * 3 / 8, 2);
        rs_select_kernel<false, V><<<nchunks, RS_THREADS, 0, st>>>(T3, QL, n, k, b0, b1, pb, P, cc, nullptr, d_prev, nullptr, nullptr, nullptr);
        rs_scan_chunks_kernel<<<1, 1024, 0, st>>>(cc, cb, nchunks, (uint32_t *)(d_prev + 1));
        BFQ_CUDA(cudaGetLastError());
    }
    fetch_small(c, c->h_small, d_prev, 16);
    BFQ_CUDA(cudaStreamSynchronize(st));
    const uint64_t pk = c->h_small[0];
    const uint64_t cnt = (uint32_t)c->h_small[1];
    *has_prev = (pk >> 63) != 0;
    *prev_key = pk & ~(1ull << 63);
    *n_local = cnt;
    *row_base = below;
    if (cnt > capacity) BFQ_FAIL(BFQ_ECAPACITY, "rank %d of %d sorts %llu suffixes, capacity is %llu", rank, world, (unsigned long long)cnt, (unsigned long long)capacity);
    int flip = 0;
    if (cnt) {
        {
            ProfScope ps(c, PC_KEYS, n * 3 / 8 + cnt * 14);
            rs_select_kernel<true, V><<<nchunks, RS_THREADS, 0, st>>>(T3, QL, n, k, b0, b1, pb, P, nullptr, cb, nullptr, keys_a, pos_a, hist);
            BFQ_CUDA(cudaGetLastError());
        }
        flip = radix_passes<uint64_t, V, 16>(c, keys_a, keys_b, pos_a, pos_b, cnt, P, hist);
    }
    c->arena.release(mk);
    return flip;
}
template int radix_sort_suffix_keys_range<uint32_t>(bfq_ctx *, const uint64_t *, const uint16_t *, uint64_t, int, int, int, uint64_t *, uint64_t *,
                                                    uint32_t *, uint32_t *, uint64_t, uint64_t *, uint64_t *, uint64_t *, bool *);
template int radix_sort_suffix_keys_range<uint64_t>(bfq_ctx *, const uint64_t *, const uint16_t *, uint64_t, int, int, int, uint64_t *, uint64_t *,
                                                    uint64_t *, uint64_t *, uint64_t, uint64_t *, uint64_t *, uint64_t *, bool *);

int radix_sort_suffix_keys(bfq_ctx *c, const uint64_t *T3, const uint16_t *QL, uint64_t n, int k, uint64_t *keys_a,
                           uint64_t *keys_b, uint32_t *pos_a, uint32_t *pos_b)
{
    if (n == 0) return 0;
    if (n >= ((uint64_t)1 << 32)) BFQ_FAIL(BFQ_ETOOLARGE, "radix sort: n=%llu needs 64-bit indices", (unsigned long long)n);
    RsPasses P = plan_passes(3 * k);
    Arena::Mark mk = c->arena.mark();
    uint32_t *hist = arena_alloc<uint32_t>(c, RS_MAX_PASSES * RS_RADIX);
    BFQ_CUDA(cudaMemsetAsync(hist, 0, RS_MAX_PASSES * RS_RADIX * sizeof(uint32_t), c->stream));
    {
        ProfScope ps(c, PC_KEYS, n * (8 + 4 + 2) + n * 3 / 8);
        int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 8, div_up(n, RS_THREADS * 4));
        bool six8 = P.num == 6;
        for (int q = 0; q < P.num; q++) six8 = six8 && P.bits[q] == 8 && P.shift[q] == 8 * q;
        if (six8) rs_suffix_keys_kernel<true><<<grid, RS_THREADS, 0, c->stream>>>(T3, QL, (uint32_t)n, k, P, keys_a, pos_a, hist);
        else rs_suffix_keys_kernel<false><<<grid, RS_THREADS, 0, c->stream>>>(T3, QL, (uint32_t)n, k, P, keys_a, pos_a, hist);
        BFQ_CUDA(cudaGetLastError());
    }
    const int flip = radix_passes<uint64_t, uint32_t, RS_ITEMS_MAIN>(c, keys_a, keys_b, pos_a, pos_b, n, P, hist);
    c->arena.release(mk);
    return flip;
}

// exact number of keys in [0, n) whose digit (shift, bits) equals d
static uint32_t count_digit(uint64_t n, int shift, int bits, uint32_t d)
{
    const uint64_t period = (uint64_t)1 << (shift + bits), run = (uint64_t)1 << shift;
    const uint64_t full = n / period, rem = n % period;
    const uint64_t lo = (uint64_t)d << shift;
    uint64_t c = full * run;
    if (rem > lo) c += std::min(run, rem - lo);
    return (uint32_t)c;
}

// Last step of the permutation sort: after the passes on the high key bits, bucket b = keys [b << PL_BITS, (b + 1) << PL_BITS)
// sits exactly at elements [b << PL_BITS, ...) -- the keys are a permutation of 0..n-1 -- so one block puts every value of a
// bucket at smem[key & mask] and writes the bucket out in order.  No ranking, no key output: a pure stream.
constexpr int PL_BITS = 15;
constexpr int PL_THREADS = 512;
// VERIFY (keys that come from outside: bfq_dist_place): every key must lie in this block's bucket and no slot may be
// written twice -- then the keys of the bucket are exactly its positions; *bad is raised otherwise.
template <bool VERIFY>
__global__ void __launch_bounds__(PL_THREADS)
rs_place_kernel(const uint32_t *__restrict__ keys, const uint16_t *__restrict__ vals, uint32_t n, uint16_t *__restrict__ out,
                uint32_t *__restrict__ bad)
{
    extern __shared__ __align__(16) uint16_t pl_smem[];
    __shared__ uint32_t s_seen[VERIFY ? (1 << PL_BITS) / 32 : 1];
    const uint32_t base = blockIdx.x << PL_BITS;
    const uint32_t cnt = min(1u << PL_BITS, n - base);
    if constexpr (VERIFY) {
        for (uint32_t i = threadIdx.x; i < (1u << PL_BITS) / 32; i += PL_THREADS) s_seen[i] = 0;
        __syncthreads();
    }
    for (uint32_t i = threadIdx.x; i < cnt; i += PL_THREADS) {
        const uint32_t k = __ldcs(keys + base + i), slot = k & ((1u << PL_BITS) - 1u);
        if constexpr (VERIFY) {
            if ((k >> PL_BITS) != blockIdx.x || slot >= cnt || (atomicOr(&s_seen[slot >> 5], 1u << (slot & 31)) >> (slot & 31)) & 1u) {
                atomicOr(bad, 1u);
                continue;
            }
        }
        pl_smem[slot] = __ldcs(vals + base + i);
    }
    __syncthreads();
    if (cnt == (1u << PL_BITS)) {
        const uint4 *s = reinterpret_cast<const uint4 *>(pl_smem);
        uint4 *o = reinterpret_cast<uint4 *>(out + base);                       // base * 2 bytes is a multiple of 16
        for (uint32_t i = threadIdx.x; i < (1u << PL_BITS) / 8; i += PL_THREADS) __stcs(o + i, s[i]);
    } else {
        for (uint32_t i = threadIdx.x; i < cnt; i += PL_THREADS) out[base + i] = pl_smem[i];
    }
}

__global__ void __launch_bounds__(256)
rs_key_range_kernel(const uint32_t *__restrict__ keys, uint32_t n, uint32_t *__restrict__ bad)
{
    bool b = false;
    for (uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (uint64_t)gridDim.x * 256) b |= keys[i] >= n;
    if (__any_sync(0xffffffffu, b) && (threadIdx.x & 31) == 0) atomicOr(bad, 2u);
}

bool radix_permutation_generates(uint64_t n)
{
    return n > ((uint64_t)1 << PL_BITS) && n < ((uint64_t)1 << 32);      // at least one pass above the placement bits
}

int radix_sort_permutation(bfq_ctx *c, uint32_t *keys_a, uint32_t *keys_b, uint16_t *vals_a, uint16_t *vals_b, uint64_t n,
                           const PermSource *src, bool verify)
{
    if (src && !radix_permutation_generates(n)) BFQ_FAIL(BFQ_ECUDA, "internal error: no radix pass to generate the pairs in");
    if (n == 0) return 0;
    if (n >= ((uint64_t)1 << 32)) BFQ_FAIL(BFQ_ETOOLARGE, "radix sort: n=%llu needs 64-bit indices", (unsigned long long)n);
    int bits = 1;
    while (((uint64_t)1 << bits) < n) bits++;
    const bool aligned = ((reinterpret_cast<uintptr_t>(vals_a) | reinterpret_cast<uintptr_t>(vals_b)) & 15) == 0;
    const int low = aligned ? std::min(bits, PL_BITS) : 0;          // bits finished by the placement pass
    RsPasses P = plan_passes(bits - low);
    for (int p = 0; p < P.num; p++) P.shift[p] += low;
    Arena::Mark mk = c->arena.mark();
    int flip = 0;
    uint32_t *d_bad = nullptr;
    if (verify) {
        // the passes scatter with digit counts computed from n alone: a key >= n would be written out of bounds.  Checked
        // before anything moves; duplicates (which stay inside the arrays) are caught by the placement kernel.
        if (!low) BFQ_FAIL(BFQ_EINVAL, "pair lists must be 16-byte aligned");
        d_bad = arena_alloc<uint32_t>(c, 4);
        BFQ_CUDA(cudaMemsetAsync(d_bad, 0, 16, c->stream));
        ProfScope ps(c, PC_UNPERMUTE, n * 4);
        rs_key_range_kernel<<<(int)std::min<uint64_t>((uint64_t)c->num_sms * 8, div_up(n, 256)), 256, 0, c->stream>>>(keys_a, (uint32_t)n, d_bad);
        BFQ_CUDA(cudaGetLastError());
        fetch_small(c, c->h_small, d_bad, 4);
        BFQ_CUDA(cudaStreamSynchronize(c->stream));
        if (((const uint32_t *)c->h_small)[0])
            BFQ_FAIL(BFQ_EINVAL, "the pair list holds a position outside [0, %llu): it is not the set of text positions of this rank", (unsigned long long)n);
    }
    if (P.num) {
        uint32_t *hist = arena_alloc<uint32_t>(c, RS_MAX_PASSES * RS_RADIX);
        std::vector<uint32_t> h(RS_MAX_PASSES * RS_RADIX, 0);
        for (int p = 0; p < P.num; p++)
            for (uint32_t d = 0; d < (1u << P.bits[p]); d++) h[p * RS_RADIX + d] = count_digit(n, P.shift[p], P.bits[p], d);
        // pageable source: the copy is staged by the runtime before the call returns
        BFQ_CUDA(cudaMemcpyAsync(hist, h.data(), h.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
        flip = radix_passes<uint32_t, uint16_t, 16>(c, keys_a, keys_b, vals_a, vals_b, n, P, hist, src, PC_UNPERMUTE);
    }
    if (low) {
        const size_t smem = sizeof(uint16_t) << PL_BITS;
        BFQ_CUDA(cudaFuncSetAttribute(rs_place_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        BFQ_CUDA(cudaFuncSetAttribute(rs_place_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ProfScope ps(c, PC_UNPERMUTE, n * (4 + 2 + 2));
        if (verify)
            rs_place_kernel<true><<<div_up(n, (uint64_t)1 << PL_BITS), PL_THREADS, smem, c->stream>>>(
                flip ? keys_b : keys_a, flip ? vals_b : vals_a, (uint32_t)n, flip ? vals_a : vals_b, d_bad);
        else
            rs_place_kernel<false><<<div_up(n, (uint64_t)1 << PL_BITS), PL_THREADS, smem, c->stream>>>(
                flip ? keys_b : keys_a, flip ? vals_b : vals_a, (uint32_t)n, flip ? vals_a : vals_b, nullptr);
        BFQ_CUDA(cudaGetLastError());
        flip ^= 1;                                                  // only the values are in order now (the keys are known: 0..n-1)
    }
    if (verify) {
        fetch_small(c, c->h_small, d_bad, 4);
        BFQ_CUDA(cudaStreamSynchronize(c->stream));
        if (((const uint32_t *)c->h_small)[0]) {
            c->arena.release(mk);
            BFQ_FAIL(BFQ_EINVAL, "the pair list is not a permutation of the rank's %llu text positions (a position is missing or doubled)", (unsigned long long)n);
        }
    }
    c->arena.release(mk);
    return flip;
}

int radix_sort_pairs64(bfq_ctx *c, uint64_t *ka, uint64_t *kb, uint32_t *va, uint32_t *vb, uint64_t n, int bits)
{
    return radix_sort_pairs<uint64_t, 16>(c, ka, kb, va, vb, n, bits);
}
int radix_sort_pairs64(bfq_ctx *c, uint64_t *ka, uint64_t *kb, uint64_t *va, uint64_t *vb, uint64_t n, int bits)
{
    return radix_sort_pairs<uint64_t, 16, uint64_t>(c, ka, kb, va, vb, n, bits);
}
int radix_sort_pairs32(bfq_ctx *c, uint32_t *ka, uint32_t *kb, uint32_t *va, uint32_t *vb, uint64_t n, int bits)
{
    return radix_sort_pairs<uint32_t, 16>(c, ka, kb, va, vb, n, bits);
}
int radix_sort_pairs32_16(bfq_ctx *c, uint32_t *ka, uint32_t *kb, uint16_t *va, uint16_t *vb, uint64_t n, int bits)
{
    return radix_sort_pairs<uint32_t, 16, uint16_t>(c, ka, kb, va, vb, n, bits);
}

} // namespace bfq

// ------------------------------------------------------------------------------ sort without any allocation
// bfq_sort_pairs_noalloc (include/bfq_gpu.h): the same kernels as radix_sort_pairs64, for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch the passes need -- the
// per-tile status words of the decoupled look-back, the tile counter and the digit histograms -- lives in ONE file-scope
// device array of fixed size instead of the context's arena, so every GPU that loads the library holds it until exit
// and every call shares it: it keeps the last call's contents (what must start at zero is cleared by stream-ordered
// memsets below), and calls that overlap on the GPU race on it.
namespace bfq {

constexpr uint64_t NA_MAX_PAIRS = BFQ_NOALLOC_MAX_PAIRS;
constexpr int NA_ITEMS = 16;
constexpr uint32_t NA_TILE = RS_THREADS * NA_ITEMS;
constexpr uint32_t NA_MAX_TILES = (uint32_t)((NA_MAX_PAIRS + NA_TILE - 1) / NA_TILE);
constexpr size_t NA_STATUS_WORDS = (size_t)NA_MAX_TILES * RS_RADIX;         // u32 status words (n < 2^30)
constexpr size_t NA_COUNTER_WORDS = 64;                                     // tile counter (one word used; keeps hist aligned)
constexpr size_t NA_HIST_WORDS = (size_t)RS_MAX_PASSES * RS_RADIX;
constexpr size_t NA_WORDS = NA_STATUS_WORDS + NA_COUNTER_WORDS + NA_HIST_WORDS;   // 1 MiB + 256 B + 8 KiB
static_assert(NA_MAX_PAIRS < ((uint64_t)1 << 30), "32-bit status words");

static __device__ __align__(256) uint32_t g_noalloc_scratch[NA_WORDS];

// In host code the array's name is only a stand-in for the symbol, not a device address: the address is fetched with
// cudaGetSymbolAddress, once per device and with that device current, and handed to the kernels as an argument.
constexpr int NA_MAX_DEVICES = 64;
static std::mutex g_noalloc_mutex;
static uint32_t *g_noalloc_ptr[NA_MAX_DEVICES] = {nullptr};

static int noalloc_scratch(uint32_t **out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { set_error("bfq_sort_pairs_noalloc: cudaGetDevice -> %s", cudaGetErrorString(e)); return BFQ_ENODEVICE; }
    if (dev < 0 || dev >= NA_MAX_DEVICES) { set_error("bfq_sort_pairs_noalloc: device %d out of range", dev); return BFQ_EINVAL; }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[dev]) {
        uint32_t *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            set_error("bfq_sort_pairs_noalloc: cudaGetSymbolAddress on device %d -> %s", dev, cudaGetErrorString(e));
            return BFQ_ECUDA;
        }
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return BFQ_OK;
}

} // namespace bfq

extern "C" int bfq_sort_pairs_noalloc(uint64_t *d_keys_a, uint64_t *d_keys_b, uint32_t *d_vals_a, uint32_t *d_vals_b,
                                      uint64_t n, int bits, void *cuda_stream, int *result_in_b)
{
    using namespace bfq;
    if (result_in_b) *result_in_b = 0;
    if (bits < 0 || bits > 64) { set_error("bfq_sort_pairs_noalloc: bits must be 0..64"); return BFQ_EINVAL; }
    if (n > NA_MAX_PAIRS) {
        set_error("bfq_sort_pairs_noalloc: %llu pairs exceed the fixed scratch of this entry point (at most %llu); "
                  "use a context-based call for larger inputs", (unsigned long long)n, (unsigned long long)NA_MAX_PAIRS);
        return BFQ_ECAPACITY;
    }
    if (n == 0 || bits == 0) return BFQ_OK;                 // nothing to order: the input (a) is the result
    if (!d_keys_a || !d_keys_b || !d_vals_a || !d_vals_b) { set_error("bfq_sort_pairs_noalloc: null pointer"); return BFQ_EINVAL; }
    uint32_t *scratch = nullptr;
    const int rc = noalloc_scratch(&scratch);
    if (rc != BFQ_OK) return rc;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    uint32_t *status = scratch, *tile_counter = scratch + NA_STATUS_WORDS, *hist = tile_counter + NA_COUNTER_WORDS;
    const RsPasses P = plan_passes(bits);
    const uint32_t tiles = div_up(n, NA_TILE);
    cudaError_t e;
#define NA_CUDA(call)                                                                                        \
    do {                                                                                                     \
        e = (call);                                                                                          \
        if (e != cudaSuccess) { set_error("bfq_sort_pairs_noalloc: %s -> %s", #call, cudaGetErrorString(e)); return BFQ_ECUDA; } \
    } while (0)
    // the array keeps the previous call's contents: the histograms are accumulated, so they start at zero
    NA_CUDA(cudaMemsetAsync(hist, 0, NA_HIST_WORDS * sizeof(uint32_t), st));
    {
        const int grid = (int)std::min<uint64_t>((uint64_t)148 * 8, div_up(n, RS_THREADS));   // grid-stride kernel: 8 blocks per B200 SM
        rs_hist_kernel<uint64_t><<<grid, RS_THREADS, 0, st>>>(d_keys_a, n, P, hist);
        rs_scan_hist_kernel<<<P.num, RS_RADIX, 0, st>>>(hist);
        NA_CUDA(cudaGetLastError());
    }
    uint64_t *kin = d_keys_a, *kout = d_keys_b;
    uint32_t *vin = d_vals_a, *vout = d_vals_b;
    const size_t dyn = (size_t)NA_TILE * sizeof(uint64_t);   // 32 KiB: below the opt-in threshold, no attribute call needed
    for (int p = 0; p < P.num; p++) {
        // status words (flag 0 = "not published yet") and the tile counter of THIS pass: the words of the used tiles and
        // the counter are contiguous only when all tiles are used, so they are cleared separately
        NA_CUDA(cudaMemsetAsync(status, 0, (size_t)tiles * RS_RADIX * sizeof(uint32_t), st));
        NA_CUDA(cudaMemsetAsync(tile_counter, 0, NA_COUNTER_WORDS * sizeof(uint32_t), st));
        rs_onesweep_kernel<uint64_t, uint32_t, uint32_t, NA_ITEMS><<<tiles, RS_THREADS, dyn, st>>>(
            kin, kout, vin, vout, (uint32_t)n, P.shift[p], P.bits[p], hist + p * RS_RADIX, status, tile_counter);
        NA_CUDA(cudaGetLastError());
        std::swap(kin, kout);
        std::swap(vin, vout);
    }
#undef NA_CUDA
    if (result_in_b) *result_in_b = P.num & 1;
    return BFQ_OK;
}
