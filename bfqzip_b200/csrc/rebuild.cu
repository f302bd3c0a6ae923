// rebuild.cu -- part C on the device: every read is an independent inverse-LF walk over the
// walk lines built by smooth.cu, emitting (output base, stored quality) backwards.
//
// Replaces: src_ext_mem/decode.cpp decodeBCRmultipleReverse
// :499-686, RecoverNsymbolsReverseByVector :691-966 (one fseek+fread pair per symbol there,
// one 64-byte line fetch here), transposeCycFile :409-496 (no analogue: reads are written in
// place), convertFromCycFileToFastaOrFastq :968-1152 (strip, reverse).
// Read g starts at row g (decode.cpp:578-582); at most max_len symbols are recovered
// (decode.cpp:605); the walk stops at the end-marker (decode.cpp:890-891).
#include "common.cuh"
#include <algorithm>

namespace bfq {

struct WalkHdr { uint32_t cnt[5]; uint32_t plane[3]; };

// off_known != NULL: read lengths are known (fused pipeline) -> symbols go straight to their
// final place.  Otherwise they go to a [nreads x max_len] scratch in walk order and lens[g]
// records how many were found; assemble_kernel reverses and compacts them.
__global__ void __launch_bounds__(256)
walk_kernel(const uint8_t *__restrict__ lines, const uint32_t *__restrict__ super, const uint8_t *__restrict__ obase, uint32_t nreads,
            int max_len, uint64_t code2byte, const uint64_t *__restrict__ off_known,
            uint8_t *__restrict__ out_b, uint8_t *__restrict__ out_q, uint32_t *__restrict__ lens, uint32_t g0 = 0)
{
    const uint32_t g = g0 + blockIdx.x * 256u + threadIdx.x;
    if (g >= nreads) return;
    uint64_t row = g;
    uint64_t wpos;            // next write index
    if (off_known) wpos = off_known[g + 1];
    else wpos = (uint64_t)g * (uint32_t)max_len;
    // Known lengths: the read is written backwards into its final place.  Bytes are gathered in two
    // 8-byte shift registers and leave as one aligned 64-bit store per 8 symbols (single-byte stores
    // scattered over the kernel's lifetime cost a 32-byte read-modify-write in DRAM each).
    uint64_t accb = 0, accq = 0;
    int cnt = 0;
    const bool wide = ((reinterpret_cast<uintptr_t>(out_b) | reinterpret_cast<uintptr_t>(out_q)) & 7) == 0;
    int k = 0;
    for (; k < max_len; k++) {
        // one 32-byte sector per step: codes + in-block counts in its first half, the row's quality in the second
        const WalkSector S = ws_load(lines, row);
        const uint8_t q = *ws_q(lines, row);
        const uint32_t c = ws_code(S, (uint32_t)(row & 15));
        if (c == 0) break;                                   // end-marker: read complete
        uint8_t base = (uint8_t)(code2byte >> (8 * c));
        if (q & 0x80) base = obase[row];                     // noise-reduced row (BWT_MOD, bfq_ext.cpp:555)
        if (off_known) {
            wpos--;
            accb = (accb << 8) | base;
            accq = (accq << 8) | (uint64_t)(q & 0x7F);
            cnt++;
            if ((wpos & 7) == 0 || !wide) {
                if (cnt == 8) {
                    *reinterpret_cast<uint64_t *>(out_b + wpos) = accb;
                    *reinterpret_cast<uint64_t *>(out_q + wpos) = accq;
                } else {
                    for (int t = 0; t < cnt; t++) { out_b[wpos + t] = (uint8_t)(accb >> (8 * t)); out_q[wpos + t] = (uint8_t)(accq >> (8 * t)); }
                }
                cnt = 0;
            }
        } else { out_b[wpos] = base; out_q[wpos] = q & 0x7F; wpos++; }
        row = ws_lf(S, super, row, c);
    }
    if (off_known)
        for (int t = 0; t < cnt; t++) { out_b[wpos + t] = (uint8_t)(accb >> (8 * t)); out_q[wpos + t] = (uint8_t)(accq >> (8 * t)); }
    if (lens) lens[g] = (uint32_t)k;
}

// ---- exclusive scan u32 -> u64 (three small kernels; R elements, a few MB)
__global__ void __launch_bounds__(256) scan_reduce_kernel(const uint32_t *in, uint32_t n, unsigned long long *blk)
{
    __shared__ unsigned long long s[8];
    uint32_t i0 = blockIdx.x * 1024u + threadIdx.x * 4u;
    unsigned long long v = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) if (i0 + k < n) v += in[i0 + k];
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) { unsigned long long t = 0; for (int k = 0; k < 8; k++) t += s[k]; blk[blockIdx.x] = t; }
}
__global__ void __launch_bounds__(1024) scan_blocks_kernel(unsigned long long *blk, uint32_t nblk, unsigned long long *total)
{
    __shared__ unsigned long long s_warp[32];
    __shared__ unsigned long long s_carry;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < nblk; base += 1024) {
        uint32_t i = base + tid;
        unsigned long long v = i < nblk ? blk[i] : 0, x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { unsigned long long y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) s_warp[w] = x;
        __syncthreads();
        unsigned long long add = 0;
        for (int j = 0; j < w; j++) add += s_warp[j];
        unsigned long long carry = s_carry;
        if (i < nblk) blk[i] = carry + add + x - v;
        __syncthreads();
        if (tid == 1023) s_carry = carry + add + x;
        __syncthreads();
    }
    if (tid == 0) *total = s_carry;
}
__global__ void __launch_bounds__(256) scan_down_kernel(const uint32_t *in, uint32_t n, const unsigned long long *blk,
                                                       unsigned long long *out, const unsigned long long *total)
{
    __shared__ unsigned long long s[8];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    uint32_t i0 = blockIdx.x * 1024u + tid * 4u;
    uint32_t v[4]; unsigned long long sum = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) { v[k] = (i0 + k < n) ? in[i0 + k] : 0; sum += v[k]; }
    unsigned long long x = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { unsigned long long y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) s[w] = x;
    __syncthreads();
    unsigned long long add = blk[blockIdx.x];
    for (int j = 0; j < w; j++) add += s[j];
    unsigned long long run = add + x - sum;
#pragma unroll
    for (int k = 0; k < 4; k++) { if (i0 + k < n) out[i0 + k] = run; run += v[k]; }
    if (blockIdx.x == 0 && tid == 0) out[n] = *total;
}

void scan_u32_u64_dev(bfq_ctx *c, const uint32_t *in, uint64_t n, unsigned long long *out)
{
    if (n >= ((uint64_t)1 << 32)) BFQ_FAIL(BFQ_ETOOLARGE, "scan of %llu elements", (unsigned long long)n);
    const uint32_t nblk = div_up(n, 1024);
    unsigned long long *blk = arena_alloc<unsigned long long>(c, (size_t)nblk + 2);
    scan_reduce_kernel<<<nblk, 256, 0, c->stream>>>(in, (uint32_t)n, blk);
    scan_blocks_kernel<<<1, 1024, 0, c->stream>>>(blk, nblk, blk + nblk);
    scan_down_kernel<<<nblk, 256, 0, c->stream>>>(in, (uint32_t)n, blk, out, blk + nblk);
    BFQ_CUDA(cudaGetLastError());
}

// one warp per read: reverse the walk-order scratch into the final position
__global__ void __launch_bounds__(256)
assemble_kernel(const uint8_t *__restrict__ sb, const uint8_t *__restrict__ sq, const uint32_t *__restrict__ lens,
                const unsigned long long *__restrict__ off, uint32_t nreads, int max_len,
                uint8_t *__restrict__ out_b, uint8_t *__restrict__ out_q)
{
    const int lane = threadIdx.x & 31;
    const uint32_t warp0 = (blockIdx.x * 256u + threadIdx.x) >> 5, nwarps = (gridDim.x * 256u) >> 5;
    for (uint32_t g = warp0; g < nreads; g += nwarps) {
        const uint32_t len = lens[g];
        const uint64_t src = (uint64_t)g * (uint32_t)max_len, dst = off[g];
        for (uint32_t j = lane; j < len; j += 32) {
            out_b[dst + j] = sb[src + (len - 1 - j)];
            out_q[dst + j] = sq[src + (len - 1 - j)];
        }
    }
}

// ---------------------------------------------------------------- flat mode: rebuild without the LF walk
// Row i of the BWT holds the symbol at text position SA[i] - 1.  (text position, base | quality << 8) pairs are
// sorted on the position -- the keys are exactly 0..n-1, so the passes are pure streaming (radix_sort_permutation)
// -- and the reads are then copied out without their terminator slots.  At large n this replaces one dependent
// 64-byte random fetch per base (DRAM sector / TLB bound) with ~64 streamed bytes per row.
__global__ void __launch_bounds__(256)
emit_pairs_kernel(const uint32_t *__restrict__ SA, const uint8_t *__restrict__ bwt, const uint8_t *__restrict__ qflat,
                  const uint8_t *__restrict__ obase, uint32_t n, uint32_t *__restrict__ keys, uint16_t *__restrict__ vals,
                  uint32_t n_total)
{
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i >= n) return;
    const uint32_t p = SA[i];
    const uint8_t q = qflat[i];
    const uint8_t base = (q & 0x80) ? obase[i] : bwt[i];
    keys[i] = p ? p - 1 : n_total - 1;                 // read starts land on a terminator slot (never copied out)
    vals[i] = (uint16_t)(base | ((uint32_t)(q & 0x7F) << 8));
}

// one warp per read: text-order pairs (terminator slots included) -> the two output arrays
// reads [read_lo, read_lo + nreads): T starts at text position t_lo, the outputs at base offset off[read_lo]
__global__ void __launch_bounds__(256)
compact_reads_kernel(const uint16_t *__restrict__ T, const uint64_t *__restrict__ off, uint64_t nreads,
                     uint8_t *__restrict__ out_b, uint8_t *__restrict__ out_q, uint64_t read_lo = 0, uint64_t t_lo = 0)
{
    const int lane = threadIdx.x & 31;
    const uint64_t warp0 = ((uint64_t)blockIdx.x * 256 + threadIdx.x) >> 5, nwarps = ((uint64_t)gridDim.x * 256) >> 5;
    const uint64_t o_lo = read_lo ? off[read_lo] : 0;
    for (uint64_t d = read_lo + warp0; d < read_lo + nreads; d += nwarps) {
        const uint64_t s = off[d], len = off[d + 1] - s, src = s + d - t_lo, dst = s - o_lo;
        for (uint64_t j = lane; j < len; j += 32) {
            const uint16_t v = T[src + j];
            out_b[dst + j] = (uint8_t)v;
            out_q[dst + j] = (uint8_t)(v >> 8);
        }
    }
}

void unpermute_dev(bfq_ctx *c, const SmoothResult &S, const uint8_t *d_bwt, const uint32_t *d_sa,
                   const uint64_t *d_off, uint64_t nreads, uint8_t *d_out_bases, uint8_t *d_out_quals)
{
    cudaStream_t st = c->stream;
    const uint64_t n = S.n;
    if (n == 0 || nreads == 0) return;
    Arena::Mark mk = c->arena.mark();
    uint32_t *ka = arena_alloc<uint32_t>(c, n + 16), *kb = arena_alloc<uint32_t>(c, n + 16);
    uint16_t *va = arena_alloc<uint16_t>(c, n + 16), *vb = arena_alloc<uint16_t>(c, n + 16);
    // the first sort pass reads the smoothed rows itself whenever the sort has one; tiny collections emit the pairs first
    const bool fused = radix_permutation_generates(n);
    const PermSource src{d_sa, d_bwt, S.lines, S.obase, (uint32_t)n};
    if (!fused) {
        ProfScope ps(c, PC_WALK, n * (4 + 3 + 6));
        emit_pairs_kernel<<<div_up(n, 256), 256, 0, st>>>(d_sa, d_bwt, S.lines, S.obase, (uint32_t)n, ka, va, (uint32_t)n);
        BFQ_CUDA(cudaGetLastError());
    }
    const int flip = radix_sort_permutation(c, ka, kb, va, vb, n, fused ? &src : nullptr);
    {
        ProfScope ps(c, PC_ASSEMBLE, n * 2 + (n - nreads) * 2);
        int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 16, div_up(nreads * 32, 256));
        compact_reads_kernel<<<std::max(grid, 1), 256, 0, st>>>(flip ? vb : va, d_off, nreads, d_out_bases, d_out_quals);
        BFQ_CUDA(cudaGetLastError());
    }
    c->arena.release(mk);
}

// ---------------------------------------------------------------- single-collection mode over several GPUs
// Every smoothed row goes to the rank that rebuilds its text position.  A row becomes (destination, position relative to
// the destination's first text position | value << 32); ONE radix pass on the destination groups them, then the two
// halves are split into the arrays the all-to-all sends.  Relative positions fit 32 bits even when the collection does not.
constexpr int DIST_MAX_WORLD = 255;

template <typename P>
__global__ void __launch_bounds__(256)
emit_dest_pairs_kernel(const P *__restrict__ SA, const uint8_t *__restrict__ bwt, const uint8_t *__restrict__ qflat,
                       const uint8_t *__restrict__ obase, uint32_t n, uint64_t n_total, const uint64_t *__restrict__ bounds,
                       int world, uint64_t *__restrict__ dest, uint64_t *__restrict__ packed)
{
    __shared__ uint64_t s_b[DIST_MAX_WORLD + 1];
    for (int d = threadIdx.x; d <= world; d += 256) s_b[d] = bounds[d];
    __syncthreads();
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i >= n) return;
    const uint64_t p = (uint64_t)SA[i];
    const uint64_t tp = p ? p - 1 : n_total - 1;         // read starts land on a terminator slot (never copied out)
    int lo = 0, hi = world;                               // last d with s_b[d] <= tp
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (s_b[mid] <= tp) lo = mid; else hi = mid; }
    const uint8_t q = qflat[i];
    const uint8_t base = (q & 0x80) ? obase[i] : bwt[i];
    dest[i] = (uint64_t)lo;
    packed[i] = (tp - s_b[lo]) | ((uint64_t)(base | ((uint32_t)(q & 0x7F) << 8)) << 32);
}

__global__ void __launch_bounds__(256)
split_pairs_kernel(const uint64_t *__restrict__ packed, uint32_t n, uint32_t *__restrict__ keys, uint16_t *__restrict__ vals)
{
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i >= n) return;
    const uint64_t v = packed[i];
    keys[i] = (uint32_t)v;
    vals[i] = (uint16_t)(v >> 32);
}

__global__ void dest_counts_kernel(const uint64_t *__restrict__ dest, uint32_t n, int world, unsigned long long *__restrict__ lower)
{
    const int d = threadIdx.x;
    if (d > world) return;
    uint32_t lo = 0, hi = n;                      // first index with dest >= d
    while (lo < hi) { const uint32_t mid = lo + ((hi - lo) >> 1); if (dest[mid] < (uint64_t)d) lo = mid + 1; else hi = mid; }
    lower[d] = lo;
}

void emit_sorted_pairs_dev(bfq_ctx *c, const SmoothResult &S, const uint8_t *d_bwt, const void *d_sa, bool sa_wide, uint64_t n_total,
                           const uint64_t *text_bounds, int world, uint32_t *d_keys, uint16_t *d_vals, uint64_t *send_counts)
{
    cudaStream_t st = c->stream;
    const uint64_t n = S.n;
    for (int d = 0; d < world; d++) send_counts[d] = 0;
    if (n == 0) return;
    if (world > DIST_MAX_WORLD) BFQ_FAIL(BFQ_EINVAL, "at most %d ranks", DIST_MAX_WORLD);
    for (int d = 0; d < world; d++)
        if (text_bounds[d + 1] < text_bounds[d] || text_bounds[d + 1] - text_bounds[d] >= ((uint64_t)1 << 32))
            BFQ_FAIL(BFQ_EINVAL, "text_bounds must be non-decreasing with fewer than 2^32 positions per rank");
    if (text_bounds[0] != 0 || text_bounds[world] != n_total) BFQ_FAIL(BFQ_EINVAL, "text_bounds must cover [0, rows of the collection)");
    Arena::Mark mk = c->arena.mark();
    uint64_t *da = arena_alloc<uint64_t>(c, n + 16), *db = arena_alloc<uint64_t>(c, n + 16);
    uint64_t *pa = arena_alloc<uint64_t>(c, n + 16), *pb = arena_alloc<uint64_t>(c, n + 16);
    uint64_t *d_bounds = arena_alloc<uint64_t>(c, 2 * (world + 1));
    unsigned long long *d_lower = (unsigned long long *)(d_bounds + world + 1);
    BFQ_CUDA(cudaMemcpyAsync(d_bounds, text_bounds, (world + 1) * 8, cudaMemcpyHostToDevice, st));
    {
        ProfScope ps(c, PC_WALK, n * (4 + 3 + 16));
        if (sa_wide)
            emit_dest_pairs_kernel<uint64_t><<<div_up(n, 256), 256, 0, st>>>((const uint64_t *)d_sa, d_bwt, S.lines, S.obase, (uint32_t)n, n_total, d_bounds, world, da, pa);
        else
            emit_dest_pairs_kernel<uint32_t><<<div_up(n, 256), 256, 0, st>>>((const uint32_t *)d_sa, d_bwt, S.lines, S.obase, (uint32_t)n, n_total, d_bounds, world, da, pa);
        BFQ_CUDA(cudaGetLastError());
    }
    int bits = 1;
    while ((1 << bits) < world) bits++;
    const int flip = world > 1 ? radix_sort_pairs64(c, da, db, pa, pb, n, bits) : 0;
    {
        ProfScope ps(c, PC_ASSEMBLE, n * (8 + 6), 2);
        split_pairs_kernel<<<div_up(n, 256), 256, 0, st>>>(flip ? pb : pa, (uint32_t)n, d_keys, d_vals);
        dest_counts_kernel<<<1, 256, 0, st>>>(flip ? db : da, (uint32_t)n, world, d_lower);
        BFQ_CUDA(cudaGetLastError());
    }
    std::vector<uint64_t> lower(world + 1);
    BFQ_CUDA(cudaMemcpyAsync(lower.data(), d_lower, (world + 1) * 8, cudaMemcpyDeviceToHost, st));
    BFQ_CUDA(cudaStreamSynchronize(st));
    for (int d = 0; d < world; d++) send_counts[d] = lower[d + 1] - lower[d];
    c->arena.release(mk);
}

void place_pairs_dev(bfq_ctx *c, uint32_t *d_keys, uint16_t *d_vals, uint64_t cnt, uint64_t t_lo, const uint64_t *d_off,
                     uint64_t read_lo, uint64_t read_hi, uint8_t *d_out_bases, uint8_t *d_out_quals)
{
    cudaStream_t st = c->stream;
    if (cnt == 0 || read_hi <= read_lo) return;
    // the caller's numbers must describe one range of whole reads: cnt text positions from t_lo = off[read_lo] + read_lo on
    fetch_small(c, c->h_small, d_off + read_lo, 8);
    fetch_small(c, c->h_small + 1, d_off + read_hi, 8);
    BFQ_CUDA(cudaStreamSynchronize(st));
    if (c->h_small[0] + read_lo != t_lo || c->h_small[1] + read_hi - t_lo != cnt)
        BFQ_FAIL(BFQ_EINVAL, "reads [%llu, %llu) cover text positions [%llu, %llu), not %llu positions from %llu on", (unsigned long long)read_lo,
                 (unsigned long long)read_hi, (unsigned long long)(c->h_small[0] + read_lo), (unsigned long long)(c->h_small[1] + read_hi),
                 (unsigned long long)cnt, (unsigned long long)t_lo);
    if (cnt >= ((uint64_t)1 << 32)) BFQ_FAIL(BFQ_ETOOLARGE, "%llu positions for one rank; a rank indexes its positions with 32 bits", (unsigned long long)cnt);
    Arena::Mark mk = c->arena.mark();
    uint32_t *kb = arena_alloc<uint32_t>(c, cnt + 16);
    uint16_t *vb = arena_alloc<uint16_t>(c, cnt + 16);
    const int flip = radix_sort_permutation(c, d_keys, kb, d_vals, vb, cnt, nullptr, /*verify=*/true);
    {
        const uint64_t nreads = read_hi - read_lo;
        ProfScope ps(c, PC_ASSEMBLE, cnt * 2 + (cnt - nreads) * 2);
        int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 16, div_up(nreads * 32, 256));
        compact_reads_kernel<<<std::max(grid, 1), 256, 0, st>>>(flip ? vb : d_vals, d_off, nreads, d_out_bases, d_out_quals, read_lo, t_lo);
        BFQ_CUDA(cudaGetLastError());
    }
    BFQ_CUDA(cudaStreamSynchronize(st));
    c->arena.release(mk);
}

void rebuild_dev(bfq_ctx *c, const SmoothResult &S, const bfq_params *P, const uint64_t *d_off_known,
                 uint8_t *d_out_bases, uint8_t *d_out_quals, uint64_t *d_out_off, uint64_t *total_out)
{
    cudaStream_t st = c->stream;
    const uint64_t R = S.nreads;
    if (R >= ((uint64_t)1 << 32)) BFQ_FAIL(BFQ_ETOOLARGE, "too many reads for 32-bit read indices (dataTypeNSeq is uint, parameters.h:82)");
    if (total_out) *total_out = 0;
    if (R == 0) {
        if (d_out_off) BFQ_CUDA(cudaMemsetAsync(d_out_off, 0, 8, st));
        return;
    }
    uint64_t c2b = 0;
    for (int k = 0; k < 8; k++) c2b |= (uint64_t)S.alpha.code2byte[k] << (8 * k);
    if (d_off_known) {
        bfq_ctx::HostSink *sink = c->sink;
        // Host-buffer call: the walk runs one GPU-full of reads at a time (a read is one thread, so a slice of resident
        // size costs nothing) and every slice's reads leave for the host while the next slice walks.
        const uint64_t wave = (uint64_t)c->num_sms * 2048;
        const bool sliced = sink && sink->out_b && sink->out_q && sink->off && R >= 2 * wave;
        const uint64_t step = sliced ? wave : R;
        for (uint64_t g0 = 0; g0 < R; g0 += step) {
            const uint64_t g1 = std::min(R, g0 + step);
            {
                ProfScope ps(c, PC_WALK, (S.n - R) * (32 + 2) * (g1 - g0) / R);
                walk_kernel<<<div_up(g1 - g0, 256), 256, 0, st>>>(S.lines, S.super, S.obase, (uint32_t)g1, P->max_len, c2b, d_off_known,
                                                                 d_out_bases, d_out_quals, nullptr, (uint32_t)g0);
                BFQ_CUDA(cudaGetLastError());
            }
            if (sliced) {
                // walk_kernel writes exactly the bytes of its own read, so [off[g0], off[g1]) is final here
                const uint64_t lo = sink->off[g0], hi = sink->off[g1];
                BFQ_CUDA(cudaEventRecord(c->ev_a, st));
                BFQ_CUDA(cudaStreamWaitEvent(c->copy_stream, c->ev_a, 0));
                BFQ_CUDA(cudaMemcpyAsync(sink->out_b + lo, d_out_bases + lo, hi - lo, cudaMemcpyDeviceToHost, c->copy_stream));
                BFQ_CUDA(cudaMemcpyAsync(sink->out_q + lo, d_out_quals + lo, hi - lo, cudaMemcpyDeviceToHost, c->copy_stream));
                sink->out_copied = true;
            }
        }
        return;
    }
    Arena::Mark mk = c->arena.mark();
    uint8_t *sb = arena_alloc<uint8_t>(c, R * (uint64_t)P->max_len);
    uint8_t *sq = arena_alloc<uint8_t>(c, R * (uint64_t)P->max_len);
    uint32_t *lens = arena_alloc<uint32_t>(c, R);
    const uint32_t nblk = div_up(R, 1024);
    unsigned long long *blk = arena_alloc<unsigned long long>(c, nblk + 1);
    {
        ProfScope ps(c, PC_WALK, (S.n - R) * (32 + 2));
        walk_kernel<<<div_up(R, 256), 256, 0, st>>>(S.lines, S.super, S.obase, (uint32_t)R, P->max_len, c2b, nullptr, sb, sq, lens);
        BFQ_CUDA(cudaGetLastError());
    }
    {
        ProfScope ps(c, PC_ASSEMBLE, (S.n - R) * 4, 4);
        scan_reduce_kernel<<<nblk, 256, 0, st>>>(lens, (uint32_t)R, blk);
        scan_blocks_kernel<<<1, 1024, 0, st>>>(blk, nblk, blk + nblk);
        scan_down_kernel<<<nblk, 256, 0, st>>>(lens, (uint32_t)R, blk, (unsigned long long *)d_out_off, blk + nblk);
        int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 16, div_up(R * 32, 256));
        assemble_kernel<<<grid, 256, 0, st>>>(sb, sq, lens, (const unsigned long long *)d_out_off, (uint32_t)R,
                                              P->max_len, d_out_bases, d_out_quals);
        BFQ_CUDA(cudaGetLastError());
    }
    if (total_out) {
        fetch_small(c, c->h_small, blk + nblk, 8);
        BFQ_CUDA(cudaStreamSynchronize(st));
        *total_out = c->h_small[0];
    } else {
        BFQ_CUDA(cudaStreamSynchronize(st));
    }
    c->arena.release(mk);
}

} // namespace bfq
