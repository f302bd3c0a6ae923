// rle.cu -- quality-stream pre-transform for the host entropy coders.
//
// After smoothing, the quality stream of step 4 (BFQzip_ext.py:224-244: `sed -n 4~4p OUT.fq > OUT.fq.qs`, one line of
// qualities per read) is dominated by runs of the cluster's new quality (the constant of M=2, the max / mean of the other
// modes).  The reference hands that stream to `7z a -mm=PPMd` / `bsc` as it is (BFQzip_ext.py:248-272).  This transform
// writes it as two byte streams of equal length -- run VALUES and run LENGTHS -- which a byte-oriented coder models
// separately; it is invertible and changes no contract (the .fq and the component streams stay as they are).
// The stream is newline-separated lines of at most 255 symbols (reads are limited to 255, parameters.h:68) and a newline
// ends every run, so a run is at most 255 bytes and its length fits one byte; a longer run is refused (BFQ_EINVAL).
//
// Streaming byte work, HBM-bound: encode reads the stream twice (count, write: 2 B per byte) and writes 6 B per run;
// no tensor cores.  There is no counterpart kernel in the reference (its step 5 is host code and stays there).
#include "common.cuh"
#include <algorithm>

namespace bfq {

constexpr int RL_THREADS = 256;
constexpr int RL_PER = 16;                               // bytes per thread
constexpr int RL_TILE = RL_THREADS * RL_PER;

// bit j of the result: byte i0 + j starts a run (differs from the byte before it, or is the first byte of the stream)
__device__ __forceinline__ uint32_t rle_heads(const uint8_t *__restrict__ q, uint64_t n, uint64_t i0, uint8_t b[RL_PER])
{
    if (i0 >= n) return 0u;
    uint8_t prev = i0 ? q[i0 - 1] : 0;
    if (i0 + RL_PER <= n && ((reinterpret_cast<uintptr_t>(q) + i0) & 15) == 0) {
        const uint4 v = *reinterpret_cast<const uint4 *>(q + i0);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < RL_PER; j++) b[j] = (uint8_t)(w[j >> 2] >> (8 * (j & 3)));
    } else {
#pragma unroll
        for (int j = 0; j < RL_PER; j++) b[j] = i0 + j < n ? q[i0 + j] : 0;
    }
    uint32_t m = 0;
#pragma unroll
    for (int j = 0; j < RL_PER; j++) {
        if (i0 + j < n && (b[j] != prev || (i0 + j) == 0)) m |= 1u << j;
        prev = b[j];
    }
    return m;
}

__global__ void __launch_bounds__(RL_THREADS)
rle_count_kernel(const uint8_t *__restrict__ q, uint64_t n, uint32_t *__restrict__ counts)
{
    __shared__ uint32_t s_w[RL_THREADS / 32];
    uint8_t b[RL_PER];
    const uint32_t m = rle_heads(q, n, (uint64_t)blockIdx.x * RL_TILE + (uint64_t)threadIdx.x * RL_PER, b);
    const uint32_t c = __reduce_add_sync(0xffffffffu, __popc(m));
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int i = 0; i < RL_THREADS / 32; i++) t += s_w[i];
        counts[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(RL_THREADS)
rle_write_kernel(const uint8_t *__restrict__ q, uint64_t n, const unsigned long long *__restrict__ base,
                 uint8_t *__restrict__ vals, uint32_t *__restrict__ pos)
{
    __shared__ uint32_t s_w[RL_THREADS / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint8_t b[RL_PER];
    const uint64_t i0 = (uint64_t)blockIdx.x * RL_TILE + (uint64_t)threadIdx.x * RL_PER;
    uint32_t m = rle_heads(q, n, i0, b);
    const uint32_t c = __popc(m);
    uint32_t x = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) s_w[w] = x;
    __syncthreads();
    uint32_t add = 0;
    for (int i = 0; i < w; i++) add += s_w[i];
    unsigned long long r = base[blockIdx.x] + add + x - c;        // index of this thread's first run
    while (m) {
        const int j = __ffs(m) - 1;
        m &= m - 1;
        vals[r] = b[j];
        pos[r] = (uint32_t)(i0 + j);
        r++;
    }
}

__global__ void __launch_bounds__(256)
rle_len_kernel(const uint32_t *__restrict__ pos, uint64_t nruns, uint64_t n, uint8_t *__restrict__ lens, uint32_t *__restrict__ bad)
{
    const uint64_t r = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (r >= nruns) return;
    const uint64_t len = (r + 1 < nruns ? (uint64_t)pos[r + 1] : n) - pos[r];
    if (len > 255) atomicMax(bad, (uint32_t)min(len, (uint64_t)0xFFFFFFFFu));
    lens[r] = (uint8_t)len;
}

// d_q: n stream bytes (device).  Returns the number of runs; d_vals / d_lens hold `capacity` bytes each.
uint64_t rle_encode_dev(bfq_ctx *c, const uint8_t *d_q, uint64_t n, uint8_t *d_vals, uint8_t *d_lens, uint64_t capacity)
{
    if (n == 0) return 0;
    if (n >= ((uint64_t)1 << 32)) BFQ_FAIL(BFQ_ETOOLARGE, "quality stream of %llu bytes; run positions are 32-bit", (unsigned long long)n);
    cudaStream_t st = c->stream;
    Arena::Mark mk = c->arena.mark();
    const uint32_t nblk = div_up(n, RL_TILE);
    uint32_t *counts = arena_alloc<uint32_t>(c, nblk + 4);
    unsigned long long *base = arena_alloc<unsigned long long>(c, nblk + 4);
    uint32_t *bad = arena_alloc<uint32_t>(c, 4);
    BFQ_CUDA(cudaMemsetAsync(bad, 0, 16, st));
    {
        ProfScope ps(c, PC_ASSEMBLE, n, 1);
        rle_count_kernel<<<nblk, RL_THREADS, 0, st>>>(d_q, n, counts);
        BFQ_CUDA(cudaGetLastError());
    }
    scan_u32_u64_dev(c, counts, nblk, base);
    fetch_small(c, c->h_small, base + nblk, 8);
    BFQ_CUDA(cudaStreamSynchronize(st));
    const uint64_t nruns = c->h_small[0];
    if (nruns > capacity) BFQ_FAIL(BFQ_ECAPACITY, "the stream has %llu runs, the output buffers hold %llu", (unsigned long long)nruns, (unsigned long long)capacity);
    uint32_t *pos = arena_alloc<uint32_t>(c, nruns + 4);
    {
        ProfScope ps(c, PC_ASSEMBLE, n + 10 * nruns, 2);
        rle_write_kernel<<<nblk, RL_THREADS, 0, st>>>(d_q, n, base, d_vals, pos);
        rle_len_kernel<<<div_up(nruns, 256), 256, 0, st>>>(pos, nruns, n, d_lens, bad);
        BFQ_CUDA(cudaGetLastError());
    }
    fetch_small(c, c->h_small, bad, 4);
    BFQ_CUDA(cudaStreamSynchronize(st));
    c->arena.release(mk);
    const uint32_t longest = ((const uint32_t *)c->h_small)[0];
    if (longest)
        BFQ_FAIL(BFQ_EINVAL, "a run of %u equal bytes: the transform is defined on newline-separated lines of at most 255 symbols "
                 "(run lengths are one byte)", longest);
    return nruns;
}

// ---------------------------------------------------------------- decode (the inverse, for the reader's side and the round-trip tests)
constexpr int RD_RUNS = 1024;                            // runs per block

__global__ void __launch_bounds__(256)
rle_sum_kernel(const uint8_t *__restrict__ lens, uint64_t nruns, uint32_t *__restrict__ sums)
{
    __shared__ uint32_t s_w[8];
    uint32_t t = 0;
    for (int k = 0; k < RD_RUNS / 256; k++) {
        const uint64_t r = (uint64_t)blockIdx.x * RD_RUNS + k * 256 + threadIdx.x;
        if (r < nruns) t += lens[r];
    }
    t = __reduce_add_sync(0xffffffffu, t);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) { uint32_t a = 0; for (int i = 0; i < 8; i++) a += s_w[i]; sums[blockIdx.x] = a; }
}

// thread t of a block owns RD_RUNS / 256 consecutive runs; the block scans their lengths and every run is written out
__global__ void __launch_bounds__(256)
rle_expand_kernel(const uint8_t *__restrict__ vals, const uint8_t *__restrict__ lens, uint64_t nruns,
                  const unsigned long long *__restrict__ base, uint8_t *__restrict__ out)
{
    constexpr int PER = RD_RUNS / 256;
    __shared__ uint32_t s_w[8];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint64_t r0 = (uint64_t)blockIdx.x * RD_RUNS + (uint64_t)threadIdx.x * PER;
    uint32_t l[PER], c = 0;
#pragma unroll
    for (int k = 0; k < PER; k++) { l[k] = r0 + k < nruns ? lens[r0 + k] : 0u; c += l[k]; }
    uint32_t x = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) s_w[w] = x;
    __syncthreads();
    uint32_t add = 0;
    for (int i = 0; i < w; i++) add += s_w[i];
    unsigned long long o = base[blockIdx.x] + add + x - c;
#pragma unroll
    for (int k = 0; k < PER; k++) {
        if (l[k]) {
            const uint8_t v = vals[r0 + k];
            for (uint32_t j = 0; j < l[k]; j++) out[o + j] = v;
            o += l[k];
        }
    }
}

// returns the stream length; d_out holds `capacity` bytes
uint64_t rle_decode_dev(bfq_ctx *c, const uint8_t *d_vals, const uint8_t *d_lens, uint64_t nruns, uint8_t *d_out, uint64_t capacity)
{
    if (nruns == 0) return 0;
    if (nruns >= ((uint64_t)1 << 32)) BFQ_FAIL(BFQ_ETOOLARGE, "%llu runs", (unsigned long long)nruns);
    cudaStream_t st = c->stream;
    Arena::Mark mk = c->arena.mark();
    const uint32_t nblk = div_up(nruns, RD_RUNS);
    uint32_t *sums = arena_alloc<uint32_t>(c, nblk + 4);
    unsigned long long *base = arena_alloc<unsigned long long>(c, nblk + 4);
    {
        ProfScope ps(c, PC_ASSEMBLE, nruns, 1);
        rle_sum_kernel<<<nblk, 256, 0, st>>>(d_lens, nruns, sums);
        BFQ_CUDA(cudaGetLastError());
    }
    scan_u32_u64_dev(c, sums, nblk, base);
    fetch_small(c, c->h_small, base + nblk, 8);
    BFQ_CUDA(cudaStreamSynchronize(st));
    const uint64_t n = c->h_small[0];
    if (n > capacity) BFQ_FAIL(BFQ_ECAPACITY, "the runs expand to %llu bytes, the output buffer holds %llu", (unsigned long long)n, (unsigned long long)capacity);
    {
        ProfScope ps(c, PC_ASSEMBLE, 2 * nruns + n, 1);
        rle_expand_kernel<<<nblk, 256, 0, st>>>(d_vals, d_lens, nruns, base, d_out);
        BFQ_CUDA(cudaGetLastError());
    }
    BFQ_CUDA(cudaStreamSynchronize(st));
    c->arena.release(mk);
    return n;
}

} // namespace bfq

using namespace bfq;

extern "C" {

int bfq_qs_rle_encode(bfq_ctx *ctx, const uint8_t *qs, uint64_t n, uint8_t *vals, uint8_t *lens, uint64_t capacity, uint64_t *nruns)
{
    Arena::Mark mk{0, 0}; bool marked = false;            // scratch is released on the error path too
    try {
        if (!ctx || !nruns || (n && (!qs || !vals || !lens))) BFQ_FAIL(BFQ_EINVAL, "bfq_qs_rle_encode: null pointer");
        *nruns = 0;
        BFQ_CUDA(cudaSetDevice(ctx->device));
        if (n == 0) return BFQ_OK;
        mk = ctx->arena.mark(); marked = true;
        uint8_t *d_q = arena_alloc<uint8_t>(ctx, n + 64);
        const uint64_t cap = std::min<uint64_t>(capacity, n);
        uint8_t *d_v = arena_alloc<uint8_t>(ctx, cap + 64), *d_l = arena_alloc<uint8_t>(ctx, cap + 64);
        BFQ_CUDA(cudaMemcpyAsync(d_q, qs, n, cudaMemcpyHostToDevice, ctx->stream));
        const uint64_t r = rle_encode_dev(ctx, d_q, n, d_v, d_l, cap);
        BFQ_CUDA(cudaMemcpyAsync(vals, d_v, r, cudaMemcpyDeviceToHost, ctx->stream));
        BFQ_CUDA(cudaMemcpyAsync(lens, d_l, r, cudaMemcpyDeviceToHost, ctx->stream));
        BFQ_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->arena.release(mk);
        *nruns = r;
        return BFQ_OK;
    } catch (bfq::Error &e) { if (ctx) { cudaStreamSynchronize(ctx->stream); if (marked) ctx->arena.release(mk); } cudaGetLastError(); return e.code; }
}

int bfq_qs_rle_decode(bfq_ctx *ctx, const uint8_t *vals, const uint8_t *lens, uint64_t nruns, uint8_t *qs, uint64_t capacity, uint64_t *n)
{
    Arena::Mark mk{0, 0}; bool marked = false;
    try {
        if (!ctx || !n || (nruns && (!vals || !lens || !qs))) BFQ_FAIL(BFQ_EINVAL, "bfq_qs_rle_decode: null pointer");
        *n = 0;
        BFQ_CUDA(cudaSetDevice(ctx->device));
        if (nruns == 0) return BFQ_OK;
        mk = ctx->arena.mark(); marked = true;
        uint8_t *d_v = arena_alloc<uint8_t>(ctx, nruns + 64), *d_l = arena_alloc<uint8_t>(ctx, nruns + 64);
        uint8_t *d_o = arena_alloc<uint8_t>(ctx, capacity + 64);
        BFQ_CUDA(cudaMemcpyAsync(d_v, vals, nruns, cudaMemcpyHostToDevice, ctx->stream));
        BFQ_CUDA(cudaMemcpyAsync(d_l, lens, nruns, cudaMemcpyHostToDevice, ctx->stream));
        const uint64_t len = rle_decode_dev(ctx, d_v, d_l, nruns, d_o, capacity);
        BFQ_CUDA(cudaMemcpyAsync(qs, d_o, len, cudaMemcpyDeviceToHost, ctx->stream));
        BFQ_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->arena.release(mk);
        *n = len;
        return BFQ_OK;
    } catch (bfq::Error &e) { if (ctx) { cudaStreamSynchronize(ctx->stream); if (marked) ctx->arena.release(mk); } cudaGetLastError(); return e.code; }
}

} // extern "C"
