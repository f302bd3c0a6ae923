// fastq.cu -- the FASTQ text on the device side of the pipe.
//
// Ingest replaces the host loaders of step 1 (external/gsufsort/lib/file.c:182-275: getline + malloc per read):
// the file contents go to the device as they are, a newline scan finds the lines, line 2 / line 4 of every 4-line
// record become the base and quality arrays, empty reads are dropped as cat_all drops them (lib/utils.c:229-236) and a
// trailing partial record is ignored.  Emit replaces the record writer of step 3
// (src_ext_mem/decode.cpp:1080-1127): "@\nBASES\n+\nQUALS\n" per read, assembled on the device.
#include "common.cuh"
#include <algorithm>

namespace bfq {

constexpr int FQ_THREADS = 256;
constexpr int FQ_BYTES = 64;                       // text bytes per thread
constexpr int FQ_CHUNK = FQ_THREADS * FQ_BYTES;    // 16 KiB per block

__device__ __forceinline__ uint64_t newline_mask64(const uint8_t *__restrict__ text, uint64_t base, uint64_t size)
{
    uint64_t m = 0;
    if (base + FQ_BYTES <= size && ((reinterpret_cast<uintptr_t>(text) + base) & 15) == 0) {
        const uint4 *v = reinterpret_cast<const uint4 *>(text + base);
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint4 x = __ldg(v + q);
            const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (int k = 0; k < 4; k++) {
                // byte == '\n' test without branches: zero bytes of (w ^ 0x0A0A0A0A)
                const uint32_t y = w[k] ^ 0x0A0A0A0Au;
                const uint32_t z = ~(((y & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | y | 0x7F7F7F7Fu);      // 0x80 in every zero byte
                const uint64_t bits = ((z >> 7) & 1u) | ((z >> 14) & 2u) | ((z >> 21) & 4u) | ((z >> 28) & 8u);
                m |= bits << (16 * q + 4 * k);
            }
        }
    } else {
        for (int j = 0; j < FQ_BYTES; j++)
            if (base + j < size && text[base + j] == '\n') m |= 1ull << j;
    }
    return m;
}

// pass 1: newlines per 16 KiB chunk; pass 2 (WRITE): their positions, in order
template <bool WRITE>
__global__ void __launch_bounds__(FQ_THREADS)
newline_kernel(const uint8_t *__restrict__ text, uint64_t size, uint32_t *__restrict__ chunk_count,
               const unsigned long long *__restrict__ chunk_base, uint64_t *__restrict__ nl)
{
    __shared__ uint32_t s_warp[FQ_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint64_t base = (uint64_t)blockIdx.x * FQ_CHUNK + (uint64_t)tid * FQ_BYTES;
    uint64_t m = base < size ? newline_mask64(text, base, size) : 0ull;
    const uint32_t mine = __popcll(m);
    uint32_t x = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) s_warp[w] = x;
    __syncthreads();
    uint32_t add = 0, total = 0;
    for (int i = 0; i < FQ_THREADS / 32; i++) { if (i < w) add += s_warp[i]; total += s_warp[i]; }
    if (!WRITE) {
        if (tid == 0) chunk_count[blockIdx.x] = total;
        return;
    }
    uint64_t dst = chunk_base[blockIdx.x] + add + x - mine;
    while (m) {
        const int j = __ffsll((long long)m) - 1;
        nl[dst++] = base + j;
        m &= m - 1;
    }
}

// per record: length of line 2 (0 = dropped) and whether line 4 has the same length
__global__ void __launch_bounds__(256)
record_len_kernel(const uint8_t *__restrict__ text, uint64_t size, const uint64_t *__restrict__ nl, uint64_t nlines, uint64_t nrec,
                  uint32_t *__restrict__ len, uint32_t *__restrict__ keep, unsigned long long *__restrict__ bad)
{
    const uint64_t g = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (g >= nrec) return;
    auto line_end = [&](uint64_t i) -> uint64_t { return i < nlines ? nl[i] : size; };      // the last line may lack its '\n'
    const uint64_t b0 = nl[4 * g] + 1, b1 = line_end(4 * g + 1);
    const uint64_t q0 = nl[4 * g + 2] + 1, q1 = line_end(4 * g + 3);
    uint64_t bl = b1 - b0, ql = q1 - q0;
    if (bl && text[b1 - 1] == '\r') bl--;
    if (ql && text[q1 - 1] == '\r') ql--;
    if (bl != ql) atomicMin(bad, (unsigned long long)g);
    len[g] = (uint32_t)bl;
    keep[g] = bl ? 1u : 0u;
}

// one warp per record: its two payload lines go to their place in the flat arrays
__global__ void __launch_bounds__(256)
record_copy_kernel(const uint8_t *__restrict__ text, const uint64_t *__restrict__ nl, uint64_t nrec,
                   const uint32_t *__restrict__ len, const unsigned long long *__restrict__ off_all,
                   const unsigned long long *__restrict__ idx_all, uint8_t *__restrict__ bases, uint8_t *__restrict__ quals,
                   uint64_t *__restrict__ off)
{
    const int lane = threadIdx.x & 31;
    const uint64_t warp0 = ((uint64_t)blockIdx.x * 256 + threadIdx.x) >> 5, nwarps = ((uint64_t)gridDim.x * 256) >> 5;
    for (uint64_t g = warp0; g < nrec; g += nwarps) {
        const uint32_t l = len[g];
        if (l == 0) continue;
        const uint64_t dst = off_all[g], b0 = nl[4 * g] + 1, q0 = nl[4 * g + 2] + 1;
        if (lane == 0) off[idx_all[g]] = dst;
        for (uint32_t j = lane; j < l; j += 32) { bases[dst + j] = text[b0 + j]; quals[dst + j] = text[q0 + j]; }
    }
    if (warp0 == 0 && lane == 0) off[idx_all[nrec]] = off_all[nrec];       // totals of the two scans: off[nreads] = bases
}

void fastq_ingest_dev(bfq_ctx *c, const uint8_t *d_text, uint64_t size, DeviceReads *out)
{
    cudaStream_t st = c->stream;
    *out = DeviceReads();
    if (size == 0) { out->off = arena_alloc<uint64_t>(c, 1); BFQ_CUDA(cudaMemsetAsync(out->off, 0, 8, st)); return; }
    const uint64_t nchunks64 = div_up(size, FQ_CHUNK);
    if (nchunks64 >= ((uint64_t)1 << 31)) BFQ_FAIL(BFQ_ETOOLARGE, "FASTQ text of %llu bytes", (unsigned long long)size);
    const uint32_t nchunks = (uint32_t)nchunks64;
    ProfScope ps(c, PC_PACK, 3 * size, 8);
    uint32_t *cc = arena_alloc<uint32_t>(c, nchunks + 4);
    unsigned long long *cb = arena_alloc<unsigned long long>(c, (size_t)nchunks + 4);
    newline_kernel<false><<<nchunks, FQ_THREADS, 0, st>>>(d_text, size, cc, nullptr, nullptr);
    scan_u32_u64_dev(c, cc, nchunks, cb);
    fetch_small(c, c->h_small, cb + nchunks, 8);
    // does the text end with a newline?  (else the last line ends at `size`)
    fetch_small(c, c->h_small + 1, d_text + ((size - 1) & ~(uint64_t)3), 4);
    BFQ_CUDA(cudaStreamSynchronize(st));
    const uint64_t nnl = c->h_small[0];
    const uint8_t last = ((const uint8_t *)(c->h_small + 1))[(size - 1) & 3];
    const uint64_t nlines_total = nnl + (last == '\n' ? 0 : 1);
    const uint64_t nrec = nlines_total / 4;
    if (nrec >= ((uint64_t)1 << 32) - 1) BFQ_FAIL(BFQ_ETOOLARGE, "%llu records", (unsigned long long)nrec);
    out->off = arena_alloc<uint64_t>(c, nrec + 2);
    if (nrec == 0) { BFQ_CUDA(cudaMemsetAsync(out->off, 0, 8, st)); return; }
    uint64_t *nl = arena_alloc<uint64_t>(c, nnl + 4);
    newline_kernel<true><<<nchunks, FQ_THREADS, 0, st>>>(d_text, size, nullptr, cb, nl);
    uint32_t *len = arena_alloc<uint32_t>(c, nrec + 4), *keep = arena_alloc<uint32_t>(c, nrec + 4);
    unsigned long long *off_all = arena_alloc<unsigned long long>(c, nrec + 4), *idx_all = arena_alloc<unsigned long long>(c, nrec + 4);
    unsigned long long *bad = arena_alloc<unsigned long long>(c, 4);
    BFQ_CUDA(cudaMemsetAsync(bad, 0xFF, 8, st));
    record_len_kernel<<<div_up(nrec, 256), 256, 0, st>>>(d_text, size, nl, nnl, nrec, len, keep, bad);
    scan_u32_u64_dev(c, len, nrec, off_all);
    scan_u32_u64_dev(c, keep, nrec, idx_all);
    fetch_small(c, c->h_small, off_all + nrec, 8);
    fetch_small(c, c->h_small + 1, idx_all + nrec, 8);
    fetch_small(c, c->h_small + 2, bad, 8);
    BFQ_CUDA(cudaStreamSynchronize(st));
    const uint64_t total = c->h_small[0], nreads = c->h_small[1], badrec = c->h_small[2];
    if (badrec != ~0ull)
        BFQ_FAIL(BFQ_EINVAL, "FASTQ record %llu: bases and qualities differ in length", (unsigned long long)badrec);
    out->bases = arena_alloc<uint8_t>(c, total + 64);
    out->quals = arena_alloc<uint8_t>(c, total + 64);
    out->nreads = nreads; out->total = total;
    int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 16, div_up(nrec * 32, 256));
    record_copy_kernel<<<std::max(grid, 1), 256, 0, st>>>(d_text, nl, nrec, len, off_all, idx_all, out->bases, out->quals, out->off);
    BFQ_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------- emit
// record g starts at 2 * off[g] + 6 * g: '@' '\n' bases '\n' '+' '\n' quals '\n'
__global__ void __launch_bounds__(256)
fastq_emit_kernel(const uint8_t *__restrict__ bases, const uint8_t *__restrict__ quals, const uint64_t *__restrict__ off,
                  uint64_t nreads, uint8_t *__restrict__ text, uint8_t *__restrict__ dna, uint8_t *__restrict__ qsl)
{
    const int lane = threadIdx.x & 31;
    const uint64_t warp0 = ((uint64_t)blockIdx.x * 256 + threadIdx.x) >> 5, nwarps = ((uint64_t)gridDim.x * 256) >> 5;
    for (uint64_t g = warp0; g < nreads; g += nwarps) {
        const uint64_t s = off[g], l = off[g + 1] - s;
        uint8_t *o = text + 2 * s + 6 * g;
        if (lane == 0) { o[0] = '@'; o[1] = '\n'; o[2 + l] = '\n'; o[3 + l] = '+'; o[4 + l] = '\n'; o[5 + 2 * l] = '\n'; }
        for (uint64_t j = lane; j < l; j += 32) { o[2 + j] = bases[s + j]; o[5 + l + j] = quals[s + j]; }
        if (dna) {
            // the component streams of step 4 (BFQzip_ext.py:224-244: sed -n 2~4p / 4~4p): one line per read
            for (uint64_t j = lane; j < l; j += 32) { dna[s + g + j] = bases[s + j]; qsl[s + g + j] = quals[s + j]; }
            if (lane == 0) { dna[s + g + l] = '\n'; qsl[s + g + l] = '\n'; }
        }
    }
}

void fastq_emit_dev(bfq_ctx *c, const uint8_t *d_bases, const uint8_t *d_quals, const uint64_t *d_off, uint64_t nreads,
                    uint64_t total, uint8_t *d_text, uint8_t *d_dna, uint8_t *d_qs)
{
    if (nreads == 0) return;
    if ((d_dna == nullptr) != (d_qs == nullptr)) BFQ_FAIL(BFQ_EINVAL, "the two component streams come together");
    ProfScope ps(c, PC_ASSEMBLE, 4 * total + 6 * nreads);
    int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 16, div_up(nreads * 32, 256));
    fastq_emit_kernel<<<std::max(grid, 1), 256, 0, c->stream>>>(d_bases, d_quals, d_off, nreads, d_text, d_dna, d_qs);
    BFQ_CUDA(cudaGetLastError());
}

} // namespace bfq
