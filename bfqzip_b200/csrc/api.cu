// api.cu -- the extern "C" boundary (include/bfq_gpu.h): argument checks, host<->device
// staging, stage orchestration.  All real work is in the CUDA kernels of the other files.
#include "common.cuh"
#include <algorithm>

namespace bfq {

void byte_histogram(bfq_ctx *c, const uint8_t *d_data, uint64_t n, uint64_t hist_out[256]);

static void begin_call(bfq_ctx *c)
{
    if (!c) BFQ_FAIL(BFQ_EINVAL, "null context");
    BFQ_CUDA(cudaSetDevice(c->device));
    c->arena.release(Arena::Mark{0, 0});
    c->has_last = false;
}

// min / max read length from device offsets (also rejects non-monotone offsets)
__global__ void __launch_bounds__(256)
len_range_kernel(const uint64_t *__restrict__ off, uint64_t nreads, unsigned long long *__restrict__ out /*[min,max,bad]*/)
{
    unsigned long long mn = ~0ull, mx = 0, bad = 0;
    for (uint64_t g = (uint64_t)blockIdx.x * 256 + threadIdx.x; g < nreads; g += (uint64_t)gridDim.x * 256) {
        unsigned long long a = off[g], b = off[g + 1];
        if (b < a) { bad = 1; continue; }
        mn = min(mn, b - a);
        mx = max(mx, b - a);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        bad |= __shfl_xor_sync(0xffffffffu, bad, o);
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(&out[0], mn); atomicMax(&out[1], mx); if (bad) atomicOr(&out[2], 1ull); }
}

static void check_reads_dev(bfq_ctx *c, const uint64_t *d_off, uint64_t nreads, uint64_t total_bases,
                            uint64_t *min_len, uint64_t *max_len)
{
    unsigned long long *d = (unsigned long long *)((char *)c->d_small + 3072);
    unsigned long long init[4] = {~0ull, 0, 0, 0};
    memcpy((char *)c->h_small + 3072, init, sizeof init);
    BFQ_CUDA(cudaMemcpyAsync(d, (char *)c->h_small + 3072, 32, cudaMemcpyHostToDevice, c->stream));
    if (nreads) {
        int grid = (int)std::min<uint64_t>((uint64_t)c->num_sms * 8, div_up(nreads, 256));
        len_range_kernel<<<grid, 256, 0, c->stream>>>(d_off, nreads, d);
        c->launches++;
    }
    fetch_small(c, (char *)c->h_small + 3072, d, 32);
    fetch_small(c, (char *)c->h_small + 3072 + 32, d_off + nreads, 8);
    fetch_small(c, (char *)c->h_small + 3072 + 40, d_off, 8);
    BFQ_CUDA(cudaStreamSynchronize(c->stream));
    const unsigned long long *h = (const unsigned long long *)((char *)c->h_small + 3072);
    if (h[2]) BFQ_FAIL(BFQ_EINVAL, "read offsets are not non-decreasing");
    if (h[5] != 0 || h[4] != total_bases) BFQ_FAIL(BFQ_EINVAL, "off[0] must be 0 and off[nreads] must equal the number of bases");
    if (nreads && h[0] == 0)
        BFQ_FAIL(BFQ_EINVAL, "empty reads are not supported: the reference drops them (external/gsufsort/lib/utils.c:229-236); remove them first");
    *min_len = nreads ? h[0] : 0;
    *max_len = h[1];
}

__global__ void __launch_bounds__(256)
export_rows_kernel(const uint8_t *__restrict__ lines, const uint8_t *__restrict__ obase, uint64_t n,
                   uint64_t code2byte, uint8_t *__restrict__ qs_out, uint8_t *__restrict__ ob_out,
                   const uint8_t *__restrict__ flat_bwt)
{
    const uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    if (flat_bwt) {                                   // flat mode: lines is the per-row stored quality
        const uint8_t q = lines[i];
        qs_out[i] = q & 0x7F;
        ob_out[i] = (q & 0x80) ? obase[i] : flat_bwt[i];
        return;
    }
    const uint32_t c = ws_code(ws_load(lines, i), (uint32_t)(i & 15));
    const uint8_t q = *ws_q(lines, i);
    qs_out[i] = q & 0x7F;
    ob_out[i] = (q & 0x80) ? obase[i] : (uint8_t)(code2byte >> (8 * c));
}

static void remember(bfq_ctx *c, const SmoothResult &S, const uint8_t *d_bwt = nullptr)
{
    c->last_flat_bwt = S.flat ? d_bwt : nullptr;
    c->last_lines = S.lines; c->last_obase = S.obase; c->last_n = S.n;
    memcpy(c->last_code2byte, S.alpha.code2byte, 8);
    c->has_last = true;
}

// fused device pipeline; all pointers device.  Returns through stats.
static void pipeline_dev(bfq_ctx *c, const uint8_t *d_bases, const uint8_t *d_quals, const uint64_t *d_off,
                         uint64_t nreads, uint64_t total_bases, const bfq_params *P, uint8_t *d_bwt,
                         uint8_t *d_qs, uint8_t *d_lcp, uint8_t *d_out_bases, uint8_t *d_out_quals,
                         bfq_stats *stats, uint8_t *h_bwt = nullptr, uint8_t *h_qs = nullptr, uint8_t *h_lcp = nullptr)
{
    uint64_t mn, mx;
    check_reads_dev(c, d_off, nreads, total_bases, &mn, &mx);
    const uint64_t N = total_bases + nreads;
    if (stats) memset(stats, 0, sizeof *stats);
    if (N == 0) return;
    // The fused calls return the reads under the CALLER's offsets.  The reference truncates a read longer than -l to its
    // last -l symbols (decode.cpp:605), which changes every later offset: that case is served by bfq_build_ebwt +
    // bfq_smooth_rebuild (which returns the new offsets); here it is refused instead of returning shifted reads.
    if (mx > (uint64_t)P->max_len)
        BFQ_FAIL(BFQ_EINVAL, "a read has %llu symbols but max_len (-l) is %d: the fused pipeline keeps the caller's offsets and cannot truncate; "
                 "raise max_len (<= 255) or use bfq_build_ebwt + bfq_smooth_rebuild", (unsigned long long)mx, P->max_len);
    if (!d_bwt) d_bwt = arena_alloc<uint8_t>(c, N + 64);
    if (!d_qs) d_qs = arena_alloc<uint8_t>(c, N + 64);
    if (!d_lcp) d_lcp = arena_alloc<uint8_t>(c, N + 64);
    // Flat mode (no walk lines, no LF walk): the group kernel also writes the suffix array and the packed text is
    // kept, so part B reads neighbours from the text and part C is a streaming sort back into text order.
    // auto: from flat_min_rows rows on; a device-resident call (no host sink: nothing leaves while the walk runs, so the walk's
    // overlap with the copies does not count) already from flat_min_rows_dev on, where the walk lines outgrow L2
    const uint64_t auto_min = (uint64_t)(c->sink ? c->opt_flat_min_rows : std::min(c->opt_flat_min_rows, c->opt_flat_min_rows_dev));
    const bool flat = c->opt_rebuild == 1 || (c->opt_rebuild == 0 && N >= auto_min);
    uint32_t *d_sa = flat ? arena_alloc<uint32_t>(c, N + 16) : nullptr;
    uint64_t *d_t3 = flat ? arena_alloc<uint64_t>(c, packed_text_words(N)) : nullptr;
    {
        Arena::Mark mk = c->arena.mark();
        build_ebwt_dev(c, d_bases, d_quals, d_off, nreads, total_bases, mx, P->eos_byte, d_bwt, d_qs, d_lcp, d_sa, d_t3);
        c->arena.release(mk);        // sort scratch is dead (stream-ordered) once the gather is queued
    }
    c->trace("part A done", c->stream);
    if (h_bwt || h_qs || h_lcp) {
        // the three files are final: copy them out on the side stream while parts B and C run
        BFQ_CUDA(cudaEventRecord(c->ev_a, c->stream));
        BFQ_CUDA(cudaStreamWaitEvent(c->copy_stream, c->ev_a, 0));
        if (!(c->sink && c->sink->files_copied)) {             // else: they left slice by slice during the group kernel
            if (h_bwt) BFQ_CUDA(cudaMemcpyAsync(h_bwt, d_bwt, N, cudaMemcpyDeviceToHost, c->copy_stream));
            if (h_qs) BFQ_CUDA(cudaMemcpyAsync(h_qs, d_qs, N, cudaMemcpyDeviceToHost, c->copy_stream));
            if (h_lcp) BFQ_CUDA(cudaMemcpyAsync(h_lcp, d_lcp, N, cudaMemcpyDeviceToHost, c->copy_stream));
        }
        BFQ_CUDA(cudaEventRecord(c->ev_copy, c->copy_stream));
        c->trace("files on host (copy stream)", c->copy_stream);
    }
    SmoothResult S;
    {
        // what part A knows about its own outputs: bwt = the reads' bases + one end-marker per read; qs = their qualities + the
        // end-marker byte at the same rows (main.c:848-867).  Saves two passes over 2 N bytes and two synchronisations.
        uint64_t bh[256], qh[256];
        const bool known = c->has_qual_hist;
        if (known) {
            memcpy(bh, c->base_hist, sizeof bh); memcpy(qh, c->qual_hist, sizeof qh);
            bh[P->eos_byte] += nreads; qh[P->eos_byte] += nreads;
        }
        smooth_dev(c, d_bwt, d_qs, d_lcp, N, P, &S, stats, d_sa, d_t3, nullptr, false, known ? bh : nullptr, known ? qh : nullptr);
    }
    c->trace("part B done", c->stream);
    if (S.nreads != nreads) BFQ_FAIL(BFQ_ECUDA, "internal error: %llu end-markers for %llu reads", (unsigned long long)S.nreads, (unsigned long long)nreads);
    if (flat) unpermute_dev(c, S, d_bwt, d_sa, d_off, nreads, d_out_bases, d_out_quals);
    else rebuild_dev(c, S, P, d_off, d_out_bases, d_out_quals, nullptr, nullptr);
    c->trace("part C done", c->stream);
    if (c->sink && c->sink->out_copied) { BFQ_CUDA(cudaEventRecord(c->ev_copy, c->copy_stream)); c->trace("reads on host (copy stream)", c->copy_stream); }
    if (h_bwt || h_qs || h_lcp || (c->sink && c->sink->out_copied))
        BFQ_CUDA(cudaStreamWaitEvent(c->stream, c->ev_copy, 0));                             // side copies join the main stream
    BFQ_CUDA(cudaStreamSynchronize(c->stream));
    remember(c, S, d_bwt);
}

} // namespace bfq

using namespace bfq;

#define API_BEGIN try { begin_call(ctx);
#define API_END   return BFQ_OK; } catch (bfq::Error &e__) { if (ctx) { cudaStreamSynchronize(ctx->stream); if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream); } cudaGetLastError(); return e__.code; } \
                  catch (std::exception &ex__) { bfq::set_error("exception: %s", ex__.what()); return BFQ_ENOMEM; }

static uint64_t host_check_off(const uint64_t *off, uint64_t nreads, uint64_t *max_len = nullptr)
{
    uint64_t mx = 0;
    if (off[0] != 0) BFQ_FAIL(BFQ_EINVAL, "off[0] must be 0");
    for (uint64_t g = 0; g < nreads; g++) {
        if (off[g + 1] < off[g]) BFQ_FAIL(BFQ_EINVAL, "read offsets are not non-decreasing (read %llu)", (unsigned long long)g);
        if (off[g + 1] == off[g])
            BFQ_FAIL(BFQ_EINVAL, "read %llu is empty: the reference drops empty reads (external/gsufsort/lib/utils.c:229-236); remove them first", (unsigned long long)g);
        mx = std::max(mx, off[g + 1] - off[g]);
    }
    if (max_len) *max_len = mx;
    return off[nreads];
}

template <typename T> static T *to_device(bfq_ctx *c, const T *h, uint64_t count, uint64_t pad = 16)
{
    T *d = arena_alloc<T>(c, count + pad);
    if (count) BFQ_CUDA(cudaMemcpyAsync(d, h, count * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    return d;
}

extern "C" {

int bfq_build_ebwt(bfq_ctx *ctx, const uint8_t *bases, const uint8_t *quals, const uint64_t *off,
                   uint64_t nreads, uint8_t eos_byte, uint8_t *bwt, uint8_t *qs, uint8_t *lcp)
{
    API_BEGIN
    if (!off || !bwt || !qs || !lcp || (nreads && (!bases || !quals))) BFQ_FAIL(BFQ_EINVAL, "bfq_build_ebwt: null pointer");
    uint64_t mxl = 0;
    const uint64_t total = host_check_off(off, nreads, &mxl), N = total + nreads;
    if (N == 0) return BFQ_OK;
    uint8_t *d_b = to_device(ctx, bases, total), *d_q = to_device(ctx, quals, total);
    uint64_t *d_o = to_device(ctx, off, nreads + 1);
    uint8_t *d_bwt = arena_alloc<uint8_t>(ctx, N + 64), *d_qs = arena_alloc<uint8_t>(ctx, N + 64), *d_lcp = arena_alloc<uint8_t>(ctx, N + 64);
    build_ebwt_dev(ctx, d_b, d_q, d_o, nreads, total, mxl, eos_byte, d_bwt, d_qs, d_lcp, nullptr);
    BFQ_CUDA(cudaMemcpyAsync(bwt, d_bwt, N, cudaMemcpyDeviceToHost, ctx->stream));
    BFQ_CUDA(cudaMemcpyAsync(qs, d_qs, N, cudaMemcpyDeviceToHost, ctx->stream));
    BFQ_CUDA(cudaMemcpyAsync(lcp, d_lcp, N, cudaMemcpyDeviceToHost, ctx->stream));
    BFQ_CUDA(cudaStreamSynchronize(ctx->stream));
    API_END
}

int bfq_smooth_rebuild(bfq_ctx *ctx, const uint8_t *bwt, const uint8_t *qs, const uint8_t *lcp, uint64_t n,
                       const bfq_params *params, uint8_t *out_bases, uint8_t *out_quals, uint64_t *out_off,
                       uint64_t off_capacity, bfq_stats *stats)
{
    API_BEGIN
    if (!params || !out_off || (n && (!bwt || !qs || !lcp || !out_bases || !out_quals))) BFQ_FAIL(BFQ_EINVAL, "bfq_smooth_rebuild: null pointer");
    if (stats) memset(stats, 0, sizeof *stats);
    if (n == 0) { if (off_capacity < 1) BFQ_FAIL(BFQ_ECAPACITY, "out_off too small"); out_off[0] = 0; return BFQ_OK; }
    uint8_t *d_bwt = to_device(ctx, bwt, n, 64), *d_qs = to_device(ctx, qs, n, 64), *d_lcp = to_device(ctx, lcp, n, 64);
    SmoothResult S;
    smooth_dev(ctx, d_bwt, d_qs, d_lcp, n, params, &S, stats);
    if (S.nreads + 1 > off_capacity) BFQ_FAIL(BFQ_ECAPACITY, "out_off holds %llu entries, %llu needed", (unsigned long long)off_capacity, (unsigned long long)(S.nreads + 1));
    uint8_t *d_ob = arena_alloc<uint8_t>(ctx, n + 64), *d_oq = arena_alloc<uint8_t>(ctx, n + 64);
    uint64_t *d_oo = arena_alloc<uint64_t>(ctx, S.nreads + 1);
    uint64_t total = 0;
    rebuild_dev(ctx, S, params, nullptr, d_ob, d_oq, d_oo, &total);
    if (total) {
        BFQ_CUDA(cudaMemcpyAsync(out_bases, d_ob, total, cudaMemcpyDeviceToHost, ctx->stream));
        BFQ_CUDA(cudaMemcpyAsync(out_quals, d_oq, total, cudaMemcpyDeviceToHost, ctx->stream));
    }
    BFQ_CUDA(cudaMemcpyAsync(out_off, d_oo, (S.nreads + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    BFQ_CUDA(cudaStreamSynchronize(ctx->stream));
    remember(ctx, S);
    API_END
}

int bfq_pipeline(bfq_ctx *ctx, const uint8_t *bases, const uint8_t *quals, const uint64_t *off, uint64_t nreads,
                 const bfq_params *params, uint8_t *bwt, uint8_t *qs, uint8_t *lcp, uint8_t *out_bases,
                 uint8_t *out_quals, bfq_stats *stats)
{
    API_BEGIN
    if (!params || !off || (nreads && (!bases || !quals || !out_bases || !out_quals))) BFQ_FAIL(BFQ_EINVAL, "bfq_pipeline: null pointer");
    if (stats) memset(stats, 0, sizeof *stats);
    if (nreads == 0) { host_check_off(off, nreads); return BFQ_OK; }
    // off[nreads] is the size the caller vouches for; the full validation of the offsets runs on the host while the
    // reads are already on their way to the device
    const uint64_t total = off[nreads], N = total + nreads;
    ctx->trace("call start", ctx->stream);
    uint8_t *d_b = to_device(ctx, bases, total), *d_q = to_device(ctx, quals, total);
    uint64_t *d_o = to_device(ctx, off, nreads + 1);
    ctx->trace("inputs on device", ctx->stream);
    host_check_off(off, nreads);
    uint8_t *d_bwt = arena_alloc<uint8_t>(ctx, N + 64), *d_qs = arena_alloc<uint8_t>(ctx, N + 64), *d_lcp = arena_alloc<uint8_t>(ctx, N + 64);
    uint8_t *d_ob = arena_alloc<uint8_t>(ctx, total + 64), *d_oq = arena_alloc<uint8_t>(ctx, total + 64);
    bfq_ctx::HostSink sink;
    sink.bwt = bwt; sink.qs = qs; sink.lcp = lcp; sink.out_b = out_bases; sink.out_q = out_quals; sink.off = off;
    struct SinkScope { bfq_ctx *c; ~SinkScope() { c->sink = nullptr; } } scope{ctx};
    ctx->sink = &sink;
    pipeline_dev(ctx, d_b, d_q, d_o, nreads, total, params, d_bwt, d_qs, d_lcp, d_ob, d_oq, stats, bwt, qs, lcp);
    if (!sink.out_copied) {
        BFQ_CUDA(cudaMemcpyAsync(out_bases, d_ob, total, cudaMemcpyDeviceToHost, ctx->stream));
        BFQ_CUDA(cudaMemcpyAsync(out_quals, d_oq, total, cudaMemcpyDeviceToHost, ctx->stream));
        BFQ_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    ctx->trace("call end", ctx->stream);
    ctx->trace_dump();
    API_END
}

// ---- FASTQ text in / FASTQ text out: parsing and record assembly on the device
int bfq_build_ebwt_fastq(bfq_ctx *ctx, const uint8_t *text, uint64_t size, uint8_t eos_byte, uint8_t *bwt, uint8_t *qs,
                         uint8_t *lcp, uint64_t capacity, uint64_t *nreads_out, uint64_t *n_out)
{
    API_BEGIN
    if (!nreads_out || !n_out || (size && (!text || !bwt || !qs || !lcp))) BFQ_FAIL(BFQ_EINVAL, "bfq_build_ebwt_fastq: null pointer");
    *nreads_out = 0; *n_out = 0;
    uint8_t *d_text = to_device(ctx, text, size, 64);
    DeviceReads R;
    fastq_ingest_dev(ctx, d_text, size, &R);
    const uint64_t N = R.total + R.nreads;
    *nreads_out = R.nreads; *n_out = N;
    if (N == 0) { BFQ_CUDA(cudaStreamSynchronize(ctx->stream)); return BFQ_OK; }
    if (N > capacity) BFQ_FAIL(BFQ_ECAPACITY, "the collection has %llu rows, the output buffers hold %llu", (unsigned long long)N, (unsigned long long)capacity);
    uint64_t mn, mx;
    check_reads_dev(ctx, R.off, R.nreads, R.total, &mn, &mx);
    uint8_t *d_bwt = arena_alloc<uint8_t>(ctx, N + 64), *d_qs = arena_alloc<uint8_t>(ctx, N + 64), *d_lcp = arena_alloc<uint8_t>(ctx, N + 64);
    build_ebwt_dev(ctx, R.bases, R.quals, R.off, R.nreads, R.total, mx, eos_byte, d_bwt, d_qs, d_lcp, nullptr);
    BFQ_CUDA(cudaMemcpyAsync(bwt, d_bwt, N, cudaMemcpyDeviceToHost, ctx->stream));
    BFQ_CUDA(cudaMemcpyAsync(qs, d_qs, N, cudaMemcpyDeviceToHost, ctx->stream));
    BFQ_CUDA(cudaMemcpyAsync(lcp, d_lcp, N, cudaMemcpyDeviceToHost, ctx->stream));
    BFQ_CUDA(cudaStreamSynchronize(ctx->stream));
    API_END
}

int bfq_smooth_rebuild_fastq(bfq_ctx *ctx, const uint8_t *bwt, const uint8_t *qs, const uint8_t *lcp, uint64_t n,
                             const bfq_params *params, uint8_t *fastq_out, uint64_t capacity, uint64_t *fastq_size,
                             uint8_t *dna_out, uint8_t *qs_out, bfq_stats *stats)
{
    API_BEGIN
    if (!params || !fastq_size || (n && (!bwt || !qs || !lcp || !fastq_out))) BFQ_FAIL(BFQ_EINVAL, "bfq_smooth_rebuild_fastq: null pointer");
    if (stats) memset(stats, 0, sizeof *stats);
    *fastq_size = 0;
    if (n == 0) return BFQ_OK;
    uint8_t *d_bwt = to_device(ctx, bwt, n, 64), *d_qs = to_device(ctx, qs, n, 64), *d_lcp = to_device(ctx, lcp, n, 64);
    SmoothResult S;
    smooth_dev(ctx, d_bwt, d_qs, d_lcp, n, params, &S, stats);
    uint8_t *d_ob = arena_alloc<uint8_t>(ctx, n + 64), *d_oq = arena_alloc<uint8_t>(ctx, n + 64);
    uint64_t *d_oo = arena_alloc<uint64_t>(ctx, S.nreads + 1);
    uint64_t total = 0;
    rebuild_dev(ctx, S, params, nullptr, d_ob, d_oq, d_oo, &total);
    const uint64_t fsz = 2 * total + 6 * S.nreads;
    if (fsz > capacity) BFQ_FAIL(BFQ_ECAPACITY, "the FASTQ text takes %llu bytes, the buffer holds %llu", (unsigned long long)fsz, (unsigned long long)capacity);
    if ((dna_out == nullptr) != (qs_out == nullptr)) BFQ_FAIL(BFQ_EINVAL, "bfq_smooth_rebuild_fastq: dna_out and qs_out come together");
    uint8_t *d_text = arena_alloc<uint8_t>(ctx, fsz + 64);
    uint8_t *d_dna = dna_out ? arena_alloc<uint8_t>(ctx, n + 64) : nullptr, *d_qsl = dna_out ? arena_alloc<uint8_t>(ctx, n + 64) : nullptr;
    fastq_emit_dev(ctx, d_ob, d_oq, d_oo, S.nreads, total, d_text, d_dna, d_qsl);
    if (fsz) BFQ_CUDA(cudaMemcpyAsync(fastq_out, d_text, fsz, cudaMemcpyDeviceToHost, ctx->stream));
    if (dna_out) {          // the rebuilt reads keep their lengths: total + nreads == n
        BFQ_CUDA(cudaMemcpyAsync(dna_out, d_dna, total + S.nreads, cudaMemcpyDeviceToHost, ctx->stream));
        BFQ_CUDA(cudaMemcpyAsync(qs_out, d_qsl, total + S.nreads, cudaMemcpyDeviceToHost, ctx->stream));
    }
    BFQ_CUDA(cudaStreamSynchronize(ctx->stream));
    *fastq_size = fsz;
    remember(ctx, S);
    API_END
}

int bfq_pipeline_fastq(bfq_ctx *ctx, const uint8_t *text, uint64_t size, const bfq_params *params, uint8_t *bwt, uint8_t *qs,
                       uint8_t *lcp, uint64_t capacity, uint8_t *fastq_out, uint64_t fastq_capacity, uint64_t *fastq_size,
                       uint64_t *nreads_out, uint64_t *n_out, bfq_stats *stats)
{
    API_BEGIN
    if (!params || !fastq_size || !nreads_out || !n_out || (size && (!text || !fastq_out))) BFQ_FAIL(BFQ_EINVAL, "bfq_pipeline_fastq: null pointer");
    if (stats) memset(stats, 0, sizeof *stats);
    *fastq_size = 0; *nreads_out = 0; *n_out = 0;
    uint8_t *d_text = to_device(ctx, text, size, 64);
    DeviceReads R;
    fastq_ingest_dev(ctx, d_text, size, &R);
    const uint64_t N = R.total + R.nreads, fsz = 2 * R.total + 6 * R.nreads;
    *nreads_out = R.nreads; *n_out = N;
    if (N == 0) { BFQ_CUDA(cudaStreamSynchronize(ctx->stream)); return BFQ_OK; }
    if ((bwt || qs || lcp) && N > capacity) BFQ_FAIL(BFQ_ECAPACITY, "the collection has %llu rows, the output buffers hold %llu", (unsigned long long)N, (unsigned long long)capacity);
    if (fsz > fastq_capacity) BFQ_FAIL(BFQ_ECAPACITY, "the FASTQ text takes %llu bytes, the buffer holds %llu", (unsigned long long)fsz, (unsigned long long)fastq_capacity);
    uint8_t *d_ob = arena_alloc<uint8_t>(ctx, R.total + 64), *d_oq = arena_alloc<uint8_t>(ctx, R.total + 64);
    // the input text is dead once the reads are parsed, but it sits below them on the arena stack: the output text gets its own block
    uint8_t *d_out = arena_alloc<uint8_t>(ctx, fsz + 64);
    pipeline_dev(ctx, R.bases, R.quals, R.off, R.nreads, R.total, params, nullptr, nullptr, nullptr, d_ob, d_oq, stats, bwt, qs, lcp);
    fastq_emit_dev(ctx, d_ob, d_oq, R.off, R.nreads, R.total, d_out);
    BFQ_CUDA(cudaMemcpyAsync(fastq_out, d_out, fsz, cudaMemcpyDeviceToHost, ctx->stream));
    BFQ_CUDA(cudaStreamSynchronize(ctx->stream));
    *fastq_size = fsz;
    API_END
}

int bfq_pipeline_dev(bfq_ctx *ctx, const uint8_t *d_bases, const uint8_t *d_quals, const uint64_t *d_off,
                     uint64_t nreads, uint64_t total_bases, const bfq_params *params, uint8_t *d_bwt,
                     uint8_t *d_qs, uint8_t *d_lcp, uint8_t *d_out_bases, uint8_t *d_out_quals, bfq_stats *stats)
{
    API_BEGIN
    if (!params || !d_off || (nreads && (!d_bases || !d_quals || !d_out_bases || !d_out_quals))) BFQ_FAIL(BFQ_EINVAL, "bfq_pipeline_dev: null pointer");
    pipeline_dev(ctx, d_bases, d_quals, d_off, nreads, total_bases, params, d_bwt, d_qs, d_lcp, d_out_bases, d_out_quals, stats);
    API_END
}

int bfq_dist_build(bfq_ctx *ctx, const uint8_t *d_bases, const uint8_t *d_quals, const uint64_t *d_off, uint64_t nreads,
                   uint64_t total_bases, const bfq_params *params, int rank, int world, uint64_t capacity,
                   const uint64_t *text_bounds, uint8_t *d_bwt, uint8_t *d_qs, uint8_t *d_lcp, uint32_t *d_keys,
                   uint16_t *d_vals, uint64_t *n_local, uint64_t *row_base, uint64_t *send_counts, bfq_stats *stats)
{
    API_BEGIN
    if (!params || !d_off || !d_bases || !d_quals || !text_bounds || !d_bwt || !d_qs || !d_lcp || !d_keys || !d_vals || !n_local ||
        !row_base || !send_counts)
        BFQ_FAIL(BFQ_EINVAL, "bfq_dist_build: null pointer");
    if (world < 1 || rank < 0 || rank >= world || world > 255) BFQ_FAIL(BFQ_EINVAL, "bfq_dist_build: rank %d of %d", rank, world);
    // rows of different ranks never share a cluster only if a cluster needs more context than the range cut (4 symbols)
    if (params->K < 4) BFQ_FAIL(BFQ_EINVAL, "single-collection mode over several GPUs needs -k >= 4 (ranges are cut on 4 symbols)");
    bfq_ctx *c = ctx;
    uint64_t mn, mx;
    c->trace("dist_build: call start", c->stream);
    check_reads_dev(c, d_off, nreads, total_bases, &mn, &mx);
    c->trace("dist_build: reads checked", c->stream);
    if (mx > (uint64_t)params->max_len) BFQ_FAIL(BFQ_EINVAL, "a read has %llu symbols, -l is %d", (unsigned long long)mx, params->max_len);
    const uint64_t N = total_bases + nreads;
    if (stats) memset(stats, 0, sizeof *stats);
    *n_local = 0; *row_base = 0;
    for (int d = 0; d < world; d++) send_counts[d] = 0;
    if (N == 0) return BFQ_OK;
    KeyRange R;
    R.rank = rank; R.world = world; R.capacity = capacity;
    // text positions take 64 bits from 2^32 rows on; the rows of one rank stay 32-bit indexed
    const bool wide = c->opt_wide_positions || N >= ((uint64_t)1 << 32) - 64;
    if (capacity >= ((uint64_t)1 << 32) - 64) BFQ_FAIL(BFQ_ETOOLARGE, "capacity of %llu rows; one GPU indexes its rows with 32 bits", (unsigned long long)capacity);
    void *d_sa = wide ? (void *)arena_alloc<uint64_t>(c, capacity + 16) : (void *)arena_alloc<uint32_t>(c, capacity + 16);
    uint64_t *d_t3 = arena_alloc<uint64_t>(c, packed_text_words(N));
    {
        Arena::Mark mk = c->arena.mark();
        if (wide) build_ebwt_dev_wide(c, d_bases, d_quals, d_off, nreads, total_bases, mx, params->eos_byte, d_bwt, d_qs, d_lcp, (uint64_t *)d_sa, d_t3, &R);
        else build_ebwt_dev(c, d_bases, d_quals, d_off, nreads, total_bases, mx, params->eos_byte, d_bwt, d_qs, d_lcp, (uint32_t *)d_sa, d_t3, &R);
        c->arena.release(mk);
    }
    *n_local = R.n_local; *row_base = R.row_base;
    c->trace("dist_build: slice of the eBWT built", c->stream);
    if (R.n_local) {
        uint64_t ah[256];
        memcpy(ah, c->base_hist, sizeof ah);
        ah[params->eos_byte] += nreads;
        SmoothResult S;
        smooth_dev(c, d_bwt, d_qs, d_lcp, R.n_local, params, &S, stats, d_sa, d_t3, ah, wide);
        c->trace("dist_build: slice smoothed", c->stream);
        emit_sorted_pairs_dev(c, S, d_bwt, d_sa, wide, N, text_bounds, world, d_keys, d_vals, send_counts);
        c->trace("dist_build: pairs grouped by destination", c->stream);
    }
    if (stats) { stats->n = N; stats->nreads = nreads; }
    BFQ_CUDA(cudaStreamSynchronize(c->stream));
    c->trace_dump();
    API_END
}

int bfq_dist_place(bfq_ctx *ctx, uint32_t *d_keys, uint16_t *d_vals, uint64_t count, uint64_t text_lo, const uint64_t *d_off,
                   uint64_t read_lo, uint64_t read_hi, uint8_t *d_out_bases, uint8_t *d_out_quals)
{
    API_BEGIN
    if (count && (!d_keys || !d_vals || !d_off || !d_out_bases || !d_out_quals)) BFQ_FAIL(BFQ_EINVAL, "bfq_dist_place: null pointer");
    place_pairs_dev(ctx, d_keys, d_vals, count, text_lo, d_off, read_lo, read_hi, d_out_bases, d_out_quals);
    API_END
}

int bfq_export_rows(bfq_ctx *ctx, uint8_t *qs_out, uint8_t *obwt_out)
{
    try {
        if (!ctx || !qs_out || !obwt_out) BFQ_FAIL(BFQ_EINVAL, "bfq_export_rows: null pointer");
        if (!ctx->has_last) BFQ_FAIL(BFQ_EINVAL, "bfq_export_rows: no smooth/pipeline result is held by this context");
        BFQ_CUDA(cudaSetDevice(ctx->device));
        const uint64_t n = ctx->last_n;
        Arena::Mark mk = ctx->arena.mark();
        uint8_t *dq = arena_alloc<uint8_t>(ctx, n + 64), *db = arena_alloc<uint8_t>(ctx, n + 64);
        uint64_t c2b = 0;
        for (int k = 0; k < 8; k++) c2b |= (uint64_t)ctx->last_code2byte[k] << (8 * k);
        export_rows_kernel<<<div_up(n, 256), 256, 0, ctx->stream>>>(ctx->last_lines, ctx->last_obase, n, c2b, dq, db, ctx->last_flat_bwt);
        ctx->launches++;
        BFQ_CUDA(cudaMemcpyAsync(qs_out, dq, n, cudaMemcpyDeviceToHost, ctx->stream));
        BFQ_CUDA(cudaMemcpyAsync(obwt_out, db, n, cudaMemcpyDeviceToHost, ctx->stream));
        BFQ_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->arena.release(mk);
        return BFQ_OK;
    } catch (bfq::Error &e) { return e.code; }
}

int bfq_debug_sort_pairs(bfq_ctx *ctx, uint64_t *keys, uint32_t *vals, uint64_t n, int bits)
{
    API_BEGIN
    if (!keys || !vals) BFQ_FAIL(BFQ_EINVAL, "null pointer");
    if (bits < 0 || bits > 64) BFQ_FAIL(BFQ_EINVAL, "bits must be 0..64");
    uint64_t *ka = to_device(ctx, keys, n), *kb = arena_alloc<uint64_t>(ctx, n + 16);
    uint32_t *va = to_device(ctx, vals, n), *vb = arena_alloc<uint32_t>(ctx, n + 16);
    int flip = radix_sort_pairs64(ctx, ka, kb, va, vb, n, bits);
    BFQ_CUDA(cudaMemcpyAsync(keys, flip ? kb : ka, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    BFQ_CUDA(cudaMemcpyAsync(vals, flip ? vb : va, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    BFQ_CUDA(cudaStreamSynchronize(ctx->stream));
    API_END
}

int bfq_debug_suffix_array(bfq_ctx *ctx, const uint8_t *bases, const uint64_t *off, uint64_t nreads, uint32_t *sa)
{
    API_BEGIN
    if (!off || !sa || (nreads && !bases)) BFQ_FAIL(BFQ_EINVAL, "null pointer");
    uint64_t mxl = 0;
    const uint64_t total = host_check_off(off, nreads, &mxl), N = total + nreads;
    if (N == 0) return BFQ_OK;
    uint8_t *d_b = to_device(ctx, bases, total);
    uint64_t *d_o = to_device(ctx, off, nreads + 1);
    uint32_t *d_sa = arena_alloc<uint32_t>(ctx, N + 16);
    build_ebwt_dev(ctx, d_b, d_b, d_o, nreads, total, mxl, '#', nullptr, nullptr, nullptr, d_sa);
    BFQ_CUDA(cudaMemcpyAsync(sa, d_sa, N * 4, cudaMemcpyDeviceToHost, ctx->stream));
    BFQ_CUDA(cudaStreamSynchronize(ctx->stream));
    API_END
}

} // extern "C"
