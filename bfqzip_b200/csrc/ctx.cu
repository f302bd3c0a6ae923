// ctx.cu -- context lifetime, device arena, error string, profiling scopes.
#include "common.cuh"
#include <stdarg.h>
#include <algorithm>

namespace bfq {

L2Pin::L2Pin(bfq_ctx *c, const void *ptr, size_t bytes)
{
    if (!bytes || c->l2_persist_max == 0 || c->l2_window_max == 0) return;
    const size_t want = std::min(c->l2_persist_max, bytes + (bytes >> 3));
    if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) != cudaSuccess) { cudaGetLastError(); return; }
    cudaStreamAttrValue av;
    memset(&av, 0, sizeof av);
    av.accessPolicyWindow.base_ptr = const_cast<void *>(ptr);
    av.accessPolicyWindow.num_bytes = std::min(bytes, c->l2_window_max);
    av.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)want / (double)av.accessPolicyWindow.num_bytes);
    av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    st = c->stream;
    on = cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &av) == cudaSuccess;
    if (!on) cudaGetLastError();
}

L2Pin::~L2Pin()
{
    if (!on) return;
    cudaStreamAttrValue av;
    memset(&av, 0, sizeof av);
    cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &av);
    cudaCtxResetPersistingL2Cache();
    cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, 0);   // give the set-aside back
}

__global__ void fetch_small_kernel(uint32_t *__restrict__ host_dst, const uint32_t *__restrict__ src, uint32_t words)
{
    for (uint32_t i = threadIdx.x; i < words; i += blockDim.x) host_dst[i] = src[i];
}

void fetch_small(bfq_ctx *c, void *host_dst, const void *dev_src, size_t bytes)
{
    if (!c->h_small_dev || (bytes & 3) || (char *)host_dst < (char *)c->h_small || (char *)host_dst + bytes > (char *)c->h_small + 4096) {
        BFQ_CUDA(cudaMemcpyAsync(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost, c->stream));
        return;
    }
    uint32_t *dst = (uint32_t *)((char *)c->h_small_dev + ((char *)host_dst - (char *)c->h_small));
    fetch_small_kernel<<<1, 256, 0, c->stream>>>(dst, (const uint32_t *)dev_src, (uint32_t)(bytes / 4));
    c->launches++;
    BFQ_CUDA(cudaGetLastError());
}

static thread_local char g_err[1024] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}
const char *get_error() { return g_err; }

// ------------------------------------------------------------------------- arena
void *Arena::alloc(size_t bytes)
{
    bytes = (bytes + 255) & ~(size_t)255;
    if (bytes == 0) bytes = 256;
    // first slab (from the current top) with room
    for (size_t i = 0; i < slabs.size(); i++) {
        Slab &s = slabs[i];
        bool later_used = false;
        for (size_t j = i + 1; j < slabs.size(); j++) if (slabs[j].used) later_used = true;
        if (later_used) continue;                 // keep stack discipline across slabs
        if (s.size - s.used >= bytes) {
            void *p = s.base + s.used;
            s.used += bytes;
            return p;
        }
    }
    size_t sz = std::max(bytes, min_slab);
    char *base = nullptr;
    cudaError_t e = cudaMalloc(&base, sz);
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("device arena: cudaMalloc(%zu) failed: %s (arena holds %zu bytes)", sz,
                  cudaGetErrorString(e), total());
        throw Error{BFQ_ENOMEM};
    }
    slabs.push_back(Slab{base, sz, bytes});
    return base;
}

Arena::Mark Arena::mark() const
{
    // top = last slab with used > 0
    for (size_t i = slabs.size(); i-- > 0;)
        if (slabs[i].used) return Mark{i, slabs[i].used};
    return Mark{0, 0};
}

void Arena::release(Mark m)
{
    for (size_t i = 0; i < slabs.size(); i++) {
        if (i > m.slab) slabs[i].used = 0;
        else if (i == m.slab) slabs[i].used = std::min(slabs[i].used, m.used);
    }
}

void Arena::destroy()
{
    for (auto &s : slabs) cudaFree(s.base);
    slabs.clear();
}

// ------------------------------------------------------------------------- profiling
static cudaEvent_t get_event(bfq_ctx *c)
{
    if (!c->event_pool.empty()) {
        cudaEvent_t e = c->event_pool.back();
        c->event_pool.pop_back();
        return e;
    }
    cudaEvent_t e;
    BFQ_CUDA(cudaEventCreate(&e));
    return e;
}

ProfScope::ProfScope(bfq_ctx *ctx, ProfClass klass, uint64_t bytes, uint64_t nlaunch) : c(ctx), k(klass)
{
    c->launches += nlaunch;
    c->prof[k].launches += nlaunch;
    c->prof[k].bytes += bytes;
    if (c->opt_profile) {
        e0 = get_event(c);
        e1 = get_event(c);
        cudaEventRecord(e0, c->stream);
    }
}
ProfScope::~ProfScope()
{
    if (e0) {
        cudaEventRecord(e1, c->stream);
        c->prof[k].pending.push_back({e0, e1});
    }
}

void prof_collect(bfq_ctx *c)
{
    for (int k = 0; k < PC_COUNT; k++) {
        for (auto &pr : c->prof[k].pending) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess) c->prof[k].ms += ms;
            c->event_pool.push_back(pr.first);
            c->event_pool.push_back(pr.second);
        }
        c->prof[k].pending.clear();
    }
}

} // namespace bfq

using namespace bfq;

extern "C" {

int bfq_mode_M(void) { return BFQ_M; }
int bfq_mode_B(void) { return BFQ_B; }
const char *bfq_last_error(void) { return bfq::get_error(); }

void bfq_params_default(bfq_params *p)
{
    // bfq_ext.cpp:69-96, 1026-1031
    p->K = 16; p->m = 2; p->qthr = 20; p->max_len = 250; p->fthr = 40.0f;
    p->eos_byte = '#'; p->default_q = '>';
}

int bfq_ctx_create(bfq_ctx **out, int device)
{
    if (!out) { set_error("bfq_ctx_create: null out"); return BFQ_EINVAL; }
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        set_error("no CUDA device available (%s); this library has no CPU path",
                  e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
        return BFQ_ENODEVICE;
    }
    if (device < 0 || device >= ndev) { set_error("device %d out of range (%d devices)", device, ndev); return BFQ_EINVAL; }
    bfq_ctx *c = new bfq_ctx();
    try {
        c->device = device;
        BFQ_CUDA(cudaSetDevice(device));
        cudaDeviceProp prop;
        BFQ_CUDA(cudaGetDeviceProperties(&prop, device));
        if (prop.major < 10) {
            set_error("device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
            delete c;
            return BFQ_ENODEVICE;
        }
        c->num_sms = prop.multiProcessorCount;
        c->l2_persist_max = prop.persistingL2CacheMaxSize > 0 ? (size_t)prop.persistingL2CacheMaxSize : 0;
        c->l2_window_max = prop.accessPolicyMaxWindowSize > 0 ? (size_t)prop.accessPolicyMaxWindowSize : 0;
        BFQ_CUDA(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
        c->stream = c->own_stream;
        BFQ_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
        BFQ_CUDA(cudaEventCreateWithFlags(&c->ev_a, cudaEventDisableTiming));
        BFQ_CUDA(cudaEventCreateWithFlags(&c->ev_copy, cudaEventDisableTiming));
        BFQ_CUDA(cudaHostAlloc(&c->h_small, 4096, cudaHostAllocMapped));
        if (cudaHostGetDevicePointer(&c->h_small_dev, c->h_small, 0) != cudaSuccess) { cudaGetLastError(); c->h_small_dev = nullptr; }
        BFQ_CUDA(cudaMalloc(&c->d_small, 4096));
    } catch (Error &er) {
        delete c;
        return er.code == BFQ_ECUDA ? BFQ_ENODEVICE : er.code;
    }
    *out = c;
    return BFQ_OK;
}

void bfq_ctx_destroy(bfq_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    prof_collect(c);
    for (auto e : c->event_pool) cudaEventDestroy(e);
    c->arena.destroy();
    if (c->h_small) cudaFreeHost(c->h_small);
    if (c->d_small) cudaFree(c->d_small);
    if (c->ev_a) cudaEventDestroy(c->ev_a);
    if (c->ev_copy) cudaEventDestroy(c->ev_copy);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

int bfq_ctx_set_option(bfq_ctx *c, const char *key, int64_t value)
{
    if (!c || !key) { set_error("set_option: null argument"); return BFQ_EINVAL; }
    std::string k(key);
    if (k == "initial_k") { if (value < 0 || value > 21) { set_error("initial_k must be 0..21"); return BFQ_EINVAL; } c->opt_initial_k = (int)value; }
    else if (k == "long_cluster") { if (value < 4) { set_error("long_cluster must be >= 4"); return BFQ_EINVAL; } c->opt_long_cluster = value; }
    else if (k == "group_cap") { if (value < 1) { set_error("group_cap must be >= 1"); return BFQ_EINVAL; } c->opt_group_cap = value; }
    else if (k == "batch_min_groups") { if (value < 0) { set_error("batch_min_groups must be >= 0"); return BFQ_EINVAL; } c->opt_batch_min_groups = value; }
    else if (k == "rebuild") { if (value < 0 || value > 2) { set_error("rebuild must be 0 (auto), 1 (un-permute) or 2 (LF walk)"); return BFQ_EINVAL; } c->opt_rebuild = (int)value; }
    else if (k == "flat_min_rows") { if (value < 0) { set_error("flat_min_rows must be >= 0"); return BFQ_EINVAL; } c->opt_flat_min_rows = value; }
    else if (k == "flat_min_rows_dev") { if (value < 0) { set_error("flat_min_rows_dev must be >= 0"); return BFQ_EINVAL; } c->opt_flat_min_rows_dev = value; }
    else if (k == "r11_cap") { if (value < 0) { set_error("r11_cap must be >= 0"); return BFQ_EINVAL; } c->opt_r11_cap = value; }
    else if (k == "bulk_tile") c->opt_bulk_tile = value ? 1 : 0;
    else if (k == "profile") c->opt_profile = value ? 1 : 0;
    else if (k == "trace") c->opt_trace = value ? 1 : 0;
    else if (k == "wide_positions") c->opt_wide_positions = value ? 1 : 0;
    else if (k == "host_slices") c->opt_host_slices = (int)std::max<int64_t>(1, std::min<int64_t>(value, 64));
    else { set_error("unknown option '%s'", key); return BFQ_EINVAL; }
    return BFQ_OK;
}

int bfq_ctx_set_stream(bfq_ctx *c, void *cuda_stream)
{
    if (!c) { set_error("set_stream: null context"); return BFQ_EINVAL; }
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    prof_collect(c);
    c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
    return BFQ_OK;
}

int bfq_prof_get(bfq_ctx *c, const char *klass, double *ms, uint64_t *launches, uint64_t *bytes)
{
    if (!c || !klass) { set_error("prof_get: null argument"); return BFQ_EINVAL; }
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    prof_collect(c);
    for (int k = 0; k < PC_COUNT; k++)
        if (!strcmp(klass, kProfNames[k])) {
            if (ms) *ms = c->prof[k].ms;
            if (launches) *launches = c->prof[k].launches;
            if (bytes) *bytes = c->prof[k].bytes;
            return BFQ_OK;
        }
    set_error("unknown profile class '%s'", klass);
    return BFQ_EINVAL;
}

void bfq_prof_reset(bfq_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    prof_collect(c);
    for (int k = 0; k < PC_COUNT; k++) { c->prof[k].ms = 0; c->prof[k].launches = 0; c->prof[k].bytes = 0; }
}

uint64_t bfq_launch_count(bfq_ctx *c) { return c ? c->launches : 0; }

} // extern "C"
