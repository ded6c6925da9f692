// ctx.cu -- context lifecycle, arena, staging, stage timing, raw memory helpers (include/pcp_cuda.h).
#include "common.cuh"
#include <cstring>
#include <cstdlib>
#include <algorithm>
#include <cstdarg>

static thread_local std::string g_last_error;

void pcp_set_global_error(const char* msg) { g_last_error = msg ? msg : ""; }

int pcp_fail(pcp_ctx* ctx, int code, const char* fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf;
    g_last_error = buf;
    return code;
}

int pcp_arena_alloc(pcp_ctx* ctx, size_t bytes, void** out)
{
    *out = nullptr;
    cudaError_t e = cudaMallocFromPoolAsync(out, bytes, ctx->pool, ctx->stream);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return pcp_fail(ctx, e == cudaErrorMemoryAllocation ? PCP_ERR_NOMEM : PCP_ERR_CUDA,
                        "device allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
    }
    return PCP_OK;
}

void pcp_arena_free(pcp_ctx* ctx, void* p)
{
    if (p) cudaFreeAsync(p, ctx->stream);
}

int pcp_scratch_get(pcp_ctx* ctx, size_t bytes, void** out)
{
    if (bytes > ctx->scratch_bytes) {
        if (ctx->scratch) {
            cudaStreamSynchronize(ctx->stream);
            cudaFree(ctx->scratch);
            ctx->scratch = nullptr, ctx->scratch_bytes = 0;
        }
        const size_t want = bytes + bytes / 8;
        cudaError_t e = cudaMalloc(&ctx->scratch, want);
        if (e != cudaSuccess) {
            cudaGetLastError();
            ctx->scratch = nullptr;
            return pcp_fail(ctx, e == cudaErrorMemoryAllocation ? PCP_ERR_NOMEM : PCP_ERR_CUDA,
                            "device allocation of %zu bytes failed: %s", want, cudaGetErrorString(e));
        }
        ctx->scratch_bytes = want;
    }
    *out = ctx->scratch;
    return PCP_OK;
}

pcp_ptr_kind pcp_classify(pcp_ctx* ctx, const void* p)
{
    cudaPointerAttributes at;
    cudaError_t e = cudaPointerGetAttributes(&at, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return PCP_PTR_HOST; // very old drivers report unregistered host memory as an error
    }
    if (at.type == cudaMemoryTypeDevice) return at.device == ctx->device ? PCP_PTR_DEVICE : PCP_PTR_BAD;
    if (at.type == cudaMemoryTypeManaged) return PCP_PTR_DEVICE;
    return PCP_PTR_HOST;
}

int InStage::open(pcp_ctx* c, const void* p, size_t bytes)
{
    ctx = c;
    if (!p || bytes == 0) {
        dptr = p;
        return PCP_OK;
    }
    pcp_ptr_kind k = pcp_classify(c, p);
    if (k == PCP_PTR_BAD) return pcp_fail(c, PCP_ERR_INVALID, "pointer belongs to another device");
    if (k == PCP_PTR_DEVICE) {
        dptr = p;
        return PCP_OK;
    }
    PCP_TRY(pcp_arena_alloc(c, bytes, &owned));
    // small inputs (a few coordinates, offsets) do not take a place in the copy engine's queue, where a bulk upload on
    // a copy lane may be ahead of them
    if (bytes <= PCP_READBACK_MAX) PCP_TRY(pcp_writeto(c, owned, p, bytes));
    else PCP_CUDA_TRY(c, cudaMemcpyAsync(owned, p, bytes, cudaMemcpyHostToDevice, c->stream));
    dptr = owned;
    return PCP_OK;
}

int OutStage::open(pcp_ctx* c, void* p, size_t nbytes, bool preload)
{
    ctx = c;
    user = p;
    bytes = nbytes;
    if (!p || nbytes == 0) {
        dptr = p;
        return PCP_OK;
    }
    pcp_ptr_kind k = pcp_classify(c, p);
    if (k == PCP_PTR_BAD) return pcp_fail(c, PCP_ERR_INVALID, "pointer belongs to another device");
    if (k == PCP_PTR_DEVICE) {
        dptr = p;
        return PCP_OK;
    }
    PCP_TRY(pcp_arena_alloc(c, nbytes, &owned));
    if (preload) {
        if (nbytes <= PCP_READBACK_MAX) PCP_TRY(pcp_writeto(c, owned, p, nbytes));
        else PCP_CUDA_TRY(c, cudaMemcpyAsync(owned, p, nbytes, cudaMemcpyHostToDevice, c->stream));
    }
    dptr = owned;
    return PCP_OK;
}

int OutStage::commit(size_t nbytes_override)
{
    if (!owned || !user) return PCP_OK;
    size_t nb = nbytes_override == (size_t)-1 ? bytes : nbytes_override;
    if (nb > bytes) nb = bytes;
    if (nb && nb <= PCP_READBACK_MAX) return pcp_readback(ctx, user, owned, nb); // (waits)
    if (nb) PCP_CUDA_TRY(ctx, cudaMemcpyAsync(user, owned, nb, cudaMemcpyDeviceToHost, ctx->stream));
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

// ---- stage timing --------------------------------------------------------------------------------------
static cudaEvent_t take_event(pcp_ctx* ctx)
{
    if (!ctx->ev_free.empty()) {
        cudaEvent_t e = ctx->ev_free.back();
        ctx->ev_free.pop_back();
        return e;
    }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}

StageTimer::StageTimer(pcp_ctx* c, const char* name, double bytes) : ctx(c)
{
    if (!c->profile) return;
    for (size_t i = 0; i < c->stages.size(); i++)
        if (c->stages[i].name == name) stage = (int)i;
    if (stage < 0) {
        pcp_stage_rec r;
        r.name = name;
        c->stages.push_back(r);
        stage = (int)c->stages.size() - 1;
    }
    c->stages[stage].bytes += bytes;
    c->stages[stage].launches++;
    a = take_event(c);
    b = take_event(c);
    cudaEventRecord(a, c->stream);
}

StageTimer::~StageTimer()
{
    if (stage < 0) return;
    cudaEventRecord(b, ctx->stream);
    ctx->pending.push_back({stage, a, b});
    if (ctx->pending.size() > 4096) pcp_profile_resolve(ctx);
}

void pcp_profile_resolve(pcp_ctx* ctx)
{
    for (auto& p : ctx->pending) {
        cudaEventSynchronize(p.b);
        float ms = 0;
        if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) ctx->stages[p.stage].ms += ms;
        ctx->ev_free.push_back(p.a);
        ctx->ev_free.push_back(p.b);
    }
    ctx->pending.clear();
}

extern "C" {

const char* pcp_version(void) { return "pcp-b200 0.1 (sm_100a)"; }

int pcp_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int pcp_create(pcp_ctx** out, int cuda_device)
{
    if (!out) return pcp_fail(nullptr, PCP_ERR_INVALID, "pcp_create: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return pcp_fail(nullptr, PCP_ERR_CUDA, "pcp_create: no CUDA device (%s); this library has no CPU fallback",
                        e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    }
    if (cuda_device < 0 || cuda_device >= n)
        return pcp_fail(nullptr, PCP_ERR_INVALID, "pcp_create: device %d out of range [0,%d)", cuda_device, n);
    pcp_ctx* c = new pcp_ctx();
    c->device = cuda_device;
    // on any failure below: the message goes to the thread's error slot, what was created is released
    auto fail = [&](cudaError_t e, const char* what) {
        pcp_fail(nullptr, PCP_ERR_CUDA, "pcp_create: %s failed: %s", what, cudaGetErrorString(e));
        if (c->pool) cudaMemPoolDestroy(c->pool);
        if (c->own_stream) cudaStreamDestroy(c->own_stream);
        delete c;
        return PCP_ERR_CUDA;
    };
    cudaError_t e2;
    if ((e2 = cudaSetDevice(cuda_device)) != cudaSuccess) return fail(e2, "cudaSetDevice");
    cudaDeviceProp prop;
    if ((e2 = cudaGetDeviceProperties(&prop, cuda_device)) != cudaSuccess) return fail(e2, "cudaGetDeviceProperties");
    c->sm_count = prop.multiProcessorCount;
    if ((e2 = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking)) != cudaSuccess) return fail(e2, "cudaStreamCreate");
    c->stream = c->own_stream;
    cudaMemPoolProps pp;
    memset(&pp, 0, sizeof(pp));
    pp.allocType = cudaMemAllocationTypePinned;
    pp.handleTypes = cudaMemHandleTypeNone;
    pp.location.type = cudaMemLocationTypeDevice;
    pp.location.id = cuda_device;
    if ((e2 = cudaMemPoolCreate(&c->pool, &pp)) != cudaSuccess) return fail(e2, "cudaMemPoolCreate");
    uint64_t thr = UINT64_MAX; // keep freed blocks cached: steady-state calls do no driver allocation
    if ((e2 = cudaMemPoolSetAttribute(c->pool, cudaMemPoolAttrReleaseThreshold, &thr)) != cudaSuccess) return fail(e2, "cudaMemPoolSetAttribute");
    *out = c;
    return PCP_OK;
}

void pcp_destroy(pcp_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    pcp_profile_resolve(ctx);
    for (auto e : ctx->ev_free) cudaEventDestroy(e);
    for (int l = 0; l < PCP_COPY_LANES; l++) {
        if (!ctx->lane[l]) continue;
        cudaStreamSynchronize(ctx->lane[l]);
        cudaStreamDestroy(ctx->lane[l]);
        cudaEventDestroy(ctx->lane_in[l]);
        cudaEventDestroy(ctx->lane_out[l]);
    }
    if (ctx->rb_host) cudaFreeHost(ctx->rb_host);
    if (ctx->wb_host) cudaFreeHost(ctx->wb_host);
    if (ctx->scratch) cudaFree(ctx->scratch);
    if (ctx->pool) cudaMemPoolDestroy(ctx->pool);
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    delete ctx;
}

const char* pcp_last_error(const pcp_ctx* ctx) { return ctx ? ctx->err.c_str() : g_last_error.c_str(); }

int pcp_set_stream(pcp_ctx* ctx, void* cuda_stream)
{
    if (!ctx) return PCP_ERR_INVALID;
    cudaStreamSynchronize(ctx->stream);
    ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
    return PCP_OK;
}

int pcp_synchronize(pcp_ctx* ctx)
{
    if (!ctx) return PCP_ERR_INVALID;
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

int pcp_trim(pcp_ctx* ctx)
{
    if (!ctx) return PCP_ERR_INVALID;
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    PCP_CUDA_TRY(ctx, cudaMemPoolTrimTo(ctx->pool, 0));
    if (ctx->scratch) cudaFree(ctx->scratch);
    ctx->scratch = nullptr, ctx->scratch_bytes = 0;
    return PCP_OK;
}

int pcp_profile_enable(pcp_ctx* ctx, int on)
{
    if (!ctx) return PCP_ERR_INVALID;
    ctx->profile = on != 0;
    return PCP_OK;
}

int pcp_profile_reset(pcp_ctx* ctx)
{
    if (!ctx) return PCP_ERR_INVALID;
    pcp_profile_resolve(ctx);
    ctx->stages.clear();
    return PCP_OK;
}

int pcp_profile_get(pcp_ctx* ctx, pcp_stage_time* out, int cap, int* n)
{
    if (!ctx || !n) return PCP_ERR_INVALID;
    pcp_profile_resolve(ctx);
    *n = (int)ctx->stages.size();
    for (int i = 0; i < *n && i < cap && out; i++) {
        memset(&out[i], 0, sizeof(out[i]));
        strncpy(out[i].name, ctx->stages[i].name.c_str(), sizeof(out[i].name) - 1);
        out[i].ms = ctx->stages[i].ms;
        out[i].launches = ctx->stages[i].launches;
        out[i].bytes = ctx->stages[i].bytes;
    }
    return PCP_OK;
}

int64_t pcp_launch_count(const pcp_ctx* ctx) { return ctx ? ctx->launches : 0; }

int pcp_dev_alloc(pcp_ctx* ctx, size_t bytes, void** dptr)
{
    if (!ctx || !dptr) return PCP_ERR_INVALID;
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    PCP_CUDA_TRY(ctx, cudaMalloc(dptr, bytes ? bytes : 1));
    return PCP_OK;
}

int pcp_dev_free(pcp_ctx* ctx, void* dptr)
{
    if (!ctx) return PCP_ERR_INVALID;
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    PCP_CUDA_TRY(ctx, cudaFree(dptr));
    return PCP_OK;
}

int pcp_host_alloc_pinned(pcp_ctx* ctx, size_t bytes, void** hptr)
{
    if (!ctx || !hptr) return PCP_ERR_INVALID;
    PCP_CUDA_TRY(ctx, cudaHostAlloc(hptr, bytes ? bytes : 1, cudaHostAllocDefault));
    return PCP_OK;
}

int pcp_host_free_pinned(pcp_ctx* ctx, void* hptr)
{
    if (!ctx) return PCP_ERR_INVALID;
    PCP_CUDA_TRY(ctx, cudaFreeHost(hptr));
    return PCP_OK;
}

int pcp_memcpy(pcp_ctx* ctx, void* dst, const void* src, size_t bytes)
{
    if (!ctx) return PCP_ERR_INVALID;
    if (!bytes) return PCP_OK;
    PCP_CUDA_TRY(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, ctx->stream));
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

int pcp_memcpy_async(pcp_ctx* ctx, void* dst, const void* src, size_t bytes)
{
    if (!ctx) return PCP_ERR_INVALID;
    if (!bytes) return PCP_OK;
    PCP_CUDA_TRY(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, ctx->stream));
    return PCP_OK;
}

static __global__ void readback_kernel(unsigned char* __restrict__ dst, const unsigned char* __restrict__ src, unsigned bytes)
{
    for (unsigned i = threadIdx.x; i < bytes; i += blockDim.x) dst[i] = src[i];
}

} // extern "C"

int pcp_writeto(pcp_ctx* ctx, void* dev_dst, const void* host_src, size_t bytes)
{
    if (!bytes) return PCP_OK;
    if (bytes > PCP_READBACK_MAX) {
        PCP_CUDA_TRY(ctx, cudaMemcpyAsync(dev_dst, host_src, bytes, cudaMemcpyHostToDevice, ctx->stream));
        PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
        return PCP_OK;
    }
    if (!ctx->wb_host) PCP_CUDA_TRY(ctx, cudaHostAlloc(&ctx->wb_host, PCP_READBACK_MAX, cudaHostAllocMapped));
    void* dalias = nullptr;
    PCP_CUDA_TRY(ctx, cudaHostGetDevicePointer(&dalias, ctx->wb_host, 0));
    memcpy(ctx->wb_host, host_src, bytes);
    readback_kernel<<<1, 256, 0, ctx->stream>>>(static_cast<unsigned char*>(dev_dst), static_cast<const unsigned char*>(dalias), (unsigned)bytes);
    PCP_CUDA_TRY(ctx, cudaGetLastError());
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream)); // the staging word is free again
    return PCP_OK;
}

int pcp_readback(pcp_ctx* ctx, void* host_dst, const void* dev_src, size_t bytes)
{
    if (!bytes) return PCP_OK;
    if (bytes > PCP_READBACK_MAX) {
        PCP_CUDA_TRY(ctx, cudaMemcpyAsync(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
        PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
        return PCP_OK;
    }
    if (!ctx->rb_host) PCP_CUDA_TRY(ctx, cudaHostAlloc(&ctx->rb_host, PCP_READBACK_MAX, cudaHostAllocMapped));
    void* dalias = nullptr;
    PCP_CUDA_TRY(ctx, cudaHostGetDevicePointer(&dalias, ctx->rb_host, 0));
    readback_kernel<<<1, 256, 0, ctx->stream>>>(static_cast<unsigned char*>(dalias), static_cast<const unsigned char*>(dev_src),
                                                (unsigned)bytes);
    PCP_CUDA_TRY(ctx, cudaGetLastError());
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    memcpy(host_dst, ctx->rb_host, bytes);
    return PCP_OK;
}

extern "C" {

static int lane_open(pcp_ctx* ctx, int lane)
{
    if (lane < 0 || lane >= PCP_COPY_LANES) return pcp_fail(ctx, PCP_ERR_INVALID, "copy lane %d out of range", lane);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (!ctx->lane[lane]) {
        PCP_CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->lane[lane], cudaStreamNonBlocking));
        PCP_CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->lane_in[lane], cudaEventDisableTiming));
        PCP_CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->lane_out[lane], cudaEventDisableTiming));
    }
    return PCP_OK;
}

int pcp_copy_async(pcp_ctx* ctx, int lane, void* dst, const void* src, size_t bytes)
{
    if (!ctx || (bytes && (!dst || !src))) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_copy_async: bad argument");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_TRY(lane_open(ctx, lane));
    PCP_CUDA_TRY(ctx, cudaEventRecord(ctx->lane_in[lane], ctx->stream));
    PCP_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->lane[lane], ctx->lane_in[lane], 0));
    // In pieces: a copy engine serves one copy at a time, and the few-byte read-backs of the compute stream
    // (counts, cell totals) must not queue behind a gigabyte transfer.
    static const size_t piece = ((size_t)(getenv("PCP_COPY_PIECE_MB") ? std::max(1, atoi(getenv("PCP_COPY_PIECE_MB"))) : 32)) << 20;
    for (size_t off = 0; off < bytes; off += piece)
        PCP_CUDA_TRY(ctx, cudaMemcpyAsync(static_cast<char*>(dst) + off, static_cast<const char*>(src) + off,
                                          std::min(piece, bytes - off), cudaMemcpyDefault, ctx->lane[lane]));
    return PCP_OK;
}

int pcp_copy_fence(pcp_ctx* ctx, int lane)
{
    if (!ctx) return PCP_ERR_INVALID;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_TRY(lane_open(ctx, lane));
    PCP_CUDA_TRY(ctx, cudaEventRecord(ctx->lane_out[lane], ctx->lane[lane]));
    PCP_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, ctx->lane_out[lane], 0));
    return PCP_OK;
}

int pcp_copy_wait(pcp_ctx* ctx, int lane)
{
    if (!ctx) return PCP_ERR_INVALID;
    cudaStream_t s = nullptr;
    {
        std::lock_guard<std::recursive_mutex> lock(ctx->mu);
        PCP_TRY(lane_open(ctx, lane));
        s = ctx->lane[lane];
    }
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    return PCP_OK;
}

} // extern "C"
