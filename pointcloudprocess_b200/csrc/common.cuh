// common.cuh -- context, stream-ordered arena, host/device pointer staging, stage timing.
// Internal to libpcp_cuda.so; the public surface is include/pcp_cuda.h.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/pcp_cuda.h"

#define PCP_SM_COUNT_B200 148

struct pcp_stage_rec {
    std::string name;
    double ms = 0;
    int64_t launches = 0;
    double bytes = 0;
};

struct pcp_ctx {
    int device = 0;
    int sm_count = PCP_SM_COUNT_B200;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    cudaMemPool_t pool = nullptr;
    std::string err;
    std::recursive_mutex mu; // recursive: entry points compose (pcp_features_rpca calls pcp_features_knn_ex)
    bool profile = false;
    std::vector<pcp_stage_rec> stages;
    int64_t launches = 0;
    // pending stage events (resolved lazily at pcp_profile_get / next sync)
    struct pending_ev {
        int stage;
        cudaEvent_t a, b;
    };
    std::vector<pending_ev> pending;
    std::vector<cudaEvent_t> ev_free;
    // copy lanes (pcp_copy_async): streams created on first use, one reusable event per direction
    void* rb_host = nullptr; // mapped pinned scratch of pcp_readback
    void* wb_host = nullptr; // ... and of pcp_writeto
    // one grow-only device buffer kept between calls (pcp_scratch_get): the neighbour rows of the kNN kernels.  A
    // 2 GB block allocated and freed on every call made the pool's other allocations slow and erratic (sharded
    // steps: 4 ms -> 7..37 ms for the same remove_duplicate); kept here it is allocated once.
    void* scratch = nullptr;
    size_t scratch_bytes = 0;
    double hint_h = 0, hint_areal = 0; // index.cu: cell size and areal density of the previous index
    cudaStream_t lane[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t lane_in[4] = {nullptr, nullptr, nullptr, nullptr};  // compute -> lane
    cudaEvent_t lane_out[4] = {nullptr, nullptr, nullptr, nullptr}; // lane -> compute
};

int pcp_fail(pcp_ctx* ctx, int code, const char* fmt, ...);
// A few bytes from device memory to the host, ordered on the context's stream, and a wait for them.  A kernel
// writes them into mapped host memory: a cudaMemcpy would queue behind whatever bulk download a copy lane
// has in flight on the device-to-host copy engine (pcp_copy_async).
#define PCP_READBACK_MAX 65536
int pcp_readback(pcp_ctx* ctx, void* host_dst, const void* dev_src, size_t bytes);
// The mirror: a few bytes from the host into device memory, ordered on the context's stream, without a place in the
// host-to-device copy engine's queue (a kernel reads them from mapped host memory).  Waits for the kernel.
int pcp_writeto(pcp_ctx* ctx, void* dev_dst, const void* host_src, size_t bytes);
void pcp_set_global_error(const char* msg);

#define PCP_CUDA_TRY(ctx, expr)                                                                              \
    do {                                                                                                     \
        cudaError_t _e = (expr);                                                                             \
        if (_e != cudaSuccess)                                                                               \
            return pcp_fail((ctx), PCP_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),     \
                            __FILE__, __LINE__);                                                             \
    } while (0)

#define PCP_TRY(expr)                                                                                        \
    do {                                                                                                     \
        int _s = (expr);                                                                                     \
        if (_s != PCP_OK) return _s;                                                                         \
    } while (0)

// after a kernel launch
#define PCP_LAUNCH_CHECK(ctx)                                                                                \
    do {                                                                                                     \
        (ctx)->launches++;                                                                                   \
        PCP_CUDA_TRY((ctx), cudaGetLastError());                                                             \
    } while (0)

// ---- stream-ordered arena ----------------------------------------------------------------------------
int pcp_arena_alloc(pcp_ctx* ctx, size_t bytes, void** out);
// the context's persistent scratch, grown to at least `bytes` (contents undefined; valid until the next call)
int pcp_scratch_get(pcp_ctx* ctx, size_t bytes, void** out);
void pcp_arena_free(pcp_ctx* ctx, void* p);

// RAII device buffer from the arena (freed stream-ordered at scope exit)
template <typename T> struct DevBuf {
    pcp_ctx* ctx = nullptr;
    T* p = nullptr;
    size_t n = 0;
    DevBuf() {}
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    int alloc(pcp_ctx* c, size_t count)
    {
        release();
        ctx = c;
        n = count;
        void* q = nullptr;
        int s = pcp_arena_alloc(c, (count ? count : 1) * sizeof(T), &q);
        p = static_cast<T*>(q);
        return s;
    }
    void release()
    {
        if (p) pcp_arena_free(ctx, p);
        p = nullptr;
        n = 0;
    }
    T* detach()
    {
        T* q = p;
        p = nullptr;
        return q;
    }
};

// ---- pointer staging -----------------------------------------------------------------------------------
enum pcp_ptr_kind { PCP_PTR_HOST = 0, PCP_PTR_DEVICE = 1, PCP_PTR_BAD = 2 };
pcp_ptr_kind pcp_classify(pcp_ctx* ctx, const void* p);

// Input that must be readable on the device: device pointers pass through, host data is copied in.
struct InStage {
    pcp_ctx* ctx = nullptr;
    const void* dptr = nullptr;
    void* owned = nullptr;
    ~InStage()
    {
        if (owned) pcp_arena_free(ctx, owned);
    }
    int open(pcp_ctx* c, const void* p, size_t bytes);
};

// Output written on the device: device pointers are written in place, host outputs get a device
// scratch that commit() copies back (synchronises the stream).
struct OutStage {
    pcp_ctx* ctx = nullptr;
    void* user = nullptr;
    void* dptr = nullptr;
    void* owned = nullptr;
    size_t bytes = 0;
    ~OutStage()
    {
        if (owned) pcp_arena_free(ctx, owned);
    }
    int open(pcp_ctx* c, void* p, size_t nbytes, bool preload = false);
    int commit(size_t nbytes_override = (size_t)-1);
};

// ---- stage timing --------------------------------------------------------------------------------------
struct StageTimer {
    pcp_ctx* ctx;
    int stage = -1;
    cudaEvent_t a = nullptr, b = nullptr;
    StageTimer(pcp_ctx* c, const char* name, double bytes = 0);
    ~StageTimer();
};
void pcp_profile_resolve(pcp_ctx* ctx);

static inline unsigned pcp_blocks_for(size_t n, unsigned threads) { return (unsigned)((n + threads - 1) / threads); }

// 256-bit global loads and stores (LDG.E.256 / STG.E.256, sm_100): one instruction and one L1 wavefront for a 32-byte
// record; the address must be 32-byte aligned.
struct pcp_d4 {
    double a, b, c, d;
};
__device__ __forceinline__ pcp_d4 pcp_ld256(const void* p)
{
    pcp_d4 v;
    asm volatile("ld.global.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(v.a), "=d"(v.b), "=d"(v.c), "=d"(v.d) : "l"(p));
    return v;
}
__device__ __forceinline__ void pcp_st256(void* p, const pcp_d4& v)
{
    asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" : : "l"(p), "d"(v.a), "d"(v.b), "d"(v.c), "d"(v.d) : "memory");
}
