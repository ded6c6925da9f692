// sort.cu -- round-1 implementation: CUB's onesweep radix sort and decoupled-lookback scan (library code,
// counted as such in DESIGN.md); the hand-written sm_100a radix sort replaces it in a later round.
#include "sort.cuh"
#include <cub/cub.cuh>

template <typename K>
static int sort_pairs(pcp_ctx* ctx, K** keys, K** keys_alt, uint32_t** vals, uint32_t** vals_alt, size_t n, int bits)
{
    if (n == 0) return PCP_OK;
    if (n > (size_t)INT32_MAX) return pcp_fail(ctx, PCP_ERR_RANGE, "sort: more than 2^31-1 items");
    if (bits < 1) bits = 1;
    cub::DoubleBuffer<K> dk(*keys, *keys_alt);
    cub::DoubleBuffer<uint32_t> dv(*vals, *vals_alt);
    size_t tmp_bytes = 0;
    PCP_CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, (int)n, 0, bits, ctx->stream));
    DevBuf<char> tmp;
    PCP_TRY(tmp.alloc(ctx, tmp_bytes));
    PCP_CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, dk, dv, (int)n, 0, bits, ctx->stream));
    ctx->launches += 1 + (bits + 7) / 8;
    K* kc = dk.Current();
    uint32_t* vc = dv.Current();
    if (kc != *keys) {
        *keys_alt = *keys;
        *keys = kc;
    }
    if (vc != *vals) {
        *vals_alt = *vals;
        *vals = vc;
    }
    return PCP_OK;
}

int pcp_sort_pairs_u64(pcp_ctx* ctx, unsigned long long** keys, unsigned long long** keys_alt, uint32_t** vals, uint32_t** vals_alt,
                       size_t n, int bits)
{
    return sort_pairs<unsigned long long>(ctx, keys, keys_alt, vals, vals_alt, n, bits);
}
int pcp_sort_pairs_u32(pcp_ctx* ctx, uint32_t** keys, uint32_t** keys_alt, uint32_t** vals, uint32_t** vals_alt,
                       size_t n, int bits)
{
    return sort_pairs<uint32_t>(ctx, keys, keys_alt, vals, vals_alt, n, bits);
}

template <typename T> static int excl_scan(pcp_ctx* ctx, const T* in, T* out, size_t n)
{
    if (n == 0) return PCP_OK;
    size_t tmp_bytes = 0;
    PCP_CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, in, out, (int)n, ctx->stream));
    DevBuf<char> tmp;
    PCP_TRY(tmp.alloc(ctx, tmp_bytes));
    PCP_CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, in, out, (int)n, ctx->stream));
    ctx->launches += 2;
    return PCP_OK;
}
int pcp_exclusive_scan_u32(pcp_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n) { return excl_scan(ctx, in, out, n); }
int pcp_exclusive_scan_u64(pcp_ctx* ctx, const unsigned long long* in, unsigned long long* out, size_t n) { return excl_scan(ctx, in, out, n); }
int pcp_exclusive_scan_f64(pcp_ctx* ctx, const double* in, double* out, size_t n) { return excl_scan(ctx, in, out, n); }
