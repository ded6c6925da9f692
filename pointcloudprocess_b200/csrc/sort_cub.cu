// sort_cub.cu -- CUB's radix sort behind PCP_SORT_CUB=1: LIBRARY code, kept only as the A/B reference that
// tests/test_gpu_sort.py compares the hand-written sort (sort.cu) with.  Not on the product path.
#include "sort.cuh"
#include <cub/cub.cuh>

template <typename K>
static int cub_sort_pairs(pcp_ctx* ctx, K** keys, K** keys_alt, uint32_t** vals, uint32_t** vals_alt, size_t n, int bits)
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

int pcp_cub_sort_pairs_u64(pcp_ctx* ctx, unsigned long long** keys, unsigned long long** keys_alt, uint32_t** vals,
                           uint32_t** vals_alt, size_t n, int bits)
{
    return cub_sort_pairs<unsigned long long>(ctx, keys, keys_alt, vals, vals_alt, n, bits);
}
int pcp_cub_sort_pairs_u32(pcp_ctx* ctx, uint32_t** keys, uint32_t** keys_alt, uint32_t** vals, uint32_t** vals_alt, size_t n,
                           int bits)
{
    return cub_sort_pairs<uint32_t>(ctx, keys, keys_alt, vals, vals_alt, n, bits);
}
