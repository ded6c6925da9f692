// sort.cuh -- key/value radix sort used to rebuild the voxel order and the Morton search structure.
#pragma once
#include "common.cuh"

// Sorts n (key, value) pairs by the low `bits` bits of key (stable LSD radix sort).
// Both key and value buffers are double-buffered: on return *keys / *vals point at the sorted data
// (which may be either of the two buffers passed in).
// iota_vals: the values are the indices 0..n-1 in order and need not have been written by the caller (the first
// pass generates them).  n < 2^30.
int pcp_sort_pairs_u64(pcp_ctx* ctx, unsigned long long** keys, unsigned long long** keys_alt, uint32_t** vals, uint32_t** vals_alt,
                       size_t n, int bits, bool iota_vals = false);
int pcp_sort_pairs_u32(pcp_ctx* ctx, uint32_t** keys, uint32_t** keys_alt, uint32_t** vals, uint32_t** vals_alt,
                       size_t n, int bits, bool iota_vals = false);
// exclusive prefix sum of n uint32 into n+1 outputs is done by callers via pcp_exclusive_scan_u32
int pcp_exclusive_scan_u32(pcp_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n);
int pcp_exclusive_scan_u64(pcp_ctx* ctx, const unsigned long long* in, unsigned long long* out, size_t n);
// plain (non-reproducible-order) prefix sum of doubles: only used for predictions, never for results
int pcp_exclusive_scan_f64(pcp_ctx* ctx, const double* in, double* out, size_t n);
