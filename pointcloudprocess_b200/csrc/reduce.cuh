// reduce.cuh -- shared declarations for reductions over strided clouds.
#pragma once
#include "common.cuh"

struct pcp_bbox {
    double mn[3];
    double mx[3];               // true max over finite points (-DBL_MAX when empty); DBL_MIN seed applied by callers
    unsigned long long count;   // finite points
};

// dpts is a device pointer. d_out is a device pointer (no sync) / h_out a host pointer (syncs).
int pcp_bbox_device(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, pcp_bbox* d_out);
int pcp_bbox_host(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, pcp_bbox* h_out);

// Serial-order sum of x,y,z over the finite points (compute3DCentroid, point_cloud_helper.h:192-230),
// bit-identical to the left-to-right double accumulation.  d_sum3: 3 doubles on the device.
int pcp_serial_sum_device(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, double* d_sum3, int carry = 0);
