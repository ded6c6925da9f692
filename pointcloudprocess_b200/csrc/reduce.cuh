// reduce.cuh -- shared declarations for reductions over strided clouds.
#pragma once
#include "common.cuh"

struct pcp_bbox {
    double mn[3];
    double mx[3];               // true max over finite points (-DBL_MAX when empty); DBL_MIN seed applied by callers
    unsigned long long count;   // finite points
    double sum[3];              // x, y, z sums in TREE order: a prediction of the serial-order sums, never a result
};

// dpts is a device pointer. d_out is a device pointer (no sync) / h_out a host pointer (syncs).
int pcp_bbox_device(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, pcp_bbox* d_out);
int pcp_bbox_host(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, pcp_bbox* h_out);

// Serial-order sum of x,y,z over the finite points (compute3DCentroid, point_cloud_helper.h:192-230),
// bit-identical to the left-to-right double accumulation.  d_sum3: 3 doubles on the device.
int pcp_serial_sum_device(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, double* d_sum3, int carry = 0);

// The serial-order sum in two steps, for clouds whose point order runs over several chunks / GPUs: `prepare` does the
// parallel part (the transfer records of the chunk) under a PREDICTED carry (the approximate sum of everything before
// the chunk: only the binades matter), `finish` walks the records with the TRUE carry (carry3 in: sum of the points
// before the chunk, out: including the chunk) -- a few tens of microseconds, the only sequential step of a chain of
// chunks.  The cloud must stay in place between the two calls.
struct pcp_serial_plan;
int pcp_serial_sum_prepare(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, const double* d_pred3, pcp_serial_plan** out);
int pcp_serial_sum_finish(pcp_ctx* ctx, pcp_serial_plan* plan, double* d_carry3);
void pcp_serial_sum_release(pcp_ctx* ctx, pcp_serial_plan* plan);
