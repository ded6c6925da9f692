// knn.cu -- host side of the neighbour-search entry points of include/pcp_cuda.h:
// pcp_features_knn (the calculate_plan_parameter path, calculate_feature.h:15) and pcp_knn
// (KdTreeFLANN::nearestKSearch, kd_tree.h:814-845).
#include "knn_warp.cuh"
#include "knn_far.cuh"
#include "knn_wide.cuh"
#include <algorithm>
#include <climits>
#include <cstdlib>

namespace {

__global__ void __launch_bounds__(256) props_fallback_kernel(void* props, double* eig3, float* lps3, int32_t* knn_idx,
                                                             int k, size_t n)
{
    size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (j >= n) return;
    pcp_plane pl;
    pcp_store_props(props, j, false, pl);
    pcp_store_extra(eig3, lps3, j, false, pl);
    if (knn_idx)
        for (int e = 0; e < k; e++) knn_idx[j * k + e] = -1;
}

// qpts != NULL: also the squared distances (the expression of the search kernels, bit for bit: L2_Simple order, no FMA)
__global__ void __launch_bounds__(256) knn_export_kernel(const pcp_rec* __restrict__ rec,
                                                         const uint32_t* __restrict__ pos, size_t total, int k,
                                                         int32_t* __restrict__ idx, int32_t* __restrict__ found,
                                                         const char* __restrict__ qpts = nullptr, size_t qstride = 0,
                                                         double* __restrict__ d2 = nullptr)
{
    size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= total) return;
    uint32_t p = pos[t];
    idx[t] = p == 0xffffffffu ? -1 : (int32_t)rec[p].idx;
    if (d2) {
        double v = __longlong_as_double(0x7ff0000000000000ll);
        if (p != 0xffffffffu) {
            const double* qp = reinterpret_cast<const double*>(qpts + (t / k) * qstride);
            const pcp_rec r = pcp_load_rec(rec, p);
            v = pcp_dist2(qp[0], qp[1], qp[2], r.x, r.y, r.z);
        }
        d2[t] = v;
    }
    if (found && (t % k) == 0) {
        int c = 0;
        for (int e = 0; e < k; e++) c += pos[t + e] != 0xffffffffu;
        found[t / k] = c;
    }
}

template <int KPL> int launch_wknn(pcp_ctx* ctx, const pcp_wknn_params& P, uint32_t expected)
{
    int blocks = ctx->sm_count * 4;
    if (expected && (expected + 3) / 4 < (uint32_t)blocks) blocks = (int)((expected + 3) / 4);
    if (blocks < 1) blocks = 1;
    knn_warp_kernel<KPL><<<blocks, 128, 0, ctx->stream>>>(P);
    PCP_LAUNCH_CHECK(ctx);
    return PCP_OK;
}

template <int CAP, int GS, int CAND, bool FLAT, bool DEFER = false> int launch_fknn_t(pcp_ctx* ctx, const pcp_fknn_params& P)
{
    if constexpr (!DEFER) {
        if (P.nbr_out) return launch_fknn_t<CAP, GS, CAND, FLAT, true>(ctx, P);
    }
    const size_t smem = PCP_FK_WARPS * pcp_fd_smem<CAP, GS, CAND, DEFER>::bytes(P.k);
    PCP_CUDA_TRY(ctx, cudaFuncSetAttribute(knn_fast_kernel<CAP, GS, CAND, FLAT, DEFER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    PCP_CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, knn_fast_kernel<CAP, GS, CAND, FLAT, DEFER>, 32 * PCP_FK_WARPS, smem));
    if (per_sm < 1) per_sm = 1;
    // persistent warps: every resident slot of every SM is filled once, items come from the work counter
    unsigned blocks = (unsigned)(ctx->sm_count * per_sm);
    const unsigned need = pcp_blocks_for(P.nitem, PCP_FK_WARPS);
    if (blocks > need) blocks = need;
    if (P.nitem_ptr) blocks = (unsigned)(ctx->sm_count * per_sm); // item count known on the device only
    if (P.dbg) fprintf(stderr, "[pcp knn] dense kernel: %zu bytes of shared memory per block, %d blocks per SM, %u blocks\n", smem, per_sm, blocks);
    knn_fast_kernel<CAP, GS, CAND, FLAT, DEFER><<<blocks, 32 * PCP_FK_WARPS, smem, ctx->stream>>>(P);
    PCP_LAUNCH_CHECK(ctx);
    return PCP_OK;
}

// list capacity CAP >= k plus room for the candidates of the boundary bucket
int launch_fknn(pcp_ctx* ctx, const pcp_fknn_params& P)
{
    if (P.kk <= 24 && getenv("PCP_FK_NOFLAT")) return launch_fknn_t<32, 8, 320, false>(ctx, P);
    if (P.kk <= 24) return launch_fknn_t<32, 8, 320, true>(ctx, P);
    if (P.kk <= 52) return launch_fknn_t<64, 8, 480, true>(ctx, P);
    return launch_fknn_t<160, 8, 480, true>(ctx, P);
}

// the third pass: leaves with open queries, over the box of their bound spheres (knn_wide.cuh)
template <int CAP, int GS, int CAND> int launch_wide_t(pcp_ctx* ctx, const pcp_fknn_params& P)
{
    const size_t smem = PCP_FK_WARPS * pcp_wide_layout<CAP, GS, CAND>::bytes();
    PCP_CUDA_TRY(ctx, cudaFuncSetAttribute(knn_wide_kernel<CAP, GS, CAND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    PCP_CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, knn_wide_kernel<CAP, GS, CAND>, 32 * PCP_FK_WARPS, smem));
    if (per_sm < 1) return pcp_fail(ctx, PCP_ERR_CUDA, "knn_wide_kernel does not fit an SM");
    knn_wide_kernel<CAP, GS, CAND><<<ctx->sm_count * per_sm, 32 * PCP_FK_WARPS, smem, ctx->stream>>>(P);
    PCP_LAUNCH_CHECK(ctx);
    return PCP_OK;
}
int launch_wide(pcp_ctx* ctx, const pcp_fknn_params& P)
{
    if (P.kk <= 24) return launch_wide_t<32, 8, 256>(ctx, P);
    return launch_wide_t<64, 8, 480>(ctx, P);
}

// the second pass over the items the first one found too dense to stage (k <= 52)
#ifndef PCP_FK_CAND2
#define PCP_FK_CAND2 800
#endif
int launch_fknn_dense2(pcp_ctx* ctx, const pcp_fknn_params& P)
{
    if (P.kk <= 24) return launch_fknn_t<32, 8, PCP_FK_CAND2, true>(ctx, P);
    return launch_fknn_t<64, 8, 1008, true>(ctx, P); // k <= 52: the first pass stages 480
}

template <int CAP, bool DEFER = false> int launch_far_t(pcp_ctx* ctx, const pcp_fknn_params& P, const pcp_far_args& A)
{
    if constexpr (!DEFER) {
        if (P.nbr_out) return launch_far_t<CAP, true>(ctx, P, A);
    }
    const size_t smem = PCP_FK_WARPS * pcp_far_smem<CAP>::bytes(P.k);
    PCP_CUDA_TRY(ctx, cudaFuncSetAttribute(knn_far_kernel<CAP, DEFER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    PCP_CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, knn_far_kernel<CAP, DEFER>, 32 * PCP_FK_WARPS, smem));
    if (per_sm < 1) per_sm = 1;
    knn_far_kernel<CAP, DEFER><<<ctx->sm_count * per_sm, 32 * PCP_FK_WARPS, smem, ctx->stream>>>(P, A);
    PCP_LAUNCH_CHECK(ctx);
    return PCP_OK;
}

int launch_far(pcp_ctx* ctx, const pcp_fknn_params& P, const pcp_far_args& A)
{
    if (P.kk <= 40) return launch_far_t<64>(ctx, P, A);
    if (P.kk <= 100) return launch_far_t<128>(ctx, P, A);
    return launch_far_t<160>(ctx, P, A);
}

// a handful of queries: a block of 8 warps per query (the longest query is the kernel's time)
#define PCP_BKNN_W 8
#define PCP_BKNN_MAXQ 512
template <int KPL> int launch_bknn(pcp_ctx* ctx, const pcp_wknn_params& P, uint32_t expected)
{
    const size_t smem = (size_t)(PCP_BKNN_W - 1) * P.kk * sizeof(pcp_bent);
    int blocks = (int)std::min<uint32_t>(std::max<uint32_t>(expected, 1u), (uint32_t)ctx->sm_count * 2u);
    knn_block_kernel<KPL, PCP_BKNN_W><<<blocks, 32 * PCP_BKNN_W, smem, ctx->stream>>>(P);
    PCP_LAUNCH_CHECK(ctx);
    return PCP_OK;
}

int run_wknn(pcp_ctx* ctx, const pcp_wknn_params& P, uint32_t expected)
{
    if (expected && expected <= PCP_BKNN_MAXQ && !P.dbg && !getenv("PCP_KNN_NOBLOCK")) {
        if (P.k <= 32) return launch_bknn<1>(ctx, P, expected);
        if (P.k <= 64) return launch_bknn<2>(ctx, P, expected);
        if (P.k <= 128) return launch_bknn<4>(ctx, P, expected);
    }
    if (P.k <= 32) return launch_wknn<1>(ctx, P, expected);
    if (P.k <= 64) return launch_wknn<2>(ctx, P, expected);
    if (P.k <= 128) return launch_wknn<4>(ctx, P, expected);
    return pcp_fail(ctx, PCP_ERR_RANGE, "kNN: k > 128 is not supported");
}

} // namespace

extern "C" int pcp_features_knn(const pcp_index* ix, int k, void* props, double* eig3, float* lps3, int32_t* knn_idx)
{
    return pcp_features_knn_ex(ix, k, -1, props, eig3, lps3, knn_idx, nullptr);
}

extern "C" int pcp_features_knn_ex(const pcp_index* ix, int k, int64_t query_limit, void* props, double* eig3, float* lps3,
                                   int32_t* knn_idx, float* kth_d2)
{
    if (!ix || !props || k < 1) return pcp_fail(ix ? ix->ctx : nullptr, PCP_ERR_INVALID, "pcp_features_knn: bad argument");
    pcp_ctx* ctx = ix->ctx;
    if (k > 128) return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_features_knn: k > 128 is not supported");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const size_t n0 = ix->n_orig, nv = ix->n;
    if (n0 == 0) return PCP_OK;
    OutStage o_props, o_eig, o_lps, o_idx;
    PCP_TRY(o_props.open(ctx, props, n0 * 48));
    PCP_TRY(o_eig.open(ctx, eig3, n0 * 24));
    PCP_TRY(o_lps.open(ctx, lps3, n0 * 12));
    PCP_TRY(o_idx.open(ctx, knn_idx, n0 * (size_t)k * 4));
    OutStage o_kth;
    PCP_TRY(o_kth.open(ctx, kth_d2, n0 * 4));
    if (o_kth.dptr) PCP_CUDA_TRY(ctx, cudaMemsetAsync(o_kth.dptr, 0, n0 * 4, ctx->stream));
    const long long qlimit = query_limit < 0 ? LLONG_MAX : (long long)query_limit;
    const int kk = (int)(nv < (size_t)k ? nv : (size_t)k);
    if (nv < n0 || kk == 0) {
        props_fallback_kernel<<<pcp_blocks_for(n0, 256), 256, 0, ctx->stream>>>(
            o_props.dptr, (double*)o_eig.dptr, (float*)o_lps.dptr, (int32_t*)o_idx.dptr, k, n0);
        PCP_LAUNCH_CHECK(ctx);
    }
    if (nv > 0 && kk > 0) {
        DevBuf<uint32_t> fail_list, counters;
        DevBuf<float> fail_u;
        // (room for every query once, plus the ones the re-rank kernel puts back after the third pass had struck them:
        // at most its list's capacity, nv / 16)
        const size_t nfar_max = nv + nv / 16 + 1024;
        PCP_TRY(fail_list.alloc(ctx, nfar_max));
        PCP_TRY(fail_u.alloc(ctx, nfar_max));
        DevBuf<uint32_t> fail2;
        PCP_TRY(fail2.alloc(ctx, nfar_max));
        PCP_TRY(counters.alloc(ctx, 12));
        PCP_CUDA_TRY(ctx, cudaMemsetAsync(counters.p, 0, 48, ctx->stream));
        uint32_t nfail = 0;
        if (kk <= 128) {
            // work items = octree leaves whose capacity follows k (index.cu; 24 for k = 20)
            // (a cache inside the index, rebuilt under the context lock when k changes)
            PCP_TRY(pcp_index_build_items(ctx, const_cast<pcp_index*>(ix), pcp_item_cmax(kk)));
            pcp_fknn_params P;
            P.rec = ix->rec;
            P.fpt = ix->fpt;
            P.htab = ix->htab;
            P.hmask = ix->hmask;
            P.ltab = ix->ltab;
            P.lmask = ix->lmask;
            P.cinfo = ix->cinfo;
            P.ckey = ix->ckey;
            P.cstart = ix->cstart;
            P.ncell = ix->ncell;
            P.item_cell = ix->item_cell;
            P.item_first = ix->item_first;
            P.item_meta = ix->item_meta;
            P.nitem = ix->nitem;
            P.g = ix->grid;
            P.k = k;
            P.kk = kk;
            P.eps0 = ix->eps0;
            P.epsd = ix->epsd;
            P.query_limit = qlimit;
            P.kth_d2 = (float*)o_kth.dptr;
            P.props = o_props.dptr;
            P.eig3 = (double*)o_eig.dptr;
            P.lps3 = (float*)o_lps.dptr;
            P.knn_idx = (int32_t*)o_idx.dptr;
            P.fail_list = fail_list.p;
            P.fail_u = fail_u.p;
            P.fail_count = counters.p;
            P.work_counter = counters.p + 2;
            P.dbg = nullptr;
            DevBuf<unsigned long long> dbg;
            unsigned long long hfast6 = 0, hfast7 = 0;
            const bool want_dbg = getenv("PCP_KNN_DEBUG") != nullptr;
            if (want_dbg) {
                PCP_TRY(dbg.alloc(ctx, 32));
                PCP_CUDA_TRY(ctx, cudaMemsetAsync(dbg.p, 0, 32 * 8, ctx->stream));
                P.dbg = dbg.p;
            }
            // deferred features: the search kernels only write neighbour rows; one thread-per-query kernel fits the planes
            struct {
                uint32_t* p = nullptr;
            } rows;
            // (the search kernels address a row by a 32-bit offset)
            const bool defer = !getenv("PCP_FK_FUSED") && nv * (unsigned long long)k < (1ull << 32);
            if (defer) {
                void* q = nullptr;
                PCP_TRY(pcp_scratch_get(ctx, nv * (size_t)k * 4, &q)); // kept by the context between calls
                rows.p = static_cast<uint32_t*>(q);
                // (no pass over the rows: the dense kernels mark every row they leave without an answer)
                P.nbr_out = rows.p;
            }
            // lists the FP32 search cannot order go to knn_rerank_kernel (about 1 % of the queries; room for 1 in 16)
            DevBuf<uint32_t> rr_list;
            const int rr_cap_words = kk <= 24 ? 32 : 64; // the list capacity CAP of launch_fknn's instances
            if (defer && kk <= 52 && !getenv("PCP_FK_NORERANK")) {
                const size_t cap = std::max<size_t>(nv / 16, 1024);
                PCP_TRY(rr_list.alloc(ctx, cap * (size_t)(rr_cap_words + PCP_RR_HEAD)));
                P.rr.list = rr_list.p;
                P.rr.count = counters.p + 5;
                P.rr.cap = (uint32_t)cap;
            }
            DevBuf<uint2> wide_list;
            const bool wide = defer && kk <= 52 && !getenv("PCP_FK_NOWIDE");
            if (wide) {
                PCP_TRY(wide_list.alloc(ctx, nv)); // a record holds at least one of the far list's entries
                P.wide_list = wide_list.p;
                P.wide_count = counters.p + 8;
                P.wide_work = counters.p + 9;
            }
            DevBuf<uint32_t> dense_list;
            const bool two_pass = kk <= 52 && !getenv("PCP_FK_ONEPASS");
            if (two_pass) {
                PCP_TRY(dense_list.alloc(ctx, ix->nitem));
                P.dense_list = dense_list.p;
                P.dense_count = counters.p + 6;
            }
            {
                StageTimer t(ctx, "knn.fast_kernel", 72.0 * nv);
                PCP_TRY(launch_fknn(ctx, P));
            }
            if (two_pass) {
                // items whose block holds more than 320 points (a sparse leaf beside a dense surface): the same
                // kernel with room for PCP_FK_CAND2 staged candidates and fewer resident warps
                StageTimer t(ctx, "knn.fast_kernel_dense", 0);
                pcp_fknn_params P2 = P;
                P2.item_list = dense_list.p;
                P2.nitem_ptr = counters.p + 6;
                P2.dense_list = nullptr;
                P2.dense_count = nullptr;
                P2.work_counter = counters.p + 7;
                PCP_TRY(launch_fknn_dense2(ctx, P2));
            }
            if (wide) {
                StageTimer t(ctx, "knn.wide_kernel");
                PCP_TRY(launch_wide(ctx, P));
            }
            if (P.rr.list) {
                StageTimer t(ctx, "knn.rerank_kernel");
                if (rr_cap_words == 32) knn_rerank_kernel<32><<<ctx->sm_count * 8, 128, 0, ctx->stream>>>(P);
                else knn_rerank_kernel<64><<<ctx->sm_count * 8, 128, 0, ctx->stream>>>(P);
                PCP_LAUNCH_CHECK(ctx);
            }
            if (want_dbg) {
                unsigned long long h[16];
                PCP_TRY(pcp_readback(ctx, h, dbg.p, 16 * 8));
                PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
                hfast6 = h[6], hfast7 = h[7];
                fprintf(stderr, "[pcp knn] h=%.4f items=%llu staged=%llu queries=%llu cand/item=%.1f settled=%llu beyond=%llu "
                                "ambiguous=%llu too_dense: items=%llu queries=%llu too_sparse queries=%llu\n", ix->grid.h,
                        h[0], h[1], h[2], h[0] ? (double)h[3] / h[0] : 0.0, h[5], h[6], h[7], h[4], h[8], h[9]);
            }
            {
                // everything the dense kernel left open: sparse cells, far neighbours, FP32 ambiguities
                pcp_far_args A;
                A.list = fail_list.p;
                A.listU = fail_u.p;
                A.count = counters.p;
                A.work_counter = counters.p + 3;
                A.fail_list = fail2.p;
                A.fail_count = counters.p + 4;
                StageTimer t(ctx, "knn.far_kernel");
                PCP_TRY(launch_far(ctx, P, A));
            }
            if (defer) {
                StageTimer t(ctx, "knn.features_kernel", 72.0 * nv);
                features_from_list_kernel<<<pcp_blocks_for(nv, 128), 128, 0, ctx->stream>>>(
                    ix->rec, nullptr, nullptr, rows.p, k, kk, o_props.dptr, (double*)o_eig.dptr, (float*)o_lps.dptr,
                    (int32_t*)o_idx.dptr, (float*)o_kth.dptr, (uint32_t)nv);
                PCP_LAUNCH_CHECK(ctx);
            }
            uint32_t hc[5];
            PCP_TRY(pcp_readback(ctx, hc, counters.p, 20));
            PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
            nfail = hc[4];
            if (want_dbg) {
                unsigned long long h[32];
                PCP_TRY(pcp_readback(ctx, h, dbg.p, 32 * 8));
                PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
                fprintf(stderr, "[pcp knn] far queries=%u attempts=%llu (block too sparse %llu) cells/attempt=%.1f searches=%llu "
                                "cand/search=%.1f selected cells/search=%.1f uncertified=%llu tightened=%llu streaming-kernel queries=%u\n",
                        hc[0], h[10], h[16], h[10] ? (double)h[11] / h[10] : 0.0, h[12], h[12] ? (double)h[13] / h[12] : 0.0,
                        h[12] ? (double)h[14] / h[12] : 0.0, h[15], h[17], hc[4]);
                fprintf(stderr, "[pcp knn] wide pass: leaves=%llu (no bounded query %llu, cells hold too many or too few points %llu, box over capacity %llu) "
                                "queries=%llu settled=%llu points in cells/leaf=%.1f staged/leaf=%.1f\n", h[24], h[25], h[26], h[27], h[28],
                        h[29], h[24] ? (double)h[31] / h[24] : 0.0, h[24] ? (double)h[30] / h[24] : 0.0);
                fprintf(stderr, "[pcp knn] far searches with bound >= certified radius: %llu, of which certified anyway %llu\n", h[6] - hfast6, h[7] - hfast7);
                fprintf(stderr, "[pcp knn] far failures: overflow=%llu (cells %llu) beyond=%llu uncert=%llu no-bound=%llu mean level=%.2f\n",
                        h[18], h[23], h[19], h[20], h[21], hc[4] ? (double)h[22] / hc[4] : 0.0);
            }
            if (nfail) PCP_TRY(pcp_writeto(ctx, counters.p, &hc[4], 4));
        } else {
            // k too large for the cell kernel's shared-memory lists: every query takes the warp kernel
            nfail = (uint32_t)nv;
            // qlist = identity over sorted positions
            // (reuse fail_list; filled by a trivial kernel below)
        }
        if (nfail) {
            DevBuf<uint32_t> nbr;
            PCP_TRY(nbr.alloc(ctx, (size_t)nfail * k));
            pcp_wknn_params W;
            memset(&W, 0, sizeof(W));
            W.rec = ix->rec;
            W.ckey = ix->ckey;
            W.cstart = ix->cstart;
            W.ncell = ix->ncell;
            W.g = ix->grid;
    W.htab = ix->htab, W.hmask = ix->hmask, W.ltab = ix->ltab, W.lmask = ix->lmask;
            W.k = k;
            W.kk = kk;
            W.qcount = nfail;
            W.out_pos = nbr.p;
            W.work_counter = counters.p + 1;
            DevBuf<uint32_t> ident;
            if (kk > 128) {
                // identity query list with start level 0
                PCP_TRY(ident.alloc(ctx, nv));
                std::vector<uint32_t> h(nv);
                for (size_t i = 0; i < nv; i++) h[i] = (uint32_t)i;
                PCP_CUDA_TRY(ctx, cudaMemcpyAsync(ident.p, h.data(), nv * 4, cudaMemcpyHostToDevice, ctx->stream));
                PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
                PCP_CUDA_TRY(ctx, cudaMemcpyAsync(counters.p, &nfail, 4, cudaMemcpyHostToDevice, ctx->stream));
                W.qlist = ident.p;
            } else {
                W.qlist = fail2.p;
            }
            {
                StageTimer t(ctx, "knn.warp_kernel", 72.0 * nfail); // bytes/72 = queries that fell back
                PCP_TRY(run_wknn(ctx, W, nfail));
            }
            {
                StageTimer t(ctx, "knn.fallback_features");
                features_from_list_kernel<<<pcp_blocks_for(nfail, 128), 128, 0, ctx->stream>>>(
                    ix->rec, W.qlist, counters.p, nbr.p, k, kk, o_props.dptr, (double*)o_eig.dptr,
                    (float*)o_lps.dptr, (int32_t*)o_idx.dptr, (float*)o_kth.dptr);
                PCP_LAUNCH_CHECK(ctx);
            }
        }
    }
    PCP_TRY(o_props.commit());
    PCP_TRY(o_eig.commit());
    PCP_TRY(o_lps.commit());
    PCP_TRY(o_idx.commit());
    PCP_TRY(o_kth.commit());
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

extern "C" int pcp_knn(const pcp_index* ix, const void* q, size_t nq, size_t qstride, int k, int32_t* idx, double* d2,
                       int32_t* found)
{
    if (!ix || k < 1 || qstride < 24 || (nq && (!q || !idx)))
        return pcp_fail(ix ? ix->ctx : nullptr, PCP_ERR_INVALID, "pcp_knn: bad argument");
    pcp_ctx* ctx = ix->ctx;
    if (k > 128) return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_knn: k > 128 is not supported");
    if (nq > 0xfffffff0ull) return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_knn: too many queries in one call");
    if (nq == 0) return PCP_OK;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in;
    OutStage o_idx, o_d2, o_found;
    PCP_TRY(in.open(ctx, q, nq * qstride));
    PCP_TRY(o_idx.open(ctx, idx, nq * (size_t)k * 4));
    PCP_TRY(o_d2.open(ctx, d2, nq * (size_t)k * 8));
    PCP_TRY(o_found.open(ctx, found, nq * 4));
    DevBuf<uint32_t> pos, counter, fail;
    PCP_TRY(pos.alloc(ctx, nq * (size_t)k));
    PCP_TRY(counter.alloc(ctx, 8));
    PCP_CUDA_TRY(ctx, cudaMemsetAsync(counter.p, 0, 32, ctx->stream));
    const int kk = (int)(ix->n < (size_t)k ? ix->n : (size_t)k);
    pcp_wknn_params W;
    memset(&W, 0, sizeof(W));
    W.rec = ix->rec;
    W.ckey = ix->ckey;
    W.cstart = ix->cstart;
    W.ncell = ix->ncell;
    W.g = ix->grid;
    W.htab = ix->htab, W.hmask = ix->hmask, W.ltab = ix->ltab, W.lmask = ix->lmask;
    W.k = k;
    W.kk = kk;
    W.qcount = (uint32_t)nq;
    W.qpts = static_cast<const char*>(in.dptr);
    W.qstride = qstride;
    W.out_pos = pos.p;
    W.out_d2 = (double*)o_d2.dptr;
    W.work_counter = counter.p + 3;
    DevBuf<unsigned long long> dbg;
    if (getenv("PCP_KNN_DEBUG")) {
        PCP_TRY(dbg.alloc(ctx, 17));
        PCP_CUDA_TRY(ctx, cudaMemsetAsync(dbg.p, 0, 17 * 8, ctx->stream));
        W.dbg = dbg.p;
    }
    // Many queries at once (the boundary queries of a neighbouring shard, closest-point batches): the far kernel's
    // pruned search, as for the cloud's own left-over queries -- the streaming kernel scans whole 3x3x3 blocks of level
    // cells and is kept for what the far kernel cannot settle (list overflows on mass ties, non-finite queries).
    static const bool ext_stream_only = getenv("PCP_KNN_STREAM") != nullptr;
    const bool use_far = !ext_stream_only && nq >= 64 && kk > 0 && ix->cinfo && nq * (unsigned long long)k < (1ull << 32);
    bool export_d2 = false;
    {
        StageTimer t(ctx, "knn.query_kernel");
        uint32_t nstream = (uint32_t)nq;
        if (use_far) {
            PCP_TRY(fail.alloc(ctx, nq));
            const uint32_t nq32 = (uint32_t)nq;
            PCP_TRY(pcp_writeto(ctx, counter.p, &nq32, 4));
            pcp_fknn_params P;
            P.rec = ix->rec, P.fpt = ix->fpt, P.htab = ix->htab, P.hmask = ix->hmask, P.ltab = ix->ltab, P.lmask = ix->lmask;
            P.cinfo = ix->cinfo, P.ckey = ix->ckey, P.cstart = ix->cstart, P.ncell = ix->ncell;
            P.item_cell = P.item_first = P.item_meta = nullptr, P.nitem = 0;
            P.g = ix->grid, P.k = k, P.kk = kk, P.eps0 = ix->eps0, P.epsd = ix->epsd, P.query_limit = LLONG_MAX;
            P.kth_d2 = nullptr, P.props = nullptr, P.eig3 = nullptr, P.lps3 = nullptr, P.knn_idx = nullptr;
            P.fail_list = nullptr, P.fail_u = nullptr, P.fail_count = nullptr, P.work_counter = nullptr, P.dbg = nullptr;
            P.nbr_out = pos.p; // rows of sorted positions, one per query
            pcp_far_args A;
            A.list = nullptr, A.listU = nullptr;
            A.count = counter.p;
            A.work_counter = counter.p + 1;
            A.fail_list = fail.p;
            A.fail_count = counter.p + 2;
            A.ext_q = static_cast<const char*>(in.dptr);
            A.ext_stride = qstride;
            PCP_TRY(launch_far(ctx, P, A));
            uint32_t hc[3];
            PCP_TRY(pcp_readback(ctx, hc, counter.p, 12));
            nstream = hc[2];
            W.qsel = fail.p;
            W.qcount = nstream;
            W.out_d2 = nullptr; // the export computes every distance
            export_d2 = o_d2.dptr != nullptr;
        }
        if (nstream) PCP_TRY(run_wknn(ctx, W, nstream));
        knn_export_kernel<<<pcp_blocks_for(nq * (size_t)k, 256), 256, 0, ctx->stream>>>(
            ix->rec, pos.p, nq * (size_t)k, k, (int32_t*)o_idx.dptr, (int32_t*)o_found.dptr,
            export_d2 ? static_cast<const char*>(in.dptr) : nullptr, qstride, export_d2 ? (double*)o_d2.dptr : nullptr);
        PCP_LAUNCH_CHECK(ctx);
    }
    if (W.dbg) {
        unsigned long long h[17];
        PCP_TRY(pcp_readback(ctx, h, dbg.p, 17 * 8));
        fprintf(stderr, "[pcp knn] %zu external queries (%u by the streaming kernel), settled at level:", nq, W.qcount);
        for (int l = 0; l < 16; l++)
            if (h[l]) fprintf(stderr, " %d:%llu", l, h[l]);
        fprintf(stderr, "; %.0f points scanned per streamed query (top level %d)\n", W.qcount ? (double)h[16] / (double)W.qcount : 0.0,
                ix->grid.top_level);
    }
    PCP_TRY(o_idx.commit());
    PCP_TRY(o_d2.commit());
    PCP_TRY(o_found.commit());
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}
