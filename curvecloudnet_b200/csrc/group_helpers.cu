// Two grouping helpers of the reference's op API that no shipped configuration reaches (last row):
//   ccn_knn_1d_group       knn_1d_group_subset   (src/models/utils/point_ops.py:114-140)
//   ccn_group_first_k      group_first_k_values  (src/models/utils/point_ops.py:384-428; its one caller is the scanning
//                                                 simulator's padded curve layout, curve_cloud.py:223)
// Both are index bookkeeping: one thread per query / one CTA per group, no atomics, results in the reference's order.
#include "common.cuh"

namespace {

// Neighbours along the point ORDER: candidates c, c-1, c+1, c-2, c+2, ... (point_ops.py:119-123), kept while they are inside
// the array and on the query's curve (global curve id), at most knn of them - the query itself counts (cumsum <= knn, :133).
__global__ void knn_1d_group_kernel(const int64_t* __restrict__ glob, int64_t n, const int64_t* __restrict__ idxs,
                                    int64_t nq, int knn, int32_t* __restrict__ idx_out, int32_t* __restrict__ cnt_out) {
    ccn_pdl_wait();
    const int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const int W = 2 * knn + 1;
    const int64_t c = idxs[q];
    int32_t* row = idx_out + q * W;
    int kept = 0;
    if (c >= 0 && c < n) {
        const int64_t gc = glob[c];
        for (int s = 0; s < W && kept < knn; ++s) {
            const int64_t off = s == 0 ? 0 : ((s & 1) ? -(int64_t)((s + 1) >> 1) : (int64_t)(s >> 1));
            const int64_t j = c + off;
            if (j >= 0 && j < n && glob[j] == gc) row[kept++] = (int32_t)j;
        }
    }
    for (int s = kept; s < W; ++s) row[s] = -1;
    cnt_out[q] = kept;
}

// CTA g walks the whole id vector in order and keeps the first kmax rows whose id is g (a stable selection: the order of
// torch.sort(batch, stable=True), point_ops.py:406); rows it does not fill are zeroed; count_out[g] = min(#rows of g, kmax).
constexpr int kGroupThreads = 256;

__global__ void __launch_bounds__(kGroupThreads) group_first_k_kernel(
    const float* __restrict__ values, int64_t ldv, const int64_t* __restrict__ batch, int64_t n, int d, int kmax,
    float* __restrict__ out, int32_t* __restrict__ count_out) {
    ccn_pdl_wait();
    __shared__ int warp_total[kGroupThreads / 32];
    const int g = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float* dst = out + (int64_t)g * kmax * d;
    int taken = 0;
    for (int64_t i0 = 0; i0 < n && taken < kmax; i0 += kGroupThreads) {
        const int64_t i = i0 + tid;
        const bool hit = i < n && batch[i] == (int64_t)g;
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        if (lane == 0) warp_total[warp] = __popc(m);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kGroupThreads / 32; ++w) {
            before += w < warp ? warp_total[w] : 0;
            total += warp_total[w];
        }
        const int pos = taken + before + __popc(m & ((1u << lane) - 1u));
        if (hit && pos < kmax)
            for (int c = 0; c < d; ++c) dst[(int64_t)pos * d + c] = values[i * ldv + c];
        taken += total;
        __syncthreads();
    }
    taken = min(taken, kmax);
    for (int64_t t = (int64_t)taken * d + tid; t < (int64_t)kmax * d; t += kGroupThreads) dst[t] = 0.f;
    if (tid == 0) count_out[g] = taken;
}

}  // namespace

extern "C" int ccn_knn_1d_group(const int64_t* curve_glob, int64_t n, const int64_t* idxs, int64_t nq, int32_t knn,
                                int32_t* idx_out, int32_t* cnt_out, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && nq >= 0 && knn >= 1 && idx_out && cnt_out);
    if (nq == 0) return CCN_OK;
    CCN_REQUIRE(curve_glob && idxs);
    ccn_launch(knn_1d_group_kernel, dim3(ccn_blocks_exact(nq, 256)), dim3(256), 0, (cudaStream_t)stream, curve_glob, n, idxs, nq,
               (int)knn, idx_out, cnt_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_group_first_k(const float* values, int64_t ldv, const int64_t* batch, int64_t n, int32_t d,
                                 int32_t n_groups, int32_t kmax, float* out, int32_t* count_out, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && d >= 1 && ldv >= d && n_groups >= 0 && kmax >= 1 && out && count_out);
    if (n_groups == 0) return CCN_OK;
    CCN_REQUIRE(n == 0 || (values && batch));
    ccn_launch(group_first_k_kernel, dim3((unsigned)n_groups), dim3(kGroupThreads), 0, (cudaStream_t)stream, values, ldv, batch, n,
               (int)d, (int)kmax, out, count_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
