// Curve grouping of a raw sweep (reference src/data/nuscenes_dataset.py:91-118 ==
// src/data/kitti_dataset.py:73-92 `_get_curves`): stable sort of the points by beam id, a new curve wherever the beam
// changes or the edge to the previous point is longer than thresh * sqrt(|p.xy|), curve id = running count of splits.
// The reference does this per sample on the CPU inside the DataLoader (torch.sort + boolean arithmetic + cumsum).
// Here: counting sort by beam (one CTA per beam compacts its points in acquisition order = stable), one element-wise
// split kernel (edge length in f64 like the reference, radius and threshold in correctly rounded f32), one scan.
#include "common.cuh"

namespace {

__global__ void beam_hist_kernel(const int64_t* __restrict__ beam, int64_t n, int n_beams, int32_t* __restrict__ count,
                                 int32_t* __restrict__ flags) {
    ccn_pdl_wait();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = beam[i];
        if (b < 0 || b >= n_beams) { atomicOr(flags, 1); continue; }
        atomicAdd(count + b, 1);
    }
}

// CTA b: positions of the points of beam b, in acquisition order, start at sum(count[< b]).
__global__ void __launch_bounds__(1024) beam_rank_kernel(const int64_t* __restrict__ beam, int64_t n,
                                                         const int32_t* __restrict__ count, int64_t* __restrict__ reorder,
                                                         int64_t* __restrict__ inv) {
    ccn_pdl_wait();
    __shared__ int scan_smem[33];
    __shared__ int base_s;
    const int b = blockIdx.x;
    if (count[b] == 0) return;
    int part = 0;
    for (int j = threadIdx.x; j < b; j += blockDim.x) part += count[j];
    int total;
    ccn_block_excl_scan<int>(part, scan_smem, &total);
    if (threadIdx.x == 0) base_s = total;
    __syncthreads();
    int running = base_s;
    for (int64_t i0 = 0; i0 < n; i0 += blockDim.x) {
        const int64_t i = i0 + threadIdx.x;
        const int mine = (i < n && beam[i] == b) ? 1 : 0;
        int tot;
        const int ex = ccn_block_excl_scan<int>(mine, scan_smem, &tot);
        if (mine) {
            const int64_t pos = running + ex;
            reorder[pos] = i;
            inv[i] = pos;
        }
        running += tot;
    }
}

// single-beam sweeps (KITTI): the stable sort is the identity
__global__ void iota_i64_kernel(int64_t n, int64_t* __restrict__ a, int64_t* __restrict__ b) {
    ccn_pdl_wait();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        a[i] = i;
        b[i] = i;
    }
}

// sorted_points[j] = points[reorder[j]]; split[j] = beam changes or |p_j - p_{j-1}| (f64) > thresh * sqrt(|p_j.xy|) (f32)
__global__ void curve_split_kernel(const float* __restrict__ points, int64_t ldp, const int64_t* __restrict__ beam,
                                   const int64_t* __restrict__ reorder, int64_t n, float thresh,
                                   float* __restrict__ sorted_points, int32_t* __restrict__ split) {
    ccn_pdl_wait();
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = reorder[j];
        const float x = points[i * ldp], y = points[i * ldp + 1], z = points[i * ldp + 2];
        sorted_points[3 * j] = x; sorted_points[3 * j + 1] = y; sorted_points[3 * j + 2] = z;
        int s = 0;
        if (j > 0) {
            const int64_t q = reorder[j - 1];
            if (beam[i] != beam[q]) {
                s = 1;
            } else {
                const double ex = (double)x - (double)points[q * ldp], ey = (double)y - (double)points[q * ldp + 1],
                             ez = (double)z - (double)points[q * ldp + 2];
                const double en = sqrt(fma(ez, ez, fma(ey, ey, ex * ex)));
                const float rad = __fsqrt_rn(__fmaf_rn(y, y, __fmul_rn(x, x)));       // torch.linalg.norm(p[:, :2])
                const float thr = __fmul_rn(thresh, __fsqrt_rn(rad));
                s = en > (double)thr ? 1 : 0;
            }
        }
        split[j] = s;
    }
}

// curve_idxs[j] = inclusive running sum of split (single CTA, chunked block scan); *n_curves = last + 1
__global__ void __launch_bounds__(1024) curve_cumsum_kernel(const int32_t* __restrict__ split, int64_t n,
                                                            int64_t* __restrict__ cid, int32_t* __restrict__ n_curves) {
    ccn_pdl_wait();
    __shared__ int scan_smem[33];
    int running = 0;
    for (int64_t i0 = 0; i0 < n; i0 += blockDim.x) {
        const int64_t i = i0 + threadIdx.x;
        const int v = i < n ? split[i] : 0;
        int tot;
        const int ex = ccn_block_excl_scan<int>(v, scan_smem, &tot);
        if (i < n) cid[i] = (int64_t)(running + ex + v);
        running += tot;
    }
    if (threadIdx.x == 0 && n_curves) *n_curves = n > 0 ? running + 1 : 0;
}

}  // namespace

extern "C" int ccn_get_curves(const float* points, int64_t ldp, const int64_t* beam, int64_t n, int32_t n_beams,
                              float curve_thresh, int32_t* scratch, int64_t* reorder, int64_t* inv_reorder,
                              float* sorted_points, int64_t* curve_idxs, int32_t* counts_out, ccn_stream_t stream) {
    // scratch: int32 [n_beams + n]; counts_out: int32 [2] = {number of curves, error flags (1: beam id out of range)}
    CCN_REQUIRE(n >= 0 && n < 0x7fffffff && n_beams > 0 && n_beams <= 65535 && ldp >= 3);
    CCN_REQUIRE(points && beam && scratch && reorder && inv_reorder && sorted_points && curve_idxs && counts_out);
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(scratch, 0, (size_t)n_beams * sizeof(int32_t), s);
    if (e == cudaSuccess) e = cudaMemsetAsync(counts_out, 0, 2 * sizeof(int32_t), s);
    if (e != cudaSuccess) return (int)e;
    if (n == 0) return CCN_OK;
    int32_t* count = scratch;
    int32_t* split = scratch + n_beams;
    ccn_launch(beam_hist_kernel, dim3(ccn_grid(n, 256, 148 * 8)), dim3(256), 0, s, beam, n, n_beams, count, counts_out + 1);
    CCN_CHECK_LAUNCH();
    if (n_beams == 1) ccn_launch(iota_i64_kernel, dim3(ccn_grid(n, 256, 148 * 8)), dim3(256), 0, s, n, reorder, inv_reorder);
    else ccn_launch(beam_rank_kernel, dim3(n_beams), dim3(1024), 0, s, beam, n, count, reorder, inv_reorder);
    CCN_CHECK_LAUNCH();
    ccn_launch(curve_split_kernel, dim3(ccn_grid(n, 256, 148 * 8)), dim3(256), 0, s, points, ldp, beam, reorder, n, curve_thresh,
                                                                  sorted_points, split);
    CCN_CHECK_LAUNCH();
    ccn_launch(curve_cumsum_kernel, dim3(1), dim3(1024), 0, s, split, n, curve_idxs, counts_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
