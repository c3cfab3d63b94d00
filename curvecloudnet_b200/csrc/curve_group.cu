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

// ---------------------------------------------------------------------------------------------------------------------
// Kortx scans (scanning_simulator/kortx/generate_kortx.py:38-75, 94-117): a long structured-light scan (f64 xyz, a raw curve
// counter per point) -> `fix_curve_idxs` (a curve start directly followed by another start is dropped, ids = running count
// of the starts that remain) -> `cut_curves` (new curve where the id grows or the f64 edge is longer than thresh, ids from 0)
// -> `samples_per_scan` windows of npoints consecutive points, each scaled by 1 / (2 max |coordinate|) with its curve ids
// renumbered from 0 (torch.unique of a non-decreasing sequence).  The reference does this once per file on the CPU.
// ---------------------------------------------------------------------------------------------------------------------
namespace {

__global__ void kortx_start_kernel(const int64_t* __restrict__ raw, int64_t n, int32_t* __restrict__ start) {
    ccn_pdl_wait();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const bool e = i == 0 || raw[i] - raw[i - 1] > 0;
        const bool e_next = i + 1 < n && raw[i + 1] - raw[i] > 0;
        start[i] = (e && !e_next) ? 1 : 0;                     // adjacent end points: the first of the two is dropped (:100-102)
    }
}

// inclusive running sum of v[0..n) (single CTA, chunked): out[i] = sum v[0..i]
__global__ void __launch_bounds__(1024) incl_scan_i32_kernel(const int32_t* __restrict__ v, int64_t n, int32_t* __restrict__ out) {
    ccn_pdl_wait();
    __shared__ int scan_smem[33];
    int running = 0;
    for (int64_t i0 = 0; i0 < n; i0 += blockDim.x) {
        const int64_t i = i0 + threadIdx.x;
        const int x = i < n ? v[i] : 0;
        int tot;
        const int ex = ccn_block_excl_scan<int>(x, scan_smem, &tot);
        if (i < n) out[i] = running + ex + x;
        running += tot;
    }
}

__global__ void kortx_cut_kernel(const double* __restrict__ p, const int32_t* __restrict__ fixed, int64_t n, double thresh,
                                 int32_t* __restrict__ cut) {
    ccn_pdl_wait();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int c = 0;
        if (i > 0) {
            const double ex = p[3 * i] - p[3 * i - 3], ey = p[3 * i + 1] - p[3 * i - 2], ez = p[3 * i + 2] - p[3 * i - 1];
            const double len = sqrt(fma(ez, ez, fma(ey, ey, ex * ex)));
            c = (fixed[i] - fixed[i - 1] > 0 || len > thresh) ? 1 : 0;
        }
        cut[i] = c;
    }
}

__global__ void i32_to_i64_kernel(const int32_t* __restrict__ a, int64_t n, int64_t* __restrict__ out) {
    ccn_pdl_wait();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = a[i];
}

// CTA w: window of npoints consecutive points starting at w * interval
__global__ void __launch_bounds__(256) kortx_window_kernel(const double* __restrict__ p, const int64_t* __restrict__ curve,
                                                           const int64_t* __restrict__ seg, int npoints, int64_t interval,
                                                           double* __restrict__ pos_out, int64_t* __restrict__ curve_out,
                                                           int64_t* __restrict__ seg_out) {
    ccn_pdl_wait();
    __shared__ double smax[8];
    __shared__ int scan_smem[33];
    const int w = blockIdx.x;
    const int64_t base = w * interval;
    double m = 0.0;
    for (int t = threadIdx.x; t < 3 * npoints; t += blockDim.x) m = fmax(m, fabs(p[3 * base + t]));
    for (int o = 16; o; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (ccn_lane() == 0) smax[ccn_warp()] = m;
    __syncthreads();
    m = smax[0];
    for (int k = 1; k < (int)(blockDim.x >> 5); ++k) m = fmax(m, smax[k]);
    const double div = m * 2.0;                                                      // max |coordinate| * 2 (:58)
    for (int t = threadIdx.x; t < 3 * npoints; t += blockDim.x) pos_out[(int64_t)w * 3 * npoints + t] = p[3 * base + t] / div;
    int running = 0;
    for (int i0 = 0; i0 < npoints; i0 += blockDim.x) {
        const int i = i0 + threadIdx.x;
        const int ch = (i < npoints && i > 0 && curve[base + i] != curve[base + i - 1]) ? 1 : 0;
        int tot;
        const int ex = ccn_block_excl_scan<int>(ch, scan_smem, &tot);
        if (i < npoints) {
            curve_out[(int64_t)w * npoints + i] = running + ex + ch;
            if (seg_out) seg_out[(int64_t)w * npoints + i] = seg[base + i];
        }
        running += tot;
    }
}

}  // namespace

extern "C" int ccn_kortx_curves(const double* points, const int64_t* raw_curve, int64_t n, double cut_thresh,
                                int32_t* scratch, int64_t* curve_out, ccn_stream_t stream) {
    // scratch: int32 [2 n]
    CCN_REQUIRE(n >= 0 && n < 0x7fffffff);
    if (n == 0) return CCN_OK;
    CCN_REQUIRE(points && raw_curve && scratch && curve_out);
    cudaStream_t s = (cudaStream_t)stream;
    int32_t* a = scratch;
    int32_t* b = scratch + n;
    const unsigned g = ccn_grid(n, 256, 148 * 8);
    ccn_launch(kortx_start_kernel, dim3(g), dim3(256), 0, s, raw_curve, n, a);
    CCN_CHECK_LAUNCH();
    ccn_launch(incl_scan_i32_kernel, dim3(1), dim3(1024), 0, s, (const int32_t*)a, n, b);       // b = fix_curve_idxs
    CCN_CHECK_LAUNCH();
    ccn_launch(kortx_cut_kernel, dim3(g), dim3(256), 0, s, points, (const int32_t*)b, n, cut_thresh, a);
    CCN_CHECK_LAUNCH();
    ccn_launch(incl_scan_i32_kernel, dim3(1), dim3(1024), 0, s, (const int32_t*)a, n, b);       // b = cut_curves (cut[0] = 0)
    CCN_CHECK_LAUNCH();
    ccn_launch(i32_to_i64_kernel, dim3(g), dim3(256), 0, s, (const int32_t*)b, n, curve_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_kortx_windows(const double* points, const int64_t* curve, const int64_t* segmentation, int64_t n,
                                 int32_t npoints, int32_t n_windows, double* pos_out, int64_t* curve_out, int64_t* seg_out,
                                 ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && npoints >= 1 && n_windows >= 1 && n >= npoints);
    CCN_REQUIRE(points && curve && pos_out && curve_out && (!seg_out || segmentation));
    const int64_t interval = (n - npoints) / n_windows;                                 // generate_kortx.py:40
    ccn_launch(kortx_window_kernel, dim3((unsigned)n_windows), dim3(256), 0, (cudaStream_t)stream, points, curve, segmentation,
               (int)npoints, interval, pos_out, curve_out, seg_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
