// Scanning simulator, stages 2 and 3: scan lines over the rasterised view -> polylines -> curve cloud.
//
//   ccn_sim_scan_lines     ScanLineGenerator.generate_linear_scan_lines, one attempt of its while loop
//                          (scanning_simulator/utils/scanlines.py:24-76 + sample_line_points :133-144): `ts` samples along each of
//                          L lines per mesh, truncated to pixels; a sample is kept when it is the first of its line on that
//                          pixel (:41-50, np.unique over (u, v, mesh, line)), the pixel was hit by the rasteriser (:53-57) and
//                          the mesh has not reached N points yet (:60-66); kept samples are appended in (line, sample) order.
//   ccn_sim_scan_finish    :78-88 line ids compressed to 0..n_lines-1 per mesh, and the pixel look-up of
//                          SampleMeshes._sample_curves (scanning_simulator/utils/sampling.py:76-80).
//   ccn_sim_curve_split    CurveClouds.compute_3d_curve_idxs (scanning_simulator/utils/curve_cloud.py:40-47).
//   ccn_sim_anti_alias     CurveClouds._anti_alias (curve_cloud.py:168-190): 7-tap filter along every curve, renormalised at the ends.
//   ccn_sim_curvature      estimate_curvature_and_grads + weighted_regression_curvature + solve_weighted_parametric
//                          (scanning_simulator/utils/curvature_approx.py:14-146).
// One CTA per mesh for the order-dependent stages (prefix sums over a mesh's lines / points live in shared memory), one thread
// per point for the filters and the per-point regression.  Integer / index results are bit-exact against the reference's
// python (tests/golden/simulator.pt); the float64 stages agree to summation order.
#include "common.cuh"

namespace {

constexpr int kScanThreads = 1024;
constexpr int kMaxLines = 1024;          // lines per mesh and attempt (the reference uses NLINES = 100)

// pixel of sample k on the segment b0 -> b1: trunc((b1 - b0) * t_k + b0), t_k = k / ts + offset.  f32_math: float32 like
// torch's int64 * float32 promotion for the integer borders of direction 'random'; else float64 (float64 borders of
// 'parallel' / 'grid' promote the product to float64).
__device__ __forceinline__ void line_pixel(const double* __restrict__ bp, int k, int ts, float offset, int f32_math, int& u, int& v) {
    const float tk = __fadd_rn(__fdiv_rn((float)k, (float)ts), offset);
    if (f32_math) {
        const float u0 = (float)bp[0], v0 = (float)bp[1];
        u = (int)__fadd_rn(__fmul_rn((float)(bp[2] - bp[0]), tk), u0);
        v = (int)__fadd_rn(__fmul_rn((float)(bp[3] - bp[1]), tk), v0);
    } else {
        u = (int)__dadd_rn(__dmul_rn(bp[2] - bp[0], (double)tk), bp[0]);
        v = (int)__dadd_rn(__dmul_rn(bp[3] - bp[1], (double)tk), bp[1]);
    }
}

__device__ __forceinline__ bool sample_kept(const double* __restrict__ bp, int k, int ts, float offset, int f32_math, int res,
                                            const uint8_t* __restrict__ mask, int& u, int& v) {
    line_pixel(bp, k, ts, offset, f32_math, u, v);
    if (k > 0) {
        int pu, pv;
        line_pixel(bp, k - 1, ts, offset, f32_math, pu, pv);
        if (pu == u && pv == v) return false;          // the line is monotone in both coordinates: repeats are consecutive
    }
    if ((unsigned)u >= (unsigned)res || (unsigned)v >= (unsigned)res) return false;
    return mask[(int64_t)u * res + v] != 0;
}

// one CTA per mesh: (1) every warp counts the kept samples of its lines, (2) exclusive scan over the lines on top of the
// mesh's running total, (3) the warps walk their lines again and write the kept samples to their final positions (< N)
__global__ void __launch_bounds__(kScanThreads) sim_scan_lines_kernel(const double* __restrict__ border, int L, int ts, float offset,
                                                                      int f32_math, int res, const uint8_t* __restrict__ masks,
                                                                      int N, int line_base, int32_t* __restrict__ npoints,
                                                                      int64_t* __restrict__ uv) {
    ccn_pdl_wait();
    __shared__ int cnt[kMaxLines];
    __shared__ int start[kMaxLines];
    __shared__ int scan_tmp[33];
    __shared__ int base_s;
    const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    const uint8_t* mask = masks + (int64_t)b * res * res;
    if (threadIdx.x == 0) base_s = npoints[b];
    for (int l = warp; l < L; l += nwarp) {
        const double* bp = border + ((int64_t)b * L + l) * 4;
        int c = 0;
        for (int k0 = 0; k0 < ts; k0 += 32) {
            const int k = k0 + lane;
            int u, v;
            const bool keep = k < ts && sample_kept(bp, k, ts, offset, f32_math, res, mask, u, v);
            c += __popc(__ballot_sync(0xffffffffu, keep));
        }
        if (lane == 0) cnt[l] = c;
    }
    __syncthreads();
    const int base = base_s;
    {
        const int mine = threadIdx.x < L ? cnt[threadIdx.x] : 0;
        int total;
        const int ex = ccn_block_excl_scan<int>(mine, scan_tmp, &total);
        if (threadIdx.x < L) start[threadIdx.x] = base + ex;
        if (threadIdx.x == 0) npoints[b] = min(N, base + total);
    }
    __syncthreads();
    for (int l = warp; l < L; l += nwarp) {
        int pos = start[l];
        if (pos >= N || cnt[l] == 0) continue;
        const double* bp = border + ((int64_t)b * L + l) * 4;
        for (int k0 = 0; k0 < ts && pos < N; k0 += 32) {
            const int k = k0 + lane;
            int u = 0, v = 0;
            const bool keep = k < ts && sample_kept(bp, k, ts, offset, f32_math, res, mask, u, v);
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            const int p = pos + __popc(m & ((1u << lane) - 1u));
            if (keep && p < N) {
                int64_t* o = uv + ((int64_t)b * N + p) * 3;
                o[0] = u; o[1] = v; o[2] = line_base + l;
            }
            pos += __popc(m);
        }
    }
}

// one CTA per mesh: line id -> rank among the mesh's lines that own points (they appear in increasing order), pixel look-up
__global__ void __launch_bounds__(kScanThreads) sim_scan_finish_kernel(int64_t* __restrict__ uv, int N, int res,
                                                                       const float* __restrict__ coords, const float* __restrict__ normals,
                                                                       float* __restrict__ points, float* __restrict__ pnormals) {
    ccn_pdl_wait();
    __shared__ int scan_tmp[33];
    __shared__ long long last_raw;          // raw line id of the previous chunk's last point (its slot already holds a rank)
    const int b = blockIdx.x;
    int64_t* o = uv + (int64_t)b * N * 3;
    int carry = 0;
    for (int p0 = 0; p0 < N; p0 += blockDim.x) {
        const int p = p0 + threadIdx.x;
        int first = 0;
        int64_t line = 0;
        if (p < N) {
            line = o[3 * p + 2];
            const int64_t prev = threadIdx.x == 0 ? (p == 0 ? -1 : (int64_t)last_raw) : o[3 * (p - 1) + 2];
            first = prev != line ? 1 : 0;
        }
        int total;
        const int ex = ccn_block_excl_scan<int>(first, scan_tmp, &total);      // ends with a barrier: every read of o[.., 2] above precedes the writes
        if (p < N && threadIdx.x == blockDim.x - 1) last_raw = line;
        if (p < N) {
            o[3 * p + 2] = carry + ex + first - 1;
            if (points) {
                const int64_t pix = ((int64_t)b * res + o[3 * p]) * res + o[3 * p + 1];
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    points[((int64_t)b * N + p) * 3 + c] = coords[3 * pix + c];
                    pnormals[((int64_t)b * N + p) * 3 + c] = normals[3 * pix + c];
                }
            }
        }
        carry += total;
        __syncthreads();
    }
}

// ---- stage 3: CurveClouds ------------------------------------------------------------------------------------------------
// one CTA per mesh: a new curve starts after every edge longer than `thresh` (float64 norm of the float32 points);
// curve_idx = running count of such edges; cstart / cend (per curve id): its point range [cstart, cend) - curves are contiguous
__global__ void __launch_bounds__(kScanThreads) sim_curve_split_kernel(const float* __restrict__ points, int N, double thresh,
                                                                       int32_t* __restrict__ curve_idx, int32_t* __restrict__ cstart,
                                                                       int32_t* __restrict__ cend, int32_t* __restrict__ ncurves) {
    ccn_pdl_wait();
    __shared__ int scan_tmp[33];
    const int b = blockIdx.x;
    const float* P = points + (int64_t)b * N * 3;
    int carry = 0;
    for (int p0 = 0; p0 < N; p0 += blockDim.x) {
        const int p = p0 + threadIdx.x;
        int split = 0;
        if (p < N && p > 0) {
            const double dx = (double)P[3 * p] - (double)P[3 * (p - 1)], dy = (double)P[3 * p + 1] - (double)P[3 * (p - 1) + 1];
            const double dz = (double)P[3 * p + 2] - (double)P[3 * (p - 1) + 2];
            split = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz))) > thresh ? 1 : 0;
        }
        int total;
        const int ex = ccn_block_excl_scan<int>(split, scan_tmp, &total);
        if (p < N) {
            const int c = carry + ex + split;                        // inclusive count: the edge INTO point p
            curve_idx[(int64_t)b * N + p] = c;
            if (p == 0 || split) cstart[(int64_t)b * N + c] = p;
            if (split) cend[(int64_t)b * N + c - 1] = p;
            if (p == N - 1) { cend[(int64_t)b * N + c] = N; ncurves[b] = c + 1; }
        }
        carry += total;
    }
}

// out[i] = sum_k w[k] x[i + k - 3] / sum_k w[k] over the taps that stay inside point i's curve (conv1d 'same' on the zero-padded
// curve divided by the same filter over its 0/1 mask), float64; one thread per (point, component)
__global__ void sim_anti_alias_kernel(const float* __restrict__ vals, const int32_t* __restrict__ curve_idx, int B, int N,
                                      double* __restrict__ out) {
    ccn_pdl_wait();
    const double w[7] = {1.0, 2.0, 3.0, 3.0, 3.0, 2.0, 2.0};       // CurveClouds.ANTI_ALIAS_KERNEL (curve_cloud.py:11)
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= (int64_t)B * N * 3) return;
    const int c = (int)(t % 3);
    const int64_t bp = t / 3;
    const int p = (int)(bp % N);
    const int64_t b0 = bp - p;                                      // first point of the mesh
    const int cid = curve_idx[bp];
    double num = 0.0, den = 0.0;
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        const int q = p + k - 3;
        if (q < 0 || q >= N || curve_idx[b0 + q] != cid) continue;
        num = __dadd_rn(num, __dmul_rn(w[k], (double)vals[(b0 + q) * 3 + c]));
        den = __dadd_rn(den, w[k]);
    }
    out[t] = num / den;
}

// one CTA per mesh: arc length of every point from the start of its curve, mean length of the within-curve edges
__global__ void __launch_bounds__(kScanThreads) sim_geodesic_kernel(const double* __restrict__ points, const int32_t* __restrict__ curve_idx,
                                                                    const int32_t* __restrict__ cstart, int N, double* __restrict__ raw,
                                                                    double* __restrict__ geo, double* __restrict__ mean_edge) {
    ccn_pdl_wait();
    __shared__ double scan_tmp[33];
    __shared__ double red[2][32];
    const int b = blockIdx.x;
    const double* P = points + (int64_t)b * N * 3;
    const int32_t* C = curve_idx + (int64_t)b * N;
    double carry = 0.0, vsum = 0.0;
    int vcnt = 0;
    for (int p0 = 0; p0 < N; p0 += blockDim.x) {
        const int p = p0 + threadIdx.x;
        double len = 0.0;
        if (p < N && p > 0) {
            const double dx = P[3 * p] - P[3 * (p - 1)], dy = P[3 * p + 1] - P[3 * (p - 1) + 1], dz = P[3 * p + 2] - P[3 * (p - 1) + 2];
            len = sqrt(dx * dx + dy * dy + dz * dz);
            if (C[p] == C[p - 1]) { vsum += len; ++vcnt; }
        }
        double total;
        const double ex = ccn_block_excl_scan<double>(len, scan_tmp, &total);
        if (p < N) raw[(int64_t)b * N + p] = carry + ex + len;      // cumulative length up to and including the edge into p
        carry += total;
    }
    __syncthreads();
    // geodesic -= geodesic[curve start]
    for (int p = threadIdx.x; p < N; p += blockDim.x)
        geo[(int64_t)b * N + p] = raw[(int64_t)b * N + p] - raw[(int64_t)b * N + cstart[(int64_t)b * N + C[p]]];
    // mean within-curve edge length
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double cnt = (double)vcnt;
    for (int o = 16; o; o >>= 1) { vsum += __shfl_xor_sync(0xffffffffu, vsum, o); cnt += __shfl_xor_sync(0xffffffffu, cnt, o); }
    if (lane == 0) { red[0][warp] = vsum; red[1][warp] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0, c = 0.0;
        for (int w2 = 0; w2 < (int)(blockDim.x >> 5); ++w2) { s += red[0][w2]; c += red[1][w2]; }
        mean_edge[b] = s / c;                                       // no within-curve edge: NaN, like torch's mean of an empty tensor
    }
}

constexpr int kMaxKnn = 16;

// 3 x 3 system M p = r for three right-hand sides, LU with partial pivoting (torch.linalg.solve = LAPACK gesv)
__device__ __forceinline__ void solve3(double M[3][3], double r[3][3]) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        int piv = c;
        double best = fabs(M[c][c]);
#pragma unroll
        for (int i = c + 1; i < 3; ++i)
            if (fabs(M[i][c]) > best) { best = fabs(M[i][c]); piv = i; }
        if (piv != c) {
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const double a = M[c][j]; M[c][j] = M[piv][j]; M[piv][j] = a;
                const double b = r[c][j]; r[c][j] = r[piv][j]; r[piv][j] = b;
            }
        }
#pragma unroll
        for (int i = c + 1; i < 3; ++i) {
            const double f = M[i][c] / M[c][c];
#pragma unroll
            for (int j = 0; j < 3; ++j) { M[i][j] -= f * M[c][j]; r[i][j] -= f * r[c][j]; }
        }
    }
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        r[2][j] = r[2][j] / M[2][2];
        r[1][j] = (r[1][j] - M[1][2] * r[2][j]) / M[1][1];
        r[0][j] = (r[0][j] - M[0][1] * r[1][j] - M[0][2] * r[2][j]) / M[0][0];
    }
}

__device__ __forceinline__ double nan0(double x) { return x == x ? x : 0.0; }     // remove_nans (curvature_approx.py:150-159)

// one thread per point: k nearest points of its own curve (Euclidean, itself included), Gaussian weights in arc length,
// weighted quadratic fit p(s) = c0 + c1 s + c2 s^2 of the centred neighbours -> curvature vector and d(position)/d(space)
__global__ void __launch_bounds__(128) sim_curvature_kernel(const double* __restrict__ points, const int32_t* __restrict__ curve_idx,
                                                            const int32_t* __restrict__ cstart, const int32_t* __restrict__ cend,
                                                            const double* __restrict__ geo, const double* __restrict__ mean_edge,
                                                            int B, int N, int k, double kernel_width, double alpha,
                                                            double* __restrict__ curvature, double* __restrict__ gradients) {
    ccn_pdl_wait();
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= (int64_t)B * N) return;
    const int b = (int)(t / N), p = (int)(t - (int64_t)b * N);
    const int64_t b0 = (int64_t)b * N;
    const double* P = points + b0 * 3;
    const int cid = curve_idx[t], lo = cstart[b0 + cid], hi = cend[b0 + cid];
    const double px = P[3 * p], py = P[3 * p + 1], pz = P[3 * p + 2];
    // torch_cluster.knn: the k smallest squared distances inside the curve (ties: the lower index first)
    double nd[kMaxKnn];
    int ni[kMaxKnn];
    int cnt = 0;
    for (int q = lo; q < hi; ++q) {
        const double dx = P[3 * q] - px, dy = P[3 * q + 1] - py, dz = P[3 * q + 2] - pz;
        const double d = dx * dx + dy * dy + dz * dz;
        if (cnt == k && !(d < nd[k - 1])) continue;
        int s = cnt < k ? cnt : k - 1;
        while (s > 0 && d < nd[s - 1]) { nd[s] = nd[s - 1]; ni[s] = ni[s - 1]; --s; }
        nd[s] = d; ni[s] = q;
        if (cnt < k) ++cnt;
    }
    const double g0 = geo[t];
    const double scale = kernel_width * mean_edge[b];
    const double s2 = scale * scale;
    // Gaussian weights, normalised over all k slots: the empty slots of a short curve point at point 0 of the mesh
    // (to_batch_padded pads with index 0, point_ops.py:358-381) and take part in the normalisation only
    double wsum = 0.0, w[kMaxKnn], T[kMaxKnn];
    for (int j = 0; j < cnt; ++j) {
        T[j] = geo[b0 + ni[j]] - g0;
        w[j] = exp(-(T[j] * T[j]) / s2);
        wsum += w[j];
    }
    if (cnt < k) {
        const double Tp = geo[b0] - g0;
        wsum += (double)(k - cnt) * exp(-(Tp * Tp) / s2);
    }
    wsum = fmax(wsum, 1e-5);
    double M[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}, r[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    for (int j = 0; j < cnt; ++j) {
        const double wj = w[j] / wsum;
        const double a[3] = {1.0, T[j], T[j] * T[j]};
        const int q = ni[j];
        const double x[3] = {P[3 * q] - px, P[3 * q + 1] - py, P[3 * q + 2] - pz};
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int c = 0; c < 3; ++c) { M[i][c] += a[i] * wj * a[c]; r[i][c] += a[i] * wj * x[c]; }
    }
    // lI = alpha diag(1, 0, 0) + EPS I is a FLOAT32 tensor in the reference (curvature_approx.py:117-121: python float x int64
    // tensor, EPS x torch.eye), added to the float64 normal matrix
    M[0][0] += (double)__fadd_rn((float)alpha, 1e-5f);
    M[1][1] += (double)1e-5f; M[2][2] += (double)1e-5f;
    solve3(M, r);
    int min_pnts = (int)((double)k / 2.5);
    if (min_pnts < 5) min_pnts = 5;
    const double keep = cnt >= min_pnts ? 1.0 : 0.0;
    double vel[3], acc[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        vel[c] = r[1][c] * keep;
        acc[c] = 2.0 * (rint((r[2][c] * keep) * 1e6) / 1e6);        // 2 * torch.round(., decimals=6)
    }
    const double vn = sqrt(vel[0] * vel[0] + vel[1] * vel[1] + vel[2] * vel[2]);
    double tan[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) tan[c] = nan0(vel[c] / vn);
    const double at = acc[0] * tan[0] + acc[1] * tan[1] + acc[2] * tan[2];
    double na[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) na[c] = acc[c] - tan[c] * at;
    const double nn = sqrt(na[0] * na[0] + na[1] * na[1] + na[2] * na[2]);
    const double cx = vel[1] * acc[2] - vel[2] * acc[1], cy = vel[2] * acc[0] - vel[0] * acc[2], cz = vel[0] * acc[1] - vel[1] * acc[0];
    const double cr = sqrt(cx * cx + cy * cy + cz * cz);
    const double kappa = nan0(exp(log(cr) - 3.0 * log(vn)));
#pragma unroll
    for (int c = 0; c < 3; ++c) curvature[t * 3 + c] = nan0(na[c] / nn) * kappa;
    if (gradients) {
#pragma unroll
        for (int f = 0; f < 3; ++f)
#pragma unroll
            for (int c = 0; c < 3; ++c) gradients[t * 9 + f * 3 + c] = nan0(vel[f] * tan[c]);
    }
}

}  // namespace

extern "C" int ccn_sim_scan_lines(const double* border, int32_t B, int32_t L, int32_t ts, float offset, int32_t f32_math,
                                  int32_t res, const uint8_t* masks, int32_t N, int32_t line_base, int32_t* npoints,
                                  int64_t* uv, ccn_stream_t stream) {
    CCN_REQUIRE(B >= 0 && L > 0 && L <= kMaxLines && ts > 0 && res > 0 && N > 0 && line_base >= 0 && border && masks && npoints && uv);
    if (B == 0) return CCN_OK;
    ccn_launch(sim_scan_lines_kernel, dim3(B), dim3(kScanThreads), 0, (cudaStream_t)stream, border, (int)L, (int)ts, offset,
               (int)f32_math, (int)res, masks, (int)N, (int)line_base, npoints, uv);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_sim_scan_finish(int64_t* uv, int32_t B, int32_t N, int32_t res, const float* coords, const float* normals,
                                   float* points_or_null, float* point_normals_or_null, ccn_stream_t stream) {
    CCN_REQUIRE(B >= 0 && N > 0 && res > 0 && uv);
    CCN_REQUIRE(!points_or_null || (coords && normals && point_normals_or_null));
    if (B == 0) return CCN_OK;
    ccn_launch(sim_scan_finish_kernel, dim3(B), dim3(kScanThreads), 0, (cudaStream_t)stream, uv, (int)N, (int)res, coords, normals,
               points_or_null, point_normals_or_null);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_sim_curve_split(const float* points, int32_t B, int32_t N, double thresh, int32_t* curve_idx, int32_t* cstart,
                                   int32_t* cend, int32_t* ncurves, ccn_stream_t stream) {
    CCN_REQUIRE(B >= 0 && N > 0 && points && curve_idx && cstart && cend && ncurves);
    if (B == 0) return CCN_OK;
    ccn_launch(sim_curve_split_kernel, dim3(B), dim3(kScanThreads), 0, (cudaStream_t)stream, points, (int)N, thresh, curve_idx, cstart,
               cend, ncurves);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_sim_anti_alias(const float* vals, const int32_t* curve_idx, int32_t B, int32_t N, double* out, ccn_stream_t stream) {
    CCN_REQUIRE(B >= 0 && N > 0 && vals && curve_idx && out);
    if (B == 0) return CCN_OK;
    ccn_launch(sim_anti_alias_kernel, dim3(ccn_blocks_exact((int64_t)B * N * 3, 256)), dim3(256), 0, (cudaStream_t)stream, vals, curve_idx,
               (int)B, (int)N, out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_sim_curvature(const double* points, const int32_t* curve_idx, const int32_t* cstart, const int32_t* cend, int32_t B,
                                 int32_t N, int32_t k, double kernel_width, double hinge_reg, double* geo_raw_scratch, double* geo_scratch,
                                 double* mean_edge_scratch, double* curvature, double* gradients_or_null, ccn_stream_t stream) {
    CCN_REQUIRE(B >= 0 && N > 0 && k >= 1 && k <= kMaxKnn && points && curve_idx && cstart && cend && geo_raw_scratch && geo_scratch &&
                mean_edge_scratch && curvature);
    if (B == 0) return CCN_OK;
    cudaStream_t s = (cudaStream_t)stream;
    ccn_launch(sim_geodesic_kernel, dim3(B), dim3(kScanThreads), 0, s, points, curve_idx, cstart, (int)N, geo_raw_scratch, geo_scratch,
               mean_edge_scratch);
    CCN_CHECK_LAUNCH();
    ccn_launch(sim_curvature_kernel, dim3(ccn_blocks_exact((int64_t)B * N, 128)), dim3(128), 0, s, points, curve_idx, cstart, cend,
               (const double*)geo_scratch, (const double*)mean_edge_scratch, (int)B, (int)N, (int)k, kernel_width, hinge_reg, curvature,
               gradients_or_null);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// CurveClouds.compute_intersections (scanning_simulator/utils/curve_cloud.py:114-166) + line_seg_dists (:285-331), ONE curve
// cloud (the reference's second scatter_min only broadcasts for B = 1, which is how generate_shapenet.py calls it):
// every pair of polyline edges (i, j): mua_ij = parameter on edge i of the point closest to the LINE through edge j (0 for
// parallel pairs, clipped to [0, 1]); dist(i, j) = |closest_ij - closest_ji|; pairs with an edge that joins two curves or
// with both edges on one curve count as 1.0 (= 100 * thresh).  min over the rows of a curve, then over the columns of a
// curve (lowest index among equals) -> curve pairs closer than thresh -> the two edges -> the nearer end point of each.
// All in f64 like upstream, products and sums rounded separately.
// ---------------------------------------------------------------------------------------------------------------------
namespace {

__device__ __forceinline__ double dot3_rn(double ax, double ay, double az, double bx, double by, double bz) {
    return __dadd_rn(__dadd_rn(__dmul_rn(ax, bx), __dmul_rn(ay, by)), __dmul_rn(az, bz));
}

// mua of edge a against the line of edge b, and the closest point on a
__device__ __forceinline__ double seg_mua(const double* __restrict__ p, int a, int b, double* cx, double* cy, double* cz) {
    const double p1x = p[3 * a], p1y = p[3 * a + 1], p1z = p[3 * a + 2];
    const double ux = __dsub_rn(p[3 * a + 3], p1x), uy = __dsub_rn(p[3 * a + 4], p1y), uz = __dsub_rn(p[3 * a + 5], p1z);
    const double p3x = p[3 * b], p3y = p[3 * b + 1], p3z = p[3 * b + 2];
    const double vx = __dsub_rn(p[3 * b + 3], p3x), vy = __dsub_rn(p[3 * b + 4], p3y), vz = __dsub_rn(p[3 * b + 5], p3z);
    const double wx = __dsub_rn(p1x, p3x), wy = __dsub_rn(p1y, p3y), wz = __dsub_rn(p1z, p3z);
    const double d1343 = dot3_rn(wx, wy, wz, vx, vy, vz), d4321 = dot3_rn(ux, uy, uz, vx, vy, vz);
    const double d1321 = dot3_rn(wx, wy, wz, ux, uy, uz), d4343 = dot3_rn(vx, vy, vz, vx, vy, vz);
    const double d2121 = dot3_rn(ux, uy, uz, ux, uy, uz);
    const double denom = __dsub_rn(__dmul_rn(d2121, d4343), __dmul_rn(d4321, d4321));
    const double num = __dsub_rn(__dmul_rn(d1343, d4321), __dmul_rn(d1321, d4343));
    double mua = denom < 1e-8 ? 0.0 : __ddiv_rn(num, denom);
    mua = fmin(fmax(mua, 0.0), 1.0);
    *cx = __dadd_rn(p1x, __dmul_rn(mua, ux));
    *cy = __dadd_rn(p1y, __dmul_rn(mua, uy));
    *cz = __dadd_rn(p1z, __dmul_rn(mua, uz));
    return mua;
}

__device__ __forceinline__ double seg_pair_dist(const double* __restrict__ p, const int32_t* __restrict__ cidx, int i, int j,
                                                double masked) {
    if (cidx[i + 1] != cidx[i] || cidx[j + 1] != cidx[j] || cidx[i] == cidx[j]) return masked;
    double ax, ay, az, bx, by, bz;
    seg_mua(p, i, j, &ax, &ay, &az);
    seg_mua(p, j, i, &bx, &by, &bz);
    const double dx = __dsub_rn(ax, bx), dy = __dsub_rn(ay, by), dz = __dsub_rn(az, bz);
    return sqrt(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz)));
}

// cptr[g] = first point of curve g (curve ids are 0, 1, 2, ... along the points); cptr[n_curves] = n
__global__ void sim_curve_ptr_kernel(const int32_t* __restrict__ cidx, int n, int n_curves, int32_t* __restrict__ cptr) {
    ccn_pdl_wait();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        if (i == 0 || cidx[i] != cidx[i - 1]) cptr[cidx[i]] = i;
    if (blockIdx.x == 0 && threadIdx.x == 0) cptr[n_curves] = n;
}

// (g, j): min over the edges i of curve g of dist(i, j), lowest i among equals (scatter_min over the rows, :145)
__global__ void sim_isect_rows_kernel(const double* __restrict__ p, const int32_t* __restrict__ cidx,
                                      const int32_t* __restrict__ cptr, int n_edges, int n_groups, double masked,
                                      double* __restrict__ min1, int32_t* __restrict__ arg1) {
    ccn_pdl_wait();
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= (int64_t)n_groups * n_edges) return;
    const int g = (int)(t / n_edges), j = (int)(t - (int64_t)g * n_edges);
    const int lo = cptr[g], hi = min(cptr[g + 1], n_edges);          // edges of curve g = its points that have a successor
    double best = 0.0;
    int at = -1;
    for (int i = lo; i < hi; ++i) {
        const double d = seg_pair_dist(p, cidx, i, j, masked);
        if (at < 0 || d < best) { best = d; at = i; }
    }
    min1[t] = best;
    arg1[t] = at;
}

// (g, h): min over the edges j of curve h of min1[g][j], lowest j among equals (scatter_min over the columns, :146);
// flag = closer than thresh
__global__ void sim_isect_cols_kernel(const double* __restrict__ min1, const int32_t* __restrict__ cptr, int n_edges,
                                      int n_groups, double thresh, int32_t* __restrict__ arg2, int32_t* __restrict__ flag,
                                      int32_t* __restrict__ slot) {
    ccn_pdl_wait();
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_groups * n_groups) return;
    const int g = t / n_groups, h = t - g * n_groups;
    const int lo = cptr[h], hi = min(cptr[h + 1], n_edges);
    const bool g_has_edges = cptr[g] < min(cptr[g + 1], n_edges);    // a last curve of one point owns no edge: not a group upstream
    double best = 0.0;
    int at = -1;
    for (int j = lo; j < hi; ++j) {
        const double d = min1[(int64_t)g * n_edges + j];
        if (at < 0 || d < best) { best = d; at = j; }
    }
    arg2[t] = at;
    const int f = (g_has_edges && at >= 0 && best < thresh) ? 1 : 0;
    flag[t] = f;
    slot[t] = f;
}

// curve pair k (row-major over (g, h)): the two edges, each moved to its end point when the closest approach lies in the
// edge's second half (:153-160; the second look-up uses the already shifted first index, as upstream)
__global__ void sim_isect_fill_kernel(const double* __restrict__ p, const int32_t* __restrict__ arg1,
                                      const int32_t* __restrict__ arg2, const int32_t* __restrict__ flag,
                                      const int32_t* __restrict__ slot, int n_edges, int n_groups, int64_t n_out,
                                      int64_t* __restrict__ pairs, int32_t* __restrict__ err) {
    ccn_pdl_wait();
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_groups * n_groups || !flag[t]) return;
    const int g = t / n_groups;
    int e2 = arg2[t];
    int e1 = arg1[(int64_t)g * n_edges + e2];
    double cx, cy, cz;
    if (seg_mua(p, e1, e2, &cx, &cy, &cz) > 0.5) e1 += 1;
    if (e1 >= n_edges) { atomicOr(err, 1); e1 = n_edges - 1; }      // upstream indexes past the matrix here and raises
    if (seg_mua(p, e2, e1, &cx, &cy, &cz) > 0.5) e2 += 1;
    const int k = slot[t];
    pairs[k] = e1;
    pairs[n_out + k] = e2;
}

}  // namespace

extern "C" int ccn_sim_intersections_count(const double* points, const int32_t* curve_idx, int32_t n, int32_t n_curves,
                                           double thresh, int32_t* cptr_scratch, double* min1_scratch, int32_t* arg1_scratch,
                                           int32_t* arg2_scratch, int32_t* flag_scratch, int32_t* slot_scan,
                                           ccn_stream_t stream) {
    // scratch: cptr n_curves + 1 ints; min1 / arg1 n_curves * (n - 1); arg2 / flag n_curves^2; slot_scan n_curves^2 + 1 ints:
    // on return the exclusive scan of the pair flags, the number of intersections in its last entry
    CCN_REQUIRE(n >= 2 && n_curves >= 1 && points && curve_idx && cptr_scratch && min1_scratch && arg1_scratch && arg2_scratch &&
                flag_scratch && slot_scan);
    cudaStream_t st = (cudaStream_t)stream;
    const int n_edges = n - 1;
    ccn_launch(sim_curve_ptr_kernel, dim3(ccn_grid(n, 256, 148 * 4)), dim3(256), 0, st, curve_idx, (int)n, (int)n_curves, cptr_scratch);
    CCN_CHECK_LAUNCH();
    const int64_t work = (int64_t)n_curves * n_edges;
    ccn_launch(sim_isect_rows_kernel, dim3(ccn_blocks_exact(work, 128)), dim3(128), 0, st, points, curve_idx,
               (const int32_t*)cptr_scratch, n_edges, (int)n_curves, 100.0 * thresh, min1_scratch, arg1_scratch);
    CCN_CHECK_LAUNCH();
    ccn_launch(sim_isect_cols_kernel, dim3(ccn_blocks_exact((int64_t)n_curves * n_curves, 128)), dim3(128), 0, st,
               (const double*)min1_scratch, (const int32_t*)cptr_scratch, n_edges, (int)n_curves, thresh, arg2_scratch, flag_scratch,
               slot_scan);
    CCN_CHECK_LAUNCH();
    return ccn_excl_scan_i32(slot_scan, (int64_t)n_curves * n_curves, nullptr, stream);
}

extern "C" int ccn_sim_intersections_fill(const double* points, int32_t n, int32_t n_curves, const int32_t* arg1_scratch,
                                          const int32_t* arg2_scratch, const int32_t* flag_scratch, const int32_t* slot_scan,
                                          int64_t n_out, int64_t* pairs_out, int32_t* err_out, ccn_stream_t stream) {
    // pairs_out (2, n_out) int64: point indices (first row: on the row curve, second: on the column curve), curve pairs in
    // row-major order; err_out: 1 when a shifted first index left the edge range (upstream raises an IndexError there)
    CCN_REQUIRE(n >= 2 && n_curves >= 1 && n_out >= 0 && points && arg1_scratch && arg2_scratch && flag_scratch && slot_scan &&
                err_out);
    cudaError_t e = cudaMemsetAsync(err_out, 0, sizeof(int32_t), (cudaStream_t)stream);
    if (e != cudaSuccess) return (int)e;
    if (n_out == 0) return CCN_OK;
    CCN_REQUIRE(pairs_out);
    ccn_launch(sim_isect_fill_kernel, dim3(ccn_blocks_exact((int64_t)n_curves * n_curves, 128)), dim3(128), 0, (cudaStream_t)stream,
               points, arg1_scratch, arg2_scratch, flag_scratch, slot_scan, (int)(n - 1), (int)n_curves, n_out, pairs_out, err_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// Label transfer of the dataset generators (scanning_simulator/shapenet_seg/generate_shapenet.py:88-89,
// kortx/generate_shapenet.py: `argmin(cdist(points.double(), pos.double()), dim=1)`, then `y[closest]`): for every sampled
// point the nearest labelled point in f64 (lowest index among equals) and its label.  Warp per query, the labelled cloud
// streamed by the lanes; squared distances (monotone in the distance cdist compares).
// ---------------------------------------------------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256) sim_nearest_kernel(const double* __restrict__ q, int64_t nq, const double* __restrict__ r,
                                                          int64_t nr, const int64_t* __restrict__ labels,
                                                          int64_t* __restrict__ idx_out, int64_t* __restrict__ label_out) {
    ccn_pdl_wait();
    const int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (w >= nq) return;
    const int lane = ccn_lane();
    const double x = q[3 * w], y = q[3 * w + 1], z = q[3 * w + 2];
    double best = 1e300;
    long long at = 0x7fffffffffffffffll;
    for (int64_t j = lane; j < nr; j += 32) {
        const double dx = x - r[3 * j], dy = y - r[3 * j + 1], dz = z - r[3 * j + 2];
        const double d = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
        if (d < best) { best = d; at = j; }                  // ascending j per lane: the first minimum stays
    }
    for (int o = 16; o; o >>= 1) {
        const double ob = __shfl_xor_sync(0xffffffffu, best, o);
        const long long oa = __shfl_xor_sync(0xffffffffu, at, o);
        if (ob < best || (ob == best && oa < at)) { best = ob; at = oa; }
    }
    if (lane == 0) {
        idx_out[w] = at;
        if (label_out) label_out[w] = labels[at];
    }
}
}  // namespace

extern "C" int ccn_sim_nearest_label(const double* points, int64_t nq, const double* ref_points, int64_t nr,
                                     const int64_t* ref_labels, int64_t* idx_out, int64_t* label_out, ccn_stream_t stream) {
    CCN_REQUIRE(nq >= 0 && nr >= 1 && ref_points && idx_out && (!label_out || ref_labels));
    if (nq == 0) return CCN_OK;
    CCN_REQUIRE(points);
    ccn_launch(sim_nearest_kernel, dim3(ccn_blocks_exact(nq * 32, 256)), dim3(256), 0, (cudaStream_t)stream, points, nq, ref_points, nr,
               ref_labels, idx_out, label_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
