// Training-step kernels: the backward of every op on the CurveCloudNet path plus train-mode BatchNorm
// (SURVEY.md §8(b) "curve_conv1d_sym_bwd_*, feature_diffs_bwd, gather_mlp_aggregate_bwd, interp_invdist_bwd,
// mlp_bwd, bn_stats, segment_max_bwd").  The reference gets these from torch autograd over cuBLAS / cuDNN /
// torch_scatter / pytorch3d kernels (src/run/*_seg.py `loss.backward()`); here each is one HBM-bound kernel:
//   * row gathers / scatter-adds with int32 maps (-1 = zero row / skipped row)
//   * Linear weight gradient dW = dY^T X (+ column sums for the bias), split over the row axis
//   * BatchNorm batch statistics (one pass, fp64 accumulators), normalise + activation, and their backward
//   * segmented max / mean / sigmoid-weighted / softmax-attention aggregation backward
//   * dense (K+1)-slot max backward, inverse-distance interpolation backward, curve feature-difference backward
//   * log-softmax + NLL loss forward/backward.
#include <cuda_bf16.h>

#include "common.cuh"

namespace {

// optional second output of the BatchNorm kernels: the value as split-bf16 activation planes (gemm_planes.cu)
struct PlanesOut {
    __nv_bfloat16* p;      // null: no planes
    int64_t kpad, stride;  // row stride / plane stride in elements
    int n;                 // 1 = bf16, 2 = hi + lo
};
__device__ __forceinline__ void st_plane(const PlanesOut& po, int64_t r, int c, float v) {
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    po.p[r * po.kpad + c] = h;
    if (po.n == 2) po.p[po.stride + r * po.kpad + c] = __float2bfloat16_rn(v - __bfloat162float(h));
}

#define ELT_GRID(total) ccn_grid((total), 256, 148LL * 16)
#define GRID_STRIDE(t, total) \
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < (total); t += (int64_t)gridDim.x * blockDim.x)

__device__ __forceinline__ float act_grad(float pre, int act, float slope) {
    // d act(pre) / d pre as torch defines it (relu'(0) = 0, leaky_relu'(0) = slope)
    if (act == CCN_ACT_RELU) return pre > 0.f ? 1.f : 0.f;
    if (act == CCN_ACT_LEAKY) return pre > 0.f ? 1.f : slope;
    return 1.f;
}
__device__ __forceinline__ float act_fwd(float y, int act, float slope) {
    if (act == CCN_ACT_RELU) return fmaxf(y, 0.f);
    if (act == CCN_ACT_LEAKY) return y > 0.f ? y : y * slope;
    return y;
}

// dst[r] = idx[r] >= 0 ? src[idx[r]] : 0
__global__ void gather_rows_i32_kernel(const float* __restrict__ src, int64_t lds, const int32_t* __restrict__ idx,
                                       int64_t n, int c, float* __restrict__ dst, int64_t ldd) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)c;
    GRID_STRIDE(t, total) {
        int64_t r;
        int k;
        ccn_divmod(t, c, r, k);
        const int s = idx[r];
        dst[r * ldd + k] = s >= 0 ? src[(int64_t)s * lds + k] : 0.f;
    }
}

// dst[idx[r]] += src[r]   (rows with idx < 0 are skipped; dst is accumulated into)
template <typename IdxT>
__global__ void scatter_add_rows_kernel(const float* __restrict__ src, int64_t lds, const IdxT* __restrict__ idx,
                                        int64_t n, int c, float* __restrict__ dst, int64_t ldd) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)c;
    GRID_STRIDE(t, total) {
        int64_t r;
        int k;
        ccn_divmod(t, c, r, k);
        const int64_t d = (int64_t)idx[r];
        if (d >= 0) atomicAdd(dst + d * ldd + k, src[r * lds + k]);
    }
}

// float4 forms (c % 4 == 0, 16-byte aligned rows): one thread per (row, 4 channels); the scatter uses one 16-byte reduction
__global__ void gather_rows_i32_v4_kernel(const float* __restrict__ src, int64_t lds, const int32_t* __restrict__ idx,
                                          int64_t n, int c4, float* __restrict__ dst, int64_t ldd) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)c4;
    GRID_STRIDE(t, total) {
        int64_t r;
        int k;
        ccn_divmod(t, c4, r, k);
        const int s = idx[r];
        *reinterpret_cast<float4*>(dst + r * ldd + 4 * k) =
            s >= 0 ? *reinterpret_cast<const float4*>(src + (int64_t)s * lds + 4 * k) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

template <typename IdxT>
__global__ void scatter_add_rows_v4_kernel(const float* __restrict__ src, int64_t lds, const IdxT* __restrict__ idx,
                                           int64_t n, int c4, float* __restrict__ dst, int64_t ldd) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)c4;
    GRID_STRIDE(t, total) {
        int64_t r;
        int k;
        ccn_divmod(t, c4, r, k);
        const int64_t d = (int64_t)idx[r];
        if (d < 0) continue;
        const float4 v = *reinterpret_cast<const float4*>(src + r * lds + 4 * k);
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + d * ldd + 4 * k), "f"(v.x), "f"(v.y),
                     "f"(v.z), "f"(v.w) : "memory");
    }
}

static inline bool rows_v4_ok(const float* a, int64_t lda, const float* b, int64_t ldb, int32_t c) {
    return (c & 3) == 0 && (lda & 3) == 0 && (ldb & 3) == 0 && ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0;
}

// dst[i] (+)= sum of the G consecutive rows i*G .. i*G+G-1 of src
__global__ void group_sum_rows_kernel(const float* __restrict__ src, int64_t lds, int64_t n, int G, int c,
                                      float* __restrict__ dst, int64_t ldd, int accumulate) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)c;
    GRID_STRIDE(t, total) {
        int64_t i;
        int k;
        ccn_divmod(t, c, i, k);
        float s = 0.f;
        for (int g = 0; g < G; ++g) s += src[(i * G + g) * lds + k];
        float* o = dst + i * ldd + k;
        *o = accumulate ? *o + s : s;
    }
}

// ------------------------------------------------------------------------------------------------------------
// Linear weight gradient: dW[n][k] += sum_m dY[m][n] X[m][k], db[n] += sum_m dY[m][n].
// CTA = 64 x 64 tile of dW over one slice of the rows; 256 threads x (4 x 4) register tile; slices meet in
// fp32 atomics (the caller zero-fills dW / db).
// ------------------------------------------------------------------------------------------------------------
constexpr int WG_T = 64, WG_R = 16;

__global__ void __launch_bounds__(256) linear_wgrad_kernel(const float* __restrict__ dY, int64_t lddy,
                                                           const float* __restrict__ X, int64_t ldx, int64_t M, int N,
                                                           int K, int64_t rows_per_split, float* __restrict__ dW,
                                                           int64_t lddw, float* __restrict__ db) {
    ccn_pdl_wait();
    __shared__ float sy[WG_R][WG_T + 4];
    __shared__ float sx[WG_R][WG_T + 4];
    const int k0 = blockIdx.x * WG_T, n0 = blockIdx.y * WG_T;
    const int64_t m_beg = (int64_t)blockIdx.z * rows_per_split;
    const int64_t m_end = min(M, m_beg + rows_per_split);
    const int tn = (threadIdx.x >> 4) * 4, tk = (threadIdx.x & 15) * 4;
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
    float bsum = 0.f;                                   // threads 0..63 of the k-tile-0 CTAs: column sums of dY
    const bool do_bias = db != nullptr && blockIdx.x == 0 && threadIdx.x < WG_T;

    for (int64_t m0 = m_beg; m0 < m_end; m0 += WG_R) {
        __syncthreads();
        for (int t = threadIdx.x; t < WG_R * WG_T; t += 256) {
            const int r = t >> 6, cidx = t & 63;
            const int64_t m = m0 + r;
            const bool in = m < m_end;
            sy[r][cidx] = (in && n0 + cidx < N) ? dY[m * lddy + n0 + cidx] : 0.f;
            sx[r][cidx] = (in && k0 + cidx < K) ? X[m * ldx + k0 + cidx] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < WG_R; ++r) {
            const float4 y4 = *reinterpret_cast<const float4*>(&sy[r][tn]);
            const float4 x4 = *reinterpret_cast<const float4*>(&sx[r][tk]);
            const float yv[4] = {y4.x, y4.y, y4.z, y4.w}, xv[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(yv[a], xv[b], acc[a][b]);
        }
        if (do_bias) {
#pragma unroll
            for (int r = 0; r < WG_R; ++r) bsum += sy[r][threadIdx.x];
        }
    }
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int n = n0 + tn + a, k = k0 + tk + b;
            if (n < N && k < K) atomicAdd(dW + (int64_t)n * lddw + k, acc[a][b]);
        }
    if (do_bias && n0 + (int)threadIdx.x < N) atomicAdd(db + n0 + threadIdx.x, bsum);
}

// ------------------------------------------------------------------------------------------------------------
// BatchNorm1d, training mode (torch.nn.BatchNorm1d inside torch_geometric MLP and fast_conv1d.py): batch
// statistics over the M rows.  sums: fp64 [2*C] = (sum z, sum z^2), zero-filled by the entry point.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bn_stats_kernel(const float* __restrict__ z, int64_t ld, int64_t M, int C,
                                                       int64_t rows_per_cta, double* __restrict__ sums) {
    ccn_pdl_wait();
    // thread = one column (c = threadIdx.x + 256*blockIdx.y handled by grid.y), rows split over grid.x; for
    // narrow C several row lanes share a column: lanes = 256 / Cw
    const int cw = C >= 256 ? 256 : C;                   // columns handled by this CTA row of threads
    const int lanes = 256 / cw;
    const int col = blockIdx.y * 256 + (threadIdx.x % cw);
    const int lane = threadIdx.x / cw;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
    if (col >= C || lane >= lanes) return;
    double s = 0.0, q = 0.0;                              // fp64: E[z^2] - mean^2 must survive the cancellation
    for (int64_t r = r0 + lane; r < r1; r += lanes) {
        const double v = (double)z[r * ld + col];
        s += v;
        q = fma(v, v, q);
    }
    atomicAdd(sums + col, s);
    atomicAdd(sums + C + col, q);
}

// ------------------------------------------------------------------------------------------------------------
// Vectorised column reductions (C % 4 == 0, 16-byte aligned rows): a thread owns 4 adjacent columns (float4 loads)
// and a row lane; two / four rows in flight per thread; the row lanes of a CTA are combined in shared memory so
// that a CTA issues ONE atomic per column and statistic.  MODE 0: float column sums (bias gradient),
// MODE 1: fp64 (sum z, sum z^2) of BatchNorm's batch statistics, MODE 2: fp64 (sum g, sum g * xhat) of its backward.
// ------------------------------------------------------------------------------------------------------------
struct ColRed {
    const float* a; int64_t lda;          // MODE 0: src, 1: z, 2: dy
    const float* z; int64_t ldz;          // MODE 2: z
    const float *mean, *invstd, *gamma, *beta;
    int act; float slope;
    int64_t M; int C; int64_t rows_per_cta;
    double* sums; float* out;
};

template <int MODE>
__global__ void __launch_bounds__(256) col_reduce_v4_kernel(const ColRed p) {
    ccn_pdl_wait();
    constexpr int NS = MODE == 0 ? 4 : 8;              // statistics per thread
    constexpr int UNR = MODE == 2 ? 2 : 4;             // rows in flight per thread
    __shared__ double red[256 * NS];
    const int c4n = p.C >> 2;
    const int cw = c4n >= 256 ? 256 : c4n;
    const int lanes = 256 / cw;
    const int g = blockIdx.y * 256 + (threadIdx.x % cw);
    const int lane = threadIdx.x / cw;
    const bool live = g < c4n && lane < lanes;
    const int64_t r0 = (int64_t)blockIdx.x * p.rows_per_cta, r1 = min(p.M, r0 + p.rows_per_cta);
    double acc[NS];
#pragma unroll
    for (int k = 0; k < NS; ++k) acc[k] = 0.0;
    if (live) {
        const int c = g << 2;
        float4 mu = make_float4(0, 0, 0, 0), is = mu, ga = mu, be = mu;
        if (MODE == 2) {
            mu = *reinterpret_cast<const float4*>(p.mean + c); is = *reinterpret_cast<const float4*>(p.invstd + c);
            ga = *reinterpret_cast<const float4*>(p.gamma + c); be = *reinterpret_cast<const float4*>(p.beta + c);
        }
        float fs[4] = {0.f, 0.f, 0.f, 0.f};            // MODE 0 accumulates in float like the scalar kernel
        for (int64_t r = r0 + lane; r < r1; r += (int64_t)lanes * UNR) {
            float4 va[UNR], vz[UNR];
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                const int64_t rr = r + (int64_t)u * lanes;
                const bool ok = rr < r1;
                va[u] = ok ? *reinterpret_cast<const float4*>(p.a + rr * p.lda + c) : make_float4(0, 0, 0, 0);
                if (MODE == 2) vz[u] = ok ? *reinterpret_cast<const float4*>(p.z + rr * p.ldz + c) : mu;
            }
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                const float a4[4] = {va[u].x, va[u].y, va[u].z, va[u].w};
                if (MODE == 0) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) fs[e] += a4[e];
                } else if (MODE == 1) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const double v = (double)a4[e];
                        acc[e] += v;
                        acc[4 + e] = fma(v, v, acc[4 + e]);
                    }
                } else {
                    const float z4[4] = {vz[u].x, vz[u].y, vz[u].z, vz[u].w};
                    const float m4[4] = {mu.x, mu.y, mu.z, mu.w}, i4[4] = {is.x, is.y, is.z, is.w};
                    const float g4[4] = {ga.x, ga.y, ga.z, ga.w}, b4[4] = {be.x, be.y, be.z, be.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float xh = (z4[e] - m4[e]) * i4[e];
                        const float gg = a4[e] * act_grad(fmaf(xh, g4[e], b4[e]), p.act, p.slope);
                        acc[e] += (double)gg;
                        acc[4 + e] += (double)gg * (double)xh;
                    }
                }
            }
        }
        if (MODE == 0) {
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[e] = (double)fs[e];
        }
    }
#pragma unroll
    for (int k = 0; k < NS; ++k) red[k * 256 + threadIdx.x] = acc[k];
    __syncthreads();
    if (live && lane == 0) {
        const int c = g << 2;
#pragma unroll
        for (int k = 0; k < NS; ++k) {
            double t = acc[k];
            for (int l = 1; l < lanes; ++l) t += red[k * 256 + l * cw + (threadIdx.x % cw)];
            if (MODE == 0) atomicAdd(p.out + c + k, (float)t);
            else atomicAdd(p.sums + (k < 4 ? c + k : p.C + c + (k - 4)), t);
        }
    }
}

static inline bool col_reduce_v4_ok(const float* a, int64_t lda, int32_t C) {
    return (C & 3) == 0 && (lda & 3) == 0 && (reinterpret_cast<uintptr_t>(a) & 15) == 0;
}

template <int MODE>
static int launch_col_reduce_v4(ColRed p, cudaStream_t s) {
    const int c4n = p.C >> 2;
    const int gy = (c4n + 255) / 256;
    const int cw = c4n >= 256 ? 256 : c4n;
    const int lanes = 256 / cw;
    int64_t gx = (148 * 4) / gy;                                      // ~4 CTAs of 8 warps per SM
    const int64_t max_gx = (p.M + (int64_t)lanes * 8 - 1) / ((int64_t)lanes * 8);
    if (gx > max_gx) gx = max_gx;
    if (gx < 1) gx = 1;
    p.rows_per_cta = (p.M + gx - 1) / gx;
    gx = (p.M + p.rows_per_cta - 1) / p.rows_per_cta;
    ccn_launch(col_reduce_v4_kernel<MODE>, dim3(dim3((unsigned)gx, (unsigned)gy)), dim3(256), 0, s, p);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}


// mean / invstd from the sums; running statistics updated like torch (momentum, unbiased variance).
__global__ void bn_finalize_kernel(const double* __restrict__ sums, int64_t M, int C, float eps, float momentum,
                                   float* __restrict__ mean, float* __restrict__ invstd,
                                   float* __restrict__ running_mean, float* __restrict__ running_var) {
    ccn_pdl_wait();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const double mu = sums[c] / (double)M;
    double var = sums[C + c] / (double)M - mu * mu;
    if (var < 0.0) var = 0.0;
    mean[c] = (float)mu;
    invstd[c] = (float)(1.0 / sqrt(var + (double)eps));
    if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mu;
    if (running_var) {
        const double unb = M > 1 ? var * (double)M / (double)(M - 1) : var;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unb;
    }
}

// y = act(((z - mean) * invstd) * gamma + beta); gamma == nullptr: y = act(z)
__global__ void bn_apply_kernel(const float* __restrict__ z, int64_t ldz, int64_t M, int C,
                                const float* __restrict__ mean, const float* __restrict__ invstd,
                                const float* __restrict__ gamma, const float* __restrict__ beta, int act, float slope,
                                float* __restrict__ y, int64_t ldy, const PlanesOut po) {
    ccn_pdl_wait();
    const int W = po.p ? (int)po.kpad : C;               // with planes: the pad columns are written too (zeros)
    const int64_t total = M * (int64_t)W;
    GRID_STRIDE(t, total) {
        int64_t r;
        int c;
        ccn_divmod(t, W, r, c);
        float o = 0.f;
        if (c < C) {
            float v = z[r * ldz + c];
            if (gamma) v = fmaf((v - mean[c]) * invstd[c], gamma[c], beta[c]);
            o = act_fwd(v, act, slope);
            y[r * ldy + c] = o;
        }
        if (po.p) st_plane(po, r, c, o);
    }
}

// sums (fp64 [2*C]) += (sum g, sum g * xhat) with g = dy * act'(pre)
__global__ void __launch_bounds__(256) bn_bwd_reduce_kernel(const float* __restrict__ dy, int64_t lddy,
                                                            const float* __restrict__ z, int64_t ldz, int64_t M, int C,
                                                            const float* __restrict__ mean,
                                                            const float* __restrict__ invstd,
                                                            const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, int act, float slope,
                                                            int64_t rows_per_cta, double* __restrict__ sums) {
    ccn_pdl_wait();
    const int cw = C >= 256 ? 256 : C;
    const int lanes = 256 / cw;
    const int col = blockIdx.y * 256 + (threadIdx.x % cw);
    const int lane = threadIdx.x / cw;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
    if (col >= C || lane >= lanes) return;
    const float mu = mean[col], is = invstd[col], ga = gamma[col], be = beta[col];
    double s = 0.0, q = 0.0;
    for (int64_t r = r0 + lane; r < r1; r += lanes) {
        const float xh = (z[r * ldz + col] - mu) * is;
        const float g = dy[r * lddy + col] * act_grad(fmaf(xh, ga, be), act, slope);
        s += (double)g;
        q += (double)g * (double)xh;
    }
    atomicAdd(sums + col, s);
    atomicAdd(sums + C + col, q);
}

// batch_stats: dz = gamma * invstd * (g - sum_g / M - xhat * sum_gx / M); running stats: dz = gamma * invstd * g;
// gamma == nullptr (no norm): dz = dy * act'(z).  CTA 0 also writes dgamma = sum_gx, dbeta = sum_g.
__global__ void bn_bwd_apply_kernel(const float* __restrict__ dy, int64_t lddy, const float* __restrict__ z,
                                    int64_t ldz, int64_t M, int C, const float* __restrict__ mean,
                                    const float* __restrict__ invstd, const float* __restrict__ gamma,
                                    const float* __restrict__ beta, int act, float slope, int batch_stats,
                                    const double* __restrict__ sums, float* __restrict__ dz, int64_t lddz,
                                    float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dzsum,
                                    const PlanesOut po) {
    ccn_pdl_wait();
    if (gamma && blockIdx.x == 0) {
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            if (dbeta) dbeta[c] = (float)sums[c];
            if (dgamma) dgamma[c] = (float)sums[C + c];
            // column sums of dz (the bias gradient of the Linear in front): identically zero under batch statistics
            // (sum xhat = 0), gamma * invstd * sum g on running statistics
            if (dzsum) dzsum[c] = batch_stats ? 0.f : gamma[c] * invstd[c] * (float)sums[c];
        }
    }
    const float inv_m = 1.f / (float)M;
    const int W = po.p ? (int)po.kpad : C;
    const int64_t total = M * (int64_t)W;
    GRID_STRIDE(t, total) {
        int64_t r;
        int c;
        ccn_divmod(t, W, r, c);
        float out = 0.f;
        if (c < C) {
            const float zv = z[r * ldz + c];
            if (!gamma) {
                out = dy[r * lddy + c] * act_grad(zv, act, slope);
            } else {
                const float is = invstd[c], ga = gamma[c];
                const float xh = (zv - mean[c]) * is;
                const float g = dy[r * lddy + c] * act_grad(fmaf(xh, ga, beta[c]), act, slope);
                out = batch_stats ? ga * is * (g - (float)sums[c] * inv_m - xh * ((float)sums[C + c] * inv_m)) : ga * is * g;
            }
            dz[r * lddz + c] = out;
        }
        if (po.p) st_plane(po, r, c, out);
    }
}


// ---- float4 variants of the two BatchNorm element-wise kernels (C % 4 == 0, 16-byte aligned rows): no index
// division, 16-byte loads / stores, 8-byte plane stores.  Thread = (row lane, group of 4 columns).
__device__ __forceinline__ void st_plane4(const PlanesOut& po, int64_t r, int c, const float (&v)[4]) {
    __nv_bfloat162 h0 = __floats2bfloat162_rn(v[0], v[1]), h1 = __floats2bfloat162_rn(v[2], v[3]);
    uint2 u;
    u.x = *reinterpret_cast<uint32_t*>(&h0);
    u.y = *reinterpret_cast<uint32_t*>(&h1);
    *reinterpret_cast<uint2*>(po.p + r * po.kpad + c) = u;
    if (po.n == 2) {
        const float2 f0 = __bfloat1622float2(h0), f1 = __bfloat1622float2(h1);
        __nv_bfloat162 l0 = __floats2bfloat162_rn(v[0] - f0.x, v[1] - f0.y), l1 = __floats2bfloat162_rn(v[2] - f1.x, v[3] - f1.y);
        u.x = *reinterpret_cast<uint32_t*>(&l0);
        u.y = *reinterpret_cast<uint32_t*>(&l1);
        *reinterpret_cast<uint2*>(po.p + po.stride + r * po.kpad + c) = u;
    }
}

__global__ void __launch_bounds__(256) bn_apply_v4_kernel(const float* __restrict__ z, int64_t ldz, int64_t M, int C,
                                                          const float* __restrict__ mean, const float* __restrict__ invstd,
                                                          const float* __restrict__ gamma, const float* __restrict__ beta,
                                                          int act, float slope, float* __restrict__ y, int64_t ldy,
                                                          const PlanesOut po) {
    ccn_pdl_wait();
    const int G = (po.p ? (int)po.kpad : C) >> 2;
    const int tg = G < 256 ? G : 256, rows_pb = 256 / tg;
    const int rl = threadIdx.x / tg, cg0 = threadIdx.x - rl * tg;
    if (rl >= rows_pb) return;
    for (int64_t r = (int64_t)blockIdx.x * rows_pb + rl; r < M; r += (int64_t)gridDim.x * rows_pb) {
        for (int cg = cg0; cg < G; cg += tg) {
            const int c = cg << 2;
            float o[4] = {0.f, 0.f, 0.f, 0.f};
            if (c < C) {
                const float4 zv = *reinterpret_cast<const float4*>(z + r * ldz + c);
                float v[4] = {zv.x, zv.y, zv.z, zv.w};
                if (gamma) {
                    const float4 mu = *reinterpret_cast<const float4*>(mean + c), is = *reinterpret_cast<const float4*>(invstd + c);
                    const float4 ga = *reinterpret_cast<const float4*>(gamma + c), be = *reinterpret_cast<const float4*>(beta + c);
                    v[0] = fmaf((v[0] - mu.x) * is.x, ga.x, be.x); v[1] = fmaf((v[1] - mu.y) * is.y, ga.y, be.y);
                    v[2] = fmaf((v[2] - mu.z) * is.z, ga.z, be.z); v[3] = fmaf((v[3] - mu.w) * is.w, ga.w, be.w);
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) o[j] = act_fwd(v[j], act, slope);
                *reinterpret_cast<float4*>(y + r * ldy + c) = make_float4(o[0], o[1], o[2], o[3]);
            }
            if (po.p) st_plane4(po, r, c, o);
        }
    }
}

__global__ void __launch_bounds__(256) bn_bwd_apply_v4_kernel(
    const float* __restrict__ dy, int64_t lddy, const float* __restrict__ z, int64_t ldz, int64_t M, int C,
    const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ gamma,
    const float* __restrict__ beta, int act, float slope, int batch_stats, const double* __restrict__ sums,
    float* __restrict__ dz, int64_t lddz, float* __restrict__ dgamma, float* __restrict__ dbeta,
    float* __restrict__ dzsum, const PlanesOut po) {
    ccn_pdl_wait();
    if (gamma && blockIdx.x == 0) {
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            if (dbeta) dbeta[c] = (float)sums[c];
            if (dgamma) dgamma[c] = (float)sums[C + c];
            if (dzsum) dzsum[c] = batch_stats ? 0.f : gamma[c] * invstd[c] * (float)sums[c];
        }
    }
    const float inv_m = 1.f / (float)M;
    const int G = (po.p ? (int)po.kpad : C) >> 2;
    const int tg = G < 256 ? G : 256, rows_pb = 256 / tg;
    const int rl = threadIdx.x / tg, cg0 = threadIdx.x - rl * tg;
    if (rl >= rows_pb) return;
    for (int cg = cg0; cg < G; cg += tg) {
        const int c = cg << 2;
        float mu[4] = {0, 0, 0, 0}, is[4] = {1, 1, 1, 1}, ga[4] = {1, 1, 1, 1}, be[4] = {0, 0, 0, 0}, s0[4] = {0, 0, 0, 0},
              s1[4] = {0, 0, 0, 0};
        if (gamma && c < C) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                mu[j] = mean[c + j]; is[j] = invstd[c + j]; ga[j] = gamma[c + j]; be[j] = beta[c + j];
                s0[j] = (float)sums[c + j] * inv_m; s1[j] = (float)sums[C + c + j] * inv_m;
            }
        }
        for (int64_t r = (int64_t)blockIdx.x * rows_pb + rl; r < M; r += (int64_t)gridDim.x * rows_pb) {
            float o[4] = {0.f, 0.f, 0.f, 0.f};
            if (c < C) {
                const float4 zv4 = *reinterpret_cast<const float4*>(z + r * ldz + c);
                const float4 dy4 = *reinterpret_cast<const float4*>(dy + r * lddy + c);
                const float zv[4] = {zv4.x, zv4.y, zv4.z, zv4.w}, dv[4] = {dy4.x, dy4.y, dy4.z, dy4.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (!gamma) {
                        o[j] = dv[j] * act_grad(zv[j], act, slope);
                    } else {
                        const float xh = (zv[j] - mu[j]) * is[j];
                        const float g = dv[j] * act_grad(fmaf(xh, ga[j], be[j]), act, slope);
                        o[j] = batch_stats ? ga[j] * is[j] * (g - s0[j] - xh * s1[j]) : ga[j] * is[j] * g;
                    }
                }
                *reinterpret_cast<float4*>(dz + r * lddz + c) = make_float4(o[0], o[1], o[2], o[3]);
            }
            if (po.p) st_plane4(po, r, c, o);
        }
    }
}

static inline unsigned bn_v4_grid(int64_t M, int W) {
    const int G = W >> 2, tg = G < 256 ? G : 256, rows_pb = 256 / tg;
    int64_t b = (M + rows_pb - 1) / rows_pb;
    const int64_t cap = 148LL * 16;
    return (unsigned)(b < 1 ? 1 : (b > cap ? cap : b));
}
static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// ------------------------------------------------------------------------------------------------------------
// Backward of ccn_seg_aggregate (agg_ops.cu seg_aggregate_kernel; point_conv.py:71-93, dgcnn.py:240-244).
// One thread per (segment, channel); the forward is recomputed from msg / att.
//   max       : the gradient goes to the first edge (lowest index) attaining the maximum
//   mean/sum  : dmsg = dout / max(count, 1)   |   dmsg = dout
//   weighted  : s = sigmoid(a); dmsg = s dout; datt = s (1 - s) msg dout   (normalised form: see the branch)
//   attend    : s = softmax_seg(a); dmsg = s dout; datt = s (msg - sum_e s_e msg_e) dout
// ------------------------------------------------------------------------------------------------------------
__global__ void seg_aggregate_bwd_kernel(const float* __restrict__ msg, int64_t ldm, const float* __restrict__ att,
                                         int64_t lda, const int32_t* __restrict__ row_ptr, int64_t nq, int c, int mode,
                                         const float* __restrict__ dout, int64_t ldo, float* __restrict__ dmsg,
                                         int64_t lddm, float* __restrict__ datt, int64_t ldda) {
    ccn_pdl_wait();
    const int64_t total = nq * (int64_t)c;
    GRID_STRIDE(t, total) {
        int64_t q;
        int k;
        ccn_divmod(t, c, q, k);
        const int b = row_ptr[q], e = row_ptr[q + 1];
        if (e <= b) continue;
        const float g = dout[q * ldo + k];
        if (mode == CCN_AGG_MAX) {
            float r = msg[(int64_t)b * ldm + k];
            int arg = b;
            for (int j = b + 1; j < e; ++j) {
                const float v = msg[(int64_t)j * ldm + k];
                if (v > r) { r = v; arg = j; }
            }
            for (int j = b; j < e; ++j) dmsg[(int64_t)j * lddm + k] = j == arg ? g : 0.f;
        } else if (mode == CCN_AGG_MEAN || mode == CCN_AGG_SUM) {
            const float v = mode == CCN_AGG_MEAN ? g / (float)(e - b) : g;
            for (int j = b; j < e; ++j) dmsg[(int64_t)j * lddm + k] = v;
        } else if (mode == CCN_AGG_WSUM) {
            for (int j = b; j < e; ++j) {
                const float s = 1.f / (1.f + expf(-att[(int64_t)j * lda + k]));
                dmsg[(int64_t)j * lddm + k] = s * g;
                datt[(int64_t)j * ldda + k] = s * (1.f - s) * msg[(int64_t)j * ldm + k] * g;
            }
        } else if (mode == CCN_AGG_WSUM_NORM) {
            // out = sum_j m_j w_j / D, w = sigmoid(a), D = clamp(sum w, 1e-3) (dgcnn.py:190-197): the clamp passes the
            // gradient of the sum only where it is not active
            float tot = 0.f, o = 0.f;
            for (int j = b; j < e; ++j) tot += 1.f / (1.f + expf(-att[(int64_t)j * lda + k]));
            const bool free_sum = tot > 1e-3f;
            const float den = fmaxf(tot, 1e-3f);
            for (int j = b; j < e; ++j)
                o += msg[(int64_t)j * ldm + k] * ((1.f / (1.f + expf(-att[(int64_t)j * lda + k]))) / den);
            for (int j = b; j < e; ++j) {
                const float s = 1.f / (1.f + expf(-att[(int64_t)j * lda + k]));
                const float m = msg[(int64_t)j * ldm + k];
                dmsg[(int64_t)j * lddm + k] = (s / den) * g;
                datt[(int64_t)j * ldda + k] = s * (1.f - s) * ((free_sum ? m - o : m) / den) * g;
            }
        } else {
            float m = -INFINITY;
            for (int j = b; j < e; ++j) m = fmaxf(m, att[(int64_t)j * lda + k]);
            float s = 0.f;
            for (int j = b; j < e; ++j) s += expf(att[(int64_t)j * lda + k] - m);
            const float inv = 1.f / (s + 1e-16f);
            float o = 0.f;
            for (int j = b; j < e; ++j) o += msg[(int64_t)j * ldm + k] * (expf(att[(int64_t)j * lda + k] - m) * inv);
            for (int j = b; j < e; ++j) {
                const float p = expf(att[(int64_t)j * lda + k] - m) * inv;
                dmsg[(int64_t)j * lddm + k] = p * g;
                datt[(int64_t)j * ldda + k] = p * (msg[(int64_t)j * ldm + k] - o) * g;
            }
        }
    }
}

// Backward of ccn_group_max_dense: slot 0 = self loop, slot 1+k valid iff idx[i][k] >= 0 (else the constant -100).
__global__ void group_max_dense_bwd_kernel(const float* __restrict__ feat, int64_t ldf, const int32_t* __restrict__ idx,
                                           int64_t n, int K, int c, const float* __restrict__ dout, int64_t ldo,
                                           float* __restrict__ dfeat, int64_t lddf) {
    ccn_pdl_wait();
    const int G = K + 1;
    const int64_t total = n * (int64_t)c;
    GRID_STRIDE(t, total) {
        int64_t i;
        int k;
        ccn_divmod(t, c, i, k);
        float r = feat[(i * G) * ldf + k];
        int arg = 0;
        for (int g = 1; g < G; ++g) {
            const bool valid = idx[i * K + g - 1] >= 0;
            const float v = valid ? feat[(i * G + g) * ldf + k] : -100.f;
            if (v > r) { r = v; arg = valid ? g : -1; }
        }
        const float go = dout[i * ldo + k];
        for (int g = 0; g < G; ++g) dfeat[(i * G + g) * lddf + k] = g == arg ? go : 0.f;
    }
}

// Backward of ccn_interp_blend w.r.t. x: dx[nbr] += (w / sum w) dout  (dx zero-filled by the caller)
__global__ void interp_blend_bwd_kernel(const int32_t* __restrict__ nbr, const float* __restrict__ wgt, int from_d2,
                                        int64_t n, int knn, int c, const float* __restrict__ dout, int64_t ldo,
                                        float* __restrict__ dx, int64_t lddx) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)c;
    GRID_STRIDE(t, total) {
        int64_t i;
        int k;
        ccn_divmod(t, c, i, k);
        float den = 0.f;
        for (int s = 0; s < knn; ++s) {
            if (nbr[i * knn + s] < 0) continue;
            float w = wgt[i * knn + s];
            if (from_d2) w = __fdiv_rn(1.f, fmaxf(w, 1e-16f));
            den += w;
        }
        const float g = dout[i * ldo + k] / den;
        for (int s = 0; s < knn; ++s) {
            const int j = nbr[i * knn + s];
            if (j < 0) continue;
            float w = wgt[i * knn + s];
            if (from_d2) w = __fdiv_rn(1.f, fmaxf(w, 1e-16f));
            atomicAdd(dx + (int64_t)j * lddx + k, w * g);
        }
    }
}

// Backward of ccn_curve_feat_prep w.r.t. x (the position columns carry no gradient):
// out = [f, |d|], d_i = (ef_i + eb_i) / den_i; gather form, no atomics.
__global__ void curve_feat_prep_bwd_kernel(const float* __restrict__ x, int64_t ldx, int cx, int c_in, int with_diff,
                                           const int32_t* __restrict__ ord, int64_t n, const float* __restrict__ dout,
                                           int64_t ldo, float* __restrict__ dx, int64_t lddx) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)cx;
    GRID_STRIDE(t, total) {
        int64_t i;
        int c;
        ccn_divmod(t, cx, i, c);
        float g = dout[i * ldo + c];
        if (with_diff) {
            auto same = [&](int64_t a, int64_t b) { return a >= 0 && b < n && ord[a] == ord[b]; };
            // sign(d_r) * dout_diff_r / den_r of row r
            auto sg = [&](int64_t r) -> float {
                const bool vb = same(r - 1, r), vf = same(r, r + 1);
                const float v = x[r * ldx + c];
                const float eb = vb ? __fsub_rn(v, x[(r - 1) * ldx + c]) : 0.f;
                const float ef = vf ? __fsub_rn(x[(r + 1) * ldx + c], v) : 0.f;
                const float d = __fadd_rn(ef, eb);
                const float den = fmaxf((vf ? 1.f : 0.f) + (vb ? 1.f : 0.f), 1.f);
                const float s = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
                return s * dout[r * ldo + c_in + c] / den;
            };
            const bool vb = same(i - 1, i), vf = same(i, i + 1);
            if (vb) g += sg(i - 1);                               // row i-1's forward difference reads +x_i
            if (vf) g -= sg(i + 1);                               // row i+1's backward difference reads -x_i
            if (vb != vf) g += sg(i) * (vb ? 1.f : -1.f);         // own row: (+1 if vb) + (-1 if vf)
        }
        dx[i * lddx + c] = g;
    }
}

// rel[e] = (pos_src[ecol[e]] - pos_dst[erow[e]]) (/ radius): the positional input columns of PointNetConv2's message
__global__ void edge_rel_pos_kernel(const float* __restrict__ pos_src, const float* __restrict__ pos_dst,
                                    const int32_t* __restrict__ erow, const int32_t* __restrict__ ecol, int64_t E,
                                    float radius, float* __restrict__ rel) {
    ccn_pdl_wait();
    const int64_t total = E * 3;
    GRID_STRIDE(t, total) {
        int64_t e;
        int a;
        ccn_divmod(t, 3, e, a);
        float r = __fsub_rn(pos_src[3 * (int64_t)ecol[e] + a], pos_dst[3 * (int64_t)erow[e] + a]);
        if (radius != 0.f) r = __fdiv_rn(r, radius);
        rel[t] = r;
    }
}

// ------------------------------------------------------------------------------------------------------------
// log_softmax + NLL (F.nll_loss(F.log_softmax(out, 1), y, weight, ignore_index), reduction 'mean':
// src/run/shapenet_seg.py:182-186, nuscenes_seg.py:229-249).  acc: fp64 [2] = (sum w nll, sum w), zero-filled.
// ------------------------------------------------------------------------------------------------------------
__global__ void nll_fwd_kernel(const float* __restrict__ logits, int64_t ld, const int64_t* __restrict__ target,
                               const float* __restrict__ cw, int64_t n, int C, int64_t ignore, float* __restrict__ lse,
                               double* __restrict__ acc, float* __restrict__ perpoint) {
    ccn_pdl_wait();
    double ls = 0.0, ws = 0.0;
    GRID_STRIDE(i, n) {
        const float* row = logits + i * ld;
        float m = -INFINITY;
        for (int c = 0; c < C; ++c) m = fmaxf(m, row[c]);
        float s = 0.f;
        for (int c = 0; c < C; ++c) s += expf(row[c] - m);
        const float l = m + logf(s);
        lse[i] = l;
        const int64_t y = target[i];
        if (perpoint) perpoint[i] = (y >= 0 && y < C) ? l - row[y] : 0.f;      // F.nll_loss(reduction="none"), unweighted
        if (y != ignore && y >= 0 && y < C) {
            const float w = cw ? cw[y] : 1.f;
            ls += (double)(w * (l - row[y]));
            ws += (double)w;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        ls += __shfl_down_sync(0xffffffffu, ls, o);
        ws += __shfl_down_sync(0xffffffffu, ws, o);
    }
    if ((threadIdx.x & 31) == 0 && (ls != 0.0 || ws != 0.0)) {
        atomicAdd(acc, ls);
        atomicAdd(acc + 1, ws);
    }
}

__global__ void nll_finish_kernel(const double* __restrict__ acc, float* __restrict__ loss) {
    ccn_pdl_wait();
    loss[0] = (float)(acc[0] / acc[1]);
}

// dlogits[i][c] = gscale * w_y / sum_w * (softmax[i][c] - [c == y])
__global__ void nll_bwd_kernel(const float* __restrict__ logits, int64_t ld, const int64_t* __restrict__ target,
                               const float* __restrict__ cw, int64_t n, int C, int64_t ignore,
                               const float* __restrict__ lse, const double* __restrict__ acc,
                               const float* __restrict__ gscale, float* __restrict__ dlogits, int64_t ldd) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)C;
    const float gs = gscale[0] / (float)acc[1];
    GRID_STRIDE(t, total) {
        int64_t i;
        int c;
        ccn_divmod(t, C, i, c);
        const int64_t y = target[i];
        float g = 0.f;
        if (y != ignore && y >= 0 && y < C) {
            const float w = cw ? cw[y] : 1.f;
            g = gs * w * (expf(logits[i * ld + c] - lse[i]) - (c == y ? 1.f : 0.f));
        }
        dlogits[i * ldd + c] = g;
    }
}

// y = u >= p ? x / (1 - p) : 0   (F.dropout with the uniform draws u supplied by the caller; its own backward)
__global__ void dropout_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ u, int64_t n, int c,
                               float p, float* __restrict__ y, int64_t ldy) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)c;
    const float sc = 1.f / (1.f - p);
    GRID_STRIDE(t, total) {
        int64_t r;
        int k;
        ccn_divmod(t, c, r, k);
        y[r * ldy + k] = u[t] >= p ? x[r * ldx + k] * sc : 0.f;
    }
}

}  // namespace

// bias gradient: out[c] += sum_r src[r][c] (called by ccn_col_sum, gemm_wgrad.cu)
int ccn_col_sum_v4(const float* src, int64_t ld, int64_t M, int32_t C, float* out, cudaStream_t s) {
    ColRed p = {};
    p.a = src; p.lda = ld; p.M = M; p.C = C; p.out = out;
    return launch_col_reduce_v4<0>(p, s);
}

// ------------------------------------------------------------------------------------------------------------
// C-ABI (include/ccn_b200.h, "Training step")
// ------------------------------------------------------------------------------------------------------------
#define S_ ((cudaStream_t)stream)

extern "C" int ccn_gather_rows_i32(const float* src, int64_t lds, const int32_t* idx, int64_t n, int32_t c, float* dst,
                                   int64_t ldd, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && c > 0 && src && idx && dst && lds >= c && ldd >= c);
    if (n == 0) return CCN_OK;
    if (rows_v4_ok(src, lds, dst, ldd, c)) {
        ccn_launch(gather_rows_i32_v4_kernel, dim3(ELT_GRID(n * (c >> 2))), dim3(256), 0, S_, src, lds, idx, n, c >> 2, dst, ldd);
        CCN_CHECK_LAUNCH();
        return CCN_OK;
    }
    ccn_launch(gather_rows_i32_kernel, dim3(ELT_GRID(n * c)), dim3(256), 0, S_, src, lds, idx, n, c, dst, ldd);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_scatter_add_rows_i32(const float* src, int64_t lds, const int32_t* idx, int64_t n, int32_t c,
                                        float* dst, int64_t ldd, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && c > 0 && src && idx && dst && lds >= c && ldd >= c);
    if (n == 0) return CCN_OK;
    if (rows_v4_ok(src, lds, dst, ldd, c)) {
        ccn_launch(scatter_add_rows_v4_kernel<int32_t>, dim3(ELT_GRID(n * (c >> 2))), dim3(256), 0, S_, src, lds, idx, n, c >> 2, dst, ldd);
        CCN_CHECK_LAUNCH();
        return CCN_OK;
    }
    ccn_launch(scatter_add_rows_kernel<int32_t>, dim3(ELT_GRID(n * c)), dim3(256), 0, S_, src, lds, idx, n, c, dst, ldd);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_scatter_add_rows_i64(const float* src, int64_t lds, const int64_t* idx, int64_t n, int32_t c,
                                        float* dst, int64_t ldd, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && c > 0 && src && idx && dst && lds >= c && ldd >= c);
    if (n == 0) return CCN_OK;
    if (rows_v4_ok(src, lds, dst, ldd, c)) {
        ccn_launch(scatter_add_rows_v4_kernel<int64_t>, dim3(ELT_GRID(n * (c >> 2))), dim3(256), 0, S_, src, lds, idx, n, c >> 2, dst, ldd);
        CCN_CHECK_LAUNCH();
        return CCN_OK;
    }
    ccn_launch(scatter_add_rows_kernel<int64_t>, dim3(ELT_GRID(n * c)), dim3(256), 0, S_, src, lds, idx, n, c, dst, ldd);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_group_sum_rows(const float* src, int64_t lds, int64_t n, int32_t G, int32_t c, float* dst,
                                  int64_t ldd, int32_t accumulate, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && G >= 1 && c > 0 && src && dst && lds >= c && ldd >= c);
    if (n == 0) return CCN_OK;
    ccn_launch(group_sum_rows_kernel, dim3(ELT_GRID(n * c)), dim3(256), 0, S_, src, lds, n, G, c, dst, ldd, accumulate);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_linear_wgrad(const float* dY, int64_t lddy, const float* X, int64_t ldx, int64_t M, int32_t N,
                                int32_t K, float* dW, int64_t lddw, float* db, ccn_stream_t stream) {
    CCN_REQUIRE(M >= 0 && N > 0 && K > 0 && dY && X && dW && lddy >= N && ldx >= K && lddw >= K);
    if (M == 0) return CCN_OK;
    const int tk = (K + WG_T - 1) / WG_T, tn = (N + WG_T - 1) / WG_T;
    // enough row slices for ~4 CTAs per SM, at least 256 rows each
    int64_t splits = (4 * 148 + (int64_t)tk * tn - 1) / ((int64_t)tk * tn);
    const int64_t max_splits = (M + 255) / 256;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    if (splits > 65535) splits = 65535;
    int64_t rps = (M + splits - 1) / splits;
    rps = (rps + WG_R - 1) / WG_R * WG_R;
    splits = (M + rps - 1) / rps;
    dim3 grid(tk, tn, (unsigned)splits);
    ccn_launch(linear_wgrad_kernel, dim3(grid), dim3(256), 0, S_, dY, lddy, X, ldx, M, N, K, rps, dW, lddw, db);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

static inline void bn_grid(int64_t M, int32_t C, dim3* grid, int64_t* rows_per_cta) {
    const int gy = (C + 255) / 256;
    int64_t gx = (148 * 8) / gy;
    const int64_t max_gx = (M + 63) / 64;
    if (gx > max_gx) gx = max_gx;
    if (gx < 1) gx = 1;
    *rows_per_cta = (M + gx - 1) / gx;
    gx = (M + *rows_per_cta - 1) / *rows_per_cta;
    *grid = dim3((unsigned)gx, (unsigned)gy);
}

extern "C" int ccn_bn_stats(const float* z, int64_t ldz, int64_t M, int32_t C, float eps, float momentum, double* sums,
                            float* mean, float* invstd, float* running_mean, float* running_var, ccn_stream_t stream) {
    CCN_REQUIRE(M > 0 && C > 0 && z && sums && mean && invstd && ldz >= C);
    cudaError_t e = cudaMemsetAsync(sums, 0, 2 * (size_t)C * sizeof(double), S_);
    if (e != cudaSuccess) return (int)e;
    // measured inside the training step (tools/profile_train.py): for the fp64 statistics the float4 kernel wins on
    // narrow layers (many row lanes per CTA), the one-column-per-thread kernel on wide ones; the bias-gradient and the
    // BatchNorm-backward reductions are faster vectorised at every width
    if (C <= 128 && col_reduce_v4_ok(z, ldz, C)) {
        ColRed p = {};
        p.a = z; p.lda = ldz; p.M = M; p.C = C; p.sums = sums;
        const int rc = launch_col_reduce_v4<1>(p, S_);
        if (rc != CCN_OK) return rc;
    } else {
        dim3 grid;
        int64_t rpc;
        bn_grid(M, C, &grid, &rpc);
        ccn_launch(bn_stats_kernel, dim3(grid), dim3(256), 0, S_, z, ldz, M, C, rpc, sums);
        CCN_CHECK_LAUNCH();
    }
    ccn_launch(bn_finalize_kernel, dim3((C + 255) / 256), dim3(256), 0, S_, sums, M, C, eps, momentum, mean, invstd, running_mean,
                                                        running_var);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

static inline bool planes_ok(const void* planes, int64_t kpad, int64_t pstride, int32_t npl, int32_t C) {
    return !planes || (kpad >= C && (npl == 1 || (npl == 2 && pstride > 0)));
}

extern "C" int ccn_bn_apply(const float* z, int64_t ldz, int64_t M, int32_t C, const float* mean, const float* invstd,
                            const float* gamma, const float* beta, int32_t act, float slope, float* y, int64_t ldy,
                            void* y_planes, int64_t yp_ld, int64_t yp_plane_stride, int32_t n_planes,
                            ccn_stream_t stream) {
    CCN_REQUIRE(M >= 0 && C > 0 && z && y && ldz >= C && ldy >= C && (!gamma || (mean && invstd && beta)));
    CCN_REQUIRE(planes_ok(y_planes, yp_ld, yp_plane_stride, n_planes, C));
    if (M == 0) return CCN_OK;
    const PlanesOut po = {(__nv_bfloat16*)y_planes, yp_ld, yp_plane_stride, n_planes};
    if ((C & 3) == 0 && (ldz & 3) == 0 && (ldy & 3) == 0 && al16(z) && al16(y) && (!gamma || (al16(mean) && al16(invstd) && al16(gamma) && al16(beta))) &&
        (!y_planes || ((yp_ld & 3) == 0 && (yp_plane_stride & 3) == 0 && (reinterpret_cast<uintptr_t>(y_planes) & 7) == 0))) {
        ccn_launch(bn_apply_v4_kernel, dim3(bn_v4_grid(M, (int)(y_planes ? yp_ld : C))), dim3(256), 0, S_, z, ldz, M, C, mean, invstd, gamma, beta,
                                                                                     act, slope, y, ldy, po);
        CCN_CHECK_LAUNCH();
        return CCN_OK;
    }
    ccn_launch(bn_apply_kernel, dim3(ELT_GRID(M * (y_planes ? yp_ld : C))), dim3(256), 0, S_, z, ldz, M, C, mean, invstd, gamma, beta, act,
                                                                          slope, y, ldy, po);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_bn_backward(const float* dy, int64_t lddy, const float* z, int64_t ldz, int64_t M, int32_t C,
                               const float* mean, const float* invstd, const float* gamma, const float* beta,
                               int32_t act, float slope, int32_t batch_stats, double* sums, float* dz, int64_t lddz,
                               float* dgamma, float* dbeta, float* dzsum, void* dz_planes, int64_t dzp_ld,
                               int64_t dzp_plane_stride, int32_t n_planes, ccn_stream_t stream) {
    CCN_REQUIRE(M >= 0 && C > 0 && dy && z && dz && lddy >= C && ldz >= C && lddz >= C);
    CCN_REQUIRE(!gamma || (mean && invstd && beta && sums));
    CCN_REQUIRE(planes_ok(dz_planes, dzp_ld, dzp_plane_stride, n_planes, C) && (!dzsum || gamma));
    if (M == 0) return CCN_OK;
    if (gamma) {
        cudaError_t e = cudaMemsetAsync(sums, 0, 2 * (size_t)C * sizeof(double), S_);
        if (e != cudaSuccess) return (int)e;
        if (col_reduce_v4_ok(dy, lddy, C) && col_reduce_v4_ok(z, ldz, C) && al16(mean) && al16(invstd) && al16(gamma) &&
            al16(beta)) {
            ColRed p = {};
            p.a = dy; p.lda = lddy; p.z = z; p.ldz = ldz; p.mean = mean; p.invstd = invstd; p.gamma = gamma; p.beta = beta;
            p.act = act; p.slope = slope; p.M = M; p.C = C; p.sums = sums;
            const int rc = launch_col_reduce_v4<2>(p, S_);
            if (rc != CCN_OK) return rc;
        } else {
            dim3 grid;
            int64_t rpc;
            bn_grid(M, C, &grid, &rpc);
            ccn_launch(bn_bwd_reduce_kernel, dim3(grid), dim3(256), 0, S_, dy, lddy, z, ldz, M, C, mean, invstd, gamma, beta, act, slope, rpc,
                                                       sums);
            CCN_CHECK_LAUNCH();
        }
    }
    const PlanesOut po = {(__nv_bfloat16*)dz_planes, dzp_ld, dzp_plane_stride, n_planes};
    if ((C & 3) == 0 && (ldz & 3) == 0 && (lddy & 3) == 0 && (lddz & 3) == 0 && al16(z) && al16(dy) && al16(dz) &&
        (!dz_planes || ((dzp_ld & 3) == 0 && (dzp_plane_stride & 3) == 0 && (reinterpret_cast<uintptr_t>(dz_planes) & 7) == 0))) {
        ccn_launch(bn_bwd_apply_v4_kernel, dim3(bn_v4_grid(M, (int)(dz_planes ? dzp_ld : C))), dim3(256), 0, S_, 
            dy, lddy, z, ldz, M, C, mean, invstd, gamma, beta, act, slope, batch_stats, sums, dz, lddz, dgamma, dbeta, dzsum, po);
        CCN_CHECK_LAUNCH();
        return CCN_OK;
    }
    ccn_launch(bn_bwd_apply_kernel, dim3(ELT_GRID(M * (dz_planes ? dzp_ld : C))), dim3(256), 0, S_, 
        dy, lddy, z, ldz, M, C, mean, invstd, gamma, beta, act, slope, batch_stats, sums, dz, lddz, dgamma, dbeta, dzsum, po);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_seg_aggregate_bwd(const float* msg, int64_t ldm, const float* att, int64_t lda,
                                     const int32_t* row_ptr, int64_t nq, int32_t c, int32_t mode, const float* dout,
                                     int64_t ldo, float* dmsg, int64_t lddm, float* datt, int64_t ldda,
                                     ccn_stream_t stream) {
    CCN_REQUIRE(nq >= 0 && c > 0 && row_ptr && dout && dmsg && mode >= 0 && mode <= 5);
    CCN_REQUIRE((mode != CCN_AGG_WSUM && mode != CCN_AGG_ATTEND && mode != CCN_AGG_WSUM_NORM) || (att && datt && msg));
    CCN_REQUIRE(mode != CCN_AGG_MAX || msg);
    if (nq == 0) return CCN_OK;
    ccn_launch(seg_aggregate_bwd_kernel, dim3(ELT_GRID(nq * c)), dim3(256), 0, S_, msg, ldm, att, lda, row_ptr, nq, c, mode, dout, ldo,
                                                               dmsg, lddm, datt, ldda);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_group_max_dense_bwd(const float* feat, int64_t ldf, const int32_t* idx, int64_t n, int32_t K,
                                       int32_t c, const float* dout, int64_t ldo, float* dfeat, int64_t lddf,
                                       ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && K >= 0 && c > 0 && feat && dout && dfeat && (K == 0 || idx));
    if (n == 0) return CCN_OK;
    ccn_launch(group_max_dense_bwd_kernel, dim3(ELT_GRID(n * c)), dim3(256), 0, S_, feat, ldf, idx, n, K, c, dout, ldo, dfeat, lddf);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_interp_blend_bwd(const int32_t* nbr, const float* wgt, int32_t from_d2, int64_t n, int32_t knn,
                                    int32_t c, const float* dout, int64_t ldo, float* dx, int64_t lddx,
                                    ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && knn >= 1 && c > 0 && nbr && wgt && dout && dx);
    if (n == 0) return CCN_OK;
    ccn_launch(interp_blend_bwd_kernel, dim3(ELT_GRID(n * c)), dim3(256), 0, S_, nbr, wgt, from_d2, n, knn, c, dout, ldo, dx, lddx);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_curve_feat_prep_bwd(const float* x, int64_t ldx, int32_t cx, int32_t with_xyz, int32_t with_diff,
                                       const int32_t* curve_ord, int64_t n, const float* dout, int64_t ldo, float* dx,
                                       int64_t lddx, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && cx > 0 && x && dout && dx && (!with_diff || curve_ord));
    if (n == 0) return CCN_OK;
    const int c_in = cx + (with_xyz ? 3 : 0);
    ccn_launch(curve_feat_prep_bwd_kernel, dim3(ELT_GRID(n * cx)), dim3(256), 0, S_, x, ldx, cx, c_in, with_diff, curve_ord, n, dout, ldo,
                                                                 dx, lddx);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_edge_rel_pos(const float* pos_src, const float* pos_dst, const int32_t* erow, const int32_t* ecol,
                                int64_t E, float normalize_radius, float* rel, ccn_stream_t stream) {
    CCN_REQUIRE(E >= 0 && pos_src && pos_dst && erow && ecol && rel);
    if (E == 0) return CCN_OK;
    ccn_launch(edge_rel_pos_kernel, dim3(ELT_GRID(E * 3)), dim3(256), 0, S_, pos_src, pos_dst, erow, ecol, E, normalize_radius, rel);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_nll_loss_fwd(const float* logits, int64_t ld, const int64_t* target, const float* class_weight,
                                int64_t n, int32_t C, int64_t ignore_index, float* lse, double* acc, float* loss,
                                float* perpoint, ccn_stream_t stream) {
    CCN_REQUIRE(n > 0 && C > 0 && logits && target && lse && acc && loss && ld >= C);
    cudaError_t e = cudaMemsetAsync(acc, 0, 2 * sizeof(double), S_);
    if (e != cudaSuccess) return (int)e;
    ccn_launch(nll_fwd_kernel, dim3(ccn_grid(n, 128, 148 * 8)), dim3(128), 0, S_, logits, ld, target, class_weight, n, C, ignore_index, lse,
                                                              acc, perpoint);
    CCN_CHECK_LAUNCH();
    ccn_launch(nll_finish_kernel, dim3(1), dim3(1), 0, S_, acc, loss);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_nll_loss_bwd(const float* logits, int64_t ld, const int64_t* target, const float* class_weight,
                                int64_t n, int32_t C, int64_t ignore_index, const float* lse, const double* acc,
                                const float* gscale, float* dlogits, int64_t ldd, ccn_stream_t stream) {
    CCN_REQUIRE(n > 0 && C > 0 && logits && target && lse && acc && gscale && dlogits && ld >= C && ldd >= C);
    ccn_launch(nll_bwd_kernel, dim3(ELT_GRID(n * C)), dim3(256), 0, S_, logits, ld, target, class_weight, n, C, ignore_index, lse, acc,
                                                    gscale, dlogits, ldd);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_dropout(const float* x, int64_t ldx, const float* u, int64_t n, int32_t c, float p, float* y,
                           int64_t ldy, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && c > 0 && x && u && y && p >= 0.f && p < 1.f);
    if (n == 0) return CCN_OK;
    ccn_launch(dropout_kernel, dim3(ELT_GRID(n * c)), dim3(256), 0, S_, x, ldx, u, n, c, p, y, ldy);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
