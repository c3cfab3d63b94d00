// Symmetric 1-D convolution along curves (reference
// src/models/modules/fast_conv1d.py:11-205).
//
// The reference inserts zero rows between curves, transposes, runs cuDNN conv1d over one long
// "image", applies BatchNorm + LeakyReLU as separate kernels and compacts with a boolean mask
// (three host syncs).  Here the padded signal exists only as an index map (row2pt); every layer is
// one kernel: tap-shifted FMA over a shared-memory tile of the signal, BatchNorm (eval) folded into
// the mirrored weights, LeakyReLU in the epilogue, pad rows skipped on the last layer.
#include <cuda_bf16.h>

#include "common.cuh"

namespace {

// out[i] = [x[i], pos[i], |(fwd + bwd) / max(nf + nb, 1)|]   (fast_conv1d.py:42-45,64-66,190-205)
__global__ void curve_feat_prep_kernel(const float* __restrict__ x, int64_t ldx, int cx,
                                       const float* __restrict__ pos, int with_xyz, int with_diff,
                                       const int32_t* __restrict__ ord, int64_t n,
                                       float* __restrict__ out, int64_t ldo) {
    ccn_pdl_wait();
    const int c_in = cx + (with_xyz ? 3 : 0);
    const int64_t total = n * (int64_t)c_in;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t i;
        int c;
        ccn_divmod(t, c_in, i, c);
        auto feat = [&](int64_t r) -> float { return c < cx ? x[r * ldx + c] : pos[r * 3 + (c - cx)]; };
        const float v = feat(i);
        out[i * ldo + c] = v;
        if (with_diff) {
            const int o = ord[i];
            const bool vb = i > 0 && ord[i - 1] == o;
            const bool vf = i + 1 < n && ord[i + 1] == o;
            const float eb = vb ? __fsub_rn(v, feat(i - 1)) : 0.f;
            const float ef = vf ? __fsub_rn(feat(i + 1), v) : 0.f;
            const float den = fmaxf((vf ? 1.f : 0.f) + (vb ? 1.f : 0.f), 1.f);
            out[i * ldo + c_in + c] = fabsf(__fdiv_rn(__fadd_rn(ef, eb), den));
        }
    }
}

// feat_prep + row gather + plane packing in one pass: the bf16 planes of the padded signal rows [x, pos, |diff|] that the
// first tensor-core conv layer reads, without the fp32 feature matrix (n x 2 c_in), its gathered copy (rows x 2 c_in) and
// the pack kernel's read of it.  Same fp32 values as curve_feat_prep_kernel, same split as ccn_pack_planes: bit-identical.
// One thread per (row, 8 output columns); pad rows (row2pt < 0) and pad columns are zeros.
__global__ void curve_feat_planes_kernel(const float* __restrict__ x, int64_t ldx, int cx, const float* __restrict__ pos,
                                         int with_xyz, int with_diff, const int32_t* __restrict__ ord, int64_t n,
                                         const int32_t* __restrict__ row2pt, int64_t rows, __nv_bfloat16* __restrict__ planes,
                                         int64_t kpad, int64_t plane_stride, int n_planes) {
    ccn_pdl_wait();
    const int c_in = cx + (with_xyz ? 3 : 0), C = with_diff ? 2 * c_in : c_in;
    const int groups = (int)(kpad >> 3);
    const int64_t total = rows * groups;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t r;
        int g;
        ccn_divmod(t, groups, r, g);
        const int64_t i = row2pt[r];
        float y[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) y[e] = 0.f;
        if (i >= 0) {
            const int o = with_diff ? ord[i] : 0;
            const bool vb = with_diff && i > 0 && ord[i - 1] == o;
            const bool vf = with_diff && i + 1 < n && ord[i + 1] == o;
            const float den = fmaxf((vf ? 1.f : 0.f) + (vb ? 1.f : 0.f), 1.f);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int col = 8 * g + e;
                if (col >= C) break;
                const int c = col < c_in ? col : col - c_in;
                auto feat = [&](int64_t q) -> float { return c < cx ? x[q * ldx + c] : pos[q * 3 + (c - cx)]; };
                const float v = feat(i);
                if (col < c_in) { y[e] = v; continue; }
                const float eb = vb ? __fsub_rn(v, feat(i - 1)) : 0.f;
                const float ef = vf ? __fsub_rn(feat(i + 1), v) : 0.f;
                y[e] = fabsf(__fdiv_rn(__fadd_rn(ef, eb), den));
            }
        }
        __nv_bfloat162 h[4], l[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            h[e] = __floats2bfloat162_rn(y[2 * e], y[2 * e + 1]);
            const float2 f = __bfloat1622float2(h[e]);
            l[e] = __floats2bfloat162_rn(y[2 * e] - f.x, y[2 * e + 1] - f.y);
        }
        __nv_bfloat16* op = planes + r * kpad + 8 * g;
        *reinterpret_cast<uint4*>(op) = *reinterpret_cast<uint4*>(h);
        if (n_planes == 2) *reinterpret_cast<uint4*>(op + plane_stride) = *reinterpret_cast<uint4*>(l);
    }
}

__global__ void row_map_kernel(const int32_t* __restrict__ ord, int64_t n, int pad, int32_t* __restrict__ row2pt) {
    ccn_pdl_wait();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) row2pt[i + (int64_t)pad * (ord[i] + 1)] = (int32_t)i;
}

constexpr int kTR = 64;        // output rows per CTA
constexpr int kRPT = 8;        // rows per thread (8 row groups x 32 output channels = 256 threads)
constexpr int kCK = 32;        // input channels staged per round
constexpr int kMaxHalf = 4;    // kernel sizes up to 9

// w: (ksize, cin, cout) with BatchNorm scale folded in, bias: (cout) with BatchNorm shift folded in.
template <int HALF>
__global__ void __launch_bounds__(256) curve_conv_layer_kernel(
    const float* __restrict__ in, int64_t ld_in, const float* __restrict__ w, const float* __restrict__ bias,
    float* __restrict__ out, int64_t ld_out, const int32_t* __restrict__ row2pt, int64_t rows,
    int cin, int cout, int in_mapped, int out_mapped, float slope) {
    ccn_pdl_wait();
    constexpr int K = 2 * HALF + 1;
    constexpr int TROWS = kTR + 2 * HALF;
    __shared__ float tile[TROWS][kCK];
    __shared__ int32_t src_row[TROWS];      // input row of every staged row, -1 = zero row

    const int64_t r0 = (int64_t)blockIdx.x * kTR;
    const int co = blockIdx.y * 32 + (threadIdx.x & 31);
    const int rg = threadIdx.x >> 5;
    for (int t = threadIdx.x; t < TROWS; t += blockDim.x) {
        const int64_t r = r0 - HALF + t;
        int32_t s = -1;
        if (r >= 0 && r < rows) s = in_mapped ? row2pt[r] : (int32_t)r;
        src_row[t] = s;
    }
    float acc[kRPT];
#pragma unroll
    for (int j = 0; j < kRPT; ++j) acc[j] = 0.f;

    for (int c0 = 0; c0 < cin; c0 += kCK) {
        __syncthreads();
        for (int t = threadIdx.x; t < TROWS * kCK; t += blockDim.x) {
            const int tr = t / kCK, tc = t - tr * kCK;
            const int32_t s = src_row[tr];
            const int c = c0 + tc;
            tile[tr][tc] = (s >= 0 && c < cin) ? in[(int64_t)s * ld_in + c] : 0.f;
        }
        __syncthreads();
        const int cend = min(kCK, cin - c0);
        if (co < cout) {
            for (int ci = 0; ci < cend; ++ci) {
                float v[kRPT + 2 * HALF];
#pragma unroll
                for (int j = 0; j < kRPT + 2 * HALF; ++j) v[j] = tile[rg * kRPT + j][ci];
#pragma unroll
                for (int tap = 0; tap < K; ++tap) {
                    const float wv = w[((int64_t)tap * cin + (c0 + ci)) * cout + co];
#pragma unroll
                    for (int j = 0; j < kRPT; ++j) acc[j] = fmaf(wv, v[j + tap], acc[j]);
                }
            }
        }
    }
    if (co >= cout) return;
    const float b = bias[co];
#pragma unroll
    for (int j = 0; j < kRPT; ++j) {
        const int64_t r = r0 + rg * kRPT + j;
        if (r >= rows) break;
        int64_t o = r;
        if (out_mapped) {
            o = row2pt[r];
            if (o < 0) continue;
        }
        float y = acc[j] + b;
        y = y > 0.f ? y : y * slope;
        out[o * ld_out + co] = y;
    }
}

}  // namespace

extern "C" int ccn_curve_feat_prep(const float* x, int64_t ldx, int32_t cx, const float* pos, int32_t with_xyz,
                                   int32_t with_diff, const int32_t* curve_ord, int64_t n, float* out, int64_t ldo,
                                   ccn_stream_t stream) {
    const int c_in = cx + (with_xyz ? 3 : 0);
    CCN_REQUIRE(n >= 0 && c_in > 0 && out && (cx == 0 || x) && (!with_xyz || pos) && (!with_diff || curve_ord));
    CCN_REQUIRE(ldo >= (with_diff ? 2 : 1) * c_in);
    if (n == 0) return CCN_OK;
    ccn_launch(curve_feat_prep_kernel, dim3(ccn_grid(n * c_in, 256)), dim3(256), 0, (cudaStream_t)stream, x, ldx, cx, pos, with_xyz,
                                                                                     with_diff, curve_ord, n, out, ldo);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_curve_feat_planes(const float* x, int64_t ldx, int32_t cx, const float* pos, int32_t with_xyz, int32_t with_diff,
                                     const int32_t* curve_ord, int64_t n, const int32_t* row2pt, int64_t rows, void* planes,
                                     int64_t kpad, int64_t plane_stride, int32_t n_planes, ccn_stream_t stream) {
    const int c_in = cx + (with_xyz ? 3 : 0);
    CCN_REQUIRE(n >= 0 && rows >= 0 && c_in > 0 && planes && row2pt && (cx == 0 || x) && (!with_xyz || pos) && (!with_diff || curve_ord));
    CCN_REQUIRE((kpad & 7) == 0 && kpad >= (with_diff ? 2 : 1) * c_in && (n_planes == 1 || (n_planes == 2 && (plane_stride & 7) == 0)) &&
                (reinterpret_cast<uintptr_t>(planes) & 15) == 0);
    if (rows == 0) return CCN_OK;
    ccn_launch(curve_feat_planes_kernel, dim3(ccn_grid(rows * (kpad >> 3), 256, 148LL * 32)), dim3(256), 0, (cudaStream_t)stream, x, ldx,
               (int)cx, pos, (int)with_xyz, (int)with_diff, curve_ord, n, row2pt, rows, (__nv_bfloat16*)planes, kpad, plane_stride,
               (int)n_planes);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_curve_row_map(const int32_t* curve_ord, int64_t n, int32_t n_curves, int32_t pad,
                                 int32_t* row2pt, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && n_curves >= 0 && pad >= 0 && row2pt && (n == 0 || curve_ord));
    const int64_t rows = n + (int64_t)(n_curves + 1) * pad;
    cudaError_t e = cudaMemsetAsync(row2pt, 0xff, rows * sizeof(int32_t), (cudaStream_t)stream);
    if (e != cudaSuccess) return (int)e;
    if (n == 0) return CCN_OK;
    ccn_launch(row_map_kernel, dim3(ccn_blocks_exact(n, 256)), dim3(256), 0, (cudaStream_t)stream, curve_ord, n, pad, row2pt);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_curve_conv_layer(const float* in, int64_t ld_in, const float* w, const float* bias, float* out,
                                    int64_t ld_out, const int32_t* row2pt, int64_t rows, int32_t cin, int32_t cout,
                                    int32_t ksize, int32_t in_mapped, int32_t out_mapped, float slope,
                                    ccn_stream_t stream) {
    CCN_REQUIRE(in && w && bias && out && rows >= 0 && cin > 0 && cout > 0);
    CCN_REQUIRE((ksize & 1) && ksize >= 1 && ksize <= 2 * kMaxHalf + 1);
    CCN_REQUIRE(!(in_mapped || out_mapped) || row2pt);
    if (rows == 0) return CCN_OK;
    dim3 grid(ccn_blocks_exact(rows, kTR), (cout + 31) / 32);
    cudaStream_t s = (cudaStream_t)stream;
#define CCN_LAUNCH_CONV(H)                                                                                     \
    ccn_launch(curve_conv_layer_kernel<H>, dim3(grid), dim3(256), 0, s, in, ld_in, w, bias, out, ld_out, row2pt, rows, cin, cout, \
                                                    in_mapped, out_mapped, slope)
    switch (ksize / 2) {
        case 0: CCN_LAUNCH_CONV(0); break;
        case 1: CCN_LAUNCH_CONV(1); break;
        case 2: CCN_LAUNCH_CONV(2); break;
        case 3: CCN_LAUNCH_CONV(3); break;
        default: CCN_LAUNCH_CONV(4); break;
    }
#undef CCN_LAUNCH_CONV
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
