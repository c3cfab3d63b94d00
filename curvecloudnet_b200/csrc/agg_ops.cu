// Gather / segmented aggregation kernels around the shared MLPs.
// They replace the reference's materialised x[col] gathers, torch_scatter atomics
// (scatter_max/add/mean, reference src/models/modules/point_conv.py:71-93), torch_geometric
// segment softmax, frnn_gather (third_party/FRNN/frnn/frnn.py:351-397) and the dense masked max of
// StaticEdgeConv.forward_fast (src/models/modules/dgcnn.py:158-207).  All sums run in edge order
// (deterministic), one thread per (segment, channel) with channels on the fast axis.
#include <cuda_bf16.h>

#include "common.cuh"

namespace {

__device__ __forceinline__ float apply_act(float y, int act, float slope) {
    if (act == CCN_ACT_RELU) return fmaxf(y, 0.f);
    if (act == CCN_ACT_LEAKY) return y > 0.f ? y : y * slope;
    return y;
}

__global__ void gather_rows_kernel(const float* __restrict__ src, int64_t lds, const int64_t* __restrict__ idx,
                                   int64_t n, int c, float* __restrict__ dst, int64_t ldd) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)c;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t i;
        int k;
        ccn_divmod(t, c, i, k);
        const int64_t s = idx ? idx[i] : i;
        dst[i * ldd + k] = src[s * lds + k];
    }
}

__global__ void gather_rows_v4_kernel(const float* __restrict__ src, int64_t lds, const int64_t* __restrict__ idx,
                                      int64_t n, int c4, float* __restrict__ dst, int64_t ldd) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)c4;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t i;
        int k;
        ccn_divmod(t, c4, i, k);
        const int64_t s = idx ? idx[i] : i;
        *reinterpret_cast<float4*>(dst + i * ldd + 4 * k) = *reinterpret_cast<const float4*>(src + s * lds + 4 * k);
    }
}

__global__ void gather_i64_kernel(const int64_t* __restrict__ src, const int64_t* __restrict__ idx, int64_t n,
                                  int64_t* __restrict__ dst) {
    ccn_pdl_wait();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[idx[i]];
}

// CSR from a dense (nq, K) neighbour table whose valid entries lead every row.
__global__ void dense_to_csr_kernel(const int32_t* __restrict__ idx, const int32_t* __restrict__ row_ptr, int64_t nq,
                                    int K, int32_t* __restrict__ erow, int32_t* __restrict__ ecol) {
    ccn_pdl_wait();
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= nq * K) return;
    int64_t q;
    int k;
    ccn_divmod(t, K, q, k);
    const int b = row_ptr[q], cnt = row_ptr[q + 1] - b;
    if (k < cnt) {
        erow[b + k] = (int32_t)q;
        ecol[b + k] = idx[t];
    }
}

__global__ void csr_rows_kernel(const int32_t* __restrict__ row_ptr, int64_t nq, int32_t* __restrict__ erow) {
    ccn_pdl_wait();
    const int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (q >= nq) return;
    for (int e = row_ptr[q]; e < row_ptr[q + 1]; ++e) erow[e] = (int32_t)q;
}

// First layer of PointNetConv2's local_nn, decomposed (point_conv.py:60-69):
//   W [x_j ; (pos_j - pos_i)/r] + b = H[j] + Wp (pos_j - pos_i)/r + b,   H = x W_x^T computed per SOURCE point.
__global__ void edge_layer1_kernel(const float* __restrict__ H, int64_t ldh, const float* __restrict__ wp,
                                   const float* __restrict__ bias, const float* __restrict__ pos_src,
                                   const float* __restrict__ pos_dst, const int32_t* __restrict__ erow,
                                   const int32_t* __restrict__ ecol, int64_t E, int c, float inv_norm_is_div,
                                   float radius, int act, float slope, float* __restrict__ out, int64_t ldo) {
    ccn_pdl_wait();
    const int64_t total = E * (int64_t)c;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t e;
        int k;
        ccn_divmod(t, c, e, k);
        const int j = ecol[e], i = erow[e];
        float rx = __fsub_rn(pos_src[3 * j], pos_dst[3 * i]);
        float ry = __fsub_rn(pos_src[3 * j + 1], pos_dst[3 * i + 1]);
        float rz = __fsub_rn(pos_src[3 * j + 2], pos_dst[3 * i + 2]);
        if (inv_norm_is_div != 0.f) { rx = __fdiv_rn(rx, radius); ry = __fdiv_rn(ry, radius); rz = __fdiv_rn(rz, radius); }
        float y = bias ? bias[k] : 0.f;
        if (H) y += H[(int64_t)j * ldh + k];
        y = fmaf(wp[k], rx, y);
        y = fmaf(wp[c + k], ry, y);
        y = fmaf(wp[2 * c + k], rz, y);
        out[e * ldo + k] = apply_act(y, act, slope);
    }
}

// Segmented aggregation over CSR rows.  mode: 0 max (empty -> 0), 1 mean, 2 sum sigmoid(att)*msg,
// 3 sum softmax_row(att)*msg with exp(a - max)/(sum + 1e-16) (torch_geometric.utils.softmax).
__global__ void seg_aggregate_kernel(const float* __restrict__ msg, int64_t ldm, const float* __restrict__ att,
                                     int64_t lda, const int32_t* __restrict__ row_ptr, int64_t nq, int c, int mode,
                                     float* __restrict__ out, int64_t ldo) {
    ccn_pdl_wait();
    const int64_t total = nq * (int64_t)c;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t q;
        int k;
        ccn_divmod(t, c, q, k);
        const int b = row_ptr[q], e = row_ptr[q + 1];
        float r = 0.f;
        if (mode == CCN_AGG_MAX) {
            if (e > b) {
                r = msg[(int64_t)b * ldm + k];
                for (int j = b + 1; j < e; ++j) r = fmaxf(r, msg[(int64_t)j * ldm + k]);
            }
        } else if (mode == CCN_AGG_MEAN || mode == CCN_AGG_SUM) {
            for (int j = b; j < e; ++j) r += msg[(int64_t)j * ldm + k];
            if (mode == CCN_AGG_MEAN) r = r / fmaxf((float)(e - b), 1.f);
        } else if (mode == CCN_AGG_WSUM) {
            for (int j = b; j < e; ++j) {
                const float a = att[(int64_t)j * lda + k];
                r += msg[(int64_t)j * ldm + k] * (1.f / (1.f + expf(-a)));
            }
        } else if (mode == CCN_AGG_WSUM_NORM) {
            // dense SGCNN 'weighted-sum' (dgcnn.py:190-197): w = sigmoid(att) / clamp(sum w, 1e-3)
            float tot = 0.f;
            for (int j = b; j < e; ++j) tot += 1.f / (1.f + expf(-att[(int64_t)j * lda + k]));
            const float den = fmaxf(tot, 1e-3f);
            for (int j = b; j < e; ++j)
                r += msg[(int64_t)j * ldm + k] * ((1.f / (1.f + expf(-att[(int64_t)j * lda + k]))) / den);
        } else {
            float m = -INFINITY;
            for (int j = b; j < e; ++j) m = fmaxf(m, att[(int64_t)j * lda + k]);
            float s = 0.f;
            for (int j = b; j < e; ++j) s += expf(att[(int64_t)j * lda + k] - m);
            const float inv = 1.f / (s + 1e-16f);
            for (int j = b; j < e; ++j) r += msg[(int64_t)j * ldm + k] * (expf(att[(int64_t)j * lda + k] - m) * inv);
        }
        out[q * ldo + k] = r;
    }
}

// StaticEdgeConv.forward_fast first layer, decomposed (dgcnn.py:166-173):
//   W [x_j ; x_i - x_j] + b = (Wa - Wb) x_j + (Wb x_i + b) = P[j] + Q[i];  x_j = 0 for missing neighbours.
// Row layout: (i, 0) = self loop, (i, 1+k) = idx[i][k].
__global__ void sgcnn_layer1_kernel(const float* __restrict__ PQ, int64_t ldpq, const int32_t* __restrict__ idx,
                                    int64_t n, int K, int c, int act, float slope, float* __restrict__ out,
                                    int64_t ldo) {
    ccn_pdl_wait();
    const int G = K + 1;
    const int64_t total = n * (int64_t)G * c;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t row;
        int k;
        ccn_divmod(t, c, row, k);
        const int64_t i = row / G;
        const int g = (int)(row - i * G);
        const int j = g == 0 ? (int)i : idx[i * K + g - 1];
        float y = PQ[i * ldpq + c + k];
        if (j >= 0) y += PQ[(int64_t)j * ldpq + k];
        out[row * ldo + k] = apply_act(y, act, slope);
    }
}

__global__ void edge_pair_layer1_kernel(const float* __restrict__ PQ, int64_t ldpq, const int32_t* __restrict__ erow,
                                        const int32_t* __restrict__ ecol, int64_t E, int c, int act, float slope,
                                        float* __restrict__ out, int64_t ldo) {
    ccn_pdl_wait();
    const int64_t total = E * (int64_t)c;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t e;
        int k;
        ccn_divmod(t, c, e, k);
        const float y = PQ[(int64_t)ecol[e] * ldpq + k] + PQ[(int64_t)erow[e] * ldpq + c + k];
        out[e * ldo + k] = apply_act(y, act, slope);
    }
}

// max over the K+1 rows of every point, invalid rows count as -100 (dgcnn.py:187-189).
__global__ void group_max_dense_kernel(const float* __restrict__ feat, int64_t ldf, const int32_t* __restrict__ idx,
                                       int64_t n, int K, int c, float* __restrict__ out, int64_t ldo) {
    ccn_pdl_wait();
    const int G = K + 1;
    const int64_t total = n * (int64_t)c;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t i;
        int k;
        ccn_divmod(t, c, i, k);
        float r = feat[(i * G) * ldf + k];
        for (int g = 1; g < G; ++g) {
            const float v = idx[i * K + g - 1] >= 0 ? feat[(i * G + g) * ldf + k] : -100.f;
            r = fmaxf(r, v);
        }
        out[i * ldo + k] = r;
    }
}

// ---- float4 forms of the two reductions above (c % 4 == 0, 16-byte aligned rows): a thread owns 4 adjacent channels
// of one target; every channel sees exactly the scalar kernel's operation order, so results are bit-identical.
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 max4(float4 a, float4 b) {
    return make_float4(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z), fmaxf(a.w, b.w));
}

// fp32 result AND its activation planes from the producing kernel (the consumer is a Linear: saves the pack launch):
// 4 channels -> 4 bf16 hi (+ 4 bf16 lo), one 8-byte store per plane
__device__ __forceinline__ void st_planes4(const float4 y, __nv_bfloat16* o, int64_t plane_stride, int n_planes) {
    const __nv_bfloat162 h0 = __floats2bfloat162_rn(y.x, y.y), h1 = __floats2bfloat162_rn(y.z, y.w);
    uint2 hv;
    hv.x = *reinterpret_cast<const uint32_t*>(&h0);
    hv.y = *reinterpret_cast<const uint32_t*>(&h1);
    *reinterpret_cast<uint2*>(o) = hv;
    if (n_planes == 2) {
        const float2 f0 = __bfloat1622float2(h0), f1 = __bfloat1622float2(h1);
        const __nv_bfloat162 l0 = __floats2bfloat162_rn(y.x - f0.x, y.y - f0.y), l1 = __floats2bfloat162_rn(y.z - f1.x, y.w - f1.y);
        uint2 lv;
        lv.x = *reinterpret_cast<const uint32_t*>(&l0);
        lv.y = *reinterpret_cast<const uint32_t*>(&l1);
        *reinterpret_cast<uint2*>(o + plane_stride) = lv;
    }
}

__global__ void group_max_dense_v4_kernel(const float* __restrict__ feat, int64_t ldf, const int32_t* __restrict__ idx,
                                          int64_t n, int K, int c4, float* __restrict__ out, int64_t ldo,
                                          __nv_bfloat16* __restrict__ planes, int64_t kpad, int64_t pstride, int npl) {
    ccn_pdl_wait();
    const int G = K + 1;
    const int64_t total = n * (int64_t)c4;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t i;
        int k4;
        ccn_divmod(t, c4, i, k4);
        const float* base = feat + (i * G) * ldf + 4 * k4;
        float4 r = ld4(base);
        const float4 skip = make_float4(-100.f, -100.f, -100.f, -100.f);
#pragma unroll 4
        for (int g = 1; g < G; ++g) {
            const float4 v = idx[i * K + g - 1] >= 0 ? ld4(base + (int64_t)g * ldf) : skip;
            r = max4(r, v);
        }
        *reinterpret_cast<float4*>(out + i * ldo + 4 * k4) = r;
        if (planes) st_planes4(r, planes + i * kpad + 4 * k4, pstride, npl);
    }
}

__global__ void seg_aggregate_v4_kernel(const float* __restrict__ msg, int64_t ldm, const float* __restrict__ att,
                                        int64_t lda, const int32_t* __restrict__ row_ptr, int64_t nq, int c4, int mode,
                                        float* __restrict__ out, int64_t ldo, __nv_bfloat16* __restrict__ planes,
                                        int64_t kpad, int64_t pstride, int npl) {
    ccn_pdl_wait();
    const int64_t total = nq * (int64_t)c4;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t q;
        int k4;
        ccn_divmod(t, c4, q, k4);
        const int b = row_ptr[q], e = row_ptr[q + 1];
        const float* m0 = msg + 4 * k4;
        const float* a0 = att ? att + 4 * k4 : nullptr;
        float r[4] = {0.f, 0.f, 0.f, 0.f};
        if (mode == CCN_AGG_MAX) {
            if (e > b) {
                float4 v = ld4(m0 + (int64_t)b * ldm);
#pragma unroll 4
                for (int j = b + 1; j < e; ++j) v = max4(v, ld4(m0 + (int64_t)j * ldm));
                r[0] = v.x; r[1] = v.y; r[2] = v.z; r[3] = v.w;
            }
        } else if (mode == CCN_AGG_MEAN || mode == CCN_AGG_SUM) {
#pragma unroll 4
            for (int j = b; j < e; ++j) {
                const float4 v = ld4(m0 + (int64_t)j * ldm);
                r[0] += v.x; r[1] += v.y; r[2] += v.z; r[3] += v.w;
            }
            if (mode == CCN_AGG_MEAN) {
                const float d = fmaxf((float)(e - b), 1.f);
#pragma unroll
                for (int l = 0; l < 4; ++l) r[l] = r[l] / d;
            }
        } else if (mode == CCN_AGG_WSUM) {
#pragma unroll 2
            for (int j = b; j < e; ++j) {
                const float4 a = ld4(a0 + (int64_t)j * lda), v = ld4(m0 + (int64_t)j * ldm);
                const float av[4] = {a.x, a.y, a.z, a.w}, mv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int l = 0; l < 4; ++l) r[l] += mv[l] * (1.f / (1.f + expf(-av[l])));
            }
        } else if (mode == CCN_AGG_WSUM_NORM) {
            float tot[4] = {0.f, 0.f, 0.f, 0.f};
            for (int j = b; j < e; ++j) {
                const float4 a = ld4(a0 + (int64_t)j * lda);
                tot[0] += 1.f / (1.f + expf(-a.x)); tot[1] += 1.f / (1.f + expf(-a.y));
                tot[2] += 1.f / (1.f + expf(-a.z)); tot[3] += 1.f / (1.f + expf(-a.w));
            }
            for (int j = b; j < e; ++j) {
                const float4 a = ld4(a0 + (int64_t)j * lda), v = ld4(m0 + (int64_t)j * ldm);
                const float av[4] = {a.x, a.y, a.z, a.w}, mv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int l = 0; l < 4; ++l) r[l] += mv[l] * ((1.f / (1.f + expf(-av[l]))) / fmaxf(tot[l], 1e-3f));
            }
        } else {
            float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY}, sum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
            for (int j = b; j < e; ++j) {
                const float4 a = ld4(a0 + (int64_t)j * lda);
                mx[0] = fmaxf(mx[0], a.x); mx[1] = fmaxf(mx[1], a.y); mx[2] = fmaxf(mx[2], a.z); mx[3] = fmaxf(mx[3], a.w);
            }
#pragma unroll 2
            for (int j = b; j < e; ++j) {
                const float4 a = ld4(a0 + (int64_t)j * lda);
                sum[0] += expf(a.x - mx[0]); sum[1] += expf(a.y - mx[1]); sum[2] += expf(a.z - mx[2]); sum[3] += expf(a.w - mx[3]);
            }
            float inv[4];
#pragma unroll
            for (int l = 0; l < 4; ++l) inv[l] = 1.f / (sum[l] + 1e-16f);
#pragma unroll 2
            for (int j = b; j < e; ++j) {
                const float4 a = ld4(a0 + (int64_t)j * lda), v = ld4(m0 + (int64_t)j * ldm);
                r[0] += v.x * (expf(a.x - mx[0]) * inv[0]); r[1] += v.y * (expf(a.y - mx[1]) * inv[1]);
                r[2] += v.z * (expf(a.z - mx[2]) * inv[2]); r[3] += v.w * (expf(a.w - mx[3]) * inv[3]);
            }
        }
        const float4 rv = make_float4(r[0], r[1], r[2], r[3]);
        *reinterpret_cast<float4*>(out + q * ldo + 4 * k4) = rv;
        if (planes) st_planes4(rv, planes + q * kpad + 4 * k4, pstride, npl);
    }
}

// y[i] = sum_k w x[nbr] / sum_k w   (point_ops.py:336-341, 350-355); w = 1/clamp(d2, 1e-16) when from_d2.
__global__ void interp_blend_kernel(const float* __restrict__ x, int64_t ldx, const int32_t* __restrict__ nbr,
                                    const float* __restrict__ wgt, int from_d2, int64_t n, int knn, int c,
                                    float* __restrict__ out, int64_t ldo) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)c;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t i;
        int k;
        ccn_divmod(t, c, i, k);
        float num = 0.f, den = 0.f;
        for (int s = 0; s < knn; ++s) {
            const int j = nbr[i * knn + s];
            if (j < 0) continue;
            float w = wgt[i * knn + s];
            if (from_d2) w = __fdiv_rn(1.f, fmaxf(w, 1e-16f));
            num = __fadd_rn(num, __fmul_rn(x[(int64_t)j * ldx + k], w));
            den = __fadd_rn(den, w);
        }
        out[i * ldo + k] = __fdiv_rn(num, den);
    }
}


__global__ void interp_blend_v4_kernel(const float* __restrict__ x, int64_t ldx, const int32_t* __restrict__ nbr,
                                       const float* __restrict__ wgt, int from_d2, int64_t n, int knn, int c4,
                                       float* __restrict__ out, int64_t ldo, __nv_bfloat16* __restrict__ planes,
                                       int64_t kpad, int64_t pstride, int npl) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)c4;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t i;
        int k4;
        ccn_divmod(t, c4, i, k4);
        float num[4] = {0.f, 0.f, 0.f, 0.f}, den = 0.f;
        for (int s = 0; s < knn; ++s) {
            const int j = nbr[i * knn + s];
            if (j < 0) continue;
            float w = wgt[i * knn + s];
            if (from_d2) w = __fdiv_rn(1.f, fmaxf(w, 1e-16f));
            const float4 v = ld4(x + (int64_t)j * ldx + 4 * k4);
            num[0] = __fadd_rn(num[0], __fmul_rn(v.x, w)); num[1] = __fadd_rn(num[1], __fmul_rn(v.y, w));
            num[2] = __fadd_rn(num[2], __fmul_rn(v.z, w)); num[3] = __fadd_rn(num[3], __fmul_rn(v.w, w));
            den = __fadd_rn(den, w);
        }
        const float4 rv = make_float4(__fdiv_rn(num[0], den), __fdiv_rn(num[1], den), __fdiv_rn(num[2], den), __fdiv_rn(num[3], den));
        *reinterpret_cast<float4*>(out + i * ldo + 4 * k4) = rv;
        if (planes) st_planes4(rv, planes + i * kpad + 4 * k4, pstride, npl);
    }
}

// float4 forms of the fp32-output per-edge first layers (training keeps fp32 activations in front of BatchNorm)
__global__ void edge_layer1_v4_kernel(const float* __restrict__ H, int64_t ldh, const float* __restrict__ wp,
                                      const float* __restrict__ bias, const float* __restrict__ pos_src,
                                      const float* __restrict__ pos_dst, const int32_t* __restrict__ erow,
                                      const int32_t* __restrict__ ecol, int64_t E, int c, float inv_norm_is_div,
                                      float radius, int act, float slope, float* __restrict__ out, int64_t ldo) {
    ccn_pdl_wait();
    const int c4 = c >> 2;
    const int64_t total = E * (int64_t)c4;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t e;
        int k;
        ccn_divmod(t, c4, e, k);
        k <<= 2;
        const int j = ecol[e], i = erow[e];
        float rx = __fsub_rn(pos_src[3 * j], pos_dst[3 * i]);
        float ry = __fsub_rn(pos_src[3 * j + 1], pos_dst[3 * i + 1]);
        float rz = __fsub_rn(pos_src[3 * j + 2], pos_dst[3 * i + 2]);
        if (inv_norm_is_div != 0.f) { rx = __fdiv_rn(rx, radius); ry = __fdiv_rn(ry, radius); rz = __fdiv_rn(rz, radius); }
        float4 y = bias ? ld4(bias + k) : make_float4(0.f, 0.f, 0.f, 0.f);
        if (H) {
            const float4 h = ld4(H + (int64_t)j * ldh + k);
            y.x += h.x; y.y += h.y; y.z += h.z; y.w += h.w;
        }
        const float4 wx = ld4(wp + k), wy = ld4(wp + c + k), wz = ld4(wp + 2 * c + k);
        y.x = apply_act(fmaf(wz.x, rz, fmaf(wy.x, ry, fmaf(wx.x, rx, y.x))), act, slope);
        y.y = apply_act(fmaf(wz.y, rz, fmaf(wy.y, ry, fmaf(wx.y, rx, y.y))), act, slope);
        y.z = apply_act(fmaf(wz.z, rz, fmaf(wy.z, ry, fmaf(wx.z, rx, y.z))), act, slope);
        y.w = apply_act(fmaf(wz.w, rz, fmaf(wy.w, ry, fmaf(wx.w, rx, y.w))), act, slope);
        *reinterpret_cast<float4*>(out + e * ldo + k) = y;
    }
}

__global__ void edge_pair_layer1_v4_kernel(const float* __restrict__ PQ, int64_t ldpq, const int32_t* __restrict__ erow,
                                           const int32_t* __restrict__ ecol, int64_t E, int c, int act, float slope,
                                           float* __restrict__ out, int64_t ldo) {
    ccn_pdl_wait();
    const int c4 = c >> 2;
    const int64_t total = E * (int64_t)c4;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t e;
        int k;
        ccn_divmod(t, c4, e, k);
        k <<= 2;
        const float4 a = ld4(PQ + (int64_t)ecol[e] * ldpq + k), b = ld4(PQ + (int64_t)erow[e] * ldpq + c + k);
        *reinterpret_cast<float4*>(out + e * ldo + k) =
            make_float4(apply_act(a.x + b.x, act, slope), apply_act(a.y + b.y, act, slope), apply_act(a.z + b.z, act, slope),
                        apply_act(a.w + b.w, act, slope));
    }
}

// ---- planes-output variants of the per-edge first layers (the consumer is always the next Linear of the MLP):
// one thread per (row, 8 channels); requires c % 8 == 0 and 16-byte aligned fp32 rows.
struct F8 { float v[8]; };
__device__ __forceinline__ F8 ld8(const float* p) {
    F8 r;
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
    return r;
}
__device__ __forceinline__ void st_planes8(const F8& y, __nv_bfloat16* o, int64_t plane_stride, int n_planes) {
    __nv_bfloat162 h[4], l[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        h[e] = __floats2bfloat162_rn(y.v[2 * e], y.v[2 * e + 1]);
        const float2 f = __bfloat1622float2(h[e]);
        l[e] = __floats2bfloat162_rn(y.v[2 * e] - f.x, y.v[2 * e + 1] - f.y);
    }
    *reinterpret_cast<uint4*>(o) = *reinterpret_cast<uint4*>(h);
    if (n_planes == 2) *reinterpret_cast<uint4*>(o + plane_stride) = *reinterpret_cast<uint4*>(l);
}

__global__ void sgcnn_layer1_planes_kernel(const float* __restrict__ PQ, int64_t ldpq, const int32_t* __restrict__ idx,
                                           int64_t n, int K, int c, int act, float slope,
                                           __nv_bfloat16* __restrict__ planes, int64_t kpad, int64_t plane_stride,
                                           int n_planes) {
    ccn_pdl_wait();
    const int G = K + 1, cg = c >> 3;
    const int64_t total = n * (int64_t)G * cg;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t row;
        int k;
        ccn_divmod(t, cg, row, k);
        k <<= 3;
        const int64_t i = row / G;
        const int g = (int)(row - i * G);
        const int j = g == 0 ? (int)i : idx[i * K + g - 1];
        F8 y = ld8(PQ + i * ldpq + c + k);
        if (j >= 0) {
            const F8 pj = ld8(PQ + (int64_t)j * ldpq + k);
#pragma unroll
            for (int e = 0; e < 8; ++e) y.v[e] += pj.v[e];
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) y.v[e] = apply_act(y.v[e], act, slope);
        st_planes8(y, planes + row * kpad + k, plane_stride, n_planes);
    }
}

__global__ void edge_pair_layer1_planes_kernel(const float* __restrict__ PQ, int64_t ldpq, const int32_t* __restrict__ erow,
                                               const int32_t* __restrict__ ecol, int64_t E, int c, int act, float slope,
                                               __nv_bfloat16* __restrict__ planes, int64_t kpad, int64_t plane_stride,
                                               int n_planes) {
    ccn_pdl_wait();
    const int cg = c >> 3;
    const int64_t total = E * (int64_t)cg;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t e;
        int k;
        ccn_divmod(t, cg, e, k);
        k <<= 3;
        F8 y = ld8(PQ + (int64_t)ecol[e] * ldpq + k);
        const F8 q = ld8(PQ + (int64_t)erow[e] * ldpq + c + k);
#pragma unroll
        for (int u = 0; u < 8; ++u) y.v[u] = apply_act(y.v[u] + q.v[u], act, slope);
        st_planes8(y, planes + e * kpad + k, plane_stride, n_planes);
    }
}

__global__ void edge_layer1_planes_kernel(const float* __restrict__ H, int64_t ldh, const float* __restrict__ wp,
                                          const float* __restrict__ bias, const float* __restrict__ pos_src,
                                          const float* __restrict__ pos_dst, const int32_t* __restrict__ erow,
                                          const int32_t* __restrict__ ecol, int64_t E, int c, float inv_norm_is_div,
                                          float radius, int act, float slope, __nv_bfloat16* __restrict__ planes,
                                          int64_t kpad, int64_t plane_stride, int n_planes) {
    ccn_pdl_wait();
    const int cg = c >> 3;
    const int64_t total = E * (int64_t)cg;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t e;
        int k;
        ccn_divmod(t, cg, e, k);
        k <<= 3;
        const int j = ecol[e], i = erow[e];
        float rx = __fsub_rn(pos_src[3 * j], pos_dst[3 * i]);
        float ry = __fsub_rn(pos_src[3 * j + 1], pos_dst[3 * i + 1]);
        float rz = __fsub_rn(pos_src[3 * j + 2], pos_dst[3 * i + 2]);
        if (inv_norm_is_div != 0.f) { rx = __fdiv_rn(rx, radius); ry = __fdiv_rn(ry, radius); rz = __fdiv_rn(rz, radius); }
        F8 y;
        if (bias) y = ld8(bias + k);
        else {
#pragma unroll
            for (int u = 0; u < 8; ++u) y.v[u] = 0.f;
        }
        if (H) {
            const F8 h = ld8(H + (int64_t)j * ldh + k);
#pragma unroll
            for (int u = 0; u < 8; ++u) y.v[u] += h.v[u];
        }
        const F8 wx = ld8(wp + k), wy = ld8(wp + c + k), wz = ld8(wp + 2 * c + k);
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            float v = fmaf(wx.v[u], rx, y.v[u]);
            v = fmaf(wy.v[u], ry, v);
            v = fmaf(wz.v[u], rz, v);
            y.v[u] = apply_act(v, act, slope);
        }
        st_planes8(y, planes + e * kpad + k, plane_stride, n_planes);
    }
}

}  // namespace

#define CCN_ELT_GRID(total) ccn_grid((total), 256, 148LL * 32)
static inline bool ccn_al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int ccn_gather_rows_f32(const float* src, int64_t lds, const int64_t* idx, int64_t n, int32_t c,
                                   float* dst, int64_t ldd, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && c > 0 && src && dst && lds >= c && ldd >= c);
    if (n == 0) return CCN_OK;
    if ((c & 3) == 0 && (lds & 3) == 0 && (ldd & 3) == 0 && ccn_al16(src) && ccn_al16(dst)) {
        ccn_launch(gather_rows_v4_kernel, dim3(CCN_ELT_GRID(n * (c >> 2))), dim3(256), 0, (cudaStream_t)stream, src, lds, idx, n, c >> 2, dst, ldd);
        CCN_CHECK_LAUNCH();
        return CCN_OK;
    }
    ccn_launch(gather_rows_kernel, dim3(CCN_ELT_GRID(n * c)), dim3(256), 0, (cudaStream_t)stream, src, lds, idx, n, c, dst, ldd);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_gather_i64(const int64_t* src, const int64_t* idx, int64_t n, int64_t* dst, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && src && idx && dst);
    if (n == 0) return CCN_OK;
    ccn_launch(gather_i64_kernel, dim3(ccn_blocks_exact(n, 256)), dim3(256), 0, (cudaStream_t)stream, src, idx, n, dst);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

namespace {
// pos / batch / curve of the selected points in one pass (pos[idx], batch[idx], curve_idxs[idx] of pointnet2.py:69-72)
__global__ void subset_gather_kernel(const float* __restrict__ pos, int64_t ldp, const int64_t* __restrict__ batch,
                                     const int64_t* __restrict__ curve, const int64_t* __restrict__ idx, int64_t n,
                                     float* __restrict__ posq, int64_t* __restrict__ batchq, int64_t* __restrict__ curveq,
                                     __nv_bfloat16* __restrict__ planes, int64_t pstride, int npl) {
    ccn_pdl_wait();
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int64_t i = idx[t];
    const float x = pos[i * ldp], y = pos[i * ldp + 1], z = pos[i * ldp + 2];
    posq[3 * t] = x; posq[3 * t + 1] = y; posq[3 * t + 2] = z;
    batchq[t] = batch[i];
    if (curve) curveq[t] = curve[i];
    if (planes) {
        // activation planes of the (n, 3) position matrix (kpad = 16: xyz + 13 zero columns), ccn_pack_planes layout
        const __nv_bfloat162 h0 = __floats2bfloat162_rn(x, y), h1 = __floats2bfloat162_rn(z, 0.f);
        __nv_bfloat16* o = planes + t * 16;
        *reinterpret_cast<uint4*>(o) = make_uint4(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1), 0u, 0u);
        *reinterpret_cast<uint4*>(o + 8) = make_uint4(0u, 0u, 0u, 0u);
        if (npl == 2) {
            const float2 f0 = __bfloat1622float2(h0), f1 = __bfloat1622float2(h1);
            const __nv_bfloat162 l0 = __floats2bfloat162_rn(x - f0.x, y - f0.y), l1 = __floats2bfloat162_rn(z - f1.x, 0.f);
            o += pstride;
            *reinterpret_cast<uint4*>(o) = make_uint4(*reinterpret_cast<const uint32_t*>(&l0), *reinterpret_cast<const uint32_t*>(&l1), 0u, 0u);
            *reinterpret_cast<uint4*>(o + 8) = make_uint4(0u, 0u, 0u, 0u);
        }
    }
}
}  // namespace

extern "C" int ccn_subset_gather(const float* pos, int64_t ldp, const int64_t* batch, const int64_t* curve,
                                 const int64_t* idx, int64_t n, float* posq, int64_t* batchq, int64_t* curveq,
                                 void* pos_planes, int64_t plane_stride, int32_t n_planes, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && ldp >= 3);
    if (n == 0) return CCN_OK;
    CCN_REQUIRE(pos && batch && idx && posq && batchq && (!curve || curveq));
    CCN_REQUIRE(!pos_planes || (ccn_al16(pos_planes) && (n_planes == 1 || (n_planes == 2 && (plane_stride & 7) == 0))));
    ccn_launch(subset_gather_kernel, dim3(ccn_blocks_exact(n, 256)), dim3(256), 0, (cudaStream_t)stream, 
        pos, ldp, batch, curve, idx, n, posq, batchq, curveq, (__nv_bfloat16*)pos_planes, plane_stride, n_planes);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_dense_to_csr(const int32_t* idx, const int32_t* row_ptr, int64_t nq, int32_t K, int32_t* erow,
                                int32_t* ecol, ccn_stream_t stream) {
    CCN_REQUIRE(nq >= 0 && K >= 1 && idx && row_ptr && erow && ecol);
    if (nq == 0) return CCN_OK;
    ccn_launch(dense_to_csr_kernel, dim3(ccn_blocks_exact(nq * K, 256)), dim3(256), 0, (cudaStream_t)stream, idx, row_ptr, nq, K, erow, ecol);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_csr_rows(const int32_t* row_ptr, int64_t nq, int32_t* erow, ccn_stream_t stream) {
    CCN_REQUIRE(nq >= 0 && row_ptr && erow);
    if (nq == 0) return CCN_OK;
    ccn_launch(csr_rows_kernel, dim3(ccn_blocks_exact(nq, 256)), dim3(256), 0, (cudaStream_t)stream, row_ptr, nq, erow);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_edge_layer1(const float* H, int64_t ldh, const float* wp, const float* bias, const float* pos_src,
                               const float* pos_dst, const int32_t* erow, const int32_t* ecol, int64_t E, int32_t c,
                               float normalize_radius, int32_t act, float slope, float* out, int64_t ldo,
                               ccn_stream_t stream) {
    CCN_REQUIRE(E >= 0 && c > 0 && wp && pos_src && pos_dst && erow && ecol && out && ldo >= c);
    if (E == 0) return CCN_OK;
    if ((c & 3) == 0 && (ldo & 3) == 0 && ccn_al16(out) && ccn_al16(wp) && (!bias || ccn_al16(bias)) &&
        (!H || ((ldh & 3) == 0 && ccn_al16(H)))) {
        ccn_launch(edge_layer1_v4_kernel, dim3(CCN_ELT_GRID(E * (c >> 2))), dim3(256), 0, (cudaStream_t)stream, 
            H, ldh, wp, bias, pos_src, pos_dst, erow, ecol, E, c, normalize_radius > 0.f ? 1.f : 0.f, normalize_radius, act,
            slope, out, ldo);
        CCN_CHECK_LAUNCH();
        return CCN_OK;
    }
    ccn_launch(edge_layer1_kernel, dim3(CCN_ELT_GRID(E * c)), dim3(256), 0, (cudaStream_t)stream, 
        H, ldh, wp, bias, pos_src, pos_dst, erow, ecol, E, c, normalize_radius > 0.f ? 1.f : 0.f, normalize_radius, act,
        slope, out, ldo);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_edge_pair_layer1(const float* PQ, int64_t ldpq, const int32_t* erow, const int32_t* ecol, int64_t E,
                                    int32_t c, int32_t act, float slope, float* out, int64_t ldo, ccn_stream_t stream) {
    CCN_REQUIRE(E >= 0 && c > 0 && PQ && erow && ecol && out && ldpq >= 2 * c && ldo >= c);
    if (E == 0) return CCN_OK;
    if ((c & 3) == 0 && (ldpq & 3) == 0 && (ldo & 3) == 0 && ccn_al16(PQ) && ccn_al16(out)) {
        ccn_launch(edge_pair_layer1_v4_kernel, dim3(CCN_ELT_GRID(E * (c >> 2))), dim3(256), 0, (cudaStream_t)stream, PQ, ldpq, erow, ecol, E, c, act,
                                                                                                  slope, out, ldo);
        CCN_CHECK_LAUNCH();
        return CCN_OK;
    }
    ccn_launch(edge_pair_layer1_kernel, dim3(CCN_ELT_GRID(E * c)), dim3(256), 0, (cudaStream_t)stream, PQ, ldpq, erow, ecol, E, c, act,
                                                                                   slope, out, ldo);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_seg_aggregate(const float* msg, int64_t ldm, const float* att, int64_t lda, const int32_t* row_ptr,
                                 int64_t nq, int32_t c, int32_t mode, float* out, int64_t ldo, ccn_stream_t stream) {
    CCN_REQUIRE(nq >= 0 && c > 0 && row_ptr && out && mode >= 0 && mode <= 5);
    CCN_REQUIRE((mode != CCN_AGG_WSUM && mode != CCN_AGG_ATTEND && mode != CCN_AGG_WSUM_NORM) || att);
    if (nq == 0) return CCN_OK;
    if ((c & 3) == 0 && (ldm & 3) == 0 && (ldo & 3) == 0 && ccn_al16(msg) && ccn_al16(out) &&
        (!att || ((lda & 3) == 0 && ccn_al16(att)))) {
        ccn_launch(seg_aggregate_v4_kernel, dim3(CCN_ELT_GRID(nq * (c >> 2))), dim3(256), 0, (cudaStream_t)stream, msg, ldm, att, lda, row_ptr, nq,
                                                                                                c >> 2, mode, out, ldo, nullptr, 0, 0, 0);
        CCN_CHECK_LAUNCH();
        return CCN_OK;
    }
    ccn_launch(seg_aggregate_kernel, dim3(CCN_ELT_GRID(nq * c)), dim3(256), 0, (cudaStream_t)stream, msg, ldm, att, lda, row_ptr, nq, c,
                                                                                 mode, out, ldo);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_sgcnn_layer1(const float* PQ, int64_t ldpq, const int32_t* idx, int64_t n, int32_t K, int32_t c,
                                int32_t act, float slope, float* out, int64_t ldo, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && K >= 0 && c > 0 && PQ && out && ldpq >= 2 * c && ldo >= c && (K == 0 || idx));
    if (n == 0) return CCN_OK;
    ccn_launch(sgcnn_layer1_kernel, dim3(CCN_ELT_GRID(n * (K + 1) * c)), dim3(256), 0, (cudaStream_t)stream, PQ, ldpq, idx, n, K, c, act,
                                                                                        slope, out, ldo);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_group_max_dense(const float* feat, int64_t ldf, const int32_t* idx, int64_t n, int32_t K,
                                   int32_t c, float* out, int64_t ldo, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && K >= 0 && c > 0 && feat && out);
    if (n == 0) return CCN_OK;
    if ((c & 3) == 0 && (ldf & 3) == 0 && (ldo & 3) == 0 && ccn_al16(feat) && ccn_al16(out)) {
        ccn_launch(group_max_dense_v4_kernel, dim3(CCN_ELT_GRID(n * (c >> 2))), dim3(256), 0, (cudaStream_t)stream, feat, ldf, idx, n, K, c >> 2,
                                                                                                 out, ldo, nullptr, 0, 0, 0);
        CCN_CHECK_LAUNCH();
        return CCN_OK;
    }
    ccn_launch(group_max_dense_kernel, dim3(CCN_ELT_GRID(n * c)), dim3(256), 0, (cudaStream_t)stream, feat, ldf, idx, n, K, c, out, ldo);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_interp_blend(const float* x, int64_t ldx, const int32_t* nbr, const float* wgt, int32_t from_d2,
                                int64_t n, int32_t knn, int32_t c, float* out, int64_t ldo, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && knn >= 1 && c > 0 && x && nbr && wgt && out);
    if (n == 0) return CCN_OK;
    if ((c & 3) == 0 && (ldx & 3) == 0 && (ldo & 3) == 0 && ccn_al16(x) && ccn_al16(out)) {
        ccn_launch(interp_blend_v4_kernel, dim3(CCN_ELT_GRID(n * (c >> 2))), dim3(256), 0, (cudaStream_t)stream, x, ldx, nbr, wgt, from_d2, n, knn,
                                                                                              c >> 2, out, ldo, nullptr, 0, 0, 0);
        CCN_CHECK_LAUNCH();
        return CCN_OK;
    }
    ccn_launch(interp_blend_kernel, dim3(CCN_ELT_GRID(n * c)), dim3(256), 0, (cudaStream_t)stream, x, ldx, nbr, wgt, from_d2, n, knn, c,
                                                                               out, ldo);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

#define CCN_PLANES_OK(planes, kpad, pstride, npl, c) \
    ((planes) && ccn_al16(planes) && ((kpad) & 7) == 0 && (kpad) >= (c) && ((c) & 7) == 0 && ((npl) == 1 || ((npl) == 2 && ((pstride) & 7) == 0)))

// fp32 result + activation planes in one launch (float4 kernels; c % 16 == 0 so that the planes carry no pad columns)
#define CCN_BOTH_OK(planes, kpad, pstride, npl, c) \
    (CCN_PLANES_OK(planes, kpad, pstride, npl, c) && ((c) & 15) == 0 && (kpad) == (c))

extern "C" int ccn_seg_aggregate_both(const float* msg, int64_t ldm, const float* att, int64_t lda, const int32_t* row_ptr,
                                      int64_t nq, int32_t c, int32_t mode, float* out, int64_t ldo, void* planes,
                                      int64_t kpad, int64_t plane_stride, int32_t n_planes, ccn_stream_t stream) {
    CCN_REQUIRE(nq >= 0 && c > 0 && row_ptr && out && mode >= 0 && mode <= 5);
    CCN_REQUIRE((mode != CCN_AGG_WSUM && mode != CCN_AGG_ATTEND && mode != CCN_AGG_WSUM_NORM) || att);
    CCN_REQUIRE(CCN_BOTH_OK(planes, kpad, plane_stride, n_planes, c) && (ldm & 3) == 0 && (ldo & 3) == 0 && ccn_al16(msg) &&
                ccn_al16(out) && (!att || ((lda & 3) == 0 && ccn_al16(att))));
    if (nq == 0) return CCN_OK;
    ccn_launch(seg_aggregate_v4_kernel, dim3(CCN_ELT_GRID(nq * (c >> 2))), dim3(256), 0, (cudaStream_t)stream, 
        msg, ldm, att, lda, row_ptr, nq, c >> 2, mode, out, ldo, (__nv_bfloat16*)planes, kpad, plane_stride, n_planes);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_group_max_dense_both(const float* feat, int64_t ldf, const int32_t* idx, int64_t n, int32_t K, int32_t c,
                                        float* out, int64_t ldo, void* planes, int64_t kpad, int64_t plane_stride,
                                        int32_t n_planes, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && K >= 0 && c > 0 && feat && out && (K == 0 || idx));
    CCN_REQUIRE(CCN_BOTH_OK(planes, kpad, plane_stride, n_planes, c) && (ldf & 3) == 0 && (ldo & 3) == 0 && ccn_al16(feat) &&
                ccn_al16(out));
    if (n == 0) return CCN_OK;
    ccn_launch(group_max_dense_v4_kernel, dim3(CCN_ELT_GRID(n * (c >> 2))), dim3(256), 0, (cudaStream_t)stream, 
        feat, ldf, idx, n, K, c >> 2, out, ldo, (__nv_bfloat16*)planes, kpad, plane_stride, n_planes);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_interp_blend_both(const float* x, int64_t ldx, const int32_t* nbr, const float* wgt, int32_t from_d2,
                                     int64_t n, int32_t knn, int32_t c, float* out, int64_t ldo, void* planes, int64_t kpad,
                                     int64_t plane_stride, int32_t n_planes, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && knn >= 1 && c > 0 && x && nbr && wgt && out);
    CCN_REQUIRE(CCN_BOTH_OK(planes, kpad, plane_stride, n_planes, c) && (ldx & 3) == 0 && (ldo & 3) == 0 && ccn_al16(x) &&
                ccn_al16(out));
    if (n == 0) return CCN_OK;
    ccn_launch(interp_blend_v4_kernel, dim3(CCN_ELT_GRID(n * (c >> 2))), dim3(256), 0, (cudaStream_t)stream, 
        x, ldx, nbr, wgt, from_d2, n, knn, c >> 2, out, ldo, (__nv_bfloat16*)planes, kpad, plane_stride, n_planes);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_sgcnn_layer1_planes(const float* PQ, int64_t ldpq, const int32_t* idx, int64_t n, int32_t K, int32_t c,
                                       int32_t act, float slope, void* planes, int64_t kpad, int64_t plane_stride,
                                       int32_t n_planes, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && K >= 0 && c > 0 && PQ && ldpq >= 2 * c && (K == 0 || idx));
    CCN_REQUIRE(CCN_PLANES_OK(planes, kpad, plane_stride, n_planes, c) && kpad == c && ccn_al16(PQ) && (ldpq & 3) == 0);
    if (n == 0) return CCN_OK;
    ccn_launch(sgcnn_layer1_planes_kernel, dim3(CCN_ELT_GRID(n * (K + 1) * (c >> 3))), dim3(256), 0, (cudaStream_t)stream, 
        PQ, ldpq, idx, n, K, c, act, slope, (__nv_bfloat16*)planes, kpad, plane_stride, n_planes);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_edge_pair_layer1_planes(const float* PQ, int64_t ldpq, const int32_t* erow, const int32_t* ecol,
                                           int64_t E, int32_t c, int32_t act, float slope, void* planes, int64_t kpad,
                                           int64_t plane_stride, int32_t n_planes, ccn_stream_t stream) {
    CCN_REQUIRE(E >= 0 && c > 0 && PQ && erow && ecol && ldpq >= 2 * c);
    CCN_REQUIRE(CCN_PLANES_OK(planes, kpad, plane_stride, n_planes, c) && kpad == c && ccn_al16(PQ) && (ldpq & 3) == 0);
    if (E == 0) return CCN_OK;
    ccn_launch(edge_pair_layer1_planes_kernel, dim3(CCN_ELT_GRID(E * (c >> 3))), dim3(256), 0, (cudaStream_t)stream, 
        PQ, ldpq, erow, ecol, E, c, act, slope, (__nv_bfloat16*)planes, kpad, plane_stride, n_planes);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_edge_layer1_planes(const float* H, int64_t ldh, const float* wp, const float* bias,
                                      const float* pos_src, const float* pos_dst, const int32_t* erow,
                                      const int32_t* ecol, int64_t E, int32_t c, float normalize_radius, int32_t act,
                                      float slope, void* planes, int64_t kpad, int64_t plane_stride, int32_t n_planes,
                                      ccn_stream_t stream) {
    CCN_REQUIRE(E >= 0 && c > 0 && wp && pos_src && pos_dst && erow && ecol);
    CCN_REQUIRE(CCN_PLANES_OK(planes, kpad, plane_stride, n_planes, c) && kpad == c && ccn_al16(wp) &&
                (!bias || ccn_al16(bias)) && (!H || (ccn_al16(H) && (ldh & 3) == 0)));
    if (E == 0) return CCN_OK;
    ccn_launch(edge_layer1_planes_kernel, dim3(CCN_ELT_GRID(E * (c >> 3))), dim3(256), 0, (cudaStream_t)stream, 
        H, ldh, wp, bias, pos_src, pos_dst, erow, ecol, E, c, normalize_radius > 0.f ? 1.f : 0.f, normalize_radius, act,
        slope, (__nv_bfloat16*)planes, kpad, plane_stride, n_planes);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
