// fp32 Linear layer  Y = act(A W^T + b)  on the FP32 pipe (the 1e-4 parity path of the shared MLPs,
// SURVEY.md §8(a) a21; reference torch_geometric.nn.MLP -> nn.Linear + BatchNorm1d + activation,
// call sites src/models/base.py:32, src/models/modules/mlp.py:16-22).  BatchNorm (eval) is folded
// into W/b by the host, the activation runs in the epilogue.  The bf16 tensor-core path
// (gemm_tc.cu) is the throughput path; this kernel is what it is checked against.
#include "common.cuh"

namespace {

constexpr int kBK = 16;

template <int TM, int TN>
__global__ void __launch_bounds__(256) linear_f32_kernel(const float* __restrict__ A, int64_t lda,
                                                         const float* __restrict__ W, int64_t ldw,
                                                         const float* __restrict__ bias, float* __restrict__ Y,
                                                         int64_t ldy, int M, int N, int K, int act, float slope) {
    ccn_pdl_wait();
    constexpr int BM = 16 * TM, BN = 16 * TN;
    constexpr int PA = BM / 16, PB = BN / 16;     // rows loaded per thread per k-tile
    __shared__ __align__(16) float As[2][kBK][BM + 4];
    __shared__ __align__(16) float Bs[2][kBK][BN + 4];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int lk = tid & 15, lr = tid >> 4;       // loader: k within tile, first row
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    float ra[PA], rb[PB];
    auto load_global = [&](int k0) {
        const int k = k0 + lk;
#pragma unroll
        for (int p = 0; p < PA; ++p) {
            const int m = m0 + lr + 16 * p;
            ra[p] = (m < M && k < K) ? A[(int64_t)m * lda + k] : 0.f;
        }
#pragma unroll
        for (int p = 0; p < PB; ++p) {
            const int n = n0 + lr + 16 * p;
            rb[p] = (n < N && k < K) ? W[(int64_t)n * ldw + k] : 0.f;
        }
    };
    auto store_smem = [&](int buf) {
#pragma unroll
        for (int p = 0; p < PA; ++p) As[buf][lk][lr + 16 * p] = ra[p];
#pragma unroll
        for (int p = 0; p < PB; ++p) Bs[buf][lk][lr + 16 * p] = rb[p];
    };

    const int nk = (K + kBK - 1) / kBK;
    load_global(0);
    store_smem(0);
    __syncthreads();
    for (int t = 0; t < nk; ++t) {
        const int buf = t & 1;
        if (t + 1 < nk) load_global((t + 1) * kBK);
#pragma unroll
        for (int k = 0; k < kBK; ++k) {
            float a[TM], b[TN];
#pragma unroll
            for (int i = 0; i < TM; i += 4) *(float4*)&a[i] = *(const float4*)&As[buf][k][ty * TM + i];
#pragma unroll
            for (int j = 0; j < TN; j += 4) *(float4*)&b[j] = *(const float4*)&Bs[buf][k][tx * TN + j];
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (t + 1 < nk) {
            store_smem(buf ^ 1);
            __syncthreads();
        }
    }

    const bool vec = (ldy % 4 == 0) && ((reinterpret_cast<uintptr_t>(Y) & 15) == 0);
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int m = m0 + ty * TM + i;
        if (m >= M) break;
#pragma unroll
        for (int j = 0; j < TN; j += 4) {
            const int n = n0 + tx * TN + j;
            float v[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                float y = acc[i][j + e] + ((bias && n + e < N) ? bias[n + e] : 0.f);
                if (act == CCN_ACT_RELU) y = fmaxf(y, 0.f);
                else if (act == CCN_ACT_LEAKY) y = y > 0.f ? y : y * slope;
                v[e] = y;
            }
            if (vec && n + 3 < N) {
                *(float4*)&Y[(int64_t)m * ldy + n] = make_float4(v[0], v[1], v[2], v[3]);
            } else {
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if (n + e < N) Y[(int64_t)m * ldy + n + e] = v[e];
            }
        }
    }
}

}  // namespace

extern "C" int ccn_linear_f32(const float* A, int64_t lda, const float* W, int64_t ldw, const float* bias, float* Y,
                              int64_t ldy, int64_t M, int32_t N, int32_t K, int32_t act, float slope,
                              ccn_stream_t stream) {
    CCN_REQUIRE(M >= 0 && M < 0x7fffffff && N > 0 && K > 0 && A && W && Y && lda >= K && ldw >= K && ldy >= N);
    CCN_REQUIRE(act >= 0 && act <= 2);
    if (M == 0) return CCN_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t big_tiles = ((M + 127) / 128) * ((N + 127) / 128);
    if (big_tiles >= 148) {
        dim3 grid((N + 127) / 128, (unsigned)((M + 127) / 128));
        ccn_launch(linear_f32_kernel<8, 8>, dim3(grid), dim3(256), 0, s, A, lda, W, ldw, bias, Y, ldy, (int)M, N, K, act, slope);
    } else {
        dim3 grid((N + 63) / 64, (unsigned)((M + 63) / 64));
        ccn_launch(linear_f32_kernel<4, 4>, dim3(grid), dim3(256), 0, s, A, lda, W, ldw, bias, Y, ldy, (int)M, N, K, act, slope);
    }
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
