// Shared-MLP Linear layer on the 5th-gen tensor cores (tcgen05 + TMEM + TMA), sm_100a only.
//   Y[M,N] = act(A[M,K] W[N,K]^T + bias)       A, Y fp32 in HBM; W pre-split to bf16 by ccn_split_bf16
// (replaces torch_geometric.nn.MLP -> cuBLAS SGEMM + BatchNorm + act
//  kernels, reference src/models/base.py:32, src/models/modules/mlp.py:16-22.)
//
// Design (one CTA = one 128 x BN output tile, cta_group::1):
//   warps 0-3  A producers: coalesced fp32 loads of a 128 x 64 chunk, convert to bf16 (hi, and the
//              bf16 residual lo in x3 mode), st.shared in the SWIZZLE_128B K-major UMMA layout,
//              fence.proxy.async + mbarrier arrive.  The same warps run the epilogue: tcgen05.ld of
//              their 32 TMEM lanes, bias + activation, 128-byte row stores.
//   warp 4     one thread: TMA (cp.async.bulk.tensor.2d, SWIZZLE_128B) of the BN x 64 bf16 weight tile(s)
//   warp 5     one thread: tcgen05.mma kind::f16 (bf16 x bf16 -> fp32 in TMEM), tcgen05.commit to free stages
// Modes: 1 = plain bf16 (2e-2 budget of BASELINE.json), 3 = split-bf16 "x3" (A_hi W_hi + A_lo W_hi +
//   A_hi W_lo, ~fp32 accuracy for the 1e-4 budget) - three MMAs per K step into the same accumulator.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace {

using namespace ccn_tc;

constexpr int kBM = 128;
constexpr int kBK = 64;                  // bf16 elements per K chunk = 128 B = one swizzle row
constexpr int kABytes = kBM * kBK * 2;   // 16 KB
constexpr int kThreads = 192;

struct __align__(8) Bf16x4 { __nv_bfloat162 a, b; };

// MODE 1: stage = [A_hi | W_hi];  MODE 3: stage = [A_hi | A_lo | W_hi | W_lo]
template <int MODE>
__global__ void __launch_bounds__(kThreads) linear_tc_kernel(
    const float* __restrict__ A, int64_t lda, const __grid_constant__ CUtensorMap w_hi,
    const __grid_constant__ CUtensorMap w_lo, const float* __restrict__ bias, float* __restrict__ Y, int64_t ldy,
    int M, int N, int K, int BN, int stages, int act, float slope) {
    ccn_pdl_wait();
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * 8 + 1];     // full[8], empty[8], tmem_full
    __shared__ uint32_t tmem_base_slot;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * kBM, n0 = blockIdx.y * BN;
    const int nk = (K + kBK - 1) / kBK;
    const int wbytes = BN * kBK * 2;
    const int stage_bytes = (MODE == 3 ? 2 : 1) * (kABytes + wbytes);
    const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[8]), tfull = smem_u32(&bars[16]);
    const int mma_n = min(BN, ((N - n0) + 15) & ~15);
    uint32_t tmem_cols = 32;
    while ((int)tmem_cols < BN) tmem_cols <<= 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(full0 + 8 * s, 4 + 1);       // 4 producer warps + the TMA thread's expect_tx arrive
            mbar_init(empty0 + 8 * s, 1);          // one tcgen05.commit
        }
        mbar_init(tfull, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(&tmem_base_slot)), "r"(tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_slot;

    if (warp < 4) {
        // ------------------------------------------------------------------ A producers
        const int c4 = threadIdx.x & 15;            // float4 column group: columns 4*c4 .. 4*c4+3
        const int rsub = threadIdx.x >> 4;          // 0..7; rows rsub + 8*i
        const bool vec = ((lda & 3) == 0) && ((reinterpret_cast<uintptr_t>(A) & 15) == 0);
        for (int kc = 0; kc < nk; ++kc) {
            const int s = kc % stages;
            mbar_wait(empty0 + 8 * s, ((kc / stages) & 1) ^ 1);
            const uint32_t a_hi = smem0 + s * stage_bytes;
            const uint32_t a_lo = a_hi + kABytes;
            const int col = kc * kBK + 4 * c4;
            float4 v[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int r = m0 + rsub + 8 * i;
                float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < M) {
                    const float* p = A + (int64_t)r * lda + col;
                    if (vec && col + 3 < K) {
                        t = *reinterpret_cast<const float4*>(p);
                    } else {
                        if (col < K) t.x = p[0];
                        if (col + 1 < K) t.y = p[1];
                        if (col + 2 < K) t.z = p[2];
                        if (col + 3 < K) t.w = p[3];
                    }
                }
                v[i] = t;
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int r = rsub + 8 * i;
                // SWIZZLE_128B: 16-byte chunk index XOR (row mod 8); this thread's 8 bytes sit in chunk c4/2
                const uint32_t off = r * 128 + ((((c4 >> 1) ^ (r & 7)) << 4) | ((c4 & 1) << 3));
                Bf16x4 h;
                h.a = __floats2bfloat162_rn(v[i].x, v[i].y);
                h.b = __floats2bfloat162_rn(v[i].z, v[i].w);
                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(a_hi + off), "r"(*reinterpret_cast<uint32_t*>(&h.a)),
                             "r"(*reinterpret_cast<uint32_t*>(&h.b)) : "memory");
                if (MODE == 3) {
                    const float2 fa = __bfloat1622float2(h.a), fb = __bfloat1622float2(h.b);
                    Bf16x4 l;
                    l.a = __floats2bfloat162_rn(v[i].x - fa.x, v[i].y - fa.y);
                    l.b = __floats2bfloat162_rn(v[i].z - fb.x, v[i].w - fb.y);
                    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(a_lo + off),
                                 "r"(*reinterpret_cast<uint32_t*>(&l.a)), "r"(*reinterpret_cast<uint32_t*>(&l.b)) : "memory");
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> async proxy (UMMA)
            __syncwarp();
            if (lane == 0) mbar_arrive(full0 + 8 * s);
        }
        // ------------------------------------------------------------------ epilogue
        mbar_wait(tfull, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int row = m0 + warp * 32 + lane;
        const bool vecy = ((ldy & 3) == 0) && ((reinterpret_cast<uintptr_t>(Y) & 15) == 0);
        for (int c0 = 0; c0 < mma_n; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
            if (row < M) {
                float* yrow = Y + (int64_t)row * ldy + n0 + c0;
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    float o[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int n = n0 + c0 + j + e;
                        float y = __uint_as_float(v[j + e]) + ((bias && n < N) ? bias[n] : 0.f);
                        if (act == CCN_ACT_RELU) y = fmaxf(y, 0.f);
                        else if (act == CCN_ACT_LEAKY) y = y > 0.f ? y : y * slope;
                        o[e] = y;
                    }
                    const int n = n0 + c0 + j;
                    if (vecy && n + 3 < N) {
                        *reinterpret_cast<float4*>(yrow + j) = make_float4(o[0], o[1], o[2], o[3]);
                    } else {
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            if (n + e < N) yrow[j + e] = o[e];
                    }
                }
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    } else if (warp == 4) {
        // ------------------------------------------------------------------ TMA weight producer
        if (lane == 0) {
            for (int kc = 0; kc < nk; ++kc) {
                const int s = kc % stages;
                mbar_wait(empty0 + 8 * s, ((kc / stages) & 1) ^ 1);
                const uint32_t w_hi_s = smem0 + s * stage_bytes + (MODE == 3 ? 2 : 1) * kABytes;
                mbar_arrive_expect_tx(full0 + 8 * s, (MODE == 3 ? 2 : 1) * wbytes);
                tma_load_2d(w_hi_s, &w_hi, kc * kBK, n0, full0 + 8 * s);
                if (MODE == 3) tma_load_2d(w_hi_s + wbytes, &w_lo, kc * kBK, n0, full0 + 8 * s);
            }
        }
    } else {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            // instruction descriptor: D=f32, A=B=bf16, K-major both, N>>3 at bit 17, M>>4 at bit 24
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(mma_n >> 3) << 17) |
                                   ((uint32_t)(kBM >> 4) << 24);
            for (int kc = 0; kc < nk; ++kc) {
                const int s = kc % stages;
                mbar_wait(full0 + 8 * s, (kc / stages) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t a_hi = smem0 + s * stage_bytes;
                const uint32_t a_lo = a_hi + kABytes;
                const uint32_t w_hi_s = a_hi + (MODE == 3 ? 2 : 1) * kABytes;
                const uint32_t w_lo_s = w_hi_s + wbytes;
#pragma unroll
                for (int k = 0; k < kBK / 16; ++k) {
                    const uint32_t ko = k * 32;     // 16 bf16 = 32 bytes inside the 128-byte swizzle row
                    umma_bf16(tmem_base, umma_desc_sw128(a_hi + ko), umma_desc_sw128(w_hi_s + ko), idesc,
                              (kc | k) ? 1u : 0u);
                    if (MODE == 3) {
                        umma_bf16(tmem_base, umma_desc_sw128(a_lo + ko), umma_desc_sw128(w_hi_s + ko), idesc, 1u);
                        umma_bf16(tmem_base, umma_desc_sw128(a_hi + ko), umma_desc_sw128(w_lo_s + ko), idesc, 1u);
                    }
                }
                umma_commit(empty0 + 8 * s);        // frees the stage once these MMAs have read it
            }
            umma_commit(tfull);                     // accumulator complete
        }
    }
    __syncthreads();
    if (warp == 4) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
    }
}

// fp32 (N, K) -> bf16 hi (N, Kpad) and residual lo = bf16(w - hi), zero padded along K.
__global__ void split_bf16_kernel(const float* __restrict__ w, int64_t ldw, int n, int k, int kpad,
                                  __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo) {
    ccn_pdl_wait();
    const int64_t total = (int64_t)n * kpad;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = t / kpad;
        const int c = (int)(t - r * kpad);
        const float v = c < k ? w[r * ldw + c] : 0.f;
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        hi[t] = h;
        if (lo) lo[t] = __float2bfloat16_rn(v - __bfloat162float(h));
    }
}

}  // namespace

extern "C" int ccn_split_bf16(const float* w, int64_t ldw, int32_t n, int32_t k, int32_t kpad, void* hi, void* lo,
                              ccn_stream_t stream) {
    CCN_REQUIRE(w && hi && n > 0 && k > 0 && kpad >= k && (kpad % kBK) == 0 && ldw >= k);
    ccn_launch(split_bf16_kernel, dim3(ccn_grid((int64_t)n * kpad, 256)), dim3(256), 0, (cudaStream_t)stream, 
        w, ldw, n, k, kpad, (__nv_bfloat16*)hi, (__nv_bfloat16*)lo);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

// Host-side TMA descriptor of a bf16 (n, kpad) row-major matrix: box = 64 (K) x box_n rows, SWIZZLE_128B.
// libcuda is resolved at run time through the runtime's driver entry point (the library must load, and export its
// symbols, on hosts without a driver; nothing here links -lcuda).
int ccn_encode_bf16_map(void* tmap_out_host, const void* base, int64_t n, int64_t kpad, int32_t box_n) {
    typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeTiled encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
        if (e != cudaSuccess) return (int)e;
        if (q != cudaDriverEntryPointSuccess || !fn) return CCN_EUNSUPPORTED;
        encode = (EncodeTiled)fn;
    }
    cuuint64_t dims[2] = {(cuuint64_t)kpad, (cuuint64_t)n};
    cuuint64_t strides[1] = {(cuuint64_t)kpad * 2};
    cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)box_n};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = encode((CUtensorMap*)tmap_out_host, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims,
                        strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? CCN_OK : 1000 + (int)r;
}

extern "C" int ccn_weight_tensor_map(void* tmap_out_host, const void* w_bf16, int32_t n, int32_t kpad, int32_t box_n) {
    CCN_REQUIRE(tmap_out_host && w_bf16 && n > 0 && kpad > 0 && (kpad % 8) == 0 && box_n >= 16 && box_n <= 256);
    return ccn_encode_bf16_map(tmap_out_host, w_bf16, n, kpad, box_n);
}

extern "C" int ccn_linear_tc(const float* A, int64_t lda, const void* w_hi_tmap_host, const void* w_lo_tmap_host,
                             const float* bias, float* Y, int64_t ldy, int64_t M, int32_t N, int32_t K, int32_t box_n,
                             int32_t mode, int32_t act, float slope, ccn_stream_t stream) {
    CCN_REQUIRE(A && w_hi_tmap_host && Y && M >= 0 && M < 0x7fffffff && N > 0 && K > 0 && lda >= K && ldy >= N);
    CCN_REQUIRE((mode == 1 || (mode == 3 && w_lo_tmap_host)) && act >= 0 && act <= 2);
    CCN_REQUIRE(box_n == 32 || box_n == 64 || box_n == 128 || box_n == 256);
    if (M == 0) return CCN_OK;
    const int per_stage = (mode == 3 ? 2 : 1) * (kABytes + box_n * kBK * 2);
    // large grids: keep two CTAs per SM resident (epilogue of one overlaps the main loop of the other);
    // skinny problems (at most one CTA per SM anyway): spend the shared memory on a deeper pipeline
    const int64_t tiles = ((M + kBM - 1) / kBM) * ((N + box_n - 1) / box_n);
    const int budget = (tiles > 148 ? 100 : 200) * 1024;
    int stages = (budget - 1024) / per_stage;
    if (stages > 8) stages = 8;
    if (stages < 2) stages = 2;
    const int nk = (K + kBK - 1) / kBK;
    if (stages > nk) stages = nk < 2 ? 2 : nk;
    const int smem = stages * per_stage + 1024;
    const CUtensorMap* hi = (const CUtensorMap*)w_hi_tmap_host;
    const CUtensorMap* lo = mode == 3 ? (const CUtensorMap*)w_lo_tmap_host : hi;
    dim3 grid((unsigned)((M + kBM - 1) / kBM), (unsigned)((N + box_n - 1) / box_n));
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e;
    if (mode == 3) {
        static bool attr3 = false;
        if (!attr3) {
            e = cudaFuncSetAttribute(linear_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
            if (e != cudaSuccess) return (int)e;
            attr3 = true;
        }
        ccn_launch(linear_tc_kernel<3>, dim3(grid), dim3(kThreads), smem, s, A, lda, *hi, *lo, bias, Y, ldy, (int)M, N, K, box_n, stages, act, slope);
    } else {
        static bool attr1 = false;
        if (!attr1) {
            e = cudaFuncSetAttribute(linear_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
            if (e != cudaSuccess) return (int)e;
            attr1 = true;
        }
        ccn_launch(linear_tc_kernel<1>, dim3(grid), dim3(kThreads), smem, s, A, lda, *hi, *lo, bias, Y, ldy, (int)M, N, K, box_n, stages, act, slope);
    }
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
