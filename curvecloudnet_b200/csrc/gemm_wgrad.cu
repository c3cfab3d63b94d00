// Linear / curve-conv WEIGHT GRADIENT on tcgen05 (sm_100a only):   dW[N,K] += sum_m dY[m,N]^T X[m + x_row_off, K]
// (SURVEY.md §8(b) mlp_bwd / curve_conv1d_sym_bwd_weight; the reference gets it from cuBLAS / cuDNN through autograd.)
//
// The reduction runs over the ROW axis of both operands, so both are "MN-major" for the tensor core: a TMA box of
// 64 columns x 64 rows of a bf16 activation plane (SWIZZLE_128B) is exactly one canonical MN-major UMMA block
// (64 MN elements = one 128-byte line, 8 K-rows per 1024-byte swizzle atom, SBO = 1024, LBO = distance between the
// 64-column blocks).  No transposed copy of dY or X is ever made.
//   MMA M (128)      <- 128 columns of dY  = rows of dW
//   MMA N (BKc<=256) <- BKc columns of X   = columns of dW
//   MMA K (16 x 4)   <- 64 rows of dY / X per pipeline stage
// Grid: (dW tiles) x (row splits); every CTA accumulates its row range in TMEM (fp32) and adds its tile into dW with
// vector reductions (red.global.add.v4.f32).  x3 mode: dY_hi X_hi + dY_lo X_hi + dY_hi X_lo (split-bf16, fp32 class).
// 192 threads: warp 0 = TMA producer, warp 1 = MMA issuer, warps 2-5 = epilogue.
#include "common.cuh"
#include "tc_common.cuh"

namespace {

using namespace ccn_tc;

constexpr int kRows = 64;                  // reduction rows per stage
constexpr int kBoxBytes = 64 * kRows * 2;  // one 64-column x 64-row bf16 box
constexpr int kWM = 128;                   // dW rows per tile
constexpr int kThreads = 192;
constexpr int kMaxStages = 6;
constexpr int kSmemLimit = 227 * 1024;

struct WgradParams {
    float* dW;
    int64_t lddw;
    int M, N, K;
    int BKc;              // dW columns per tile (multiple of 64, <= 256)
    int tiles_k, tiles;   // tiles = tiles_n * tiles_k
    int chunks, cps;      // 64-row chunks in total / per split
    int stages;
    int x_row_off;        // X row read for dY row m is m + x_row_off (curve-conv taps); out-of-range rows read as zero
};

// MN-major, SWIZZLE_128B shared-memory matrix descriptor
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t saddr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)(lbo_bytes >> 4) << 16;         // leading byte offset: next 64-element block along M/N
    d |= (uint64_t)(1024 >> 4) << 32;              // stride byte offset: next group of 8 K-rows
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;                        // SWIZZLE_128B
    return d;
}

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(kThreads, 1) wgrad_tc_kernel(
    const __grid_constant__ CUtensorMap dy_hi, const __grid_constant__ CUtensorMap dy_lo,
    const __grid_constant__ CUtensorMap x_hi, const __grid_constant__ CUtensorMap x_lo, const WgradParams p) {
    ccn_pdl_wait();
    constexpr int P = MODE == 3 ? 2 : 1;
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * kMaxStages + 1];     // full[S], empty[S], tfull
    __shared__ uint32_t tmem_base_slot;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x % p.tiles, split = blockIdx.x / p.tiles;
    const int n0 = (tile / p.tiles_k) * kWM, k0 = (tile % p.tiles_k) * p.BKc;
    const int c_beg = split * p.cps, c_end = min(p.chunks, c_beg + p.cps);
    const int nb = p.BKc >> 6;                                     // 64-column boxes of X per plane
    const int a_plane = 2 * kBoxBytes, b_plane = nb * kBoxBytes;
    const int stage_bytes = P * (a_plane + b_plane);
    const uint32_t ring0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[kMaxStages]);
    const uint32_t tfull = smem_u32(&bars[2 * kMaxStages]);
    uint32_t tmem_cols = 32;
    while ((int)tmem_cols < p.BKc) tmem_cols <<= 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 1);
        }
        mbar_init(tfull, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(smem_u32(&tmem_base_slot), tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;

    if (c_end > c_beg) {
        if (warp == 0) {
            // converged warp, one elected lane issues; ring slot / phase carried (see gemm_planes.cu)
            {
                int s = 0;
                uint32_t ph = 1;
                for (int c = c_beg; c < c_end; ++c) {
                    mbar_wait(empty0 + 8 * s, ph);
                    const uint32_t dst = ring0 + s * stage_bytes;
                    const uint32_t fb = full0 + 8 * s;
                    const int m0 = c * kRows;
                    if (elect_one()) {
                        mbar_arrive_expect_tx(fb, (uint32_t)stage_bytes);
#pragma unroll
                        for (int pl = 0; pl < P; ++pl) {
                            const CUtensorMap* my = pl ? &dy_lo : &dy_hi;
                            const CUtensorMap* mx = pl ? &x_lo : &x_hi;
                            tma_load_2d(dst + pl * a_plane, my, n0, m0, fb);
                            tma_load_2d(dst + pl * a_plane + kBoxBytes, my, n0 + 64, m0, fb);
                            const uint32_t bdst = dst + P * a_plane + pl * b_plane;
                            for (int j = 0; j < nb; ++j) tma_load_2d(bdst + j * kBoxBytes, mx, k0 + 64 * j, m0 + p.x_row_off, fb);
                        }
                    }
                    __syncwarp();
                    if (++s == p.stages) { s = 0; ph ^= 1u; }
                }
            }
        } else if (warp == 1) {
            {
                // D = f32, A = B = bf16, both MN-major (bits 15, 16), N >> 3 at bit 17, M >> 4 at bit 24
                const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) |
                                       ((uint32_t)(p.BKc >> 3) << 17) | ((uint32_t)(kWM >> 4) << 24);
                int s = 0;
                uint32_t ph = 0;
                bool first = true;
                for (int c = c_beg; c < c_end; ++c) {
                    mbar_wait(full0 + 8 * s, ph);
                    tc_fence_after();
                    const uint32_t a_hi = ring0 + s * stage_bytes, a_lo = a_hi + a_plane;
                    const uint32_t b_hi = a_hi + P * a_plane, b_lo = b_hi + b_plane;
                    if (elect_one()) {
#pragma unroll
                        for (int ks = 0; ks < kRows / 16; ++ks) {
                            const uint32_t ko = ks * 16 * 128;             // 16 K-rows of 128 bytes
                            const uint64_t da = umma_desc_mn_sw128(a_hi + ko, kBoxBytes);
                            const uint64_t db = umma_desc_mn_sw128(b_hi + ko, kBoxBytes);
                            umma_bf16(tmem_base, da, db, idesc, (first && ks == 0) ? 0u : 1u);
                            if (P == 2) {
                                umma_bf16(tmem_base, umma_desc_mn_sw128(a_lo + ko, kBoxBytes), db, idesc, 1u);
                                umma_bf16(tmem_base, da, umma_desc_mn_sw128(b_lo + ko, kBoxBytes), idesc, 1u);
                            }
                        }
                        umma_commit(empty0 + 8 * s);
                    }
                    __syncwarp();
                    first = false;
                    if (++s == p.stages) { s = 0; ph ^= 1u; }
                }
                if (elect_one()) umma_commit(tfull);
                __syncwarp();
            }
        } else {
            const int q = warp & 3;
            mbar_wait(tfull, 0);
            tc_fence_after();
            const int row = n0 + q * 32 + lane;
            const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16);
            const bool vec = ((p.lddw & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.dW) & 15) == 0);
            for (int c0 = 0; c0 < p.BKc; c0 += 32) {
                if (k0 + c0 >= p.K) break;                                  // warp-uniform
                uint32_t v[32];
                tmem_ld32(tbase + (uint32_t)c0, v);
                if (row < p.N) {
                    float* o = p.dW + (int64_t)row * p.lddw + k0 + c0;
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        const int k = k0 + c0 + j;
                        if (vec && k + 3 < p.K) {
                            red_add_v4(o + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                       __uint_as_float(v[j + 3]));
                        } else {
#pragma unroll
                            for (int e = 0; e < 4; ++e)
                                if (k + e < p.K) atomicAdd(o + j + e, __uint_as_float(v[j + e]));
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, tmem_cols);
    }
}

// out[c] += sum_r src[r][c]   (bias gradient)
__global__ void __launch_bounds__(256) col_sum_kernel(const float* __restrict__ src, int64_t ld, int64_t M, int C,
                                                      int64_t rows_per_cta, float* __restrict__ out) {
    ccn_pdl_wait();
    const int cw = C >= 256 ? 256 : C;
    const int lanes = 256 / cw;
    const int col = blockIdx.y * 256 + (threadIdx.x % cw);
    const int lane = threadIdx.x / cw;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
    if (col >= C || lane >= lanes) return;
    float s = 0.f;
    for (int64_t r = r0 + lane; r < r1; r += lanes) s += src[r * ld + col];
    atomicAdd(out + col, s);
}

}  // namespace

int ccn_col_sum_v4(const float* src, int64_t ld, int64_t M, int32_t C, float* out, cudaStream_t s);   // train_ops.cu

extern "C" int ccn_col_sum(const float* src, int64_t ld, int64_t M, int32_t C, float* out, ccn_stream_t stream) {
    CCN_REQUIRE(M >= 0 && C > 0 && src && out && ld >= C);
    if (M == 0) return CCN_OK;
    if ((C & 3) == 0 && (ld & 3) == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0)
        return ccn_col_sum_v4(src, ld, M, C, out, (cudaStream_t)stream);
    const int gy = (C + 255) / 256;
    int64_t gx = (148 * 8) / gy;
    const int64_t max_gx = (M + 255) / 256;
    if (gx > max_gx) gx = max_gx;
    if (gx < 1) gx = 1;
    const int64_t rpc = (M + gx - 1) / gx;
    gx = (M + rpc - 1) / rpc;
    ccn_launch(col_sum_kernel, dim3(dim3((unsigned)gx, (unsigned)gy)), dim3(256), 0, (cudaStream_t)stream, src, ld, M, C, rpc, out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_linear_wgrad_tc(const void* dy_planes, int64_t dy_ld, int64_t dy_plane_stride, const void* x_planes,
                                   int64_t x_ld, int64_t x_plane_stride, int64_t M, int32_t N, int32_t K,
                                   int32_t x_row_off, int64_t x_rows, int32_t mode, float* dW, int64_t lddw,
                                   ccn_stream_t stream) {
    CCN_REQUIRE(dy_planes && x_planes && dW && M >= 0 && M < 0x7fffffff && N > 0 && K > 0 && lddw >= K);
    CCN_REQUIRE(mode == 1 || mode == 3);
    CCN_REQUIRE(dy_ld >= N && x_ld >= K && (dy_ld & 7) == 0 && (x_ld & 7) == 0 && x_rows > 0);
    CCN_REQUIRE((reinterpret_cast<uintptr_t>(dy_planes) & 15) == 0 && (reinterpret_cast<uintptr_t>(x_planes) & 15) == 0);
    CCN_REQUIRE(mode == 1 || ((dy_plane_stride & 7) == 0 && (x_plane_stride & 7) == 0));
    if (M == 0) return CCN_OK;
    const int P = mode == 3 ? 2 : 1;
    WgradParams p;
    p.dW = dW; p.lddw = lddw; p.M = (int)M; p.N = N; p.K = K; p.x_row_off = x_row_off;
    int bkc = (K + 63) / 64 * 64;
    if (bkc > 256) bkc = 256;
    p.BKc = bkc;
    p.tiles_k = (K + bkc - 1) / bkc;
    const int tiles_n = (N + kWM - 1) / kWM;
    p.tiles = tiles_n * p.tiles_k;
    p.chunks = (int)((M + kRows - 1) / kRows);
    int splits = 148 / p.tiles;
    if (splits < 1) splits = 1;
    if (splits > p.chunks) splits = p.chunks;
    p.cps = (p.chunks + splits - 1) / splits;
    splits = (p.chunks + p.cps - 1) / p.cps;
    const int stage_bytes = P * (2 * kBoxBytes + (bkc / 64) * kBoxBytes);
    int stages = (kSmemLimit - 2560) / stage_bytes;
    if (stages > kMaxStages) stages = kMaxStages;
    if (stages < 2) return CCN_EUNSUPPORTED;
    p.stages = stages;
    const int smem = stages * stage_bytes + 1024;
    alignas(64) CUtensorMap maps[4];
    const __nv_bfloat16* dyb = (const __nv_bfloat16*)dy_planes;
    const __nv_bfloat16* xb = (const __nv_bfloat16*)x_planes;
    int rc = ccn_encode_bf16_map(&maps[0], dyb, M, dy_ld, kRows);
    if (rc == CCN_OK) rc = ccn_encode_bf16_map(&maps[1], mode == 3 ? dyb + dy_plane_stride : dyb, M, dy_ld, kRows);
    if (rc == CCN_OK) rc = ccn_encode_bf16_map(&maps[2], xb, x_rows, x_ld, kRows);
    if (rc == CCN_OK) rc = ccn_encode_bf16_map(&maps[3], mode == 3 ? xb + x_plane_stride : xb, x_rows, x_ld, kRows);
    if (rc != CCN_OK) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e;
    const unsigned grid = (unsigned)(p.tiles * splits);
    if (mode == 3) {
        static bool attr3 = false;
        if (!attr3) {
            e = cudaFuncSetAttribute(wgrad_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit - 1536);
            if (e != cudaSuccess) return (int)e;
            attr3 = true;
        }
        ccn_launch(wgrad_tc_kernel<3>, dim3(grid), dim3(kThreads), smem, s, maps[0], maps[1], maps[2], maps[3], p);
    } else {
        static bool attr1 = false;
        if (!attr1) {
            e = cudaFuncSetAttribute(wgrad_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit - 1536);
            if (e != cudaSuccess) return (int)e;
            attr1 = true;
        }
        ccn_launch(wgrad_tc_kernel<1>, dim3(grid), dim3(kThreads), smem, s, maps[0], maps[1], maps[2], maps[3], p);
    }
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
