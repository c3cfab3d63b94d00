// Shared-MLP Linear layer on tcgen05 with BOTH operands delivered by TMA (sm_100a only).
//   Y[M,N] = act(A[M,K] W[N,K]^T + bias)
// (replaces torch_geometric.nn.MLP -> cuBLAS SGEMM + BatchNorm + activation
//  kernels, reference src/models/base.py:32, src/models/modules/mlp.py:16-22.)
//
// Activation format between layers: "split-bf16 planes" - an (M, K) activation is stored as bf16 plane(s)
// (planes, M, kpad), plane 0 = bf16(x), plane 1 = bf16(x - plane 0) (x3 mode only), kpad % 8 == 0 and pad
// columns zero.  Both planes together cost the 4 bytes/element of fp32, but they are TMA-loadable straight
// into the SWIZZLE_128B K-major UMMA layout: no conversion warps, no register staging, the load pipeline is
// as deep as shared memory allows.  Producers (this kernel's epilogue, ccn_pack_planes, the edge-layer
// kernels) write planes; the last layer of a chain writes fp32.
//
// One persistent CTA per SM, 320 threads:
//   warp 0   lane 0: TMA producer.  Weights (bf16 hi [+ lo]) are loaded ONCE per CTA when the CTA's N-slice of W
//            fits in shared memory next to a >= 3-stage ring ("resident": every big-M layer of the path), else
//            they stream through the ring with A.  A tiles: 128 rows x 64 bf16 per plane per stage.
//   warp 1   lane 0: tcgen05.mma kind::f16 issuer, fp32 accumulators in TMEM, DOUBLE-BUFFERED (2 x BN columns)
//            so the epilogue of tile t overlaps the MMAs of tile t+1.  x3 mode: A_hi W_hi + A_lo W_hi + A_hi W_lo.
//   warps 2-9 epilogue: tcgen05.ld (32 lanes x 32 columns), bias + activation, then fp32 rows and/or bf16 planes;
//            two warps per TMEM lane quarter, alternating 32-column chunks.
// Tile order: tile t -> (m = t / tiles_n, n = t % tiles_n); the grid is a multiple of tiles_n in resident mode, so
// a CTA keeps one N-slice and the CTAs sharing an A tile run concurrently (second read served by L2).
#include <cstdlib>
#include <type_traits>

#include "common.cuh"
#include "tc_common.cuh"

namespace {

using namespace ccn_tc;

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kATile = kBM * kBK * 2;     // 16 KB per plane per stage
constexpr int kThreads = 320;          // warp 0: TMA producer, warp 1: MMA issuer, warps 2..9: epilogue
constexpr int kEpiThreads = kThreads - 64;
constexpr int kMaxStages = 8;
constexpr int kSmemLimit = 227 * 1024;

struct PlanesParams {
    const float* bias;
    float* Y;               // fp32 output or null
    int64_t ldy;
    __nv_bfloat16* Yp;      // planes output or null
    int64_t yp_ld;          // row stride (elements) of an output plane
    int64_t yp_plane;       // plane stride (elements)
    int yp_cols;            // columns to write per output plane row (N rounded up to 16)
    int M, N, K;
    int BN, tiles_m, tiles_n;
    int stages, resident;
    int act;
    float slope;
    int nseg;               // A = [A_0 | A_1 | ...] along K: every segment is its own planes tensor
    int sc1, sc2, sc3, nk;  // first 64-wide K chunk of segments 1..3 (= nk for absent segments), total chunks
    int taps, cpt;          // taps > 0: curve convolution - K = taps x (cpt chunks of segment 0), tap t reads rows m + t - taps/2
    int b_mn;               // 1: the weight is given as (K, N) row-major (Y = A W, the data gradient with the FORWARD weight):
                            //    B tiles are MN-major UMMA operands, one 64 x 64 TMA box per 64 output columns
};

constexpr int kBoxBytes = 64 * 64 * 2;    // one 64 (N) x 64 (K) bf16 box = one MN-major SWIZZLE_128B block

// MN-major, SWIZZLE_128B shared-memory matrix descriptor (see gemm_wgrad.cu): LBO = distance between 64-column blocks,
// SBO = 1024 = 8 K-rows of 128 bytes
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t saddr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)(lbo_bytes >> 4) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

// one W tile (BN output columns x 64 reduction elements) of plane `m`: K-major = one box, MN-major = BN / 64 boxes
__device__ __forceinline__ void load_w_tile(uint32_t dst, const CUtensorMap* m, int kc, int n0, int BN, int b_mn,
                                            uint32_t bar) {
    if (!b_mn) {
        tma_load_2d(dst, m, kc * kBK, n0, bar);
    } else {
        for (int j = 0; j < BN / 64; ++j) tma_load_2d(dst + j * kBoxBytes, m, n0 + 64 * j, kc * kBK, bar);
    }
}

__device__ __forceinline__ void st_global_256(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t e,
                                              uint32_t f, uint32_t g, uint32_t h) {
    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d),
                 "r"(e), "r"(f), "r"(g), "r"(h) : "memory");
}

constexpr int kMaxSeg = 4;
struct AMaps { CUtensorMap hi[kMaxSeg], lo[kMaxSeg]; };

template <int MODE, bool MAPPED>
__global__ void __launch_bounds__(kThreads, 1) linear_planes_kernel(
    const __grid_constant__ AMaps am, const __grid_constant__ CUtensorMap w_hi,
    const __grid_constant__ CUtensorMap w_lo, const PlanesParams p, const int32_t* __restrict__ out_map) {
    // out_map (MAPPED only): output row m lands in row out_map[m] of Y / Yp, dropped when negative (valid-row selection of
    // the last curve-conv layer).  It is a kernel parameter of its own and MAPPED a template parameter on purpose: growing
    // PlanesParams by 8 bytes, or testing the pointer at run time, changed ptxas' allocation of the epilogue (103 -> 88
    // registers) and cost the big-M layers 45 % of their throughput (measured A/B on one box, profiles/r01h_ab.log).
    constexpr int P = MODE == 3 ? 2 : 1;
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * kMaxStages + 5];   // full[S], empty[S], wfull, tfull[2], tempty[2]
    __shared__ uint32_t tmem_base_slot;
    __shared__ float bias_s[256];                                // bias of the current N-slice (epilogue warps)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int BN = p.BN, stages = p.stages;
    const int nk = p.nk;
    const int wtile = BN * kBK * 2;                      // bytes of one W tile (one plane, one K chunk)
    const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t w_res = smem0;                        // resident weights: [nk][P][BN x 64]
    const uint32_t ring0 = smem0 + (p.resident ? (uint32_t)(nk * P * wtile) : 0u);
    const int stage_bytes = P * kATile + (p.resident ? 0 : P * wtile);
    const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[kMaxStages]);
    const uint32_t wfull = smem_u32(&bars[2 * kMaxStages]);
    const uint32_t tfull0 = smem_u32(&bars[2 * kMaxStages + 1]), tempty0 = smem_u32(&bars[2 * kMaxStages + 3]);
    uint32_t tmem_cols = 32;
    while ((int)tmem_cols < 2 * BN) tmem_cols <<= 1;

    const int total_tiles = p.tiles_m * p.tiles_n;

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 1);
        }
        mbar_init(wfull, 1);
        for (int b = 0; b < 2; ++b) {
            mbar_init(tfull0 + 8 * b, 1);
            mbar_init(tempty0 + 8 * b, kEpiThreads / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(smem_u32(&tmem_base_slot), tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;
    ccn_pdl_wait();      // barrier init + TMEM allocation overlapped the previous kernel's tail; global memory from here on

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        // converged warp, one elected lane issues (see the MMA issuer below); ring slot and phase are carried, not divided out
        {
            if (p.resident) {
                const int n0 = (blockIdx.x % p.tiles_n) * BN;
                if (elect_one()) {
                    mbar_arrive_expect_tx(wfull, (uint32_t)(nk * P * wtile));
                    for (int kc = 0; kc < nk; ++kc) {
                        load_w_tile(w_res + (kc * P) * wtile, &w_hi, kc, n0, BN, p.b_mn, wfull);
                        if (P == 2) load_w_tile(w_res + (kc * P + 1) * wtile, &w_lo, kc, n0, BN, p.b_mn, wfull);
                    }
                }
                __syncwarp();
            }
            int slot = 0;
            uint32_t phase = 1;                                         // parity to wait for on `empty`: slots start free
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
                const int m0 = (t / p.tiles_n) * kBM, n0 = (t % p.tiles_n) * BN;
                for (int kc = 0; kc < nk; ++kc) {
                    const CUtensorMap *mh, *ml;
                    int ka;                                             // column inside the segment's planes
                    int roff = 0;                                       // row shift of a convolution tap
                    if (p.taps) {
                        const int tap = kc / p.cpt;
                        mh = &am.hi[0]; ml = &am.lo[0]; ka = (kc - tap * p.cpt) * kBK; roff = tap - (p.taps >> 1);
                    }
                    else if (kc < p.sc1) { mh = &am.hi[0]; ml = &am.lo[0]; ka = kc * kBK; }
                    else if (kc < p.sc2) { mh = &am.hi[1]; ml = &am.lo[1]; ka = (kc - p.sc1) * kBK; }
                    else if (kc < p.sc3) { mh = &am.hi[2]; ml = &am.lo[2]; ka = (kc - p.sc2) * kBK; }
                    else { mh = &am.hi[3]; ml = &am.lo[3]; ka = (kc - p.sc3) * kBK; }
                    mbar_wait(empty0 + 8 * slot, phase);
                    const uint32_t dst = ring0 + slot * stage_bytes;
                    if (elect_one()) {
                        mbar_arrive_expect_tx(full0 + 8 * slot, (uint32_t)stage_bytes);
                        tma_load_2d(dst, mh, ka, m0 + roff, full0 + 8 * slot);      // rows outside [0, M) arrive as zeros
                        if (P == 2) tma_load_2d(dst + kATile, ml, ka, m0 + roff, full0 + 8 * slot);
                        if (!p.resident) {
                            load_w_tile(dst + P * kATile, &w_hi, kc, n0, BN, p.b_mn, full0 + 8 * slot);
                            if (P == 2) load_w_tile(dst + P * kATile + wtile, &w_lo, kc, n0, BN, p.b_mn, full0 + 8 * slot);
                        }
                    }
                    __syncwarp();
                    if (++slot == stages) { slot = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        // The whole warp runs the control flow (converged); ONE lane, chosen by elect.sync, issues.  Under `if (lane == 0)`
        // ptxas cannot prove the region single-threaded and wraps every tcgen05.mma in a per-lane loop
        // (ELECT ... @P0 BRA.U.ANY, ~7 extra instructions per MMA on the thread that bounds the narrow-tile layers).
        {
            if (p.resident) {
                mbar_wait(wfull, 0);
                tc_fence_after();
            }
            int slot = 0;
            uint32_t phase = 0;                                         // parity to wait for on `full`
            int it = 0;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
                const int n0 = (t % p.tiles_n) * BN;
                const int mma_n = min(BN, ((p.N - n0) + 15) & ~15);
                // instruction descriptor: D=f32, A=B=bf16, K-major both, N>>3 at bit 17, M>>4 at bit 24
                const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(mma_n >> 3) << 17) |
                                       ((uint32_t)(kBM >> 4) << 24) | (p.b_mn ? (1u << 16) : 0u);   // bit 16: B is MN-major
                const int buf = it & 1;
                mbar_wait(tempty0 + 8 * buf, ((it >> 1) & 1) ^ 1);     // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t acc = tmem_base + (uint32_t)(buf * BN);
                for (int kc = 0; kc < nk; ++kc) {
                    const int s = slot;
                    mbar_wait(full0 + 8 * s, phase);
                    tc_fence_after();
                    const uint32_t a_hi_s = ring0 + s * stage_bytes;
                    const uint32_t a_lo_s = a_hi_s + kATile;
                    const uint32_t w_hi_s = p.resident ? w_res + (kc * P) * wtile : a_hi_s + P * kATile;
                    const uint32_t w_lo_s = w_hi_s + wtile;
                    // The issuing thread is the bottleneck of the narrow-tile layers (12 MMAs per chunk, ~14 SASS instructions of
                    // descriptor arithmetic each = 0.65 us per chunk whatever the tile width, tools/bench_linear2.py small): the
                    // descriptors of a stage are built ONCE, a k-step of 16 bf16 = 32 bytes = +2 in their 16-byte address field
                    // (tiles are 1024-byte aligned: no carry out of the field)
                    const uint64_t da_hi = umma_desc_sw128(a_hi_s), da_lo = umma_desc_sw128(a_lo_s);
                    if (elect_one()) {
                    if (!p.b_mn) {
                        const uint64_t dw_hi = umma_desc_sw128(w_hi_s), dw_lo = umma_desc_sw128(w_lo_s);
#pragma unroll
                        for (int k = 0; k < kBK / 16; ++k) {
                            umma_bf16(acc, da_hi + 2 * k, dw_hi + 2 * k, idesc, (kc | k) ? 1u : 0u);
                            if (P == 2) {
                                umma_bf16(acc, da_lo + 2 * k, dw_hi + 2 * k, idesc, 1u);
                                umma_bf16(acc, da_hi + 2 * k, dw_lo + 2 * k, idesc, 1u);
                            }
                        }
                    } else {
#pragma unroll
                        for (int k = 0; k < kBK / 16; ++k) {
                            // MN-major B: 16 reduction rows of 128 bytes further down the box
                            const uint64_t dwh = umma_desc_mn_sw128(w_hi_s + k * 2048, kBoxBytes);
                            umma_bf16(acc, da_hi + 2 * k, dwh, idesc, (kc | k) ? 1u : 0u);
                            if (P == 2) {
                                const uint64_t dwl = umma_desc_mn_sw128(w_lo_s + k * 2048, kBoxBytes);
                                umma_bf16(acc, da_lo + 2 * k, dwh, idesc, 1u);
                                umma_bf16(acc, da_hi + 2 * k, dwl, idesc, 1u);
                            }
                        }
                    }
                    umma_commit(empty0 + 8 * s);        // stage reusable once these MMAs have read it
                    }
                    __syncwarp();
                    if (++slot == stages) { slot = 0; phase ^= 1u; }
                }
                if (elect_one()) umma_commit(tfull0 + 8 * buf);          // accumulator of this tile complete
                __syncwarp();
            }
        }
    } else {
        // ------------------------------------------------------------------ epilogue (warps 2..9)
        // Two warps per TMEM lane quarter (= per scheduler), taking alternate 32-column chunks of the accumulator: with one
        // warp per scheduler the chunk loop (tcgen05.ld -> bias / activation / bf16 split -> 32-byte stores, ~10 dependent
        // instructions per element) ran unhidden and bounded every K <= 256 layer at ~3.5 us per 128 x 128 tile against
        // 1.4 us of MMA (ncu source view, profiles/r02q_linear_epilogue.txt)
        const int q = warp & 3;                         // TMEM lane quarter this warp may access
        const int half = (warp - 2) >> 2;               // which of the quarter's two warps
        const bool vecy = p.Y && ((p.ldy & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.Y) & 31) == 0);
        int it = 0, bias_n0 = -1;
        const float* bs = bias_s;
        const int et = threadIdx.x - 64;                // 0..255 over the eight epilogue warps
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
            const int m0 = (t / p.tiles_n) * kBM, n0 = (t % p.tiles_n) * BN;
            const int mma_n = min(BN, ((p.N - n0) + 15) & ~15);
            const int buf = it & 1;
            if (n0 != bias_n0) {                        // once per CTA in resident mode
                asm volatile("bar.sync 1, 256;" ::: "memory");      // everyone is done with the previous slice
                for (int j = et; j < BN; j += kEpiThreads) bias_s[j] = (p.bias && n0 + j < p.N) ? __ldg(p.bias + n0 + j) : 0.f;
                asm volatile("bar.sync 1, 256;" ::: "memory");
                bias_n0 = n0;
            }
            mbar_wait(tfull0 + 8 * buf, (it >> 1) & 1);
            tc_fence_after();
            const int row = m0 + q * 32 + lane;
            const int orow = MAPPED ? (row < p.M ? __ldg(out_map + row) : -1) : row;
            const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BN);
            // The activation is a compile-time constant of the chunk loop (switch below): ptxas only unswitched the
            // run-time test on its own while the parameter struct stayed small, and the branchy form (two uniform branches
            // per element, 150 BRA in the kernel) ran the HBM-bound layers 1.6-1.8x slower (profiles/r01h_ab.log).
            auto chunks = [&](auto act_tag) {
            constexpr int ACT = decltype(act_tag)::value;
            auto emit = [&](const uint32_t (&v)[32], const int c0) {
                float y[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    float o = __uint_as_float(v[j]) + bs[c0 + j];
                    if (ACT == CCN_ACT_RELU) o = fmaxf(o, 0.f);
                    else if (ACT == CCN_ACT_LEAKY) o = o > 0.f ? o : o * p.slope;
                    y[j] = o;
                }
                if (MAPPED ? orow >= 0 : row < p.M) {
                    if (p.Y) {
                        float* yrow = p.Y + (int64_t)orow * p.ldy + n0 + c0;
#pragma unroll
                        for (int j = 0; j < 32; j += 8) {
                            const int n = n0 + c0 + j;
                            if (vecy && n + 7 < p.N) {          // one full 32-byte sector per store
                                st_global_256(yrow + j, __float_as_uint(y[j]), __float_as_uint(y[j + 1]),
                                              __float_as_uint(y[j + 2]), __float_as_uint(y[j + 3]), __float_as_uint(y[j + 4]),
                                              __float_as_uint(y[j + 5]), __float_as_uint(y[j + 6]), __float_as_uint(y[j + 7]));
                            } else {
#pragma unroll
                                for (int e = 0; e < 8; ++e)
                                    if (n + e < p.N) yrow[j + e] = y[j + e];
                            }
                        }
                    }
                    if (p.Yp) {
                        __nv_bfloat16* hrow = p.Yp + (int64_t)orow * p.yp_ld + n0 + c0;
#pragma unroll
                        for (int j = 0; j < 32; j += 16) {
                            if (n0 + c0 + j >= p.yp_cols) continue;    // yp_cols % 16 == 0: whole groups only
                            uint32_t h[8], l[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                const __nv_bfloat162 hv = __floats2bfloat162_rn(y[j + 2 * e], y[j + 2 * e + 1]);
                                h[e] = *reinterpret_cast<const uint32_t*>(&hv);
                                if (P == 2) {
                                    const float2 f = __bfloat1622float2(hv);
                                    const __nv_bfloat162 lv = __floats2bfloat162_rn(y[j + 2 * e] - f.x, y[j + 2 * e + 1] - f.y);
                                    l[e] = *reinterpret_cast<const uint32_t*>(&lv);
                                }
                            }
                            st_global_256(hrow + j, h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);
                            if (P == 2) st_global_256(hrow + p.yp_plane + j, l[0], l[1], l[2], l[3], l[4], l[5], l[6], l[7]);
                        }
                    }
                }
            };
            for (int c0 = half * 32; c0 < mma_n; c0 += 64) {
                uint32_t v[32];
                tmem_ld32(tbase + (uint32_t)c0, v);
                emit(v, c0);
            }
            };
            if (p.act == CCN_ACT_RELU) chunks(std::integral_constant<int, CCN_ACT_RELU>{});
            else if (p.act == CCN_ACT_LEAKY) chunks(std::integral_constant<int, CCN_ACT_LEAKY>{});
            else chunks(std::integral_constant<int, CCN_ACT_NONE>{});
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty0 + 8 * buf);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, tmem_cols);
    }
}

// Up to four fp32 column blocks (each optionally row-gathered through an int64 index) -> bf16 planes.
// One thread per (row, group of 8 output columns); pad columns up to kpad are written as zeros.
struct PackSrc {
    const float* p[4];
    const int64_t* idx[4];
    int64_t ld[4];
    int c[4];
};

__global__ void pack_planes_kernel(const PackSrc s, int64_t n, int ctot, __nv_bfloat16* __restrict__ planes,
                                   int64_t kpad, int64_t plane_stride, int n_planes) {
    ccn_pdl_wait();
    const int groups = (int)(kpad >> 3);
    const int64_t total = n * groups;
    const int e0 = s.c[0], e1 = e0 + s.c[1], e2 = e1 + s.c[2];
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t i;
        int g;
        ccn_divmod(t, groups, i, g);
        float v[8];
        bool vec = false;
        const int c0 = g * 8;
        if (c0 + 8 <= ctot) {
            // all 8 columns inside one source block and 16-byte aligned: two float4 loads
            const int k = c0 < e0 ? 0 : c0 < e1 ? 1 : c0 < e2 ? 2 : 3;
            const int kend = k == 0 ? e0 : k == 1 ? e1 : k == 2 ? e2 : ctot;
            if (c0 + 8 <= kend) {
                const int cc = c0 - (k == 0 ? 0 : k == 1 ? e0 : k == 2 ? e1 : e2);
                const int64_t r = s.idx[k] ? s.idx[k][i] : i;
                const float* src = s.p[k] + r * s.ld[k] + cc;
                if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
                    const float4 a = *reinterpret_cast<const float4*>(src), b = *reinterpret_cast<const float4*>(src + 4);
                    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
                    vec = true;
                }
            }
        }
        if (!vec) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int c = c0 + e;
                float x = 0.f;
                if (c < ctot) {
                    const int k = c < e0 ? 0 : c < e1 ? 1 : c < e2 ? 2 : 3;
                    const int cc = c - (k == 0 ? 0 : k == 1 ? e0 : k == 2 ? e1 : e2);
                    const int64_t r = s.idx[k] ? s.idx[k][i] : i;
                    x = s.p[k][r * s.ld[k] + cc];
                }
                v[e] = x;
            }
        }
        __nv_bfloat162 h[4], l[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            h[e] = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
            const float2 f = __bfloat1622float2(h[e]);
            l[e] = __floats2bfloat162_rn(v[2 * e] - f.x, v[2 * e + 1] - f.y);
        }
        __nv_bfloat16* o = planes + i * kpad + g * 8;
        *reinterpret_cast<uint4*>(o) = *reinterpret_cast<uint4*>(h);
        if (n_planes == 2) *reinterpret_cast<uint4*>(o + plane_stride) = *reinterpret_cast<uint4*>(l);
    }
}

}  // namespace

extern "C" int ccn_pack_planes(const float* s0, int64_t ld0, int32_t c0, const int64_t* i0, const float* s1, int64_t ld1,
                               int32_t c1, const int64_t* i1, const float* s2, int64_t ld2, int32_t c2,
                               const int64_t* i2, const float* s3, int64_t ld3, int32_t c3, const int64_t* i3, int64_t n,
                               void* planes, int64_t kpad, int64_t plane_stride, int32_t n_planes, ccn_stream_t stream) {
    const int ctot = c0 + c1 + c2 + c3;
    CCN_REQUIRE(n >= 0 && planes && c0 > 0 && s0 && c1 >= 0 && c2 >= 0 && c3 >= 0 && (c1 == 0 || s1) && (c2 == 0 || s2) &&
                (c3 == 0 || s3));
    CCN_REQUIRE(kpad >= ctot && (kpad & 7) == 0 && (n_planes == 1 || n_planes == 2) &&
                (reinterpret_cast<uintptr_t>(planes) & 15) == 0 && (n_planes == 1 || (plane_stride & 7) == 0));
    if (n == 0) return CCN_OK;
    PackSrc s;
    s.p[0] = s0; s.p[1] = s1; s.p[2] = s2; s.p[3] = s3;
    s.idx[0] = i0; s.idx[1] = i1; s.idx[2] = i2; s.idx[3] = i3;
    s.ld[0] = ld0; s.ld[1] = ld1; s.ld[2] = ld2; s.ld[3] = ld3;
    s.c[0] = c0; s.c[1] = c1; s.c[2] = c2; s.c[3] = c3;
    ccn_launch(pack_planes_kernel, dim3(ccn_grid(n * (kpad >> 3), 256, 148LL * 32)), dim3(256), 0, (cudaStream_t)stream, 
        s, n, ctot, (__nv_bfloat16*)planes, kpad, plane_stride, n_planes);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

static int launch_planes(const void* const* ap, const int64_t* ald, const int64_t* aps, const int32_t* ak, int32_t taps,
                         int32_t b_mn, const int32_t* out_map, const void* w_hi_tmap_host, const void* w_lo_tmap_host,
                         const float* bias, float* Y,
                         int64_t ldy, void* y_planes, int64_t yp_ld, int64_t yp_plane_stride, int64_t M, int32_t N,
                         int32_t box_n, int32_t mode, int32_t act, float slope, ccn_stream_t stream) {
    const void* a0 = ap[0];
    const int32_t k0 = ak[0];
    CCN_REQUIRE(a0 && k0 > 0 && w_hi_tmap_host && (Y || y_planes) && M >= 0 && M < 0x7fffffff && N > 0);
    CCN_REQUIRE((mode == 1 || (mode == 3 && w_lo_tmap_host)) && act >= 0 && act <= 2);
    CCN_REQUIRE(box_n == 32 || box_n == 64 || box_n == 128 || box_n == 256);
    CCN_REQUIRE(!b_mn || (box_n >= 64 && taps == 0));
    CCN_REQUIRE(!Y || ldy >= N);
    const int yp_cols = (N + 15) & ~15;
    CCN_REQUIRE(!y_planes || (yp_ld >= yp_cols && (yp_ld & 15) == 0 && (reinterpret_cast<uintptr_t>(y_planes) & 31) == 0 &&
                              (mode == 1 || (yp_plane_stride & 15) == 0)));
    PlanesParams p;
    p.nseg = 0;
    int chunk0[kMaxSeg + 1] = {0, 0, 0, 0, 0};
    int K = 0;
    for (int i = 0; i < kMaxSeg; ++i) {
        if (ak[i] <= 0) break;
        CCN_REQUIRE(ap[i] && ald[i] >= ak[i] && (ald[i] & 7) == 0 && (reinterpret_cast<uintptr_t>(ap[i]) & 15) == 0 &&
                    (mode == 1 || (aps[i] & 7) == 0));
        chunk0[i + 1] = chunk0[i] + (ak[i] + kBK - 1) / kBK;
        K += ak[i];
        p.nseg = i + 1;
    }
    for (int i = p.nseg; i < kMaxSeg; ++i) {
        CCN_REQUIRE(ak[i] <= 0);
        chunk0[i + 1] = chunk0[p.nseg];
    }
    p.sc1 = chunk0[1]; p.sc2 = chunk0[2]; p.sc3 = chunk0[3]; p.nk = chunk0[p.nseg];
    p.taps = 0; p.cpt = 0; p.b_mn = b_mn;
    if (taps > 0) {
        CCN_REQUIRE(p.nseg == 1 && (taps & 1) && taps <= 15);
        p.taps = taps; p.cpt = chunk0[1]; p.nk = taps * p.cpt;
    }
    if (M == 0) return CCN_OK;
    const int P = mode == 3 ? 2 : 1;
    p.bias = bias; p.Y = Y; p.ldy = ldy; p.Yp = (__nv_bfloat16*)y_planes; p.yp_ld = yp_ld; p.yp_plane = yp_plane_stride;
    p.yp_cols = yp_cols; p.M = (int)M; p.N = N; p.K = K; p.BN = box_n;
    p.tiles_m = (int)((M + kBM - 1) / kBM);
    p.tiles_n = (N + box_n - 1) / box_n;
    p.act = act; p.slope = slope;
    const int nk = p.nk;
    const int wtile = box_n * kBK * 2;
    const int64_t total = (int64_t)p.tiles_m * p.tiles_n;
    const int budget = kSmemLimit - 2560;        // dynamic limit (kSmemLimit - 1536; 1.2 KB is static) minus the 1024-byte alignment slack
    // resident weights when the CTA's N-slice fits next to a >= 3-stage A ring and a CTA sees more than one tile
    const int64_t wres = (int64_t)nk * P * wtile;
    int grid = (int)(total < 148 ? total : 148);
    p.resident = 0;
    if (total > 148 && p.tiles_n <= 148 && wres + 3 * P * kATile <= budget) {
        p.resident = 1;
        grid = (148 / p.tiles_n) * p.tiles_n;
    }
    const int stage_bytes = P * kATile + (p.resident ? 0 : P * wtile);
    int stages = (int)((budget - (p.resident ? wres : 0)) / stage_bytes);
    if (stages > kMaxStages) stages = kMaxStages;
    if (stages < 2) return CCN_EUNSUPPORTED;
    p.stages = stages;
    const int smem = (int)((p.resident ? wres : 0) + (int64_t)stages * stage_bytes + 1024);
    alignas(64) AMaps am;
    for (int i = 0; i < kMaxSeg; ++i) {
        if (i >= p.nseg) {                        // unused slots repeat segment 0 (a valid descriptor): no driver call
            am.hi[i] = am.hi[0];
            am.lo[i] = am.lo[0];
            continue;
        }
        int rc = ccn_encode_bf16_map(&am.hi[i], ap[i], M, ald[i], kBM);
        if (rc != CCN_OK) return rc;
        if (mode == 3) {
            rc = ccn_encode_bf16_map(&am.lo[i], (const void*)((const __nv_bfloat16*)ap[i] + aps[i]), M, ald[i], kBM);
            if (rc != CCN_OK) return rc;
        } else {
            am.lo[i] = am.hi[i];
        }
    }
    const CUtensorMap* whi = (const CUtensorMap*)w_hi_tmap_host;
    const CUtensorMap* wlo = mode == 3 ? (const CUtensorMap*)w_lo_tmap_host : whi;
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e;
    static bool attr[4] = {false, false, false, false};
    const int which = (mode == 3 ? 2 : 0) + (out_map ? 1 : 0);
    if (!attr[which]) {
        switch (which) {
            case 0: e = cudaFuncSetAttribute(linear_planes_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit - 1536); break;
            case 1: e = cudaFuncSetAttribute(linear_planes_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit - 1536); break;
            case 2: e = cudaFuncSetAttribute(linear_planes_kernel<3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit - 1536); break;
            default: e = cudaFuncSetAttribute(linear_planes_kernel<3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit - 1536); break;
        }
        if (e != cudaSuccess) return (int)e;
        attr[which] = true;
    }
    switch (which) {
        case 0: ccn_launch(linear_planes_kernel<1, false>, dim3(grid), dim3(kThreads), smem, s, am, *whi, *wlo, p, out_map); break;
        case 1: ccn_launch(linear_planes_kernel<1, true>, dim3(grid), dim3(kThreads), smem, s, am, *whi, *wlo, p, out_map); break;
        case 2: ccn_launch(linear_planes_kernel<3, false>, dim3(grid), dim3(kThreads), smem, s, am, *whi, *wlo, p, out_map); break;
        default: ccn_launch(linear_planes_kernel<3, true>, dim3(grid), dim3(kThreads), smem, s, am, *whi, *wlo, p, out_map); break;
    }
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_linear_planes(const void* a0, int64_t ld0, int64_t ps0, int32_t k0, const void* a1, int64_t ld1,
                                 int64_t ps1, int32_t k1, const void* a2, int64_t ld2, int64_t ps2, int32_t k2,
                                 const void* a3, int64_t ld3, int64_t ps3, int32_t k3, const void* w_hi_tmap_host,
                                 const void* w_lo_tmap_host, const float* bias, float* Y, int64_t ldy, void* y_planes,
                                 int64_t yp_ld, int64_t yp_plane_stride, int64_t M, int32_t N, int32_t box_n,
                                 int32_t mode, int32_t act, float slope, ccn_stream_t stream) {
    const void* ap[kMaxSeg] = {a0, a1, a2, a3};
    const int64_t ald[kMaxSeg] = {ld0, ld1, ld2, ld3}, aps[kMaxSeg] = {ps0, ps1, ps2, ps3};
    const int32_t ak[kMaxSeg] = {k0, k1, k2, k3};
    return launch_planes(ap, ald, aps, ak, 0, 0, nullptr, w_hi_tmap_host, w_lo_tmap_host, bias, Y, ldy, y_planes, yp_ld,
                         yp_plane_stride, M, N, box_n, mode, act, slope, stream);
}

// Y[M, N] = A[M, K] W[K, N] with W given as a (K, N) row-major split-bf16 weight - the data gradient dX = dY W computed
// from the FORWARD weight's planes (no transposed copy, no second split): the tensor maps must have a 64 x 64 box
// (ccn_weight_tensor_map(..., box_n = 64) on the (K rows, N-padded cols) planes); box_n here is the OUTPUT tile width.
extern "C" int ccn_linear_planes_bt(const void* a, int64_t ld, int64_t ps, int32_t k, const void* w_hi_tmap_host,
                                    const void* w_lo_tmap_host, float* Y, int64_t ldy, void* y_planes, int64_t yp_ld,
                                    int64_t yp_plane_stride, int64_t M, int32_t N, int32_t box_n, int32_t mode,
                                    ccn_stream_t stream) {
    const void* ap[kMaxSeg] = {a, nullptr, nullptr, nullptr};
    const int64_t ald[kMaxSeg] = {ld, 0, 0, 0}, aps[kMaxSeg] = {ps, 0, 0, 0};
    const int32_t ak[kMaxSeg] = {k, 0, 0, 0};
    return launch_planes(ap, ald, aps, ak, 0, 1, nullptr, w_hi_tmap_host, w_lo_tmap_host, nullptr, Y, ldy, y_planes, yp_ld,
                         yp_plane_stride, M, N, box_n, mode, CCN_ACT_NONE, 0.f, stream);
}

// 'same' zero-padded 1-D convolution along the row axis as ONE GEMM: K = taps x cin, tap t of output row m reads
// input row m + t - taps/2 (a TMA row coordinate; rows outside the signal arrive as zeros).  out_row_map (optional,
// int32 [M]): output row m is stored at row out_row_map[m], dropped when negative (the stack's last layer keeps the
// curve points and drops the pad rows).  W: (N, taps x cin) in the
// ccn_split_bf16 layout with every tap's cin columns starting at a multiple of 64.
extern "C" int ccn_conv_planes(const void* a, int64_t ld, int64_t ps, int32_t cin, int32_t taps,
                               const void* w_hi_tmap_host, const void* w_lo_tmap_host, const float* bias, float* Y,
                               int64_t ldy, void* y_planes, int64_t yp_ld, int64_t yp_plane_stride, int64_t M, int32_t N,
                               int32_t box_n, int32_t mode, int32_t act, float slope, const int32_t* out_row_map,
                               ccn_stream_t stream) {
    const void* ap[kMaxSeg] = {a, nullptr, nullptr, nullptr};
    const int64_t ald[kMaxSeg] = {ld, 0, 0, 0}, aps[kMaxSeg] = {ps, 0, 0, 0};
    const int32_t ak[kMaxSeg] = {cin, 0, 0, 0};
    CCN_REQUIRE(taps >= 1);
    return launch_planes(ap, ald, aps, ak, taps, 0, out_row_map, w_hi_tmap_host, w_lo_tmap_host, bias, Y, ldy, y_planes, yp_ld,
                         yp_plane_stride, M, N, box_n, mode, act, slope, stream);
}
