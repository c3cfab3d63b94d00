// One kernel per edge-MLP step: gather -> layer 1 (decomposed) -> layers 2..n -> max / mean / attention aggregate, the
// per-edge intermediates never leave the SM (sm_100a only; north_star (f)).
//
//   ccn_sgcnn_fused      StaticEdgeConv.forward_fast (reference src/models/modules/dgcnn.py:158-207): rows (i, g) of the
//                        dense K+1 neighbour table, layer 1 = act(P[j] + Q[i]) from the per-POINT products PQ, the remaining
//                        layers of `nn`, max over the K+1 rows with missing neighbours at -100 (dgcnn.py:187-189).
//   ccn_pointconv_fused  PointNetConv2.message/aggregate (src/models/modules/point_conv.py:60-93) and the sparse
//                        StaticEdgeConv.forward_slow (dgcnn.py:209-246) over a target-sorted CSR edge list: layer 1 =
//                        act(H[j] + Wp (pos_j - pos_i)/r + b) or act(P[j] + Q[i]), the remaining layers of local_nn, then
//                        max | mean | sum sigmoid(att) msg | sum softmax_target(att) msg with att = attend_nn(msg) - the two
//                        attention layers run in the same kernel, msg stays in TMEM meanwhile.
//
// Before: layer-1 kernel (writes E x C planes) -> one ccn_linear_planes launch per layer (reads and re-writes E x C) ->
// aggregation kernel (reads E x C again): ~585 MB through L2/HBM for the 374 k-row level-1 SGCNN step against ~11 MB of
// compulsory bytes (N x C in, N x C out, the neighbour table, the weights).  Here a CTA owns a tile of <= 128 edge rows
// that are WHOLE neighbour groups (6 groups of K+1 = 21 rows, or the CSR segments whose first edge falls into a stride
// of S = 129 - maxdeg rows):
//   warps 2-5 (128 threads)  form the layer-1 rows straight into the SWIZZLE_128B K-major A tile in shared memory as
//                            bf16 (hi, lo) planes; after every layer read the fp32 accumulator from TMEM (tcgen05.ld),
//                            add the folded-BatchNorm bias, apply the activation, split to bf16 (hi, lo) and store it IN
//                            PLACE as the next layer's A operand; after the last layer transpose 32-column chunks
//                            through shared memory and reduce every neighbour group / CSR segment in edge order.
//   warp 0 lane 0            TMA producer: streams the weight tiles (<= 128 output rows x 64 k, hi [+ lo]) of all layers
//                            through a ring, running ahead across layers and tiles (all layers' weights live in ONE
//                            bf16 matrix, one tensor map per plane, 64 x 64 boxes).
//   warp 1 lane 0            tcgen05.mma kind::f16 issuer: A from the resident tile, B from the ring, fp32 accumulators
//                            in TMEM (x3 mode: A_hi W_hi + A_lo W_hi + A_hi W_lo, same order as ccn_linear_planes, so a
//                            fused step is bit-identical to the unfused chain).
// The chain inside a tile is serial (layer l+1 needs all of layer l); CTAs are independent, so where shared memory allows
// two CTAs share an SM and one's MMAs overlap the other's epilogue.  Attention: msg accumulates in TMEM columns [0, 256),
// the other layers in [256, 512).
#include <type_traits>

#include "common.cuh"
#include "tc_common.cuh"

namespace {

using namespace ccn_tc;

constexpr int kRows = 128;
constexpr int kChunkBytes = kRows * 128;      // one 128-row x 64-column bf16 plane chunk (SWIZZLE_128B)
constexpr int kThreads = 320;                // warp 0 TMA, warp 1 MMA, warps 2-9 row workers
constexpr int kWorkers = 256;
constexpr int kSegTable = 264;                // segment offsets staged per tile (tiles with more, i.e. runs of empty targets, read row_ptr)
constexpr int kU = 2;                         // layer-1 items in flight per worker thread (registers: two CTAs per SM need <= 102)
constexpr int kMaxStages = 6;
constexpr int kMaxWidth = 256;                // widest activation kept resident / widest layer output
constexpr int kTable = kRows * 32;            // floats of one transposition table: 128 rows x 32 columns, column XOR (row & 31)

enum { FORM_DENSE = 0, FORM_PAIR = 1, FORM_POINT = 2, FORM_ROWS = 3 };   // FORM_ROWS: per-point MLP chain, A rows = input planes
constexpr int kAggNone = 100;                 // FORM_ROWS: no aggregation, the last layer's rows are the result

struct ChainLayerDev {
    int K, N, w_row, act;
    float slope;
    const float* bias;
};

struct ChainParams {
    int L;
    ChainLayerDev ly[CCN_CHAIN_MAX_LAYERS];
    int stages, res_chunks;        // weight ring depth; resident activation chunks (max width / 64)
    int w_plane_bytes;             // bytes of one plane of a ring stage: min(128, widest layer) rows x 128 B
    int n_tiles;
    int c1, act0;                  // width and activation of the (decomposed) first layer
    float slope0;
    // FORM_DENSE: rows (i, g), g = 0 self loop, g >= 1 neighbour idx[i][g-1] (-1 = missing)
    const float* PQ;               // (n, 2 c1): [P | Q]
    int64_t ldpq;
    const int32_t* idx;
    int n, Knb, G, gpt;            // points, neighbours per point, rows per group (K+1), groups per tile
    // CSR forms: rows = edges, segments = targets
    const int32_t *row_ptr, *erow, *ecol;
    const int4* tiles;             // per tile: {first target, targets, first edge, edges} (csr_tiles_kernel)
    int nq, E, S;                  // targets, edges, tile stride in edge rows (129 - max degree)
    // FORM_POINT
    const float* H;                // (n_src, c1) or null
    int64_t ldh;
    const float *wp, *bias0, *pos_src, *pos_dst;   // (3, c1), (c1), (n_src, 3), (nq, 3)
    float radius;                  // > 0: relative positions divided by it
    // FORM_ROWS: the chain's input activations [A_0 | A_1 | ...] as up to four split-bf16 planes tensors (M rows; segment s
    // has seg_k[s] columns, row stride seg_ld[s] with zero pad columns, and occupies the 64-column chunks from seg_c0[s] on)
    const __nv_bfloat16* seg_p[4];
    int64_t seg_ld[4], seg_ps[4];
    int seg_k[4], seg_c0[4], nsegs;
    int M;
    // aggregation
    int agg, msg_layer;            // CCN_AGG_*; layer whose output is the message (attention modes), else -1
    float* out;
    int64_t ldo;
    __nv_bfloat16* planes;
    int64_t kpad, pstride;
    int npl;
};

__device__ __forceinline__ float act_apply(float y, int act, float slope) {
    if (act == CCN_ACT_RELU) return fmaxf(y, 0.f);
    if (act == CCN_ACT_LEAKY) return y > 0.f ? y : y * slope;
    return y;
}

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// 8 fp32 values -> 8 bf16 hi (+ 8 bf16 lo = bf16(x - hi)) stored as one 16-byte piece per plane
template <int P>
__device__ __forceinline__ void store_piece(uint32_t addr_hi, const float (&y)[8]) {
    uint32_t h[4], l[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const __nv_bfloat162 hv = __floats2bfloat162_rn(y[2 * e], y[2 * e + 1]);
        h[e] = *reinterpret_cast<const uint32_t*>(&hv);
        if (P == 2) {
            const float2 f = __bfloat1622float2(hv);
            const __nv_bfloat162 lv = __floats2bfloat162_rn(y[2 * e] - f.x, y[2 * e + 1] - f.y);
            l[e] = *reinterpret_cast<const uint32_t*>(&lv);
        }
    }
    st_shared_v4(addr_hi, h[0], h[1], h[2], h[3]);
    if (P == 2) st_shared_v4(addr_hi + kChunkBytes, l[0], l[1], l[2], l[3]);
}

// 1-D bulk async copy global -> shared (TMA engine), completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

__device__ __forceinline__ void workers_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

template <int P, int FORM>
__global__ void __launch_bounds__(kThreads, 2) edge_chain_kernel(const __grid_constant__ CUtensorMap w_hi,
                                                                 const __grid_constant__ CUtensorMap w_lo,
                                                                 const __grid_constant__ ChainParams p) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * kMaxStages + 2];     // w_full[S], w_empty[S], a_ready, acc_full
    __shared__ uint32_t tmem_base_slot;
    __shared__ float bias_s[CCN_CHAIN_MAX_LAYERS][kMaxWidth];
    __shared__ float wp_s[4][kMaxWidth];                            // FORM_POINT: Wp rows 0..2, bias of layer 1
    __shared__ int32_t row_i[kRows], row_j[kRows];                  // target / source of every tile row (-1: none)
    __shared__ float row_rel[kRows][3];                             // FORM_POINT: (pos_j - pos_i) [/ r]
    __shared__ int32_t seg_s[kSegTable];                            // CSR forms: first tile row of every segment (+ end)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool att_mode = p.msg_layer >= 0;
    const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t res0 = smem0;                                    // resident activations [chunk][plane][128 x 128 B]
    const uint32_t ring0 = smem0 + (uint32_t)(p.res_chunks * P * kChunkBytes);
    const int stage_bytes = P * p.w_plane_bytes;                    // one W tile: <= 128 rows x 64 k per plane
    const int stages = p.stages;
    const uint32_t wfull0 = smem_u32(&bars[0]), wempty0 = smem_u32(&bars[kMaxStages]);
    const uint32_t a_ready = smem_u32(&bars[2 * kMaxStages]), acc_full = smem_u32(&bars[2 * kMaxStages + 1]);
    int maxn = 32;
    for (int l = 0; l < p.L; ++l) maxn = max(maxn, p.ly[l].N);
    uint32_t tmem_cols = 32;
    while ((int)tmem_cols < maxn) tmem_cols <<= 1;
    if (att_mode) tmem_cols = 512;
    const uint32_t chain_col = att_mode ? 256u : 0u;

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(wfull0 + 8 * s, 1);
            mbar_init(wempty0 + 8 * s, 1);
        }
        mbar_init(a_ready, 8);
        mbar_init(acc_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    ccn_pdl_wait();      // global memory (biases included: the optimizer's kernels may precede a forward) from here on
    for (int t = threadIdx.x; t < p.L * kMaxWidth; t += kThreads) {
        const int l = t / kMaxWidth, c = t % kMaxWidth;
        bias_s[l][c] = (p.ly[l].bias && c < p.ly[l].N) ? __ldg(p.ly[l].bias + c) : 0.f;
    }
    if (FORM == FORM_POINT) {
        for (int t = threadIdx.x; t < 4 * kMaxWidth; t += kThreads) {
            const int r = t / kMaxWidth, c = t % kMaxWidth;
            float v = 0.f;
            if (c < p.c1) v = r < 3 ? __ldg(p.wp + r * p.c1 + c) : (p.bias0 ? __ldg(p.bias0 + c) : 0.f);
            wp_s[r][c] = v;
        }
    }
    if (warp == 0) tmem_alloc(smem_u32(&tmem_base_slot), tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer: the weight stream
        // (converged warp, one elected lane issues; ring slot / phase carried - see gemm_planes.cu)
        {
            int s = 0;
            uint32_t ph = 1;
            for (int t = blockIdx.x; t < p.n_tiles; t += gridDim.x) {
                for (int l = 0; l < p.L; ++l) {
                    const int N = p.ly[l].N, nk = p.ly[l].K >> 6, wrow = p.ly[l].w_row;
                    for (int n0 = 0; n0 < N; n0 += 128) {
                        const int mma_n = (min(128, N - n0) + 63) & ~63;     // whole 64-row boxes (rows past the matrix: zeros)
                        for (int kc = 0; kc < nk; ++kc) {
                            mbar_wait(wempty0 + 8 * s, ph);
                            const uint32_t dst = ring0 + s * stage_bytes;
                            if (elect_one()) {
                                mbar_arrive_expect_tx(wfull0 + 8 * s, (uint32_t)(P * mma_n * 128));
                                for (int b = 0; b < mma_n; b += 64) {
                                    tma_load_2d(dst + b * 128, &w_hi, kc * 64, wrow + n0 + b, wfull0 + 8 * s);
                                    if (P == 2) tma_load_2d(dst + p.w_plane_bytes + b * 128, &w_lo, kc * 64, wrow + n0 + b, wfull0 + 8 * s);
                                }
                            }
                            __syncwarp();
                            if (++s == stages) { s = 0; ph ^= 1u; }
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        {
            uint32_t aphase = 0, ph = 0;
            int s = 0;
            for (int t = blockIdx.x; t < p.n_tiles; t += gridDim.x) {
                for (int l = 0; l < p.L; ++l) {
                    const int N = p.ly[l].N, nk = p.ly[l].K >> 6;
                    mbar_wait(a_ready, aphase);                      // the layer's A operand is complete in shared memory
                    aphase ^= 1;
                    tc_fence_after();
                    const uint32_t acc = tmem_base + ((att_mode && l == p.msg_layer) ? 0u : chain_col);
                    for (int n0 = 0; n0 < N; n0 += 128) {
                        const int mma_n = (min(128, N - n0) + 15) & ~15;
                        // instruction descriptor: D=f32, A=B=bf16, K-major both, N>>3 at bit 17, M>>4 at bit 24
                        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(mma_n >> 3) << 17) |
                                               ((uint32_t)(kRows >> 4) << 24);
                        for (int kc = 0; kc < nk; ++kc) {
                            mbar_wait(wfull0 + 8 * s, ph);
                            tc_fence_after();
                            const uint32_t a_hi = res0 + (uint32_t)(kc * P) * kChunkBytes, a_lo = a_hi + kChunkBytes;
                            const uint32_t b_hi = ring0 + s * stage_bytes, b_lo = b_hi + p.w_plane_bytes;
                            // descriptors once per chunk; a k-step of 16 bf16 = 32 bytes = +2 in the 16-byte address field
                            const uint64_t da_hi = umma_desc_sw128(a_hi), da_lo = umma_desc_sw128(a_lo);
                            const uint64_t db_hi = umma_desc_sw128(b_hi), db_lo = umma_desc_sw128(b_lo);
                            if (elect_one()) {
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    umma_bf16(acc + n0, da_hi + 2 * k, db_hi + 2 * k, idesc, (kc | k) ? 1u : 0u);
                                    if (P == 2) {
                                        umma_bf16(acc + n0, da_lo + 2 * k, db_hi + 2 * k, idesc, 1u);
                                        umma_bf16(acc + n0, da_hi + 2 * k, db_lo + 2 * k, idesc, 1u);
                                    }
                                }
                                umma_commit(wempty0 + 8 * s);       // ring slot reusable once these MMAs have read it
                            }
                            __syncwarp();
                            if (++s == stages) { s = 0; ph ^= 1u; }
                        }
                    }
                    if (elect_one()) umma_commit(acc_full);         // accumulator complete, A operand no longer read
                    __syncwarp();
                }
            }
        }
    } else {
        // ------------------------------------------------------------------ row workers (warps 2..9)
        const int wt = threadIdx.x - 64;                            // 0..255
        const int q = warp & 3;                                     // TMEM lane quarter this warp may access
        const int half = (warp - 2) >> 2;                           // warps w and w + 4 share a lane quarter: they split the columns
        const int myrow = q * 32 + lane;
        const uint32_t lane_base = (uint32_t)(q * 32) << 16;
        float* T = reinterpret_cast<float*>(smem_raw + (smem0 - smem_u32(smem_raw)));   // two transposition tables alias the resident tile
        uint32_t accphase = 0;
        const int g8n = p.c1 >> 3;                                  // 8-channel groups of a layer-1 row
        // row metadata of the NEXT tile is fetched while the current one computes (registers of threads 0..127)
        int4 hdr = make_int4(0, 0, 0, 0);
        int nri = -1, nrj = -1;
        auto fetch_hdr = [&](int t) {
            if (FORM == FORM_DENSE) {
                const int q0 = t * p.gpt, nseg = max(0, min(p.gpt, p.n - q0));
                return make_int4(q0, nseg, 0, nseg * p.G);
            }
            if (FORM == FORM_ROWS) return make_int4(t * kRows, 0, 0, max(0, min(kRows, p.M - t * kRows)));
            return t < p.n_tiles ? __ldg(p.tiles + t) : make_int4(0, 0, 0, 0);
        };
        auto fetch_rows = [&](const int4& h) {
            nri = -1; nrj = -1;
            if (FORM == FORM_ROWS) return;
            if (wt < min(h.w, kRows)) {
                if (FORM == FORM_DENSE) {
                    const int sg = wt / p.G, g = wt - sg * p.G;
                    nri = h.x + sg;
                    nrj = g == 0 ? nri : __ldg(p.idx + (int64_t)nri * p.Knb + g - 1);
                } else {
                    nri = __ldg(p.erow + h.z + wt);
                    nrj = __ldg(p.ecol + h.z + wt);
                }
            }
        };
        hdr = fetch_hdr(blockIdx.x);
        fetch_rows(hdr);
        for (int t = blockIdx.x; t < p.n_tiles; t += gridDim.x) {
            // ---- tile header: which rows, which segments
            const int q0 = hdr.x, nseg = hdr.y, e0 = hdr.z, nrows = min(hdr.w, kRows);
            if (wt < kRows) {
                row_i[wt] = nri;
                row_j[wt] = nrj;
                if (FORM == FORM_POINT && nri >= 0) {
                    float rx = __fsub_rn(__ldg(p.pos_src + 3 * (int64_t)nrj), __ldg(p.pos_dst + 3 * (int64_t)nri));
                    float ry = __fsub_rn(__ldg(p.pos_src + 3 * (int64_t)nrj + 1), __ldg(p.pos_dst + 3 * (int64_t)nri + 1));
                    float rz = __fsub_rn(__ldg(p.pos_src + 3 * (int64_t)nrj + 2), __ldg(p.pos_dst + 3 * (int64_t)nri + 2));
                    if (p.radius > 0.f) { rx = __fdiv_rn(rx, p.radius); ry = __fdiv_rn(ry, p.radius); rz = __fdiv_rn(rz, p.radius); }
                    row_rel[wt][0] = rx; row_rel[wt][1] = ry; row_rel[wt][2] = rz;
                }
            }
            if (FORM != FORM_DENSE && nseg < kSegTable)
                for (int sg = wt; sg <= nseg; sg += kWorkers) seg_s[sg] = __ldg(p.row_ptr + q0 + sg) - e0;
            workers_sync();
            hdr = fetch_hdr(t + gridDim.x);                         // next tile's header: in flight during the formation

            if (FORM == FORM_ROWS) {
                // per-point MLP chain: the input planes ARE the A operand - 16-byte pieces copied into the swizzled tile
                for (int sgm = 0; sgm < p.nsegs; ++sgm) {
                    const int nchunks = (p.seg_k[sgm] + 63) >> 6, pieces = nchunks * 8, have = (int)(p.seg_ld[sgm] >> 3);
                    const __nv_bfloat16* sp = p.seg_p[sgm];
                    for (int base = wt; base < kRows * pieces; base += kWorkers * 4) {
                        uint4 hv[4], lv[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int item = base + u * kWorkers;
                            const int row = item / pieces, g8 = item - row * pieces;
                            hv[u] = make_uint4(0, 0, 0, 0); lv[u] = hv[u];
                            if (item < kRows * pieces && row < nrows && g8 < have) {      // pad columns of the planes are zero
                                const __nv_bfloat16* src = sp + (int64_t)(q0 + row) * p.seg_ld[sgm] + 8 * g8;
                                hv[u] = __ldg(reinterpret_cast<const uint4*>(src));
                                if (P == 2) lv[u] = __ldg(reinterpret_cast<const uint4*>(src + p.seg_ps[sgm]));
                            }
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int item = base + u * kWorkers;
                            if (item >= kRows * pieces) break;
                            const int row = item / pieces, g8 = item - row * pieces;
                            const uint32_t a = res0 + (uint32_t)((p.seg_c0[sgm] + (g8 >> 3)) * P) * kChunkBytes + row * 128 +
                                               (((g8 & 7) ^ (row & 7)) << 4);
                            st_shared_v4(a, hv[u].x, hv[u].y, hv[u].z, hv[u].w);
                            if (P == 2) st_shared_v4(a + kChunkBytes, lv[u].x, lv[u].y, lv[u].z, lv[u].w);
                        }
                    }
                }
            }
            // ---- layer 1 (decomposed): rows formed straight into the A tile, kU items (row, 8 channels) per thread in flight
            for (int base = wt; FORM != FORM_ROWS && base < kRows * g8n; base += kWorkers * kU) {
                float4 va[kU], vb[kU], vc[kU], vd[kU];
                int rowv[kU], g8v[kU], iv[kU], jv[kU];
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    const int item = base + u * kWorkers;
                    const int row = item < kRows * g8n ? item / g8n : 0;
                    rowv[u] = row;
                    g8v[u] = item - row * g8n;
                    iv[u] = item < kRows * g8n ? row_i[row] : -1;
                    jv[u] = row_j[row];
                }
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    if (FORM == FORM_DENSE || FORM == FORM_PAIR) {
                        const float4* qrow = reinterpret_cast<const float4*>(p.PQ + (int64_t)max(iv[u], 0) * p.ldpq + p.c1 + 8 * g8v[u]);
                        const float4* prow = reinterpret_cast<const float4*>(p.PQ + (int64_t)max(jv[u], 0) * p.ldpq + 8 * g8v[u]);
                        va[u] = __ldg(qrow); vb[u] = __ldg(qrow + 1);
                        vc[u] = __ldg(prow); vd[u] = __ldg(prow + 1);
                    } else if (p.H) {
                        const float4* hrow = reinterpret_cast<const float4*>(p.H + (int64_t)max(jv[u], 0) * p.ldh + 8 * g8v[u]);
                        vc[u] = __ldg(hrow); vd[u] = __ldg(hrow + 1);
                    }
                }
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    if (base + u * kWorkers >= kRows * g8n) break;
                    const int row = rowv[u], g8 = g8v[u];
                    float y[8];
                    if (iv[u] < 0) {
#pragma unroll
                        for (int e = 0; e < 8; ++e) y[e] = 0.f;
                    } else {
                        if (FORM == FORM_DENSE || FORM == FORM_PAIR) {
                            y[0] = va[u].x; y[1] = va[u].y; y[2] = va[u].z; y[3] = va[u].w;
                            y[4] = vb[u].x; y[5] = vb[u].y; y[6] = vb[u].z; y[7] = vb[u].w;
                            if (jv[u] >= 0) {
                                y[0] += vc[u].x; y[1] += vc[u].y; y[2] += vc[u].z; y[3] += vc[u].w;
                                y[4] += vd[u].x; y[5] += vd[u].y; y[6] += vd[u].z; y[7] += vd[u].w;
                            }
                        } else {
                            const float rx = row_rel[row][0], ry = row_rel[row][1], rz = row_rel[row][2];
#pragma unroll
                            for (int e = 0; e < 8; ++e) y[e] = wp_s[3][8 * g8 + e];
                            if (p.H) {
                                y[0] += vc[u].x; y[1] += vc[u].y; y[2] += vc[u].z; y[3] += vc[u].w;
                                y[4] += vd[u].x; y[5] += vd[u].y; y[6] += vd[u].z; y[7] += vd[u].w;
                            }
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                float v = fmaf(wp_s[0][8 * g8 + e], rx, y[e]);
                                v = fmaf(wp_s[1][8 * g8 + e], ry, v);
                                y[e] = fmaf(wp_s[2][8 * g8 + e], rz, v);
                            }
                        }
#pragma unroll
                        for (int e = 0; e < 8; ++e) y[e] = act_apply(y[e], p.act0, p.slope0);
                    }
                    const int kc = g8 >> 3, piece = g8 & 7;
                    store_piece<P>(res0 + (uint32_t)(kc * P) * kChunkBytes + row * 128 + ((piece ^ (row & 7)) << 4), y);
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic-proxy writes -> async proxy (UMMA)
            __syncwarp();
            if (lane == 0) mbar_arrive(a_ready);
            const bool row_masked = FORM == FORM_DENSE && myrow < nrows && row_j[myrow] < 0;   // missing neighbour: -100 (dgcnn.py:187-189)
            fetch_rows(hdr);                                        // next tile's rows: in flight during the MMAs

            for (int l = 0; l < p.L; ++l) {
                mbar_wait(acc_full, accphase);
                accphase ^= 1;
                tc_fence_after();
                const int N = p.ly[l].N, act = p.ly[l].act;
                const float slope = p.ly[l].slope;
                const uint32_t tacc = tmem_base + lane_base + ((att_mode && l == p.msg_layer) ? 0u : chain_col);
                if (l + 1 < p.L) {
                    // ---- epilogue: accumulator -> bias + activation -> bf16 planes, in place, as the next layer's A operand
                    for (int c0 = 32 * half; c0 < N; c0 += 64) {
                        uint32_t v[32];
                        tmem_ld32(tacc + (uint32_t)c0, v);
#pragma unroll
                        for (int pc = 0; pc < 4; ++pc) {
                            float y[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e)
                                y[e] = act_apply(__uint_as_float(v[8 * pc + e]) + bias_s[l][c0 + 8 * pc + e], act, slope);
                            const int kc = c0 >> 6, piece = ((c0 & 63) >> 3) + pc;
                            store_piece<P>(res0 + (uint32_t)(kc * P) * kChunkBytes + myrow * 128 + ((piece ^ (myrow & 7)) << 4), y);
                        }
                    }
                    tc_fence_before();
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(a_ready);
                    continue;
                }
                if (FORM == FORM_ROWS) {
                    // ---- per-point chain: the last layer's rows are the result (fp32 rows and / or activation planes)
                    const int64_t orow = (int64_t)q0 + myrow;
                    const int ncols = p.planes ? (int)p.kpad : N;                    // plane rows are written to their padded width (zeros)
                    for (int c0 = 32 * half; c0 < ncols; c0 += 64) {
                        uint32_t v[32];
                        tmem_ld32(tacc + (uint32_t)c0, v);
                        if (myrow >= nrows) continue;
                        float y[32];
#pragma unroll
                        for (int e = 0; e < 32; ++e) y[e] = act_apply(__uint_as_float(v[e]) + bias_s[l][c0 + e], act, slope);
                        if (p.out) {
                            float* yrow = p.out + orow * p.ldo + c0;
                            if (c0 + 32 <= N && (p.ldo & 3) == 0) {
#pragma unroll
                                for (int e = 0; e < 32; e += 4) *reinterpret_cast<float4*>(yrow + e) = make_float4(y[e], y[e + 1], y[e + 2], y[e + 3]);
                            } else {
#pragma unroll
                                for (int e = 0; e < 32; ++e)
                                    if (c0 + e < N) yrow[e] = y[e];
                            }
                        }
                        if (p.planes) {
                            __nv_bfloat16* hrow = p.planes + orow * p.kpad + c0;
#pragma unroll
                            for (int e = 0; e < 32; e += 8) {
                                if (c0 + e >= (int)p.kpad) break;                    // kpad % 16 == 0: whole 8-column pieces
                                uint32_t h[4], lo[4];
#pragma unroll
                                for (int k2 = 0; k2 < 4; ++k2) {
                                    const __nv_bfloat162 hv2 = __floats2bfloat162_rn(y[e + 2 * k2], y[e + 2 * k2 + 1]);
                                    h[k2] = *reinterpret_cast<const uint32_t*>(&hv2);
                                    const float2 f = __bfloat1622float2(hv2);
                                    const __nv_bfloat162 lv2 = __floats2bfloat162_rn(y[e + 2 * k2] - f.x, y[e + 2 * k2 + 1] - f.y);
                                    lo[k2] = *reinterpret_cast<const uint32_t*>(&lv2);
                                }
                                *reinterpret_cast<uint4*>(hrow + e) = make_uint4(h[0], h[1], h[2], h[3]);
                                if (p.npl == 2) *reinterpret_cast<uint4*>(hrow + p.pstride + e) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                            }
                        }
                    }
                    tc_fence_before();
                    workers_sync();               // the tile (shared row tables, TMEM) is free for the next one
                    continue;
                }
                // ---- last layer: aggregate every segment of the tile in edge order.  Per pass two 128 x 32 tables: the two
                // warp halves load columns [c0, c0+32) and [c0+32, c0+64) (max / mean / sum), or the attention logits and the
                // message of the same 32 columns (attention modes)
                const uint32_t tmsg = tmem_base + lane_base;
                const int ml = att_mode ? p.msg_layer : l;
                const int swz = myrow & 31;
                for (int c0 = 0; c0 < N; c0 += att_mode ? 32 : 64) {
                    uint32_t v[32];
                    float* Tw = T + half * kTable + myrow * 32;
                    if (!att_mode || half == 0) {
                        const int cc = c0 + (att_mode ? 0 : 32 * half);
                        tmem_ld32(tacc + (uint32_t)cc, v);
#pragma unroll
                        for (int e = 0; e < 32; ++e) {
                            float y = act_apply(__uint_as_float(v[e]) + bias_s[l][cc + e], act, slope);
                            if (row_masked) y = -100.f;
                            Tw[e ^ swz] = y;
                        }
                    } else {
                        tmem_ld32(tmsg + (uint32_t)c0, v);
#pragma unroll
                        for (int e = 0; e < 32; ++e)
                            Tw[e ^ swz] = act_apply(__uint_as_float(v[e]) + bias_s[ml][c0 + e], p.ly[ml].act, p.ly[ml].slope);
                    }
                    workers_sync();
                    const int c = wt & 31;
                    const int tab = att_mode ? 0 : (wt >> 7);
                    const int s0 = att_mode ? (wt >> 5) : ((wt >> 5) & 3), sstep = att_mode ? 8 : 4;
                    float* Ta = T + tab * kTable;                  // last layer's rows (attention logits in the attention modes)
                    const float* Tm = T + kTable;                  // message rows (attention modes)
                    const int col = c0 + 32 * tab + c;
                    for (int sg = s0; sg < nseg; sg += sstep) {
                        int lo, hi;
                        if (FORM == FORM_DENSE) { lo = sg * p.G; hi = lo + p.G; }
                        else if (nseg < kSegTable) { lo = seg_s[sg]; hi = seg_s[sg + 1]; }
                        else { lo = __ldg(p.row_ptr + q0 + sg) - e0; hi = __ldg(p.row_ptr + q0 + sg + 1) - e0; }
                        float r = 0.f;
                        if (p.agg == CCN_AGG_MAX) {
                            if (hi > lo) {
                                r = Ta[lo * 32 + (c ^ (lo & 31))];
#pragma unroll 4
                                for (int j = lo + 1; j < hi; ++j) r = fmaxf(r, Ta[j * 32 + (c ^ (j & 31))]);
                            }
                        } else if (p.agg == CCN_AGG_MEAN || p.agg == CCN_AGG_SUM) {
                            for (int j = lo; j < hi; ++j) r += Ta[j * 32 + (c ^ (j & 31))];
                            if (p.agg == CCN_AGG_MEAN) r = r / fmaxf((float)(hi - lo), 1.f);
                        } else if (p.agg == CCN_AGG_WSUM) {
                            for (int j = lo; j < hi; ++j)
                                r += Tm[j * 32 + (c ^ (j & 31))] * (1.f / (1.f + expf(-Ta[j * 32 + (c ^ (j & 31))])));
                        } else {
                            float m = -INFINITY;
                            for (int j = lo; j < hi; ++j) m = fmaxf(m, Ta[j * 32 + (c ^ (j & 31))]);
                            float sum = 0.f;
                            for (int j = lo; j < hi; ++j) {          // exp(a - max) once per element, kept in the table (own column)
                                const float ev = expf(Ta[j * 32 + (c ^ (j & 31))] - m);
                                Ta[j * 32 + (c ^ (j & 31))] = ev;
                                sum += ev;
                            }
                            const float inv = 1.f / (sum + 1e-16f);
                            for (int j = lo; j < hi; ++j) r += Tm[j * 32 + (c ^ (j & 31))] * (Ta[j * 32 + (c ^ (j & 31))] * inv);
                        }
                        const int64_t orow = q0 + sg;
                        p.out[orow * p.ldo + col] = r;
                        if (p.planes) {
                            const __nv_bfloat16 h = __float2bfloat16_rn(r);
                            p.planes[orow * p.kpad + col] = h;
                            if (p.npl == 2) p.planes[p.pstride + orow * p.kpad + col] = __float2bfloat16_rn(r - __bfloat162float(h));
                        }
                    }
                    workers_sync();
                }
                tc_fence_before();
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

// CSR tiling: tile t owns the targets whose first edge lies in [S t, S (t+1)) - at most S - 1 + max_degree = 128 edge rows,
// whole segments only; targets without edges belong to the tile of their offset.  One thread per tile, two binary searches.
__global__ void csr_tiles_kernel(const int32_t* __restrict__ row_ptr, int nq, int S, int n_tiles, int4* __restrict__ tiles) {
    ccn_pdl_wait();
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tiles) return;
    const int lo_v = S * t, hi_v = lo_v + S;
    int a = 0, b = nq + 1;
    while (a < b) { const int m = (a + b) >> 1; if (__ldg(row_ptr + m) < lo_v) a = m + 1; else b = m; }
    const int qa = min(a, nq);
    b = nq + 1;
    while (a < b) { const int m = (a + b) >> 1; if (__ldg(row_ptr + m) < hi_v) a = m + 1; else b = m; }
    const int qb = min(a, nq);
    const int e0 = __ldg(row_ptr + qa);
    tiles[t] = make_int4(qa, qb - qa, e0, __ldg(row_ptr + qb) - e0);
}

// ---------------------------------------------------------------------------------------------------------------------
// Dense SGCNN step whose `nn` has ONE layer after the decomposed first (levels 2..5 of the driving configs: 128 -> 128 ...
// 1024 -> 1024): a pipelined GEMM whose A operand is FORMED on the fly and whose epilogue is the masked group max.
//   warps 2-5   form act0(P[j] + Q[i]) for one 64-channel chunk of the tile's 128 rows straight into the ring stage
//               (SWIZZLE_128B bf16 hi/lo planes), running ahead of the MMAs by the depth of the ring;
//   warp 0      TMA: the matching W chunk (BN output rows x 64 k) into the same stage;
//   warp 1      tcgen05.mma into one of TWO TMEM accumulators, so that
//   warps 6-9   the epilogue of tile t (bias, -100 on missing neighbours, max over every group of K+1 rows through a
//               128 x 32 transposition table) overlaps the formation and MMAs of tile t+1.
// A CTA owns one (row tile, BN-column slice) at a time (persistent over the tile list), so wide layers spread over many CTAs.
// Same arithmetic and order as the unfused chain (ccn_sgcnn_layer1_planes -> ccn_linear_planes -> ccn_group_max_dense).
constexpr int kStreamThreads = 320;
constexpr int kMaskSlots = 16;               // tiles a former may run ahead of the epilogue (ring depth / chunks per tile + 2 accumulators)

struct StreamParams {
    const float* PQ;
    int64_t ldpq;
    const int32_t* idx;
    int n, Knb, G, gpt;
    int c1, act0;
    float slope0;
    int N, act;                    // the layer: c1 -> N
    float slope;
    const float* bias;
    int BN, n_slices, tiles_m, stages;
    int w_plane_bytes;             // BN * 128
    float* out;
    int64_t ldo;
    __nv_bfloat16* planes;
    int64_t kpad, pstride;
    int npl;
};

template <int P>
__global__ void __launch_bounds__(kStreamThreads, 1) sgcnn_stream_kernel(const __grid_constant__ CUtensorMap w_hi,
                                                                         const __grid_constant__ CUtensorMap w_lo,
                                                                         const __grid_constant__ StreamParams p) {
    ccn_pdl_wait();
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * kMaxStages + 4];     // full[S], empty[S], tfull[2], tempty[2]
    __shared__ uint32_t tmem_base_slot;
    __shared__ float bias_s[256];
    __shared__ int32_t row_i[2][kRows], row_j[2][kRows];           // per-tile row metadata of the formers (double buffered)
    __shared__ uint8_t row_mask[kMaskSlots][kRows];                 // 1: missing neighbour (-100 in the max), per tile in flight
    __shared__ __align__(8) uint64_t gbars[2];                      // row gathers of the two staging buffers

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int stages = p.stages, nk = p.c1 >> 6, BN = p.BN;
    const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const int a_bytes = P * kChunkBytes, stage_bytes = a_bytes + P * p.w_plane_bytes;
    const uint32_t table0 = smem0 + (uint32_t)(stages * stage_bytes);              // epilogue's transposition table
    const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[kMaxStages]);
    const uint32_t tfull0 = smem_u32(&bars[2 * kMaxStages]), tempty0 = smem_u32(&bars[2 * kMaxStages + 2]);
    uint32_t tmem_cols = 32;
    while ((int)tmem_cols < 2 * BN) tmem_cols <<= 1;
    const int total = p.tiles_m * p.n_slices;

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(full0 + 8 * s, 5);           // 4 former warps + the TMA thread's expect_tx arrive
            mbar_init(empty0 + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(tfull0 + 8 * b, 1);
            mbar_init(tempty0 + 8 * b, 4);
            mbar_init(smem_u32(&gbars[b]), 128);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(smem_u32(&tmem_base_slot), tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA: weight chunks
        if (lane == 0) {
            uint32_t cnt = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                const int n0 = (t % p.n_slices) * BN;
                const int mma_n = min(BN, p.N - n0);
                for (int kc = 0; kc < nk; ++kc, ++cnt) {
                    const int s = cnt % stages;
                    mbar_wait(empty0 + 8 * s, ((cnt / stages) & 1) ^ 1);
                    const uint32_t dst = smem0 + s * stage_bytes + a_bytes;
                    mbar_arrive_expect_tx(full0 + 8 * s, (uint32_t)(P * mma_n * 128));
                    for (int b = 0; b < mma_n; b += 64) {
                        tma_load_2d(dst + b * 128, &w_hi, kc * 64, n0 + b, full0 + 8 * s);
                        if (P == 2) tma_load_2d(dst + p.w_plane_bytes + b * 128, &w_lo, kc * 64, n0 + b, full0 + 8 * s);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            uint32_t cnt = 0;
            int it = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
                const int n0 = (t % p.n_slices) * BN;
                const int mma_n = min(BN, p.N - n0);
                const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(mma_n >> 3) << 17) | ((uint32_t)(kRows >> 4) << 24);
                const int buf = it & 1;
                mbar_wait(tempty0 + 8 * buf, ((it >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t acc = tmem_base + (uint32_t)(buf * BN);
                for (int kc = 0; kc < nk; ++kc, ++cnt) {
                    const int s = cnt % stages;
                    mbar_wait(full0 + 8 * s, (cnt / stages) & 1);
                    tc_fence_after();
                    const uint32_t a_hi = smem0 + s * stage_bytes, a_lo = a_hi + kChunkBytes;
                    const uint32_t b_hi = a_hi + a_bytes, b_lo = b_hi + p.w_plane_bytes;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t ko = k * 32;
                        umma_bf16(acc, umma_desc_sw128(a_hi + ko), umma_desc_sw128(b_hi + ko), idesc, (kc | k) ? 1u : 0u);
                        if (P == 2) {
                            umma_bf16(acc, umma_desc_sw128(a_lo + ko), umma_desc_sw128(b_hi + ko), idesc, 1u);
                            umma_bf16(acc, umma_desc_sw128(a_hi + ko), umma_desc_sw128(b_lo + ko), idesc, 1u);
                        }
                    }
                    umma_commit(empty0 + 8 * s);
                }
                umma_commit(tfull0 + 8 * buf);
            }
        }
    } else if (warp < 6) {
        // ------------------------------------------------------------------ formers (warps 2..5)
        // The P rows (one per tile row: the neighbour's product) and the Q rows (one per group: the point's own product) of a
        // 64-channel chunk are GATHERED by the TMA engine (one 256-byte cp.async.bulk per row, issued one chunk ahead into a
        // double-buffered fp32 staging area), so the formation itself only converts shared memory to shared memory.
        const int ft = threadIdx.x - 64;                            // 0..127
        const uint32_t stageP0 = table0 + kTable * 4;               // [2][128 rows][64 f32]
        const uint32_t stageQ0 = stageP0 + 2 * kRows * 256;         // [2][gpt rows][64 f32]
        const uint32_t gbar0 = smem_u32(&gbars[0]);
        // sequence of (tile, chunk) steps of this CTA
        const int my_tiles = blockIdx.x < total ? (total - 1 - blockIdx.x) / gridDim.x + 1 : 0;
        const int n_steps = my_tiles * nk;
        int ri_cur = -1, rj_cur = -1, ri_nxt = -1, rj_nxt = -1;
        auto meta_of = [&](int tile_iter, int& ri, int& rj) {
            const int t = blockIdx.x + tile_iter * gridDim.x;
            ri = -1; rj = -1;
            if (tile_iter >= my_tiles) return;
            const int q0 = (t / p.n_slices) * p.gpt, nseg = max(0, min(p.gpt, p.n - q0));
            if (ft < nseg * p.G) {
                const int sg = ft / p.G, g = ft - sg * p.G;
                ri = q0 + sg;
                rj = g == 0 ? ri : __ldg(p.idx + (int64_t)ri * p.Knb + g - 1);
            }
        };
        auto issue = [&](int step, int ri, int rj) {                // gathers of step -> staging buffer step & 1
            const int kc = step % nk, b = step & 1;
            const int t = blockIdx.x + (step / nk) * gridDim.x;
            const int q0 = (t / p.n_slices) * p.gpt, nseg = max(0, min(p.gpt, p.n - q0));
            uint32_t bytes = 0;
            if (rj >= 0) bytes += 256;
            if (ft < nseg) bytes += 256;
            mbar_arrive_tx(gbar0 + 8 * b, bytes);
            if (rj >= 0) bulk_g2s(stageP0 + (uint32_t)(b * kRows + ft) * 256, p.PQ + (int64_t)rj * p.ldpq + kc * 64, 256, gbar0 + 8 * b);
            if (ft < nseg)
                bulk_g2s(stageQ0 + (uint32_t)(b * p.gpt + ft) * 256, p.PQ + (int64_t)(q0 + ft) * p.ldpq + p.c1 + kc * 64, 256, gbar0 + 8 * b);
        };
        meta_of(0, ri_cur, rj_cur);
        if (n_steps > 0) issue(0, ri_cur, rj_cur);
        uint32_t cnt = 0;
        for (int step = 0; step < n_steps; ++step, ++cnt) {
            const int it = step / nk, kc = step - it * nk, b = step & 1, mb = it & 1;
            if (kc == 0) {
                row_i[mb][ft] = ri_cur;
                row_j[mb][ft] = rj_cur;
                row_mask[it & (kMaskSlots - 1)][ft] = (ri_cur >= 0 && rj_cur < 0) ? 1 : 0;
                meta_of(it + 1, ri_nxt, rj_nxt);                    // next tile's neighbour ids: in flight during this tile
            }
            asm volatile("bar.sync 2, 128;" ::: "memory");          // metadata visible; everyone has converted step - 1 (buffer b ^ 1 free)
            if (step + 1 < n_steps) {
                const bool next_tile = kc + 1 == nk;
                issue(step + 1, next_tile ? ri_nxt : ri_cur, next_tile ? rj_nxt : rj_cur);
            }
            mbar_wait(gbar0 + 8 * b, (step >> 1) & 1);              // this step's rows have landed
            const int s = cnt % stages;
            mbar_wait(empty0 + 8 * s, ((cnt / stages) & 1) ^ 1);
            const uint32_t a0 = smem0 + s * stage_bytes;
#pragma unroll 2
            for (int item = ft; item < kRows * 8; item += 128) {
                const int row = item >> 3, g8 = item & 7;
                const int i = row_i[mb][row], j = row_j[mb][row];
                float y[8];
                if (i < 0) {
#pragma unroll
                    for (int e = 0; e < 8; ++e) y[e] = 0.f;
                } else {
                    const int sg = row / p.G;
                    const uint32_t qa = stageQ0 + (uint32_t)(b * p.gpt + sg) * 256 + g8 * 32;
                    const float4 a = lds128(qa), c = lds128(qa + 16);
                    y[0] = a.x; y[1] = a.y; y[2] = a.z; y[3] = a.w; y[4] = c.x; y[5] = c.y; y[6] = c.z; y[7] = c.w;
                    if (j >= 0) {
                        const uint32_t pa = stageP0 + (uint32_t)(b * kRows + row) * 256 + g8 * 32;
                        const float4 d = lds128(pa), f = lds128(pa + 16);
                        y[0] += d.x; y[1] += d.y; y[2] += d.z; y[3] += d.w; y[4] += f.x; y[5] += f.y; y[6] += f.z; y[7] += f.w;
                    }
#pragma unroll
                    for (int e = 0; e < 8; ++e) y[e] = act_apply(y[e], p.act0, p.slope0);
                }
                store_piece<P>(a0 + row * 128 + ((g8 ^ (row & 7)) << 4), y);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(full0 + 8 * s);
            if (kc + 1 == nk) { ri_cur = ri_nxt; rj_cur = rj_nxt; }
        }
    } else {
        // ------------------------------------------------------------------ epilogue (warps 6..9)
        const int et = threadIdx.x - 192;                           // 0..127
        const int q = warp & 3;
        const int myrow = q * 32 + lane;
        const uint32_t lane_base = (uint32_t)(q * 32) << 16;
        float* T = reinterpret_cast<float*>(smem_raw + (table0 - smem_u32(smem_raw)));
        const int swz = myrow & 31;
        int it = 0, bias_n0 = -1;
        for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
            const int m_tile = t / p.n_slices, n0 = (t % p.n_slices) * BN;
            const int mma_n = min(BN, p.N - n0);
            const int q0 = m_tile * p.gpt, nseg = max(0, min(p.gpt, p.n - q0));
            const int buf = it & 1;
            if (n0 != bias_n0) {
                asm volatile("bar.sync 3, 128;" ::: "memory");
                for (int j = et; j < BN; j += 128) bias_s[j] = (p.bias && n0 + j < p.N) ? __ldg(p.bias + n0 + j) : 0.f;
                asm volatile("bar.sync 3, 128;" ::: "memory");
                bias_n0 = n0;
            }
            mbar_wait(tfull0 + 8 * buf, (it >> 1) & 1);
            tc_fence_after();
            const bool masked = row_mask[it & (kMaskSlots - 1)][myrow] != 0;
            const uint32_t tacc = tmem_base + lane_base + (uint32_t)(buf * BN);
            for (int c0 = 0; c0 < mma_n; c0 += 32) {
                uint32_t v[32];
                tmem_ld32(tacc + (uint32_t)c0, v);
#pragma unroll
                for (int e = 0; e < 32; ++e) {
                    float y = act_apply(__uint_as_float(v[e]) + bias_s[c0 + e], p.act, p.slope);
                    if (masked) y = -100.f;
                    T[myrow * 32 + (e ^ swz)] = y;
                }
                asm volatile("bar.sync 3, 128;" ::: "memory");
                const int c = et & 31;
                for (int sg = et >> 5; sg < nseg; sg += 4) {
                    const int lo = sg * p.G, hi = lo + p.G;
                    float r = T[lo * 32 + (c ^ (lo & 31))];
#pragma unroll 4
                    for (int j = lo + 1; j < hi; ++j) r = fmaxf(r, T[j * 32 + (c ^ (j & 31))]);
                    const int64_t orow = q0 + sg;
                    const int col = n0 + c0 + c;
                    p.out[orow * p.ldo + col] = r;
                    if (p.planes) {
                        const __nv_bfloat16 h = __float2bfloat16_rn(r);
                        p.planes[orow * p.kpad + col] = h;
                        if (p.npl == 2) p.planes[p.pstride + orow * p.kpad + col] = __float2bfloat16_rn(r - __bfloat162float(h));
                    }
                }
                asm volatile("bar.sync 3, 128;" ::: "memory");
            }
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

// ---- host side -------------------------------------------------------------------------------------------------
constexpr int kSmemLimit = 227 * 1024;
constexpr int kStaticSmem = 14 * 1024;          // bias / Wp / row tables / barriers (ptxas reports ~12.8 KB)
constexpr int kHalfSm = 113 * 1024;             // two CTAs per SM: 228 KB / 2 minus the per-CTA reservation

int fill_chain(ChainParams& p, const ccn_chain_t* ch, int c1, bool att, int& smem_bytes, int& P, bool rows_form = false) {
    CCN_REQUIRE(ch && ch->n_layers >= 1 && ch->n_layers <= CCN_CHAIN_MAX_LAYERS && ch->w_hi_tmap);
    CCN_REQUIRE(ch->mode == 1 || (ch->mode == 3 && ch->w_lo_tmap));
    P = ch->mode == 3 ? 2 : 1;
    p.L = ch->n_layers;
    int width = c1, maxw = c1, maxn = 64;
    CCN_REQUIRE(c1 >= 64 && c1 <= kMaxWidth && (c1 & 63) == 0);
    for (int l = 0; l < p.L; ++l) {
        // the LAST layer may have any width (its rows end the stacked weight matrix: the TMA boxes past them arrive as zeros)
        const bool last_any = rows_form && l + 1 == p.L;
        CCN_REQUIRE(ch->k[l] == width && ch->n[l] >= (last_any ? 1 : 64) && ch->n[l] <= kMaxWidth && (last_any || (ch->n[l] & 63) == 0));
        CCN_REQUIRE(ch->act[l] >= 0 && ch->act[l] <= 2 && ch->w_row[l] >= 0 && (ch->w_row[l] & 63) == 0);
        p.ly[l].K = ch->k[l]; p.ly[l].N = ch->n[l]; p.ly[l].w_row = ch->w_row[l]; p.ly[l].act = ch->act[l];
        p.ly[l].slope = ch->slope[l]; p.ly[l].bias = ch->bias[l];
        width = ch->n[l];
        if (width > maxn) maxn = width;
        if (l + 1 < p.L) maxw = width > maxw ? width : maxw;
    }
    p.res_chunks = maxw / 64;
    int res_bytes = p.res_chunks * P * kChunkBytes;
    const int table_bytes = 2 * kTable * 4;                           // the aggregation's transposition tables alias the tile
    if (res_bytes < table_bytes) {
        p.res_chunks = (table_bytes + P * kChunkBytes - 1) / (P * kChunkBytes);
        res_bytes = p.res_chunks * P * kChunkBytes;
    }
    p.w_plane_bytes = (maxn < 128 ? ((maxn + 63) & ~63) : 128) * 128;
    const int stage = P * p.w_plane_bytes;
    int stages = (kSmemLimit - kStaticSmem - 1024 - res_bytes) / stage;
    if (stages > kMaxStages) stages = kMaxStages;
    if (stages < 2) return CCN_EUNSUPPORTED;
    // narrow chains: keep the footprint under half an SM so that two CTAs are co-resident (one's MMAs overlap the other's epilogue)
    if (!att && res_bytes + 3 * stage + 1024 + kStaticSmem <= kHalfSm) {
        const int fit = (kHalfSm - kStaticSmem - 1024 - res_bytes) / stage;
        if (stages > fit) stages = fit;
    }
    p.stages = stages;
    smem_bytes = res_bytes + stages * stage + 1024;
    return CCN_OK;
}

template <int P, int FORM>
int launch_chain(const ccn_chain_t* ch, const ChainParams& p, int smem, cudaStream_t s) {
    static bool attr = false;
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(edge_chain_kernel<P, FORM>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit - kStaticSmem);
        if (e != cudaSuccess) return (int)e;
        attr = true;
    }
    const int per_sm = smem + kStaticSmem <= kHalfSm && p.msg_layer < 0 ? 2 : 1;
    const int grid = p.n_tiles < 148 * per_sm ? p.n_tiles : 148 * per_sm;
    const CUtensorMap* whi = (const CUtensorMap*)ch->w_hi_tmap;
    const CUtensorMap* wlo = P == 2 ? (const CUtensorMap*)ch->w_lo_tmap : whi;
    ccn_launch(edge_chain_kernel<P, FORM>, dim3(grid), dim3(kThreads), smem, s, *whi, *wlo, p);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

template <int P>
int launch_stream_t(const StreamParams& p, const ccn_chain_t* ch, int smem, int grid, cudaStream_t s) {
    static bool attr = false;
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(sgcnn_stream_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit - 8 * 1024);
        if (e != cudaSuccess) return (int)e;
        attr = true;
    }
    const CUtensorMap* whi = (const CUtensorMap*)ch->w_hi_tmap;
    const CUtensorMap* wlo = P == 2 ? (const CUtensorMap*)ch->w_lo_tmap : whi;
    ccn_launch(sgcnn_stream_kernel<P>, dim3(grid), dim3(kStreamThreads), smem, s, *whi, *wlo, p);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

int launch_stream(const float* PQ, int64_t ldpq, const int32_t* idx, int64_t n, int32_t K, int32_t c1, int32_t act0, float slope0,
                  const ccn_chain_t* ch, float* out, int64_t ldo, void* planes, int64_t kpad, int64_t plane_stride, int32_t n_planes,
                  cudaStream_t s) {
    CCN_REQUIRE(ch->w_hi_tmap && (ch->mode == 1 || (ch->mode == 3 && ch->w_lo_tmap)));
    const int P = ch->mode == 3 ? 2 : 1;
    const int N = ch->n[0];
    CCN_REQUIRE(ch->k[0] == c1 && c1 >= 64 && c1 <= 1024 && (c1 & 63) == 0 && N >= 64 && N <= 1024 && (N & 63) == 0);
    CCN_REQUIRE(ch->act[0] >= 0 && ch->act[0] <= 2 && ch->w_row[0] == 0);
    CCN_REQUIRE(ldo >= N && (!planes || (kpad >= N && (n_planes == 1 || n_planes == 2))));
    CCN_REQUIRE((ldpq & 3) == 0 && (reinterpret_cast<uintptr_t>(PQ) & 15) == 0);       // 16-byte aligned rows for the bulk gathers
    if (n == 0) return CCN_OK;
    StreamParams p = {};
    p.PQ = PQ; p.ldpq = ldpq; p.idx = idx; p.n = (int)n; p.Knb = K; p.G = K + 1; p.gpt = kRows / (K + 1);
    p.c1 = c1; p.act0 = act0; p.slope0 = slope0;
    p.N = N; p.act = ch->act[0]; p.slope = ch->slope[0]; p.bias = ch->bias[0];
    p.BN = N < 128 ? N : 128;
    p.n_slices = (N + p.BN - 1) / p.BN;
    p.tiles_m = (int)((n + p.gpt - 1) / p.gpt);
    p.w_plane_bytes = p.BN * 128;
    const int stage = P * kChunkBytes + P * p.w_plane_bytes;
    const int table = kTable * 4 + 2 * kRows * 256 + 2 * p.gpt * 256;      // transposition table + fp32 row staging (P rows, Q rows)
    int stages = (kSmemLimit - 8 * 1024 - 1024 - table) / stage;
    if (stages > kMaxStages) stages = kMaxStages;
    if (stages < 2) return CCN_EUNSUPPORTED;
    p.stages = stages;
    p.out = out; p.ldo = ldo; p.planes = (__nv_bfloat16*)planes; p.kpad = kpad; p.pstride = plane_stride; p.npl = n_planes;
    const int smem = stages * stage + table + 1024;
    const int64_t total = (int64_t)p.tiles_m * p.n_slices;
    // a CTA keeps one column slice (tile t -> slice t % n_slices): the grid is a multiple of n_slices
    int grid = (int)(total < 148 ? total : (148 / p.n_slices) * p.n_slices);
    if (grid < 1) grid = 1;
    return P == 2 ? launch_stream_t<2>(p, ch, smem, grid, s) : launch_stream_t<1>(p, ch, smem, grid, s);
}

}  // namespace

extern "C" int ccn_sgcnn_fused(const float* PQ, int64_t ldpq, const int32_t* idx, int64_t n, int32_t K, int32_t c1,
                               int32_t act0, float slope0, const ccn_chain_t* chain_host, float* out, int64_t ldo,
                               void* planes, int64_t kpad, int64_t plane_stride, int32_t n_planes, ccn_stream_t stream) {
    CCN_REQUIRE(PQ && idx && out && n >= 0 && n < 0x7fffffff && K >= 1 && K + 1 <= kRows && act0 >= 0 && act0 <= 2);
    CCN_REQUIRE(ldpq >= 2 * c1 && (ldpq & 3) == 0 && (reinterpret_cast<uintptr_t>(PQ) & 15) == 0);
    if (chain_host && chain_host->n_layers == 1)
        return launch_stream(PQ, ldpq, idx, n, K, c1, act0, slope0, chain_host, out, ldo, planes, kpad, plane_stride, n_planes,
                             (cudaStream_t)stream);
    ChainParams p = {};
    int smem = 0, P = 1;
    const int rc = fill_chain(p, chain_host, c1, false, smem, P);
    if (rc != CCN_OK) return rc;
    const int nout = p.ly[p.L - 1].N;
    CCN_REQUIRE(ldo >= nout && (!planes || (kpad >= nout && (n_planes == 1 || n_planes == 2))));
    if (n == 0) return CCN_OK;
    p.c1 = c1; p.act0 = act0; p.slope0 = slope0;
    p.PQ = PQ; p.ldpq = ldpq; p.idx = idx; p.n = (int)n; p.Knb = K; p.G = K + 1; p.gpt = kRows / (K + 1);
    p.n_tiles = (int)((n + p.gpt - 1) / p.gpt);
    p.agg = CCN_AGG_MAX; p.msg_layer = -1;
    p.out = out; p.ldo = ldo; p.planes = (__nv_bfloat16*)planes; p.kpad = kpad; p.pstride = plane_stride; p.npl = n_planes;
    return P == 2 ? launch_chain<2, FORM_DENSE>(chain_host, p, smem, (cudaStream_t)stream)
                  : launch_chain<1, FORM_DENSE>(chain_host, p, smem, (cudaStream_t)stream);
}

extern "C" int ccn_pointconv_fused(const float* H, int64_t ldh, const float* wp, const float* bias0, const float* pos_src,
                                   const float* pos_dst, const float* PQ, int64_t ldpq, const int32_t* row_ptr,
                                   const int32_t* erow, const int32_t* ecol, int64_t nq, int64_t E, int32_t max_degree,
                                   int32_t c1, float radius, int32_t act0, float slope0, const ccn_chain_t* chain_host,
                                   int32_t agg, int32_t msg_layer, int32_t* tile_scratch, float* out, int64_t ldo,
                                   void* planes, int64_t kpad, int64_t plane_stride, int32_t n_planes, ccn_stream_t stream) {
    const bool pair = PQ != nullptr;
    CCN_REQUIRE(row_ptr && erow && ecol && out && nq >= 0 && nq < 0x7fffffff && E >= 0 && E < 0x7fffffff);
    CCN_REQUIRE(tile_scratch && (reinterpret_cast<uintptr_t>(tile_scratch) & 15) == 0);
    CCN_REQUIRE(max_degree >= 1 && max_degree <= kRows && act0 >= 0 && act0 <= 2 && agg >= 0 && agg <= 4);
    if (pair) CCN_REQUIRE(ldpq >= 2 * c1 && (ldpq & 3) == 0 && (reinterpret_cast<uintptr_t>(PQ) & 15) == 0);
    else CCN_REQUIRE(wp && pos_src && pos_dst && (!H || ((ldh & 3) == 0 && ldh >= c1 && (reinterpret_cast<uintptr_t>(H) & 15) == 0)));
    ChainParams p = {};
    int smem = 0, P = 1;
    const bool att = agg == CCN_AGG_WSUM || agg == CCN_AGG_ATTEND;
    const int rc = fill_chain(p, chain_host, c1, att, smem, P);
    if (rc != CCN_OK) return rc;
    CCN_REQUIRE(att ? (msg_layer >= 0 && msg_layer < p.L - 1 && p.ly[msg_layer].N == p.ly[p.L - 1].N) : msg_layer < 0);
    const int nout = p.ly[p.L - 1].N;
    CCN_REQUIRE(ldo >= nout && (!planes || (kpad >= nout && (n_planes == 1 || n_planes == 2))));
    if (nq == 0) return CCN_OK;
    p.c1 = c1; p.act0 = act0; p.slope0 = slope0;
    p.PQ = PQ; p.ldpq = ldpq; p.H = H; p.ldh = ldh; p.wp = wp; p.bias0 = bias0; p.pos_src = pos_src; p.pos_dst = pos_dst;
    p.radius = radius;
    p.row_ptr = row_ptr; p.erow = erow; p.ecol = ecol; p.nq = (int)nq; p.E = (int)E; p.S = kRows + 1 - max_degree;
    p.n_tiles = (int)(E / p.S) + 1;
    p.agg = agg; p.msg_layer = att ? msg_layer : -1;
    p.out = out; p.ldo = ldo; p.planes = (__nv_bfloat16*)planes; p.kpad = kpad; p.pstride = plane_stride; p.npl = n_planes;
    cudaStream_t s = (cudaStream_t)stream;
    p.tiles = reinterpret_cast<const int4*>(tile_scratch);
    ccn_launch(csr_tiles_kernel, dim3((p.n_tiles + 127) / 128), dim3(128), 0, s, row_ptr, p.nq, p.S, p.n_tiles, reinterpret_cast<int4*>(tile_scratch));
    CCN_CHECK_LAUNCH();
    if (pair) return P == 2 ? launch_chain<2, FORM_PAIR>(chain_host, p, smem, s) : launch_chain<1, FORM_PAIR>(chain_host, p, smem, s);
    return P == 2 ? launch_chain<2, FORM_POINT>(chain_host, p, smem, s) : launch_chain<1, FORM_POINT>(chain_host, p, smem, s);
}

// Per-point MLP chain: Y = layer_n(... layer_1(A)) for 2+ consecutive Linear (+ folded BatchNorm + activation) layers of a shared
// MLP (torch_geometric.nn.MLP, reference src/models/base.py:32, mlp.py:16-22, skip_connect.py:12-15) in ONE launch: the
// 128-row activation tile stays in shared memory between the layers (edge_chain_kernel, FORM_ROWS).  A: the input activations
// as split-bf16 planes (k = chain->k[0], a multiple of 64, <= 256); every layer width but the last a multiple of 64, all <= 256.
// Same MMA / bias / activation order as consecutive ccn_linear_planes launches: bit-identical to them.
extern "C" int ccn_mlp_chain(const void* a0, int64_t ld0, int64_t ps0, int32_t k0, const void* a1, int64_t ld1, int64_t ps1,
                             int32_t k1, const void* a2, int64_t ld2, int64_t ps2, int32_t k2, const void* a3, int64_t ld3,
                             int64_t ps3, int32_t k3, int64_t M, const ccn_chain_t* chain_host, float* Y, int64_t ldy,
                             void* y_planes, int64_t yp_ld, int64_t yp_plane_stride, ccn_stream_t stream) {
    const void* ap[4] = {a0, a1, a2, a3};
    const int64_t ald[4] = {ld0, ld1, ld2, ld3}, aps[4] = {ps0, ps1, ps2, ps3};
    const int32_t ak[4] = {k0, k1, k2, k3};
    CCN_REQUIRE(a0 && k0 > 0 && M >= 0 && M < 0x7fffffff && (Y || y_planes));
    ChainParams p = {};
    int c1 = 0;
    for (int i = 0; i < 4; ++i) {
        if (ak[i] <= 0) break;
        CCN_REQUIRE(ap[i] && ald[i] >= ((ak[i] + 7) & ~7) && (ald[i] & 7) == 0 && (reinterpret_cast<uintptr_t>(ap[i]) & 15) == 0);
        p.seg_p[i] = (const __nv_bfloat16*)ap[i]; p.seg_ld[i] = ald[i]; p.seg_ps[i] = aps[i]; p.seg_k[i] = ak[i];
        p.seg_c0[i] = c1 >> 6;
        c1 += (ak[i] + 63) & ~63;
        p.nsegs = i + 1;
    }
    int smem = 0, P = 1;
    const int rc = fill_chain(p, chain_host, c1, false, smem, P, true);
    if (rc != CCN_OK) return rc;
    const int nout = p.ly[p.L - 1].N;
    for (int i = 0; i < p.nsegs; ++i) CCN_REQUIRE(P == 1 || (aps[i] & 7) == 0);
    CCN_REQUIRE(!Y || ldy >= nout);
    CCN_REQUIRE(!y_planes || (yp_ld == ((nout + 15) & ~15) && (reinterpret_cast<uintptr_t>(y_planes) & 15) == 0 &&
                              (P == 1 || (yp_plane_stride & 7) == 0)));
    if (M == 0) return CCN_OK;
    p.c1 = c1; p.act0 = CCN_ACT_NONE; p.slope0 = 0.f;
    p.M = (int)M;
    p.n_tiles = (int)((M + kRows - 1) / kRows);
    p.agg = kAggNone; p.msg_layer = -1;
    p.out = Y; p.ldo = ldy; p.planes = (__nv_bfloat16*)y_planes;
    p.kpad = y_planes ? yp_ld : 0;                         // row stride = columns written per plane row (pad columns: zeros)
    p.pstride = yp_plane_stride; p.npl = P;
    cudaStream_t s = (cudaStream_t)stream;
    return P == 2 ? launch_chain<2, FORM_ROWS>(chain_host, p, smem, s) : launch_chain<1, FORM_ROWS>(chain_host, p, smem, s);
}
