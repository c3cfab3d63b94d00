// Sampling and grouping ALONG curves:
//   ccn_curve_fps            CurveFPS.forward            reference src/models/modules/fps_ops.py:16-39
//   ccn_radius_1d_group_*    radius_1d_group_subset      reference src/models/utils/point_ops.py:143-193
//   ccn_knn_1d_superset      knn_1d_group_superset       reference src/models/utils/point_ops.py:196-260
// Index decisions reproduce the reference's CPU float arithmetic operation by operation
// (oracle/ccn_point_ops.py header): linalg.norm = sqrt(fma(z,z,fma(y,y,x*x))), cumsum = f64
// running sum rounded per element, scalar/tensor = reciprocal * scalar, % = fmod + sign fix.
#include "common.cuh"

namespace {

constexpr int kItems = 4;

__device__ __forceinline__ float edge_to_prev(const float* __restrict__ pos, const int32_t* __restrict__ ord,
                                              int64_t i) {
    if (i <= 0 || ord[i] != ord[i - 1]) return 0.f;
    const float* a = pos + 3 * i;
    const float* b = a - 3;
    return ccn_norm3(__fsub_rn(a[0], b[0]), __fsub_rn(a[1], b[1]), __fsub_rn(a[2], b[2]));
}

__device__ __forceinline__ float curve_fps_bucket(const float* __restrict__ run, const int32_t* __restrict__ ord,
                                                  const int32_t* __restrict__ cptr, int64_t i, float u,
                                                  float arclen) {
    const int32_t st = cptr[ord[i]];
    float s = __fsub_rn(run[i], run[st]);
    // ((start * 117) * u) % arclen: int64 -> f32, f32 multiply, fmod with the sign of the divisor
    float ph = __fmul_rn(__ll2float_rn((long long)st * 117LL), u);
    float m = fmodf(ph, arclen);
    if (m != 0.f && ((arclen < 0.f) != (m < 0.f))) m = __fadd_rn(m, arclen);
    s = __fadd_rn(s, m);
    return rintf(__fdiv_rn(s, arclen));
}

// Four launches over 4096-point tiles (a single CTA serialised over the whole batch: 0.42 ms at 6 x 32 k points):
//   1. per tile: f64 sum of the edge lengths                       -> tile_tot[t]   (head of the idx_out buffer)
//   2. per tile: carry = tile_tot[0] + tile_tot[1] + ... in order, f64 scan of the tile -> run[i] = (float)prefix
//   3. per tile: number of kept points (bucket changes / curve starts)               -> tile_cnt[t] (tail of run_scratch)
//   4. per tile: ordered compaction of the kept indices into idx_out
// The f64 additions happen in exactly the order of the former one-CTA loop (chunk totals added one after the other, the
// same 1024 x 4 block scan inside a chunk), so run[] - and with it every index decision - is bit-identical to it.
constexpr int kTile = 1024 * kItems;

__global__ void __launch_bounds__(1024) cfps_tile_sum_kernel(const float* __restrict__ pos, const int32_t* __restrict__ ord,
                                                             int64_t n, double* __restrict__ tile_tot) {
    ccn_pdl_wait();
    __shared__ double smd[33];
    const int64_t i0 = (int64_t)blockIdx.x * kTile + (int64_t)threadIdx.x * kItems;
    double local = 0.0;
#pragma unroll
    for (int k = 0; k < kItems; ++k) local += (i0 + k < n) ? (double)edge_to_prev(pos, ord, i0 + k) : 0.0;
    double total;
    ccn_block_excl_scan<double>(local, smd, &total);
    if (threadIdx.x == 0) tile_tot[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024) cfps_run_kernel(const float* __restrict__ pos, const int32_t* __restrict__ ord,
                                                        int64_t n, const double* __restrict__ tile_tot,
                                                        float* __restrict__ run) {
    ccn_pdl_wait();
    __shared__ double smd[33];
    __shared__ double s_carry;
    if (threadIdx.x == 0) {
        double c = 0.0;
        for (int t = 0; t < (int)blockIdx.x; ++t) c += tile_tot[t];
        s_carry = c;
    }
    __syncthreads();
    const int64_t i0 = (int64_t)blockIdx.x * kTile + (int64_t)threadIdx.x * kItems;
    double e[kItems], local = 0.0;
#pragma unroll
    for (int k = 0; k < kItems; ++k) {
        e[k] = (i0 + k < n) ? (double)edge_to_prev(pos, ord, i0 + k) : 0.0;
        local += e[k];
    }
    double total;
    double acc = s_carry + ccn_block_excl_scan<double>(local, smd, &total);
#pragma unroll
    for (int k = 0; k < kItems; ++k) {
        acc += e[k];
        if (i0 + k < n) run[i0 + k] = (float)acc;
    }
}

// keep flags of the kItems points of this thread (a point is kept when its arclength bucket differs from its
// predecessor's or when it starts a curve)
__device__ __forceinline__ int cfps_flags(const float* __restrict__ run, const int32_t* __restrict__ ord,
                                          const int32_t* __restrict__ cptr, int64_t n, int64_t i0, float u, float arclen,
                                          int (&f)[kItems]) {
    int local = 0;
    float qprev = 0.f;
    if (i0 > 0 && i0 < n) qprev = curve_fps_bucket(run, ord, cptr, i0 - 1, u, arclen);
#pragma unroll
    for (int k = 0; k < kItems; ++k) {
        const int64_t i = i0 + k;
        f[k] = 0;
        if (i < n) {
            const float q = curve_fps_bucket(run, ord, cptr, i, u, arclen);
            f[k] = (i == 0) || (q != qprev) || (cptr[ord[i]] == (int32_t)i);
            qprev = q;
        }
        local += f[k];
    }
    return local;
}

__global__ void __launch_bounds__(1024) cfps_count_kernel(const float* __restrict__ run, const int32_t* __restrict__ ord,
                                                          const int32_t* __restrict__ cptr, int64_t n,
                                                          const float* __restrict__ u_ptr, float arclen,
                                                          int32_t* __restrict__ tile_cnt) {
    ccn_pdl_wait();
    __shared__ int smi[33];
    int f[kItems];
    const int local = cfps_flags(run, ord, cptr, n, (int64_t)blockIdx.x * kTile + (int64_t)threadIdx.x * kItems, *u_ptr, arclen, f);
    int total;
    ccn_block_excl_scan<int>(local, smi, &total);
    if (threadIdx.x == 0) tile_cnt[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024) cfps_compact_kernel(const float* __restrict__ run, const int32_t* __restrict__ ord,
                                                            const int32_t* __restrict__ cptr, int64_t n,
                                                            const float* __restrict__ u_ptr, float arclen,
                                                            const int32_t* __restrict__ tile_cnt,
                                                            int64_t* __restrict__ idx_out, int32_t* __restrict__ count_out) {
    ccn_pdl_wait();
    __shared__ int smi[33];
    int part = 0;
    for (int t = threadIdx.x; t < (int)blockIdx.x; t += blockDim.x) part += tile_cnt[t];
    int carry;
    ccn_block_excl_scan<int>(part, smi, &carry);
    int f[kItems];
    const int64_t i0 = (int64_t)blockIdx.x * kTile + (int64_t)threadIdx.x * kItems;
    const int local = cfps_flags(run, ord, cptr, n, i0, *u_ptr, arclen, f);
    int total;
    int o = carry + ccn_block_excl_scan<int>(local, smi, &total);
#pragma unroll
    for (int k = 0; k < kItems; ++k)
        if (f[k]) idx_out[o++] = i0 + k;
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) *count_out = carry + total;
}

// Per-curve statistics for radius_1d_group_subset (point_ops.py:149-167): one WARP per curve.  Lanes
// load 32 edge norms at a time; the f32 sum still runs strictly in index order (shuffle broadcast of
// lane 0, 1, 2, ...), as the oracle's scatter_add does.  stats[0] = max k_c, stats[1] = max #points (float bits).
__global__ void __launch_bounds__(256) radius_1d_curve_stats_kernel(
    const float* __restrict__ pos, const int32_t* __restrict__ ord, const int32_t* __restrict__ cptr,
    int32_t n_curves, float radius, float* __restrict__ kc, int32_t* __restrict__ stats) {
    ccn_pdl_wait();
    const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (c >= n_curves) return;
    const int lane = ccn_lane();
    const int32_t b = cptr[c], e = cptr[c + 1];
    float len = 0.f;
    for (int32_t i0 = b + 1; i0 < e; i0 += 32) {
        const int32_t i = i0 + lane;
        const float en = i < e ? edge_to_prev(pos, ord, i) : 0.f;
        const int m = min(32, e - i0);
        for (int l = 0; l < m; ++l) len = __fadd_rn(len, __shfl_sync(0xffffffffu, en, l));
    }
    if (lane != 0) return;
    const float cnt = (float)(e - b);
    const float avg = __fdiv_rn(len, cnt);
    float k = ceilf(__fmul_rn(__frcp_rn(avg), radius));
    if (isinf(k)) k = 1.f;
    kc[c] = k;
    if (!(k != k)) atomicMax(&stats[0], __float_as_int(fmaxf(k, 0.f)));
    atomicMax(&stats[1], __float_as_int(cnt));
}

// pass 0: count per query; pass 1: fill col[row_ptr[q] ...].  Candidate order 0,-1,+1,-2,+2,...
template <bool FILL>
__global__ void radius_1d_group_kernel(const int64_t* __restrict__ idxs, int64_t nq,
                                       const int32_t* __restrict__ ord, const int64_t* __restrict__ curve_local,
                                       int64_t n, int32_t n_curves, const float* __restrict__ kc,
                                       const int32_t* __restrict__ stats, int32_t* __restrict__ row_ptr,
                                       int32_t* __restrict__ col) {
    ccn_pdl_wait();
    const int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const int64_t p = idxs[q];
    const int o = ord[p];
    // K = int(min(max k_c, max #points)) (point_ops.py:165-167)
    const int K = (int)fminf(__int_as_float(stats[0]), __int_as_float(stats[1]));
    // point_ops.py:185 looks the per-curve k up with the LOCAL curve id of the query (reference quirk)
    const int64_t lc = curve_local[p];
    const float kq = (lc >= 0 && lc < n_curves) ? kc[lc] : 0.f;
    int cnt = 0;
    int32_t* dst = FILL ? col + row_ptr[q] : nullptr;
    if ((float)(cnt + 1) <= kq) {            // offset 0 (the query itself)
        if (FILL) dst[cnt] = (int32_t)p;
        ++cnt;
    }
    bool left = true, right = true;
    for (int d = 1; d <= K && (left || right); ++d) {
        if (left) {
            const int64_t c = p - d;
            left = c >= 0 && ord[c] == o;
            if (left && (float)(cnt + 1) <= kq) {
                if (FILL) dst[cnt] = (int32_t)c;
                ++cnt;
            }
        }
        if (right) {
            const int64_t c = p + d;
            right = c < n && ord[c] == o;
            if (right && (float)(cnt + 1) <= kq) {
                if (FILL) dst[cnt] = (int32_t)c;
                ++cnt;
            }
        }
        if (!((float)(cnt + 1) <= kq)) break;
    }
    if (!FILL) row_ptr[q] = cnt;
}

// Exclusive scan of an int32 array in place (single CTA, running carry); total -> *total_out,
// also written to a[n] (CSR terminator).  16 items per thread and round (the rounds are a load -> block scan -> store latency
// chain: 17 k row counts are 2 rounds instead of 5, the 262 k cell counts of a level-1 grid 16 instead of 64); the next
// round's items are loaded before this round's scan.
constexpr int kScanItems = 16;
__global__ void __launch_bounds__(1024) excl_scan_i32_kernel(int32_t* __restrict__ a, int64_t n,
                                                             int32_t* __restrict__ total_out) {
    ccn_pdl_wait();
    __shared__ int sm[33];
    int carry = 0;
    const bool vec = (reinterpret_cast<uintptr_t>(a) & 15) == 0;
    int v[kScanItems], nx[kScanItems];
    auto load = [&](int64_t base, int (&dst)[kScanItems]) {
        const int64_t i0 = base + (int64_t)threadIdx.x * kScanItems;
        if (vec && i0 + kScanItems <= n) {
#pragma unroll
            for (int k = 0; k < kScanItems; k += 4) {
                const int4 q = *reinterpret_cast<const int4*>(a + i0 + k);
                dst[k] = q.x; dst[k + 1] = q.y; dst[k + 2] = q.z; dst[k + 3] = q.w;
            }
        } else {
#pragma unroll
            for (int k = 0; k < kScanItems; ++k) dst[k] = (i0 + k < n) ? a[i0 + k] : 0;
        }
    };
    if (n > 0) load(0, v);
    for (int64_t base = 0; base < n; base += 1024 * kScanItems) {
        const int64_t i0 = base + (int64_t)threadIdx.x * kScanItems;
        const bool more = base + 1024 * kScanItems < n;
        if (more) load(base + 1024 * kScanItems, nx);          // rows of the next round: not written by this one
        int local = 0;
#pragma unroll
        for (int k = 0; k < kScanItems; ++k) local += v[k];
        int total;
        int run = carry + ccn_block_excl_scan<int>(local, sm, &total);
        if (vec && i0 + kScanItems <= n) {
#pragma unroll
            for (int k = 0; k < kScanItems; k += 4) {
                int4 q;
                q.x = run; run += v[k];
                q.y = run; run += v[k + 1];
                q.z = run; run += v[k + 2];
                q.w = run; run += v[k + 3];
                *reinterpret_cast<int4*>(a + i0 + k) = q;
            }
        } else {
#pragma unroll
            for (int k = 0; k < kScanItems; ++k) {
                if (i0 + k < n) a[i0 + k] = run;
                run += v[k];
            }
        }
        carry += total;
        if (more) {
#pragma unroll
            for (int k = 0; k < kScanItems; ++k) v[k] = nx[k];
        }
    }
    if (threadIdx.x == 0) {
        a[n] = carry;
        if (total_out) *total_out = carry;
    }
}

// knn_1d_group_superset: per fine point the <= knn nearest of the 2*knn+3 along-curve sampled
// candidates, ordered by (Euclidean distance, candidate order); weights 1/clamp(d^2, 1e-16)
// (point_ops.py:344-355).  nbr: (n, knn) sampled ORDINAL or -1, wgt: (n, knn).
constexpr int kMaxKnn1d = 8;
__global__ void knn_1d_superset_kernel(const float* __restrict__ pos, const int32_t* __restrict__ ord, int64_t n,
                                       const int64_t* __restrict__ idxs, int64_t nq, int knn,
                                       int32_t* __restrict__ nbr, float* __restrict__ wgt) {
    ccn_pdl_wait();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    // a = #sampled positions <= i  (cumsum of the hit mask, point_ops.py:203-205)
    int64_t lo = 0, hi = nq;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (idxs[mid] <= i) lo = mid + 1; else hi = mid;
    }
    const int64_t a = lo;
    const int o = ord[i];
    const float px = pos[3 * i], py = pos[3 * i + 1], pz = pos[3 * i + 2];
    float bd[kMaxKnn1d];
    int32_t bi[kMaxKnn1d];
    int have = 0;
    const int W = 2 * (knn + 1) + 1;
    for (int t = 0; t < W; ++t) {
        const int m = (t + 1) >> 1;
        const int64_t c = a + ((t & 1) ? -m : m);       // 0,-1,+1,-2,+2,...
        if (c < 0 || c >= nq) continue;
        const int64_t p = idxs[c];
        if (ord[p] != o) continue;
        const float d = ccn_norm3(__fsub_rn(pos[3 * p], px), __fsub_rn(pos[3 * p + 1], py),
                                  __fsub_rn(pos[3 * p + 2], pz));
        // stable insertion: strictly smaller distances move ahead, equal ones keep candidate order
        int s = have < knn ? have : knn;
        while (s > 0 && d < bd[s - 1]) --s;
        if (s >= knn) continue;
        const int last = have < knn ? have : knn - 1;
        for (int j = last; j > s; --j) { bd[j] = bd[j - 1]; bi[j] = bi[j - 1]; }
        bd[s] = d; bi[s] = (int32_t)c;
        if (have < knn) ++have;
    }
    for (int j = 0; j < knn; ++j) {
        if (j < have) {
            nbr[i * knn + j] = bi[j];
            const float d2 = __fmul_rn(bd[j], bd[j]);
            wgt[i * knn + j] = __fdiv_rn(1.f, fmaxf(d2, 1e-16f));
        } else {
            nbr[i * knn + j] = -1;
            wgt[i * knn + j] = 0.f;
        }
    }
}

}  // namespace

extern "C" int ccn_curve_fps(const float* pos, const int32_t* curve_ord, const int32_t* curve_ptr, int64_t n,
                             const float* u, float arclen, float* run_scratch, int64_t* idx_out, int32_t* count_out,
                             ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && n < 0x7fffffff && arclen != 0.f && u && run_scratch && idx_out && count_out);
    CCN_REQUIRE(n == 0 || (pos && curve_ord && curve_ptr));
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0) {
        const cudaError_t e = cudaMemsetAsync(count_out, 0, sizeof(int32_t), s);
        return e == cudaSuccess ? CCN_OK : (int)e;
    }
    const int tiles = (int)((n + kTile - 1) / kTile);
    double* tile_tot = reinterpret_cast<double*>(idx_out);                    // consumed before idx_out is written
    int32_t* tile_cnt = reinterpret_cast<int32_t*>(run_scratch + n);          // the tail of run_scratch
    ccn_launch(cfps_tile_sum_kernel, dim3(tiles), dim3(1024), 0, s, pos, curve_ord, n, tile_tot);
    CCN_CHECK_LAUNCH();
    ccn_launch(cfps_run_kernel, dim3(tiles), dim3(1024), 0, s, pos, curve_ord, n, tile_tot, run_scratch);
    CCN_CHECK_LAUNCH();
    ccn_launch(cfps_count_kernel, dim3(tiles), dim3(1024), 0, s, run_scratch, curve_ord, curve_ptr, n, u, arclen, tile_cnt);
    CCN_CHECK_LAUNCH();
    ccn_launch(cfps_compact_kernel, dim3(tiles), dim3(1024), 0, s, run_scratch, curve_ord, curve_ptr, n, u, arclen, tile_cnt, idx_out, count_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_radius_1d_group_count(const float* pos, const int32_t* curve_ord, const int32_t* curve_ptr,
                                         const int64_t* curve_local, int64_t n, int32_t n_curves,
                                         const int64_t* idxs, int64_t nq, float radius, float* kc_scratch,
                                         int32_t* stats_scratch, int32_t* row_ptr, int32_t* total_out,
                                         ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && nq >= 0 && n_curves >= 0 && kc_scratch && stats_scratch && row_ptr && total_out);
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(stats_scratch, 0, 2 * sizeof(int32_t), s);
    if (e != cudaSuccess) return (int)e;
    if (n_curves > 0) {
        ccn_launch(radius_1d_curve_stats_kernel, dim3(ccn_blocks_exact((int64_t)n_curves * 32, 256)), dim3(256), 0, s, 
            pos, curve_ord, curve_ptr, n_curves, radius, kc_scratch, stats_scratch);
        CCN_CHECK_LAUNCH();
    }
    if (nq > 0) {
        ccn_launch(radius_1d_group_kernel<false>, dim3(ccn_blocks_exact(nq, 128)), dim3(128), 0, s, 
            idxs, nq, curve_ord, curve_local, n, n_curves, kc_scratch, stats_scratch, row_ptr, nullptr);
        CCN_CHECK_LAUNCH();
    }
    ccn_launch(excl_scan_i32_kernel, dim3(1), dim3(1024), 0, s, row_ptr, nq, total_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

// The two halves of ccn_radius_1d_group_count for callers that overlap them: the per-curve statistics depend on the level's
// positions and curve layout only (not on the sampled queries), so the native executor computes them on its geometry stream
// while the main stream is still busy with the step before.
extern "C" int ccn_radius_1d_curve_stats(const float* pos, const int32_t* curve_ord, const int32_t* curve_ptr, int32_t n_curves,
                                         float radius, float* kc_scratch, int32_t* stats_scratch, ccn_stream_t stream) {
    CCN_REQUIRE(n_curves >= 0 && kc_scratch && stats_scratch);
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(stats_scratch, 0, 2 * sizeof(int32_t), s);
    if (e != cudaSuccess) return (int)e;
    if (n_curves > 0) {
        ccn_launch(radius_1d_curve_stats_kernel, dim3(ccn_blocks_exact((int64_t)n_curves * 32, 256)), dim3(256), 0, s, pos, curve_ord,
                   curve_ptr, n_curves, radius, kc_scratch, stats_scratch);
        CCN_CHECK_LAUNCH();
    }
    return CCN_OK;
}

extern "C" int ccn_radius_1d_group_count_ready(const int32_t* curve_ord, const int64_t* curve_local, int64_t n, int32_t n_curves,
                                               const int64_t* idxs, int64_t nq, const float* kc_ready, const int32_t* stats_ready,
                                               int32_t* row_ptr, int32_t* total_out, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && nq >= 0 && n_curves >= 0 && kc_ready && stats_ready && row_ptr && total_out);
    cudaStream_t s = (cudaStream_t)stream;
    if (nq > 0) {
        ccn_launch(radius_1d_group_kernel<false>, dim3(ccn_blocks_exact(nq, 128)), dim3(128), 0, s, idxs, nq, curve_ord, curve_local, n,
                   n_curves, kc_ready, stats_ready, row_ptr, nullptr);
        CCN_CHECK_LAUNCH();
    }
    ccn_launch(excl_scan_i32_kernel, dim3(1), dim3(1024), 0, s, row_ptr, nq, total_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_radius_1d_group_fill(const int32_t* curve_ord, const int64_t* curve_local, int64_t n,
                                        int32_t n_curves, const int64_t* idxs, int64_t nq, const float* kc_scratch,
                                        const int32_t* stats_scratch, const int32_t* row_ptr, int32_t* col,
                                        ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && nq >= 0 && kc_scratch && stats_scratch && row_ptr && (col || nq == 0));
    if (nq == 0) return CCN_OK;
    ccn_launch(radius_1d_group_kernel<true>, dim3(ccn_blocks_exact(nq, 128)), dim3(128), 0, (cudaStream_t)stream, 
        idxs, nq, curve_ord, curve_local, n, n_curves, kc_scratch, stats_scratch, const_cast<int32_t*>(row_ptr), col);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_excl_scan_i32(int32_t* a, int64_t n, int32_t* total_out, ccn_stream_t stream) {
    CCN_REQUIRE(a && n >= 0);
    ccn_launch(excl_scan_i32_kernel, dim3(1), dim3(1024), 0, (cudaStream_t)stream, a, n, total_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_knn_1d_superset(const float* pos, const int32_t* curve_ord, int64_t n, const int64_t* idxs,
                                   int64_t nq, int32_t knn, int32_t* nbr, float* wgt, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && nq >= 0 && knn >= 1 && knn <= kMaxKnn1d && nbr && wgt);
    if (n == 0) return CCN_OK;
    ccn_launch(knn_1d_superset_kernel, dim3(ccn_blocks_exact(n, 128)), dim3(128), 0, (cudaStream_t)stream, pos, curve_ord, n, idxs, nq,
                                                                                      knn, nbr, wgt);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
