// Neighbour search:
//   ccn_grid_build + ccn_frnn_knn   fixed-radius kNN  (frnn.frnn_grid_points, reference
//                                   third_party/FRNN/frnn/frnn.py:210-347, kernels
//                                   third_party/FRNN/frnn/csrc/grid/{insert_points,counting_sort,find_nbrs}.cu)
//   ccn_knn_exact                   exact kNN         (pytorch3d.ops.knn_points, call site point_ops.py:91)
//   ccn_ball_query                  first-K in radius (pytorch3d.ops.ball_query, call site point_ops.py:81)
//
// Not a translation of FRNN: the reference bins into a dense <=64^3 grid whose cell size grows
// with the scene extent (frnn.py:55,78-79), sorts queries too, scans cell counts per cloud from a
// python loop and keeps K candidates per THREAD in registers.  Here one hashed grid with cell size
// ~r covers all clouds of the batch (count -> scan -> fill, 3 launches, no host sync), and one WARP
// per query keeps the K best (d2, index) keys as a sorted list distributed over its lanes, so the
// result is the unique total-order answer regardless of the (atomic) fill order inside a cell.
#include "common.cuh"

namespace {

constexpr unsigned long long kKeyMax = 0xffffffffffffffffull;

__device__ __forceinline__ unsigned cell_hash(int b, int cx, int cy, int cz, unsigned mask) {
    unsigned h = (unsigned)cx * 73856093u ^ (unsigned)cy * 19349663u ^ (unsigned)cz * 83492791u ^
                 (unsigned)b * 2654435761u;
    h ^= h >> 15;
    return h & mask;
}

__device__ __forceinline__ int cell_of(float v, float inv_cell) { return __float2int_rd(__fmul_rn(v, inv_cell)); }

__global__ void grid_count_kernel(const float* __restrict__ pos, const int64_t* __restrict__ batch, int64_t n,
                                  float inv_cell, unsigned mask, int32_t* __restrict__ cell_cnt,
                                  int32_t* __restrict__ slot) {
    ccn_pdl_wait();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned h = cell_hash((int)batch[i], cell_of(pos[3 * i], inv_cell), cell_of(pos[3 * i + 1], inv_cell),
                                 cell_of(pos[3 * i + 2], inv_cell), mask);
    slot[i] = atomicAdd(&cell_cnt[h], 1);
}

__global__ void grid_fill_kernel(const float* __restrict__ pos, const int64_t* __restrict__ batch, int64_t n,
                                 float inv_cell, unsigned mask, const int32_t* __restrict__ cell_off,
                                 const int32_t* __restrict__ slot, float4* __restrict__ sorted) {
    ccn_pdl_wait();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = pos[3 * i], y = pos[3 * i + 1], z = pos[3 * i + 2];
    const unsigned h = cell_hash((int)batch[i], cell_of(x, inv_cell), cell_of(y, inv_cell), cell_of(z, inv_cell), mask);
    sorted[cell_off[h] + slot[i]] = make_float4(x, y, z, __int_as_float((int)i));
}

// Sorted K-best list distributed over the lanes of a warp: lane l holds the l-th smallest key.
struct WarpTopK {
    unsigned long long held;
    __device__ __forceinline__ void init() { held = kKeyMax; }
    __device__ __forceinline__ unsigned long long worst(int K) const { return __shfl_sync(0xffffffffu, held, K - 1); }
    // every lane passes its own candidate (valid or not); candidates are merged one at a time
    __device__ __forceinline__ void offer(bool valid, unsigned long long key, int K) {
        unsigned long long thr = worst(K);
        unsigned m = __ballot_sync(0xffffffffu, valid && key < thr);
        const int lane = ccn_lane();
        while (m) {
            const int src = __ffs(m) - 1;
            m &= m - 1;
            const unsigned long long ck = __shfl_sync(0xffffffffu, key, src);
            if (ck < thr) {
                const int p = __popc(__ballot_sync(0xffffffffu, held < ck));
                const unsigned long long up = __shfl_up_sync(0xffffffffu, held, 1);
                if (lane == p) held = ck;
                else if (lane > p) held = up;
                thr = worst(K);
            }
        }
    }
};

// One warp per query.  p2 points come from the hashed grid (27 cells around the query's cell).
__global__ void __launch_bounds__(256) frnn_knn_kernel(
    const float* __restrict__ p1, const int64_t* __restrict__ batch1, int64_t n1,
    const float4* __restrict__ sorted, const int32_t* __restrict__ cell_off, const int32_t* __restrict__ sptr2,
    float inv_cell, unsigned mask, float r2, int K, int32_t* __restrict__ idx_out, float* __restrict__ d2_out,
    int32_t* __restrict__ cnt_out) {
    ccn_pdl_wait();
    const int64_t q = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (q >= n1) return;
    const int lane = ccn_lane();
    const float qx = p1[3 * q], qy = p1[3 * q + 1], qz = p1[3 * q + 2];
    const int b = (int)batch1[q];
    const int lo2 = sptr2[b], hi2 = sptr2[b + 1];
    const int cx = cell_of(qx, inv_cell), cy = cell_of(qy, inv_cell), cz = cell_of(qz, inv_cell);
    // lane l < 27 owns neighbour cell l
    int tx = 0, ty = 0, tz = 0, beg = 0, len = 0;
    if (lane < 27) {
        tx = cx + lane % 3 - 1; ty = cy + (lane / 3) % 3 - 1; tz = cz + lane / 9 - 1;
        const unsigned h = cell_hash(b, tx, ty, tz, mask);
        beg = cell_off[h];
        len = cell_off[h + 1] - beg;
    }
    // two different cells of the 27 may share a bucket: every cell filters by exact cell match,
    // so a bucket visited twice contributes each point once
    WarpTopK top;
    top.init();
    for (int c = 0; c < 27; ++c) {
        const int cb = __shfl_sync(0xffffffffu, beg, c), cl = __shfl_sync(0xffffffffu, len, c);
        const int ctx = __shfl_sync(0xffffffffu, tx, c), cty = __shfl_sync(0xffffffffu, ty, c),
                  ctz = __shfl_sync(0xffffffffu, tz, c);
        for (int j0 = 0; j0 < cl; j0 += 32) {
            const int j = j0 + lane;
            bool ok = false;
            unsigned long long key = kKeyMax;
            if (j < cl) {
                const float4 p = sorted[cb + j];
                const int pi = __float_as_int(p.w);
                if (pi >= lo2 && pi < hi2 && cell_of(p.x, inv_cell) == ctx && cell_of(p.y, inv_cell) == cty &&
                    cell_of(p.z, inv_cell) == ctz) {
                    const float d2 = ccn_sqdist3(qx, qy, qz, p.x, p.y, p.z);
                    ok = d2 <= r2;
                    key = ccn_pack_key(d2, pi);
                }
            }
            top.offer(ok, key, K);
        }
    }
    const bool have = top.held != kKeyMax;
    if (lane < K) {
        idx_out[q * K + lane] = have ? (int32_t)(top.held & 0xffffffffu) : -1;
        if (d2_out) d2_out[q * K + lane] = have ? __uint_as_float((unsigned)(top.held >> 32)) : -1.f;
    }
    if (cnt_out) {
        const int cnt = __popc(__ballot_sync(0xffffffffu, have && lane < K));
        if (lane == 0) cnt_out[q] = cnt;
    }
}

// Exact kNN: brute force over the query's own cloud, same warp list.  Unfilled slots: -1, or the
// first point of the cloud when zero_fill (pytorch3d pads idx with 0 where lengths2 < K).
__global__ void __launch_bounds__(256) knn_exact_kernel(
    const float* __restrict__ p1, const int64_t* __restrict__ batch1, int64_t n1, const float* __restrict__ p2,
    const int32_t* __restrict__ sptr2, int K, int zero_fill, int32_t* __restrict__ idx_out,
    float* __restrict__ d2_out, int32_t* __restrict__ cnt_out) {
    ccn_pdl_wait();
    const int64_t q = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (q >= n1) return;
    const int lane = ccn_lane();
    const float qx = p1[3 * q], qy = p1[3 * q + 1], qz = p1[3 * q + 2];
    const int b = (int)batch1[q];
    const int lo2 = sptr2[b], hi2 = sptr2[b + 1];
    WarpTopK top;
    top.init();
    for (int j0 = lo2; j0 < hi2; j0 += 32) {
        const int j = j0 + lane;
        bool ok = false;
        unsigned long long key = kKeyMax;
        if (j < hi2) {
            const float d2 = ccn_sqdist3(qx, qy, qz, p2[3 * j], p2[3 * j + 1], p2[3 * j + 2]);
            ok = true;
            key = ccn_pack_key(d2, j);
        }
        top.offer(ok, key, K);
    }
    if (lane < K) {
        const bool have = top.held != kKeyMax;
        int32_t id = have ? (int32_t)(top.held & 0xffffffffu) : (zero_fill && hi2 > lo2 ? lo2 : -1);
        idx_out[q * K + lane] = id;
        // filled slots carry the true distance to the cloud's first point: that is what the reference's
        // interpolation recomputes from the padded index (point_ops.py:333-336)
        if (d2_out)
            d2_out[q * K + lane] = have ? __uint_as_float((unsigned)(top.held >> 32))
                                        : (id >= 0 ? ccn_sqdist3(qx, qy, qz, p2[3 * id], p2[3 * id + 1], p2[3 * id + 2]) : 0.f);
    }
    if (cnt_out) {
        const bool filled = top.held != kKeyMax || (zero_fill && hi2 > lo2);
        const int cnt = __popc(__ballot_sync(0xffffffffu, filled && lane < K));
        if (lane == 0) cnt_out[q] = cnt;
    }
}


// Exact kNN for small K (the K=3 of feature propagation, point_ops.py:331): 8 lanes per query, 32 queries per CTA,
// the source cloud streamed through shared memory in SoA tiles shared by all queries of the CTA, the K best
// (d2, index) keys of every lane in registers, merged across the 8 lanes by a shuffle butterfly at the end.
// Same total order as knn_exact_kernel, so the two agree bit for bit.
constexpr int kSmallTile = 2048;

template <int K>
__device__ __forceinline__ void topk_insert(unsigned long long (&k)[K], unsigned long long c) {
    if (c < k[K - 1]) {
        k[K - 1] = c;
#pragma unroll
        for (int i = K - 1; i > 0; --i) {
            const unsigned long long a = k[i - 1], b = k[i];
            const bool sw = b < a;
            k[i - 1] = sw ? b : a;
            k[i] = sw ? a : b;
        }
    }
}

template <int K>
__global__ void __launch_bounds__(256) knn_small_kernel(
    const float* __restrict__ p1, const int64_t* __restrict__ batch1, int64_t n1, const float* __restrict__ p2,
    const int32_t* __restrict__ sptr2, int Kout, int zero_fill, int32_t* __restrict__ idx_out,
    float* __restrict__ d2_out, int32_t* __restrict__ cnt_out) {
    ccn_pdl_wait();
    __shared__ float sx[kSmallTile], sy[kSmallTile], sz[kSmallTile];
    __shared__ int s_lo, s_hi;
    const int sub = threadIdx.x & 7;
    const int64_t q = (int64_t)blockIdx.x * 32 + (threadIdx.x >> 3);
    const bool live = q < n1;
    float qx = 0.f, qy = 0.f, qz = 0.f;
    int lo2 = 0, hi2 = 0;
    if (live) {
        qx = p1[3 * q]; qy = p1[3 * q + 1]; qz = p1[3 * q + 2];
        const int b = (int)batch1[q];
        lo2 = sptr2[b]; hi2 = sptr2[b + 1];
    }
    if (threadIdx.x == 0) {                 // queries are batch-sorted: the CTA's source range is one interval
        const int64_t qa = (int64_t)blockIdx.x * 32, qb = min(qa + 31, n1 - 1);
        s_lo = sptr2[(int)batch1[qa]];
        s_hi = sptr2[(int)batch1[qb] + 1];
    }
    __syncthreads();
    const int blo = s_lo, bhi = s_hi;
    unsigned long long best[K];
#pragma unroll
    for (int i = 0; i < K; ++i) best[i] = kKeyMax;
    float worst = INFINITY;
    for (int t0 = blo; t0 < bhi; t0 += kSmallTile) {
        const int tn = min(kSmallTile, bhi - t0);
        __syncthreads();
        for (int j = threadIdx.x; j < tn; j += blockDim.x) {
            sx[j] = p2[3 * (int64_t)(t0 + j)]; sy[j] = p2[3 * (int64_t)(t0 + j) + 1]; sz[j] = p2[3 * (int64_t)(t0 + j) + 2];
        }
        __syncthreads();
        const int ja = max(lo2 - t0, 0), jb = min(hi2 - t0, tn);
        for (int j = ja + sub; j < jb; j += 8) {
            const float d2 = ccn_sqdist3(qx, qy, qz, sx[j], sy[j], sz[j]);
            if (d2 <= worst) {
                topk_insert<K>(best, ccn_pack_key(d2, t0 + j));
                worst = __uint_as_float((unsigned)(best[K - 1] >> 32));     // +inf bits while the list is not full
                if (best[K - 1] == kKeyMax) worst = INFINITY;
            }
        }
    }
    // merge the 8 lanes of a query (all lanes end with the full result)
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
        unsigned long long other[K];
#pragma unroll
        for (int i = 0; i < K; ++i) other[i] = __shfl_xor_sync(0xffffffffu, best[i], o);
#pragma unroll
        for (int i = 0; i < K; ++i) topk_insert<K>(best, other[i]);
    }
    if (!live || sub != 0) return;
    int cnt = 0;
#pragma unroll
    for (int i = 0; i < K; ++i) {
        if (i >= Kout) break;
        const bool have = best[i] != kKeyMax;
        const int32_t id = have ? (int32_t)(best[i] & 0xffffffffu) : (zero_fill && hi2 > lo2 ? lo2 : -1);
        idx_out[q * Kout + i] = id;
        if (d2_out)
            d2_out[q * Kout + i] = have ? __uint_as_float((unsigned)(best[i] >> 32))
                                        : (id >= 0 ? ccn_sqdist3(qx, qy, qz, p2[3 * id], p2[3 * id + 1], p2[3 * id + 2]) : 0.f);
        cnt += (have || (zero_fill && hi2 > lo2)) ? 1 : 0;
    }
    if (cnt_out) cnt_out[q] = cnt;
}

// ball_query: first K points of the query's cloud in INDEX order with d2 < r2 (strict).
__global__ void __launch_bounds__(256) ball_query_kernel(
    const float* __restrict__ p1, const int64_t* __restrict__ batch1, int64_t n1, const float* __restrict__ p2,
    const int32_t* __restrict__ sptr2, float r2, int K, int32_t* __restrict__ idx_out,
    int32_t* __restrict__ cnt_out) {
    ccn_pdl_wait();
    const int64_t q = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (q >= n1) return;
    const int lane = ccn_lane();
    const float qx = p1[3 * q], qy = p1[3 * q + 1], qz = p1[3 * q + 2];
    const int b = (int)batch1[q];
    const int lo2 = sptr2[b], hi2 = sptr2[b + 1];
    int cnt = 0;
    for (int j0 = lo2; j0 < hi2 && cnt < K; j0 += 32) {
        const int j = j0 + lane;
        bool hit = false;
        if (j < hi2) hit = ccn_sqdist3(qx, qy, qz, p2[3 * j], p2[3 * j + 1], p2[3 * j + 2]) < r2;
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        const int at = cnt + __popc(m & ((1u << lane) - 1u));
        if (hit && at < K) idx_out[q * K + at] = j;
        cnt += __popc(m);
    }
    cnt = min(cnt, K);
    for (int t = cnt + lane; t < K; t += 32) idx_out[q * K + t] = -1;
    if (cnt_out && lane == 0) cnt_out[q] = cnt;
}

// Neighbour search in FEATURE space (DynamicEdgeConv, dgcnn.py:64-89: the graph is rebuilt from the D-dimensional rows
// [x, pos] of every layer).  One warp per query, brute force over the query's own cloud; squared distance accumulated
// over the D columns in order, every operation rounded (no FMA) - oracle/tp_ops.py::_sqdist.
//   mode 0: the FRNN contract (fast_knn with r = 0.25, point_ops.py:84-88): d2 <= r2, the K smallest by (d2, index)
//   mode 1: ball_query (point_ops.py:81): d2 < r2, the first K in index order
__device__ __forceinline__ float feat_sqdist(const float* __restrict__ a, const float* __restrict__ b, int dim) {
    float t = __fsub_rn(a[0], b[0]);
    float acc = __fmul_rn(t, t);
    for (int d = 1; d < dim; ++d) {
        t = __fsub_rn(a[d], b[d]);
        acc = __fadd_rn(acc, __fmul_rn(t, t));
    }
    return acc;
}

__global__ void __launch_bounds__(256) feat_neighbours_kernel(
    const float* __restrict__ x1, int64_t ld1, const int64_t* __restrict__ batch1, int64_t n1,
    const float* __restrict__ x2, int64_t ld2, const int32_t* __restrict__ sptr2, int dim, float r2, int K, int mode,
    int32_t* __restrict__ idx_out, float* __restrict__ d2_out, int32_t* __restrict__ cnt_out) {
    ccn_pdl_wait();
    const int64_t q = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (q >= n1) return;
    const int lane = ccn_lane();
    const float* qrow = x1 + q * ld1;
    const int b = (int)batch1[q];
    const int lo2 = sptr2[b], hi2 = sptr2[b + 1];
    if (mode == 0) {
        WarpTopK top;
        top.init();
        for (int j0 = lo2; j0 < hi2; j0 += 32) {
            const int j = j0 + lane;
            bool ok = false;
            unsigned long long key = kKeyMax;
            if (j < hi2) {
                const float d2 = feat_sqdist(qrow, x2 + (int64_t)j * ld2, dim);
                ok = d2 <= r2;
                key = ccn_pack_key(d2, j);
            }
            top.offer(ok, key, K);
        }
        const bool have = top.held != kKeyMax;
        if (lane < K) {
            idx_out[q * K + lane] = have ? (int32_t)(top.held & 0xffffffffu) : -1;
            if (d2_out) d2_out[q * K + lane] = have ? __uint_as_float((unsigned)(top.held >> 32)) : -1.f;
        }
        if (cnt_out) {
            const int cnt = __popc(__ballot_sync(0xffffffffu, have && lane < K));
            if (lane == 0) cnt_out[q] = cnt;
        }
        return;
    }
    int cnt = 0;
    for (int j0 = lo2; j0 < hi2 && cnt < K; j0 += 32) {
        const int j = j0 + lane;
        bool hit = false;
        if (j < hi2) hit = feat_sqdist(qrow, x2 + (int64_t)j * ld2, dim) < r2;
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        const int at = cnt + __popc(m & ((1u << lane) - 1u));
        if (hit && at < K) idx_out[q * K + at] = j;
        cnt += __popc(m);
    }
    cnt = min(cnt, K);
    for (int t = cnt + lane; t < K; t += 32) idx_out[q * K + t] = -1;
    if (cnt_out && lane == 0) cnt_out[q] = cnt;
}

}  // namespace

extern "C" int64_t ccn_grid_num_cells(int64_t n2) {
    int64_t m = 1024;
    while (m < 2 * n2 && m < (1LL << 26)) m <<= 1;
    return m;
}

extern "C" int ccn_excl_scan_i32(int32_t* a, int64_t n, int32_t* total_out, ccn_stream_t stream);

extern "C" int ccn_grid_build(const float* pos2, const int64_t* batch2, int64_t n2, float radius, int64_t num_cells,
                              int32_t* cell_off, int32_t* slot_scratch, float* sorted_xyzi, ccn_stream_t stream) {
    CCN_REQUIRE(n2 >= 0 && n2 < 0x7fffffff && radius > 0.f && cell_off && slot_scratch && sorted_xyzi);
    CCN_REQUIRE(num_cells >= 2 && (num_cells & (num_cells - 1)) == 0 && num_cells <= (1LL << 30));
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(cell_off, 0, (num_cells + 1) * sizeof(int32_t), s);
    if (e != cudaSuccess) return (int)e;
    const float inv_cell = 1.f / (radius * 1.0001f);
    const unsigned mask = (unsigned)(num_cells - 1);
    if (n2 > 0) {
        ccn_launch(grid_count_kernel, dim3(ccn_blocks_exact(n2, 256)), dim3(256), 0, s, pos2, batch2, n2, inv_cell, mask, cell_off,
                                                                   slot_scratch);
        CCN_CHECK_LAUNCH();
    }
    int rc = ccn_excl_scan_i32(cell_off, num_cells, nullptr, stream);
    if (rc) return rc;
    if (n2 > 0) {
        ccn_launch(grid_fill_kernel, dim3(ccn_blocks_exact(n2, 256)), dim3(256), 0, s, pos2, batch2, n2, inv_cell, mask, cell_off,
                                                                  slot_scratch, (float4*)sorted_xyzi);
        CCN_CHECK_LAUNCH();
    }
    return CCN_OK;
}

extern "C" int ccn_frnn_knn(const float* pos1, const int64_t* batch1, int64_t n1, const float* sorted_xyzi,
                            const int32_t* cell_off, int64_t num_cells, const int32_t* sample_ptr2, float radius,
                            int32_t K, int32_t* idx_out, float* d2_out, int32_t* cnt_out, ccn_stream_t stream) {
    CCN_REQUIRE(n1 >= 0 && K >= 1 && K <= 32 && radius > 0.f && idx_out && sorted_xyzi && cell_off && sample_ptr2);
    CCN_REQUIRE(num_cells >= 2 && (num_cells & (num_cells - 1)) == 0);
    if (n1 == 0) return CCN_OK;
    const float inv_cell = 1.f / (radius * 1.0001f);
    const float r2 = radius * radius;      // f32 product, as frnn.py computes r2s = rs * rs
    ccn_launch(frnn_knn_kernel, dim3(ccn_blocks_exact(n1 * 32, 256)), dim3(256), 0, (cudaStream_t)stream, 
        pos1, batch1, n1, (const float4*)sorted_xyzi, cell_off, sample_ptr2, inv_cell, (unsigned)(num_cells - 1), r2, K,
        idx_out, d2_out, cnt_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_knn_exact(const float* pos1, const int64_t* batch1, int64_t n1, const float* pos2,
                             const int32_t* sample_ptr2, int32_t K, int32_t zero_fill, int32_t* idx_out,
                             float* d2_out, int32_t* cnt_out, ccn_stream_t stream) {
    CCN_REQUIRE(n1 >= 0 && K >= 1 && K <= 32 && idx_out && sample_ptr2);
    if (n1 == 0) return CCN_OK;
    if (K <= 4) {
        const unsigned blocks = ccn_blocks_exact(n1, 32);
        if (K <= 2)
            ccn_launch(knn_small_kernel<2>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, pos1, batch1, n1, pos2, sample_ptr2, K, zero_fill,
                                                                          idx_out, d2_out, cnt_out);
        else if (K == 3)
            ccn_launch(knn_small_kernel<3>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, pos1, batch1, n1, pos2, sample_ptr2, K, zero_fill,
                                                                          idx_out, d2_out, cnt_out);
        else
            ccn_launch(knn_small_kernel<4>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, pos1, batch1, n1, pos2, sample_ptr2, K, zero_fill,
                                                                          idx_out, d2_out, cnt_out);
        CCN_CHECK_LAUNCH();
        return CCN_OK;
    }
    ccn_launch(knn_exact_kernel, dim3(ccn_blocks_exact(n1 * 32, 256)), dim3(256), 0, (cudaStream_t)stream, 
        pos1, batch1, n1, pos2, sample_ptr2, K, zero_fill, idx_out, d2_out, cnt_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_ball_query(const float* pos1, const int64_t* batch1, int64_t n1, const float* pos2,
                              const int32_t* sample_ptr2, float radius, int32_t K, int32_t* idx_out,
                              int32_t* cnt_out, ccn_stream_t stream) {
    CCN_REQUIRE(n1 >= 0 && K >= 1 && idx_out && sample_ptr2);
    if (n1 == 0) return CCN_OK;
    const float r2 = radius * radius;
    ccn_launch(ball_query_kernel, dim3(ccn_blocks_exact(n1 * 32, 256)), dim3(256), 0, (cudaStream_t)stream, 
        pos1, batch1, n1, pos2, sample_ptr2, r2, K, idx_out, cnt_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_feat_neighbours(const float* x1, int64_t ld1, const int64_t* batch1, int64_t n1, const float* x2,
                                   int64_t ld2, const int32_t* sample_ptr2, int32_t dim, float radius, int32_t K,
                                   int32_t mode, int32_t* idx_out, float* d2_out, int32_t* cnt_out, ccn_stream_t stream) {
    CCN_REQUIRE(n1 >= 0 && dim >= 1 && ld1 >= dim && ld2 >= dim && K >= 1 && (mode == 0 || mode == 1) && idx_out && sample_ptr2);
    CCN_REQUIRE(mode == 1 || K <= 32);
    CCN_REQUIRE(mode == 0 || !d2_out);                      // the ball query keeps index order and reports no distances
    if (n1 == 0) return CCN_OK;
    CCN_REQUIRE(x1 && x2 && batch1);
    ccn_launch(feat_neighbours_kernel, dim3(ccn_blocks_exact(n1 * 32, 256)), dim3(256), 0, (cudaStream_t)stream,
               x1, ld1, batch1, n1, x2, ld2, sample_ptr2, (int)dim, radius * radius, (int)K, (int)mode, idx_out, d2_out, cnt_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
