// Exact K-nearest neighbours (K <= 4) through the hashed grid: ccn_knn_grid.
// Replaces pytorch3d.ops.knn_points as called for the feature-propagation interpolation (reference
// src/models/utils/point_ops.py:293-341, `knn_points(pos_y, pos_x, K=3)` at :331) - same result as ccn_knn_exact (the
// O(P1 P2) kernels of search_ops.cu: total order (d2, index), zero fill), bit for bit, at O(P1) cost:
//   1. bounding box of the sources -> cell edge h = 1.5 sqrt(A / n2), A = product of the two largest extents (laser scans
//      and object scans are surfaces: ~2 points per cell where the sampling is uniform) - on the DEVICE, no read-back;
//   2. hashed grid of the sources (count -> scan -> fill, the cloud id is part of the hash; same layout as ccn_grid_build);
//   3. one thread per query: the 27 cells around it, then the 98 cells of the next shell; the K best (d2, index) keys in
//      registers.  After the cube of half-width R cells every point NOT examined is farther than R h, so the answer is
//      final as soon as the K-th distance is within R h;
//   4. queries that are still open (sparse regions, clouds with fewer than K points, zero fill) go to a list and a
//      brute-force pass of one warp per listed query - the list length stays on the device.
#include "common.cuh"

namespace {

constexpr unsigned long long kKeyMax = 0xffffffffffffffffull;

__device__ __forceinline__ unsigned cell_hash(int b, int cx, int cy, int cz, unsigned mask) {
    unsigned h = (unsigned)cx * 73856093u ^ (unsigned)cy * 19349663u ^ (unsigned)cz * 83492791u ^ (unsigned)b * 2654435761u;
    h ^= h >> 15;
    return h & mask;
}
__device__ __forceinline__ int cell_of(float v, float inv_cell) { return __float2int_rd(__fmul_rn(v, inv_cell)); }

// order-preserving float <-> int map for atomicMin / atomicMax
__device__ __forceinline__ int f2ord(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

// meta (int32[16]): [0..2] min xyz (ordered ints), [3..5] max xyz, [6] open-query count, [8] h bits, [9] 1/h bits
__global__ void kg_init_kernel(int32_t* meta) {
    ccn_pdl_wait();
    if (threadIdx.x < 3) meta[threadIdx.x] = 0x7fffffff;
    else if (threadIdx.x < 6) meta[threadIdx.x] = (int)0x80000000;
    else if (threadIdx.x < 16) meta[threadIdx.x] = 0;
}

__global__ void kg_bbox_kernel(const float* __restrict__ pos, int64_t n, int32_t* meta) {
    ccn_pdl_wait();
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float v = pos[3 * i + a];
            lo[a] = fminf(lo[a], v);
            hi[a] = fmaxf(hi[a], v);
        }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
        if ((threadIdx.x & 31) == 0 && lo[a] <= hi[a]) {
            atomicMin(&meta[a], f2ord(lo[a]));
            atomicMax(&meta[3 + a], f2ord(hi[a]));
        }
    }
}

__global__ void kg_cell_kernel(int32_t* meta, int64_t n2, float cell_given) {
    ccn_pdl_wait();
    float h = cell_given;
    if (!(h > 0.f)) {
        float e[3];
        for (int a = 0; a < 3; ++a) e[a] = fmaxf(ord2f(meta[3 + a]) - ord2f(meta[a]), 0.f);
        // two largest extents
        const float emax = fmaxf(e[0], fmaxf(e[1], e[2])), emin = fminf(e[0], fminf(e[1], e[2]));
        const float emid = e[0] + e[1] + e[2] - emax - emin;
        const float area = emax * emid;
        if (area > 0.f) h = 1.5f * sqrtf(area / (float)n2);
        else if (emax > 0.f) h = 2.f * emax / (float)n2;
        else h = 1.f;
        h = fmaxf(h, emax * (1.f / 1048576.f));          // cell coordinates stay far inside int range
        if (!(h > 0.f) || isinf(h)) h = 1.f;
    }
    meta[8] = __float_as_int(h);
    meta[9] = __float_as_int(1.f / h);
}

__global__ void kg_count_kernel(const float* __restrict__ pos, const int64_t* __restrict__ batch, int64_t n,
                                const int32_t* __restrict__ meta, unsigned mask, int32_t* __restrict__ cell_cnt,
                                int32_t* __restrict__ slot) {
    ccn_pdl_wait();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float inv_cell = __int_as_float(meta[9]);
    const unsigned h = cell_hash((int)batch[i], cell_of(pos[3 * i], inv_cell), cell_of(pos[3 * i + 1], inv_cell),
                                 cell_of(pos[3 * i + 2], inv_cell), mask);
    slot[i] = atomicAdd(&cell_cnt[h], 1);
}

__global__ void kg_fill_kernel(const float* __restrict__ pos, const int64_t* __restrict__ batch, int64_t n,
                               const int32_t* __restrict__ meta, unsigned mask, const int32_t* __restrict__ cell_off,
                               const int32_t* __restrict__ slot, float4* __restrict__ sorted) {
    ccn_pdl_wait();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float inv_cell = __int_as_float(meta[9]);
    const float x = pos[3 * i], y = pos[3 * i + 1], z = pos[3 * i + 2];
    const unsigned h = cell_hash((int)batch[i], cell_of(x, inv_cell), cell_of(y, inv_cell), cell_of(z, inv_cell), mask);
    sorted[cell_off[h] + slot[i]] = make_float4(x, y, z, __int_as_float((int)i));
}

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
__global__ void __launch_bounds__(128) kg_search_kernel(
    const float* __restrict__ p1, const int64_t* __restrict__ batch1, int64_t n1, const float4* __restrict__ sorted,
    const int32_t* __restrict__ cell_off, const int32_t* __restrict__ sptr2, int32_t* __restrict__ meta, unsigned mask,
    int Kout, int32_t* __restrict__ open_list, int32_t* __restrict__ idx_out, float* __restrict__ d2_out,
    int32_t* __restrict__ cnt_out) {
    ccn_pdl_wait();
    const int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (q >= n1) return;
    const float h = __int_as_float(meta[8]), inv_cell = __int_as_float(meta[9]);
    const float qx = p1[3 * q], qy = p1[3 * q + 1], qz = p1[3 * q + 2];
    const int b = (int)batch1[q];
    const int lo2 = sptr2[b], hi2 = sptr2[b + 1];
    const int cx = cell_of(qx, inv_cell), cy = cell_of(qy, inv_cell), cz = cell_of(qz, inv_cell);
    unsigned long long best[K];
#pragma unroll
    for (int i = 0; i < K; ++i) best[i] = kKeyMax;
    bool done = false;
    for (int R = 1; R <= 2 && !done; ++R) {
        for (int dz = -R; dz <= R; ++dz)
            for (int dy = -R; dy <= R; ++dy)
                for (int dx = -R; dx <= R; ++dx) {
                    if (R == 2 && abs(dx) < 2 && abs(dy) < 2 && abs(dz) < 2) continue;     // examined in the first pass
                    const int tx = cx + dx, ty = cy + dy, tz = cz + dz;
                    const unsigned hh = cell_hash(b, tx, ty, tz, mask);
                    const int beg = cell_off[hh], end = cell_off[hh + 1];
                    for (int j = beg; j < end; ++j) {
                        const float4 p = sorted[j];
                        const int pi = __float_as_int(p.w);
                        if (pi < lo2 || pi >= hi2 || cell_of(p.x, inv_cell) != tx || cell_of(p.y, inv_cell) != ty ||
                            cell_of(p.z, inv_cell) != tz)
                            continue;
                        topk_insert<K>(best, ccn_pack_key(ccn_sqdist3(qx, qy, qz, p.x, p.y, p.z), pi));
                    }
                }
        if (best[Kout - 1] != kKeyMax) {
            // every point outside the cube of half-width R cells is farther than R h (the query lies inside the centre cell)
            const float kth = __uint_as_float((unsigned)(best[Kout - 1] >> 32));
            const float reach = (float)R * h * 0.9999f;
            done = kth <= reach * reach;
        }
    }
    if (!done) {
        open_list[atomicAdd(&meta[6], 1)] = (int32_t)q;
        return;
    }
#pragma unroll
    for (int i = 0; i < K; ++i) {
        if (i >= Kout) break;
        idx_out[q * Kout + i] = (int32_t)(best[i] & 0xffffffffu);
        if (d2_out) d2_out[q * Kout + i] = __uint_as_float((unsigned)(best[i] >> 32));
    }
    if (cnt_out) cnt_out[q] = Kout;
}

// brute force over the listed queries: one warp per query, lane-local K-best lists merged by a shuffle butterfly
template <int K>
__global__ void __launch_bounds__(256) kg_open_kernel(
    const float* __restrict__ p1, const int64_t* __restrict__ batch1, const float* __restrict__ p2,
    const int32_t* __restrict__ sptr2, const int32_t* __restrict__ meta, const int32_t* __restrict__ open_list, int Kout,
    int zero_fill, int32_t* __restrict__ idx_out, float* __restrict__ d2_out, int32_t* __restrict__ cnt_out) {
    ccn_pdl_wait();
    const int n_open = meta[6];
    const int lane = threadIdx.x & 31;
    for (int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < n_open; w += (gridDim.x * blockDim.x) >> 5) {
        const int64_t q = open_list[w];
        const float qx = p1[3 * q], qy = p1[3 * q + 1], qz = p1[3 * q + 2];
        const int b = (int)batch1[q];
        const int lo2 = sptr2[b], hi2 = sptr2[b + 1];
        unsigned long long best[K];
#pragma unroll
        for (int i = 0; i < K; ++i) best[i] = kKeyMax;
        for (int j = lo2 + lane; j < hi2; j += 32)
            topk_insert<K>(best, ccn_pack_key(ccn_sqdist3(qx, qy, qz, p2[3 * (int64_t)j], p2[3 * (int64_t)j + 1], p2[3 * (int64_t)j + 2]), j));
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long other[K];
#pragma unroll
            for (int i = 0; i < K; ++i) other[i] = __shfl_xor_sync(0xffffffffu, best[i], o);
#pragma unroll
            for (int i = 0; i < K; ++i) topk_insert<K>(best, other[i]);
        }
        if (lane != 0) continue;
        int cnt = 0;
#pragma unroll
        for (int i = 0; i < K; ++i) {
            if (i >= Kout) break;
            const bool have = best[i] != kKeyMax;
            const int32_t id = have ? (int32_t)(best[i] & 0xffffffffu) : (zero_fill && hi2 > lo2 ? lo2 : -1);
            idx_out[q * Kout + i] = id;
            // filled slots carry the true distance to the cloud's first point (as ccn_knn_exact)
            if (d2_out)
                d2_out[q * Kout + i] = have ? __uint_as_float((unsigned)(best[i] >> 32))
                                            : (id >= 0 ? ccn_sqdist3(qx, qy, qz, p2[3 * (int64_t)id], p2[3 * (int64_t)id + 1], p2[3 * (int64_t)id + 2]) : 0.f);
            cnt += (have || (zero_fill && hi2 > lo2)) ? 1 : 0;
        }
        if (cnt_out) cnt_out[q] = cnt;
    }
}

template <int K>
int launch_search(const float* pos1, const int64_t* batch1, int64_t n1, const float* pos2, const float4* sorted,
                  const int32_t* cell_off, const int32_t* sptr2, int32_t* meta, unsigned mask, int Kout, int zero_fill,
                  int32_t* open_list, int32_t* idx_out, float* d2_out, int32_t* cnt_out, cudaStream_t s) {
    ccn_launch(kg_search_kernel<K>, dim3(ccn_blocks_exact(n1, 128)), dim3(128), 0, s, pos1, batch1, n1, sorted, cell_off, sptr2, meta, mask, Kout,
                                                                  open_list, idx_out, d2_out, cnt_out);
    CCN_CHECK_LAUNCH();
    ccn_launch(kg_open_kernel<K>, dim3(148 * 4), dim3(256), 0, s, pos1, batch1, pos2, sptr2, meta, open_list, Kout, zero_fill, idx_out, d2_out, cnt_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

}  // namespace

extern "C" int ccn_excl_scan_i32(int32_t* a, int64_t n, int32_t* total_out, ccn_stream_t stream);

extern "C" int ccn_knn_grid(const float* pos1, const int64_t* batch1, int64_t n1, const float* pos2, const int64_t* batch2,
                            int64_t n2, const int32_t* sample_ptr2, int32_t K, int32_t zero_fill, float cell,
                            int64_t num_cells, int32_t* cell_off, int32_t* slot_scratch, float* sorted_xyzi,
                            int32_t* open_scratch, int32_t* idx_out, float* d2_out, int32_t* cnt_out, ccn_stream_t stream) {
    CCN_REQUIRE(n1 >= 0 && n2 >= 0 && n2 < 0x7fffffff && n1 < 0x7fffffff && K >= 1 && K <= 4 && idx_out && sample_ptr2);
    CCN_REQUIRE(cell_off && slot_scratch && sorted_xyzi && open_scratch);
    CCN_REQUIRE(num_cells >= 2 && (num_cells & (num_cells - 1)) == 0 && num_cells <= (1LL << 30));
    if (n1 == 0) return CCN_OK;
    cudaStream_t s = (cudaStream_t)stream;
    int32_t* meta = open_scratch;                 // 16 ints, then the open-query list (n1 ints)
    int32_t* open_list = open_scratch + 16;
    const unsigned mask = (unsigned)(num_cells - 1);
    ccn_launch(kg_init_kernel, dim3(1), dim3(32), 0, s, meta);
    CCN_CHECK_LAUNCH();
    cudaError_t e = cudaMemsetAsync(cell_off, 0, (num_cells + 1) * sizeof(int32_t), s);
    if (e != cudaSuccess) return (int)e;
    if (n2 > 0) {
        if (!(cell > 0.f)) {
            ccn_launch(kg_bbox_kernel, dim3(ccn_grid(n2, 256, 148LL * 4)), dim3(256), 0, s, pos2, n2, meta);
            CCN_CHECK_LAUNCH();
        }
        ccn_launch(kg_cell_kernel, dim3(1), dim3(1), 0, s, meta, n2, cell);
        CCN_CHECK_LAUNCH();
        ccn_launch(kg_count_kernel, dim3(ccn_blocks_exact(n2, 256)), dim3(256), 0, s, pos2, batch2, n2, meta, mask, cell_off, slot_scratch);
        CCN_CHECK_LAUNCH();
    } else {
        ccn_launch(kg_cell_kernel, dim3(1), dim3(1), 0, s, meta, 1, 1.f);
        CCN_CHECK_LAUNCH();
    }
    const int rc = ccn_excl_scan_i32(cell_off, num_cells, nullptr, stream);
    if (rc) return rc;
    if (n2 > 0) {
        ccn_launch(kg_fill_kernel, dim3(ccn_blocks_exact(n2, 256)), dim3(256), 0, s, pos2, batch2, n2, meta, mask, cell_off, slot_scratch,
                                                                (float4*)sorted_xyzi);
        CCN_CHECK_LAUNCH();
    }
    const float4* sorted = (const float4*)sorted_xyzi;
    if (K <= 2) return launch_search<2>(pos1, batch1, n1, pos2, sorted, cell_off, sample_ptr2, meta, mask, K, zero_fill, open_list, idx_out, d2_out, cnt_out, s);
    if (K == 3) return launch_search<3>(pos1, batch1, n1, pos2, sorted, cell_off, sample_ptr2, meta, mask, K, zero_fill, open_list, idx_out, d2_out, cnt_out, s);
    return launch_search<4>(pos1, batch1, n1, pos2, sorted, cell_off, sample_ptr2, meta, mask, K, zero_fill, open_list, idx_out, d2_out, cnt_out, s);
}
