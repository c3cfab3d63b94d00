// Curve grouping of an A2D2 frame (reference src/data/audi_dataset.py:156-243 `_get_curves` and
// `_partition_curves_at_discontinuities`).  A2D2 gives no beam ids, so the reference approximates scan lines per sensor (five
// lidars) on the CPU: kNN graph (torch_cluster.knn, k = 3..4) -> edges shorter than thresh * sqrt(|p.xy|) -> connected
// components (scipy) -> points ordered by (component, timestamp) (np.lexsort) -> curves cut again where both the edge to the
// previous and to the second-previous point are long -> per-sensor offsets -> np.unique compaction.
// Here, for all sensors at once: stable counting sort by sensor, ccn_knn_exact with the sensors as "clouds", a lock-free
// union-find (larger root hooks under smaller, so a component's root is its lowest point = scipy's numbering order), one
// 64-bit key (component | time offset | position) ranked inside the sensor's segment, an element-wise split kernel (f64
// edge lengths, f32 thresholds, as upstream), a per-sensor scan and a presence-bitmap compaction.
#include "common.cuh"

namespace {

constexpr int kMaxSensors = 16;
constexpr int kMaxK = 4;
constexpr int kPosBits = 17, kDtBits = 30;       // key = component (17) | timestamp - min (30) | position (17)

struct SensorCfg {
    int knn[kMaxSensors];
    float thresh[kMaxSensors];       // CURVE_THRESHES[id] in f32 (the python scalar joins an f32 tensor, audi_dataset.py:179)
    float thresh_cut[kMaxSensors];   // f32(0.9 * CURVE_THRESHES[id]) (:196, the product is formed in double)
};

__global__ void sensor_hist_kernel(const int64_t* __restrict__ lidar_id, const int64_t* __restrict__ timestamp, int64_t n,
                                   int n_sensors, int32_t* __restrict__ count, long long* __restrict__ tmin,
                                   int32_t* __restrict__ flags) {
    ccn_pdl_wait();
    long long lo = 0x7fffffffffffffffll;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t s = lidar_id[i];
        if (s < 0 || s >= n_sensors) { atomicOr(flags, 1); continue; }
        atomicAdd(count + s, 1);
        lo = min(lo, (long long)timestamp[i]);
    }
    for (int o = 16; o; o >>= 1) lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    if (ccn_lane() == 0 && lo != 0x7fffffffffffffffll) atomicMin(tmin, lo);
}

// sptr[s] = first position of sensor s in the sensor-major order (exclusive scan of the counts), one thread
__global__ void sensor_ptr_kernel(const int32_t* __restrict__ count, int n_sensors, int32_t* __restrict__ sptr) {
    ccn_pdl_wait();
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        int run = 0;
        for (int s = 0; s < n_sensors; ++s) { sptr[s] = run; run += count[s]; }
        sptr[n_sensors] = run;
    }
}

// CTA s: the points of sensor s in file order (the boolean-mask selection of audi_dataset.py:166-169)
__global__ void __launch_bounds__(1024) sensor_rank_kernel(const int64_t* __restrict__ lidar_id, int64_t n,
                                                           const int32_t* __restrict__ sptr, int32_t* __restrict__ order0) {
    ccn_pdl_wait();
    __shared__ int scan_smem[33];
    const int s = blockIdx.x;
    int running = sptr[s];
    if (sptr[s + 1] == running) return;
    for (int64_t i0 = 0; i0 < n; i0 += blockDim.x) {
        const int64_t i = i0 + threadIdx.x;
        const int mine = (i < n && lidar_id[i] == s) ? 1 : 0;
        int tot;
        const int ex = ccn_block_excl_scan<int>(mine, scan_smem, &tot);
        if (mine) order0[running + ex] = (int32_t)i;
        running += tot;
    }
}

__global__ void sensor_gather_kernel(const float* __restrict__ points, int64_t ldp, const int64_t* __restrict__ lidar_id,
                                     const int64_t* __restrict__ timestamp, const int32_t* __restrict__ order0, int64_t n,
                                     const long long* __restrict__ tmin, float* __restrict__ pts, int64_t* __restrict__ sensor,
                                     int32_t* __restrict__ dt, int32_t* __restrict__ parent, int32_t* __restrict__ flags) {
    ccn_pdl_wait();
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = order0[j];
        pts[3 * j] = points[i * ldp]; pts[3 * j + 1] = points[i * ldp + 1]; pts[3 * j + 2] = points[i * ldp + 2];
        sensor[j] = lidar_id[i];
        const long long d = (long long)timestamp[i] - *tmin;
        if (d < 0 || d >= (1ll << kDtBits)) atomicOr(flags, 2);
        dt[j] = (int32_t)d;
        parent[j] = (int32_t)j;
    }
}

__device__ __forceinline__ int uf_find(volatile int32_t* parent, int x) {
    while (true) {
        const int p = parent[x];
        if (p == x) return x;
        x = p;
    }
}

// one thread per (point, neighbour slot): keep the edge when it is shorter than thresh * sqrt(|p_query.xy|), then union
__global__ void a2d2_union_kernel(const float* __restrict__ pts, const int64_t* __restrict__ sensor,
                                  const int32_t* __restrict__ nbr, int64_t n, SensorCfg cfg, int32_t* parent) {
    ccn_pdl_wait();
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= n * kMaxK) return;
    const int64_t i = t / kMaxK;
    const int slot = (int)(t - i * kMaxK);
    const int s = (int)sensor[i];
    if (slot >= cfg.knn[s]) return;
    const int j = nbr[i * kMaxK + slot];
    if (j < 0 || j == i) return;
    const float x = pts[3 * i], y = pts[3 * i + 1], z = pts[3 * i + 2];
    const float len = ccn_norm3(__fsub_rn(x, pts[3 * j]), __fsub_rn(y, pts[3 * j + 1]), __fsub_rn(z, pts[3 * j + 2]));
    const float rad = __fsqrt_rn(__fmaf_rn(y, y, __fmul_rn(x, x)));
    if (!(len < __fmul_rn(cfg.thresh[s], __fsqrt_rn(rad)))) return;
    int a = uf_find(parent, (int)i), b = uf_find(parent, j);
    while (a != b) {
        if (a < b) { const int tmp = a; a = b; b = tmp; }
        const int old = atomicCAS(parent + a, a, b);            // hook the larger root under the smaller
        if (old == a) break;
        a = uf_find(parent, old);
        b = uf_find(parent, b);
    }
}

__global__ void a2d2_root_kernel(int32_t* parent, int64_t n, int32_t* __restrict__ root, int32_t* __restrict__ is_root) {
    ccn_pdl_wait();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int r = uf_find(parent, (int)i);
        root[i] = r;
        is_root[i] = r == i ? 1 : 0;
    }
}

// comp_rank = exclusive scan of is_root (ascending lowest point = the order scipy numbers components in)
__global__ void a2d2_key_kernel(const int32_t* __restrict__ root, const int32_t* __restrict__ comp_rank,
                                const int32_t* __restrict__ dt, int64_t n, unsigned long long* __restrict__ keys,
                                int32_t* __restrict__ comp) {
    ccn_pdl_wait();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = comp_rank[root[i]];
        comp[i] = c;
        keys[i] = ((unsigned long long)c << (kDtBits + kPosBits)) | ((unsigned long long)(unsigned)dt[i] << kPosBits) |
                  (unsigned long long)i;
    }
}

// rank of every key among the keys of the sensors its CTA touches (keys are ordered sensor-major: the component rank grows
// with the position of the component's lowest point, and sensors are contiguous segments)
__global__ void __launch_bounds__(256) a2d2_rank_kernel(const unsigned long long* __restrict__ keys,
                                                        const int64_t* __restrict__ sensor, const int32_t* __restrict__ sptr,
                                                        int64_t n, int32_t* __restrict__ rank) {
    ccn_pdl_wait();
    __shared__ unsigned long long tile[256];
    const int64_t i0 = blockIdx.x * 256ll;
    const int64_t i = i0 + threadIdx.x;
    const int lo = sptr[(int)sensor[i0]];
    const int hi = sptr[(int)sensor[min(i0 + 255, n - 1)] + 1];
    const unsigned long long mine = i < n ? keys[i] : 0ull;
    int r = 0;
    for (int t0 = lo; t0 < hi; t0 += 256) {
        __syncthreads();
        tile[threadIdx.x] = (t0 + threadIdx.x < hi) ? keys[t0 + threadIdx.x] : ~0ull;
        __syncthreads();
#pragma unroll 8
        for (int j = 0; j < 256; ++j) r += tile[j] < mine;
    }
    if (i < n) rank[i] = lo + r;
}

__global__ void a2d2_scatter_kernel(const int32_t* __restrict__ rank, const int32_t* __restrict__ order0,
                                    const float* __restrict__ pts, const int32_t* __restrict__ comp,
                                    const int64_t* __restrict__ sensor, int64_t n, int64_t* __restrict__ reorder,
                                    float* __restrict__ sorted_points, int32_t* __restrict__ comp_sorted,
                                    int32_t* __restrict__ sensor_sorted) {
    ccn_pdl_wait();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int r = rank[i];
        reorder[r] = order0[i];
        sorted_points[3 * r] = pts[3 * i]; sorted_points[3 * r + 1] = pts[3 * i + 1]; sorted_points[3 * r + 2] = pts[3 * i + 2];
        comp_sorted[r] = comp[i];
        sensor_sorted[r] = (int32_t)sensor[i];
    }
}

__device__ __forceinline__ double edge_len_f64(const float* __restrict__ a, const float* __restrict__ b) {
    const double ex = (double)a[0] - (double)b[0], ey = (double)a[1] - (double)b[1], ez = (double)a[2] - (double)b[2];
    return sqrt(fma(ez, ez, fma(ey, ey, ex * ex)));
}

// _partition_curves_at_discontinuities (audi_dataset.py:214-237), per sensor segment: a cut where the component changes, or
// where the edges to the previous AND to the second-previous point (10 where there is none) both exceed the threshold
__global__ void a2d2_split_kernel(const float* __restrict__ p, const int32_t* __restrict__ comp,
                                  const int32_t* __restrict__ sensor, const int32_t* __restrict__ sptr, int64_t n,
                                  SensorCfg cfg, int32_t* __restrict__ split) {
    ccn_pdl_wait();
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const int s = sensor[r];
        const int64_t k = r - sptr[s];
        if (sptr[s + 1] - sptr[s] == 1) { split[r] = 1; continue; }      // single point: its component id 0 stays (:195)
        const double e1 = k >= 1 ? edge_len_f64(p + 3 * r, p + 3 * (r - 1)) : 10.0;
        const double e2 = k >= 2 ? edge_len_f64(p + 3 * r, p + 3 * (r - 2)) : 10.0;
        const float x = p[3 * r], y = p[3 * r + 1];
        const float rad = __fsqrt_rn(__fmaf_rn(y, y, __fmul_rn(x, x)));
        const double thr = (double)__fmul_rn(cfg.thresh_cut[s], __fsqrt_rn(rad));
        const bool cut3d = e1 > thr && e2 > thr;
        const bool edge = k >= 1 && comp[r] != comp[r - 1];
        split[r] = (cut3d || edge) ? 1 : 0;
    }
}

// CTA s: local id = inclusive running sum of split - 1 over the sensor's segment; sensor_max[s] = its last (largest) id
__global__ void __launch_bounds__(1024) a2d2_local_id_kernel(const int32_t* __restrict__ split, const int32_t* __restrict__ sptr,
                                                             int32_t* __restrict__ local_id, int32_t* __restrict__ sensor_max) {
    ccn_pdl_wait();
    __shared__ int scan_smem[33];
    const int s = blockIdx.x;
    const int lo = sptr[s], hi = sptr[s + 1];
    int running = 0;
    for (int i0 = lo; i0 < hi; i0 += blockDim.x) {
        const int i = i0 + threadIdx.x;
        const int v = i < hi ? split[i] : 0;
        int tot;
        const int ex = ccn_block_excl_scan<int>(v, scan_smem, &tot);
        if (i < hi) local_id[i] = running + ex + v - 1;
        running += tot;
    }
    if (threadIdx.x == 0) sensor_max[s] = hi > lo ? running - 1 : 0;
}

// id = local + sum of the max ids of the sensors before (total_curves += max, NOT max + 1, audi_dataset.py:197,205: the first
// curve of a sensor shares its id with the last curve of the sensor before - kept); mark presence[id + 1]
__global__ void a2d2_offset_kernel(const int32_t* __restrict__ local_id, const int32_t* __restrict__ sensor,
                                   const int32_t* __restrict__ sensor_max, const int32_t* __restrict__ sptr, int n_sensors,
                                   int64_t n, int32_t* __restrict__ gid, int32_t* __restrict__ presence) {
    ccn_pdl_wait();
    __shared__ int off[kMaxSensors + 1];
    if (threadIdx.x == 0) {
        int run = 0;
        for (int s = 0; s < n_sensors; ++s) { off[s] = run; if (sptr[s + 1] > sptr[s]) run += sensor_max[s]; }
    }
    __syncthreads();
    const int n_valid = sptr[n_sensors];        // < n only when a lidar id was out of range (flag 1): those slots hold no point
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        if (r >= n_valid) { gid[r] = -1; continue; }
        const int g = local_id[r] + off[sensor[r]];
        gid[r] = g;
        presence[g + 1] = 1;
    }
}

__global__ void a2d2_compact_kernel(const int32_t* __restrict__ gid, const int32_t* __restrict__ presence_scan, int64_t n,
                                    int64_t* __restrict__ curve_idxs) {
    ccn_pdl_wait();
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x)
        curve_idxs[r] = presence_scan[gid[r] + 1];
}

// scratch layout (bytes, every block 16-byte aligned)
struct Layout {
    size_t count, sptr, tmin, order0, pts, sensor, dt, parent, nbr, root, is_root, keys, comp, rank, comp_sorted,
        sensor_sorted, split, local_id, sensor_max, gid, presence, total;
};

Layout make_layout(int64_t n, int n_sensors) {
    Layout L;
    size_t at = 0;
    auto take = [&](size_t bytes) { const size_t here = at; at += (bytes + 15) & ~(size_t)15; return here; };
    const size_t m = (size_t)(n > 0 ? n : 1);
    L.count = take(4 * (size_t)(n_sensors + 1));
    L.sptr = take(4 * (size_t)(n_sensors + 2));
    L.tmin = take(8);
    L.order0 = take(4 * m);
    L.pts = take(12 * m);
    L.sensor = take(8 * m);
    L.dt = take(4 * m);
    L.parent = take(4 * m);
    L.nbr = take(4 * kMaxK * m);
    L.root = take(4 * m);
    L.is_root = take(4 * (m + 2));
    L.keys = take(8 * m);
    L.comp = take(4 * m);
    L.rank = take(4 * m);
    L.comp_sorted = take(4 * m);
    L.sensor_sorted = take(4 * m);
    L.split = take(4 * m);
    L.local_id = take(4 * m);
    L.sensor_max = take(4 * (size_t)(n_sensors + 1));
    L.gid = take(4 * m);
    L.presence = take(4 * (m + 4));
    L.total = at;
    return L;
}

}  // namespace

extern "C" int64_t ccn_a2d2_scratch_bytes(int64_t n, int32_t n_sensors) {
    if (n < 0 || n_sensors < 1 || n_sensors > kMaxSensors) return -1;
    return (int64_t)make_layout(n, n_sensors).total;
}

extern "C" int ccn_a2d2_curves(const float* points, int64_t ldp, const int64_t* lidar_id, const int64_t* timestamp, int64_t n,
                               int32_t n_sensors, const int32_t* knn_host, const float* thresh_host, void* scratch,
                               int64_t* reorder, float* sorted_points, int64_t* curve_idxs, int32_t* counts_out,
                               ccn_stream_t stream) {
    // knn_host / thresh_host: n_sensors HOST values (KNN, CURVE_THRESHES of audi_dataset.py:19-20).  counts_out: int32 [2] =
    // {number of curves, error flags: 1 lidar id out of range, 2 a timestamp more than 2^30 ticks after the first}
    CCN_REQUIRE(n >= 0 && n < (1ll << kPosBits) && n_sensors >= 1 && n_sensors <= kMaxSensors && ldp >= 3);
    CCN_REQUIRE(knn_host && thresh_host && scratch && reorder && sorted_points && curve_idxs && counts_out);
    SensorCfg cfg;
    for (int s = 0; s < kMaxSensors; ++s) {
        cfg.knn[s] = s < n_sensors ? knn_host[s] : 0;
        cfg.thresh[s] = s < n_sensors ? thresh_host[s] : 0.f;
        cfg.thresh_cut[s] = s < n_sensors ? (float)(0.9 * (double)thresh_host[s]) : 0.f;
        CCN_REQUIRE(cfg.knn[s] >= 0 && cfg.knn[s] <= kMaxK);
    }
    cudaStream_t st = (cudaStream_t)stream;
    const Layout L = make_layout(n, n_sensors);
    char* base = (char*)scratch;
    CCN_REQUIRE(((uintptr_t)base & 15) == 0);
    int32_t* count = (int32_t*)(base + L.count);
    int32_t* sptr = (int32_t*)(base + L.sptr);
    long long* tmin = (long long*)(base + L.tmin);
    int32_t* order0 = (int32_t*)(base + L.order0);
    float* pts = (float*)(base + L.pts);
    int64_t* sensor = (int64_t*)(base + L.sensor);
    int32_t* dt = (int32_t*)(base + L.dt);
    int32_t* parent = (int32_t*)(base + L.parent);
    int32_t* nbr = (int32_t*)(base + L.nbr);
    int32_t* root = (int32_t*)(base + L.root);
    int32_t* is_root = (int32_t*)(base + L.is_root);
    unsigned long long* keys = (unsigned long long*)(base + L.keys);
    int32_t* comp = (int32_t*)(base + L.comp);
    int32_t* rank = (int32_t*)(base + L.rank);
    int32_t* comp_sorted = (int32_t*)(base + L.comp_sorted);
    int32_t* sensor_sorted = (int32_t*)(base + L.sensor_sorted);
    int32_t* split = (int32_t*)(base + L.split);
    int32_t* local_id = (int32_t*)(base + L.local_id);
    int32_t* sensor_max = (int32_t*)(base + L.sensor_max);
    int32_t* gid = (int32_t*)(base + L.gid);
    int32_t* presence = (int32_t*)(base + L.presence);

    cudaError_t e = cudaMemsetAsync(count, 0, 4 * (size_t)(n_sensors + 1), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(tmin, 0x7f, 8, st);                  // 0x7f7f...: above every real timestamp
    if (e == cudaSuccess) e = cudaMemsetAsync(counts_out, 0, 2 * sizeof(int32_t), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(presence, 0, 4 * (size_t)(n + 4), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(order0, 0, 4 * (size_t)(n > 0 ? n : 1), st);   // a lidar id out of range (flag 1) leaves holes
    if (e != cudaSuccess) return (int)e;
    if (n == 0) return CCN_OK;
    const unsigned eg = ccn_grid(n, 256, 148 * 8);
    ccn_launch(sensor_hist_kernel, dim3(eg), dim3(256), 0, st, lidar_id, timestamp, n, (int)n_sensors, count, tmin, counts_out + 1);
    CCN_CHECK_LAUNCH();
    ccn_launch(sensor_ptr_kernel, dim3(1), dim3(32), 0, st, (const int32_t*)count, (int)n_sensors, sptr);
    CCN_CHECK_LAUNCH();
    ccn_launch(sensor_rank_kernel, dim3((unsigned)n_sensors), dim3(1024), 0, st, lidar_id, n, (const int32_t*)sptr, order0);
    CCN_CHECK_LAUNCH();
    ccn_launch(sensor_gather_kernel, dim3(eg), dim3(256), 0, st, points, ldp, lidar_id, timestamp, (const int32_t*)order0, n,
               (const long long*)tmin, pts, sensor, dt, parent, counts_out + 1);
    CCN_CHECK_LAUNCH();
    int rc = ccn_knn_exact(pts, sensor, n, pts, sptr, kMaxK, 0, nbr, nullptr, nullptr, stream);
    if (rc != CCN_OK) return rc;
    ccn_launch(a2d2_union_kernel, dim3(ccn_blocks_exact(n * kMaxK, 256)), dim3(256), 0, st, (const float*)pts,
               (const int64_t*)sensor, (const int32_t*)nbr, n, cfg, parent);
    CCN_CHECK_LAUNCH();
    ccn_launch(a2d2_root_kernel, dim3(eg), dim3(256), 0, st, parent, n, root, is_root);
    CCN_CHECK_LAUNCH();
    rc = ccn_excl_scan_i32(is_root, n, nullptr, stream);
    if (rc != CCN_OK) return rc;
    ccn_launch(a2d2_key_kernel, dim3(eg), dim3(256), 0, st, (const int32_t*)root, (const int32_t*)is_root, (const int32_t*)dt, n,
               keys, comp);
    CCN_CHECK_LAUNCH();
    ccn_launch(a2d2_rank_kernel, dim3(ccn_blocks_exact(n, 256)), dim3(256), 0, st, (const unsigned long long*)keys,
               (const int64_t*)sensor, (const int32_t*)sptr, n, rank);
    CCN_CHECK_LAUNCH();
    ccn_launch(a2d2_scatter_kernel, dim3(eg), dim3(256), 0, st, (const int32_t*)rank, (const int32_t*)order0, (const float*)pts,
               (const int32_t*)comp, (const int64_t*)sensor, n, reorder, sorted_points, comp_sorted, sensor_sorted);
    CCN_CHECK_LAUNCH();
    ccn_launch(a2d2_split_kernel, dim3(eg), dim3(256), 0, st, (const float*)sorted_points, (const int32_t*)comp_sorted,
               (const int32_t*)sensor_sorted, (const int32_t*)sptr, n, cfg, split);
    CCN_CHECK_LAUNCH();
    ccn_launch(a2d2_local_id_kernel, dim3((unsigned)n_sensors), dim3(1024), 0, st, (const int32_t*)split, (const int32_t*)sptr,
               local_id, sensor_max);
    CCN_CHECK_LAUNCH();
    ccn_launch(a2d2_offset_kernel, dim3(eg), dim3(256), 0, st, (const int32_t*)local_id, (const int32_t*)sensor_sorted,
               (const int32_t*)sensor_max, (const int32_t*)sptr, (int)n_sensors, n, gid, presence);
    CCN_CHECK_LAUNCH();
    rc = ccn_excl_scan_i32(presence, n + 2, counts_out, stream);
    if (rc != CCN_OK) return rc;
    ccn_launch(a2d2_compact_kernel, dim3(eg), dim3(256), 0, st, (const int32_t*)gid, (const int32_t*)presence, n, curve_idxs);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

namespace {
// the tail of SemA2D2.get (audi_dataset.py:131-141) that is not position arithmetic: reflectance / 255 and the labels, in
// the order _get_curves produced
__global__ void a2d2_finish_kernel(const float* __restrict__ refl, const int64_t* __restrict__ labels,
                                   const int64_t* __restrict__ reorder, int64_t n, float refl_div, float* __restrict__ x_out,
                                   int64_t* __restrict__ y_out) {
    ccn_pdl_wait();
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = reorder[r];
        x_out[r] = __fdiv_rn(refl[i], refl_div);
        if (y_out) y_out[r] = labels ? labels[i] : 0;
    }
}
}  // namespace

extern "C" int ccn_a2d2_finish(const float* reflectance, const int64_t* labels, const int64_t* reorder, int64_t n,
                               float refl_div, float* x_out, int64_t* y_out, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && refl_div != 0.f);
    if (n == 0) return CCN_OK;
    CCN_REQUIRE(reflectance && reorder && x_out);
    ccn_launch(a2d2_finish_kernel, dim3(ccn_grid(n, 256, 148 * 8)), dim3(256), 0, (cudaStream_t)stream, reflectance, labels, reorder, n,
               refl_div, x_out, y_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
