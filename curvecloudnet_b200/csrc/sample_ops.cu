// Point down-sampling:
//   ccn_voxel_fps   VoxelFPS.forward                    reference src/models/modules/fps_ops.py:51-60
//   ccn_fps         pytorch3d.ops.sample_farthest_points reference call site src/models/utils/point_ops.py:57-70
// The reference voxel sampler sorts all N (batch,ix,iy,iz) rows (torch.unique(dim=0)) and runs an
// atomics scatter_min; here occupied voxels are found with a 64-bit hash table that also carries
// the per-voxel argmin as one atomicMin on a packed (score, index) key, and only the occupied
// voxels (N2 << N) are ordered.  FPS is one CTA per cloud with points and running distances in
// registers and a 2-instruction (redux.sync) warp argmax per round.
#include <mutex>

#include "common.cuh"

namespace {

constexpr unsigned long long kEmpty = 0xffffffffffffffffull;
constexpr int kCellBias = 1 << 17;

__device__ __forceinline__ unsigned long long mix64(unsigned long long k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
    return k;
}

__global__ void voxel_insert_kernel(const float* __restrict__ pos, const int64_t* __restrict__ batch, int64_t n,
                                    float voxel, const float* __restrict__ noise, unsigned long long tmask,
                                    unsigned long long* __restrict__ tkeys, unsigned long long* __restrict__ tmin,
                                    int32_t* __restrict__ err) {
    ccn_pdl_wait();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    // pos / voxel, floor, |floor - pos/voxel| + noise * voxel / 4   (fps_ops.py:52-57)
    const float vx = __fdiv_rn(pos[3 * i], voxel), vy = __fdiv_rn(pos[3 * i + 1], voxel),
                vz = __fdiv_rn(pos[3 * i + 2], voxel);
    const float fx = floorf(vx), fy = floorf(vy), fz = floorf(vz);
    const long long b = batch[i];
    const long long cx = (long long)fx + kCellBias, cy = (long long)fy + kCellBias, cz = (long long)fz + kCellBias;
    if (b < 0 || b >= 1024 || cx < 0 || cy < 0 || cz < 0 || cx >= 2 * kCellBias || cy >= 2 * kCellBias ||
        cz >= 2 * kCellBias) {
        atomicOr(err, 1);
        return;
    }
    // lexicographic (batch, ix, iy, iz) order == unsigned order of the packed key
    const unsigned long long key = ((unsigned long long)b << 54) | ((unsigned long long)cx << 36) |
                                   ((unsigned long long)cy << 18) | (unsigned long long)cz;
    float d = ccn_norm3(__fsub_rn(fx, vx), __fsub_rn(fy, vy), __fsub_rn(fz, vz));
    d = __fadd_rn(d, __fdiv_rn(__fmul_rn(noise[i], voxel), 4.f));
    const unsigned long long score = ccn_pack_key(d, (int)i);
    unsigned long long slot = mix64(key) & tmask;
    while (true) {
        const unsigned long long prev = atomicCAS(&tkeys[slot], kEmpty, key);
        if (prev == kEmpty || prev == key) {
            atomicMin(&tmin[slot], score);
            return;
        }
        slot = (slot + 1) & tmask;
    }
}

__global__ void voxel_compact_kernel(const unsigned long long* __restrict__ tkeys,
                                     const unsigned long long* __restrict__ tmin, int64_t slots,
                                     unsigned long long* __restrict__ ukeys, int32_t* __restrict__ uwin,
                                     int32_t* __restrict__ count) {
    ccn_pdl_wait();
    const int64_t s = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (s >= slots) return;
    const unsigned long long k = tkeys[s];
    if (k == kEmpty) return;
    const int at = atomicAdd(count, 1);
    ukeys[at] = k;
    uwin[at] = (int32_t)(tmin[s] & 0xffffffffu);
}

// rank[i] = #{j : key[j] < key[i]} over the unique keys; blockIdx.y walks candidate chunks.
constexpr int kRankChunk = 2048;
__global__ void __launch_bounds__(256) rank_keys_kernel(const unsigned long long* __restrict__ keys,
                                                        const int32_t* __restrict__ count_ptr,
                                                        int32_t* __restrict__ rank) {
    ccn_pdl_wait();
    __shared__ unsigned long long tile[256];
    const int n = *count_ptr;
    const int i = blockIdx.x * 256 + threadIdx.x;
    const int c0 = blockIdx.y * kRankChunk;
    if (blockIdx.x * 256 >= n || c0 >= n) return;
    const unsigned long long mine = i < n ? keys[i] : 0ull;
    int r = 0;
    const int cend = min(n, c0 + kRankChunk);
    for (int t0 = c0; t0 < cend; t0 += 256) {
        __syncthreads();
        tile[threadIdx.x] = (t0 + threadIdx.x < cend) ? keys[t0 + threadIdx.x] : kEmpty;
        __syncthreads();
#pragma unroll 8
        for (int j = 0; j < 256; ++j) r += tile[j] < mine;
    }
    if (i < n && r) atomicAdd(&rank[i], r);
}

__global__ void scatter_by_rank_kernel(const int32_t* __restrict__ rank, const int32_t* __restrict__ uwin,
                                       const int32_t* __restrict__ count_ptr, int64_t* __restrict__ idx_out) {
    ccn_pdl_wait();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < *count_ptr) idx_out[rank[i]] = uwin[i];
}

// ---------------------------------------------------------------------------------------------
// FPS: one CTA per cloud.  Thread t owns points t, t+T, t+2T, ... (ascending, so the per-thread
// running argmax keeps the lowest index among equal distances); warp argmax = redux.max on the
// distance bits + redux.min on the index of the lanes that attain it.
// ---------------------------------------------------------------------------------------------
constexpr int kFpsThreads = 512;

struct FpsPartial { unsigned bits; int idx; float x, y, z; };

// THREADS: the K selection rounds are sequential; a round = every thread's serial argmax over its own points + one block-wide
// argmax, so the block is as WIDE as the cloud allows (2 points per thread up to 2048 points).
template <int PPT, int THREADS>
__global__ void __launch_bounds__(THREADS) fps_kernel(const float* __restrict__ pos,
                                                          const int32_t* __restrict__ sptr,
                                                          const int32_t* __restrict__ ksel,
                                                          const int32_t* __restrict__ start,
                                                          float* __restrict__ closest_scratch,
                                                          int32_t* __restrict__ sel_cnt) {
    ccn_pdl_wait();
    __shared__ FpsPartial part[2][THREADS / 32];
    const int b = blockIdx.x;
    const int lo = sptr[b], len = sptr[b + 1] - lo;
    const int k = min(ksel[b], len);
    if (k <= 0) return;
    const int lane = ccn_lane(), warp = ccn_warp();
    float px[PPT > 0 ? PPT : 1], py[PPT > 0 ? PPT : 1], pz[PPT > 0 ? PPT : 1], cl[PPT > 0 ? PPT : 1];
    if (PPT > 0) {
#pragma unroll
        for (int j = 0; j < PPT; ++j) {
            const int i = threadIdx.x + j * THREADS;
            if (i < len) { px[j] = pos[3 * (lo + i)]; py[j] = pos[3 * (lo + i) + 1]; pz[j] = pos[3 * (lo + i) + 2]; }
            else { px[j] = py[j] = pz[j] = 0.f; }
            cl[j] = __int_as_float(0x7f800000);
        }
    } else {
        for (int i = threadIdx.x; i < len; i += THREADS) closest_scratch[lo + i] = __int_as_float(0x7f800000);
    }
    int sel = start[b];
    float sx = pos[3 * (lo + sel)], sy = pos[3 * (lo + sel) + 1], sz = pos[3 * (lo + sel) + 2];
    if (threadIdx.x == 0) atomicAdd(&sel_cnt[lo + sel], 1);
    for (int it = 1; it < k; ++it) {
        unsigned bbits = 0;
        int bidx = 0x7fffffff;
        float bx = 0.f, by = 0.f, bz = 0.f;
        if (PPT > 0) {
#pragma unroll
            for (int j = 0; j < PPT; ++j) {
                const int i = threadIdx.x + j * THREADS;
                if (i < len) {
                    const float d = ccn_sqdist3(px[j], py[j], pz[j], sx, sy, sz);
                    const float c = fminf(cl[j], d);
                    cl[j] = c;
                    const unsigned cb = __float_as_uint(c);
                    if (bidx == 0x7fffffff || cb > bbits) { bbits = cb; bidx = i; bx = px[j]; by = py[j]; bz = pz[j]; }
                }
            }
        } else {
            for (int i = threadIdx.x; i < len; i += THREADS) {
                const float x = pos[3 * (lo + i)], y = pos[3 * (lo + i) + 1], z = pos[3 * (lo + i) + 2];
                const float c = fminf(closest_scratch[lo + i], ccn_sqdist3(x, y, z, sx, sy, sz));
                closest_scratch[lo + i] = c;
                const unsigned cb = __float_as_uint(c);
                if (bidx == 0x7fffffff || cb > bbits) { bbits = cb; bidx = i; bx = x; by = y; bz = z; }
            }
        }
        // warp argmax (threads without points carry idx = INT_MAX and bits = 0)
        const bool has = bidx != 0x7fffffff;
        const unsigned wmax = __reduce_max_sync(0xffffffffu, has ? bbits : 0u);
        const int widx = __reduce_min_sync(0xffffffffu, (has && bbits == wmax) ? bidx : 0x7fffffff);
        const int buf = it & 1;
        if (has && bidx == widx) { part[buf][warp] = FpsPartial{wmax, widx, bx, by, bz}; }
        else if (widx == 0x7fffffff && lane == 0) { part[buf][warp] = FpsPartial{0u, 0x7fffffff, 0.f, 0.f, 0.f}; }
        __syncthreads();
        FpsPartial p = part[buf][lane < THREADS / 32 ? lane : 0];
        const bool phas = lane < THREADS / 32 && p.idx != 0x7fffffff;
        const unsigned gmax = __reduce_max_sync(0xffffffffu, phas ? p.bits : 0u);
        const int gidx = __reduce_min_sync(0xffffffffu, (phas && p.bits == gmax) ? p.idx : 0x7fffffff);
        const unsigned owner = __ballot_sync(0xffffffffu, phas && p.bits == gmax && p.idx == gidx);
        const int src = __ffs(owner) - 1;
        sel = gidx;
        sx = __shfl_sync(0xffffffffu, p.x, src);
        sy = __shfl_sync(0xffffffffu, p.y, src);
        sz = __shfl_sync(0xffffffffu, p.z, src);
        if (threadIdx.x == 0) atomicAdd(&sel_cnt[lo + sel], 1);
    }
}

// Ordered expansion of a multiplicity vector: emits i cnt[i] times, ascending (single CTA).
__global__ void __launch_bounds__(1024) expand_counts_kernel(const int32_t* __restrict__ cnt, int64_t n,
                                                             int64_t* __restrict__ idx_out,
                                                             int32_t* __restrict__ total_out) {
    ccn_pdl_wait();
    __shared__ int sm[33];
    int carry = 0;
    for (int64_t base = 0; base < n; base += 1024) {
        const int64_t i = base + threadIdx.x;
        const int c = i < n ? cnt[i] : 0;
        int total;
        int o = carry + ccn_block_excl_scan<int>(c, sm, &total);
        for (int t = 0; t < c; ++t) idx_out[o + t] = i;
        carry += total;
    }
    if (threadIdx.x == 0 && total_out) *total_out = carry;
}

}  // namespace

extern "C" int64_t ccn_voxel_table_slots(int64_t n) {
    int64_t m = 1024;
    while (m < 2 * n) m <<= 1;
    return m;
}

// Scratch layout (caller-allocated): table_keys (slots u64), table_min (slots u64), ukeys (n u64),
// uwin (n i32), rank (n i32), counts (2 i32: [0] = #voxels, [1] = error flag).
extern "C" int ccn_voxel_fps(const float* pos, const int64_t* batch, int64_t n, float voxel, const float* noise,
                             int64_t slots, uint64_t* table_keys, uint64_t* table_min, uint64_t* ukeys,
                             int32_t* uwin, int32_t* rank, int64_t* idx_out, int32_t* counts, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && n < 0x7fffffff && voxel > 0.f && slots >= 2 && (slots & (slots - 1)) == 0 && slots >= 2 * n);
    CCN_REQUIRE(table_keys && table_min && ukeys && uwin && rank && idx_out && counts);
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e;
    if ((e = cudaMemsetAsync(table_keys, 0xff, slots * 8, s)) != cudaSuccess) return (int)e;
    if ((e = cudaMemsetAsync(table_min, 0xff, slots * 8, s)) != cudaSuccess) return (int)e;
    if ((e = cudaMemsetAsync(rank, 0, (n > 0 ? n : 1) * 4, s)) != cudaSuccess) return (int)e;
    if ((e = cudaMemsetAsync(counts, 0, 2 * 4, s)) != cudaSuccess) return (int)e;
    if (n == 0) return CCN_OK;
    ccn_launch(voxel_insert_kernel, dim3(ccn_blocks_exact(n, 256)), dim3(256), 0, s, pos, batch, n, voxel, noise, (unsigned long long)(slots - 1),
                                                                (unsigned long long*)table_keys,
                                                                (unsigned long long*)table_min, counts + 1);
    CCN_CHECK_LAUNCH();
    ccn_launch(voxel_compact_kernel, dim3(ccn_blocks_exact(slots, 256)), dim3(256), 0, s, (unsigned long long*)table_keys,
                                                                     (unsigned long long*)table_min, slots,
                                                                     (unsigned long long*)ukeys, uwin, counts);
    CCN_CHECK_LAUNCH();
    dim3 grid(ccn_blocks_exact(n, 256), ccn_blocks_exact(n, kRankChunk));
    ccn_launch(rank_keys_kernel, dim3(grid), dim3(256), 0, s, (unsigned long long*)ukeys, counts, rank);
    CCN_CHECK_LAUNCH();
    ccn_launch(scatter_by_rank_kernel, dim3(ccn_blocks_exact(n, 256)), dim3(256), 0, s, rank, uwin, counts, idx_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_fps(const float* pos, const int32_t* sample_ptr, int32_t n_samples, int64_t n,
                       const int32_t* ksel, const int32_t* start, int32_t max_len, float* closest_scratch,
                       int32_t* sel_cnt_scratch, int64_t* idx_out, int32_t* total_out, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && n_samples >= 0 && sample_ptr && ksel && start && sel_cnt_scratch && idx_out);
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(sel_cnt_scratch, 0, (n > 0 ? n : 1) * 4, s);
    if (e != cudaSuccess) return (int)e;
    if (n_samples > 0 && n > 0) {
#define CCN_FPS_LAUNCH(PPT, T) ccn_launch(fps_kernel<PPT, T>, dim3(n_samples), dim3(T), 0, s, pos, sample_ptr, ksel, start, closest_scratch, sel_cnt_scratch)
        // measured (tools/bench_fps.py, profiles/r02o_fps_variants.txt): a round costs ~0.5 us whatever the split - 4 warps x 16
        // points 0.95 us, 32 warps x 2 points 0.60 us, 16 warps x 4 points 0.50 us (kept): the round is the latency chain
        // distance -> warp argmax -> shared memory -> barrier -> block argmax, not arithmetic, so a CTA cluster per cloud (a
        // slower barrier per round) cannot shorten it; at B = 1 the executor hides the sampler behind the edge steps instead
        if (max_len <= 1024) CCN_FPS_LAUNCH(2, 512);
        else if (max_len <= 2048) CCN_FPS_LAUNCH(4, 512);
        else if (max_len <= 4096) CCN_FPS_LAUNCH(8, 512);
        else if (max_len <= 8192) CCN_FPS_LAUNCH(16, 512);
        else {
            CCN_REQUIRE(closest_scratch);
            CCN_FPS_LAUNCH(0, 1024);
        }
#undef CCN_FPS_LAUNCH
        CCN_CHECK_LAUNCH();
    }
    ccn_launch(expand_counts_kernel, dim3(1), dim3(1024), 0, s, sel_cnt_scratch, n, idx_out, total_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// ccn_fps_static: ccn_fps with its scratch in a file-scope device array - no allocation anywhere on the call path (no
// counterpart in the reference; for callers under CUDA graph capture / serving loops that may not call cudaMalloc).
// Layout: [0, 4 MAX) the running squared distances (float, read only by the > 8192-point variant), [4 MAX, 8 MAX) the
// selection flags (int32, cleared by ccn_fps at every call).  One array per device that loads the library; every call on a
// device shares it, calls that overlap on the GPU race (include/ccn_b200.h).
// In HOST code `g_fps_static_scratch` is only a stand-in for the symbol, not a device address: the address is looked up once
// per device, with that device current, through cudaGetSymbolAddress.
// ---------------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_fps_static_scratch[(size_t)CCN_FPS_STATIC_MAX_POINTS * 8];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_fps_static_mutex;
void* g_fps_static_addr[kMaxDevices] = {};
}  // namespace

extern "C" int ccn_fps_static(const float* pos, const int32_t* sample_ptr, int32_t n_samples, int64_t n, const int32_t* ksel,
                              const int32_t* start, int32_t max_len, int64_t* idx_out, int32_t* total_out, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0);
    if (n > CCN_FPS_STATIC_MAX_POINTS) return CCN_ESCRATCH_LIMIT;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    CCN_REQUIRE(dev >= 0 && dev < kMaxDevices);
    void* base = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_fps_static_mutex);
        if (!g_fps_static_addr[dev]) {                       // set-up, once per device, with that device current
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_fps_static_scratch);
            if (e != cudaSuccess) return (int)e;
            if (!p) return CCN_EINVAL;
            g_fps_static_addr[dev] = p;
        }
        base = g_fps_static_addr[dev];
    }
    float* closest = reinterpret_cast<float*>(base);
    int32_t* sel_cnt = reinterpret_cast<int32_t*>(static_cast<unsigned char*>(base) + (size_t)CCN_FPS_STATIC_MAX_POINTS * 4);
    return ccn_fps(pos, sample_ptr, n_samples, n, ksel, start, max_len, closest, sel_cnt, idx_out, total_out, stream);
}

