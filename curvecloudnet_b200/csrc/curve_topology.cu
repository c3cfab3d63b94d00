// Curve / sample CSR topology of a ragged batch of scans.
// Replaces curveidx_local2global / batch2ptr (reference src/models/utils/point_ops.py:20-54),
// which the reference recomputes (with a torch.where host sync) in every curve step.
#include "common.cuh"

namespace {

constexpr int kItems = 4;

// Multi-CTA scan over the sorted (batch, curve) id vectors in four small launches (a single CTA serialised over the
// whole batch: 0.27 ms at 6 x 32 k points).  Start flags are packed in one 64-bit word: hi 32 bits count sample starts,
// lo 32 bits curve starts.
//   1. tile_flags_kernel   per 4096-point tile: sum of the packed flags -> tile_sum[t]; sortedness errors -> counts[2]
//   2. tile_scan_kernel    carry = sum of tile_sum[0..t), then the tile's own scan: ord / cptr / sptr; the last tile writes
//                          the end pointers and {n_samples, n_curves}
//   3. sample_offsets_kernel (one CTA)  soff[b] = sum over b' < b of (last curve id of b' + 1)  (point_ops.py:32-35)
//   4. glob_ids_kernel     glob[i] = curve[i] + soff[batch[i]]
// tile_sum lives in the head of the glob output buffer (8 bytes per tile <= 8 bytes per point), which step 4 overwrites.
constexpr int kTile = 1024 * kItems;

__device__ __forceinline__ unsigned long long point_flags(const int64_t* __restrict__ batch, const int64_t* __restrict__ curve,
                                                          int64_t i, int& err) {
    const int64_t b = batch[i], c = curve[i];
    bool ss = true, cs = true;
    if (i > 0) {
        const int64_t pb = batch[i - 1], pc = curve[i - 1];
        ss = b != pb;
        cs = ss || c != pc;
        if (b < pb) err |= CCN_TOPO_ERR_BATCH_UNSORTED;
        if (!ss && c < pc) err |= CCN_TOPO_ERR_CURVE_UNSORTED;
    }
    return ((unsigned long long)ss << 32) | (unsigned long long)cs;
}

__global__ void __launch_bounds__(1024) tile_flags_kernel(const int64_t* __restrict__ batch, const int64_t* __restrict__ curve,
                                                          int64_t n, unsigned long long* __restrict__ tile_sum,
                                                          int32_t* __restrict__ counts) {
    ccn_pdl_wait();
    __shared__ unsigned long long sm[33];
    const int64_t i0 = (int64_t)blockIdx.x * kTile + (int64_t)threadIdx.x * kItems;
    unsigned long long local = 0;
    int err = 0;
#pragma unroll
    for (int e = 0; e < kItems; ++e)
        if (i0 + e < n) local += point_flags(batch, curve, i0 + e, err);
    unsigned long long total;
    ccn_block_excl_scan<unsigned long long>(local, sm, &total);
    if (threadIdx.x == 0) tile_sum[blockIdx.x] = total;
    if (err) atomicOr(&counts[2], err);
}

__global__ void __launch_bounds__(1024) tile_scan_kernel(
    const int64_t* __restrict__ batch, const int64_t* __restrict__ curve, int64_t n,
    const unsigned long long* __restrict__ tile_sum, int32_t* __restrict__ ord, int32_t* __restrict__ cptr,
    int32_t* __restrict__ sptr, int32_t* __restrict__ counts, int max_samples) {
    ccn_pdl_wait();
    __shared__ unsigned long long sm[33];
    // carry of this tile = sum of the earlier tiles' totals
    unsigned long long part = 0;
    for (int t = threadIdx.x; t < (int)blockIdx.x; t += blockDim.x) part += tile_sum[t];
    unsigned long long carry;
    ccn_block_excl_scan<unsigned long long>(part, sm, &carry);
    const int64_t i0 = (int64_t)blockIdx.x * kTile + (int64_t)threadIdx.x * kItems;
    unsigned long long f[kItems], local = 0;
    int err = 0, dummy = 0;
#pragma unroll
    for (int e = 0; e < kItems; ++e) {
        f[e] = (i0 + e < n) ? point_flags(batch, curve, i0 + e, dummy) : 0ull;
        local += f[e];
    }
    unsigned long long total;
    unsigned long long run = carry + ccn_block_excl_scan<unsigned long long>(local, sm, &total);
#pragma unroll
    for (int e = 0; e < kItems; ++e) {
        const int64_t i = i0 + e;
        if (i < n) {
            run += f[e];
            const int c_ord = (int)(run & 0xffffffffu) - 1;
            const int s_ord = (int)(run >> 32) - 1;
            ord[i] = c_ord;
            if (f[e] & 1ull) cptr[c_ord] = (int32_t)i;
            if (f[e] >> 32) {
                if (s_ord < max_samples) sptr[s_ord] = (int32_t)i;
                else err |= CCN_TOPO_ERR_TOO_MANY_SAMPLES;
                if (batch[i] != (int64_t)s_ord) err |= CCN_TOPO_ERR_BATCH_NOT_DENSE;
            }
        }
    }
    if (err) atomicOr(&counts[2], err);
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
        const unsigned long long all = carry + total;
        const int n_samples = (int)(all >> 32), n_curves = (int)(all & 0xffffffffu);
        cptr[n_curves] = (int32_t)n;
        if (n_samples <= max_samples) sptr[n_samples] = (int32_t)n;
        counts[0] = n_samples;
        counts[1] = n_curves;
    }
}

__global__ void __launch_bounds__(1024) sample_offsets_kernel(const int64_t* __restrict__ curve, int64_t n,
                                                              const int32_t* __restrict__ sptr, int64_t* __restrict__ soff,
                                                              int32_t* __restrict__ counts) {
    ccn_pdl_wait();
    __shared__ long long sm[33];
    if (counts[2]) {                       // invalid topology: sptr is not trustworthy
        if (threadIdx.x == 0) counts[3] = 0;
        return;
    }
    const int n_samples = counts[0];
    long long carry = 0;
    for (int base = 0; base < n_samples; base += blockDim.x) {
        const int b = base + threadIdx.x;
        const long long v = b < n_samples ? (long long)curve[sptr[b + 1] - 1] + 1 : 0;
        long long total;
        const long long ex = carry + ccn_block_excl_scan<long long>(v, sm, &total);
        if (b < n_samples) soff[b] = ex;
        carry += total;
    }
    if (threadIdx.x == 0) {
        const long long off = n_samples == 1 ? (long long)curve[n - 1] + 1 : carry;
        counts[3] = (int32_t)(off > 0x7fffffff ? 0x7fffffff : off);   // max global id + 1
    }
}

__global__ void glob_ids_kernel(const int64_t* __restrict__ batch, const int64_t* __restrict__ curve, int64_t n,
                                const int64_t* __restrict__ soff, const int32_t* __restrict__ counts,
                                int64_t* __restrict__ glob) {
    ccn_pdl_wait();
    if (counts[2]) return;
    const bool single = counts[0] == 1;     // point_ops.py:27-28: a single sample keeps its ids unchanged
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        glob[i] = curve[i] + (single ? 0 : soff[batch[i]]);
}

// CSR pointer of a sorted int64 id vector (batch2ptr with ends) for the levels that carry no curves.
__global__ void __launch_bounds__(1024) sample_ptr_kernel(const int64_t* __restrict__ batch, int64_t n,
                                                          int32_t* __restrict__ sptr, int32_t* __restrict__ counts,
                                                          int max_samples) {
    ccn_pdl_wait();
    __shared__ int sm[33];
    int carry = 0, err = 0;
    for (int64_t base = 0; base < n; base += 1024) {
        const int64_t i = base + threadIdx.x;
        int f = 0;
        if (i < n) {
            f = (i == 0) || batch[i] != batch[i - 1];
            if (i > 0 && batch[i] < batch[i - 1]) err = CCN_TOPO_ERR_BATCH_UNSORTED;
        }
        int total;
        const int ex = carry + ccn_block_excl_scan<int>(f, sm, &total);
        if (f) {
            if (ex < max_samples) sptr[ex] = (int32_t)i;
            else err |= CCN_TOPO_ERR_TOO_MANY_SAMPLES;
            if (batch[i] != ex) err |= CCN_TOPO_ERR_BATCH_NOT_DENSE;
        }
        carry += total;
    }
    __shared__ int s_err;
    if (threadIdx.x == 0) s_err = 0;
    __syncthreads();
    if (err) atomicOr(&s_err, err);
    __syncthreads();
    if (threadIdx.x == 0) {
        if (carry <= max_samples) sptr[carry] = (int32_t)n;
        counts[0] = carry;
        counts[1] = 0;
        counts[2] = s_err;
        counts[3] = 0;
    }
}

}  // namespace

extern "C" int ccn_curve_csr(const int64_t* batch, const int64_t* curve, int64_t n, int32_t* curve_ord,
                             int32_t* curve_ptr, int32_t* sample_ptr, int64_t* sample_curve_off,
                             int64_t* curve_glob, int32_t* counts, int32_t max_samples, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && n < 0x7fffffff && max_samples > 0);
    CCN_REQUIRE(batch && curve && curve_ord && curve_ptr && sample_ptr && sample_curve_off && curve_glob && counts);
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(counts, 0, 4 * sizeof(int32_t), s);
    if (e != cudaSuccess) return (int)e;
    if (n == 0) {                            // empty batch: end pointers only (counts stay zero)
        e = cudaMemsetAsync(curve_ptr, 0, sizeof(int32_t), s);
        if (e == cudaSuccess) e = cudaMemsetAsync(sample_ptr, 0, sizeof(int32_t), s);
        return e == cudaSuccess ? CCN_OK : (int)e;
    }
    const int tiles = (int)((n + kTile - 1) / kTile);
    unsigned long long* tile_sum = reinterpret_cast<unsigned long long*>(curve_glob);
    ccn_launch(tile_flags_kernel, dim3(tiles), dim3(1024), 0, s, batch, curve, n, tile_sum, counts);
    CCN_CHECK_LAUNCH();
    ccn_launch(tile_scan_kernel, dim3(tiles), dim3(1024), 0, s, batch, curve, n, tile_sum, curve_ord, curve_ptr, sample_ptr, counts, max_samples);
    CCN_CHECK_LAUNCH();
    ccn_launch(sample_offsets_kernel, dim3(1), dim3(1024), 0, s, curve, n, sample_ptr, sample_curve_off, counts);
    CCN_CHECK_LAUNCH();
    ccn_launch(glob_ids_kernel, dim3(ccn_grid(n, 256, 148LL * 8)), dim3(256), 0, s, batch, curve, n, sample_curve_off, counts, curve_glob);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_sample_ptr(const int64_t* batch, int64_t n, int32_t* sample_ptr, int32_t* counts,
                              int32_t max_samples, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && n < 0x7fffffff && max_samples > 0 && batch && sample_ptr && counts);
    ccn_launch(sample_ptr_kernel, dim3(1), dim3(1024), 0, (cudaStream_t)stream, batch, n, sample_ptr, counts, max_samples);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

namespace {
// sample CSR of a down-sampled level: idx[0..count) selects points of a batch-sorted parent whose
// batch ids stay non-decreasing along idx (ascending idx, or voxel-key order).  Every sample keeps at
// least one point.  out = {sptr[0..B], count}.
__global__ void subset_sample_ptr_kernel(const int64_t* __restrict__ batch, const int64_t* __restrict__ idx,
                                         const int32_t* __restrict__ count_ptr, int32_t n_samples,
                                         int32_t* __restrict__ sptr) {
    ccn_pdl_wait();
    const int count = *count_ptr;
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t == 0) sptr[n_samples] = count;
    if (t >= count) return;
    const int b = (int)batch[idx[t]];
    const int pb = t == 0 ? -1 : (int)batch[idx[t - 1]];
    for (int s = pb + 1; s <= b && s < n_samples; ++s) sptr[s] = (int32_t)t;
    if (t == count - 1)
        for (int s = b + 1; s < n_samples; ++s) sptr[s] = count;
}
}  // namespace

extern "C" int ccn_subset_sample_ptr(const int64_t* batch, const int64_t* idx, const int32_t* count_ptr,
                                     int64_t capacity, int32_t n_samples, int32_t* sptr_out, ccn_stream_t stream) {
    CCN_REQUIRE(batch && idx && count_ptr && sptr_out && capacity >= 0 && n_samples >= 1);
    ccn_launch(subset_sample_ptr_kernel, dim3(ccn_blocks_exact(capacity > 0 ? capacity : 1, 256)), dim3(256), 0, (cudaStream_t)stream, 
        batch, idx, count_ptr, n_samples, sptr_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
