// Curve / sample CSR topology of a ragged batch of scans.
// Replaces curveidx_local2global / batch2ptr (reference src/models/utils/point_ops.py:20-54),
// which the reference recomputes (with a torch.where host sync) in every curve step.
#include "common.cuh"

namespace {

constexpr int kItems = 4;

// One CTA walks the sorted (batch, curve) id vectors once with a running carry:
// hi 32 bits of the scanned value count sample starts, lo 32 bits curve starts.
__global__ void __launch_bounds__(1024) curve_csr_kernel(
    const int64_t* __restrict__ batch, const int64_t* __restrict__ curve, int64_t n,
    int32_t* __restrict__ ord, int32_t* __restrict__ cptr, int32_t* __restrict__ sptr,
    int64_t* __restrict__ soff, int64_t* __restrict__ glob, int32_t* __restrict__ counts, int max_samples) {
    __shared__ unsigned long long sm[33];
    __shared__ int s_err;
    if (threadIdx.x == 0) s_err = 0;
    __syncthreads();
    unsigned long long carry = 0;
    int err = 0;
    for (int64_t base = 0; base < n; base += 1024 * kItems) {
        const int64_t i0 = base + (int64_t)threadIdx.x * kItems;
        unsigned long long f[kItems];
        unsigned long long local = 0;
#pragma unroll
        for (int e = 0; e < kItems; ++e) {
            const int64_t i = i0 + e;
            f[e] = 0;
            if (i < n) {
                const int64_t b = batch[i], c = curve[i];
                bool ss = true, cs = true;
                if (i > 0) {
                    const int64_t pb = batch[i - 1], pc = curve[i - 1];
                    ss = b != pb;
                    cs = ss || c != pc;
                    if (b < pb) err |= CCN_TOPO_ERR_BATCH_UNSORTED;
                    if (!ss && c < pc) err |= CCN_TOPO_ERR_CURVE_UNSORTED;
                }
                f[e] = ((unsigned long long)ss << 32) | (unsigned long long)cs;
            }
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
        carry += total;
    }
    if (err) atomicOr(&s_err, err);
    const int n_samples = (int)(carry >> 32), n_curves = (int)(carry & 0xffffffffu);
    if (threadIdx.x == 0) {
        cptr[n_curves] = (int32_t)n;
        if (n_samples <= max_samples) sptr[n_samples] = (int32_t)n;
    }
    __syncthreads();
    const int errs = s_err;
    if (threadIdx.x == 0) {
        // reference point_ops.py:32-35: offset of sample b = sum over b' < b of (last curve id of b' + 1)
        int64_t off = 0;
        if (!errs) {
            for (int b = 0; b < n_samples; ++b) {
                soff[b] = off;
                off += curve[sptr[b + 1] - 1] + 1;
            }
            if (n_samples == 1) off = curve[n - 1] + 1;
        }
        counts[0] = n_samples;
        counts[1] = n_curves;
        counts[2] = errs;
        counts[3] = (int32_t)(off > 0x7fffffff ? 0x7fffffff : off);   // max global id + 1
    }
    __syncthreads();
    if (errs) return;
    // point_ops.py:27-28: a single sample keeps its ids unchanged
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x)
        glob[i] = curve[i] + (n_samples == 1 ? 0 : soff[batch[i]]);
}

// CSR pointer of a sorted int64 id vector (batch2ptr with ends) for the levels that carry no curves.
__global__ void __launch_bounds__(1024) sample_ptr_kernel(const int64_t* __restrict__ batch, int64_t n,
                                                          int32_t* __restrict__ sptr, int32_t* __restrict__ counts,
                                                          int max_samples) {
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
    curve_csr_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(batch, curve, n, curve_ord, curve_ptr, sample_ptr,
                                                           sample_curve_off, curve_glob, counts, max_samples);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_sample_ptr(const int64_t* batch, int64_t n, int32_t* sample_ptr, int32_t* counts,
                              int32_t max_samples, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && n < 0x7fffffff && max_samples > 0 && batch && sample_ptr && counts);
    sample_ptr_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(batch, n, sample_ptr, counts, max_samples);
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
    subset_sample_ptr_kernel<<<ccn_blocks_exact(capacity > 0 ? capacity : 1, 256), 256, 0, (cudaStream_t)stream>>>(
        batch, idx, count_ptr, n_samples, sptr_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
