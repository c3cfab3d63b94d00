// Loss / metrics epilogue of the segmentation runs (SURVEY.md §8(f) rank 2, rank 1 helpers):
//   * Lovasz-softmax loss, forward and backward (reference src/models/utils/lovasz_losses.py:21-33,175-204 as called by
//     seg_loss, src/run/nuscenes_seg.py:229-243): per present class, errors |fg - p_c| sorted descending, Jaccard
//     gradient from running foreground counts, dot product.  The reference runs torch.sort + cumsum per class in a
//     Python loop with a host sync per class (`fg.sum() == 0`); here all classes are sorted concurrently (one CTA per
//     class, bitonic network on 64-bit (error, fg, point) keys: 4096-key tiles in shared memory, wider strides in L2).
//   * confusion histogram of fast_hist (src/utils/utils.py:4-8) with the arg-max over classes fused in
//   * the test-time-augmentation pieces of val() (nuscenes_seg.py:86-116): xy flip/scale of the positions, the
//     running sum of the 12 logit tensors, mean + log_softmax + arg-max in one pass.
#include "common.cuh"

namespace {

constexpr int kSortThreads = 1024;
constexpr int kTile = 4096;              // keys sorted per shared-memory tile (32 KB)

__device__ __forceinline__ float row_lse(const float* row, int C) {
    float m = -INFINITY;
    for (int c = 0; c < C; ++c) m = fmaxf(m, row[c]);
    float s = 0.f;
    for (int c = 0; c < C; ++c) s += expf(row[c] - m);
    return m + logf(s);
}

// keys[c][i] = ((bits(|fg - p_ic|) + 1) << 32) | (fg << 31) | i for valid points, 0 for ignored points and the
// power-of-two padding; fg_count[c] = number of valid points of class c; lse[i] kept for the backward.
__global__ void lovasz_keys_kernel(const float* __restrict__ logits, int64_t ld, const int64_t* __restrict__ target,
                                   int64_t n, int C, int64_t ignore, int64_t npad, unsigned long long* __restrict__ keys,
                                   int32_t* __restrict__ fg_count, float* __restrict__ lse) {
    ccn_pdl_wait();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < npad; i += (int64_t)gridDim.x * blockDim.x) {
        if (i >= n) {
            for (int c = 0; c < C; ++c) keys[(int64_t)c * npad + i] = 0ull;
            continue;
        }
        const int64_t y = target[i];
        const float* row = logits + i * ld;
        const float l = row_lse(row, C);
        lse[i] = l;
        const bool valid = y != ignore;
        if (valid && y >= 0 && y < C) atomicAdd(fg_count + y, 1);
        for (int c = 0; c < C; ++c) {
            unsigned long long k = 0ull;
            if (valid) {
                const float p = expf(row[c] - l);
                const unsigned fg = (y == c) ? 1u : 0u;
                const float e = fabsf((fg ? 1.f : 0.f) - p);
                k = ((unsigned long long)(__float_as_uint(e) + 1u) << 32) | ((unsigned long long)fg << 31) | (unsigned long long)i;
            }
            keys[(int64_t)c * npad + i] = k;
        }
    }
}

__device__ __forceinline__ void cmp_swap(unsigned long long& a, unsigned long long& b, bool desc) {
    if ((a < b) == desc) { const unsigned long long t = a; a = b; b = t; }
}

// One CTA per class: bitonic sort (descending) of npad keys in global memory.
__global__ void __launch_bounds__(kSortThreads) lovasz_sort_kernel(unsigned long long* __restrict__ keys, int64_t npad,
                                                                   const int32_t* __restrict__ fg_count) {
    ccn_pdl_wait();
    __shared__ unsigned long long tile[kTile];
    const int c = blockIdx.x;
    if (fg_count[c] == 0) return;                       // class absent: skipped like classes='present'
    unsigned long long* a = keys + (int64_t)c * npad;
    const int tile_n = (int)(npad < kTile ? npad : kTile);
    // phase 1: every tile fully sorted in shared memory, direction alternating as the global network needs
    for (int64_t t0 = 0; t0 < npad; t0 += tile_n) {
        for (int i = threadIdx.x; i < tile_n; i += kSortThreads) tile[i] = a[t0 + i];
        __syncthreads();
        for (int k = 2; k <= tile_n; k <<= 1) {
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int i = threadIdx.x; i < tile_n; i += kSortThreads) {
                    const int l = i ^ j;
                    if (l > i) {
                        const bool desc = (((t0 + i) & (int64_t)k) == 0);
                        unsigned long long x = tile[i], y = tile[l];
                        cmp_swap(x, y, desc);
                        tile[i] = x; tile[l] = y;
                    }
                }
                __syncthreads();
            }
        }
        for (int i = threadIdx.x; i < tile_n; i += kSortThreads) a[t0 + i] = tile[i];
        __syncthreads();
    }
    // phase 2: merges wider than a tile - strides >= tile_n in global memory, the rest of each merge in shared memory
    for (int64_t k = (int64_t)tile_n << 1; k <= npad; k <<= 1) {
        for (int64_t j = k >> 1; j >= tile_n; j >>= 1) {
            for (int64_t i = threadIdx.x; i < npad; i += kSortThreads) {
                const int64_t l = i ^ j;
                if (l > i) {
                    const bool desc = ((i & k) == 0);
                    unsigned long long x = a[i], y = a[l];
                    cmp_swap(x, y, desc);
                    a[i] = x; a[l] = y;
                }
            }
            __syncthreads();
        }
        for (int64_t t0 = 0; t0 < npad; t0 += tile_n) {
            for (int i = threadIdx.x; i < tile_n; i += kSortThreads) tile[i] = a[t0 + i];
            __syncthreads();
            const bool desc = ((t0 & k) == 0);
            for (int j = tile_n >> 1; j > 0; j >>= 1) {
                for (int i = threadIdx.x; i < tile_n; i += kSortThreads) {
                    const int l = i ^ j;
                    if (l > i) {
                        unsigned long long x = tile[i], y = tile[l];
                        cmp_swap(x, y, desc);
                        tile[i] = x; tile[l] = y;
                    }
                }
                __syncthreads();
            }
            for (int i = threadIdx.x; i < tile_n; i += kSortThreads) a[t0 + i] = tile[i];
            __syncthreads();
        }
    }
}

// One CTA per class over the sorted keys: Jaccard gradient (lovasz_grad) of every rank, written back per POINT into
// grad[c][i], and loss_c = sum_r error_r * grad_r.  acc: fp64 [2] = (sum of class losses, number of present classes).
__global__ void __launch_bounds__(kSortThreads) lovasz_scan_kernel(const unsigned long long* __restrict__ keys,
                                                                   int64_t npad, const int32_t* __restrict__ fg_count,
                                                                   float* __restrict__ grad,
                                                                   int64_t n, double* __restrict__ acc) {
    ccn_pdl_wait();
    __shared__ int scan_smem[33];
    __shared__ double red[kSortThreads / 32];
    const int c = blockIdx.x;
    const int gts_i = fg_count[c];
    if (gts_i == 0) return;
    const unsigned long long* a = keys + (int64_t)c * npad;
    float* g = grad + (int64_t)c * n;
    const float gts = (float)gts_i;
    int running = 0;                                    // foreground count before this chunk
    double part = 0.0;
    for (int64_t r0 = 0; r0 < npad; r0 += kSortThreads) {
        const int64_t r = r0 + threadIdx.x;
        const unsigned long long k = r < npad ? a[r] : 0ull;
        const int fg = (int)((k >> 31) & 1ull);
        int tot;
        const int ex = ccn_block_excl_scan<int>(k ? fg : 0, scan_smem, &tot);
        if (k) {
            const float e = __uint_as_float((unsigned)(k >> 32) - 1u);
            const int cum = running + ex + fg;                      // inclusive foreground count at rank r
            // jaccard_r = 1 - (gts - cum_fg) / (gts + (r + 1 - cum_fg)); grad_r = jaccard_r - jaccard_{r-1}
            const float jac = 1.f - (gts - (float)cum) / (gts + (float)(r + 1 - cum));
            float prev = 0.f;
            if (r > 0) {
                const int cum_p = cum - fg;
                prev = 1.f - (gts - (float)cum_p) / (gts + (float)(r - cum_p));
            }
            const float gr = jac - prev;
            g[(int64_t)(k & 0x7fffffffull)] = gr;
            part += (double)(e * gr);
        }
        running += tot;
        if (__syncthreads_and(k == 0ull)) break;        // only ignored points / padding from here on
    }
    for (int o = 16; o > 0; o >>= 1) part += __shfl_down_sync(0xffffffffu, part, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < kSortThreads / 32; ++w) s += red[w];
        atomicAdd(acc, (double)(float)s);
        atomicAdd(acc + 1, 1.0);
    }
}

__global__ void lovasz_finish_kernel(const double* __restrict__ acc, float* __restrict__ loss) {
    ccn_pdl_wait();
    loss[0] = acc[1] > 0.0 ? (float)(acc[0] / acc[1]) : 0.f;
}

// dlogits[i][k] += gscale / n_present * p_k (dP_k - sum_c dP_c p_c), dP_c = grad[c][i] * sign(p_c - fg_c)
__global__ void lovasz_bwd_kernel(const float* __restrict__ logits, int64_t ld, const int64_t* __restrict__ target,
                                  int64_t n, int C, int64_t ignore, const float* __restrict__ lse,
                                  const float* __restrict__ grad, const int32_t* __restrict__ fg_count,
                                  const double* __restrict__ acc, const float* __restrict__ gscale,
                                  float* __restrict__ dlogits, int64_t ldd, int accumulate) {
    ccn_pdl_wait();
    const float sc = acc[1] > 0.0 ? gscale[0] / (float)acc[1] : 0.f;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t y = target[i];
        const float* row = logits + i * ld;
        float* out = dlogits + i * ldd;
        if (y == ignore) {
            if (!accumulate) for (int c = 0; c < C; ++c) out[c] = 0.f;
            continue;
        }
        const float l = lse[i];
        float dot = 0.f;
        for (int c = 0; c < C; ++c) {
            if (fg_count[c] == 0) continue;
            const float p = expf(row[c] - l);
            const float fg = (y == c) ? 1.f : 0.f;
            const float d = p - fg;
            const float sg = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
            dot += grad[(int64_t)c * n + i] * sg * p;
        }
        for (int k = 0; k < C; ++k) {
            const float p = expf(row[k] - l);
            float dp = 0.f;
            if (fg_count[k] != 0) {
                const float d = p - ((y == k) ? 1.f : 0.f);
                dp = grad[(int64_t)k * n + i] * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f));
            }
            const float v = sc * p * (dp - dot);
            out[k] = accumulate ? out[k] + v : v;
        }
    }
}

// hist[label][pred] += 1 for labels in [0, ncls); pred = arg-max over the C columns of `scores` (first maximum), or
// the given prediction vector when scores == nullptr          (fast_hist, src/utils/utils.py:4-8)
__global__ void confusion_hist_kernel(const float* __restrict__ scores, int64_t ld, int C, const int64_t* __restrict__ pred_in,
                                      const int64_t* __restrict__ target, int64_t n, int ncls, int col_offset,
                                      int64_t* __restrict__ pred_out, unsigned long long* __restrict__ hist) {
    ccn_pdl_wait();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t p;
        if (scores) {
            const float* row = scores + i * ld;
            int arg = col_offset;
            float best = row[col_offset];
            for (int c = col_offset + 1; c < C; ++c)
                if (row[c] > best) { best = row[c]; arg = c; }
            p = arg;
        } else {
            p = pred_in[i];
        }
        if (pred_out) pred_out[i] = p;
        const int64_t y = target ? target[i] : -1;
        if (hist && y >= 0 && y < ncls && p >= 0 && p < ncls) atomicAdd(hist + y * ncls + p, 1ull);
    }
}

// out = [sx * x, sy * y, z]                                     (the flips / scales of val(), nuscenes_seg.py:93-104)
__global__ void scale_xy_kernel(const float* __restrict__ pos, int64_t n, float sx, float sy, float* __restrict__ out) {
    ccn_pdl_wait();
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < 3 * n; t += (int64_t)gridDim.x * blockDim.x) {
        const int a = (int)(t % 3);
        const float v = pos[t];
        out[t] = a == 0 ? v * sx : (a == 1 ? v * sy : v);
    }
}

// acc (+)= src                                                  (sum(outs) of the augmented forwards)
__global__ void accumulate_kernel(const float* __restrict__ src, int64_t lds, int64_t n, int C, float* __restrict__ acc,
                                  int64_t lda, int first) {
    ccn_pdl_wait();
    const int64_t total = n * (int64_t)C;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t r;
        int c;
        ccn_divmod(t, C, r, c);
        const float v = src[r * lds + c];
        acc[r * lda + c] = first ? v : acc[r * lda + c] + v;
    }
}

// out = log_softmax(acc * scale, dim=-1); pred = arg-max              (out = sum(outs) / 12; F.log_softmax; argmax)
__global__ void mean_logsoftmax_kernel(const float* __restrict__ acc, int64_t lda, int64_t n, int C, float scale,
                                       float* __restrict__ out, int64_t ldo, int64_t* __restrict__ pred) {
    ccn_pdl_wait();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float* row = acc + i * lda;
        float m = -INFINITY;
        int arg = 0;
        for (int c = 0; c < C; ++c) {
            const float v = row[c] * scale;
            if (v > m) { m = v; arg = c; }
        }
        float s = 0.f;
        for (int c = 0; c < C; ++c) s += expf(row[c] * scale - m);
        const float l = m + logf(s);
        for (int c = 0; c < C; ++c) out[i * ldo + c] = row[c] * scale - l;
        if (pred) pred[i] = arg;
    }
}

}  // namespace

#define S_ ((cudaStream_t)stream)
#define PT_GRID(n) ccn_grid((n), 128, 148LL * 8)

extern "C" int64_t ccn_lovasz_pad(int64_t n) {
    int64_t p = 2;
    while (p < n) p <<= 1;
    return p;
}

extern "C" int ccn_lovasz_softmax_fwd(const float* logits, int64_t ld, const int64_t* target, int64_t n, int32_t C,
                                      int64_t ignore_index, void* keys, int32_t* fg_count, float* lse, float* grad,
                                      double* acc, float* loss, ccn_stream_t stream) {
    // keys: uint64 [C * ccn_lovasz_pad(n)], fg_count: int32 [C], lse: f32 [n], grad: f32 [C * n], acc: f64 [2], loss: f32 [1]
    CCN_REQUIRE(n > 0 && n < 0x7fffffff && C > 0 && C <= 1024 && logits && target && keys && fg_count && lse && grad &&
                acc && loss && ld >= C);
    const int64_t npad = ccn_lovasz_pad(n);
    cudaError_t e = cudaMemsetAsync(fg_count, 0, (size_t)C * sizeof(int32_t), S_);
    if (e == cudaSuccess) e = cudaMemsetAsync(acc, 0, 2 * sizeof(double), S_);
    if (e == cudaSuccess) e = cudaMemsetAsync(grad, 0, (size_t)C * n * sizeof(float), S_);
    if (e != cudaSuccess) return (int)e;
    ccn_launch(lovasz_keys_kernel, dim3(PT_GRID(npad)), dim3(128), 0, S_, logits, ld, target, n, C, ignore_index, npad,
                                                      (unsigned long long*)keys, fg_count, lse);
    CCN_CHECK_LAUNCH();
    ccn_launch(lovasz_sort_kernel, dim3(C), dim3(kSortThreads), 0, S_, (unsigned long long*)keys, npad, fg_count);
    CCN_CHECK_LAUNCH();
    ccn_launch(lovasz_scan_kernel, dim3(C), dim3(kSortThreads), 0, S_, (const unsigned long long*)keys, npad, fg_count, grad, n, acc);
    CCN_CHECK_LAUNCH();
    ccn_launch(lovasz_finish_kernel, dim3(1), dim3(1), 0, S_, acc, loss);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_lovasz_softmax_bwd(const float* logits, int64_t ld, const int64_t* target, int64_t n, int32_t C,
                                      int64_t ignore_index, const float* lse, const float* grad, const int32_t* fg_count,
                                      const double* acc, const float* gscale, float* dlogits, int64_t ldd,
                                      int32_t accumulate, ccn_stream_t stream) {
    CCN_REQUIRE(n > 0 && C > 0 && logits && target && lse && grad && fg_count && acc && gscale && dlogits && ld >= C &&
                ldd >= C);
    ccn_launch(lovasz_bwd_kernel, dim3(PT_GRID(n)), dim3(128), 0, S_, logits, ld, target, n, C, ignore_index, lse, grad, fg_count, acc, gscale,
                                                  dlogits, ldd, accumulate);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_confusion_hist(const float* scores, int64_t ld, int32_t C, const int64_t* pred_in, const int64_t* target,
                                  int64_t n, int32_t ncls, int32_t col_offset, int64_t* pred_out, int64_t* hist,
                                  ccn_stream_t stream) {
    // hist: int64 [ncls * ncls], ACCUMULATED into (zero it once per evaluation)
    CCN_REQUIRE(n >= 0 && (scores || pred_in) && (!scores || (C > 0 && ld >= C && col_offset >= 0 && col_offset < C)));
    CCN_REQUIRE((!hist || (target && ncls > 0)) && (hist || pred_out));
    if (n == 0) return CCN_OK;
    ccn_launch(confusion_hist_kernel, dim3(PT_GRID(n)), dim3(128), 0, S_, scores, ld, C, pred_in, target, n, ncls, col_offset, pred_out,
                                                      (unsigned long long*)hist);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_scale_xy(const float* pos, int64_t n, float sx, float sy, float* out, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && (n == 0 || (pos && out)));
    if (n == 0) return CCN_OK;
    ccn_launch(scale_xy_kernel, dim3(ccn_grid(3 * n, 256, 148 * 8)), dim3(256), 0, S_, pos, n, sx, sy, out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_accumulate(const float* src, int64_t lds, int64_t n, int32_t C, float* acc, int64_t lda, int32_t first,
                              ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && C > 0 && (n == 0 || (src && acc)) && lds >= C && lda >= C);
    if (n == 0) return CCN_OK;
    ccn_launch(accumulate_kernel, dim3(ccn_grid(n * C, 256, 148 * 8)), dim3(256), 0, S_, src, lds, n, C, acc, lda, first);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_mean_logsoftmax(const float* acc, int64_t lda, int64_t n, int32_t C, float scale, float* out,
                                   int64_t ldo, int64_t* pred, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && C > 0 && (n == 0 || (acc && out)) && lda >= C && ldo >= C);
    if (n == 0) return CCN_OK;
    ccn_launch(mean_logsoftmax_kernel, dim3(PT_GRID(n)), dim3(128), 0, S_, acc, lda, n, C, scale, out, ldo, pred);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// Part IoU of the object-segmentation runs (src/run/shapenet_seg.py:88-99): per object, the arg-max over the score columns of
// the object's category (torch_geometric ShapeNet.seg_classes), the target mapped to its position in that part list, and
// torchmetrics' jaccard_index(num_classes = #parts, absent_score): per part intersection / union from the confusion matrix,
// absent_score where the union is empty, mean over the parts.  The reference loops over the objects on the host with one
// .item() each; here one CTA per object, one launch per batch.
// ---------------------------------------------------------------------------------------------------------------------
namespace {
constexpr int kMaxParts = 16;

__global__ void __launch_bounds__(256) part_iou_kernel(const float* __restrict__ scores, int64_t ld,
                                                       const int64_t* __restrict__ target, const int32_t* __restrict__ sptr,
                                                       const int64_t* __restrict__ category, const int32_t* __restrict__ parts,
                                                       const int32_t* __restrict__ parts_ptr, int n_categories, float absent,
                                                       float* __restrict__ iou_out, int64_t* __restrict__ pred_out,
                                                       int32_t* __restrict__ flags) {
    ccn_pdl_wait();
    __shared__ int conf[kMaxParts * kMaxParts];
    __shared__ int plist[kMaxParts];
    const int b = blockIdx.x;
    const int64_t cat = category[b];
    if (cat < 0 || cat >= n_categories) {
        if (threadIdx.x == 0) { atomicOr(flags, 1); iou_out[b] = 0.f; }
        return;
    }
    const int p0 = parts_ptr[cat], P = parts_ptr[cat + 1] - p0;
    for (int t = threadIdx.x; t < kMaxParts * kMaxParts; t += blockDim.x) conf[t] = 0;
    if (threadIdx.x < P) plist[threadIdx.x] = parts[p0 + threadIdx.x];
    __syncthreads();
    for (int i = sptr[b] + threadIdx.x; i < sptr[b + 1]; i += blockDim.x) {
        const float* row = scores + (int64_t)i * ld;
        int best = 0;
        float bv = row[plist[0]];
        for (int k = 1; k < P; ++k) {
            const float v = row[plist[k]];
            if (v > bv) { bv = v; best = k; }               // first maximum, like torch.argmax
        }
        if (pred_out) pred_out[i] = best;
        const int64_t y = target[i];
        int tk = -1;
        for (int k = 0; k < P; ++k) tk = (plist[k] == y) ? k : tk;
        if (tk < 0) { atomicOr(flags, 2); continue; }       // a label outside the category's parts (upstream reads a stale map)
        atomicAdd(&conf[tk * kMaxParts + best], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float sum = 0.f;
        for (int c = 0; c < P; ++c) {
            int row = 0, col = 0;
            for (int k = 0; k < P; ++k) { row += conf[c * kMaxParts + k]; col += conf[k * kMaxParts + c]; }
            const int inter = conf[c * kMaxParts + c], uni = row + col - inter;
            sum = __fadd_rn(sum, uni == 0 ? absent : __fdiv_rn((float)inter, (float)uni));
        }
        iou_out[b] = __fdiv_rn(sum, (float)P);
    }
}
}  // namespace

extern "C" int ccn_part_iou(const float* scores, int64_t ld, const int64_t* target, const int32_t* sample_ptr,
                            const int64_t* category, int32_t n_objects, const int32_t* parts, const int32_t* parts_ptr,
                            int32_t n_categories, int32_t max_parts, float absent_score, float* iou_out, int64_t* pred_out,
                            int32_t* flags_out, ccn_stream_t stream) {
    CCN_REQUIRE(n_objects >= 0 && n_categories >= 1 && max_parts >= 1 && max_parts <= kMaxParts && flags_out);
    cudaError_t e = cudaMemsetAsync(flags_out, 0, sizeof(int32_t), S_);
    if (e != cudaSuccess) return (int)e;
    if (n_objects == 0) return CCN_OK;
    CCN_REQUIRE(scores && target && sample_ptr && category && parts && parts_ptr && iou_out);
    ccn_launch(part_iou_kernel, dim3((unsigned)n_objects), dim3(256), 0, S_, scores, ld, target, sample_ptr, category, parts,
               parts_ptr, (int)n_categories, absent_score, iou_out, pred_out, flags_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
