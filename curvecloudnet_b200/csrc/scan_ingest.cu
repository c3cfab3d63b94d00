// Scan ingest on the device: raw sweep records -> the `Data` fields the model consumes.
// Reference: SemNuScenes.get / _load_frame / _get_curves / _training_augmentation (src/data/nuscenes_dataset.py:42-149),
// SemKITTI.get / _load_frame / _get_curves / _training_augmentation (src/data/kitti_dataset.py:40-152) - serial numpy /
// torch-CPU work inside a num_workers=0 DataLoader.  Here one C-ABI call per sweep:
//   records (n, 5) f32 = x y z intensity ring  (nuScenes .pcd.bin)   | (n, 4) f32 = x y z remission (KITTI velodyne .bin)
//   labels  u8 (nuScenes lidarseg) | i32, low 16 bits (KITTI .label) | none (test split)
//   -> ring ids, stable sort by ring + curve split (ccn_get_curves), then ONE element kernel that gathers reflectance and
//      labels through the sort, maps labels through the learning map, applies the training augmentation in the
//      reference's arithmetic (rotation in f64, flip, f32 scale, f64 translation) and the position normalisation.
#include "common.cuh"

extern "C" int ccn_get_curves(const float* points, int64_t ldp, const int64_t* beam, int64_t n, int32_t n_beams,
                              float curve_thresh, int32_t* scratch, int64_t* reorder, int64_t* inv_reorder,
                              float* sorted_points, int64_t* curve_idxs, int32_t* counts_out, ccn_stream_t stream);

namespace {

struct Augment {
    int on;                 // 0: validation / test (no augmentation)
    double c, s;            // rotation: [x y] <- [x y] [[c, s], [-s, c]]   (np.dot of the f32 columns with an f64 matrix)
    int flip;               // 0 none, 1 x, 2 y, 3 x and y
    float scale;            // xy scale (python float x f32 array: f32 product)
    double tx, ty, tz;      // translation (f32 array += f64 array: f64 sum, rounded once)
};

__global__ void ring_ids_kernel(const float* __restrict__ raw, int64_t n, int rec, int64_t* __restrict__ beam) {
    ccn_pdl_wait();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        beam[i] = rec >= 5 ? (int64_t)raw[i * rec + 4] : 0;
}

__global__ void ingest_finish_kernel(const float* __restrict__ raw, int rec, const void* __restrict__ labels,
                                     int label_bytes, uint32_t label_mask, const int32_t* __restrict__ lut, int lut_size,
                                     const int64_t* __restrict__ reorder, const float* __restrict__ sorted_points,
                                     int64_t n, Augment a, float pos_div, float* __restrict__ pos, float* __restrict__ x,
                                     int64_t* __restrict__ y, int32_t* __restrict__ flags) {
    ccn_pdl_wait();
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = reorder[j];
        x[j] = raw[i * rec + 3];
        int64_t lab = 0;
        if (label_bytes) {
            uint32_t v = label_bytes == 1 ? (uint32_t) reinterpret_cast<const uint8_t*>(labels)[i]
                                          : (uint32_t) reinterpret_cast<const int32_t*>(labels)[i];
            v &= label_mask;
            if (lut) {
                if (v >= (uint32_t)lut_size || lut[v] < 0) { atomicOr(flags, 2); v = 0; }
                else v = (uint32_t)lut[v];
            }
            lab = (int64_t)(v & 0xffu);                        // .astype(np.uint8)
        } else {
            lab = label_bytes == 0 && lut_size == -1 ? 1 : 0;  // nuScenes test split: ones; KITTI test split: zeros
        }
        y[j] = lab;
        float px = sorted_points[3 * j], py = sorted_points[3 * j + 1], pz = sorted_points[3 * j + 2];
        if (a.on) {
            const double dx = (double)px, dy = (double)py;
            px = (float)__dadd_rn(__dmul_rn(dx, a.c), __dmul_rn(dy, -a.s));
            py = (float)__dadd_rn(__dmul_rn(dx, a.s), __dmul_rn(dy, a.c));
            if (a.flip & 1) px = -px;
            if (a.flip & 2) py = -py;
            px = __fmul_rn(a.scale, px);
            py = __fmul_rn(a.scale, py);
            px = (float)__dadd_rn((double)px, a.tx);
            py = (float)__dadd_rn((double)py, a.ty);
            pz = (float)__dadd_rn((double)pz, a.tz);
        }
        pos[3 * j] = __fdiv_rn(px, pos_div);
        pos[3 * j + 1] = __fdiv_rn(py, pos_div);
        pos[3 * j + 2] = __fdiv_rn(pz, pos_div);
    }
}

}  // namespace

extern "C" int ccn_ingest_scan(const float* raw, int64_t n, int32_t rec, const void* labels, int32_t label_bytes,
                               int32_t label_mask, const int32_t* lut, int32_t lut_size, int32_t n_beams,
                               float curve_thresh, int32_t augment, double rot_c, double rot_s, int32_t flip_type,
                               float scale, double tx, double ty, double tz, float pos_div, int32_t* scratch,
                               int64_t* beam_tmp, float* sorted_tmp, int64_t* reorder, int64_t* inv_reorder, float* pos,
                               float* x, int64_t* y, int64_t* curve_idxs, int32_t* counts_out, ccn_stream_t stream) {
    // scratch: int32 [n_beams + n]; beam_tmp: int64 [n]; sorted_tmp: f32 [n, 3];
    // counts_out: int32 [2] = {number of curves, error flags (1: ring id outside [0, n_beams), 2: label not in the map)}
    // lut_size == -1 with label_bytes == 0: dummy labels of ones (nuScenes test split), else zeros (KITTI test split)
    CCN_REQUIRE((rec == 4 || rec == 5) && n >= 0 && n < 0x7fffffff && n_beams > 0);
    CCN_REQUIRE(label_bytes == 0 || label_bytes == 1 || label_bytes == 4);
    CCN_REQUIRE(label_bytes == 0 || labels);
    CCN_REQUIRE(flip_type >= 0 && flip_type <= 3 && pos_div != 0.f);
    CCN_REQUIRE(rec == 5 || n_beams == 1);
    if (n > 0) CCN_REQUIRE(raw && beam_tmp && sorted_tmp && pos && x && y);
    cudaStream_t s = (cudaStream_t)stream;
    if (n > 0) {
        ccn_launch(ring_ids_kernel, dim3(ccn_grid(n, 256, 148 * 8)), dim3(256), 0, s, raw, n, rec, beam_tmp);
        CCN_CHECK_LAUNCH();
    }
    const int rc = ccn_get_curves(raw, rec, beam_tmp, n, n_beams, curve_thresh, scratch, reorder, inv_reorder, sorted_tmp,
                                  curve_idxs, counts_out, stream);
    if (rc != CCN_OK || n == 0) return rc;
    Augment a;
    a.on = augment; a.c = rot_c; a.s = rot_s; a.flip = flip_type; a.scale = scale; a.tx = tx; a.ty = ty; a.tz = tz;
    ccn_launch(ingest_finish_kernel, dim3(ccn_grid(n, 256, 148 * 8)), dim3(256), 0, s, raw, rec, labels, label_bytes, (uint32_t)label_mask, lut, lut_size,
                                                                    reorder, sorted_tmp, n, a, pos_div, pos, x, y,
                                                                    counts_out + 1);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// Object scans: the arithmetic of CurvesInMemoryDataset.get (src/data/object_datasets.py:21-38) - pc_normalize (:61-66:
// centre on the mean, divide by the largest norm) and, for the shapenet-seg training split, torch_geometric's
// NormalizeScale (centre again, scale by 0.999999 / max |coordinate|), RandomScale (one factor) and the random shift.
// One CTA per object; sums in f64 rounded once to f32 (torch's f32 pairwise sums differ from that by an ulp or two - the
// parity test allows 1e-6), everything else in f32 like the reference.
// ---------------------------------------------------------------------------------------------------------------------
namespace {
__device__ __forceinline__ double block_sum_f64(double v, double* smem) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (ccn_lane() == 0) smem[ccn_warp()] = v;
    __syncthreads();
    double t = 0.0;
    for (int w = 0; w < (int)((blockDim.x + 31) >> 5); ++w) t += smem[w];
    return t;
}

__device__ __forceinline__ float block_max_f32(float v, float* smem) {
    for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if (ccn_lane() == 0) smem[ccn_warp()] = v;
    __syncthreads();
    float t = smem[0];
    for (int w = 1; w < (int)((blockDim.x + 31) >> 5); ++w) t = fmaxf(t, smem[w]);
    return t;
}

__global__ void __launch_bounds__(256) object_normalize_kernel(const float* __restrict__ pos, const int32_t* __restrict__ sptr,
                                                               const float* __restrict__ aug, int augment,
                                                               float* __restrict__ out) {
    ccn_pdl_wait();
    __shared__ double sd[8];
    __shared__ float sf[8];
    const int b = blockIdx.x;
    const int lo = sptr[b], hi = sptr[b + 1], n = hi - lo;
    if (n <= 0) return;
    float c[3];
    for (int k = 0; k < 3; ++k) {
        double s = 0.0;
        for (int i = lo + threadIdx.x; i < hi; i += blockDim.x) s += (double)pos[3 * (int64_t)i + k];
        c[k] = (float)(block_sum_f64(s, sd) / (double)n);
    }
    float m = 0.f;
    for (int i = lo + threadIdx.x; i < hi; i += blockDim.x)
        m = fmaxf(m, ccn_norm3(__fsub_rn(pos[3 * (int64_t)i], c[0]), __fsub_rn(pos[3 * (int64_t)i + 1], c[1]),
                               __fsub_rn(pos[3 * (int64_t)i + 2], c[2])));
    m = block_max_f32(m, sf);
    for (int i = lo + threadIdx.x; i < hi; i += blockDim.x)
        for (int k = 0; k < 3; ++k) out[3 * (int64_t)i + k] = __fdiv_rn(__fsub_rn(pos[3 * (int64_t)i + k], c[k]), m);
    if (!augment) return;
    __syncthreads();
    float c2[3];
    for (int k = 0; k < 3; ++k) {
        double s = 0.0;
        for (int i = lo + threadIdx.x; i < hi; i += blockDim.x) s += (double)out[3 * (int64_t)i + k];
        c2[k] = (float)(block_sum_f64(s, sd) / (double)n);
    }
    float am = 0.f;
    for (int i = lo + threadIdx.x; i < hi; i += blockDim.x)
        for (int k = 0; k < 3; ++k) am = fmaxf(am, fabsf(__fsub_rn(out[3 * (int64_t)i + k], c2[k])));
    am = block_max_f32(am, sf);
    const float scale = __fmul_rn(__fdiv_rn(1.f, am), 0.999999f);              // NormalizeScale: (1 / max|pos|) * 0.999999
    const float rs = aug[4 * b];                                               // RandomScale's factor
    for (int i = lo + threadIdx.x; i < hi; i += blockDim.x)
        for (int k = 0; k < 3; ++k) {
            float v = __fmul_rn(__fsub_rn(out[3 * (int64_t)i + k], c2[k]), scale);
            v = __fmul_rn(v, rs);
            out[3 * (int64_t)i + k] = __fadd_rn(v, aug[4 * b + 1 + k]);        // pos += (rand(1,3) - 0.5) * 0.05
        }
}
}  // namespace

extern "C" int ccn_object_normalize(const float* pos, const int32_t* sample_ptr, int32_t n_objects, const float* aug,
                                    float* pos_out, ccn_stream_t stream) {
    // aug: null (validation / other datasets) or (n_objects, 4) f32 = {random scale, shift x, y, z} per object
    CCN_REQUIRE(n_objects >= 0);
    if (n_objects == 0) return CCN_OK;
    CCN_REQUIRE(pos && sample_ptr && pos_out);
    ccn_launch(object_normalize_kernel, dim3((unsigned)n_objects), dim3(256), 0, (cudaStream_t)stream, pos, sample_ptr, aug,
               aug ? 1 : 0, pos_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
