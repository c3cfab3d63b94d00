// Scanning simulator, stage 1: triangle mesh -> per-pixel surface point, face normal and hit mask of a pinhole view
// (SURVEY 8(f) rank 4).  Replaces scanning_simulator/utils/minimal_rasterizer.py:14-51 `rasterize` = vertices @ R^T + t,
// Pinhole2D.project_ndc (third_party/minimal_pytorch_rasterizer/.../camera.py:38-61), estimate_normals_cuda
// (.../cuda/rasterizer_kernel.cu:80-225 z-buffer by atomicMin, :335-465 coordinates + normals of the winning face) and the
// transform back to the scene frame.  No graphics API: plain CUDA.
//
// Design (not the reference's): the reference runs one 16-thread CTA per face TWICE (depth pass, then an attribute pass in
// which every face within 1e-5 of the z-buffer writes its pixel - whichever writes last wins) over strided float stores.
//   1. sim_project_kernel   one thread per vertex: camera-space position and NDC.
//   2. sim_raster_kernel    one WARP per face: per-face set-up once (lane 0, broadcast by shuffle), the lanes sweep the
//                           bounding box row-major (coalesced key traffic), every covered pixel does ONE 64-bit atomicMin of
//                           the key (depth bits << 32 | face index): the z-buffer and the winning face in one word, and a
//                           DETERMINISTIC winner (closest face; equal depth -> lowest face index).  Faces whose box exceeds
//                           kBigBox pixels are queued instead and swept by whole CTAs (sim_raster_big_kernel).
//   3. sim_resolve_kernel   one thread per PIXEL: decodes the face, re-evaluates its perspective-correct barycentrics at the
//                           pixel centre, writes point / normal / mask in the scene frame with coalesced stores.
// Arithmetic follows the reference expression by expression in fp32 (pixel centres in double like its `j + 0.5` literal), every
// operation rounded separately (no FMA contraction) so that oracle/sim_raster.c (gcc -ffp-contract=off) is bit-identical.
#include "common.cuh"

namespace {

constexpr int kBigBox = 4096;          // bounding boxes beyond this many pixels go to the CTA-per-face kernel
constexpr unsigned long long kEmpty = ~0ull;

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float dvd(float a, float b) { return __fdiv_rn(a, b); }

// cam = v @ R^T + t ; ndc.xy = 2 ((cam.xy * f + cam.z * c) / (cam.z + 1e-9)) / res - 1 ; ndc.z = cam.z
__global__ void sim_project_kernel(const float* __restrict__ v, int64_t nv, const float* __restrict__ R,
                                   const float* __restrict__ t, int res, float* __restrict__ cam, float* __restrict__ ndc) {
    ccn_pdl_wait();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= nv) return;
    const float x = v[3 * i], y = v[3 * i + 1], z = v[3 * i + 2];
    float c[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) c[r] = add(add(add(mul(x, R[3 * r]), mul(y, R[3 * r + 1])), mul(z, R[3 * r + 2])), t[r]);
    cam[3 * i] = c[0]; cam[3 * i + 1] = c[1]; cam[3 * i + 2] = c[2];
    const float f = (float)res, cx = (float)(res / 2), zz = add(c[2], 1e-9f);
    ndc[3 * i] = sub(dvd(mul(2.f, dvd(add(mul(c[0], f), mul(c[2], cx)), zz)), f), 1.f);
    ndc[3 * i + 1] = sub(dvd(mul(2.f, dvd(add(mul(c[1], f), mul(c[2], cx)), zz)), f), 1.f);
    ndc[3 * i + 2] = c[2];
}

struct FaceSetup {
    float f[9];        // NDC vertices a, b, c: (x, y, z)
    float inv[9];      // rows of the inverse of [[ax bx cx] [ay by cy] [1 1 1]] (rasterizer_kernel.cu:127-151)
    int x0, x1, y0, y1;
    int bad;
};

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

__device__ void face_setup(const float* __restrict__ ndc, const int32_t* __restrict__ faces, int64_t face, int res, FaceSetup& s) {
    const int ai = faces[3 * face], bi = faces[3 * face + 1], ci = faces[3 * face + 2];
    float* f = s.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) { f[k] = ndc[3 * (int64_t)ai + k]; f[3 + k] = ndc[3 * (int64_t)bi + k]; f[6 + k] = ndc[3 * (int64_t)ci + k]; }
    const float eps = 1e-5f;
    s.bad = (f[2] < eps || f[5] < eps || f[8] < eps) ? 1 : 0;
    if (s.bad) return;
    float* q = s.inv;
    q[0] = sub(f[4], f[7]); q[1] = sub(f[6], f[3]); q[2] = sub(mul(f[3], f[7]), mul(f[6], f[4]));
    q[3] = sub(f[7], f[1]); q[4] = sub(f[0], f[6]); q[5] = sub(mul(f[6], f[1]), mul(f[0], f[7]));
    q[6] = sub(f[1], f[4]); q[7] = sub(f[3], f[0]); q[8] = sub(mul(f[0], f[4]), mul(f[3], f[1]));
    const float denom = add(add(mul(f[6], sub(f[1], f[4])), mul(f[0], sub(f[4], f[7]))), mul(f[3], sub(f[7], f[1])));
#pragma unroll
    for (int k = 0; k < 9; ++k) q[k] = dvd(q[k], denom);
    const float W = (float)res;
    const float mnx = fminf(fminf(f[0], f[3]), f[6]), mxx = fmaxf(fmaxf(f[0], f[3]), f[6]);
    const float mny = fminf(fminf(f[1], f[4]), f[7]), mxy = fmaxf(fmaxf(f[1], f[4]), f[7]);
    s.x0 = clampi((int)floorf(mul(dvd(add(mnx, 1.f), 2.f), W)), 0, res - 1);
    s.x1 = clampi((int)ceilf(mul(dvd(add(mxx, 1.f), 2.f), W)), 0, res - 1);
    s.y0 = clampi((int)floorf(mul(dvd(add(mny, 1.f), 2.f), W)), 0, res - 1);
    s.y1 = clampi((int)ceilf(mul(dvd(add(mxy, 1.f), 2.f), W)), 0, res - 1);
}

// pixel centre in NDC: (float)(2 (j + 0.5) / W - 1) evaluated in double (the reference's `(scalar_t)j + 0.5` is a double)
__device__ __forceinline__ float pix_ndc(int j, int res) { return (float)(2.0 * ((double)(float)j + 0.5) / (double)res - 1.0); }

// inside test + perspective-correct barycentrics of pixel (x, y) (rasterizer_kernel.cu:193-216); false: not covered
__device__ __forceinline__ bool face_eval(const float* f, const float* q, float x, float y, float& wa, float& wb, float& wc, float& z) {
    if (mul(sub(y, f[1]), sub(f[3], f[0])) > mul(sub(x, f[0]), sub(f[4], f[1])) ||
        mul(sub(y, f[4]), sub(f[6], f[3])) > mul(sub(x, f[3]), sub(f[7], f[4])) ||
        mul(sub(y, f[7]), sub(f[0], f[6])) > mul(sub(x, f[6]), sub(f[1], f[7])))
        return false;
    wa = add(add(mul(q[0], x), mul(q[1], y)), q[2]);
    wb = add(add(mul(q[3], x), mul(q[4], y)), q[5]);
    wc = add(add(mul(q[6], x), mul(q[7], y)), q[8]);
    float ws = add(add(wa, wb), wc);
    wa = dvd(wa, ws); wb = dvd(wb, ws); wc = dvd(wc, ws);
    wa = dvd(wa, f[2]); wb = dvd(wb, f[5]); wc = dvd(wc, f[8]);
    ws = add(add(wa, wb), wc);
    wa = dvd(wa, ws); wb = dvd(wb, ws); wc = dvd(wc, ws);
    z = add(add(mul(wa, f[2]), mul(wb, f[5])), mul(wc, f[8]));
    return true;
}

__device__ __forceinline__ void splat(const FaceSetup& s, int64_t face, int pix0, int npix, int stride, int res,
                                      unsigned long long* __restrict__ keys) {
    const int bw = s.x1 - s.x0 + 1;
    for (int p = pix0; p < npix; p += stride) {
        const int dy = p / bw, i = s.y0 + dy, j = s.x0 + (p - dy * bw);
        float wa, wb, wc, z;
        if (!face_eval(s.f, s.inv, pix_ndc(j, res), pix_ndc(i, res), wa, wb, wc, z)) continue;
        if (!(z == z)) continue;                                  // degenerate face: NaN depth never wins (the reference's atomicMin on NaN bits is undefined)
        const unsigned long long key = ((unsigned long long)__float_as_uint(z) << 32) | (unsigned long long)(uint32_t)face;
        atomicMin(keys + (int64_t)i * res + j, key);
    }
}

__global__ void __launch_bounds__(256) sim_raster_kernel(const float* __restrict__ ndc, const int32_t* __restrict__ faces,
                                                         int64_t nf, int res, unsigned long long* __restrict__ keys,
                                                         int32_t* __restrict__ big_list, int32_t* __restrict__ big_count) {
    ccn_pdl_wait();
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t face = warp; face < nf; face += nwarps) {
        FaceSetup s;
        face_setup(ndc, faces, face, res, s);                     // every lane: the loads are broadcast, the arithmetic uniform
        if (s.bad) continue;
        const int npix = (s.x1 - s.x0 + 1) * (s.y1 - s.y0 + 1);
        if (npix > kBigBox) {
            if (lane == 0) big_list[atomicAdd(big_count, 1)] = (int32_t)face;
            continue;
        }
        splat(s, face, lane, npix, 32, res, keys);
    }
}

__global__ void __launch_bounds__(256) sim_raster_big_kernel(const float* __restrict__ ndc, const int32_t* __restrict__ faces,
                                                             int res, unsigned long long* __restrict__ keys,
                                                             const int32_t* __restrict__ big_list,
                                                             const int32_t* __restrict__ big_count) {
    ccn_pdl_wait();
    const int n = *big_count;
    // a big face is swept by kSplit CTAs (interleaved pixels): a table top covering a quarter of the image is 260 k pixels
    constexpr int kSplit = 8;
    for (int w = blockIdx.x; w < n * kSplit; w += gridDim.x) {
        const int64_t face = big_list[w / kSplit];
        const int part = w % kSplit;
        FaceSetup s;
        face_setup(ndc, faces, face, res, s);
        const int npix = (s.x1 - s.x0 + 1) * (s.y1 - s.y0 + 1);
        splat(s, face, part * 256 + threadIdx.x, npix, 256 * kSplit, res, keys);
    }
}

// point / normal / mask of every pixel from the winning face, in the scene frame: (p - t) @ R, n @ R
__global__ void sim_resolve_kernel(const float* __restrict__ cam, const float* __restrict__ ndc, const int32_t* __restrict__ faces,
                                   const unsigned long long* __restrict__ keys, const float* __restrict__ R,
                                   const float* __restrict__ t, int res, float* __restrict__ coords, float* __restrict__ normals,
                                   uint8_t* __restrict__ mask, float* __restrict__ depth) {
    ccn_pdl_wait();
    const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (p >= (int64_t)res * res) return;
    const unsigned long long key = keys[p];
    float c[3] = {0.f, 0.f, 0.f}, n[3] = {0.f, 0.f, 0.f};
    if (key != kEmpty) {
        const int64_t face = (int64_t)(uint32_t)key;
        FaceSetup s;
        face_setup(ndc, faces, face, res, s);
        const int i = (int)(p / res), j = (int)(p - (int64_t)i * res);
        float wa, wb, wc, z;
        face_eval(s.f, s.inv, pix_ndc(j, res), pix_ndc(i, res), wa, wb, wc, z);
        const int ai = faces[3 * face], bi = faces[3 * face + 1], ci = faces[3 * face + 2];
        float A[3], B[3], C[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) { A[k] = cam[3 * (int64_t)ai + k]; B[k] = cam[3 * (int64_t)bi + k]; C[k] = cam[3 * (int64_t)ci + k]; }
#pragma unroll
        for (int k = 0; k < 3; ++k) c[k] = add(add(mul(wa, A[k]), mul(wb, B[k])), mul(wc, C[k]));
        // face normal in the camera frame (rasterizer_kernel.cu:375-387)
        const float v1x = sub(B[0], A[0]), v2x = sub(C[0], A[0]), v1y = sub(B[1], A[1]), v2y = sub(C[1], A[1]);
        const float v1z = sub(B[2], A[2]), v2z = sub(C[2], A[2]);
        n[0] = sub(mul(v1y, v2z), mul(v1z, v2y));
        n[1] = sub(mul(v1z, v2x), mul(v1x, v2z));
        n[2] = sub(mul(v1x, v2y), mul(v1y, v2x));
        const float nlen = __fsqrt_rn(add(add(mul(n[0], n[0]), mul(n[1], n[1])), mul(n[2], n[2])));
        n[0] = dvd(n[0], nlen); n[1] = dvd(n[1], nlen); n[2] = dvd(n[2], nlen);
    }
    mask[p] = c[2] > 0.f ? 1 : 0;                                  // minimal_rasterizer.py:46, before the frame change
    if (depth) depth[p] = key != kEmpty ? __uint_as_float((uint32_t)(key >> 32)) : 1e10f;
    const float d[3] = {sub(c[0], t[0]), sub(c[1], t[1]), sub(c[2], t[2])};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        coords[3 * p + k] = add(add(mul(d[0], R[k]), mul(d[1], R[3 + k])), mul(d[2], R[6 + k]));
        normals[3 * p + k] = add(add(mul(n[0], R[k]), mul(n[1], R[3 + k])), mul(n[2], R[6 + k]));
    }
}

}  // namespace

extern "C" int ccn_sim_rasterize(const float* vertices, int64_t nv, const int32_t* faces, int64_t nf, const float* R,
                                 const float* t, int32_t res, float* cam_scratch, float* ndc_scratch, uint64_t* key_scratch,
                                 int32_t* list_scratch, float* coords, float* normals, uint8_t* mask, float* depth_or_null,
                                 ccn_stream_t stream) {
    CCN_REQUIRE(nv >= 0 && nf >= 0 && nf < 0x7fffffffLL && res > 0 && res <= 16384 && R && t && key_scratch && list_scratch &&
                coords && normals && mask);
    CCN_REQUIRE(nf == 0 || (vertices && faces && cam_scratch && ndc_scratch && nv > 0));
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t npix = (int64_t)res * res;
    cudaError_t e = cudaMemsetAsync(key_scratch, 0xff, npix * 8, s);
    if (e != cudaSuccess) return (int)e;
    if ((e = cudaMemsetAsync(list_scratch, 0, 4, s)) != cudaSuccess) return (int)e;
    if (nf > 0) {
        ccn_launch(sim_project_kernel, dim3(ccn_blocks_exact(nv, 256)), dim3(256), 0, s, vertices, nv, R, t, (int)res, cam_scratch,
                   ndc_scratch);
        CCN_CHECK_LAUNCH();
        ccn_launch(sim_raster_kernel, dim3(ccn_grid(nf * 32, 256, 148LL * 16)), dim3(256), 0, s, (const float*)ndc_scratch, faces,
                   nf, (int)res, (unsigned long long*)key_scratch, list_scratch + 1, list_scratch);
        CCN_CHECK_LAUNCH();
        ccn_launch(sim_raster_big_kernel, dim3(148 * 8), dim3(256), 0, s, (const float*)ndc_scratch, faces, (int)res,
                   (unsigned long long*)key_scratch, (const int32_t*)(list_scratch + 1), (const int32_t*)list_scratch);
        CCN_CHECK_LAUNCH();
    }
    ccn_launch(sim_resolve_kernel, dim3(ccn_blocks_exact(npix, 256)), dim3(256), 0, s, (const float*)cam_scratch,
               (const float*)ndc_scratch, faces, (const unsigned long long*)key_scratch, R, t, (int)res, coords, normals, mask,
               depth_or_null);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
