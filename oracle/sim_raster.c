/* TEST INFRASTRUCTURE - CPU oracle of the scanning simulator's rasteriser.  Never imported by the product path (only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg may execute anything under oracle/).
 *
 * Restates, expression by expression in fp32:
 *   scanning_simulator/utils/minimal_rasterizer.py:14-51  rasterize(): vertices @ R^T + t, the pinhole, mask, scene frame
 *   third_party/minimal_pytorch_rasterizer/minimal_pytorch_rasterizer/camera.py:38-61  Pinhole2D.project_ndc
 *   third_party/minimal_pytorch_rasterizer/minimal_pytorch_rasterizer/cuda/rasterizer_kernel.cu:80-225  rasterize_cuda_kernel
 *   (z-buffer) and :335-465 estimate_normals_cuda_kernel (point + face normal of the closest face).
 * The reference rasteriser is CUDA-only (no CPU code path): its kernels are compiled from the reference tree into
 * oracle/_ref/ref_mpr_cuda.so (oracle/Makefile, oracle/ref_mpr_shim.cu) and run beside this restatement on the GPU box
 * (tests/test_gpu_simulator.py) - that comparison pins this file.  Differences from the reference, both deliberate:
 *   - compiled with -ffp-contract=off: every operation rounded separately (nvcc contracts a*b+c of the reference into FMAs as
 *     it sees fit, so the reference itself is only reproducible to rounding error);
 *   - where several faces lie within eps = 1e-5 of the closest depth the reference keeps whichever face's thread wrote last
 *     (a race); here the closest face wins, equal depths -> lowest face index.  A NaN depth (degenerate face) never wins.
 * Loops: faces outer, bounding-box pixels inner - O(sum of box areas), seconds for the test meshes.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    float f[9], inv[9];
    int x0, x1, y0, y1, bad;
} face_t;

static int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

static void face_setup(const float* ndc, const int32_t* faces, int64_t face, int res, face_t* s) {
    const int64_t ai = faces[3 * face], bi = faces[3 * face + 1], ci = faces[3 * face + 2];
    float* f = s->f;
    for (int k = 0; k < 3; ++k) { f[k] = ndc[3 * ai + k]; f[3 + k] = ndc[3 * bi + k]; f[6 + k] = ndc[3 * ci + k]; }
    const float eps = 1e-5f;
    s->bad = (f[2] < eps || f[5] < eps || f[8] < eps);
    if (s->bad) return;
    float* q = s->inv;
    q[0] = f[4] - f[7]; q[1] = f[6] - f[3]; q[2] = f[3] * f[7] - f[6] * f[4];
    q[3] = f[7] - f[1]; q[4] = f[0] - f[6]; q[5] = f[6] * f[1] - f[0] * f[7];
    q[6] = f[1] - f[4]; q[7] = f[3] - f[0]; q[8] = f[0] * f[4] - f[3] * f[1];
    const float denom = (f[6] * (f[1] - f[4]) + f[0] * (f[4] - f[7])) + f[3] * (f[7] - f[1]);
    for (int k = 0; k < 9; ++k) q[k] = q[k] / denom;
    const float W = (float)res;
    const float mnx = fminf(fminf(f[0], f[3]), f[6]), mxx = fmaxf(fmaxf(f[0], f[3]), f[6]);
    const float mny = fminf(fminf(f[1], f[4]), f[7]), mxy = fmaxf(fmaxf(f[1], f[4]), f[7]);
    s->x0 = clampi((int)floorf((mnx + 1.f) / 2.f * W), 0, res - 1);
    s->x1 = clampi((int)ceilf((mxx + 1.f) / 2.f * W), 0, res - 1);
    s->y0 = clampi((int)floorf((mny + 1.f) / 2.f * W), 0, res - 1);
    s->y1 = clampi((int)ceilf((mxy + 1.f) / 2.f * W), 0, res - 1);
}

static float pix_ndc(int j, int res) { return (float)(2.0 * ((double)(float)j + 0.5) / (double)res - 1.0); }

static int face_eval(const float* f, const float* q, float x, float y, float* wa, float* wb, float* wc, float* z) {
    if ((y - f[1]) * (f[3] - f[0]) > (x - f[0]) * (f[4] - f[1]) || (y - f[4]) * (f[6] - f[3]) > (x - f[3]) * (f[7] - f[4]) ||
        (y - f[7]) * (f[0] - f[6]) > (x - f[6]) * (f[1] - f[7]))
        return 0;
    float a = (q[0] * x + q[1] * y) + q[2], b = (q[3] * x + q[4] * y) + q[5], c = (q[6] * x + q[7] * y) + q[8];
    float ws = (a + b) + c;
    a = a / ws; b = b / ws; c = c / ws;
    a = a / f[2]; b = b / f[5]; c = c / f[8];
    ws = (a + b) + c;
    a = a / ws; b = b / ws; c = c / ws;
    *z = (a * f[2] + b * f[5]) + c * f[8];
    *wa = a; *wb = b; *wc = c;
    return 1;
}

/* all pointers host memory; outputs as ccn_sim_rasterize (include/ccn_b200.h).  Returns 0, -1 on bad arguments / no memory. */
int oracle_sim_rasterize(const float* vertices, int64_t nv, const int32_t* faces, int64_t nf, const float* R, const float* t,
                         int32_t res, float* coords, float* normals, uint8_t* mask, float* depth_or_null) {
    if (nv < 0 || nf < 0 || res <= 0 || !R || !t || !coords || !normals || !mask) return -1;
    const int64_t npix = (int64_t)res * res;
    float* cam = (float*)malloc(sizeof(float) * 3 * (size_t)(nv > 0 ? nv : 1));
    float* ndc = (float*)malloc(sizeof(float) * 3 * (size_t)(nv > 0 ? nv : 1));
    uint64_t* keys = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)npix);
    if (!cam || !ndc || !keys) { free(cam); free(ndc); free(keys); return -1; }
    memset(keys, 0xff, sizeof(uint64_t) * (size_t)npix);
    const float fl = (float)res, cx = (float)(res / 2);
    for (int64_t i = 0; i < nv; ++i) {
        const float x = vertices[3 * i], y = vertices[3 * i + 1], z = vertices[3 * i + 2];
        float c[3];
        for (int r = 0; r < 3; ++r) c[r] = ((x * R[3 * r] + y * R[3 * r + 1]) + z * R[3 * r + 2]) + t[r];
        cam[3 * i] = c[0]; cam[3 * i + 1] = c[1]; cam[3 * i + 2] = c[2];
        const float zz = c[2] + 1e-9f;
        ndc[3 * i] = 2.f * ((c[0] * fl + c[2] * cx) / zz) / fl - 1.f;
        ndc[3 * i + 1] = 2.f * ((c[1] * fl + c[2] * cx) / zz) / fl - 1.f;
        ndc[3 * i + 2] = c[2];
    }
    face_t s;
    for (int64_t face = 0; face < nf; ++face) {
        face_setup(ndc, faces, face, res, &s);
        if (s.bad) continue;
        for (int i = s.y0; i <= s.y1; ++i)
            for (int j = s.x0; j <= s.x1; ++j) {
                float wa, wb, wc, z;
                if (!face_eval(s.f, s.inv, pix_ndc(j, res), pix_ndc(i, res), &wa, &wb, &wc, &z)) continue;
                if (!(z == z)) continue;
                uint32_t zb;
                memcpy(&zb, &z, 4);
                const uint64_t key = ((uint64_t)zb << 32) | (uint64_t)(uint32_t)face;
                if (key < keys[(int64_t)i * res + j]) keys[(int64_t)i * res + j] = key;
            }
    }
    for (int64_t p = 0; p < npix; ++p) {
        const uint64_t key = keys[p];
        float c[3] = {0.f, 0.f, 0.f}, n[3] = {0.f, 0.f, 0.f};
        if (key != ~(uint64_t)0) {
            const int64_t face = (int64_t)(uint32_t)key;
            face_setup(ndc, faces, face, res, &s);
            const int i = (int)(p / res), j = (int)(p - (int64_t)i * res);
            float wa = 0, wb = 0, wc = 0, z = 0;
            face_eval(s.f, s.inv, pix_ndc(j, res), pix_ndc(i, res), &wa, &wb, &wc, &z);
            const float* A = cam + 3 * (int64_t)faces[3 * face];
            const float* B = cam + 3 * (int64_t)faces[3 * face + 1];
            const float* C = cam + 3 * (int64_t)faces[3 * face + 2];
            for (int k = 0; k < 3; ++k) c[k] = (wa * A[k] + wb * B[k]) + wc * C[k];
            const float v1x = B[0] - A[0], v2x = C[0] - A[0], v1y = B[1] - A[1], v2y = C[1] - A[1], v1z = B[2] - A[2], v2z = C[2] - A[2];
            n[0] = v1y * v2z - v1z * v2y;
            n[1] = v1z * v2x - v1x * v2z;
            n[2] = v1x * v2y - v1y * v2x;
            const float nlen = sqrtf((n[0] * n[0] + n[1] * n[1]) + n[2] * n[2]);
            n[0] = n[0] / nlen; n[1] = n[1] / nlen; n[2] = n[2] / nlen;
        }
        mask[p] = c[2] > 0.f ? 1 : 0;
        if (depth_or_null) {
            float z = 1e10f;
            if (key != ~(uint64_t)0) { const uint32_t zb = (uint32_t)(key >> 32); memcpy(&z, &zb, 4); }
            depth_or_null[p] = z;
        }
        const float d[3] = {c[0] - t[0], c[1] - t[1], c[2] - t[2]};
        for (int k = 0; k < 3; ++k) {
            coords[3 * p + k] = (d[0] * R[k] + d[1] * R[3 + k]) + d[2] * R[6 + k];
            normals[3 * p + k] = (n[0] * R[k] + n[1] * R[3 + k]) + n[2] * R[6 + k];
        }
    }
    free(cam); free(ndc); free(keys);
    return 0;
}
