// PolarMix of two curve-grouped sweeps on the device (reference src/data/data_utils.py:5-85
// `swap` / `rotate_copy` / `polarmix` as driven by `_apply_polarmix`, src/data/nuscenes_dataset.py:151-178 ==
// src/data/kitti_dataset.py:94-121, including the final `torch.unique(curve_idxs, return_inverse=True)`).
//
// Sweep 1 and sweep 2 arrive as ingested fields (beam-sorted xyz in sensor units, reflectance, class labels, dense
// non-decreasing curve ids).  Output, in the reference's order:
//   [ sweep-1 points outside the yaw sector ][ sweep-2 points inside it ]          (swap, if drawn)
//   [ instance points of sweep 2 ][ the same rotated by Omega_0 ][ rotated by Omega_1 ]   (rotate-paste, if drawn)
// The instance block is the reference's "concatenate per class, then stable argsort by curve id": inside one curve the
// points are ordered by (position of their class in `instance_classes`, acquisition index).  One warp per curve of
// sweep 2 emits its instance points in that order.  Curve ids follow the reference's offsets (sweep 2 + max(sweep 1),
// copies + j * max(instance ids) + max(swapped ids)) and are then densified by a change-flag scan (the sequence is
// non-decreasing, which the reference asserts).
#include <math_constants.h>

#include "common.cuh"

extern "C" int ccn_excl_scan_i32(int32_t* a, int64_t n, int32_t* total_out, ccn_stream_t stream);

namespace {

struct ClassPos { int8_t pos[256]; };      // label -> position in instance_classes, -1 = not an instance class

__device__ __forceinline__ bool in_sector(float x, float y, float lo, float hi) {
    const float yaw = -atan2f(y, x);
    return yaw > lo && yaw < hi;
}

// keep1[i] = sweep-1 point stays; take2[i] = sweep-2 point moves into sweep 1; start2[c] = first point of curve c
__global__ void pm_flags_kernel(const float* __restrict__ pos1, int64_t n1, const float* __restrict__ pos2,
                                const int64_t* __restrict__ cid2, int64_t n2, int do_swap, float lo, float hi,
                                int32_t* __restrict__ keep1, int32_t* __restrict__ take2, int32_t* __restrict__ start2) {
    ccn_pdl_wait();
    const int64_t n = n1 > n2 ? n1 : n2;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (i < n1) keep1[i] = !(do_swap && in_sector(pos1[3 * i], pos1[3 * i + 1], lo, hi));
        if (i < n2) {
            take2[i] = do_swap && in_sector(pos2[3 * i], pos2[3 * i + 1], lo, hi);
            const int64_t c = cid2[i];
            if (i == 0 || cid2[i - 1] != c) start2[c] = (int32_t)i;
            if (i == n2 - 1) start2[c + 1] = (int32_t)n2;
        }
    }
}

// one warp per curve of sweep 2: number of instance points
__global__ void pm_curve_count_kernel(const int64_t* __restrict__ y2, const int32_t* __restrict__ start2,
                                      const int32_t* __restrict__ n_curves2, ClassPos cp, int do_paste,
                                      int32_t* __restrict__ icnt, int64_t cap) {
    ccn_pdl_wait();
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5, nwarp = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t C = *n_curves2;
    for (int64_t c = warp; c < cap; c += nwarp) {
        int cnt = 0;
        if (c < C && do_paste) {
            for (int i = start2[c] + lane; i < start2[c + 1]; i += 32) cnt += cp.pos[(int)(y2[i] & 0xff)] >= 0;
            for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        }
        if (lane == 0) icnt[c] = cnt;                        // zero beyond the last curve: the scan runs over `cap`
    }
}

// exclusive scan f of 0/1.. counts with f[n] = total > 0: the last index whose count is non-zero
__device__ __forceinline__ int64_t last_selected(const int32_t* __restrict__ f, int64_t n) {
    const int32_t total = f[n];
    int64_t lo = 0, hi = n - 1;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (f[mid + 1] >= total) hi = mid; else lo = mid + 1;
    }
    return lo;
}

struct PmOut {
    float* pos; float* x; int64_t* y; int64_t* cid;
};

// kept sweep-1 points and taken sweep-2 points (stream compaction through the scanned flags)
__global__ void pm_emit_swap_kernel(const float* __restrict__ pos1, const float* __restrict__ x1,
                                    const int64_t* __restrict__ y1, const int64_t* __restrict__ cid1, int64_t n1,
                                    const float* __restrict__ pos2, const float* __restrict__ x2,
                                    const int64_t* __restrict__ y2, const int64_t* __restrict__ cid2, int64_t n2,
                                    const int32_t* __restrict__ keep1, const int32_t* __restrict__ take2, PmOut o) {
    ccn_pdl_wait();
    const int64_t n = n1 > n2 ? n1 : n2;
    const int64_t K1 = keep1[n1];
    const int64_t off2 = n1 > 0 ? cid1[n1 - 1] : 0;                    // curve_idxs2 += max(curve_idxs_orig)
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (i < n1 && keep1[i + 1] != keep1[i]) {
            const int64_t d = keep1[i];
            o.pos[3 * d] = pos1[3 * i]; o.pos[3 * d + 1] = pos1[3 * i + 1]; o.pos[3 * d + 2] = pos1[3 * i + 2];
            o.x[d] = x1[i]; o.y[d] = y1[i]; o.cid[d] = cid1[i];
        }
        if (i < n2 && take2[i + 1] != take2[i]) {
            const int64_t d = K1 + take2[i];
            o.pos[3 * d] = pos2[3 * i]; o.pos[3 * d + 1] = pos2[3 * i + 1]; o.pos[3 * d + 2] = pos2[3 * i + 2];
            o.x[d] = x2[i]; o.y[d] = y2[i]; o.cid[d] = cid2[i] + off2;
        }
    }
}

// one warp per curve of sweep 2: its instance points in (class position, acquisition index) order, three copies
__global__ void pm_emit_paste_kernel(const float* __restrict__ pos2, const float* __restrict__ x2,
                                     const int64_t* __restrict__ y2, const int64_t* __restrict__ cid2, int64_t n2,
                                     const int64_t* __restrict__ cid1, int64_t n1, const int32_t* __restrict__ start2,
                                     const int32_t* __restrict__ n_curves2, const int32_t* __restrict__ ioff,
                                     const int32_t* __restrict__ keep1, const int32_t* __restrict__ take2,
                                     ClassPos cp, int n_cls, double c0, double s0, double c1, double s1, PmOut o,
                                     int64_t cap) {
    ccn_pdl_wait();
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5, nwarp = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t C = *n_curves2;
    const int64_t I = ioff[cap];                                       // instance points in total
    if (I == 0) return;
    const int64_t K1 = keep1[n1], T2 = take2[n2];
    const int64_t off2 = n1 > 0 ? cid1[n1 - 1] : 0;
    // max curve id of the swapped sweep = id of its last element (the sequence is non-decreasing).  The last selected
    // element of a scanned flag array f (exclusive scan, f[n] = total) is the smallest i with f[i + 1] == total.
    int64_t m_out = 0;
    if (T2 > 0) m_out = cid2[last_selected(take2, n2)] + off2;
    else if (K1 > 0) m_out = cid1[last_selected(keep1, n1)];
    // max instance id: dense ids, curve c of sweep 2 carries id c + off2; the last curve that holds an instance point
    const int64_t max0 = last_selected(ioff, C) + off2;
    const int64_t base_out = K1 + T2;
    for (int64_t c = warp; c < C; c += nwarp) {
        const int a = start2[c], b = start2[c + 1];
        int64_t run = ioff[c];
        if (ioff[c + 1] == run) continue;
        for (int p = 0; p < n_cls; ++p) {
            for (int i0 = a; i0 < b; i0 += 32) {
                const int i = i0 + lane;
                const bool mine = i < b && cp.pos[(int)(y2[i] & 0xff)] == p;
                const unsigned m = __ballot_sync(0xffffffffu, mine);
                if (mine) {
                    const int64_t r = run + __popc(m & ((1u << lane) - 1u));
                    const float px = pos2[3 * i], py = pos2[3 * i + 1], pz = pos2[3 * i + 2];
                    const int64_t id = cid2[i] + off2;
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        const int64_t d = base_out + j * I + r;
                        float qx = px, qy = py;
                        if (j) {                                       // np.dot(f32 xyz, f64 rot_mat) rounded into f32
                            const double cc = j == 1 ? c0 : c1, ss = j == 1 ? s0 : s1;
                            qx = (float)__dadd_rn(__dmul_rn((double)px, cc), __dmul_rn((double)py, -ss));
                            qy = (float)__dadd_rn(__dmul_rn((double)px, ss), __dmul_rn((double)py, cc));
                        }
                        o.pos[3 * d] = qx; o.pos[3 * d + 1] = qy; o.pos[3 * d + 2] = pz;
                        o.x[d] = x2[i]; o.y[d] = y2[i];
                        o.cid[d] = id + j * max0 + m_out;
                    }
                }
                run += __popc(m);
            }
        }
    }
}

// dense curve ids (torch.unique(..., return_inverse=True) of a non-decreasing sequence): change flags -> scan
__global__ void pm_change_flags_kernel(const int64_t* __restrict__ cid, const int32_t* __restrict__ keep1, int64_t n1,
                                       const int32_t* __restrict__ take2, int64_t n2, const int32_t* __restrict__ ioff,
                                       int64_t cap, int do_paste, int32_t* __restrict__ flag, int64_t n_max,
                                       int32_t* __restrict__ n_out) {
    ccn_pdl_wait();
    const int64_t N = (int64_t)keep1[n1] + take2[n2] + (do_paste ? 3 * (int64_t)ioff[cap] : 0);
    if (blockIdx.x == 0 && threadIdx.x == 0) *n_out = (int32_t)N;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n_max; i += (int64_t)gridDim.x * blockDim.x)
        flag[i] = (i > 0 && i < N && cid[i] != cid[i - 1]) ? 1 : 0;
}

__global__ void pm_dense_ids_kernel(const int32_t* __restrict__ flag, int64_t n_max, const int32_t* __restrict__ n_out,
                                    int64_t* __restrict__ cid, int32_t* __restrict__ n_curves_out) {
    ccn_pdl_wait();
    // flag holds the EXCLUSIVE scan of the change flags: inclusive = exclusive of the next element
    const int64_t N = *n_out;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x)
        cid[i] = flag[i + 1];
    if (blockIdx.x == 0 && threadIdx.x == 0) *n_curves_out = N > 0 ? flag[N] + 1 : 0;
}

struct Augment {
    double c, s;
    int flip;
    float scale;
    double tx, ty, tz;
};

__global__ void augment_points_kernel(const float* __restrict__ in, int64_t n, int on, Augment a, float pos_div,
                                      float* __restrict__ out) {
    ccn_pdl_wait();
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
        float px = in[3 * j], py = in[3 * j + 1], pz = in[3 * j + 2];
        if (on) {
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
        out[3 * j] = __fdiv_rn(px, pos_div);
        out[3 * j + 1] = __fdiv_rn(py, pos_div);
        out[3 * j + 2] = __fdiv_rn(pz, pos_div);
    }
}

}  // namespace

extern "C" int ccn_polarmix(const float* pos1, const float* x1, const int64_t* y1, const int64_t* cid1, int64_t n1,
                            const float* pos2, const float* x2, const int64_t* y2, const int64_t* cid2, int64_t n2,
                            const int32_t* n_curves2, int32_t do_swap, float alpha, float beta, int32_t do_paste,
                            const int32_t* instance_classes, int32_t n_classes, double omega0, double omega1,
                            int32_t* scratch, float* pos_out, float* x_out, int64_t* y_out, int64_t* cid_out,
                            int32_t* counts_out, ccn_stream_t stream) {
    // instance_classes: HOST array.  Outputs hold up to n1 + 4 n2 points; counts_out: int32 [2] = {points, curves}.
    // scratch: int32 [2 n1 + 7 n2 + 8].  cid1 / cid2: dense non-decreasing ids starting at 0 (ccn_ingest_scan's output).
    CCN_REQUIRE(n1 > 0 && n2 > 0 && n1 < 0x3fffffff && n2 < 0x1fffffff && n_classes >= 0 && n_classes <= 127);
    CCN_REQUIRE(pos1 && x1 && y1 && cid1 && pos2 && x2 && y2 && cid2 && n_curves2 && scratch && counts_out);
    CCN_REQUIRE(pos_out && x_out && y_out && cid_out && (n_classes == 0 || instance_classes));
    cudaStream_t s = (cudaStream_t)stream;
    ClassPos cp;
    for (int i = 0; i < 256; ++i) cp.pos[i] = -1;
    for (int p = 0; p < n_classes; ++p) {
        CCN_REQUIRE(instance_classes[p] >= 0 && instance_classes[p] < 256);
        if (cp.pos[instance_classes[p]] < 0) cp.pos[instance_classes[p]] = (int8_t)p;
    }
    const int64_t cap = n2;                                   // curves of sweep 2 <= n2
    const int64_t n_max = n1 + 4 * n2;
    int32_t* keep1 = scratch;                                 // [n1 + 1]
    int32_t* take2 = keep1 + (n1 + 1);                        // [n2 + 1]
    int32_t* start2 = take2 + (n2 + 1);                       // [n2 + 1]
    int32_t* icnt = start2 + (n2 + 2);                        // [n2 + 1]
    int32_t* flag = icnt + (n2 + 2);                          // [n_max + 1]
    ccn_launch(pm_flags_kernel, dim3(ccn_grid(n1 > n2 ? n1 : n2, 256, 148 * 8)), dim3(256), 0, s, pos1, n1, pos2, cid2, n2, do_swap, alpha, beta,
                                                                               keep1, take2, start2);
    CCN_CHECK_LAUNCH();
    ccn_launch(pm_curve_count_kernel, dim3(ccn_grid(cap * 32, 256, 148 * 8)), dim3(256), 0, s, y2, start2, n_curves2, cp, do_paste, icnt, cap);
    CCN_CHECK_LAUNCH();
    int rc = ccn_excl_scan_i32(keep1, n1, nullptr, stream);
    if (rc == CCN_OK) rc = ccn_excl_scan_i32(take2, n2, nullptr, stream);
    if (rc == CCN_OK) rc = ccn_excl_scan_i32(icnt, cap, nullptr, stream);
    if (rc != CCN_OK) return rc;
    PmOut o{pos_out, x_out, y_out, cid_out};
    ccn_launch(pm_emit_swap_kernel, dim3(ccn_grid(n1 > n2 ? n1 : n2, 256, 148 * 8)), dim3(256), 0, s, pos1, x1, y1, cid1, n1, pos2, x2, y2, cid2,
                                                                                   n2, keep1, take2, o);
    CCN_CHECK_LAUNCH();
    if (do_paste) {
        ccn_launch(pm_emit_paste_kernel, dim3(ccn_grid(cap * 32, 256, 148 * 8)), dim3(256), 0, s, 
            pos2, x2, y2, cid2, n2, cid1, n1, start2, n_curves2, icnt, keep1, take2, cp, n_classes, cos(omega0), sin(omega0),
            cos(omega1), sin(omega1), o, cap);
        CCN_CHECK_LAUNCH();
    }
    ccn_launch(pm_change_flags_kernel, dim3(ccn_grid(n_max, 256, 148 * 8)), dim3(256), 0, s, cid_out, keep1, n1, take2, n2, icnt, cap, do_paste,
                                                                          flag, n_max, counts_out);
    CCN_CHECK_LAUNCH();
    rc = ccn_excl_scan_i32(flag, n_max, nullptr, stream);
    if (rc != CCN_OK) return rc;
    ccn_launch(pm_dense_ids_kernel, dim3(ccn_grid(n_max, 256, 148 * 8)), dim3(256), 0, s, flag, n_max, counts_out, cid_out, counts_out + 1);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}

extern "C" int ccn_augment_points(const float* pos_in, int64_t n, int32_t augment, double rot_c, double rot_s,
                                  int32_t flip_type, float scale, double tx, double ty, double tz, float pos_div,
                                  float* pos_out, ccn_stream_t stream) {
    CCN_REQUIRE(n >= 0 && flip_type >= 0 && flip_type <= 3 && pos_div != 0.f);
    if (n == 0) return CCN_OK;
    CCN_REQUIRE(pos_in && pos_out);
    Augment a;
    a.c = rot_c; a.s = rot_s; a.flip = flip_type; a.scale = scale; a.tx = tx; a.ty = ty; a.tz = tz;
    ccn_launch(augment_points_kernel, dim3(ccn_grid(n, 256, 148 * 8)), dim3(256), 0, (cudaStream_t)stream, pos_in, n, augment, a, pos_div, pos_out);
    CCN_CHECK_LAUNCH();
    return CCN_OK;
}
