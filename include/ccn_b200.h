/* libccn_b200.so - C-ABI of the B200-native (sm_100a) CurveCloudNet per-scan hot path.
 *
 * Every entry point is what the reference's FFI for this path would bind.  The reference reaches
 * native code in two ways, and each function below names the interface it replaces:
 *   (1) its vendored pybind module `frnn._C` (/root/reference/third_party/FRNN/frnn/csrc/ext.cpp:10-27:
 *       insert_points_cuda, counting_sort_cuda, find_nbrs_cuda) and `prefix_sum.prefix_sum_cuda`
 *       (third_party/FRNN/external/prefix_sum/prefix_sum.h:6-20);
 *   (2) un-vendored python packages whose CUDA kernels it calls from src/models/utils/point_ops.py
 *       and src/models/modules/{fast_conv1d,pointnet2,point_conv,dgcnn,fps_ops}.py (pytorch3d.ops, torch_scatter, torch_geometric.nn.MLP, cuDNN conv1d).
 *
 * Conventions (same as the reference's pybind functions: outputs pre-allocated by the caller,
 * frnn.py:91-103):
 *   - all pointers are DEVICE pointers; the library never allocates, never synchronises and never
 *     touches the default stream: every launch goes to `stream` (the caller passes torch's current
 *     stream, as find_nbrs.cu:596-618 does);
 *   - features are fp32 row-major with an explicit leading dimension, positions are fp32 (n,3)
 *     contiguous, reference-facing index vectors (`batch`, `curve_idxs`, sampled `idx`) are int64 as
 *     in the reference API, internal neighbour tables / CSR pointers are int32;
 *   - random draws the reference takes from torch's RNG (CurveFPS phase, VoxelFPS noise, FPS start)
 *     are INPUTS;
 *   - return 0 on success, a negative CCN_E* on bad arguments, a positive cudaError_t otherwise.
 *   - data-dependent sizes are written to small device `counts` buffers the caller reads back.
 */
#ifndef CCN_B200_H
#define CCN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* ccn_stream_t;

#define CCN_OK 0
#define CCN_EINVAL (-1)
#define CCN_EUNSUPPORTED (-2)
#define CCN_ETOPOLOGY (-3)    /* batch / curve ids not sorted (the reference asserts, point_ops.py:31,49): ccn_exec_topology_flags */
#define CCN_EVOXEL_RANGE (-4) /* VoxelFPS: voxel coordinates out of the supported range */
#define CCN_ENOMEM (-5)       /* ccn_forward_run: the arena is too small (*arena_needed = bytes used when it ran out) */
#define CCN_ESCRATCH_LIMIT (-6) /* ccn_fps_static: the input is larger than the library's built-in scratch array */

#define CCN_ACT_NONE 0
#define CCN_ACT_RELU 1
#define CCN_ACT_LEAKY 2

#define CCN_AGG_MAX 0
#define CCN_AGG_MEAN 1
#define CCN_AGG_WSUM 2   /* sum sigmoid(att) * msg */
#define CCN_AGG_ATTEND 3 /* sum softmax_target(att) * msg */
#define CCN_AGG_SUM 4
#define CCN_AGG_WSUM_NORM 5 /* sum msg * sigmoid(att) / clamp(sum sigmoid(att), 1e-3): dense SGCNN 'weighted-sum' (dgcnn.py:190-197) */

#define CCN_TOPO_ERR_BATCH_UNSORTED 1
#define CCN_TOPO_ERR_CURVE_UNSORTED 2
#define CCN_TOPO_ERR_TOO_MANY_SAMPLES 4
#define CCN_TOPO_ERR_BATCH_NOT_DENSE 8

int ccn_version(void);
/* number of kernels this library has launched in this process (bench.py reports it as gpu_launches). */
int64_t ccn_kernel_launches(void);
/* Programmatic dependent launch of every kernel of this library (cudaLaunchAttributeProgrammaticStreamSerialization + a
 * griddepcontrol.wait before the kernel's first global access): on by default (CCN_PDL=0 in the environment turns it off);
 * on = 1 / 0 sets it, on < 0 only queries.  Returns the previous setting.  Results do not depend on it. */
int ccn_set_dependent_launch(int32_t on);

/* ---- curve / sample topology -------------------------------------------------------------------
 * replaces curveidx_local2global + batch2ptr + is_sorted asserts
 * (src/models/utils/point_ops.py:20-54,464-465; fast_conv1d.py:41).
 * curve_ord[i]  dense global curve ordinal of point i; curve_ptr[0..n_curves] CSR over points;
 * sample_ptr[0..B] CSR over points; sample_curve_off[b] reference offset of sample b;
 * curve_glob[i] = reference's global id; counts = {B, n_curves, CCN_TOPO_ERR_* flags, max glob id + 1}.
 * curve_ptr needs n+1 entries, sample_ptr/sample_curve_off max_samples+1. */
int ccn_curve_csr(const int64_t* batch, const int64_t* curve, int64_t n, int32_t* curve_ord, int32_t* curve_ptr,
                  int32_t* sample_ptr, int64_t* sample_curve_off, int64_t* curve_glob, int32_t* counts,
                  int32_t max_samples, ccn_stream_t stream);
/* batch2ptr(batch, with_ends=True) for levels without curves; counts = {B, 0, flags, 0}. */
int ccn_sample_ptr(const int64_t* batch, int64_t n, int32_t* sample_ptr, int32_t* counts, int32_t max_samples,
                   ccn_stream_t stream);
/* sample CSR of the level selected by idx[0..*count_ptr) out of a batch-sorted parent (the reference gets it by
 * recomputing batch2ptr on batch[idx]); sptr_out[0..B], sptr_out[B] = *count_ptr.  capacity >= *count_ptr. */
int ccn_subset_sample_ptr(const int64_t* batch, const int64_t* idx, const int32_t* count_ptr, int64_t capacity,
                          int32_t n_samples, int32_t* sptr_out, ccn_stream_t stream);

/* ---- symmetric curve convolution (src/models/modules/fast_conv1d.py:11-205; cuDNN F.conv1d :183) ---
 * feat_prep: out[i] = [x[i] (cx), pos[i] (3 if with_xyz), |mean fwd/bwd difference along the curve| (if with_diff)]
 *            (fast_conv1d.py:42-45,64-66 and compute_feature_diffs :190-205).
 * row_map:   row2pt[r] = point of padded row r or -1; `pad` zero rows precede every curve and follow the
 *            last one (fast_conv1d.py:48-61); rows = n + (n_curves+1)*pad.
 * conv_layer: one layer over `rows` signal rows: out = leaky_relu(conv1d_same(in; w) + bias), w is
 *            (ksize, cin, cout) = mirrored weight (fast_conv1d.py:173-184) with BatchNorm(eval) folded.
 *            in_mapped: input rows are points addressed through row2pt (pad rows read zeros);
 *            out_mapped: only rows with row2pt >= 0 are written, to out[row2pt[r]]. */
/* feat_planes: feat_prep + the gather to the padded signal rows (row2pt of row_map) + ccn_pack_planes in one pass: the
 *            split-bf16 planes (rows, kpad) the first tensor-core conv layer reads; bit-identical to the three calls. */
int ccn_curve_feat_planes(const float* x, int64_t ldx, int32_t cx, const float* pos, int32_t with_xyz, int32_t with_diff,
                          const int32_t* curve_ord, int64_t n, const int32_t* row2pt, int64_t rows, void* planes,
                          int64_t kpad, int64_t plane_stride, int32_t n_planes, ccn_stream_t stream);
int ccn_curve_feat_prep(const float* x, int64_t ldx, int32_t cx, const float* pos, int32_t with_xyz,
                        int32_t with_diff, const int32_t* curve_ord, int64_t n, float* out, int64_t ldo,
                        ccn_stream_t stream);
int ccn_curve_row_map(const int32_t* curve_ord, int64_t n, int32_t n_curves, int32_t pad, int32_t* row2pt,
                      ccn_stream_t stream);
int ccn_curve_conv_layer(const float* in, int64_t ld_in, const float* w, const float* bias, float* out,
                         int64_t ld_out, const int32_t* row2pt, int64_t rows, int32_t cin, int32_t cout,
                         int32_t ksize, int32_t in_mapped, int32_t out_mapped, float slope, ccn_stream_t stream);

/* ---- sampling / grouping along curves --------------------------------------------------------------
 * curve_fps: CurveFPS.forward (src/models/modules/fps_ops.py:16-39); u = the torch.rand(1) draw (:30);
 *            idx_out ascending (capacity n), *count_out = number kept; run_scratch n + ceil(n / 4096) floats.
 * radius_1d_group_{count,fill}: radius_1d_group_subset (point_ops.py:143-193) as CSR over queries:
 *            count writes row_ptr[0..nq] and *total_out = E; fill writes col[E] (point indices, in the
 *            reference's candidate order 0,-1,+1,...).  kc_scratch n_curves floats, stats_scratch 2 ints.
 * knn_1d_superset: knn_1d_group_superset + weights of knn_interpolate_1D_pytorch3d (point_ops.py:196-260,
 *            344-355): nbr (n,knn) sampled ORDINALS (-1 padded) by (distance, candidate order), wgt = 1/clamp(d^2,1e-16). */
int ccn_curve_fps(const float* pos, const int32_t* curve_ord, const int32_t* curve_ptr, int64_t n, const float* u,
                  float arclen, float* run_scratch, int64_t* idx_out, int32_t* count_out, ccn_stream_t stream);
/* radius_1d_curve_stats + radius_1d_group_count_ready = ccn_radius_1d_group_count in two calls: the per-curve statistics
 *            (point_ops.py:149-167; kc, stats) need the level only, not the queries - a caller can compute them early, on
 *            another stream - and count_ready then does the per-query part with them. */
int ccn_radius_1d_curve_stats(const float* pos, const int32_t* curve_ord, const int32_t* curve_ptr, int32_t n_curves,
                              float radius, float* kc_scratch, int32_t* stats_scratch, ccn_stream_t stream);
int ccn_radius_1d_group_count_ready(const int32_t* curve_ord, const int64_t* curve_local, int64_t n, int32_t n_curves,
                                    const int64_t* idxs, int64_t nq, const float* kc_ready, const int32_t* stats_ready,
                                    int32_t* row_ptr, int32_t* total_out, ccn_stream_t stream);
int ccn_radius_1d_group_count(const float* pos, const int32_t* curve_ord, const int32_t* curve_ptr,
                              const int64_t* curve_local, int64_t n, int32_t n_curves, const int64_t* idxs, int64_t nq,
                              float radius, float* kc_scratch, int32_t* stats_scratch, int32_t* row_ptr,
                              int32_t* total_out, ccn_stream_t stream);
int ccn_radius_1d_group_fill(const int32_t* curve_ord, const int64_t* curve_local, int64_t n, int32_t n_curves,
                             const int64_t* idxs, int64_t nq, const float* kc_scratch, const int32_t* stats_scratch,
                             const int32_t* row_ptr, int32_t* col, ccn_stream_t stream);
int ccn_knn_1d_superset(const float* pos, const int32_t* curve_ord, int64_t n, const int64_t* idxs, int64_t nq,
                        int32_t knn, int32_t* nbr, float* wgt, ccn_stream_t stream);
/* knn_1d_group_subset (src/models/utils/point_ops.py:114-140; called by no shipped configuration): for query q the points
 * idxs[q], -1, +1, -2, +2, ... along the point order that lie inside the array and on the query's curve (curve_glob =
 * curveidx_local2global), the first knn of them, the query included.  idx_out (nq, 2 knn + 1) leading valid entries, -1
 * padded; cnt_out (nq + 2) counts (the layout ccn_dense_to_csr takes). */
int ccn_knn_1d_group(const int64_t* curve_glob, int64_t n, const int64_t* idxs, int64_t nq, int32_t knn, int32_t* idx_out,
                     int32_t* cnt_out, ccn_stream_t stream);
/* group_first_k_values (src/models/utils/point_ops.py:384-428; caller: scanning_simulator/utils/curve_cloud.py:223): the first
 * kmax rows of every group id 0 .. n_groups-1 in the order of appearance (= torch.sort(batch, stable=True)), zero padded.
 * values (n, d) fp32 rows; out (n_groups, kmax, d); count_out (n_groups) = min(rows of the group, kmax).  Ids that do not
 * occur give an all-zero block with count 0 (the reference drops them: the python wrapper does). */
int ccn_group_first_k(const float* values, int64_t ldv, const int64_t* batch, int64_t n, int32_t d, int32_t n_groups,
                      int32_t kmax, float* out, int32_t* count_out, ccn_stream_t stream);
/* in-place exclusive scan of a[0..n), a[n] = total (replaces prefix_sum.prefix_sum_cuda, prefix_sum.h:6-20). */
int ccn_excl_scan_i32(int32_t* a, int64_t n, int32_t* total_out, ccn_stream_t stream);

/* ---- point down-sampling ---------------------------------------------------------------------------
 * voxel_fps: VoxelFPS.forward (fps_ops.py:51-60); noise = the torch.rand(n) draw (:56); idx_out in
 *            lexicographic (batch, ix, iy, iz) voxel order; counts = {#voxels, error flag}.
 *            slots = ccn_voxel_table_slots(n); scratch: table_keys/table_min slots u64, ukeys n u64, uwin/rank n i32.
 * fps:       pytorch3d.ops.sample_farthest_points as called by fps_pytorch3d (point_ops.py:57-70):
 *            per cloud ksel[b] picks from local start[b], ties -> lowest index; idx_out = the selected
 *            global indices SORTED ascending (point_ops.py:69); *total_out = how many. */
int64_t ccn_voxel_table_slots(int64_t n);
int ccn_voxel_fps(const float* pos, const int64_t* batch, int64_t n, float voxel, const float* noise, int64_t slots,
                  uint64_t* table_keys, uint64_t* table_min, uint64_t* ukeys, int32_t* uwin, int32_t* rank,
                  int64_t* idx_out, int32_t* counts, ccn_stream_t stream);
int ccn_fps(const float* pos, const int32_t* sample_ptr, int32_t n_samples, int64_t n, const int32_t* ksel,
            const int32_t* start, int32_t max_len, float* closest_scratch, int32_t* sel_cnt_scratch, int64_t* idx_out,
            int32_t* total_out, ccn_stream_t stream);

/* fps_static: ccn_fps WITHOUT caller-supplied scratch and without any allocation.  No counterpart in the reference: added
 *            for callers that cannot call cudaMalloc (code under CUDA graph capture,
 *            tight serving loops, harnesses that forbid allocation).  The two scratch vectors of ccn_fps (running distances,
 *            selection flags) live in a file-scope `static __device__` array of the library: CCN_FPS_STATIC_MAX_POINTS points,
 *            8 bytes each = 2 MiB, held by every GPU that loads the library until the process exits.  Same arguments otherwise,
 *            same result as ccn_fps bit for bit.  n > CCN_FPS_STATIC_MAX_POINTS -> CCN_ESCRATCH_LIMIT, nothing is launched.
 *            The call path contains no cudaMalloc / cudaFree / allocating library routine (the array's device address is
 *            looked up once per device with cudaGetSymbolAddress; the flags are cleared with cudaMemsetAsync on `stream`).
 *            EVERY call on a device shares the one array and finds the previous call's contents in it: the entry point
 *            clears what must start at zero, and calls that overlap on the GPU (different streams, no ordering between
 *            them) RACE - order them, or use ccn_fps with scratch of your own. */
#define CCN_FPS_STATIC_MAX_POINTS 262144
int ccn_fps_static(const float* pos, const int32_t* sample_ptr, int32_t n_samples, int64_t n, const int32_t* ksel,
                   const int32_t* start, int32_t max_len, int64_t* idx_out, int32_t* total_out, ccn_stream_t stream);

/* ---- neighbour search ------------------------------------------------------------------------------
 * grid_build + frnn_knn: frnn.frnn_grid_points (third_party/FRNN/frnn/frnn.py:210-347) = _C.insert_points_cuda +
 *            prefix_sum_cuda + _C.counting_sort_cuda + _C.find_nbrs_cuda (ext.cpp:10-27): up to K points of cloud
 *            batch1[q] with d2 <= r*r, ascending by (d2, index), idx = GLOBAL row of pos2 (int32) or -1.
 *            num_cells = ccn_grid_num_cells(n2); cell_off num_cells+1 ints; slot_scratch n2 ints; sorted_xyzi 4*n2 floats.
 * knn_exact: pytorch3d.ops.knn_points (call site point_ops.py:91); zero_fill mimics its 0 padding.
 * ball_query: pytorch3d.ops.ball_query (call site point_ops.py:81): first K in index order with d2 < r*r. */
int64_t ccn_grid_num_cells(int64_t n2);
int ccn_grid_build(const float* pos2, const int64_t* batch2, int64_t n2, float radius, int64_t num_cells,
                   int32_t* cell_off, int32_t* slot_scratch, float* sorted_xyzi, ccn_stream_t stream);
int ccn_frnn_knn(const float* pos1, const int64_t* batch1, int64_t n1, const float* sorted_xyzi,
                 const int32_t* cell_off, int64_t num_cells, const int32_t* sample_ptr2, float radius, int32_t K,
                 int32_t* idx_out, float* d2_out, int32_t* cnt_out, ccn_stream_t stream);
int ccn_knn_exact(const float* pos1, const int64_t* batch1, int64_t n1, const float* pos2, const int32_t* sample_ptr2,
                  int32_t K, int32_t zero_fill, int32_t* idx_out, float* d2_out, int32_t* cnt_out, ccn_stream_t stream);
/* ccn_knn_exact for K <= 4 through a hashed grid of the sources (csrc/knn_grid.cu): same result bit for bit, O(P1) instead of
 * O(P1 P2).  cell > 0: cell edge given; <= 0: derived on the device from the sources' bounding box (1.5 sqrt(A / n2)).
 * scratch: cell_off num_cells+1 ints (num_cells = ccn_grid_num_cells(n2)), slot_scratch n2 ints, sorted_xyzi 4*n2 floats,
 * open_scratch 16 + n1 ints (queries the grid could not close - sparse regions, clouds smaller than K - are finished by a
 * brute-force pass; their count never leaves the device). */
int ccn_knn_grid(const float* pos1, const int64_t* batch1, int64_t n1, const float* pos2, const int64_t* batch2, int64_t n2,
                 const int32_t* sample_ptr2, int32_t K, int32_t zero_fill, float cell, int64_t num_cells, int32_t* cell_off,
                 int32_t* slot_scratch, float* sorted_xyzi, int32_t* open_scratch, int32_t* idx_out, float* d2_out,
                 int32_t* cnt_out, ccn_stream_t stream);
int ccn_ball_query(const float* pos1, const int64_t* batch1, int64_t n1, const float* pos2, const int32_t* sample_ptr2,
                   float radius, int32_t K, int32_t* idx_out, int32_t* cnt_out, ccn_stream_t stream);
/* Neighbour search in FEATURE space - the graph of DynamicEdgeConv (src/models/modules/dgcnn.py:64-89: DGCNNLayer :98-111,
 * DGCNNLayerRadius :114-127), which the reference rebuilds from the rows [x, pos] of every layer through
 * knn_ball_group_pytorch3d (point_ops.py:73-111).  x1 (n1, dim) / x2 (n2, dim) fp32 rows with leading dimensions ld1 / ld2;
 * a query searches the cloud sample_ptr2[batch1[q]] .. of x2; squared distance summed over the columns in order, every
 * operation rounded.  mode 0: fixed-radius kNN, d2 <= radius^2, the K <= 32 smallest by (d2, index) (fast_knn,
 * point_ops.py:84-88, frnn.frnn_grid_points on dim-D points); mode 1: d2 < radius^2, the first K in index order
 * (pytorch3d.ops.ball_query, point_ops.py:81, K = 128 there).  idx_out (n1, K) global row indices of x2, -1 padded;
 * d2_out (n1, K) squared distances, -1 padded (mode 0 only, or null); cnt_out (n1) or null.  Brute force,
 * O(n1 * cloud * dim): no shipped configuration reaches these steps. */
int ccn_feat_neighbours(const float* x1, int64_t ld1, const int64_t* batch1, int64_t n1, const float* x2, int64_t ld2,
                        const int32_t* sample_ptr2, int32_t dim, float radius, int32_t K, int32_t mode, int32_t* idx_out,
                        float* d2_out, int32_t* cnt_out, ccn_stream_t stream);

/* ---- gathers and segmented aggregation ---------------------------------------------------------------
 * replace x[idx] / torch_scatter / torch_geometric.utils.softmax / frnn_gather on the path
 * (point_conv.py:60-93, dgcnn.py:166-206, pointnet2.py:95-99, point_ops.py:336-355). */
int ccn_gather_rows_f32(const float* src, int64_t lds, const int64_t* idx, int64_t n, int32_t c, float* dst,
                        int64_t ldd, ccn_stream_t stream);
int ccn_gather_i64(const int64_t* src, const int64_t* idx, int64_t n, int64_t* dst, ccn_stream_t stream);
/* pos[idx], batch[idx], curve_idxs[idx] of a down-sampled level in one launch (src/models/modules/pointnet2.py:69-72,
 * 176-179); curve / curveq may be null.  pos_planes (optional): the activation planes of posq (n_planes x n x 16 bf16, xyz +
 * zero columns; ccn_pack_planes layout) written by the same launch - positions are concatenated into the next MLP's input. */
int ccn_subset_gather(const float* pos, int64_t ldp, const int64_t* batch, const int64_t* curve, const int64_t* idx,
                      int64_t n, float* posq, int64_t* batchq, int64_t* curveq, void* pos_planes, int64_t plane_stride,
                      int32_t n_planes, ccn_stream_t stream);
int ccn_dense_to_csr(const int32_t* idx, const int32_t* row_ptr, int64_t nq, int32_t K, int32_t* erow, int32_t* ecol,
                     ccn_stream_t stream);
int ccn_csr_rows(const int32_t* row_ptr, int64_t nq, int32_t* erow, ccn_stream_t stream);
int ccn_edge_layer1(const float* H, int64_t ldh, const float* wp, const float* bias, const float* pos_src,
                    const float* pos_dst, const int32_t* erow, const int32_t* ecol, int64_t E, int32_t c,
                    float normalize_radius, int32_t act, float slope, float* out, int64_t ldo, ccn_stream_t stream);
/* first layer of StaticEdgeConv.forward_slow's message nn([x_i, x_j - x_i]) (dgcnn.py:227-228), decomposed:
 * out[e] = act(PQ[ecol[e], 0:c] + PQ[erow[e], c:2c]),  PQ = x [Wb ; Wa - Wb]^T (+ bias on the second half). */
int ccn_edge_pair_layer1(const float* PQ, int64_t ldpq, const int32_t* erow, const int32_t* ecol, int64_t E, int32_t c,
                         int32_t act, float slope, float* out, int64_t ldo, ccn_stream_t stream);
int ccn_seg_aggregate(const float* msg, int64_t ldm, const float* att, int64_t lda, const int32_t* row_ptr, int64_t nq,
                      int32_t c, int32_t mode, float* out, int64_t ldo, ccn_stream_t stream);
int ccn_sgcnn_layer1(const float* PQ, int64_t ldpq, const int32_t* idx, int64_t n, int32_t K, int32_t c, int32_t act,
                     float slope, float* out, int64_t ldo, ccn_stream_t stream);
int ccn_group_max_dense(const float* feat, int64_t ldf, const int32_t* idx, int64_t n, int32_t K, int32_t c, float* out,
                        int64_t ldo, ccn_stream_t stream);
int ccn_interp_blend(const float* x, int64_t ldx, const int32_t* nbr, const float* wgt, int32_t from_d2, int64_t n,
                     int32_t knn, int32_t c, float* out, int64_t ldo, ccn_stream_t stream);
/* The three producers above with the result written twice in ONE launch: fp32 rows (as above) and split-bf16 activation
 * planes (layout of ccn_pack_planes; kpad == c, c % 16 == 0, 16-byte aligned fp32 rows) - the consumers of these step outputs
 * are Linear layers (skip-connect / feature-propagation MLPs, base.py:179-209), so the separate pack launch disappears. */
int ccn_seg_aggregate_both(const float* msg, int64_t ldm, const float* att, int64_t lda, const int32_t* row_ptr, int64_t nq,
                           int32_t c, int32_t mode, float* out, int64_t ldo, void* planes, int64_t kpad,
                           int64_t plane_stride, int32_t n_planes, ccn_stream_t stream);
int ccn_group_max_dense_both(const float* feat, int64_t ldf, const int32_t* idx, int64_t n, int32_t K, int32_t c, float* out,
                             int64_t ldo, void* planes, int64_t kpad, int64_t plane_stride, int32_t n_planes,
                             ccn_stream_t stream);
int ccn_interp_blend_both(const float* x, int64_t ldx, const int32_t* nbr, const float* wgt, int32_t from_d2, int64_t n,
                          int32_t knn, int32_t c, float* out, int64_t ldo, void* planes, int64_t kpad, int64_t plane_stride,
                          int32_t n_planes, ccn_stream_t stream);

/* ---- shared MLP layer (torch_geometric.nn.MLP -> nn.Linear/BatchNorm1d/act; base.py:32, mlp.py:16-22) ----
 * Y = act(A W^T + bias), W is (N, K) row-major with BatchNorm(eval) folded by the caller. */
int ccn_linear_f32(const float* A, int64_t lda, const float* W, int64_t ldw, const float* bias, float* Y, int64_t ldy,
                   int64_t M, int32_t N, int32_t K, int32_t act, float slope, ccn_stream_t stream);


/* Tensor-core path of the same layer (tcgen05.mma kind::f16, fp32 accumulation in TMEM, weights by TMA):
 *   ccn_split_bf16         W fp32 (n,k) -> bf16 hi (n,kpad) [+ residual lo], kpad % 64 == 0, zero padded;
 *   ccn_weight_tensor_map  HOST call: fills a 128-byte CUtensorMap (host memory) for such a weight, box = 64 x box_n;
 *   ccn_linear_tc          mode 1: bf16 (2e-2 budget); mode 3: split-bf16 x3 (A_hi W_hi + A_lo W_hi + A_hi W_lo, fp32-class
 *                          accuracy for the 1e-4 budget).  A and Y stay fp32 in HBM; box_n in {32, 64, 128, 256}.
 * ccn_linear_tc is the STAND-ALONE form for a caller that binds only this layer and keeps fp32 activations; the model path of
 * this library does not use it (its activations travel as split-bf16 planes: ccn_linear_planes / ccn_mlp_chain below). */
int ccn_split_bf16(const float* w, int64_t ldw, int32_t n, int32_t k, int32_t kpad, void* hi, void* lo,
                   ccn_stream_t stream);
int ccn_weight_tensor_map(void* tmap_out_host, const void* w_bf16, int32_t n, int32_t kpad, int32_t box_n);
int ccn_linear_tc(const float* A, int64_t lda, const void* w_hi_tmap_host, const void* w_lo_tmap_host,
                  const float* bias, float* Y, int64_t ldy, int64_t M, int32_t N, int32_t K, int32_t box_n, int32_t mode,
                  int32_t act, float slope, ccn_stream_t stream);

/* Tensor-core path with BOTH operands by TMA.  Activations travel between layers as "split-bf16 planes":
 * (planes, M, kpad) bf16, plane 0 = bf16(x), plane 1 = bf16(x - plane 0) (mode 3 only), kpad % 8 == 0 (% 16 for
 * planes written by ccn_linear_planes), pad columns 0.
 *   ccn_pack_planes    up to four fp32 column blocks (block k optionally row-gathered through idx_k) -> planes:
 *                      replaces torch.cat / x[idx] in front of an MLP (pointnet2.py:132-139,160-163; skip_connect.py:13;
 *                      mlp.py:18-20; dgcnn.py:163-164);
 *   ccn_linear_planes  Y = act([A_0 | A_1 | A_2 | A_3] W^T + bias): up to four K-segments, each its own planes tensor
 *                      (a_i, row stride ld_i, second plane ps_i elements further, k_i columns; k_i = 0 ends the list), so
 *                      a concatenation in front of a Linear (skip_connect.py:13, pointnet2.py:132-139) costs nothing;
 *                      W in the ccn_split_bf16 layout with every segment's columns starting at a multiple of 64,
 *                      through host-encoded tensor maps (ccn_weight_tensor_map); writes fp32 Y and/or output planes. */
int ccn_pack_planes(const float* s0, int64_t ld0, int32_t c0, const int64_t* i0, const float* s1, int64_t ld1,
                    int32_t c1, const int64_t* i1, const float* s2, int64_t ld2, int32_t c2, const int64_t* i2,
                    const float* s3, int64_t ld3, int32_t c3, const int64_t* i3, int64_t n, void* planes, int64_t kpad,
                    int64_t plane_stride, int32_t n_planes, ccn_stream_t stream);
int ccn_linear_planes(const void* a0, int64_t ld0, int64_t ps0, int32_t k0, const void* a1, int64_t ld1, int64_t ps1,
                      int32_t k1, const void* a2, int64_t ld2, int64_t ps2, int32_t k2, const void* a3, int64_t ld3,
                      int64_t ps3, int32_t k3, const void* w_hi_tmap_host, const void* w_lo_tmap_host, const float* bias,
                      float* Y, int64_t ldy, void* y_planes, int64_t yp_ld, int64_t yp_plane_stride, int64_t M, int32_t N,
                      int32_t box_n, int32_t mode, int32_t act, float slope, ccn_stream_t stream);
/* Y[M, N] = A[M, K] W[K, N] with W given as a (K, N) row-major split-bf16 weight: the data gradient dX = dY W of a Linear
 * layer computed from the FORWARD weight's bf16 planes (cuBLAS SGEMM `NN` in Linear.backward upstream) - no transposed
 * copy, no second split.  The B tiles are MN-major UMMA operands; the tensor maps must have a 64 x 64 box
 * (ccn_weight_tensor_map(..., box_n = 64) on the (K rows, N cols padded to 64) planes); box_n = output tile width
 * (64 | 128 | 256). */
int ccn_linear_planes_bt(const void* a, int64_t ld, int64_t ps, int32_t k, const void* w_hi_tmap_host,
                         const void* w_lo_tmap_host, float* Y, int64_t ldy, void* y_planes, int64_t yp_ld,
                         int64_t yp_plane_stride, int64_t M, int32_t N, int32_t box_n, int32_t mode, ccn_stream_t stream);

/* Planes-output variants of the per-edge first layers (the consumer is the next Linear of the edge MLP); c % 8 == 0,
 * kpad == c, fp32 inputs 16-byte aligned with leading dimensions % 4 == 0. */
int ccn_sgcnn_layer1_planes(const float* PQ, int64_t ldpq, const int32_t* idx, int64_t n, int32_t K, int32_t c,
                            int32_t act, float slope, void* planes, int64_t kpad, int64_t plane_stride,
                            int32_t n_planes, ccn_stream_t stream);
int ccn_edge_pair_layer1_planes(const float* PQ, int64_t ldpq, const int32_t* erow, const int32_t* ecol, int64_t E,
                                int32_t c, int32_t act, float slope, void* planes, int64_t kpad, int64_t plane_stride,
                                int32_t n_planes, ccn_stream_t stream);
int ccn_edge_layer1_planes(const float* H, int64_t ldh, const float* wp, const float* bias, const float* pos_src,
                           const float* pos_dst, const int32_t* erow, const int32_t* ecol, int64_t E, int32_t c,
                           float normalize_radius, int32_t act, float slope, void* planes, int64_t kpad,
                           int64_t plane_stride, int32_t n_planes, ccn_stream_t stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Training step: backward of the ops above + train-mode BatchNorm + the loss (csrc/train_ops.cu).  Replaces what
 * the reference obtains from torch autograd over cuBLAS/cuDNN/torch_scatter/pytorch3d (`loss.backward()` in
 * src/run/shapenet_seg.py:186-188, nuscenes_seg.py:229-252).  Gradient buffers that are accumulated into
 * (scatter targets, dW, db) are zero-filled by the CALLER.  fp64 scratch `sums`/`acc` is zero-filled by the callee.
 * --------------------------------------------------------------------------------------------------------------- */
/* dst[r] = idx[r] >= 0 ? src[idx[r]] : 0          (pad-row insertion of fast_conv1d.py:48-61; backward of a scatter) */
int ccn_gather_rows_i32(const float* src, int64_t lds, const int32_t* idx, int64_t n, int32_t c, float* dst,
                        int64_t ldd, ccn_stream_t stream);
/* dst[idx[r]] += src[r], idx < 0 skipped           (backward of x[idx] / x[col] gathers) */
int ccn_scatter_add_rows_i32(const float* src, int64_t lds, const int32_t* idx, int64_t n, int32_t c, float* dst,
                             int64_t ldd, ccn_stream_t stream);
int ccn_scatter_add_rows_i64(const float* src, int64_t lds, const int64_t* idx, int64_t n, int32_t c, float* dst,
                             int64_t ldd, ccn_stream_t stream);
/* dst[i] (+)= sum_g src[i*G + g]                   (backward of the per-point term of the dense SGCNN first layer) */
int ccn_group_sum_rows(const float* src, int64_t lds, int64_t n, int32_t G, int32_t c, float* dst, int64_t ldd,
                       int32_t accumulate, ccn_stream_t stream);
/* dW (N,K) += dY^T X, db (N) += column sums of dY (db may be NULL)      (backward of torch.nn.Linear) */
int ccn_linear_wgrad(const float* dY, int64_t lddy, const float* X, int64_t ldx, int64_t M, int32_t N, int32_t K,
                     float* dW, int64_t lddw, float* db, ccn_stream_t stream);
/* BatchNorm1d batch statistics: mean, invstd = 1/sqrt(var_biased + eps); running stats updated in place when given */
int ccn_bn_stats(const float* z, int64_t ldz, int64_t M, int32_t C, float eps, float momentum, double* sums,
                 float* mean, float* invstd, float* running_mean, float* running_var, ccn_stream_t stream);
/* y = act(((z - mean) invstd) gamma + beta); gamma == NULL: y = act(z).  y_planes != NULL: y is ALSO written as split-bf16
 * activation planes (row stride yp_ld >= C, pad columns zeroed) - the input format of the next Linear. */
int ccn_bn_apply(const float* z, int64_t ldz, int64_t M, int32_t C, const float* mean, const float* invstd,
                 const float* gamma, const float* beta, int32_t act, float slope, float* y, int64_t ldy, void* y_planes,
                 int64_t yp_ld, int64_t yp_plane_stride, int32_t n_planes, ccn_stream_t stream);
/* backward of ccn_bn_apply (+ the statistics when batch_stats != 0): dz, dgamma, dbeta; optionally the column sums of dz
 * (bias gradient of the Linear in front: exactly 0 under batch statistics) and dz as activation planes (the operand
 * of that Linear's data / weight gradient GEMMs). */
int ccn_bn_backward(const float* dy, int64_t lddy, const float* z, int64_t ldz, int64_t M, int32_t C,
                    const float* mean, const float* invstd, const float* gamma, const float* beta, int32_t act,
                    float slope, int32_t batch_stats, double* sums, float* dz, int64_t lddz, float* dgamma,
                    float* dbeta, float* dzsum, void* dz_planes, int64_t dzp_ld, int64_t dzp_plane_stride,
                    int32_t n_planes, ccn_stream_t stream);
int ccn_seg_aggregate_bwd(const float* msg, int64_t ldm, const float* att, int64_t lda, const int32_t* row_ptr,
                          int64_t nq, int32_t c, int32_t mode, const float* dout, int64_t ldo, float* dmsg,
                          int64_t lddm, float* datt, int64_t ldda, ccn_stream_t stream);
int ccn_group_max_dense_bwd(const float* feat, int64_t ldf, const int32_t* idx, int64_t n, int32_t K, int32_t c,
                            const float* dout, int64_t ldo, float* dfeat, int64_t lddf, ccn_stream_t stream);
int ccn_interp_blend_bwd(const int32_t* nbr, const float* wgt, int32_t from_d2, int64_t n, int32_t knn, int32_t c,
                         const float* dout, int64_t ldo, float* dx, int64_t lddx, ccn_stream_t stream);
int ccn_curve_feat_prep_bwd(const float* x, int64_t ldx, int32_t cx, int32_t with_xyz, int32_t with_diff,
                            const int32_t* curve_ord, int64_t n, const float* dout, int64_t ldo, float* dx,
                            int64_t lddx, ccn_stream_t stream);
/* rel (E,3) = (pos_src[ecol] - pos_dst[erow]) / radius (radius == 0: no division)      (point_conv.py:60-69) */
int ccn_edge_rel_pos(const float* pos_src, const float* pos_dst, const int32_t* erow, const int32_t* ecol, int64_t E,
                     float normalize_radius, float* rel, ccn_stream_t stream);
/* loss = sum_i w[y_i] (lse_i - logit[i][y_i]) / sum_i w[y_i] over y_i != ignore; lse (n) and acc (fp64 [2]) are kept
 * for the backward: dlogits = gscale[0] * w[y] / sum_w * (softmax - onehot) */
int ccn_nll_loss_fwd(const float* logits, int64_t ld, const int64_t* target, const float* class_weight, int64_t n,
                     int32_t C, int64_t ignore_index, float* lse, double* acc, float* loss, float* perpoint,
                     ccn_stream_t stream);   /* perpoint (n) or NULL: the unweighted per-point NLL (reduction="none") */
int ccn_nll_loss_bwd(const float* logits, int64_t ld, const int64_t* target, const float* class_weight, int64_t n,
                     int32_t C, int64_t ignore_index, const float* lse, const double* acc, const float* gscale,
                     float* dlogits, int64_t ldd, ccn_stream_t stream);
/* y = u >= p ? x / (1 - p) : 0 with caller-supplied uniform draws u (n*c); applied to the gradient it is the backward */
int ccn_dropout(const float* x, int64_t ldx, const float* u, int64_t n, int32_t c, float p, float* y, int64_t ldy,
                ccn_stream_t stream);
/* out (C) += column sums of src (M,C)                                               (bias gradient) */
int ccn_col_sum(const float* src, int64_t ld, int64_t M, int32_t C, float* out, ccn_stream_t stream);
/* dW (N,K) += dY^T X[. + x_row_off] on tcgen05, both operands as split-bf16 activation planes read by TMA as
 * MN-major UMMA blocks (no transposed copies); x_row_off != 0 serves the taps of the curve convolution; mode 1 = bf16,
 * 3 = split-bf16 x3.  dW is accumulated into with vector reductions (zero-filled by the caller). */
int ccn_linear_wgrad_tc(const void* dy_planes, int64_t dy_ld, int64_t dy_plane_stride, const void* x_planes,
                        int64_t x_ld, int64_t x_plane_stride, int64_t M, int32_t N, int32_t K, int32_t x_row_off,
                        int64_t x_rows, int32_t mode, float* dW, int64_t lddw, ccn_stream_t stream);
/* 'same' zero-padded 1-D convolution along the ROW axis of an activation-planes signal as one tcgen05 GEMM
 * (fast_conv1d.py:148-187 F.conv1d): K = taps x cin, tap t of output row m reads input row m + t - taps/2 through a
 * TMA row coordinate (rows outside [0, M) are zeros).  W: (N, taps x cin) in the ccn_split_bf16 layout, every tap's
 * cin columns starting at a multiple of 64.  out_row_map (optional int32 [M]): output row m is stored at row
 * out_row_map[m] of Y / y_planes and dropped when negative - the last layer of a conv stack keeps the curve points and
 * drops the pad rows (fast_conv1d.py:70-73).  Other arguments as ccn_linear_planes. */
int ccn_conv_planes(const void* a, int64_t ld, int64_t ps, int32_t cin, int32_t taps, const void* w_hi_tmap_host,
                    const void* w_lo_tmap_host, const float* bias, float* Y, int64_t ldy, void* y_planes, int64_t yp_ld,
                    int64_t yp_plane_stride, int64_t M, int32_t N, int32_t box_n, int32_t mode, int32_t act, float slope,
                    const int32_t* out_row_map, ccn_stream_t stream);
/* Curve grouping of a raw sweep - SemNuScenes._get_curves / SemKITTI._get_curves (src/data/nuscenes_dataset.py:91-118,
 * src/data/kitti_dataset.py:73-92): reorder = stable argsort of beam ids, inv_reorder its inverse, sorted_points (n,3) =
 * points[reorder], curve_idxs = running count of splits (beam change, or f64 edge length > curve_thresh * sqrt(|p.xy|)).
 * points rows have stride ldp >= 3 floats (xyz first).  scratch: int32 [n_beams + n].  counts_out: int32 [2] =
 * {number of curves, error flags (1 = beam id outside [0, n_beams))}. */
int ccn_get_curves(const float* points, int64_t ldp, const int64_t* beam, int64_t n, int32_t n_beams, float curve_thresh,
                   int32_t* scratch, int64_t* reorder, int64_t* inv_reorder, float* sorted_points, int64_t* curve_idxs,
                   int32_t* counts_out, ccn_stream_t stream);
/* Kortx scans (scanning_simulator/kortx/generate_kortx.py:38-75): ccn_kortx_curves = fix_curve_idxs (:94-104: a curve start
 * directly followed by another start is dropped; ids = running count of the remaining starts) then cut_curves (:107-117: new
 * curve where the id grows or the f64 edge exceeds cut_thresh - 0.2 for 'table' captures, else 1.0; ids from 0).  points (n,3)
 * f64, raw_curve (n) int64 (the file's column 4); scratch int32 [2 n]; curve_out (n) int64.
 * ccn_kortx_windows: n_windows windows of npoints consecutive points, window w starting at w * ((n - npoints) / n_windows)
 * (:40-42): pos_out (n_windows, npoints, 3) f64 = points / (2 max |coordinate| of the window) (:57-59), curve_out the ids
 * renumbered from 0 per window (torch.unique(return_inverse) of a non-decreasing sequence, :63), seg_out the labels. */
int ccn_kortx_curves(const double* points, const int64_t* raw_curve, int64_t n, double cut_thresh, int32_t* scratch,
                     int64_t* curve_out, ccn_stream_t stream);
int ccn_kortx_windows(const double* points, const int64_t* curve, const int64_t* segmentation, int64_t n, int32_t npoints,
                      int32_t n_windows, double* pos_out, int64_t* curve_out, int64_t* seg_out, ccn_stream_t stream);
/* Curve grouping of an A2D2 frame - SemA2D2._get_curves / _partition_curves_at_discontinuities
 * (src/data/audi_dataset.py:156-243).  A2D2 carries no beam ids: per lidar (lidar_id, 5 sensors) the reference builds a kNN
 * graph (torch_cluster.knn, k = KNN[id] <= 4, self included), keeps the edges shorter than CURVE_THRESHES[id] * sqrt(|p.xy|)
 * of the query, takes connected components (scipy: numbered by their lowest point), orders the sensor's points by
 * (component, timestamp) (np.lexsort, ties in file order), cuts again where the component changes or where the f64 edges to
 * the previous AND the second-previous point both exceed 0.9 * CURVE_THRESHES[id] * sqrt(|p.xy|), offsets the ids of a sensor
 * by the sum of the LARGEST ids of the sensors before (so neighbouring sensors share one id - upstream behaviour, kept) and
 * compacts the ids (np.unique).  points (n, >= 3) f32 rows of stride ldp, lidar_id / timestamp int64 (n); knn_host /
 * thresh_host: n_sensors HOST values; scratch: ccn_a2d2_scratch_bytes(n, n_sensors) bytes, 16-byte aligned.  Outputs in the
 * final order (sensor-major): reorder (n) = index into the input, sorted_points (n, 3), curve_idxs (n); counts_out int32
 * [2] = {number of curves, error flags (1: lidar id outside [0, n_sensors); 2: a timestamp more than 2^30 ticks after the
 * frame's first)}.  n < 2^17 points per frame. */
int64_t ccn_a2d2_scratch_bytes(int64_t n, int32_t n_sensors);
int ccn_a2d2_curves(const float* points, int64_t ldp, const int64_t* lidar_id, const int64_t* timestamp, int64_t n,
                    int32_t n_sensors, const int32_t* knn_host, const float* thresh_host, void* scratch, int64_t* reorder,
                    float* sorted_points, int64_t* curve_idxs, int32_t* counts_out, ccn_stream_t stream);
/* the rest of SemA2D2.get (audi_dataset.py:131-141): x_out[r] = reflectance[reorder[r]] / refl_div (255), y_out[r] =
 * labels[reorder[r]] (labels null: zeros; y_out null: skipped).  Positions go through ccn_augment_points (/ 30, then the
 * y flip and the x, y scale of _training_augmentation, :81-90). */
int ccn_a2d2_finish(const float* reflectance, const int64_t* labels, const int64_t* reorder, int64_t n, float refl_div,
                    float* x_out, int64_t* y_out, ccn_stream_t stream);

/* Scan ingest - SemNuScenes.get / _load_frame / _get_curves / _training_augmentation
 * (src/data/nuscenes_dataset.py:42-149) and SemKITTI.get / ... (src/data/kitti_dataset.py:40-152) as ONE call per sweep:
 * raw (n, rec) f32 records, rec = 5 (x y z intensity ring) or 4 (x y z remission, single sequential beam: n_beams = 1);
 * labels: label_bytes = 1 (u8 lidarseg), 4 (i32, masked with label_mask = 0xFFFF) or 0 (test split: ones when
 * lut_size == -1, else zeros), mapped through the learning map lut[lut_size] (int32, < 0 = key absent) and truncated to
 * u8 like the reference.  Outputs, in beam-sorted order: pos (n,3) = augment(xyz) / pos_div, x (n) reflectance, y (n) i64,
 * curve_idxs (n) i64, reorder / inv_reorder (n) i64.  augment != 0 applies the training augmentation with the DRAWN values
 * (rotation cos / sin in f64, flip_type 0..3, f32 xy scale, f64 translation) in the reference's arithmetic.
 * scratch: int32 [n_beams + n]; beam_tmp: int64 [n]; sorted_tmp: f32 [n,3]; counts_out: int32 [2] = {number of curves,
 * error flags (1 = ring id outside [0, n_beams), 2 = label without an entry in the learning map)}. */
int ccn_ingest_scan(const float* raw, int64_t n, int32_t rec, const void* labels, int32_t label_bytes, int32_t label_mask,
                    const int32_t* lut, int32_t lut_size, int32_t n_beams, float curve_thresh, int32_t augment,
                    double rot_c, double rot_s, int32_t flip_type, float scale, double tx, double ty, double tz,
                    float pos_div, int32_t* scratch, int64_t* beam_tmp, float* sorted_tmp, int64_t* reorder,
                    int64_t* inv_reorder, float* pos, float* x, int64_t* y, int64_t* curve_idxs, int32_t* counts_out,
                    ccn_stream_t stream);

/* PolarMix of two ingested sweeps - swap / rotate_copy / polarmix (src/data/data_utils.py:5-85) as driven by
 * _apply_polarmix (src/data/nuscenes_dataset.py:151-178, src/data/kitti_dataset.py:94-121), including the final
 * torch.unique(curve_idxs, return_inverse=True).  Inputs per sweep: xyz (n,3) in sensor units, reflectance (n), labels (n)
 * i64 in [0,256), curve ids (n) i64 dense, non-decreasing, starting at 0 (what ccn_ingest_scan emits with pos_div = 1 and
 * no augmentation); n_curves2: DEVICE int32 (counts_out[0] of sweep 2's ingest).  The DRAWN values come from the host:
 * do_swap / do_paste (the two np.random.random() < 0.5 gates), alpha / beta (yaw sector, compared in f32 like numpy
 * does), omega0 / omega1; instance_classes: HOST int32 [n_classes].  Outputs hold up to n1 + 4 n2 points:
 * [sweep 1 outside the sector][sweep 2 inside][instance points of sweep 2][rotated by omega0][rotated by omega1];
 * cid_out = dense ids; counts_out: DEVICE int32 [2] = {points, curves}.  scratch: int32 [2 n1 + 7 n2 + 8]. */
int ccn_polarmix(const float* pos1, const float* x1, const int64_t* y1, const int64_t* cid1, int64_t n1,
                 const float* pos2, const float* x2, const int64_t* y2, const int64_t* cid2, int64_t n2,
                 const int32_t* n_curves2, int32_t do_swap, float alpha, float beta, int32_t do_paste,
                 const int32_t* instance_classes, int32_t n_classes, double omega0, double omega1, int32_t* scratch,
                 float* pos_out, float* x_out, int64_t* y_out, int64_t* cid_out, int32_t* counts_out,
                 ccn_stream_t stream);
/* _training_augmentation + position normalisation on already grouped points (the tail of ccn_ingest_scan as its own
 * entry, for the PolarMix path where the mix sits between the grouping and the augmentation). */
int ccn_augment_points(const float* pos_in, int64_t n, int32_t augment, double rot_c, double rot_s, int32_t flip_type,
                       float scale, double tx, double ty, double tz, float pos_div, float* pos_out, ccn_stream_t stream);
/* Object scans - CurvesInMemoryDataset.get (src/data/object_datasets.py:21-38) for a batch of objects (points of object b:
 * sample_ptr[b] .. sample_ptr[b+1]): pc_normalize (:61-66: minus the mean, divided by the largest norm); with aug
 * ((n_objects, 4) f32 = {scale, shift xyz}) also the shapenet-seg training transform: torch_geometric NormalizeScale (centre,
 * * 0.999999 / max |coordinate|), RandomScale (one factor), pos += shift (:31-36).  The means are f64 sums rounded once. */
int ccn_object_normalize(const float* pos, const int32_t* sample_ptr, int32_t n_objects, const float* aug, float* pos_out,
                         ccn_stream_t stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Loss / metrics epilogue of the segmentation runs (csrc/loss_metrics.cu).
 * Lovasz-softmax (src/models/utils/lovasz_losses.py:21-33,175-204 via seg_loss, src/run/nuscenes_seg.py:229-243):
 *   loss = mean over PRESENT classes of dot(sorted errors, lovasz_grad(sorted fg)) on the points with
 *   target != ignore_index.  Buffers: keys uint64 [C * ccn_lovasz_pad(n)], fg_count int32 [C], lse f32 [n],
 *   grad f32 [C * n] (the Jaccard gradient of every (class, point), kept for the backward), acc f64 [2].
 * --------------------------------------------------------------------------------------------------------------- */
int64_t ccn_lovasz_pad(int64_t n);
int ccn_lovasz_softmax_fwd(const float* logits, int64_t ld, const int64_t* target, int64_t n, int32_t C,
                           int64_t ignore_index, void* keys, int32_t* fg_count, float* lse, float* grad, double* acc,
                           float* loss, ccn_stream_t stream);
int ccn_lovasz_softmax_bwd(const float* logits, int64_t ld, const int64_t* target, int64_t n, int32_t C,
                           int64_t ignore_index, const float* lse, const float* grad, const int32_t* fg_count,
                           const double* acc, const float* gscale, float* dlogits, int64_t ldd, int32_t accumulate,
                           ccn_stream_t stream);
/* fast_hist (src/utils/utils.py:4-8) with the arg-max over score columns [col_offset, C) fused: hist (ncls x ncls int64,
 * accumulated into) [target][pred]; pred_out optional; scores == NULL: predictions are given in pred_in. */
int ccn_confusion_hist(const float* scores, int64_t ld, int32_t C, const int64_t* pred_in, const int64_t* target, int64_t n,
                       int32_t ncls, int32_t col_offset, int64_t* pred_out, int64_t* hist, ccn_stream_t stream);
/* Part IoU of the object-segmentation runs (src/run/shapenet_seg.py:88-99), all objects of a batch in one launch: per object b
 * (points sample_ptr[b] .. sample_ptr[b+1]) the arg-max over the score columns parts[parts_ptr[c] .. parts_ptr[c+1]) of its
 * category c = category[b] (torch_geometric ShapeNet.seg_classes), the target mapped to its position in that list, and
 * torchmetrics.functional.jaccard_index(num_classes = #parts, absent_score): intersection / union per part from the confusion
 * matrix, absent_score where the union is empty, mean over the parts -> iou_out (n_objects) f32.  pred_out (optional): the
 * per-point arg-max position.  flags_out: 1 category out of range, 2 a label outside its category's parts.  max_parts <= 16. */
int ccn_part_iou(const float* scores, int64_t ld, const int64_t* target, const int32_t* sample_ptr, const int64_t* category,
                 int32_t n_objects, const int32_t* parts, const int32_t* parts_ptr, int32_t n_categories, int32_t max_parts,
                 float absent_score, float* iou_out, int64_t* pred_out, int32_t* flags_out, ccn_stream_t stream);
/* test-time augmentation of val() (src/run/nuscenes_seg.py:86-116): out = [sx x, sy y, z]; acc (+)= src;
 * out = log_softmax(acc * scale), pred = arg-max. */
int ccn_scale_xy(const float* pos, int64_t n, float sx, float sy, float* out, ccn_stream_t stream);
int ccn_accumulate(const float* src, int64_t lds, int64_t n, int32_t C, float* acc, int64_t lda, int32_t first,
                   ccn_stream_t stream);
int ccn_mean_logsoftmax(const float* acc, int64_t lda, int64_t n, int32_t C, float scale, float* out, int64_t ldo,
                        int64_t* pred, ccn_stream_t stream);

/* ---- fused edge-MLP steps (csrc/edge_fused.cu): gather -> layer 1 (decomposed) -> layers 2..n -> aggregate in ONE
 * kernel, per-edge intermediates stay in shared memory / TMEM (tcgen05).  replaces the chain
 *   frnn_gather + cat + nn(...) + masked max           (src/models/modules/dgcnn.py:166-189,  ccn_sgcnn_fused)
 *   message + aggregate of PointNetConv2 / forward_slow (src/models/modules/point_conv.py:60-93, dgcnn.py:209-246,
 *                                                        ccn_pointconv_fused)
 * ccn_chain_t (HOST memory) describes the layers AFTER the decomposed first one: all their weights live in one
 * bf16 matrix (rows = output channels of layer 0, 1, ... stacked at w_row[l], columns = input channels zero padded to
 * the widest K rounded up to 64) in the ccn_split_bf16 layout with tensor maps of box 64 x 64 (ccn_weight_tensor_map(...,
 * box_n = 64)); k[l] / n[l] multiples of 64, <= 256; k[0] = c1, k[l] = n[l-1]; bias[l] device pointers (BatchNorm
 * folded) or null. */
#define CCN_CHAIN_MAX_LAYERS 6
typedef struct ccn_chain {
    int32_t n_layers;
    int32_t mode;                              /* 1 = bf16, 3 = split-bf16 x3 */
    int32_t k[CCN_CHAIN_MAX_LAYERS], n[CCN_CHAIN_MAX_LAYERS], w_row[CCN_CHAIN_MAX_LAYERS], act[CCN_CHAIN_MAX_LAYERS];
    float slope[CCN_CHAIN_MAX_LAYERS];
    const float* bias[CCN_CHAIN_MAX_LAYERS];
    const void* w_hi_tmap;                     /* host CUtensorMap of the hi plane (64-byte aligned) */
    const void* w_lo_tmap;                     /* lo plane (mode 3) */
} ccn_chain_t;
/* dense SGCNN step: PQ (n, 2 c1) = [P | Q] per point, idx (n, K) int32 -1 padded; row (i, 0) = act0(P[i] + Q[i]) (self
 * loop), row (i, 1+k) = act0(P[idx[i][k]] + Q[i]) (P = 0 for a missing neighbour); chain; out[i] = max over the K+1 rows,
 * missing neighbours at -100.  out (n, ldo) fp32; planes (optional): the result's activation planes like
 * ccn_group_max_dense_both.  K + 1 <= 128. */
int ccn_sgcnn_fused(const float* PQ, int64_t ldpq, const int32_t* idx, int64_t n, int32_t K, int32_t c1, int32_t act0,
                    float slope0, const ccn_chain_t* chain_host, float* out, int64_t ldo, void* planes, int64_t kpad,
                    int64_t plane_stride, int32_t n_planes, ccn_stream_t stream);
/* CSR edge step (targets = rows of row_ptr[0..nq], edges sorted by target, erow/ecol per edge, every target has at most
 * max_degree edges): first layer act0(H[ecol] + Wp (pos_src[ecol] - pos_dst[erow]) [/ radius] + bias0) (PQ null; H may be
 * null) or act0(PQ[ecol][0:c1] + PQ[erow][c1:2c1]) (PQ given); chain; agg = CCN_AGG_*; for WSUM / ATTEND the chain is
 * local_nn's layers followed by attend_nn's, msg_layer = index of the chain layer whose output is the message.
 * tile_scratch: 4 * (E / (129 - max_degree) + 1) int32, 16-byte aligned (tile table written by a pre-pass). */
/* 2+ consecutive layers of a shared per-point MLP (torch_geometric.nn.MLP: Linear + folded BatchNorm + activation; base.py:32,
 * mlp.py:16-22, skip_connect.py:12-15) in ONE launch, the 128-row activation tile resident in shared memory between the
 * layers.  Input [A_0 | A_1 | A_2 | A_3] as in ccn_linear_planes (split-bf16 planes per K-segment; k_i = 0 ends the list);
 * every segment occupies whole 64-column chunks, sum of ceil64(k_i) <= 256; chain: the layers - the first layer's weight
 * columns laid out per padded segment (chain->k[0] = that sum), every width but the last a multiple of 64, all <= 256;
 * writes fp32 Y (ldy) and / or output planes (row stride yp_ld = last width rounded up to 16).  Bit-identical to
 * consecutive ccn_linear_planes launches. */
int ccn_mlp_chain(const void* a0, int64_t ld0, int64_t ps0, int32_t k0, const void* a1, int64_t ld1, int64_t ps1, int32_t k1,
                  const void* a2, int64_t ld2, int64_t ps2, int32_t k2, const void* a3, int64_t ld3, int64_t ps3, int32_t k3,
                  int64_t M, const ccn_chain_t* chain_host, float* Y, int64_t ldy, void* y_planes, int64_t yp_ld,
                  int64_t yp_plane_stride, ccn_stream_t stream);
int ccn_pointconv_fused(const float* H, int64_t ldh, const float* wp, const float* bias0, const float* pos_src,
                        const float* pos_dst, const float* PQ, int64_t ldpq, const int32_t* row_ptr, const int32_t* erow,
                        const int32_t* ecol, int64_t nq, int64_t E, int32_t max_degree, int32_t c1, float radius,
                        int32_t act0, float slope0, const ccn_chain_t* chain_host, int32_t agg, int32_t msg_layer,
                        int32_t* tile_scratch, float* out, int64_t ldo, void* planes, int64_t kpad, int64_t plane_stride,
                        int32_t n_planes, ccn_stream_t stream);

/* ---- native step executor (csrc/executor.cu): ModelBase.forward / _apply_steps (src/models/base.py:133-175) and the
 * step modules (src/models/modules/{fast_conv1d,pointnet2,point_conv,dgcnn,mlp,skip_connect}.py) as ONE call per forward
 * (inference, tensor-core precisions, the nuScenes / KITTI step family: conv1d-fast-v2, sa-geo, mlp, sgcnn (dense),
 * skip-connect, sa (voxel | fps, FRNN), fp, fp-geo).  It issues exactly the C-ABI calls of the python executor in the same
 * order, so the logits are bit-identical to it.
 *   ccn_exec_create   parses the PLAN - two flat arrays built by curvecloudnet_b200/executor.py from the prepared modules
 *                     (folded weights, bf16 planes, HOST tensor maps; all referenced memory stays owned by the caller and
 *                     must outlive the executor).
 *   ccn_forward_run   x (n, in_c) or null, pos (n, 3), batch / curve_idxs (n) int64 - DEVICE; arena: device scratch for every
 *                     intermediate (bump allocated, nothing survives the call); draw: the reference's random draws in its
 *                     call order - kind 0: CurveFPS phase (fps_ops.py:30) -> *(float*)out; kind 1: VoxelFPS noise
 *                     (fps_ops.py:56), n floats -> out (HOST buffer); kind 2: FPS start of a cloud of n points
 *                     (point_ops.py:64) -> return value; logits (n, ld_logits) fp32 DEVICE.  Synchronises `stream` at
 *                     the data-dependent sizes (level sizes, edge counts).  CCN_ENOMEM: enlarge the arena, call again. */
typedef struct ccn_exec ccn_exec_t;
typedef int64_t (*ccn_draw_fn)(void* user, int32_t kind, int64_t n, void* out);
int ccn_exec_create(const int64_t* iplan_host, int64_t n_i, const double* fplan_host, int64_t n_f, ccn_exec_t** out);
void ccn_exec_destroy(ccn_exec_t* exec);
int32_t ccn_exec_topology_flags(const ccn_exec_t* exec);
/* optional per-family device timing of ccn_forward_run (CUDA events on `stream` around every launch of a family, a spin of
 * spin_cycles GPU cycles queued before each so that the pair brackets the kernel, not the launch gap): out[family][4] =
 * {ms, launches, algorithmic FLOPs, algorithmic bytes} of the last forward; families: 0 Linear, 1 fused edge
 * steps, 2 curve-conv GEMMs, 3 neighbour search, 4 samplers.  n_families >= 5. */
int ccn_exec_profile(ccn_exec_t* exec, int32_t enable, int32_t spin_cycles);
/* host-side accounting of the last ccn_forward_run: out[3] = {wall ms of the call, ms of it blocked waiting for the device
 * (data-dependent sizes), number of those waits}; wall - blocked = the host's own work (launches, planning). */
int ccn_exec_host_stats(const ccn_exec_t* exec, double* out_host);
int ccn_exec_profile_read(ccn_exec_t* exec, double* out_host, int32_t n_families);
int ccn_forward_run(ccn_exec_t* exec, const float* x, int64_t ldx, const float* pos, const int64_t* batch,
                    const int64_t* curve_idxs, int64_t n, void* arena, int64_t arena_bytes, ccn_draw_fn draw,
                    void* draw_user, float* logits, int64_t ld_logits, int64_t* arena_needed_host, ccn_stream_t stream);

/* ---- scanning simulator (csrc/sim_raster.cu, csrc/sim_scan.cu): mesh -> curve cloud, plain CUDA ----
 * sim_rasterize: scanning_simulator/utils/minimal_rasterizer.py:14-51 `rasterize` = vertices @ R^T + t, Pinhole2D.project_ndc
 *            (third_party/minimal_pytorch_rasterizer/minimal_pytorch_rasterizer/camera.py:38-61; fx = fy = res, cx = cy = res/2),
 *            estimate_normals_cuda (.../cuda/rasterizer_kernel.cu:80-225 z-buffer, :335-465 point + face normal of the
 *            closest face, back-facing windings culled by the inside test :194-198), mask = z_cam > 0, then the scene frame
 *            (p - t) @ R, n @ R.  vertices (nv, 3) f32 scene frame; faces (nf, 3) i32; R (3, 3) row-major, t (3): DEVICE
 *            pointers.  Out: coords / normals (res, res, 3) f32 [row i = image y, column j = image x], mask (res, res) u8,
 *            depth_or_null (res, res) f32 camera depth (1e10 where nothing was hit).  Scratch: cam / ndc (nv, 3) f32,
 *            keys res^2 u64, list (nf + 1) i32.  Where several faces lie within 1e-5 of the closest depth the reference
 *            keeps whichever face's thread wrote last; here the closest face wins, equal depths -> lowest face index. */
int ccn_sim_rasterize(const float* vertices, int64_t nv, const int32_t* faces, int64_t nf, const float* R, const float* t,
                      int32_t res, float* cam_scratch, float* ndc_scratch, uint64_t* key_scratch, int32_t* list_scratch,
                      float* coords, float* normals, uint8_t* mask, float* depth_or_null, ccn_stream_t stream);

/* sim_scan_lines: ONE attempt of ScanLineGenerator.generate_linear_scan_lines' while loop (scanning_simulator/utils/scanlines.py:24-76,
 *            sample_line_points :133-144).  border (B, L, 2, 2) f64 = the attempt's line end points [line][end][u, v] (drawn on
 *            the host like the reference: numpy RNG); sample k of a line = trunc((b1 - b0) * (k / ts + offset) + b0), in float32
 *            when f32_math (integer borders of direction 'random') else float64 ('parallel' / 'grid'); offset = the
 *            torch.rand(1) / ts draw.  A sample is kept when it is the first of its line on its pixel, masks[b][u][v] != 0 and
 *            the mesh holds fewer than N points; kept samples go to uv (B, N, 3) i64 = (u, v, line_base + line) from position
 *            npoints[b] on, in (line, sample) order; npoints (B) i32 running totals, updated (zero them before attempt 0).
 * sim_scan_finish: scanlines.py:78-88 (line ids -> 0 .. lines-with-points - 1 per mesh, in place) and the pixel look-up of
 *            SampleMeshes._sample_curves (scanning_simulator/utils/sampling.py:76-80): points / point_normals (B, N, 3) f32 =
 *            coords / normals (B, res, res, 3) at (u, v); pass null for both to skip the look-up.  Call when every mesh has N. */
int ccn_sim_scan_lines(const double* border, int32_t B, int32_t L, int32_t ts, float offset, int32_t f32_math, int32_t res,
                       const uint8_t* masks, int32_t N, int32_t line_base, int32_t* npoints, int64_t* uv, ccn_stream_t stream);
int ccn_sim_scan_finish(int64_t* uv, int32_t B, int32_t N, int32_t res, const float* coords, const float* normals,
                        float* points_or_null, float* point_normals_or_null, ccn_stream_t stream);
/* sim_curve_split: CurveClouds.compute_3d_curve_idxs (scanning_simulator/utils/curve_cloud.py:40-47): points (B, N, 3) f32;
 *            curve_idx (B, N) i32 = number of edges longer than thresh (float64 norm; CURVE_SPLIT_THRESH = 0.01) before the
 *            point; cstart / cend (B, N) i32: point range [cstart[c], cend[c]) of curve c (first ncurves[b] entries valid).
 * sim_anti_alias: CurveClouds._anti_alias (curve_cloud.py:168-190): vals (B, N, 3) f32 -> out (B, N, 3) f64, the 7-tap
 *            ANTI_ALIAS_KERNEL [1, 2, 3, 3, 3, 2, 2] along every curve divided by the same filter over the curve's extent.
 * sim_curvature: CurveClouds.compute_curvature_new -> estimate_curvature_and_grads(points, X = points, curve ids, k)
 *            (scanning_simulator/utils/curvature_approx.py:14-146; k nearest points of the own curve, Gaussian weights of width
 *            kernel_width x mean within-curve edge length in arc length, weighted quadratic fit with hinge_reg on the constant
 *            term): points (B, N, 3) f64 -> curvature (B, N, 3) f64 (curve normal x curvature), gradients (B, N, 3, 3) f64
 *            (or null).  k <= 16.  Scratch: geo_raw / geo (B, N) f64, mean_edge (B) f64.  Float64 results agree with the
 *            reference to summation order (its torch.linalg.solve / conv1d reduce in library order). */
int ccn_sim_curve_split(const float* points, int32_t B, int32_t N, double thresh, int32_t* curve_idx, int32_t* cstart,
                        int32_t* cend, int32_t* ncurves, ccn_stream_t stream);
int ccn_sim_anti_alias(const float* vals, const int32_t* curve_idx, int32_t B, int32_t N, double* out, ccn_stream_t stream);
int ccn_sim_curvature(const double* points, const int32_t* curve_idx, const int32_t* cstart, const int32_t* cend, int32_t B,
                      int32_t N, int32_t k, double kernel_width, double hinge_reg, double* geo_raw_scratch, double* geo_scratch,
                      double* mean_edge_scratch, double* curvature, double* gradients_or_null, ccn_stream_t stream);
/* CurveClouds.compute_intersections + line_seg_dists (scanning_simulator/utils/curve_cloud.py:114-166, 285-331) for ONE curve
 * cloud (upstream's second scatter_min only broadcasts for B = 1 - how generate_shapenet.py calls it): all pairs of polyline
 * edges in f64, closest approach of edge i to the line of edge j (mua clipped to [0,1], 0 for parallel pairs), pairs across a
 * curve boundary or on one curve masked to 100 * thresh, min over the rows then over the columns of a curve (lowest index among
 * equals), curve pairs closer than thresh.  Two calls around one host read-back of the count: _count leaves the exclusive
 * scan of the pair flags in slot_scan (last entry = number of intersections), _fill writes pairs_out (2, n_out) int64 point
 * indices, each edge moved to its far end point when its mua > 0.5.  points (n,3) f64; curve_idx (n) int32 0,1,2,.. along the
 * points; scratch sizes: cptr n_curves+1, min1 / arg1 n_curves*(n-1), arg2 / flag n_curves^2, slot_scan n_curves^2+1. */
int ccn_sim_intersections_count(const double* points, const int32_t* curve_idx, int32_t n, int32_t n_curves, double thresh,
                                int32_t* cptr_scratch, double* min1_scratch, int32_t* arg1_scratch, int32_t* arg2_scratch,
                                int32_t* flag_scratch, int32_t* slot_scan, ccn_stream_t stream);
int ccn_sim_intersections_fill(const double* points, int32_t n, int32_t n_curves, const int32_t* arg1_scratch,
                               const int32_t* arg2_scratch, const int32_t* flag_scratch, const int32_t* slot_scan, int64_t n_out,
                               int64_t* pairs_out, int32_t* err_out, ccn_stream_t stream);
/* Label transfer of the dataset generators (scanning_simulator/shapenet_seg/generate_shapenet.py:88-89:
 * argmin(cdist(points.double(), pos.double()), dim=1), y[closest]): idx_out[i] = nearest of the nr labelled points to sampled
 * point i in f64 (lowest index among equals), label_out[i] = ref_labels[idx_out[i]] (label_out / ref_labels may be null). */
int ccn_sim_nearest_label(const double* points, int64_t nq, const double* ref_points, int64_t nr, const int64_t* ref_labels,
                          int64_t* idx_out, int64_t* label_out, ccn_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* CCN_B200_H */
