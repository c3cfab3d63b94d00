// Native step executor of the per-scan forward (reference src/models/base.py:133-175 `ModelBase.forward / _apply_steps`
// and the step modules of src/models/modules/{fast_conv1d,pointnet2,point_conv,dgcnn,mlp,skip_connect}.py), inference,
// tensor-core precisions: ONE C-ABI call per forward.
//
// Why: the python executor (curvecloudnet_b200/model.py + modules.py) issues ~150 ctypes calls, ~270 torch.empty and ~15
// blocking read-backs per 32 k-point scan - 5.1 ms of host time against ~4 ms of kernels (profiles/r01i_host_profile.txt),
// so the B = 1 step was host-bound.  Here the same launches are issued from C++ out of a bump-allocated arena: the host
// cost per launch is the driver's (~2-4 us), the data-dependent sizes are read back through one pinned buffer.
//
// The executor issues EXACTLY the C-ABI calls of the python path, in the same order with the same arguments (it calls
// the library's own extern "C" entry points), so its logits are bit-identical to ModelBase's python executor -
// tests/test_gpu_executor.py asserts that equality; parity against the oracle follows from the python path's tests.
//
// The model is described by a PLAN (two flat arrays, int64 and double, built once per weight version by
// curvecloudnet_b200/executor.py from the prepared python modules: folded weights, bf16 planes, TMA descriptors).
// Random draws of the reference (CurveFPS phase, VoxelFPS noise, FPS start; fps_ops.py:30,56, point_ops.py:64) come
// through a callback in the reference's call order, so seeding torch identically reproduces the python path.
#include <cuda_runtime.h>

#include <chrono>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <deque>
#include <vector>

#include "common.cuh"

namespace {

constexpr int kMaxSamples = 4096;                 // ops.MAX_SAMPLES
constexpr int64_t kSmemBudget = 227 * 1024 - 2560; // ops.SMEM_BUDGET
constexpr int kBoxMinTiles = 74;                   // ops._BOX_MIN_TILES
constexpr size_t kH2dRing = 256 * 1024;

enum StepKind { ST_CONV_V2 = 1, ST_SA_GEO = 2, ST_MLP = 3, ST_SGCNN = 4, ST_SKIP = 5, ST_SA = 6, ST_FP = 7, ST_FP_GEO = 8 };
enum { DOWN_VOXEL = 0, DOWN_FPS = 1 };
enum { DRAW_CURVE_PHASE = 0, DRAW_VOXEL_NOISE = 1, DRAW_FPS_START = 2 };

#define EX_CALL(expr)                    \
    do {                                 \
        const int rc__ = (expr);         \
        if (rc__ != CCN_OK) return rc__; \
    } while (0)
#define EX_CUDA(expr)                              \
    do {                                           \
        const cudaError_t e__ = (expr);            \
        if (e__ != cudaSuccess) return (int)e__;   \
    } while (0)

struct Linear {                // one prepared Linear use-site (a K-segmentation of a folded weight)
    int n, k_true, kpad, act, nseg;
    int seg_w[4];
    float slope;
    const float* bias;
    const void *map_hi[4], *map_lo[4];   // host tensor maps for box_n = 32, 64, 128, 256 (null: not a candidate)
    int chain;                           // >= 0: this layer and the rest of its MLP run as that ccn_mlp_chain entry
};

struct Conv {                  // one curve-convolution layer
    int taps, cin, cout, tc;   // tc: runs as ccn_conv_planes (else the FP32-pipe tap kernel)
    const float *w, *bias;     // (taps, cin, cout) folded fp32, (cout)
    int lin;                   // tc: Linear entry (n = cout, kpad = taps * ceil64(cin)); maps as above
};

struct Chain { const ccn_chain_t* desc; int n_out; };

struct Step {
    int kind;
    std::vector<int64_t> iv;   // kind-specific integers
    std::vector<double> fv;    // kind-specific reals
};

struct Planes { void* buf = nullptr; int64_t kpad = 0, pstride = 0; };

struct Feat {                  // an fp32 feature matrix and (lazily) its activation planes
    float* f = nullptr;
    int64_t ld = 0, n = 0;
    int c = 0;
    Planes pl;
};

// `ready`: recorded on the geometry stream when the entry was produced there; the main stream waits for it once (then null)
struct Grid { float radius; int64_t cells; int32_t* cell_off; float* sorted; cudaEvent_t ready; };
struct KnnMemo { int K; float radius; int32_t* idx; cudaEvent_t ready; };

struct Level {                 // one point set of the pyramid
    int64_t n = 0;
    Feat pos;                  // (n, 3) positions as a feature matrix (its planes feed the `with_xyz` concatenations)
    int64_t *batch = nullptr, *curve = nullptr;
    bool has_sptr = false, has_topo = false;
    int32_t* sptr_dev = nullptr;
    std::vector<int> sptr;     // host copy of the sample CSR (B + 1)
    int n_curves = 0, max_glob = 0;
    int32_t *curve_ord = nullptr, *curve_ptr = nullptr;
    std::vector<Grid> grids;
    std::vector<KnnMemo> knn;
    // per-curve statistics of radius_1d_group (ccn_radius_1d_curve_stats), computed ahead on the geometry stream
    float* kc = nullptr; int32_t* kstats = nullptr; float kc_radius = 0.f; cudaEvent_t kc_ready = nullptr;
};

struct Exec {
    int mode = 3, npl = 2;     // 1 = bf16, 3 = split-bf16 x3
    int n_out = 0, in_c = 0;
    std::vector<Linear> lins;
    std::vector<Conv> convs;
    std::vector<Chain> chains;
    std::vector<Step> steps;
    std::vector<int> skip_store;       // per step: its input is a skip-connection source (base.py:191-196)
    std::vector<int> out_mlp;          // Linear ids of the final MLP
    // per-forward state
    char* arena = nullptr;
    size_t cap = 0, off = 0;
    bool oom = false;
    int32_t* pinned = nullptr;         // host read-back buffer (kMaxSamples + 16 ints)
    char* h2d_ring = nullptr;          // pinned staging of small host -> device copies (sample CSRs, FPS meta, the CurveFPS draw)
    size_t h2d_off = 0;
    float* noise_host = nullptr;       // pinned staging of the VoxelFPS noise draw
    size_t noise_cap = 0;
    cudaStream_t stream = nullptr;
    // farthest-point sampling of the NEXT levels runs ahead on a side stream (one CTA per cloud, K sequential rounds: 0.3 ms on
    // one SM at B = 1) while the main stream runs the edge steps of the current level; events order the two
    cudaStream_t side = nullptr;
    std::vector<int> fps_next;          // per step: the next `sa` step (FPS or voxel sampling) on the level this step creates (-1: none)
    struct VoxelJob { int64_t* idx = nullptr; int32_t* out = nullptr; int B = 0; int64_t n1 = 0; };
    struct Pre {
        bool done = false; bool voxel = false; Level* ch = nullptr; int64_t* idx = nullptr; cudaEvent_t ready = nullptr; VoxelJob job;
        // voxel pre-samples: the parent level, and whether the read-back was already collected ahead of the step (the level is
        // then sized and gathered on the side stream, `gathered` marks the gather)
        Level* parent = nullptr; bool collected = false; cudaEvent_t gathered = nullptr;
    };
    int32_t* pinned_side = nullptr;     // read-back buffer of the side stream
    std::vector<Pre> pre;
    cudaEvent_t lvl_ready = nullptr;
    bool use_side = true;
    // CCN_EXEC_LOOKAHEAD (read at ccn_exec_create; tests run both settings against the python executor): 0 = everything on the
    // caller's stream in step order (no side / geometry streams), 1 = look-ahead streams but no early collection of voxel
    // pre-samples, early first draw or curve statistics ahead, 2 (default) = all of it
    bool early = true;
    // neighbour searches depend on positions only: as soon as a level exists, the kNN tables of the sgcnn steps that will run
    // on it and the interpolation table of the fp / fp-geo step that will return to its parent are computed on a second side
    // stream (`geo`), concurrently with the feature kernels of the main stream
    cudaStream_t geo = nullptr;
    std::vector<int> fp_partner;        // per down step: the fp / fp-geo step that pops it (-1: none)
    struct UpMemo { bool valid = false; int32_t* nbr = nullptr; float* wgt = nullptr; cudaEvent_t ready = nullptr; };
    std::vector<UpMemo> up;             // per fp step: its interpolation table, if computed ahead
    std::vector<cudaEvent_t> geo_events;
    size_t geo_used = 0;
    cudaEvent_t geo_event() {
        if (geo_used == geo_events.size()) {
            cudaEvent_t e = nullptr;
            cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
            geo_events.push_back(e);
        }
        return geo_events[geo_used++];
    }
    bool on_main() const { return stream != side && stream != geo; }
    ccn_draw_fn draw = nullptr;
    void* draw_user = nullptr;
    bool early_u_valid = false;         // the CurveFPS phase, when it is the forward's first draw, is taken at entry
    float early_u = 0.f;
    int topo_flags = 0;
    // host-side accounting of the last forward: wall time of ccn_forward_run, time blocked in stream / event synchronisation
    double host_total_ms = 0, host_blocked_ms = 0;
    int host_syncs = 0;
    int sync_stream(cudaStream_t s) {
        const auto t0 = std::chrono::steady_clock::now();
        const cudaError_t e = cudaStreamSynchronize(s);
        host_blocked_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        ++host_syncs;
        return e == cudaSuccess ? CCN_OK : (int)e;
    }
    int sync_event(cudaEvent_t ev) {
        const auto t0 = std::chrono::steady_clock::now();
        const cudaError_t e = cudaEventSynchronize(ev);
        host_blocked_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        ++host_syncs;
        return e == cudaSuccess ? CCN_OK : (int)e;
    }
    // optional per-family device timing (bench.py's roofline leg): CUDA events around every launch of a family, a short spin
    // queued before each so that the pair brackets the kernel and not the host's launch gap
    bool prof = false;
    int prof_spin = 0;
    struct ProfRec { int fam; cudaEvent_t a, b; double flops, bytes; };
    std::vector<ProfRec> prof_recs;
    std::vector<cudaEvent_t> prof_pool;
    size_t prof_used = 0;
    cudaEvent_t prof_event() {
        if (prof_used == prof_pool.size()) {
            cudaEvent_t e;
            cudaEventCreate(&e);
            prof_pool.push_back(e);
        }
        return prof_pool[prof_used++];
    }
    std::deque<Feat> feats;
    std::deque<Level> levels;

    void* alloc(size_t bytes) {
        const size_t a = (off + 255) & ~(size_t)255;
        if (a + bytes > cap) { oom = true; off = a + bytes; return arena; }   // keep counting: the caller learns the need
        off = a + bytes;
        return arena + a;
    }
    template <typename T> T* alloc_n(int64_t n) { return (T*)alloc(sizeof(T) * (size_t)(n > 0 ? n : 1)); }
    Feat* new_feat(int64_t n, int c) {
        feats.emplace_back();
        Feat& f = feats.back();
        f.f = alloc_n<float>(n * c); f.ld = c > 0 ? c : 1; f.n = n; f.c = c;
        return &f;
    }
    Planes alloc_planes(int64_t m, int k) {
        Planes p;
        p.kpad = (k + 15) / 16 * 16;
        p.pstride = (m > 0 ? m : 1) * p.kpad;
        p.buf = alloc((size_t)npl * p.pstride * 2);
        return p;
    }
    int h2d(void* dst, const void* src, size_t bytes) {   // async copy through the pinned ring (the source may be reused at once)
        if (bytes > kH2dRing) return CCN_EINVAL;
        if (h2d_off + bytes > kH2dRing) h2d_off = 0;      // a forward stages far less than the ring between two of its read-backs
        std::memcpy(h2d_ring + h2d_off, src, bytes);
        const cudaError_t e = cudaMemcpyAsync(dst, h2d_ring + h2d_off, bytes, cudaMemcpyHostToDevice, stream);
        h2d_off += (bytes + 15) & ~(size_t)15;
        return e == cudaSuccess ? CCN_OK : (int)e;
    }
    // split read-back of <= 8 ints: `begin` queues the copy and marks it with an event, the caller queues whatever does not
    // depend on the values, `end` waits for the EVENT only - the stream keeps running the work queued in between
    cudaEvent_t rb_event = nullptr;
    int readback_begin(const void* dev, int n_ints) {
        if (oom) return CCN_ENOMEM;
        if (n_ints > 8) return CCN_EINVAL;
        if (!rb_event) EX_CUDA(cudaEventCreateWithFlags(&rb_event, cudaEventDisableTiming));
        EX_CUDA(cudaMemcpyAsync(pinned + kMaxSamples + 8, dev, 4 * (size_t)n_ints, cudaMemcpyDeviceToHost, stream));
        EX_CUDA(cudaEventRecord(rb_event, stream));
        return CCN_OK;
    }
    int readback_end(const int32_t** vals) {
        *vals = pinned + kMaxSamples + 8;
        return sync_event(rb_event);
    }
    int readback(const void* dev, int n_ints) {      // blocking read of n_ints int32 into `pinned`
        if (oom) return CCN_ENOMEM;
        EX_CUDA(cudaMemcpyAsync(pinned, dev, 4 * (size_t)n_ints, cudaMemcpyDeviceToHost, stream));
        return sync_stream(stream);
    }
};

enum { FAM_LINEAR = 0, FAM_EDGE_FUSED = 1, FAM_CONV = 2, FAM_SEARCH = 3, FAM_SAMPLE = 4, FAM_COUNT = 5 };

__global__ void spin_kernel(long long cycles) {
    const long long t0 = clock64();
    while (clock64() - t0 < cycles) {}
}

struct ProfScope {            // brackets the launches issued during its lifetime
    Exec& ex;
    bool on;
    ProfScope(Exec& e, int fam, double flops, double bytes) : ex(e), on(e.prof && !e.oom) {
        if (!on) return;
        if (ex.prof_spin > 0) spin_kernel<<<1, 1, 0, ex.stream>>>(ex.prof_spin);
        Exec::ProfRec r{fam, ex.prof_event(), ex.prof_event(), flops, bytes};
        cudaEventRecord(r.a, ex.stream);
        ex.prof_recs.push_back(r);
    }
    ~ProfScope() {
        if (on) cudaEventRecord(ex.prof_recs.back().b, ex.stream);
    }
};

// ---- helpers mirroring curvecloudnet_b200/ops.py / nn.py ------------------------------------------------------------
int ensure_planes(Exec& ex, Feat* t) {               // nn.ForwardCache.planes_for
    if (t->pl.buf) return CCN_OK;
    t->pl = ex.alloc_planes(t->n, t->c);
    if (ex.oom) return CCN_OK;
    return ccn_pack_planes(t->f, t->ld, t->c, nullptr, nullptr, 0, 0, nullptr, nullptr, 0, 0, nullptr, nullptr, 0, 0, nullptr,
                           t->n, t->pl.buf, t->pl.kpad, t->pl.pstride, ex.npl, (ccn_stream_t)ex.stream);
}

int pick_box(const Linear& L, int64_t m, int npl) {  // ops.TcWeight.pick_box_planes
    const int cand[4] = {256, 128, 64, 32};
    const int64_t tiles_m = (m + 127) / 128, nk = L.kpad / 64;
    if (tiles_m >= 148) {
        int last = 32;
        for (int b : cand) {
            if (!(b == 32 || b / 2 < L.n)) continue;
            last = b;
            if (nk * npl * b * 128 + 3 * npl * 16384 <= kSmemBudget) return b;
        }
        return last;
    }
    for (int b : cand) {
        if (!(b == 32 || b / 2 < L.n)) continue;
        if (tiles_m * ((L.n + b - 1) / b) >= kBoxMinTiles) return b;
    }
    return 32;
}

inline int box_slot(int b) { return b == 32 ? 0 : b == 64 ? 1 : b == 128 ? 2 : 3; }

struct Seg { const void* buf; int64_t kpad, pstride; int k; };

// ops.linear_planes: out_kind 0 = fp32, 1 = planes, 2 = both.  y (fp32) / yp (planes) are allocated here.
// Accounting: FLOPs 2 M N K; bytes = the weights + the input rows if they are a step input (io & 1) + the output
// rows if they are a step output (io & 2) at fp32 width - the activations BETWEEN the layers of an MLP are not compulsory.
int linear_planes(Exec& ex, const Linear& L, const Seg* segs, int nseg, int64_t M, int out_kind, Feat** y_out, Planes* yp_out,
                  int io = 3) {
    Feat* Y = nullptr;
    Planes Yp;
    if (out_kind != 1) Y = ex.new_feat(M, L.n);
    if (out_kind != 0) Yp = ex.alloc_planes(M, L.n);
    if (y_out) *y_out = Y;
    if (yp_out) *yp_out = Yp;
    if (Y && out_kind == 2) Y->pl = Yp;
    if (ex.oom) return CCN_OK;
    const int box = pick_box(L, M, ex.npl);
    const int bs = box_slot(box);
    if (!L.map_hi[bs]) return CCN_EINVAL;
    Seg s[4] = {};
    for (int i = 0; i < nseg; ++i) s[i] = segs[i];
    ProfScope ps(ex, FAM_LINEAR, 2.0 * M * L.n * L.k_true,
                 (double)L.n * L.k_true * 2 * ex.npl + ((io & 1) ? 4.0 * M * L.k_true : 0.0) + ((io & 2) ? 4.0 * M * L.n : 0.0));
    return ccn_linear_planes(s[0].buf, s[0].kpad, s[0].pstride, s[0].k, s[1].buf, s[1].kpad, s[1].pstride, s[1].k, s[2].buf,
                             s[2].kpad, s[2].pstride, s[2].k, s[3].buf, s[3].kpad, s[3].pstride, s[3].k, L.map_hi[bs],
                             L.map_lo[bs], L.bias, Y ? Y->f : nullptr, L.n, Yp.buf, Yp.kpad, Yp.pstride, M, L.n, box, ex.mode,
                             L.act, L.slope, (ccn_stream_t)ex.stream);
}

int segs_of(Exec& ex, Feat* const* parts, int np, Seg* segs) {   // nn.cat_features with a cache: one K-segment per block
    for (int i = 0; i < np; ++i) {
        EX_CALL(ensure_planes(ex, parts[i]));
        segs[i] = Seg{parts[i]->pl.buf, parts[i]->pl.kpad, parts[i]->pl.pstride, parts[i]->c};
    }
    return CCN_OK;
}

// ops.mlp_chain: layers lin_ids[0..nl) of an MLP as ONE launch.  y_ext: write the fp32 result there (the logits) instead of
// allocating it.  Accounting as linear_planes: the weights of every layer, the input rows, the output rows.
int mlp_chain(Exec& ex, const Chain& ch, const int64_t* lin_ids, int nl, const Seg* segs, int nseg, int64_t M, int out_kind,
              Feat** y_out, float* y_ext = nullptr, int64_t ld_ext = 0) {
    Feat* Y = nullptr;
    Planes Yp;
    if (out_kind != 1 && !y_ext) Y = ex.new_feat(M, ch.n_out);
    if (out_kind != 0) Yp = ex.alloc_planes(M, ch.n_out);
    if (y_out) *y_out = Y;
    if (Y && out_kind == 2) Y->pl = Yp;
    if (ex.oom) return CCN_OK;
    Seg s[4] = {};
    for (int i = 0; i < nseg; ++i) s[i] = segs[i];
    double flops = 0, bytes = 0;
    for (int i = 0; i < nl; ++i) {
        const Linear& L = ex.lins[lin_ids[i]];
        flops += 2.0 * M * L.n * L.k_true;
        bytes += (double)L.n * L.k_true * 2 * ex.npl;
        if (i == 0) bytes += 4.0 * M * L.k_true;
        if (i == nl - 1) bytes += 4.0 * M * L.n;
    }
    ProfScope ps(ex, FAM_LINEAR, flops, bytes);
    return ccn_mlp_chain(s[0].buf, s[0].kpad, s[0].pstride, s[0].k, s[1].buf, s[1].kpad, s[1].pstride, s[1].k, s[2].buf, s[2].kpad,
                         s[2].pstride, s[2].k, s[3].buf, s[3].kpad, s[3].pstride, s[3].k, M, ch.desc, y_ext ? y_ext : (Y ? Y->f : nullptr),
                         y_ext ? ld_ext : ch.n_out, Yp.buf, Yp.kpad, Yp.pstride, (ccn_stream_t)ex.stream);
}

// nn.mlp_to_features / MLP.forward_from over a concatenation: hidden layers write planes, the last writes `last_kind`
int run_mlp(Exec& ex, const int64_t* lin_ids, int nl, Feat* const* parts, int np, int last_kind, Feat** out) {
    Seg segs[4];
    EX_CALL(segs_of(ex, parts, np, segs));
    const int64_t M = parts[0]->n;
    int nseg = np;
    Planes cur;
    for (int i = 0; i < nl; ++i) {
        const Linear& L = ex.lins[lin_ids[i]];
        const bool last = i == nl - 1;
        if (L.chain >= 0) return mlp_chain(ex, ex.chains[L.chain], lin_ids + i, nl - i, segs, nseg, M, last_kind, out);
        Feat* y = nullptr;
        Planes yp;
        EX_CALL(linear_planes(ex, L, segs, nseg, M, last ? last_kind : 1, &y, &yp, (i == 0 ? 1 : 0) | (last ? 2 : 0)));
        if (last) { *out = y; break; }
        cur = yp;
        segs[0] = Seg{cur.buf, cur.kpad, cur.pstride, L.n};
        nseg = 1;
    }
    return CCN_OK;
}

int level_sptr(Exec& ex, Level* lv) {                 // nn.ForwardCache.sample_ptr -> ops.sample_ptr
    if (lv->has_sptr) return CCN_OK;
    int32_t* tail = ex.alloc_n<int32_t>(kMaxSamples + 1 + 4);
    if (ex.oom) return CCN_OK;
    EX_CALL(ccn_sample_ptr(lv->batch, lv->n, tail, tail + kMaxSamples + 1, kMaxSamples, (ccn_stream_t)ex.stream));
    EX_CALL(ex.readback(tail, kMaxSamples + 1 + 4));      // sample CSR and its header in ONE round trip (4 KB)
    const int32_t* hd = ex.pinned + kMaxSamples + 1;
    const int B = hd[0], flags = hd[2];
    if (flags) { ex.topo_flags = flags; return CCN_ETOPOLOGY; }
    lv->sptr.assign(ex.pinned, ex.pinned + B + 1);
    lv->sptr_dev = tail;
    lv->has_sptr = true;
    return CCN_OK;
}

int level_topo(Exec& ex, Level* lv) {                 // nn.ForwardCache.topology -> ops.curve_csr
    if (lv->has_topo) return CCN_OK;
    const int64_t n = lv->n;
    int32_t* ord = ex.alloc_n<int32_t>(n);
    int32_t* cptr = ex.alloc_n<int32_t>(n + 1);
    int32_t* tail = ex.alloc_n<int32_t>(kMaxSamples + 1 + 4);
    int64_t* soff = ex.alloc_n<int64_t>(kMaxSamples + 1);
    int64_t* glob = ex.alloc_n<int64_t>(n);
    if (ex.oom) return CCN_OK;
    EX_CALL(ccn_curve_csr(lv->batch, lv->curve, n, ord, cptr, tail, soff, glob, tail + kMaxSamples + 1, kMaxSamples,
                          (ccn_stream_t)ex.stream));
    EX_CALL(ex.readback(tail, kMaxSamples + 1 + 4));      // one round trip
    const int32_t* hd = ex.pinned + kMaxSamples + 1;
    const int B = hd[0], n_curves = hd[1], flags = hd[2], max_glob = hd[3];
    if (flags) { ex.topo_flags = flags; return CCN_ETOPOLOGY; }
    if (!lv->has_sptr) {
        lv->sptr.assign(ex.pinned, ex.pinned + B + 1);
        lv->sptr_dev = tail;
        lv->has_sptr = true;
    }
    lv->n_curves = n_curves; lv->max_glob = max_glob; lv->curve_ord = ord; lv->curve_ptr = cptr;
    lv->has_topo = true;
    return CCN_OK;
}

// ops._finish_subset: sample CSR of the selected level + its size with ONE read-back (extra: more ints behind `count_dev`)
int finish_subset(Exec& ex, Level* parent, int64_t* idx_buf, int64_t capacity, const int32_t* count_dev, Level* child) {
    const int B = (int)parent->sptr.size() - 1;
    int32_t* out = ex.alloc_n<int32_t>(B + 1);
    if (ex.oom) return CCN_OK;
    EX_CALL(ccn_subset_sample_ptr(parent->batch, idx_buf, count_dev, capacity, B, out, (ccn_stream_t)ex.stream));
    EX_CALL(ex.readback(out, B + 1));
    child->sptr.assign(ex.pinned, ex.pinned + B + 1);
    child->sptr_dev = out;
    child->has_sptr = true;
    child->n = child->sptr[B];
    return CCN_OK;
}

// modules._subset: (pos[idx], batch[idx], curve[idx]) + the planes of the new positions
int subset_level(Exec& ex, Level* parent, const int64_t* idx, Level* child) {
    const int64_t n = child->n;
    child->pos.f = ex.alloc_n<float>(n * 3); child->pos.ld = 3; child->pos.n = n; child->pos.c = 3;
    child->batch = ex.alloc_n<int64_t>(n);
    child->curve = parent->curve ? ex.alloc_n<int64_t>(n) : nullptr;
    if (n) child->pos.pl = ex.alloc_planes(n, 3);
    if (ex.oom || n == 0) return CCN_OK;
    return ccn_subset_gather(parent->pos.f, parent->pos.ld, parent->batch, parent->curve, idx, n, child->pos.f, child->batch,
                             child->curve, child->pos.pl.buf, child->pos.pl.pstride, ex.npl, (ccn_stream_t)ex.stream);
}

// ops.frnn_knn: grid of (pos2, radius) memoised per level; self-queries (pos1 is pos2) memoised per (K, radius)
int frnn(Exec& ex, Level* q, Level* src, int K, float radius, bool want_cnt, int32_t** idx_out, int32_t** cnt_out) {
    const bool self = q == src;
    if (self && !want_cnt)
        for (KnnMemo& m : src->knn)
            if (m.K == K && m.radius == radius) {
                if (m.ready && ex.on_main()) {             // produced on the geometry stream: join it once
                    EX_CUDA(cudaStreamWaitEvent(ex.stream, m.ready, 0));
                    m.ready = nullptr;
                }
                *idx_out = m.idx;
                return CCN_OK;
            }
    Grid* g = nullptr;
    for (Grid& gg : src->grids)
        if (gg.radius == radius) g = &gg;
    if (g && g->ready && ex.on_main()) {
        EX_CUDA(cudaStreamWaitEvent(ex.stream, g->ready, 0));
        g->ready = nullptr;
    }
    if (!g) {
        Grid ng;
        ng.radius = radius;
        ng.ready = nullptr;
        ng.cells = ccn_grid_num_cells(src->n);
        ng.cell_off = ex.alloc_n<int32_t>(ng.cells + 1);
        int32_t* slot = ex.alloc_n<int32_t>(src->n);
        ng.sorted = ex.alloc_n<float>(4 * (src->n > 0 ? src->n : 1));
        ProfScope psg(ex, FAM_SEARCH, 0, 28.0 * src->n + 4.0 * ng.cells);
        if (!ex.oom)
            EX_CALL(ccn_grid_build(src->pos.f, src->batch, src->n, radius, ng.cells, ng.cell_off, slot, ng.sorted,
                                   (ccn_stream_t)ex.stream));
        src->grids.push_back(ng);
        g = &src->grids.back();
    }
    int32_t* idx = ex.alloc_n<int32_t>(q->n * K);
    int32_t* cnt = want_cnt ? ex.alloc_n<int32_t>(q->n + 2) : nullptr;
    *idx_out = idx;
    if (cnt_out) *cnt_out = cnt;
    if (ex.oom) return CCN_OK;
    ProfScope ps(ex, FAM_SEARCH, 0, 12.0 * q->n + 16.0 * src->n + 4.0 * q->n * K);
    EX_CALL(ccn_frnn_knn(q->pos.f, q->batch, q->n, g->sorted, g->cell_off, g->cells, src->sptr_dev, radius, K, idx, nullptr, cnt,
                         (ccn_stream_t)ex.stream));
    if (self && !want_cnt) src->knn.push_back(KnnMemo{K, radius, idx, nullptr});
    return CCN_OK;
}

// ops.knn_exact: brute force, or the hashed grid from ~7e8 query x source pairs on (same result bit for bit; ops.KNN_GRID_MIN_PAIRS)
int knn_exact(Exec& ex, Level* fine, Level* coarse, int knn, int32_t* nbr, float* wgt) {
    if (knn <= 4 && (double)fine->n * (double)coarse->n >= 7e8) {
        const int64_t cells = ccn_grid_num_cells(coarse->n);
        int32_t* cell_off = ex.alloc_n<int32_t>(cells + 1);
        int32_t* slot = ex.alloc_n<int32_t>(coarse->n);
        float* sorted = ex.alloc_n<float>(4 * (coarse->n > 0 ? coarse->n : 1));
        int32_t* scratch = ex.alloc_n<int32_t>(16 + fine->n);
        if (ex.oom) return CCN_OK;
        return ccn_knn_grid(fine->pos.f, fine->batch, fine->n, coarse->pos.f, coarse->batch, coarse->n, coarse->sptr_dev, knn, 1, 0.f,
                            cells, cell_off, slot, sorted, scratch, nbr, wgt, nullptr, (ccn_stream_t)ex.stream);
    }
    if (ex.oom) return CCN_OK;
    return ccn_knn_exact(fine->pos.f, fine->batch, fine->n, coarse->pos.f, coarse->sptr_dev, knn, 1, nbr, wgt, nullptr,
                         (ccn_stream_t)ex.stream);
}

struct Edges { int32_t *row_ptr, *erow, *ecol; int64_t E, nq; int max_degree; };

// PointNetConv2.forward_edges (inference): decomposed first layer, fused chain kernel or the unfused launches
// The per-POINT half of the decomposed first layer, H = x_src W_x: it does not depend on the edge list, so the sa / sa-geo steps
// queue it BEFORE they wait for the data-dependent edge count - the device works through it while the host is blocked.
int pointconv_pre(Exec& ex, const int64_t* iv, Feat* const* xparts, int nxp, Feat** H) {
    *H = nullptr;
    if (iv[0] < 0) return CCN_OK;
    Seg segs[4];
    EX_CALL(segs_of(ex, xparts, nxp, segs));
    return linear_planes(ex, ex.lins[iv[0]], segs, nxp, xparts[0]->n, 0, H, nullptr);
}

int pointconv(Exec& ex, const int64_t* iv, const double* fv, Feat* H, Level* src, Level* dst, const Edges& e, Feat** out) {
    // iv: lin_x, wp_ptr, bias0_ptr, c1, act0, agg, chain (-1: none), n_local (layers after the first), local lin ids[n_local],
    //     n_att, attend lin ids[n_att];  fv: slope0, normalize_radius (0: none)
    const Linear* lin_x = iv[0] >= 0 ? &ex.lins[iv[0]] : nullptr;
    const float* wp = (const float*)iv[1];
    const float* bias0 = (const float*)iv[2];
    const int c1 = (int)iv[3], act0 = (int)iv[4], agg = (int)iv[5], chain = (int)iv[6], n_local = (int)iv[7];
    const int64_t* local_ids = iv + 8;
    const int n_att = (int)iv[8 + n_local];
    const int64_t* att_ids = iv + 9 + n_local;
    const float slope0 = (float)fv[0], nradius = (float)fv[1];
    if ((lin_x != nullptr) != (H != nullptr) && !ex.oom) return CCN_EINVAL;
    const bool att = agg == CCN_AGG_ATTEND || agg == CCN_AGG_WSUM;
    if (chain >= 0 && e.max_degree <= 64) {
        const Chain& ch = ex.chains[chain];
        Feat* o = ex.new_feat(e.nq, ch.n_out);
        if (e.nq) o->pl = ex.alloc_planes(e.nq, ch.n_out);
        const int64_t n_tiles = e.E / (129 - e.max_degree) + 1;
        int32_t* scratch = ex.alloc_n<int32_t>(4 * n_tiles);
        *out = o;
        if (ex.oom) return CCN_OK;
        double mac = 0, wbytes = 0;
        for (int l = 0; l < ch.desc->n_layers; ++l) { mac += (double)ch.desc->k[l] * ch.desc->n[l]; }
        wbytes = mac * 2 * ex.npl;
        // gather_mlp_aggregate: N_src C 4 + E 4 + N_q (12 + C_out 4) + weights
        ProfScope ps(ex, FAM_EDGE_FUSED, 2.0 * e.E * mac, 4.0 * src->n * c1 + 4.0 * e.E + e.nq * (12.0 + 4.0 * ch.n_out) + wbytes);
        return ccn_pointconv_fused(H ? H->f : nullptr, H ? H->ld : 0, wp, bias0, src->pos.f, dst->pos.f, nullptr, 0, e.row_ptr,
                                   e.erow, e.ecol, e.nq, e.E, e.max_degree, c1, nradius, act0, slope0, ch.desc, agg,
                                   att ? n_local - 1 : -1, scratch, o->f, o->ld, o->pl.buf, o->pl.kpad, o->pl.pstride,
                                   o->pl.buf ? ex.npl : 0, (ccn_stream_t)ex.stream);
    }
    // unfused: ccn_edge_layer1_planes -> Linear launches -> ccn_seg_aggregate_both
    if (n_local < 1 || (c1 & 15) || (H && (H->ld & 3))) return CCN_EUNSUPPORTED;
    Planes msg0 = ex.alloc_planes(e.E, c1);
    if (!ex.oom)
        EX_CALL(ccn_edge_layer1_planes(H ? H->f : nullptr, H ? H->ld : 0, wp, bias0, src->pos.f, dst->pos.f, e.erow, e.ecol, e.E, c1,
                                       nradius, act0, slope0, msg0.buf, msg0.kpad, msg0.pstride, ex.npl, (ccn_stream_t)ex.stream));
    Seg seg{msg0.buf, msg0.kpad, msg0.pstride, c1};
    Feat* msg = nullptr;
    Planes msgp;
    for (int i = 0; i < n_local; ++i) {
        const Linear& L = ex.lins[local_ids[i]];
        const bool last = i == n_local - 1;
        Feat* y = nullptr;
        Planes yp;
        EX_CALL(linear_planes(ex, L, &seg, 1, e.E, last ? (att ? 2 : 0) : 1, &y, &yp, 0));
        if (last) { msg = y; msgp = yp; }
        else seg = Seg{yp.buf, yp.kpad, yp.pstride, L.n};
    }
    Feat* attf = nullptr;
    if (att) {
        seg = Seg{msgp.buf, msgp.kpad, msgp.pstride, msg->c};
        for (int i = 0; i < n_att; ++i) {
            const Linear& L = ex.lins[att_ids[i]];
            const bool last = i == n_att - 1;
            Feat* y = nullptr;
            Planes yp;
            EX_CALL(linear_planes(ex, L, &seg, 1, e.E, last ? 0 : 1, &y, &yp, 0));
            if (last) attf = y;
            else seg = Seg{yp.buf, yp.kpad, yp.pstride, L.n};
        }
    }
    const int c = msg->c;
    Feat* o = ex.new_feat(e.nq, c);
    *out = o;
    const bool both = e.nq > 0 && (c % 16) == 0;       // ops._both_ok (fp32 rows from our allocator are 256-byte aligned, ld % 4 == 0)
    if (both) o->pl = ex.alloc_planes(e.nq, c);
    if (ex.oom) return CCN_OK;
    if (both)
        return ccn_seg_aggregate_both(msg->f, msg->ld, attf ? attf->f : nullptr, attf ? attf->ld : 0, e.row_ptr, e.nq, c, agg, o->f,
                                      o->ld, o->pl.buf, o->pl.kpad, o->pl.pstride, ex.npl, (ccn_stream_t)ex.stream);
    return ccn_seg_aggregate(msg->f, msg->ld, attf ? attf->f : nullptr, attf ? attf->ld : 0, e.row_ptr, e.nq, c, agg, o->f, o->ld,
                             (ccn_stream_t)ex.stream);
}

// ops.dense_to_csr: (nq, K) table + per-row counts -> CSR (one read-back for E)
// in two halves: `begin` queues the scan of the counts and the read-back of their total, `end` waits for it and expands the
// table - the caller queues the edge-independent part of the step in between
int dense_to_csr_begin(Exec& ex, int32_t* cnt, int64_t nq) {
    if (ex.oom) return CCN_OK;
    EX_CALL(ccn_excl_scan_i32(cnt, nq, cnt + nq + 1, (ccn_stream_t)ex.stream));
    return ex.readback_begin(cnt + nq + 1, 1);
}

int dense_to_csr_end(Exec& ex, int32_t* idx, int32_t* cnt, int64_t nq, int K, Edges* e) {
    if (ex.oom) { *e = Edges{cnt, nullptr, nullptr, 0, nq, K}; return CCN_OK; }
    const int32_t* vals = nullptr;
    EX_CALL(ex.readback_end(&vals));
    const int64_t E = vals[0];
    int32_t* erow = ex.alloc_n<int32_t>(E);
    int32_t* ecol = ex.alloc_n<int32_t>(E);
    *e = Edges{cnt, erow, ecol, E, nq, K};
    if (ex.oom) return CCN_OK;
    return ccn_dense_to_csr(idx, cnt, nq, K, erow, ecol, (ccn_stream_t)ex.stream);
}

// ops.voxel_fps + ops._finish_subset, split in two: `launch` queues the sampler, the sample CSR of the kept points and the
// read-back of {CSR, error flag} into `host` on ex.stream WITHOUT waiting; `collect` (after that stream work is complete)
// sizes the new level.  One read-back instead of the python path's two.
int voxel_launch(Exec& ex, Level* lv, float voxel, int32_t* host, Exec::VoxelJob* job) {
    const int64_t n1 = lv->n > 0 ? lv->n : 1;
    const int B = (int)lv->sptr.size() - 1;
    const int64_t slots = ccn_voxel_table_slots(lv->n);
    uint64_t* table = ex.alloc_n<uint64_t>(2 * slots + n1);
    int32_t* ints = ex.alloc_n<int32_t>(2 * n1 + 2);
    int64_t* idx = ex.alloc_n<int64_t>(n1);
    float* noise = ex.alloc_n<float>(n1);
    int32_t* out = ex.alloc_n<int32_t>(B + 2);
    if ((size_t)n1 > ex.noise_cap) {
        if (ex.noise_host) cudaFreeHost(ex.noise_host);
        ex.noise_cap = (size_t)n1 * 2;
        EX_CUDA(cudaHostAlloc((void**)&ex.noise_host, 4 * ex.noise_cap, cudaHostAllocDefault));
    }
    ex.draw(ex.draw_user, DRAW_VOXEL_NOISE, lv->n, ex.noise_host);
    job->idx = idx; job->out = out; job->B = B; job->n1 = n1;
    if (ex.oom) return CCN_OK;
    EX_CUDA(cudaMemcpyAsync(noise, ex.noise_host, 4 * (size_t)lv->n, cudaMemcpyHostToDevice, ex.stream));
    int32_t* counts = ints + 2 * n1;
    {
        ProfScope ps(ex, FAM_SAMPLE, 0, 20.0 * lv->n);
        EX_CALL(ccn_voxel_fps(lv->pos.f, lv->batch, lv->n, voxel, noise, slots, table, table + slots, table + 2 * slots, ints,
                              ints + n1, idx, counts, (ccn_stream_t)ex.stream));
    }
    EX_CALL(ccn_subset_sample_ptr(lv->batch, idx, counts, n1, B, out, (ccn_stream_t)ex.stream));
    EX_CUDA(cudaMemcpyAsync(host, out, 4 * (size_t)(B + 1), cudaMemcpyDeviceToHost, ex.stream));
    EX_CUDA(cudaMemcpyAsync(host + B + 1, counts + 1, 4, cudaMemcpyDeviceToHost, ex.stream));
    return CCN_OK;
}

int voxel_collect(const int32_t* host, const Exec::VoxelJob& job, Level* ch) {
    if (host[job.B + 1]) return CCN_EVOXEL_RANGE;
    ch->sptr.assign(host, host + job.B + 1);
    ch->sptr_dev = job.out;
    ch->has_sptr = true;
    ch->n = ch->sptr[job.B];
    return CCN_OK;
}

// modules.fps_sorted + ops.fps: K = ceil(len * ratio) in f32 (point_ops.py:63), random start per cloud, selected indices sorted;
// the level sizes follow from the parent's sample CSR on the host - no read-back.  Launches go to ex.stream.
int fps_downsample(Exec& ex, Level* lv, float ratio, Level* ch, int64_t** idx_out) {
    const int B = (int)lv->sptr.size() - 1;
    std::vector<int32_t> meta(2 * (B > 0 ? B : 1));
    int64_t total = 0;
    int max_len = 0;
    ch->sptr.assign(1, 0);
    for (int b = 0; b < B; ++b) {
        const int len = lv->sptr[b + 1] - lv->sptr[b];
        meta[b] = (int32_t)std::ceil((float)len * ratio);     // torch.ceil(int64 tensor * python float) -> f32 (point_ops.py:63)
        max_len = len > max_len ? len : max_len;
    }
    for (int b = 0; b < B; ++b) {
        const int len = lv->sptr[b + 1] - lv->sptr[b];
        meta[B + b] = (int32_t)ex.draw(ex.draw_user, DRAW_FPS_START, len, nullptr);
    }
    for (int b = 0; b < B; ++b) {
        const int len = lv->sptr[b + 1] - lv->sptr[b];
        total += meta[b] < len ? meta[b] : len;
        ch->sptr.push_back((int)total);
    }
    int32_t* mdev = ex.alloc_n<int32_t>(2 * B);
    float* closest = max_len > 512 * 16 ? ex.alloc_n<float>(lv->n) : nullptr;
    int32_t* sel = ex.alloc_n<int32_t>(lv->n);
    int64_t* idx = ex.alloc_n<int64_t>(total);
    int32_t* sdev = ex.alloc_n<int32_t>(B + 1);
    ch->n = total; ch->has_sptr = true; ch->sptr_dev = sdev;
    *idx_out = idx;
    if (ex.oom) return CCN_OK;
    EX_CALL(ex.h2d(mdev, meta.data(), 4 * (size_t)(2 * B)));
    {
        ProfScope ps(ex, FAM_SAMPLE, 0, 12.0 * lv->n + 4.0 * total);
        EX_CALL(ccn_fps(lv->pos.f, lv->sptr_dev, B, lv->n, mdev, mdev + B, max_len, closest, sel, idx, nullptr, (ccn_stream_t)ex.stream));
    }
    EX_CALL(ex.h2d(sdev, ch->sptr.data(), 4 * (size_t)(B + 1)));
    return CCN_OK;
}

// Right after step `si` has created level `lv` (its gathers are queued on the main stream): sample - and gather - the levels
// of the following FPS `sa` steps on the side stream.  Their sizes are host arithmetic, their draws are the next draws of
// the reference's sequence (no step in between draws), so nothing but the overlap changes.
// `ready`: an event recorded on the main stream when the level was complete (null: record one now).
int presample_ahead(Exec& ex, int si, Level* lv, cudaEvent_t ready = nullptr) {
    if (!ex.use_side || ex.prof || !ex.side || ex.fps_next[si] < 0 || ex.oom) return CCN_OK;
    if (ex.pre[ex.fps_next[si]].done) return CCN_OK;      // already sampled as part of an earlier step's look-ahead chain
    if (!ready) {
        EX_CUDA(cudaEventRecord(ex.lvl_ready, ex.stream));
        ready = ex.lvl_ready;
    }
    EX_CUDA(cudaStreamWaitEvent(ex.side, ready, 0));
    cudaStream_t main_stream = ex.stream;
    ex.stream = ex.side;
    int rc = CCN_OK;
    for (int k = ex.fps_next[si]; k >= 0 && rc == CCN_OK && !ex.oom; k = ex.fps_next[k]) {
        ex.levels.emplace_back();
        Level* ch = &ex.levels.back();
        int64_t* idx = nullptr;
        if (ex.steps[k].iv[0] == DOWN_VOXEL) {
            // voxel sampling: the size of the new level is data dependent - queue the sampler and its read-back, stop the chain
            Exec::Pre& pr = ex.pre[k];
            rc = voxel_launch(ex, lv, (float)ex.steps[k].fv[0], ex.pinned_side, &pr.job);
            if (rc == CCN_OK && !ex.oom) {
                if (!pr.ready && cudaEventCreateWithFlags(&pr.ready, cudaEventDisableTiming) != cudaSuccess) rc = CCN_EINVAL;
                if (rc == CCN_OK && cudaEventRecord(pr.ready, ex.side) != cudaSuccess) rc = CCN_EINVAL;
                pr.done = true; pr.voxel = true; pr.ch = ch; pr.parent = lv; pr.collected = false;
            }
            break;
        }
        rc = fps_downsample(ex, lv, (float)ex.steps[k].fv[1], ch, &idx);
        if (rc == CCN_OK) rc = subset_level(ex, lv, idx, ch);
        if (rc == CCN_OK && !ex.oom) {
            Exec::Pre& pr = ex.pre[k];
            if (!pr.ready && cudaEventCreateWithFlags(&pr.ready, cudaEventDisableTiming) != cudaSuccess) rc = CCN_EINVAL;
            if (rc == CCN_OK && cudaEventRecord(pr.ready, ex.side) != cudaSuccess) rc = CCN_EINVAL;
            pr.done = true; pr.voxel = false; pr.ch = ch; pr.idx = idx;
        }
        lv = ch;
    }
    ex.stream = main_stream;
    return rc;
}

// Polled at every step boundary: a voxel pre-sample whose read-back has landed is collected NOW instead of when the main stream
// reaches its step - the new level is sized, gathered on the side stream, and the samplers of the levels after it (the FPS
// chains: 0.5 us per selected point, 300 us for level 4) start that much earlier.  Same draws in the same order: no step
// between the two points draws.
int collect_voxel_early(Exec& ex) {
    if (!ex.use_side || !ex.early || ex.prof || !ex.side || ex.oom) return CCN_OK;
    for (size_t k = 0; k < ex.pre.size(); ++k) {
        Exec::Pre& pr = ex.pre[k];
        if (!pr.done || !pr.voxel || pr.collected) continue;
        const cudaError_t q = cudaEventQuery(pr.ready);
        if (q == cudaErrorNotReady) { (void)cudaGetLastError(); continue; }
        if (q != cudaSuccess) return (int)q;
        EX_CALL(voxel_collect(ex.pinned_side, pr.job, pr.ch));
        cudaStream_t main_stream = ex.stream;
        ex.stream = ex.side;
        int rc = subset_level(ex, pr.parent, pr.job.idx, pr.ch);
        if (rc == CCN_OK && !ex.oom) {
            if (!pr.gathered && cudaEventCreateWithFlags(&pr.gathered, cudaEventDisableTiming) != cudaSuccess) rc = CCN_EINVAL;
            if (rc == CCN_OK && cudaEventRecord(pr.gathered, ex.side) != cudaSuccess) rc = CCN_EINVAL;
            if (rc == CCN_OK) {
                pr.collected = true;
                rc = presample_ahead(ex, (int)k, pr.ch);
            }
        }
        ex.stream = main_stream;
        EX_CALL(rc);
    }
    return CCN_OK;
}

// Step `si` has created level `ch` from `lv` (sampled indices `idx`): queue, on the geometry stream, (1) the interpolation
// table of the fp / fp-geo step that will pop this down step and (2) the kNN tables of the sgcnn steps that run on `ch` before
// the next level change.  The main stream joins through events at first use.
int geometry_ahead(Exec& ex, int si, Level* lv, Level* ch, const int64_t* idx, cudaEvent_t ready = nullptr) {
    if (!ex.use_side || ex.prof || !ex.geo || ex.oom || ch->n == 0) return CCN_OK;
    cudaEvent_t start = ready;
    if (!start) {
        start = ex.geo_event();
        EX_CUDA(cudaEventRecord(start, ex.stream));        // the level's positions / batch are queued on the main (or sampling) stream
    }
    EX_CUDA(cudaStreamWaitEvent(ex.geo, start, 0));
    cudaStream_t saved = ex.stream;
    ex.stream = ex.geo;
    int rc = CCN_OK;
    bool any = false;
    const int j = ex.fp_partner[si];
    if (j >= 0 && ch->has_sptr) {
        const Step& fp = ex.steps[j];
        const int knn = (int)fp.iv[0];
        int32_t* nbr = ex.alloc_n<int32_t>(lv->n * knn);
        float* wgt = ex.alloc_n<float>(lv->n * knn);
        if (!ex.oom) {
            if (fp.kind == ST_FP)
                rc = knn_exact(ex, lv, ch, knn, nbr, wgt);
            else if (lv->has_topo && idx)
                rc = ccn_knn_1d_superset(lv->pos.f, lv->curve_ord, lv->n, idx, ch->n, knn, nbr, wgt, (ccn_stream_t)ex.geo);
            else
                nbr = nullptr;
            if (rc == CCN_OK && nbr) {
                Exec::UpMemo& u = ex.up[j];
                u.valid = true; u.nbr = nbr; u.wgt = wgt; u.ready = ex.geo_event();
                if (cudaEventRecord(u.ready, ex.geo) != cudaSuccess) rc = CCN_EINVAL;
                any = true;
            }
        }
    }
    for (size_t k = (size_t)si + 1; k < ex.steps.size() && rc == CCN_OK && !ex.oom; ++k) {
        const int kind = ex.steps[k].kind;
        if (kind == ST_SA || kind == ST_SA_GEO || kind == ST_FP || kind == ST_FP_GEO) break;
        if (kind != ST_SGCNN || !ch->has_sptr) continue;
        const int K = (int)ex.steps[k].iv[1];
        const float radius = (float)ex.steps[k].fv[0];
        bool have = false;
        for (const KnnMemo& m : ch->knn) have = have || (m.K == K && m.radius == radius);
        if (have) continue;
        int32_t* nbr = nullptr;
        rc = frnn(ex, ch, ch, K, radius, false, &nbr, nullptr);
        if (rc == CCN_OK && !ex.oom) {
            cudaEvent_t e = ex.geo_event();
            if (cudaEventRecord(e, ex.geo) != cudaSuccess) { rc = CCN_EINVAL; break; }
            ch->knn.back().ready = e;
            for (Grid& g : ch->grids)
                if (g.radius == radius && !g.ready) g.ready = e;      // built just now on the geometry stream
            any = true;
        }
    }
    (void)any;
    ex.stream = saved;
    return rc;
}

}  // namespace

struct ccn_exec {
    Exec ex;
};

extern "C" int ccn_exec_create(const int64_t* iplan, int64_t n_i, const double* fplan, int64_t n_f, ccn_exec_t** out) {
    CCN_REQUIRE(iplan && fplan && out && n_i >= 8);
    ccn_exec* h = new ccn_exec();
    Exec& ex = h->ex;
    int64_t ci = 0, cf = 0;
    auto I = [&]() -> int64_t { return ci < n_i ? iplan[ci++] : 0; };
    auto F = [&]() -> double { return cf < n_f ? fplan[cf++] : 0.0; };
    const int64_t magic = I();
    if (magic != 0xCC9E0003) { delete h; return CCN_EINVAL; }
    ex.mode = (int)I(); ex.npl = ex.mode == 3 ? 2 : 1;
    ex.in_c = (int)I(); ex.n_out = (int)I();
    const int n_lin = (int)I(), n_conv = (int)I(), n_chain = (int)I(), n_step = (int)I();
    for (int i = 0; i < n_lin; ++i) {
        Linear L;
        L.n = (int)I(); L.k_true = (int)I(); L.kpad = (int)I(); L.act = (int)I(); L.nseg = (int)I();
        for (int s = 0; s < 4; ++s) L.seg_w[s] = (int)I();
        L.bias = (const float*)I();
        for (int b = 0; b < 4; ++b) L.map_hi[b] = (const void*)I();
        for (int b = 0; b < 4; ++b) L.map_lo[b] = (const void*)I();
        L.chain = (int)I();
        L.slope = (float)F();
        ex.lins.push_back(L);
    }
    for (int i = 0; i < n_conv; ++i) {
        Conv c;
        c.taps = (int)I(); c.cin = (int)I(); c.cout = (int)I(); c.tc = (int)I();
        c.w = (const float*)I(); c.bias = (const float*)I(); c.lin = (int)I();
        ex.convs.push_back(c);
    }
    for (int i = 0; i < n_chain; ++i) {
        Chain c;
        c.desc = (const ccn_chain_t*)I(); c.n_out = (int)I();
        ex.chains.push_back(c);
    }
    for (int i = 0; i < n_step; ++i) {
        Step s;
        s.kind = (int)I();
        ex.skip_store.push_back((int)I());
        const int ni = (int)I(), nf = (int)I();
        for (int k = 0; k < ni; ++k) s.iv.push_back(I());
        for (int k = 0; k < nf; ++k) s.fv.push_back(F());
        ex.steps.push_back(s);
    }
    ex.fps_next.assign(ex.steps.size(), -1);
    ex.pre.resize(ex.steps.size());
    for (size_t i = 0; i < ex.steps.size(); ++i) {
        if (ex.steps[i].kind != ST_SA && ex.steps[i].kind != ST_SA_GEO) continue;
        for (size_t j = i + 1; j < ex.steps.size(); ++j) {
            const int k = ex.steps[j].kind;
            if (k == ST_SA) { ex.fps_next[i] = (int)j; break; }
            if (k == ST_SA || k == ST_SA_GEO || k == ST_FP || k == ST_FP_GEO) break;      // another level change first
        }
    }
    ex.fp_partner.assign(ex.steps.size(), -1);
    ex.up.resize(ex.steps.size());
    {
        std::vector<int> stack;
        for (size_t i = 0; i < ex.steps.size(); ++i) {
            const int k = ex.steps[i].kind;
            if (k == ST_SA || k == ST_SA_GEO) stack.push_back((int)i);
            if ((k == ST_FP || k == ST_FP_GEO) && !stack.empty()) { ex.fp_partner[stack.back()] = (int)i; stack.pop_back(); }
        }
    }
    const int n_om = (int)I();
    for (int i = 0; i < n_om; ++i) ex.out_mlp.push_back((int)I());
    if (ci != n_i || cf != n_f) { delete h; return CCN_EINVAL; }
    if (cudaHostAlloc((void**)&ex.pinned, sizeof(int32_t) * (kMaxSamples + 16), cudaHostAllocDefault) != cudaSuccess ||
        cudaHostAlloc((void**)&ex.h2d_ring, kH2dRing, cudaHostAllocDefault) != cudaSuccess ||
        cudaHostAlloc((void**)&ex.pinned_side, sizeof(int32_t) * (kMaxSamples + 16), cudaHostAllocDefault) != cudaSuccess) {
        delete h;
        return CCN_EINVAL;
    }
    {
        const char* env = std::getenv("CCN_EXEC_LOOKAHEAD");
        const int level = env ? std::atoi(env) : 2;
        ex.use_side = level >= 1;
        ex.early = level >= 2;
    }
    if (cudaStreamCreateWithFlags(&ex.side, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&ex.lvl_ready, cudaEventDisableTiming) != cudaSuccess) {
        ex.side = nullptr;                       // no side stream: everything stays on the caller's stream
    }
    if (cudaStreamCreateWithFlags(&ex.geo, cudaStreamNonBlocking) != cudaSuccess) ex.geo = nullptr;
    *out = h;
    return CCN_OK;
}

extern "C" void ccn_exec_destroy(ccn_exec_t* h) {
    if (!h) return;
    if (h->ex.rb_event) cudaEventDestroy(h->ex.rb_event);
    if (h->ex.pinned) cudaFreeHost(h->ex.pinned);
    if (h->ex.h2d_ring) cudaFreeHost(h->ex.h2d_ring);
    if (h->ex.pinned_side) cudaFreeHost(h->ex.pinned_side);
    if (h->ex.side) cudaStreamDestroy(h->ex.side);
    if (h->ex.geo) cudaStreamDestroy(h->ex.geo);
    for (cudaEvent_t e : h->ex.geo_events) if (e) cudaEventDestroy(e);
    if (h->ex.lvl_ready) cudaEventDestroy(h->ex.lvl_ready);
    for (Exec::Pre& pr : h->ex.pre) { if (pr.ready) cudaEventDestroy(pr.ready); if (pr.gathered) cudaEventDestroy(pr.gathered); }
    for (cudaEvent_t e : h->ex.prof_pool) cudaEventDestroy(e);
    if (h->ex.noise_host) cudaFreeHost(h->ex.noise_host);
    delete h;
}

extern "C" int32_t ccn_exec_topology_flags(const ccn_exec_t* h) { return h ? h->ex.topo_flags : 0; }

// out[3] = {wall ms of the last ccn_forward_run, ms of it blocked in stream / event synchronisation, number of those waits}
extern "C" int ccn_exec_host_stats(const ccn_exec_t* h, double* out) {
    CCN_REQUIRE(h && out);
    out[0] = h->ex.host_total_ms; out[1] = h->ex.host_blocked_ms; out[2] = h->ex.host_syncs;
    return CCN_OK;
}

extern "C" int ccn_exec_profile(ccn_exec_t* h, int32_t enable, int32_t spin_cycles) {
    CCN_REQUIRE(h);
    h->ex.prof = enable != 0;
    h->ex.prof_spin = spin_cycles;
    return CCN_OK;
}

// out[family][4] = {milliseconds, launches, algorithmic FLOPs, algorithmic bytes} of the LAST forward, summed per kernel family
// (0 Linear, 1 fused edge steps, 2 curve-conv GEMMs, 3 neighbour search, 4 samplers); synchronises the device.
extern "C" int ccn_exec_profile_read(ccn_exec_t* h, double* out, int32_t n_families) {
    CCN_REQUIRE(h && out && n_families >= FAM_COUNT);
    for (int i = 0; i < 4 * n_families; ++i) out[i] = 0.0;
    for (const Exec::ProfRec& r : h->ex.prof_recs) {
        if (cudaEventSynchronize(r.b) != cudaSuccess) return CCN_EINVAL;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, r.a, r.b) != cudaSuccess) return CCN_EINVAL;
        out[4 * r.fam] += ms; out[4 * r.fam + 1] += 1; out[4 * r.fam + 2] += r.flops; out[4 * r.fam + 3] += r.bytes;
    }
    return CCN_OK;
}

namespace {

int run_forward(Exec& ex, const float* x, int64_t ldx, const float* pos, const int64_t* batch, const int64_t* curve, int64_t n,
                float* logits, int64_t ld_logits) {
    ex.feats.clear();
    ex.levels.clear();
    ex.topo_flags = 0;
    ex.prof_recs.clear();
    ex.prof_used = 0;
    for (Exec::Pre& pr : ex.pre) { pr.done = false; pr.ch = nullptr; pr.idx = nullptr; pr.parent = nullptr; pr.collected = false; }
    for (Exec::UpMemo& u : ex.up) u.valid = false;
    ex.geo_used = 0;
    ccn_stream_t st = (ccn_stream_t)ex.stream;
    // the first draw of the forward: when it is the CurveFPS phase (sa-geo before any other sampling step) it is taken NOW - the
    // callback into python (~20 us) otherwise sits between the curve conv and the sampler with the device idle
    ex.early_u_valid = false;
    for (const Step& s0 : ex.steps) {
        if (s0.kind == ST_SA) break;
        if (s0.kind == ST_SA_GEO) {
            if (s0.iv[1] && ex.use_side && ex.early && !ex.prof) {
                ex.draw(ex.draw_user, DRAW_CURVE_PHASE, 1, &ex.early_u);
                ex.early_u_valid = true;
            }
            break;
        }
    }

    ex.levels.emplace_back();
    Level* lv0 = &ex.levels.back();
    lv0->n = n;
    lv0->pos.f = const_cast<float*>(pos); lv0->pos.ld = 3; lv0->pos.n = n; lv0->pos.c = 3;
    lv0->batch = const_cast<int64_t*>(batch);
    lv0->curve = const_cast<int64_t*>(curve);
    Feat* x0 = nullptr;
    if (x) {
        ex.feats.emplace_back();
        x0 = &ex.feats.back();
        x0->f = const_cast<float*>(x); x0->ld = ldx; x0->n = n; x0->c = ex.in_c;
    }
    // hist[i] = INPUT of step i (base.py:138: the lists start with the raw input)
    struct State { Feat* x; Level* lv; };
    std::vector<State> hist;
    std::vector<int64_t*> down_idx;
    std::vector<int> prop_stack, down_stack;
    State cur{x0, lv0};
    hist.push_back(cur);

    for (size_t si = 0; si < ex.steps.size(); ++si) {
        const Step& s = ex.steps[si];
        const int64_t* iv = s.iv.data();
        const double* fv = s.fv.data();
        State outp = cur;
        int64_t* didx = nullptr;
        EX_CALL(collect_voxel_early(ex));
        if (ex.oom) return CCN_ENOMEM;
        switch (s.kind) {
            case ST_CONV_V2: {
                // iv: with_xyz, with_diff, ksize, n_layers, tc_from, conv ids[n_layers]
                const int with_xyz = (int)iv[0], with_diff = (int)iv[1], ksize = (int)iv[2], nl = (int)iv[3], tc_from = (int)iv[4];
                Level* lv = cur.lv;
                EX_CALL(level_topo(ex, lv));
                if (si + 1 < ex.steps.size() && ex.steps[si + 1].kind == ST_SA_GEO && ex.use_side && ex.early && !ex.prof && ex.geo && !ex.oom &&
                    lv->n_curves > 0 && !lv->kc) {
                    // the next step groups along the curves of THIS level: its per-curve statistics (a strictly ordered f32 sum per
                    // curve, 28 us at 32 k points) run on the geometry stream behind the topology, beside this step's kernels
                    lv->kc = ex.alloc_n<float>(lv->n_curves);
                    lv->kstats = ex.alloc_n<int32_t>(4) + 1;          // [E | max k_c, max #points]: one read-back takes all three
                    lv->kc_radius = (float)ex.steps[si + 1].fv[2];
                    if (!ex.oom) {
                        cudaEvent_t topo_done = ex.geo_event();
                        EX_CUDA(cudaEventRecord(topo_done, ex.stream));
                        EX_CUDA(cudaStreamWaitEvent(ex.geo, topo_done, 0));
                        EX_CALL(ccn_radius_1d_curve_stats(lv->pos.f, lv->curve_ord, lv->curve_ptr, lv->n_curves, lv->kc_radius, lv->kc,
                                                          lv->kstats, (ccn_stream_t)ex.geo));
                        lv->kc_ready = ex.geo_event();
                        EX_CUDA(cudaEventRecord(lv->kc_ready, ex.geo));
                    } else {
                        lv->kc = nullptr;
                    }
                }
                const int64_t np = lv->n;
                const int cx = cur.x ? cur.x->c : 0;
                const int c_in = cx + (with_xyz ? 3 : 0);
                const int pad = ksize > 1 ? (ksize / 2) * nl : 0;
                const int64_t rows = np + (int64_t)(lv->n_curves + 1) * pad;
                const int tcf = rows > 0 ? tc_from : nl;
                // every layer on the tensor cores: the signal planes are formed in ONE pass (ccn_curve_feat_planes) - the fp32
                // feature matrix, its gathered copy and the pack kernel's read of it (78 us at the decoder's 32 k x 262) are skipped
                const bool direct = tcf == 0 && nl > 0;
                Feat* feats = direct ? nullptr : ex.new_feat(np, c_in * (with_diff ? 2 : 1));
                if (!direct && !ex.oom)
                    EX_CALL(ccn_curve_feat_prep(cur.x ? cur.x->f : nullptr, cur.x ? cur.x->ld : 0, cx, lv->pos.f, with_xyz, with_diff,
                                                lv->curve_ord, np, feats->f, feats->ld, st));
                int32_t* row2pt = ex.alloc_n<int32_t>(rows);
                if (!ex.oom) EX_CALL(ccn_curve_row_map(lv->curve_ord, np, lv->n_curves, pad, row2pt, st));
                Feat* sig = feats;
                Planes sigp;
                bool have_p = false;
                if (direct) {
                    sigp = ex.alloc_planes(rows, c_in * (with_diff ? 2 : 1));
                    have_p = true;
                    if (!ex.oom)
                        EX_CALL(ccn_curve_feat_planes(cur.x ? cur.x->f : nullptr, cur.x ? cur.x->ld : 0, cx, lv->pos.f, with_xyz, with_diff,
                                                      lv->curve_ord, np, row2pt, rows, sigp.buf, sigp.kpad, sigp.pstride, ex.npl, st));
                }
                for (int li = 0; li < nl; ++li) {
                    const Conv& cv = ex.convs[iv[5 + li]];
                    const bool first = li == 0, last = li == nl - 1;
                    if (li >= tcf) {
                        if (first && !direct) {
                            Feat* g = ex.new_feat(rows, feats->c);
                            if (!ex.oom && rows)
                                EX_CALL(ccn_gather_rows_i32(feats->f, feats->ld, row2pt, rows, feats->c, g->f, g->ld, st));
                            EX_CALL(ensure_planes(ex, g));
                            sigp = g->pl; have_p = true;
                        } else if (!have_p) {
                            EX_CALL(ensure_planes(ex, sig));
                            sigp = sig->pl; have_p = true;
                        }
                        const Linear& L = ex.lins[cv.lin];
                        const int box = pick_box(L, rows, ex.npl);
                        const int bs = box_slot(box);
                        if (last) {
                            Feat* y = ex.new_feat(np, cv.cout);
                            y->pl = ex.alloc_planes(np, cv.cout);
                            ProfScope ps(ex, FAM_CONV, 2.0 * rows * cv.taps * cv.cin * cv.cout, 4.0 * np * cv.cout + 2.0 * ex.npl * cv.taps * cv.cin * cv.cout);
                            if (!ex.oom)
                                EX_CALL(ccn_conv_planes(sigp.buf, sigp.kpad, sigp.pstride, cv.cin, cv.taps, L.map_hi[bs], L.map_lo[bs],
                                                        cv.bias, y->f, cv.cout, y->pl.buf, y->pl.kpad, y->pl.pstride, rows, cv.cout,
                                                        box, ex.mode, CCN_ACT_LEAKY, 0.01f, row2pt, st));
                            sig = y;
                        } else {
                            Planes yp = ex.alloc_planes(rows, cv.cout);
                            ProfScope ps(ex, FAM_CONV, 2.0 * rows * cv.taps * cv.cin * cv.cout, 2.0 * ex.npl * cv.taps * cv.cin * cv.cout);
                            if (!ex.oom)
                                EX_CALL(ccn_conv_planes(sigp.buf, sigp.kpad, sigp.pstride, cv.cin, cv.taps, L.map_hi[bs], L.map_lo[bs],
                                                        cv.bias, nullptr, cv.cout, yp.buf, yp.kpad, yp.pstride, rows, cv.cout, box,
                                                        ex.mode, CCN_ACT_LEAKY, 0.01f, nullptr, st));
                            sigp = yp;
                        }
                        continue;
                    }
                    Feat* y = ex.new_feat(last ? np : rows, cv.cout);
                    if (!ex.oom)
                        EX_CALL(ccn_curve_conv_layer(sig->f, sig->ld, cv.w, cv.bias, y->f, cv.cout, row2pt, rows, cv.cin, cv.cout, ksize,
                                                     first, last, 0.01f, st));
                    sig = y;
                }
                outp.x = sig;
                break;
            }
            case ST_MLP:
            case ST_SKIP: {
                // MLP: iv: with_xyz, n_layers, lin ids;  SKIP: iv: num_skips, n_layers, lin ids
                const int nl = (int)iv[1];
                Feat* parts[4];
                int np = 0;
                if (s.kind == ST_MLP) {
                    if (cur.x) parts[np++] = cur.x;
                    if (iv[0] || !cur.x) parts[np++] = &cur.lv->pos;
                } else {
                    const int ns = (int)iv[0];
                    if ((int)prop_stack.size() < ns || ns + 1 > 4) return CCN_EINVAL;
                    parts[np++] = cur.x;
                    for (int k = 0; k < ns; ++k) {
                        const State& h = hist[prop_stack[prop_stack.size() - ns + k]];
                        parts[np++] = h.x ? h.x : &h.lv->pos;
                    }
                    prop_stack.resize(prop_stack.size() - ns);
                }
                Feat* y = nullptr;
                EX_CALL(run_mlp(ex, iv + 2, nl, parts, np, 2, &y));
                outp.x = y;
                break;
            }
            case ST_SGCNN: {
                // iv: with_xyz, K, lin_pq, act0, chain (-1), n_rest, rest lin ids;  fv: radius, slope0
                const int with_xyz = (int)iv[0], K = (int)iv[1], act0 = (int)iv[3], chain = (int)iv[4], n_rest = (int)iv[5];
                const float radius = (float)fv[0], slope0 = (float)fv[1];
                Level* lv = cur.lv;
                EX_CALL(level_sptr(ex, lv));
                Feat* parts[2];
                int np = 0;
                if (cur.x) parts[np++] = cur.x;
                if (with_xyz || !cur.x) parts[np++] = &lv->pos;
                Seg segs[4];
                EX_CALL(segs_of(ex, parts, np, segs));
                Feat* PQ = nullptr;
                EX_CALL(linear_planes(ex, ex.lins[iv[2]], segs, np, lv->n, 0, &PQ, nullptr));
                int32_t* nbr = nullptr;
                EX_CALL(frnn(ex, lv, lv, K, radius, false, &nbr, nullptr));
                const int c1 = PQ->c / 2;
                if (chain >= 0) {
                    const Chain& ch = ex.chains[chain];
                    Feat* o = ex.new_feat(lv->n, ch.n_out);
                    if (lv->n) o->pl = ex.alloc_planes(lv->n, ch.n_out);
                    if (!ex.oom) {
                        double mac = 0;
                        for (int l = 0; l < ch.desc->n_layers; ++l) mac += (double)ch.desc->k[l] * ch.desc->n[l];
                        // valid rows = self loops + found neighbours; the padded n (K+1) rows bound them (the count stays on the device)
                        const double rows = (double)lv->n * (K + 1);
                        ProfScope ps(ex, FAM_EDGE_FUSED, 2.0 * rows * mac,
                                     8.0 * lv->n * c1 + 4.0 * lv->n * K + lv->n * (12.0 + 4.0 * ch.n_out) + mac * 2 * ex.npl);
                        EX_CALL(ccn_sgcnn_fused(PQ->f, PQ->ld, nbr, lv->n, K, c1, act0, slope0, ch.desc, o->f, o->ld, o->pl.buf, o->pl.kpad,
                                                o->pl.pstride, o->pl.buf ? ex.npl : 0, st));
                    }
                    outp.x = o;
                    break;
                }
                if (n_rest < 1 || (c1 & 15)) return CCN_EUNSUPPORTED;
                const int64_t rows = lv->n * (K + 1);
                Planes h = ex.alloc_planes(rows, c1);
                if (!ex.oom)
                    EX_CALL(ccn_sgcnn_layer1_planes(PQ->f, PQ->ld, nbr, lv->n, K, c1, act0, slope0, h.buf, h.kpad, h.pstride, ex.npl, st));
                Seg seg{h.buf, h.kpad, h.pstride, c1};
                Feat* y = nullptr;
                for (int i = 0; i < n_rest; ++i) {
                    const Linear& L = ex.lins[iv[6 + i]];
                    const bool last = i == n_rest - 1;
                    Planes yp;
                    EX_CALL(linear_planes(ex, L, &seg, 1, rows, last ? 0 : 1, &y, &yp, 0));
                    if (!last) seg = Seg{yp.buf, yp.kpad, yp.pstride, L.n};
                }
                Feat* o = ex.new_feat(lv->n, y->c);
                const bool both = lv->n > 0 && (y->c % 16) == 0;
                if (both) o->pl = ex.alloc_planes(lv->n, y->c);
                if (!ex.oom) {
                    if (both)
                        EX_CALL(ccn_group_max_dense_both(y->f, y->ld, nbr, lv->n, K, y->c, o->f, o->ld, o->pl.buf, o->pl.kpad,
                                                         o->pl.pstride, ex.npl, st));
                    else
                        EX_CALL(ccn_group_max_dense(y->f, y->ld, nbr, lv->n, K, y->c, o->f, o->ld, st));
                }
                outp.x = o;
                break;
            }
            case ST_SA_GEO:
            case ST_SA: {
                // SA_GEO: iv: with_xyz, use_curve_fps, [pointconv block];  fv: arclen, ratio, radius, [slope0, nradius]
                // SA:     iv: down, K, [pointconv block];                  fv: voxel, ratio, radius, [slope0, nradius]
                Level* lv = cur.lv;
                const bool presampled = ex.pre[si].done;
                if (!presampled) ex.levels.emplace_back();
                Level* ch = presampled ? ex.pre[si].ch : &ex.levels.back();
                int64_t* idx = presampled ? ex.pre[si].idx : nullptr;
                Edges e{};
                const float radius = (float)fv[2];
                if (s.kind == ST_SA_GEO) {
                    EX_CALL(level_topo(ex, lv));
                    if (lv->max_glob != lv->n_curves) return CCN_EUNSUPPORTED;   // ops.radius_1d_group: curve ids with gaps
                }
                EX_CALL(level_sptr(ex, lv));
                if (ex.oom) return CCN_ENOMEM;
                const int B = (int)lv->sptr.size() - 1;
                const bool curve_fps = s.kind == ST_SA_GEO && iv[1];
                const int down = s.kind == ST_SA ? (int)iv[0] : DOWN_FPS;
                const bool early = presampled && ex.pre[si].voxel && ex.pre[si].collected;
                if (early) {
                    // collected at an earlier step boundary (collect_voxel_early): the level exists, gathered on the side stream
                    EX_CUDA(cudaStreamWaitEvent(ex.stream, ex.pre[si].gathered, 0));
                    idx = ex.pre[si].job.idx;
                } else if (presampled && ex.pre[si].voxel) {
                    // the voxel sampler ran ahead on the side stream: wait for its read-back, size the level
                    EX_CALL(ex.sync_event(ex.pre[si].ready));
                    EX_CALL(voxel_collect(ex.pinned_side, ex.pre[si].job, ch));
                    ex.pre[si].collected = true;                      // (not again at a later step boundary)
                    idx = ex.pre[si].job.idx;
                } else if (presampled) {
                    // sampled (and gathered) ahead on the side stream: the main stream joins it here
                    EX_CUDA(cudaStreamWaitEvent(ex.stream, ex.pre[si].ready, 0));
                } else if (curve_fps) {
                    // ops.curve_fps
                    float u = ex.early_u;
                    if (!ex.early_u_valid) ex.draw(ex.draw_user, DRAW_CURVE_PHASE, 1, &u);
                    ex.early_u_valid = false;
                    float* run = ex.alloc_n<float>((lv->n > 0 ? lv->n : 1) + (lv->n + 4095) / 4096);
                    idx = ex.alloc_n<int64_t>(lv->n);
                    int32_t* cnt = ex.alloc_n<int32_t>(1);
                    float* udev = ex.alloc_n<float>(1);
                    if (!ex.oom) {
                        EX_CALL(ex.h2d(udev, &u, 4));
                        ProfScope ps(ex, FAM_SAMPLE, 0, 20.0 * lv->n);
                        EX_CALL(ccn_curve_fps(lv->pos.f, lv->curve_ord, lv->curve_ptr, lv->n, udev, (float)fv[0], run, idx, cnt, st));
                        EX_CALL(finish_subset(ex, lv, idx, lv->n > 0 ? lv->n : 1, cnt, ch));
                    }
                } else if (down == DOWN_VOXEL) {
                    Exec::VoxelJob job;
                    EX_CALL(voxel_launch(ex, lv, (float)fv[0], ex.pinned, &job));
                    if (ex.oom) return CCN_ENOMEM;
                    EX_CALL(ex.sync_stream(ex.stream));
                    EX_CALL(voxel_collect(ex.pinned, job, ch));
                    idx = job.idx;
                } else {
                    EX_CALL(fps_downsample(ex, lv, (float)fv[1], ch, &idx));
                }
                if (ex.oom) return CCN_ENOMEM;
                const int64_t* pc_iv = iv + 2;
                const double* pc_fv = fv + 3;
                Feat* xparts[2];
                int nxp = 0;
                if (cur.x) xparts[nxp++] = cur.x;
                if (s.kind == ST_SA_GEO && (iv[0] || !cur.x)) xparts[nxp++] = &lv->pos;
                Feat* H = nullptr;
                if (s.kind == ST_SA_GEO) {
                    // ops.radius_1d_group
                    const int64_t nq = ch->n;
                    const bool stats_ahead = lv->kc && lv->kc_radius == radius;
                    float* kc = stats_ahead ? lv->kc : ex.alloc_n<float>(lv->n_curves);
                    int32_t* buf = ex.alloc_n<int32_t>(nq + 4);
                    int32_t* kstats = stats_ahead ? lv->kstats : buf + nq + 2;
                    int64_t E = 0;
                    int max_deg = 1;
                    if (!ex.oom) {
                        if (stats_ahead) {
                            if (lv->kc_ready) { EX_CUDA(cudaStreamWaitEvent(ex.stream, lv->kc_ready, 0)); lv->kc_ready = nullptr; }
                            EX_CALL(ccn_radius_1d_group_count_ready(lv->curve_ord, lv->curve, lv->n, lv->n_curves, idx, nq, kc, kstats, buf,
                                                                    kstats - 1, st));
                        } else
                        EX_CALL(ccn_radius_1d_group_count(lv->pos.f, lv->curve_ord, lv->curve_ptr, lv->curve, lv->n, lv->n_curves, idx, nq,
                                                          radius, kc, buf + nq + 2, buf, buf + nq + 1, st));
                        EX_CALL(ex.readback_begin(stats_ahead ? kstats - 1 : buf + nq + 1, 3));
                        EX_CALL(pointconv_pre(ex, pc_iv, xparts, nxp, &H));    // queued behind the copy: runs while the host waits
                        if (ex.oom) return CCN_ENOMEM;
                        const int32_t* vals = nullptr;
                        EX_CALL(ex.readback_end(&vals));
                        E = vals[0];
                        float kf, nf;
                        std::memcpy(&kf, &vals[1], 4);
                        std::memcpy(&nf, &vals[2], 4);
                        const float m = kf < nf ? kf : nf;
                        max_deg = m >= 1.f ? (int)m : 1;
                    }
                    int32_t* col = ex.alloc_n<int32_t>(E);
                    int32_t* erow = ex.alloc_n<int32_t>(E);
                    if (!ex.oom) {
                        EX_CALL(ccn_radius_1d_group_fill(lv->curve_ord, lv->curve, lv->n, lv->n_curves, idx, nq, kc, kstats, buf, col, st));
                        EX_CALL(ccn_csr_rows(buf, nq, erow, st));
                    }
                    e = Edges{buf, erow, col, E, nq, max_deg};
                    if (!presampled) EX_CALL(subset_level(ex, lv, idx, ch));
                } else {
                    if (!presampled || (ex.pre[si].voxel && !early)) EX_CALL(subset_level(ex, lv, idx, ch));
                }
                // The look-ahead work (next levels' samplers on the side stream, neighbour tables on the geometry stream) is
                // HOST work too - the draw callbacks, a dozen launches, for a voxel sampler 70 us of noise generation - and used to
                // run before this step's own launches, with the main stream idle meanwhile (160 us after sa-geo, 50 us after
                // every sa: tools/timeline.py).  The level is complete here: mark it, queue this step's kernels, then look ahead.
                cudaEvent_t lvl_done = nullptr;
                if (ex.use_side && !ex.prof && !ex.oom) {
                    lvl_done = ex.geo_event();
                    EX_CUDA(cudaEventRecord(lvl_done, ex.stream));
                }
                Feat* o = nullptr;
                if (s.kind == ST_SA_GEO) {
                    if (ex.oom) return CCN_ENOMEM;
                    EX_CALL(pointconv(ex, pc_iv, pc_fv, H, lv, ch, e, &o));               // no host wait inside: all queued
                    EX_CALL(presample_ahead(ex, (int)si, ch, lvl_done));
                    EX_CALL(geometry_ahead(ex, (int)si, lv, ch, idx, lvl_done));
                } else {
                    const int K = (int)iv[1];
                    int32_t *nbr = nullptr, *cnt = nullptr;
                    EX_CALL(frnn(ex, ch, lv, K, radius, true, &nbr, &cnt));
                    EX_CALL(dense_to_csr_begin(ex, cnt, ch->n));                          // scan + read-back of the edge count, queued
                    EX_CALL(pointconv_pre(ex, pc_iv, xparts, nxp, &H));                   // independent of the edges: behind the copy
                    EX_CALL(presample_ahead(ex, (int)si, ch, lvl_done));                  // while the search runs
                    EX_CALL(geometry_ahead(ex, (int)si, lv, ch, idx, lvl_done));
                    EX_CALL(dense_to_csr_end(ex, nbr, cnt, ch->n, K, &e));                // host waits for the copy's event only
                    if (ex.oom) return CCN_ENOMEM;
                    EX_CALL(pointconv(ex, pc_iv, pc_fv, H, lv, ch, e, &o));
                }
                outp.x = o;
                outp.lv = ch;
                if (s.kind == ST_SA_GEO) didx = idx;
                break;
            }
            case ST_FP:
            case ST_FP_GEO: {
                // iv: knn, with_xyz, n_layers, lin ids
                if (down_stack.empty()) return CCN_EINVAL;
                const int sidx = down_stack.back();
                down_stack.pop_back();
                const State& sk = hist[sidx];
                Level* fine = sk.lv;
                Feat* x_skip = sk.x ? sk.x : &fine->pos;
                const int knn = (int)iv[0], with_xyz = (int)iv[1], nl = (int)iv[2];
                const bool ahead = ex.up[si].valid;
                int32_t* nbr = ahead ? ex.up[si].nbr : ex.alloc_n<int32_t>(fine->n * knn);
                float* wgt = ahead ? ex.up[si].wgt : ex.alloc_n<float>(fine->n * knn);
                if (ahead) {
                    EX_CUDA(cudaStreamWaitEvent(ex.stream, ex.up[si].ready, 0));     // computed on the geometry stream at level creation
                } else if (s.kind == ST_FP) {
                    EX_CALL(level_sptr(ex, cur.lv));
                    ProfScope ps(ex, FAM_SEARCH, 0, 12.0 * fine->n + 12.0 * cur.lv->n + 8.0 * fine->n * knn);
                    EX_CALL(knn_exact(ex, fine, cur.lv, knn, nbr, wgt));
                } else {
                    EX_CALL(level_topo(ex, fine));
                    const int64_t* idx = down_idx[sidx];
                    if (!idx) return CCN_EINVAL;
                    if (!ex.oom)
                        EX_CALL(ccn_knn_1d_superset(fine->pos.f, fine->curve_ord, fine->n, idx, cur.lv->n, knn, nbr, wgt, st));
                }
                const int c = cur.x->c;
                Feat* y = ex.new_feat(fine->n, c);
                const bool both = fine->n > 0 && (c % 16) == 0 && (cur.x->ld % 4) == 0;
                if (both) y->pl = ex.alloc_planes(fine->n, c);
                if (!ex.oom) {
                    if (both)
                        EX_CALL(ccn_interp_blend_both(cur.x->f, cur.x->ld, nbr, wgt, s.kind == ST_FP, fine->n, knn, c, y->f, y->ld, y->pl.buf,
                                                      y->pl.kpad, y->pl.pstride, ex.npl, st));
                    else
                        EX_CALL(ccn_interp_blend(cur.x->f, cur.x->ld, nbr, wgt, s.kind == ST_FP, fine->n, knn, c, y->f, y->ld, st));
                }
                Feat* parts[3];
                int np = 0;
                parts[np++] = y;
                parts[np++] = x_skip;                      // modules._fp_concat: [y, x_skip, pos_skip]
                if (with_xyz) parts[np++] = &fine->pos;
                Feat* o = nullptr;
                EX_CALL(run_mlp(ex, iv + 3, nl, parts, np, 2, &o));
                outp.x = o;
                outp.lv = fine;
                break;
            }
            default:
                return CCN_EUNSUPPORTED;
        }
        if (ex.oom) return CCN_ENOMEM;
        hist.push_back(outp);
        down_idx.push_back(didx);
        if (ex.skip_store[si]) prop_stack.push_back((int)si);
        if (s.kind == ST_SA || s.kind == ST_SA_GEO) down_stack.push_back((int)si);
        cur = outp;
    }
    // final MLP (base.py:156): logits written straight into the caller's buffer
    Feat* parts[1] = {cur.x};
    Seg segs[4];
    EX_CALL(segs_of(ex, parts, 1, segs));
    int nseg = 1;
    const int64_t M = cur.x->n;
    for (size_t i = 0; i < ex.out_mlp.size(); ++i) {
        const Linear& L = ex.lins[ex.out_mlp[i]];
        const bool last = i + 1 == ex.out_mlp.size();
        if (L.chain >= 0) {
            if (ld_logits < ex.chains[L.chain].n_out) return CCN_EINVAL;
            std::vector<int64_t> ids(ex.out_mlp.begin() + i, ex.out_mlp.end());
            EX_CALL(mlp_chain(ex, ex.chains[L.chain], ids.data(), (int)ids.size(), segs, nseg, M, 0, nullptr, logits, ld_logits));
            break;
        }
        if (!last) {
            Planes yp;
            EX_CALL(linear_planes(ex, L, segs, nseg, M, 1, nullptr, &yp));
            segs[0] = Seg{yp.buf, yp.kpad, yp.pstride, L.n};
            nseg = 1;
            continue;
        }
        if (ex.oom) break;
        const int box = pick_box(L, M, ex.npl);
        const int bs = box_slot(box);
        if (!L.map_hi[bs] || ld_logits < L.n) return CCN_EINVAL;
        EX_CALL(ccn_linear_planes(segs[0].buf, segs[0].kpad, segs[0].pstride, segs[0].k, nullptr, 0, 0, 0, nullptr, 0, 0, 0, nullptr, 0, 0,
                                  0, L.map_hi[bs], L.map_lo[bs], L.bias, logits, ld_logits, nullptr, 0, 0, M, L.n, box, ex.mode, L.act,
                                  L.slope, st));
    }
    return CCN_OK;
}

}  // namespace

extern "C" int ccn_forward_run(ccn_exec_t* h, const float* x, int64_t ldx, const float* pos, const int64_t* batch,
                               const int64_t* curve_idxs, int64_t n, void* arena, int64_t arena_bytes, ccn_draw_fn draw,
                               void* draw_user, float* logits, int64_t ld_logits, int64_t* arena_needed, ccn_stream_t stream) {
    CCN_REQUIRE(h && pos && batch && curve_idxs && n > 0 && arena && arena_bytes > 0 && draw && logits);
    Exec& ex = h->ex;
    ex.arena = (char*)arena; ex.cap = (size_t)arena_bytes; ex.off = 0; ex.oom = false;
    ex.stream = (cudaStream_t)stream; ex.draw = draw; ex.draw_user = draw_user;
    ex.host_blocked_ms = 0; ex.host_syncs = 0;
    const auto t0 = std::chrono::steady_clock::now();
    const int rc = run_forward(ex, x, ldx, pos, batch, curve_idxs, n, logits, ld_logits);
    ex.host_total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (arena_needed) *arena_needed = (int64_t)ex.off;
    if (ex.oom) return CCN_ENOMEM;
    return rc;
}
