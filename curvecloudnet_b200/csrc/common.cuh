// Shared device/host helpers for libccn_b200 (sm_100a).  See include/ccn_b200.h for the C-ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ccn_b200.h"

extern unsigned long long g_ccn_kernel_launches;   // version.cu; one increment per <<<>>> (bench.py's gpu_launches)

// gemm_tc.cu: CUtensorMap (128 bytes, 64-byte aligned, HOST memory) of a bf16 (n, kpad) row-major matrix, box 64 x box_n
int ccn_encode_bf16_map(void* tmap_out_host, const void* base, int64_t n, int64_t kpad, int32_t box_n);

// ------------------------------------------------------------------------------------------
// Programmatic dependent launch.  The per-scan forward is a chain of ~400 dependent launches of 3 - 30 us kernels, and a
// plain stream-ordered launch boundary costs 3 - 4 us on B200 (tools/bench_pdl.cu, profiles/r02p_pdl_micro.txt: 3.07 us
// between two one-wave kernels, 1.87 us with the attribute below).  EVERY kernel of this library is launched through
// ccn_launch with cudaLaunchAttributeProgrammaticStreamSerialization and executes ccn_pdl_wait() (griddepcontrol.wait:
// all prerequisite grids complete, their memory visible) before it touches global memory - as its first statement, or, in
// the tensor-core kernels, after the prologue that does not depend on the previous kernel (barrier init, TMEM allocation,
// descriptor prefetch).  Since every kernel waits, completion stays transitive along the stream.  No kernel triggers its
// dependents early (measured: no gain).  ccn_set_dependent_launch(0) / CCN_PDL=0 launches without the attribute.
// ------------------------------------------------------------------------------------------
extern int g_ccn_pdl;                               // version.cu

__device__ __forceinline__ void ccn_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
static inline cudaError_t ccn_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                     Args&&... args) {
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cfg.attrs = attr;
    cfg.numAttrs = g_ccn_pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<Args&&>(args)...);
}
#endif

#define CCN_CHECK_LAUNCH()                                 \
    do {                                                   \
        cudaError_t e__ = cudaGetLastError();              \
        if (e__ != cudaSuccess) return (int)e__;           \
        ++g_ccn_kernel_launches;                           \
    } while (0)

#define CCN_REQUIRE(cond)                 \
    do {                                  \
        if (!(cond)) return CCN_EINVAL;   \
    } while (0)

static inline unsigned ccn_grid(int64_t work, int threads, int64_t cap = 148LL * 64) {
    int64_t b = (work + threads - 1) / threads;
    if (b < 1) b = 1;
    if (b > cap) b = cap;
    return (unsigned)b;
}

static inline unsigned ccn_blocks_exact(int64_t work, int threads) {
    int64_t b = (work + threads - 1) / threads;
    return (unsigned)(b < 1 ? 1 : b);
}

// ((dx*dx)+(dy*dy))+(dz*dz), every operation rounded separately (no FMA contraction): the total
// order (d2, index) that decides every neighbour index is defined on exactly this value
// (oracle/tp_ops.py sqdist3).
__device__ __forceinline__ float ccn_sqdist3(float ax, float ay, float az, float bx, float by, float bz) {
    float dx = __fsub_rn(ax, bx), dy = __fsub_rn(ay, by), dz = __fsub_rn(az, bz);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

// |v| as torch.linalg.norm(v, dim=-1) computes it for (N,3) f32 on CPU:
// sqrt(fma(z,z,fma(y,y,x*x))) with a correctly rounded sqrt (oracle/ccn_point_ops.py norm3).
__device__ __forceinline__ float ccn_norm3(float x, float y, float z) {
    return __fsqrt_rn(__fmaf_rn(z, z, __fmaf_rn(y, y, __fmul_rn(x, x))));
}

// (d2, idx) packed so that unsigned 64-bit order == lexicographic (d2, idx) order for d2 >= 0.
__device__ __forceinline__ unsigned long long ccn_pack_key(float d2, int idx) {
    return ((unsigned long long)__float_as_uint(d2) << 32) | (unsigned)idx;
}

// q = t / c, rem = t % c for a flat element index: 32-bit division whenever the index fits (a 64-bit integer division
// costs several times the rest of an element-wise kernel's body)
__device__ __forceinline__ void ccn_divmod(int64_t t, int c, int64_t& q, int& rem) {
    if ((unsigned long long)t < 0x80000000ull) {
        const unsigned tt = (unsigned)t, qq = tt / (unsigned)c;
        q = qq;
        rem = (int)(tt - qq * (unsigned)c);
    } else {
        q = t / c;
        rem = (int)(t - q * c);
    }
}

__device__ __forceinline__ int ccn_lane() { return threadIdx.x & 31; }
__device__ __forceinline__ int ccn_warp() { return threadIdx.x >> 5; }

// ------------------------------------------------------------------------------------------
// Block-wide exclusive scan (blockDim.x a multiple of 32, <= 1024).  `smem` holds 33 T's.
// Returns the exclusive prefix of `v`; *total receives the block sum (same in every thread).
// ------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T ccn_block_excl_scan(T v, T* smem, T* total) {
    const int lane = ccn_lane(), warp = ccn_warp(), nwarp = (blockDim.x + 31) >> 5;
    T incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        T t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) smem[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        T w = lane < nwarp ? smem[lane] : T(0);
        T wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            T t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        smem[lane] = wi - w;                 // exclusive prefix of the warp sums
        if (lane == 31) smem[32] = wi;       // block total
    }
    __syncthreads();
    T res = smem[warp] + (incl - v);
    *total = smem[32];
    __syncthreads();                         // smem may be reused by the caller's next round
    return res;
}
