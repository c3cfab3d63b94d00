// Micro-benchmark: what a dependent-launch boundary costs on this GPU, with and without programmatic dependent launch
// (cudaLaunchAttributeProgrammaticStreamSerialization + griddepcontrol.wait at the top of the dependent kernel).
// The per-scan forward is a chain of ~400 dependent launches of 3 - 30 us kernels; this decides whether converting the
// launches of csrc/ to PDL is worth it.   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/bench_pdl.cu -o /tmp/bench_pdl
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>   // 0: plain, 1: wait at the top, 2: launch_dependents + wait at the top
__global__ void step_kernel(const float* __restrict__ in, float* __restrict__ out, int n, int work) {
    __shared__ float stage[256];
    stage[threadIdx.x] = 0.f;                       // a prologue that does not depend on the previous kernel
    __syncthreads();
    if (MODE == 2) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (MODE >= 1) asm volatile("griddepcontrol.wait;" ::: "memory");
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float v = in[i];
        for (int k = 0; k < work; ++k) v = v * 1.0000001f + 1e-7f;
        out[i] = v + stage[threadIdx.x];
    }
}

template <int MODE>
float run_chain(float* a, float* b, int n, int work, int grid, int launches, cudaStream_t s) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(256);
    cfg.stream = s;
    cfg.attrs = attr;
    cfg.numAttrs = MODE >= 1 ? 1 : 0;
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaStreamSynchronize(s);
        cudaEventRecord(e0, s);
        for (int i = 0; i < launches; ++i) {
            const float* in = (i & 1) ? b : a;
            float* out = (i & 1) ? a : b;
            cudaLaunchKernelEx(&cfg, step_kernel<MODE>, in, out, n, work);
        }
        cudaEventRecord(e1, s);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best * 1000.f / launches;
}

int main() {
    cudaStream_t s;
    cudaStreamCreate(&s);
    const int launches = 2000;
    const int shapes[4][3] = {{148 * 256, 1, 148}, {148 * 256 * 8, 64, 148 * 8}, {1 << 22, 256, 148 * 16}, {1 << 24, 64, 148 * 32}};
    for (auto& sh : shapes) {
        float *a, *b;
        cudaMalloc(&a, sizeof(float) * sh[0]);
        cudaMalloc(&b, sizeof(float) * sh[0]);
        cudaMemset(a, 0, sizeof(float) * sh[0]);
        cudaMemset(b, 0, sizeof(float) * sh[0]);
        const float t0 = run_chain<0>(a, b, sh[0], sh[1], sh[2], launches, s);
        const float t1 = run_chain<1>(a, b, sh[0], sh[1], sh[2], launches, s);
        const float t2 = run_chain<2>(a, b, sh[0], sh[1], sh[2], launches, s);
        float chk = 0;
        cudaMemcpy(&chk, a, 4, cudaMemcpyDeviceToHost);
        printf("n=%9d work=%4d grid=%5d  plain %7.2f us/launch   pdl(wait) %7.2f   pdl(trigger+wait) %7.2f   [%g] %s\n", sh[0], sh[1],
               sh[2], t0, t1, t2, chk, cudaGetErrorString(cudaGetLastError()));
        cudaFree(a);
        cudaFree(b);
    }
    return 0;
}
