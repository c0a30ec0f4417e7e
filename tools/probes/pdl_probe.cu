// pdl_probe.cu - what does a dependent kernel launch cost inside a captured CUDA graph on this GPU, with and without
// programmatic dependent launch (cudaLaunchAttributeProgrammaticStreamSerialization + griddepcontrol.wait at the kernel top)?
// Build: nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o tools/probes/pdl_probe tools/probes/pdl_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_work(float* p, int iters, int wait) {
    if (wait) asm volatile("griddepcontrol.wait;" ::: "memory");
    float v = p[blockIdx.x * blockDim.x + threadIdx.x];
    for (int i = 0; i < iters; ++i) v = v * 1.0001f + 0.5f;
    p[blockIdx.x * blockDim.x + threadIdx.x] = v;
}

static float run(int n, int grid, int iters, bool pdl, cudaStream_t st, float* buf) {
    cudaGraph_t g;
    cudaGraphExec_t ge;
    cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
    for (int i = 0; i < n; ++i) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
        cudaLaunchKernelEx(&cfg, k_work, buf, iters, pdl ? 1 : 0);
    }
    if (cudaStreamEndCapture(st, &g) != cudaSuccess) { printf("capture failed: %s\n", cudaGetErrorString(cudaGetLastError())); return -1; }
    if (cudaGraphInstantiate(&ge, g, 0) != cudaSuccess) { printf("instantiate failed: %s\n", cudaGetErrorString(cudaGetLastError())); return -1; }
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 3; ++i) cudaGraphLaunch(ge, st);
    cudaEventRecord(a, st);
    for (int i = 0; i < 10; ++i) cudaGraphLaunch(ge, st);
    cudaEventRecord(b, st);
    cudaStreamSynchronize(st);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
    return ms / 10 * 1000.0f / n;                     // us per kernel
}

int main() {
    cudaStream_t st;
    cudaStreamCreate(&st);
    float* buf;
    cudaMalloc(&buf, 148 * 8 * 256 * sizeof(float));
    cudaMemset(buf, 0, 148 * 8 * 256 * sizeof(float));
    printf("# us per dependent kernel inside a 500-node graph (grid CTAs x 256 threads, `iters` FMAs per thread)\n");
    printf("%8s %8s %12s %12s\n", "grid", "iters", "plain", "pdl");
    const int grids[] = {1, 148, 148 * 4};
    const int its[] = {0, 2000, 20000};
    for (int g : grids)
        for (int it : its) printf("%8d %8d %12.3f %12.3f\n", g, it, run(500, g, it, false, st, buf), run(500, g, it, true, st, buf));
    return 0;
}
