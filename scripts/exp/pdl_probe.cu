// Probe: cost of a kernel boundary with and without programmatic dependent launch (PDL) on B200.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pdl_probe pdl_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
template <bool PDL>
__global__ void k_copy(const float4* __restrict__ in, float4* __restrict__ out, long long n) {
    if (PDL) { asm volatile("griddepcontrol.wait;" ::: "memory"); asm volatile("griddepcontrol.launch_dependents;"); }
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) out[i] = in[i];
}
template <bool PDL>
float run(const float4* a, float4* b, long long n, int grid, int iters, cudaStream_t st) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.stream = st;
    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
    if (PDL) { cfg.attrs = at; cfg.numAttrs = 1; }
    for (int w = 0; w < 2; ++w) {
        cudaEventRecord(e0, st);
        for (int i = 0; i < iters; ++i) { cudaLaunchKernelEx(&cfg, k_copy<PDL>, (i & 1) ? (const float4*)b : a, (i & 1) ? (float4*)a : b, n); }
        cudaEventRecord(e1, st); cudaEventSynchronize(e1);
    }
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms * 1000.f / iters;
}
int main() {
    cudaStream_t st; cudaStreamCreate(&st);
    for (long long mb : {1LL, 16LL, 104LL, 210LL}) {
        long long n = mb * 1000000 / 16; float4 *a, *b; cudaMalloc(&a, n * 16); cudaMalloc(&b, n * 16); cudaMemset(a, 0, n * 16);
        for (int grid : {148 * 4, 148 * 8}) {
            float t0 = run<false>(a, b, n, grid, 200, st), t1 = run<true>(a, b, n, grid, 200, st);
            printf("copy %4lld MB grid %5d: plain %7.2f us/kernel (%5.0f GB/s)   PDL %7.2f us/kernel (%5.0f GB/s)\n", mb, grid, t0, 2.0 * mb * 1e6 / t0 / 1e3, t1, 2.0 * mb * 1e6 / t1 / 1e3);
        }
        cudaFree(a); cudaFree(b);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
