// Can SM stores to mapped pinned host memory match the copy engine's D2H rate?  (decides whether the
// SE_MEM_HOST path should write results straight to host instead of staging + cudaMemcpyAsync)
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

// each warp writes `chunk` consecutive floats per iteration (the NoiseSup kernel writes 160 per block)
__global__ void writer(float *dst, size_t n, int chunk)
{
    const int lane = threadIdx.x & 31;
    const size_t warp = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5, nwarps = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t base = warp * chunk; base + chunk <= n; base += nwarps * chunk)
        for (int i = lane; i < chunk; i += 32) dst[base + i] = (float)i;
}
__global__ void reader(const short *src, size_t n, int chunk, int *sink)
{
    const int lane = threadIdx.x & 31;
    const size_t warp = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5, nwarps = ((size_t)gridDim.x * blockDim.x) >> 5;
    int acc = 0;
    for (size_t base = warp * chunk; base + chunk <= n; base += nwarps * chunk)
        for (int i = lane; i < chunk; i += 32) acc += src[base + i];
    if (acc == 123456789) *sink = acc;
}

int main()
{
    const size_t bytes = 2ull << 30;
    float *h = nullptr, *d = nullptr, *hd = nullptr;
    int *sink;
    CK(cudaHostAlloc(&h, bytes, cudaHostAllocMapped));
    CK(cudaHostGetDevicePointer(&hd, h, 0));
    CK(cudaMalloc(&d, bytes));
    CK(cudaMalloc(&sink, 4));
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    float ms;
    for (int rep = 0; rep < 2; rep++) {
        cudaEventRecord(a);
        CK(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost));
        cudaEventRecord(b); CK(cudaEventSynchronize(b)); cudaEventElapsedTime(&ms, a, b);
        printf("copy engine D2H      : %.1f GB/s\n", bytes / 1e9 / (ms * 1e-3));
    }
    const int grids[] = {148, 296, 592};
    for (int g : grids)
        for (int chunk : {160, 1024}) {
            writer<<<g, 256>>>(hd, bytes / 4, chunk);
            CK(cudaDeviceSynchronize());
            cudaEventRecord(a);
            writer<<<g, 256>>>(hd, bytes / 4, chunk);
            cudaEventRecord(b); CK(cudaEventSynchronize(b)); cudaEventElapsedTime(&ms, a, b);
            printf("SM stores to host, grid %3d x 256, %4d floats/warp-iter: %.1f GB/s\n", g, chunk, bytes / 1e9 / (ms * 1e-3));
        }
    for (int g : grids) {
        reader<<<g, 256>>>((const short *)hd, bytes / 2, 160, sink);
        CK(cudaDeviceSynchronize());
        cudaEventRecord(a);
        reader<<<g, 256>>>((const short *)hd, bytes / 2, 160, sink);
        cudaEventRecord(b); CK(cudaEventSynchronize(b)); cudaEventElapsedTime(&ms, a, b);
        printf("SM loads from host, grid %3d x 256, 160 int16/warp-iter: %.1f GB/s\n", g, bytes / 1e9 / (ms * 1e-3));
    }
    return 0;
}
