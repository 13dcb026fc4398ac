// Where do the CTAs of a persistent 148-CTA launch with clusters of 2 land?  Prints blockIdx -> %smid.
#include <cuda_runtime.h>
#include <cstdio>
__global__ void k(int* out) {
    extern __shared__ char s[];
    unsigned smid, nsmid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    asm volatile("mov.u32 %0, %%nsmid;" : "=r"(nsmid));
    if (threadIdx.x == 0) { out[blockIdx.x * 2] = smid; out[blockIdx.x * 2 + 1] = nsmid; }
    // keep every CTA resident until all have started, so that no SM takes two
    __nanosleep(200000);
}
int main() {
    int n = 148; int* d; cudaMalloc(&d, n * 8);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int rep = 0; rep < 2; ++rep) {
        cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3(n); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 200 * 1024;
        cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        cudaError_t e = cudaLaunchKernelEx(&cfg, k, d);
        int h[296]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        printf("launch %d (%s) nsmid=%d\nblock:smid ", rep, cudaGetErrorString(e), h[1]);
        for (int i = 0; i < n; ++i) printf("%d:%d ", i, h[2 * i]);
        printf("\n");
    }
    return 0;
}
