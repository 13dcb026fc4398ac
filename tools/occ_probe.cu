// How many 192-thread CTAs the driver places on one SM as a function of their dynamic shared memory.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/occ_probe tools/occ_probe.cu && tools/occ_probe
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(192, 2) dummy(int* out) {
    extern __shared__ unsigned char raw[];
    if (out) out[0] = raw[threadIdx.x];
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    printf("sharedMemPerMultiprocessor %zu  sharedMemPerBlockOptin %zu  reservedSharedMemPerBlock %zu  regsPerMultiprocessor %d\n",
           p.sharedMemPerMultiprocessor, p.sharedMemPerBlockOptin, p.reservedSharedMemPerBlock, p.regsPerMultiprocessor);
    cudaFuncSetAttribute(dummy, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    for (int carve = 0; carve < 2; ++carve) {
        if (carve) cudaFuncSetAttribute(dummy, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        int last = -1;
        for (int bytes = 96 * 1024; bytes <= 117 * 1024; bytes += 256) {
            int n = 0;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, dummy, 192, bytes);
            if (n != last) printf("carveout %s: from %d bytes (%.2f KiB) on: %d CTAs per SM\n", carve ? "max" : "default", bytes, bytes / 1024.0, n);
            last = n;
        }
    }
    return 0;
}
