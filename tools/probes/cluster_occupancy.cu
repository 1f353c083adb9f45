// How many clusters of a given size can be co-resident with the persistent kernel's footprint (256 threads,
// ~200 KB dynamic shared memory, 1 CTA per SM)?  Decides whether a 4-CTA cluster variant can use all 148 SMs.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256, 1) dummy(int* p) { extern __shared__ char s[]; if (p) p[0] = s[0]; }
int main() {
    const int smem = 201872;
    cudaFuncSetAttribute(dummy, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(dummy, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    for (int cs : {1, 2, 4, 8, 16}) {
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(148 / cs * cs);
        cfg.blockDim = dim3(256);
        cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        int n = -1;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&n, dummy, &cfg);
        printf("cluster %2d: max active clusters %d (%d SMs) %s\n", cs, n, n * cs, e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    printf("SMs %d, L2 %d MB\n", prop.multiProcessorCount, prop.l2CacheSize >> 20);
    return 0;
}
