// Development aid: how many clusters of each size are co-resident for a 512-thread, 128-register, ~200 KB shared-memory CTA
// (the FPS hybrid kernel's footprint): cudaOccupancyMaxActiveClusters per cluster size.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(512, 1) k(float *p) { extern __shared__ float sm[]; if (p) p[0] = sm[threadIdx.x]; }
int main()
{
    cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 221184);
    for (int smem : {196608, 147456, 98304, 0})
        for (int cs = 1; cs <= 16; ++cs) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(cs * 64); cfg.blockDim = dim3(512); cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            int n = -1;
            cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
            printf("smem %6d cluster %2d: max active clusters %3d -> %3d SMs%s\n", smem, cs, n, n * cs, e == cudaSuccess ? "" : cudaGetErrorString(e));
        }
    return 0;
}
