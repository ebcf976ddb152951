// Development micro-benchmark (round 2, voxel analysis): random-address shared-memory operations on the CTA's own shared
// memory against the same operations on a PEER CTA's shared memory (cluster of 8, DSMEM), lane-operations per clock and SM.
#include <cooperative_groups.h>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;
constexpr int kWords = 16384, kIters = 64;
__device__ __forceinline__ unsigned rng(unsigned &s) { s = s * 1664525u + 1013904223u; return s >> 8; }
template <int MODE, int CLUSTER>
__global__ void __launch_bounds__(1024) bench(unsigned long long *cycles, unsigned *sink)
{
    extern __shared__ unsigned sm[];
    for (int i = threadIdx.x; i < kWords; i += blockDim.x) sm[i] = 0;
    unsigned seed = blockIdx.x * 7919u + threadIdx.x * 104729u + 1u, acc = 0;
    unsigned *remote[8];
    if (CLUSTER > 1) { cg::cluster_group cl = cg::this_cluster(); for (int r = 0; r < CLUSTER; ++r) remote[r] = cl.map_shared_rank(sm, r); cl.sync(); }
    else { remote[0] = sm; __syncthreads(); }
    const long long t0 = clock64();
#pragma unroll 8
    for (int it = 0; it < kIters; ++it) {
        const unsigned r = rng(seed), w = r % kWords;
        unsigned *base = CLUSTER > 1 ? remote[(r >> 16) % CLUSTER] : sm;
        if (MODE == 0) atomicOr(base + w, 1u << (r & 31));
        else if (MODE == 1) acc += atomicAdd(base + w, 1u);
        else if (MODE == 2) acc += base[w];
        else if (MODE == 3) base[w] = r;
        else if (MODE == 4) { float *f = reinterpret_cast<float *>(base + (w / 3) * 3 % (kWords - 4)); f[0] = 1.f; f[1] = 2.f; f[2] = 3.f; }
    }
    if (CLUSTER > 1) cg::this_cluster().sync(); else __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
    if (acc == 0xdeadbeefu) sink[0] = acc;
}
template <int MODE, int CLUSTER> static void run(const char *name, int threads)
{
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int grid = (sms / CLUSTER) * CLUSTER;
    unsigned long long *cyc; unsigned *sink;
    cudaMalloc(&cyc, grid * 8); cudaMalloc(&sink, 4);
    cudaFuncSetAttribute(bench<MODE, CLUSTER>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWords * 4);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = kWords * 4;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CLUSTER; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    for (int rep = 0; rep < 2; ++rep) cudaLaunchKernelEx(&cfg, bench<MODE, CLUSTER>, cyc, sink);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("%s: launch failed\n", name); return; }
    unsigned long long *h = (unsigned long long *)malloc(grid * 8);
    cudaMemcpy(h, cyc, grid * 8, cudaMemcpyDeviceToHost);
    double mean = 0; for (int i = 0; i < grid; ++i) mean += (double)h[i]; mean /= grid;
    printf("%-36s cluster %d threads %4d: %.3f lane-ops/clk/SM (%.2f clk per op)\n", name, CLUSTER, threads, (double)threads * kIters / mean, mean / ((double)threads * kIters));
    free(h); cudaFree(cyc); cudaFree(sink);
}
int main()
{
    for (int threads : {256, 1024}) {
        run<0, 1>("local  atomicOr (no return)", threads); run<1, 1>("local  atomicAdd (return used)", threads);
        run<2, 1>("local  load 4 B", threads); run<3, 1>("local  store 4 B", threads); run<4, 1>("local  store 12 B", threads);
        run<0, 8>("remote atomicOr (no return)", threads); run<1, 8>("remote atomicAdd (return used)", threads);
        run<2, 8>("remote load 4 B", threads); run<3, 8>("remote store 4 B", threads); run<4, 8>("remote store 12 B", threads);
    }
    return 0;
}
