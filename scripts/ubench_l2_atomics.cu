// Development micro-benchmark (round 2, voxel analysis): chip-wide cost of the L2-side primitives a voxel accumulation can be
// built from, at the C2 shape: 64 clouds x 100 000 points in random order, ~11.7 k voxels per cloud.  Every variant touches
// every point once.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/_bin/ubench_l2_atomics scripts/ubench_l2_atomics.cu
#include <cstdio>
#include <cuda_runtime.h>
constexpr int kB = 64, kN = 100000, kV = 11700;
__device__ __forceinline__ unsigned hash32(unsigned x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }
template <int MODE>
__global__ void __launch_bounds__(256) k(const float *__restrict__ x, double *acc, unsigned *acc32, unsigned long long *acc64, float4 *accf, unsigned *sink)
{
    const long long total = (long long)kB * kN;
    unsigned s = 0;
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const int cloud = (int)(i / kN);
        const long long slot = (long long)cloud * kV + hash32((unsigned)i) % kV;
        if (MODE == 0) { atomicAdd(acc + slot * 4, 1.0); atomicAdd(acc + slot * 4 + 1, 2.0); atomicAdd(acc + slot * 4 + 2, 3.0); atomicAdd(acc + slot * 4 + 3, 1.0); }
        else if (MODE == 1) s += atomicExch(acc32 + slot, (unsigned)i);
        else if (MODE == 2) s += atomicAdd(acc32 + slot, 1u);
        else if (MODE == 3) atomicAdd(acc32 + slot, 1u);
        else if (MODE == 4) asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(accf + slot), "f"(1.f), "f"(2.f), "f"(3.f), "f"(1.f) : "memory");
        else if (MODE == 5) { unsigned long long *p = acc64 + slot * 4; atomicAdd(p, 5ull); atomicAdd(p + 1, 7ull); atomicAdd(p + 2, 9ull); atomicAdd(reinterpret_cast<unsigned *>(p + 3), 1u); }
        else if (MODE == 6) s += __ldcg(acc32 + slot);
        else if (MODE == 7) { const long long j = (long long)cloud * kN + hash32((unsigned)i * 3u + 1u) % kN; s += __float_as_uint(__ldcg(x + 3 * j)) + __float_as_uint(__ldcg(x + 3 * j + 1)) + __float_as_uint(__ldcg(x + 3 * j + 2)); }
        else if (MODE == 8) acc32[(long long)cloud * kN + hash32((unsigned)i * 5u + 2u) % kN] = (unsigned)i;
    }
    if (s == 0xdeadbeefu) sink[0] = s;
}
template <int MODE> static void run(const char *name)
{
    const size_t slots = (size_t)kB * kN;
    float *x; double *acc; unsigned *sink;
    cudaMalloc(&x, slots * 12); cudaMalloc(&acc, slots * 32); cudaMalloc(&sink, 4);
    cudaMemset(x, 0, slots * 12); cudaMemset(acc, 0, slots * 32);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        k<MODE><<<sms * 8, 256>>>(x, acc, (unsigned *)acc, (unsigned long long *)acc, (float4 *)acc, sink);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    printf("%-58s %8.1f us  (%6.1f G points/s)\n", name, best * 1e3, (double)kB * kN / best * 1e-6);
    cudaFree(x); cudaFree(acc); cudaFree(sink);
}
int main()
{
    run<0>("4 x red.f64 in one 32-byte sector (the accumulate pass)");
    run<1>("1 x atom.exch.b32 with return");
    run<2>("1 x atom.add.u32 with return");
    run<3>("1 x red.add.u32");
    run<4>("1 x red.add.v4.f32");
    run<5>("3 x red.add.u64 + 1 x red.add.u32 in one sector");
    run<6>("1 x random 4-byte L2 load (bitmap / rank look-up)");
    run<7>("1 x random 12-byte gather");
    run<8>("1 x random 4-byte store");
    return 0;
}
