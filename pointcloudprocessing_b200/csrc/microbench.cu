// microbench.cu — development micro-benchmarks of the sm_100a primitives the chain kernels are built from
// (tcgen05.mma issue rate from shared-memory operands, one CTA and CTA pair; tcgen05.ld throughput).
// Not part of the product path and not declared in include/pcp.h; driven by scripts/umma_rate.py and
// scripts/tmem_rate.py.  The numbers they produced are recorded in DESIGN.md §4.1.
#include "common.cuh"
#include "umma.cuh"

namespace pcp {
using namespace umma;
// ---------------------------------------------------------------------------------------------
// Micro-benchmark (development): issue rate of back-to-back tcgen05.mma from shared-memory operands.
// out[0] = cycles, out[1] = MMAs issued.  mode 0: cta_group::1; the operands are whatever is in smem.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(288, 1)
umma_rate_kernel(int N, int iters, int kblocks, int mode, long long *out)
{
    // mode bits: 1 = warps 4-7 stream tcgen05.ld from the idle accumulator half, 2 = warps 1-3 stream st.shared.v4,
    //            4 = warps 1-3 spin on an mbarrier that never completes, 8 = MMA thread polls an mbarrier between batches
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t done, never;
    __shared__ uint32_t tmem_base;
    __shared__ volatile int stop;
    uint8_t *base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (128 * 128 + 256 * 128) * kblocks / 4 + 8192; i += 288) reinterpret_cast<uint32_t *>(base)[i] = 0x3c003c00u;
    if (tid == 0) { mbar_init(&done, 1); mbar_init(&never, 1); stop = 0; fence_barrier_init(); }
    if (warp == 0) tmem_alloc(&tmem_base, 512);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base;
    uint8_t *scratch = base + (128 * 128 + 256 * 128) * kblocks;
    if (tid == 0) {
        const uint32_t a0 = smem_u32(base), b0 = a0 + 128 * 128 * kblocks;
        const uint32_t idesc = make_idesc_bf16(128, (uint32_t)N);
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            if (mode & 8) { (void)mbar_test(&never, 0); (void)mbar_test(&never, 0); tc_fence_after(); }
            for (int kb = 0; kb < kblocks; ++kb)
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    mma_bf16_ss(tmem, make_sdesc(a0 + kb * 16384 + k * 32), make_sdesc(b0 + kb * (N * 128) + k * 32), idesc, (kb | k) > 0);
        }
        mma_commit(&done);
        mbar_wait(&done, 0, 30);
        const long long t1 = clock64();
        out[0] = t1 - t0;
        out[1] = (long long)iters * kblocks * 4;
        stop = 1;
    } else if (warp >= 4 && warp < 8 && (mode & 1)) {
        uint32_t acc = 0;
        long long n = 0;
        const long long t0 = clock64();
        while (!stop) {
            uint32_t v0[32], v1[32], v2[32], v3[32];
            const uint32_t ta = tmem + ((uint32_t)((warp - 4) * 32) << 16) + 256;
            tmem_ld32(ta, v0); tmem_ld32(ta + 32, v1); tmem_ld32(ta + 64, v2); tmem_ld32(ta + 96, v3);
            tmem_ld_wait();
            acc += v0[0] ^ v1[1] ^ v2[2] ^ v3[3];
            ++n;
        }
        const long long t1 = clock64();
        if ((tid & 31) == 0) { out[2 + (warp - 4) * 2] = t1 - t0; out[3 + (warp - 4) * 2] = n; }
        if (acc == 0x12345678u) out[1] = 0;
    } else if (warp >= 1 && warp < 4 && (mode & 2)) {
        uint32_t i = 0;
        while (!stop) {
            *reinterpret_cast<uint4 *>(scratch + ((tid * 16 + (i++ & 7) * 4096) & 32767)) = make_uint4(i, i, i, i);
        }
    } else if (warp >= 1 && warp < 4 && (mode & 4)) {
        while (!stop) (void)mbar_try(&never, 0);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}


// Micro-benchmark (development): tcgen05.ld throughput.  out[w] = cycles for `iters` x (4 x 32x32b.x32) per warp.
__global__ void __launch_bounds__(128, 1)
tmem_ld_rate_kernel(int iters, int nwarps, int pack16, long long *out)
{
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (warp == 0) tmem_alloc(&tmem_base, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base;
    if (warp < nwarps) {
        uint32_t acc = 0;
        const uint32_t ta = tmem + ((uint32_t)(warp * 32) << 16);
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            uint32_t v0[32], v1[32], v2[32], v3[32];
            if (pack16) {
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 "
                             "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                             "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                             : "=r"(v0[0]), "=r"(v0[1]), "=r"(v0[2]), "=r"(v0[3]), "=r"(v0[4]), "=r"(v0[5]), "=r"(v0[6]), "=r"(v0[7]),
                               "=r"(v0[8]), "=r"(v0[9]), "=r"(v0[10]), "=r"(v0[11]), "=r"(v0[12]), "=r"(v0[13]), "=r"(v0[14]), "=r"(v0[15]),
                               "=r"(v0[16]), "=r"(v0[17]), "=r"(v0[18]), "=r"(v0[19]), "=r"(v0[20]), "=r"(v0[21]), "=r"(v0[22]), "=r"(v0[23]),
                               "=r"(v0[24]), "=r"(v0[25]), "=r"(v0[26]), "=r"(v0[27]), "=r"(v0[28]), "=r"(v0[29]), "=r"(v0[30]), "=r"(v0[31])
                             : "r"(ta) : "memory");
                tmem_ld_wait();
                acc += v0[3];
            } else {
                tmem_ld32(ta, v0); tmem_ld32(ta + 32, v1); tmem_ld32(ta + 64, v2); tmem_ld32(ta + 96, v3);
                tmem_ld_wait();
                acc += v0[3] ^ v1[5] ^ v2[7] ^ v3[9];
            }
        }
        const long long t1 = clock64();
        if (tid % 32 == 0) out[warp] = t1 - t0;
        if (acc == 0x12345u) out[7] = 1;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

// CTA-pair issue rate: cluster of 2, leader issues M = 256 MMAs (cta_group::2), both CTAs hold operand halves.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
umma2_rate_kernel(int N, int iters, int kblocks, long long *out)
{
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t done;
    __shared__ uint32_t tmem_base;
    uint8_t *base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t rank = cluster_ctarank();
    for (int i = tid; i < (128 * 128 + 128 * 128) * kblocks / 4; i += 128) reinterpret_cast<uint32_t *>(base)[i] = 0x3c003c00u;
    if (tid == 0) { mbar_init(&done, 1); fence_barrier_init(); }
    if (warp == 0) tmem_alloc2(&tmem_base, 512);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem = tmem_base;
    if (tid == 0 && rank == 0) {
        const uint32_t a0 = smem_u32(base), b0 = a0 + 128 * 128 * kblocks;
        const uint32_t idesc = make_idesc_bf16(256, (uint32_t)N);
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it)
            for (int kb = 0; kb < kblocks; ++kb)
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    mma2_bf16_ss(tmem, make_sdesc(a0 + kb * 16384 + k * 32), make_sdesc(b0 + kb * ((N / 2) * 128) + k * 32), idesc, (kb | k) > 0);
        mma2_commit_mc(&done, 3);
        mbar_wait(&done, 0, 31);
        out[0] = clock64() - t0;
        out[1] = (long long)iters * kblocks * 4;
    } else if (tid == 0) {
        mbar_wait(&done, 0, 32);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 0) tmem_dealloc2(tmem, 512);
}

}  // namespace pcp

using namespace pcp;

extern "C" int pcp_selftest_umma_rate(int N, int iters, int kblocks, int nctas, int mode, long long *h_out)
{
    long long *d = nullptr;
    PCP_CUDA_OK(cudaMalloc(&d, 80));
    PCP_CUDA_OK(cudaMemset(d, 0, 80));
    const size_t smem = (size_t)(128 * 128 + 256 * 128) * kblocks + 32768 + 1024;
    PCP_CUDA_OK(cudaFuncSetAttribute(umma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    umma_rate_kernel<<<nctas, 288, smem>>>(N, iters, kblocks, mode, d);
    PCP_CHECK_LAUNCH();
    PCP_CUDA_OK(cudaDeviceSynchronize());
    PCP_CUDA_OK(cudaMemcpy(h_out, d, 80, cudaMemcpyDeviceToHost));
    cudaFree(d);
    return PCP_OK;
}

extern "C" int pcp_selftest_tmem_ld_rate(int iters, int nwarps, int pack16, long long *h_out)
{
    long long *d = nullptr;
    PCP_CUDA_OK(cudaMalloc(&d, 64));
    PCP_CUDA_OK(cudaMemset(d, 0, 64));
    tmem_ld_rate_kernel<<<1, 128>>>(iters, nwarps, pack16, d);
    PCP_CHECK_LAUNCH();
    PCP_CUDA_OK(cudaDeviceSynchronize());
    PCP_CUDA_OK(cudaMemcpy(h_out, d, 64, cudaMemcpyDeviceToHost));
    cudaFree(d);
    return PCP_OK;
}


extern "C" int pcp_selftest_umma2_rate(int N, int iters, int kblocks, int nclusters, long long *h_out)
{
    long long *d = nullptr;
    PCP_CUDA_OK(cudaMalloc(&d, 80));
    PCP_CUDA_OK(cudaMemset(d, 0, 80));
    const size_t smem = (size_t)(128 * 128 + 128 * 128) * kblocks + 1024;
    PCP_CUDA_OK(cudaFuncSetAttribute(umma2_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    umma2_rate_kernel<<<2 * nclusters, 128, smem>>>(N, iters, kblocks, d);
    PCP_CHECK_LAUNCH();
    PCP_CUDA_OK(cudaDeviceSynchronize());
    PCP_CUDA_OK(cudaMemcpy(h_out, d, 80, cudaMemcpyDeviceToHost));
    cudaFree(d);
    return PCP_OK;
}
