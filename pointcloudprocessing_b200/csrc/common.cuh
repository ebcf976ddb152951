// common.cuh — shared host/device helpers for libpcp_b200.so (sm_100a only).
#pragma once
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/pcp.h"
#include "../../include/pcp_dev.h"

namespace pcp {

void set_error(const char *fmt, ...);
void count_launch(int n = 1);

#define PCP_CUDA_OK(expr)                                                                         \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess) {                                                                  \
            pcp::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return PCP_ERR_CUDA;                                                                  \
        }                                                                                         \
    } while (0)

// teardown paths cannot return an error: a failing release is reported on stderr and cleared so it does not surface later
// as somebody else's "last error"
#define PCP_RELEASE(call)                                                                        \
    do {                                                                                        \
        cudaError_t _e = (call);                                                                \
        if (_e != cudaSuccess) {                                                                \
            fprintf(stderr, "pcp: %s:%d: %s -> %s\n", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
            (void)cudaGetLastError();                                                           \
        }                                                                                       \
    } while (0)

#define PCP_CHECK_LAUNCH()                                                                      \
    do {                                                                                        \
        cudaError_t _e = cudaGetLastError();                                                    \
        if (_e != cudaSuccess) {                                                                \
            pcp::set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
            return PCP_ERR_CUDA;                                                                \
        }                                                                                       \
        pcp::count_launch();                                                                    \
    } while (0)

#define PCP_REQUIRE(cond, code, ...)      \
    do {                                  \
        if (!(cond)) {                    \
            pcp::set_error(__VA_ARGS__);  \
            return (code);                \
        }                                 \
    } while (0)

#define PCP_TRY(expr)              \
    do {                           \
        int _s = (expr);           \
        if (_s != PCP_OK) return _s; \
    } while (0)

int sm_count();

static inline cudaStream_t as_stream(pcp_stream s) { return reinterpret_cast<cudaStream_t>(s); }

// ---- device helpers ----
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Philox4x32-10 (Salmon et al., SC'11) — counter-based RNG shared by jitter / resize / dropout.
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)c[0] * 0xD2511F53ull;
        const uint64_t p1 = (uint64_t)c[2] * 0xCD9E8D57ull;
        const uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
        const uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
        const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
        c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}

}  // namespace pcp
