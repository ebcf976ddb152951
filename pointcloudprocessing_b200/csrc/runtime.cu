// runtime.cu — error reporting, launch counter, memory/stream plumbing of the C-ABI.
#include <atomic>
#include <cstdarg>
#include <cstring>

#include "common.cuh"

namespace pcp {

static thread_local char g_err[1024] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

int sm_count() {
    static int cached = 0;
    if (cached == 0) {
        int dev = 0, n = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
            cached = n;
        else
            cached = 148;
    }
    return cached;
}

}  // namespace pcp

using namespace pcp;

extern "C" const char *pcp_last_error(void) { return g_err; }
extern "C" int pcp_version(void) { return 100; }
extern "C" uint64_t pcp_launch_count(void) { return g_launches.load(); }

extern "C" int pcp_device_info(int *sm, int *cc_major, int *cc_minor, size_t *total_mem)
{
    int dev = 0;
    PCP_CUDA_OK(cudaGetDevice(&dev));
    cudaDeviceProp prop;
    PCP_CUDA_OK(cudaGetDeviceProperties(&prop, dev));
    if (sm) *sm = prop.multiProcessorCount;
    if (cc_major) *cc_major = prop.major;
    if (cc_minor) *cc_minor = prop.minor;
    if (total_mem) *total_mem = prop.totalGlobalMem;
    return PCP_OK;
}

extern "C" int pcp_malloc(void **d_ptr, size_t bytes)
{
    PCP_REQUIRE(d_ptr, PCP_ERR_INVALID, "pcp_malloc: null out pointer");
    PCP_CUDA_OK(cudaMalloc(d_ptr, bytes ? bytes : 1));
    return PCP_OK;
}
extern "C" int pcp_free(void *d_ptr) { PCP_CUDA_OK(cudaFree(d_ptr)); return PCP_OK; }
extern "C" int pcp_malloc_host(void **h_ptr, size_t bytes)
{
    PCP_REQUIRE(h_ptr, PCP_ERR_INVALID, "pcp_malloc_host: null out pointer");
    PCP_CUDA_OK(cudaMallocHost(h_ptr, bytes ? bytes : 1));
    return PCP_OK;
}
extern "C" int pcp_free_host(void *h_ptr) { PCP_CUDA_OK(cudaFreeHost(h_ptr)); return PCP_OK; }
extern "C" int pcp_memcpy_h2d(void *d_dst, const void *h_src, size_t bytes, pcp_stream stream)
{
    PCP_CUDA_OK(cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, as_stream(stream)));
    return PCP_OK;
}
extern "C" int pcp_memcpy_d2h(void *h_dst, const void *d_src, size_t bytes, pcp_stream stream)
{
    PCP_CUDA_OK(cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, as_stream(stream)));
    return PCP_OK;
}
extern "C" int pcp_memset(void *d_ptr, int value, size_t bytes, pcp_stream stream)
{
    PCP_CUDA_OK(cudaMemsetAsync(d_ptr, value, bytes, as_stream(stream)));
    return PCP_OK;
}
extern "C" int pcp_stream_sync(pcp_stream stream)
{
    PCP_CUDA_OK(cudaStreamSynchronize(as_stream(stream)));
    return PCP_OK;
}
