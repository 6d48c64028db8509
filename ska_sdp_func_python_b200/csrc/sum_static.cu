// sum_static.cu -- b2g_sumwt_static: the deterministic weight sum of b2g_sumwt (plan.cu,
// image_kernels.cu: sum_stage1 / sum_stage2) for callers that may not allocate device memory
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.
//
// b2g_sumwt takes its 1024 block partials from cudaMallocAsync; here they live in a file-scope
// __device__ array of kStaticParts doubles (8 KiB, held by every GPU that loads the library until
// the process exits).  The entry point calls no cudaMalloc / cudaFree / allocating library
// routine, no copy and no synchronisation: input and output are device pointers and the two
// kernels are only enqueued on the caller's stream.
//
// In host code the array's NAME is a host-side stand-in, not a device address (a kernel handed
// `g_sum_scratch` from the host would fault, or silently use the wrong memory where the GPU can
// reach host memory), so the device address is taken once per device with cudaGetSymbolAddress,
// that device being current, and that pointer is what the kernels get.
//
// Every call on a device shares the one array: two calls that overlap on the GPU (different
// streams, or two host threads on different streams) race on the partials.  Calls on one stream
// are ordered and safe.  Nothing has to start at zero: stage 1 always runs kStaticParts blocks and
// every block overwrites its partial, stage 2 overwrites *out.
//
// The reduction tree is the one of b2g_sumwt (1024 blocks of 256 threads over equal element
// ranges, fixed in-block tree, fixed second stage), so the two entry points agree bit for bit.
#include <cuda_runtime.h>

#include <mutex>

#include "../../include/b200_wgridder.h"

namespace b2g {
int wrapper_fail(const char *msg);  // plan.cu: sets the thread-local b2g_last_error() text
}

namespace {

constexpr int kStaticParts = 1024;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) double g_sum_scratch[kStaticParts];

std::mutex g_addr_mutex;
double *g_scratch_addr[kMaxDevices];   // device address of g_sum_scratch on device i (nullptr: not asked yet)

template <typename Tin>
__global__ void __launch_bounds__(256)
sum_static_stage1(const Tin *__restrict__ in, long long n, long long per, double *__restrict__ part)
{
    __shared__ double sh[256];
    const long long lo = (long long)blockIdx.x * per;
    const long long hi = (lo + per < n) ? lo + per : n;
    double s = 0.0;
    for (long long i = lo + threadIdx.x; i < hi; i += 256) s += (double)in[i];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if ((int)threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) part[blockIdx.x] = sh[0];
}

__global__ void __launch_bounds__(256) sum_static_stage2(const double *__restrict__ part, double *__restrict__ out)
{
    __shared__ double sh[256];
    double s = 0.0;
    for (int i = threadIdx.x; i < kStaticParts; i += 256) s += part[i];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if ((int)threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = sh[0];
}

// device address of the scratch array on the CURRENT device (looked up once per device)
int scratch_for_current_device(double **ptr)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return b2g::wrapper_fail("b2g_sumwt_static: no CUDA device available (this library has no CPU path)");
    if (dev < 0 || dev >= kMaxDevices) return b2g::wrapper_fail("b2g_sumwt_static: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_addr_mutex);
    if (!g_scratch_addr[dev]) {
        double *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_sum_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return b2g::wrapper_fail(e != cudaSuccess ? cudaGetErrorString(e)
                                                      : "b2g_sumwt_static: cudaGetSymbolAddress returned NULL");
        }
        g_scratch_addr[dev] = p;
    }
    *ptr = g_scratch_addr[dev];
    return 0;
}

}  // namespace

extern "C" int b2g_sumwt_static(const void *wgt, int dtype, int64_t n, double *out, void *stream)
{
    if (dtype != B2G_F32 && dtype != B2G_F64) return b2g::wrapper_fail("b2g_sumwt_static: dtype must be B2G_F32 or B2G_F64");
    if (!out) return b2g::wrapper_fail("b2g_sumwt_static: out is NULL");
    if (n < 0) return b2g::wrapper_fail("b2g_sumwt_static: n is negative");
    if (n > (int64_t)B2G_SUMWT_STATIC_MAX_N)
        return b2g::wrapper_fail("b2g_sumwt_static: n exceeds B2G_SUMWT_STATIC_MAX_N (2^32 weights); "
                                 "use b2g_sumwt, which allocates its scratch");
    if (n > 0 && !wgt) return b2g::wrapper_fail("b2g_sumwt_static: wgt is NULL");
    double *scratch = nullptr;
    if (int rc = scratch_for_current_device(&scratch)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    long long per = ((long long)n + kStaticParts - 1) / kStaticParts;
    if (per < 1) per = 1;
    if (dtype == B2G_F64)
        sum_static_stage1<double><<<kStaticParts, 256, 0, s>>>((const double *)wgt, (long long)n, per, scratch);
    else
        sum_static_stage1<float><<<kStaticParts, 256, 0, s>>>((const float *)wgt, (long long)n, per, scratch);
    sum_static_stage2<<<1, 256, 0, s>>>(scratch, out);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : b2g::wrapper_fail(cudaGetErrorString(e));
}
