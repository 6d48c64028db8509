// Throughput probe for the next-round design ("What comes next"): how many
// legacy mma.sync.m16n8k8 TF32 instructions per clock per SM does a B200 issue, next to FFMA2?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe/mma_probe tools/probe/mma_probe.cu
//   timeout 60 tools/probe/mma_probe
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void mma_tf32(float (&d)[4], const unsigned (&a)[4], const unsigned (&b)[2])
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int NACC>
__global__ void mma_kernel(float *out, int iters, unsigned seed)
{
    float d[NACC][4];
    unsigned a[4], b[2];
    for (int i = 0; i < 4; ++i) a[i] = __float_as_uint(1.0f + (threadIdx.x + i + seed) * 1e-3f);
    for (int i = 0; i < 2; ++i) b[i] = __float_as_uint(0.5f + (threadIdx.x + i) * 1e-3f);
    for (int k = 0; k < NACC; ++k)
        for (int i = 0; i < 4; ++i) d[k][i] = 0.f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < NACC; ++k) mma_tf32(d[k], a, b);
    }
    float s = 0.f;
    for (int k = 0; k < NACC; ++k)
        for (int i = 0; i < 4; ++i) s += d[k][i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void ffma2_kernel(float *out, int iters, float seed)
{
    float2 acc[NACC];
    const float2 x = make_float2(1.0f + seed * threadIdx.x, 0.999f);
    const float2 y = make_float2(1e-6f, 2e-6f);
    for (int k = 0; k < NACC; ++k) acc[k] = make_float2(k, -k);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < NACC; ++k) {
            asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(*reinterpret_cast<unsigned long long *>(&acc[k]))
                         : "l"(*reinterpret_cast<const unsigned long long *>(&x)),
                           "l"(*reinterpret_cast<const unsigned long long *>(&y)));
        }
    }
    float s = 0.f;
    for (int k = 0; k < NACC; ++k) s += acc[k].x + acc[k].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F> float time_ms(F launch)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    launch();
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    launch();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, a, b);
    return ms;
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const int sms = p.multiProcessorCount;
    float *out;
    cudaMalloc(&out, sizeof(float) * sms * 8 * 1024);
    const int iters = 20000;
    constexpr int NACC = 8;
    for (int warps = 4; warps <= 16; warps *= 2) {
        const int threads = 32 * warps;
        float ms = time_ms([&] { mma_kernel<NACC><<<sms, threads>>>(out, iters, 1u); });
        double inst = double(sms) * warps * iters * NACC;
        double per_clk_sm = inst / (ms * 1e-3 * clk_khz * 1e3) / sms;
        printf("mma.sync m16n8k8 tf32: %2d warps/SM  %.3f ms  %.3f inst/clk/SM  = %.1f dense TF/s (at %.0f MHz nominal)\n",
               warps, ms, per_clk_sm, inst * 16 * 8 * 8 * 2 / (ms * 1e-3) / 1e12, clk_khz / 1e3);
        ms = time_ms([&] { ffma2_kernel<NACC><<<sms, threads>>>(out, iters, 1e-7f); });
        inst = double(sms) * warps * iters * NACC;
        per_clk_sm = inst / (ms * 1e-3 * clk_khz * 1e3) / sms;
        printf("fma.rn.f32x2         : %2d warps/SM  %.3f ms  %.3f inst/clk/SM  = %.1f TF/s fp32\n", warps, ms,
               per_clk_sm, inst * 32 * 2 * 2 / (ms * 1e-3) / 1e12);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
