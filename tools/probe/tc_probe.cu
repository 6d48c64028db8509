// tc_probe.cu -- stand-alone check of the tcgen05 building blocks the tensor-core degridder uses:
//   D[128 x 16] (TMEM, fp32) = A[128 x 56] (TMEM, written by tcgen05.st, one row per thread)
//                              x B[56 x 16] (shared memory, K-major core matrices, no swizzle)
// with kind::tf32 and the 3-term split  A_hi B_hi + A_lo B_hi + A_hi B_lo.
// Prints the maximum error against an fp64 CPU product (plain TF32 would be ~1e-3, the split
// ~1e-6) and the time per 128-row tile.  Build: nvcc -gencode arch=compute_100a,code=sm_100a.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

constexpr int KP = 56, NN = 16, KSTEPS = KP / 8;
constexpr int TM_COLS = 128;          // allocation: A_hi 56 + A_lo 56 + D 16
constexpr int COL_ALO = 56, COL_D = 112;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void tm_st8(uint32_t taddr, const float *v)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr),
                 "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
                 "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
                 "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7]))
                 : "memory");
}
__device__ __forceinline__ void tm_ld16(uint32_t taddr, float *v)
{
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr)
                 : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t bdesc, uint32_t idesc, uint32_t acc)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t"
        "}\n" ::"r"(d), "r"(a), "l"(bdesc), "r"(idesc), "r"(acc), "r"(0u)
        : "memory");
}
__device__ __forceinline__ void mma_ts_masked(uint32_t d, uint32_t a, uint64_t bdesc, uint32_t idesc, uint32_t acc,
                                              uint32_t m0, uint32_t m1, uint32_t m2, uint32_t m3)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, {%5, %6, %7, %8}, p;\n\t"
        "}\n" ::"r"(d), "r"(a), "l"(bdesc), "r"(idesc), "r"(acc), "r"(m0), "r"(m1), "r"(m2), "r"(m3)
        : "memory");
}
__device__ __forceinline__ bool mbar_try(uint64_t *bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}

// B element (n, k) in floats from the start of one operand copy; k-step blocks of 512 B:
// 8-row groups 256 B apart (SBO), the two 16-byte K chunks 128 B apart (LBO)
__host__ __device__ inline int b_off(int n, int k)
{
    const int ks = k >> 3, kk = k & 7;
    return ks * 128 + (n >> 3) * 64 + (kk >> 2) * 32 + (n & 7) * 4 + (kk & 3);
}

__global__ void __launch_bounds__(128) tc_probe_kernel(const float *__restrict__ A, const float *__restrict__ B,
                                                       float *__restrict__ D, int ntiles, int split)
{
    __shared__ uint32_t tm_base_s;
    __shared__ __align__(8) uint64_t bar;
    __shared__ __align__(128) float Bs[2][KSTEPS * 128];
    const int tid = threadIdx.x, warp = tid >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tm_base_s)), "r"(TM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0) {
        // split == 2: lane 0 of every warp issues a share of the MMAs and commits -> 4 arrivals
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(split == 2 ? 4 : 1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < KP * NN; i += 128) {
        const int k = i / NN, n = i % NN;
        const float g = B[k * NN + n];
        const float hi = __uint_as_float(__float_as_uint(g) & 0xffffe000u);
        Bs[0][b_off(n, k)] = hi;
        Bs[1][b_off(n, k)] = g - hi;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic smem writes -> tensor core reads
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tm_base_s;
    const uint32_t lane_base = tm + ((uint32_t)(warp * 32) << 16);
    const int lane = tid & 31;
    // instruction descriptor: D fp32, A / B tf32, both K-major, N = 16, M = 128
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((NN >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t dbase = (uint64_t(128 >> 4) << 16) | (uint64_t(256 >> 4) << 32) | (1ull << 46);
    uint32_t parity = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const float *arow = A + ((size_t)t * 128 + tid) * KP;
#pragma unroll
        for (int c0 = 0; c0 < KP; c0 += 8) {
            float hi[8], lo[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float p = arow[c0 + i];
                hi[i] = p;                  // the tensor core reads the top 19 bits
                lo[i] = p - __uint_as_float(__float_as_uint(p) & 0xffffe000u);
            }
            tm_st8(lane_base + c0, hi);
            tm_st8(lane_base + COL_ALO + c0, lo);
        }
        if (split == 2) {             // every MMA accumulates: D starts from zero
            const float zero[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            tm_st8(lane_base + COL_D, zero);
            tm_st8(lane_base + COL_D + 8, zero);
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (split == 2) {
            if (lane == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                for (int ks = warp; ks < KSTEPS; ks += 4) {
                    const uint64_t bh = dbase | (uint64_t)((smem_u32(&Bs[0][ks * 128]) >> 4) & 0x3fff);
                    const uint64_t bl = dbase | (uint64_t)((smem_u32(&Bs[1][ks * 128]) >> 4) & 0x3fff);
                    mma_ts(tm + COL_D, tm + ks * 8, bh, idesc, 1);
                    mma_ts(tm + COL_D, tm + COL_ALO + ks * 8, bh, idesc, 1);
                    mma_ts(tm + COL_D, tm + ks * 8, bl, idesc, 1);
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
            }
        } else if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int ks = 0; ks < KSTEPS; ++ks) {
                const uint64_t bh = dbase | (uint64_t)((smem_u32(&Bs[0][ks * 128]) >> 4) & 0x3fff);
                const uint64_t bl = dbase | (uint64_t)((smem_u32(&Bs[1][ks * 128]) >> 4) & 0x3fff);
                mma_ts(tm + COL_D, tm + ks * 8, bh, idesc, ks > 0);
                if (split) {
                    mma_ts(tm + COL_D, tm + COL_ALO + ks * 8, bh, idesc, 1);
                    mma_ts(tm + COL_D, tm + ks * 8, bl, idesc, 1);
                }
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        }
        {
            const long long t0 = clock64();
            while (!mbar_try(&bar, parity))
                if (clock64() - t0 > 2000000000LL) __trap();
        }
        parity ^= 1u;
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        float out[16];
        tm_ld16(lane_base + COL_D, out);
        float4 *dst = reinterpret_cast<float4 *>(D + ((size_t)t * 128 + tid) * NN);
#pragma unroll
        for (int i = 0; i < 4; ++i) dst[i] = make_float4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]);
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();      // every thread has read D before the next tile's first MMA overwrites it
    }
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(TM_COLS) : "memory");
}

// every warp runs its own pipeline on its 32 TMEM lanes: 32-row tiles, MMAs masked to the warp's
// lane quadrant, one mbarrier per warp, no block barrier inside the loop
__global__ void __launch_bounds__(128) tc_probe_warp_kernel(const float *__restrict__ A, const float *__restrict__ B,
                                                            float *__restrict__ D, int ntiles32)
{
    __shared__ uint32_t tm_base_s;
    __shared__ __align__(8) uint64_t bar[4];
    __shared__ __align__(128) float Bs[2][KSTEPS * 128];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tm_base_s)), "r"(TM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid < 4) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[tid])) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    for (int i = tid; i < KP * NN; i += 128) {
        const int k = i / NN, n = i % NN;
        const float g = B[k * NN + n];
        const float hi = __uint_as_float(__float_as_uint(g) & 0xffffe000u);
        Bs[0][b_off(n, k)] = hi;
        Bs[1][b_off(n, k)] = g - hi;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tm_base_s;
    const uint32_t lane_base = tm + ((uint32_t)(warp * 32) << 16);
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((NN >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t dbase = (uint64_t(128 >> 4) << 16) | (uint64_t(256 >> 4) << 32) | (1ull << 46);
    const uint32_t m0 = warp == 0 ? 0u : ~0u, m1 = warp == 1 ? 0u : ~0u, m2 = warp == 2 ? 0u : ~0u, m3 = warp == 3 ? 0u : ~0u;
    uint32_t parity = 0;
    for (int t = blockIdx.x * 4 + warp; t < ntiles32; t += gridDim.x * 4) {
        const float *arow = A + ((size_t)t * 32 + lane) * KP;
#pragma unroll
        for (int c0 = 0; c0 < KP; c0 += 8) {
            float hi[8], lo[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float p = arow[c0 + i];
                hi[i] = p;
                lo[i] = p - __uint_as_float(__float_as_uint(p) & 0xffffe000u);
            }
            tm_st8(lane_base + c0, hi);
            tm_st8(lane_base + COL_ALO + c0, lo);
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int ks = 0; ks < KSTEPS; ++ks) {
                const uint64_t bh = dbase | (uint64_t)((smem_u32(&Bs[0][ks * 128]) >> 4) & 0x3fff);
                const uint64_t bl = dbase | (uint64_t)((smem_u32(&Bs[1][ks * 128]) >> 4) & 0x3fff);
                mma_ts_masked(tm + COL_D, tm + ks * 8, bh, idesc, ks > 0, m0, m1, m2, m3);
                mma_ts_masked(tm + COL_D, tm + COL_ALO + ks * 8, bh, idesc, 1, m0, m1, m2, m3);
                mma_ts_masked(tm + COL_D, tm + ks * 8, bl, idesc, 1, m0, m1, m2, m3);
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar[warp])) : "memory");
        }
        __syncwarp();
        {
            const long long t0 = clock64();
            while (!mbar_try(&bar[warp], parity))
                if (clock64() - t0 > 2000000000LL) __trap();
        }
        parity ^= 1u;
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        float out[16];
        tm_ld16(lane_base + COL_D, out);
        float4 *dst = reinterpret_cast<float4 *>(D + ((size_t)t * 32 + lane) * NN);
#pragma unroll
        for (int i = 0; i < 4; ++i) dst[i] = make_float4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]);
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(TM_COLS) : "memory");
}

int main(int argc, char **argv)
{
    const int ntiles = argc > 1 ? atoi(argv[1]) : 148 * 64;
    std::vector<float> A((size_t)ntiles * 128 * KP), B(KP * NN), D((size_t)ntiles * 128 * NN);
    srand(1234);
    for (auto &x : A) x = (float)rand() / RAND_MAX * ((rand() & 1) ? 1.f : 1e-3f);
    for (auto &x : B) x = (float)rand() / RAND_MAX - 0.5f;
    for (int k = 49; k < KP; ++k)
        for (int n = 0; n < NN; ++n) B[k * NN + n] = 0.f;
    float *dA, *dB, *dD;
    CK(cudaMalloc(&dA, A.size() * 4));
    CK(cudaMalloc(&dB, B.size() * 4));
    CK(cudaMalloc(&dD, D.size() * 4));
    CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
    for (int split = 0; split < 4; ++split) {
        CK(cudaMemset(dD, 0, D.size() * 4));
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        if (split != 2) tc_probe_kernel<<<148 * 4, 128>>>(dA, dB, dD, ntiles, split == 3 ? 2 : split);
        else tc_probe_warp_kernel<<<148 * 4, 128>>>(dA, dB, dD, ntiles * 4);
        CK(cudaDeviceSynchronize());
        cudaEventRecord(e0);
        if (split != 2) tc_probe_kernel<<<148 * 4, 128>>>(dA, dB, dD, ntiles, split == 3 ? 2 : split);
        else tc_probe_warp_kernel<<<148 * 4, 128>>>(dA, dB, dD, ntiles * 4);
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
        double maxerr = 0, maxref = 0;
        for (int t = 0; t < ntiles; t += 37)
            for (int m = 0; m < 128; ++m)
                for (int n = 0; n < NN; ++n) {
                    double ref = 0;
                    for (int k = 0; k < KP; ++k) ref += (double)A[((size_t)t * 128 + m) * KP + k] * B[k * NN + n];
                    maxerr = fmax(maxerr, fabs(ref - D[((size_t)t * 128 + m) * NN + n]));
                    maxref = fmax(maxref, fabs(ref));
                }
        printf("mode=%d (0 plain TF32, 1 split, 2 split + per-warp masked pipelines, 3 split + MMA issue shared by the four warps): max abs err %.3e (max |ref| %.3e, rel %.3e); %.3f ms for %d tiles = %.1f ns per 128-row tile per SM-slot\n",
               split, maxerr, maxref, maxerr / maxref, ms, ntiles, ms * 1e6 / ntiles * 148 * 4 / 1.0 / 1.0);
        printf("   D[0][0..3] = %g %g %g %g\n", D[0], D[1], D[2], D[3]);
    }
    return 0;
}
