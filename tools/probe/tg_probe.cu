// tg_probe.cu -- stand-alone check of the tcgen05 building blocks of the tensor-core GRIDDER:
//   D[(a,b)][(c,re/im)] (TMEM, fp32) += sum over the samples s of a tile
//                                        P[s][(a,b)] * V[s][(c,re/im)]
// i.e. the samples are the contraction index.  Both operands come from shared memory ("SS" form),
// K-major (K = samples), no swizzle, core matrices of 8 rows x 16 bytes:
//   element (row, s) at  (s / 8) * KSTEP + ((s / 4) & 1) * LBO + (row / 8) * SBO + (row % 8) * 16 + (s % 4) * 4
// with LBO = 144 bytes (not 128: thread s of a warp then writes its value of one row to bank group
// (s / 4) -- conflict free), KSTEP = 2 LBO, SBO = (tile / 8) * KSTEP.
// The 3 x TF32 split stacks the hi and lo parts of P along M: rows 0..55 = hi, 56..111 = lo, so
//   MMA 1: [P_hi; P_lo] x V_hi,  MMA 2: [P_hi; P_lo] x V_lo      (M = 128, N = 16, K = 8 each)
// and the product is D[m] + D[m + 56] (hi*hi + hi*lo + lo*hi + lo*lo).
// Prints the maximum error against fp64 and the time per 128-sample tile.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 tools/probe/tg_probe.cu -o /tmp/tg_probe
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

constexpr int KP = 56, NN = 16, TILE = 128;
constexpr int LBO = 144, KSTEP = 2 * LBO, SBO = (TILE / 8) * KSTEP;      // bytes
constexpr int A_BYTES = 16 * SBO, B_BYTES = 2 * SBO;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

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
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
__device__ __forceinline__ bool mbar_try(uint64_t *bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}

// tiles_per_run tiles accumulate into one D; then D rows are written out
__global__ void __launch_bounds__(128) tg_probe_kernel(const float *__restrict__ P, const float *__restrict__ V,
                                                       float *__restrict__ D, int nruns, int tiles_per_run)
{
    extern __shared__ __align__(128) unsigned char raw[];
    unsigned char *As = raw, *Bh = raw + A_BYTES, *Bl = Bh + B_BYTES;
    uint64_t *bar = reinterpret_cast<uint64_t *>(Bl + B_BYTES);
    uint32_t *tm_base_s = reinterpret_cast<uint32_t *>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tm_base_s)), "r"(32) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 4;" ::"r"(smem_u32(bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < (A_BYTES + 2 * B_BYTES) / 4; i += 128) reinterpret_cast<float *>(raw)[i] = 0.f;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = *tm_base_s;
    const uint32_t lane_base = tm + ((uint32_t)(warp * 32) << 16);
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((NN >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t dbase = (uint64_t(LBO >> 4) << 16) | (uint64_t(SBO >> 4) << 32) | (1ull << 46);
    const uint32_t my = (uint32_t)((tid >> 3) * KSTEP + ((tid >> 2) & 1) * LBO + (tid & 3) * 4);   // this sample's column
    uint32_t parity = 0;
    for (int run = blockIdx.x; run < nruns; run += gridDim.x) {
        for (int t = 0; t < tiles_per_run; ++t) {
            const size_t s = ((size_t)run * tiles_per_run + t) * TILE + tid;
            // one coalesced 16-byte seed per sample; the kernel values and the weighted sample are
            // derived from it in registers (the gridder computes them, it does not load them)
            const float4 seed = reinterpret_cast<const float4 *>(P)[s];
            float ku[7], kv[7], vrow[NN];
#pragma unroll
            for (int i = 0; i < 7; ++i) {
                ku[i] = seed.x + 0.1f * (float)i * seed.y;
                kv[i] = seed.y - 0.05f * (float)i * seed.x;
            }
#pragma unroll
            for (int n = 0; n < NN; ++n) vrow[n] = n < 14 ? seed.z * (0.1f * (float)(n + 1)) - seed.w : 0.f;
#pragma unroll
            for (int m = 0; m < KP; ++m) {
                const float p = m < 49 ? ku[m / 7] * kv[m % 7] : 0.f;
                const float h = tf32_hi(p);
                const int mh = m, ml = m + KP;
                *reinterpret_cast<float *>(As + my + (mh >> 3) * SBO + (mh & 7) * 16) = h;
                *reinterpret_cast<float *>(As + my + (ml >> 3) * SBO + (ml & 7) * 16) = p - h;
            }
#pragma unroll
            for (int n = 0; n < NN; ++n) {
                const float v = vrow[n];
                const float h = tf32_hi(v);
                *reinterpret_cast<float *>(Bh + my + (n >> 3) * SBO + (n & 7) * 16) = h;
                *reinterpret_cast<float *>(Bl + my + (n >> 3) * SBO + (n & 7) * 16) = v - h;
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncthreads();
            if (lane == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t ad = dbase | (uint64_t)((smem_u32(As) >> 4) & 0x3fff);
                const uint64_t bh = dbase | (uint64_t)((smem_u32(Bh) >> 4) & 0x3fff);
                const uint64_t bl = dbase | (uint64_t)((smem_u32(Bl) >> 4) & 0x3fff);
#pragma unroll
                for (int k0 = 0; k0 < TILE / 8; k0 += 4) {
                    const uint64_t off = (uint64_t)((k0 + warp) * (KSTEP >> 4));
                    // the first tile of a run starts the accumulator: its first K-steps (one per
                    // warp, in no particular order) cannot all be "overwrite", so D is zeroed by
                    // the epilogue of the previous run instead (and by the initial clear)
                    mma_ss(tm, ad + off, bh + off, idesc, 1);
                    mma_ss(tm, ad + off, bl + off, idesc, 1);
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
            }
            {
                const long long t0 = clock64();
                while (!mbar_try(bar, parity))
                    if (clock64() - t0 > 2000000000LL) __trap();
            }
            parity ^= 1u;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        float out[16];
        tm_ld16(lane_base, out);
        float4 *dst = reinterpret_cast<float4 *>(D + ((size_t)run * 128 + tid) * NN);
#pragma unroll
        for (int i = 0; i < 4; ++i) dst[i] = make_float4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]);
        // zero this thread's accumulator row for the next run
        asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(lane_base), "r"(0u) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
    }
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(32) : "memory");
}

int main(int argc, char **argv)
{
    const int nruns = argc > 1 ? atoi(argv[1]) : 148 * 16;
    const int tpr = argc > 2 ? atoi(argv[2]) : 3;
    const size_t ns = (size_t)nruns * tpr * TILE;
    std::vector<float> P(ns * 4), V(4), D((size_t)nruns * 128 * NN);
    srand(1234);
    for (size_t i = 0; i < ns * 4; ++i) P[i] = (float)rand() / RAND_MAX - 0.3f;
    float *dP, *dV, *dD;
    CK(cudaMalloc(&dP, P.size() * 4));
    CK(cudaMalloc(&dV, V.size() * 4));
    CK(cudaMalloc(&dD, D.size() * 4));
    CK(cudaMemcpy(dP, P.data(), P.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dV, V.data(), V.size() * 4, cudaMemcpyHostToDevice));
    const int smem = A_BYTES + 2 * B_BYTES + 64;
    CK(cudaFuncSetAttribute(tg_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    tg_probe_kernel<<<148 * 2, 128, smem>>>(dP, dV, dD, nruns, tpr);
    CK(cudaDeviceSynchronize());
    cudaEventRecord(e0);
    tg_probe_kernel<<<148 * 2, 128, smem>>>(dP, dV, dD, nruns, tpr);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
    double maxerr = 0, maxref = 0;
    for (int run = 0; run < nruns; run += 53)
        for (int m = 0; m < 49; ++m)
            for (int n = 0; n < 14; ++n) {
                double ref = 0;
                for (int s = 0; s < tpr * TILE; ++s) {
                    const size_t g = (size_t)run * tpr * TILE + s;
                    const float x = P[4 * g], y = P[4 * g + 1], z = P[4 * g + 2], w = P[4 * g + 3];
                    const float kua = x + 0.1f * (float)(m / 7) * y, kvb = y - 0.05f * (float)(m % 7) * x;
                    const float val = z * (0.1f * (float)(n + 1)) - w;
                    ref += (double)(kua * kvb) * val;
                }
                const double got = (double)D[((size_t)run * 128 + m) * NN + n] + D[((size_t)run * 128 + m + KP) * NN + n];
                maxerr = fmax(maxerr, fabs(ref - got));
                maxref = fmax(maxref, fabs(ref));
            }
    printf("SS gridding contraction, hi/lo stacked along M, LBO %d SBO %d: max abs err %.3e (max |ref| %.3e, rel %.3e); %.3f ms for %d tiles = %.1f ns per 128-sample tile per resident block (2 per SM)\n",
           LBO, SBO, maxerr, maxref, maxerr / maxref, ms, nruns * tpr, ms * 1e6 / (nruns * tpr) * 148 * 2);
    printf("   D[0][0..3] = %g %g %g %g\n", D[0], D[1], D[2], D[3]);
    return 0;
}
