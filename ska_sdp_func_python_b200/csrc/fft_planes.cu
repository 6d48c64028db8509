// fft_planes.cu -- hand-written pruned 2-D FFT of one fp32 w-plane, fused with the w-screen.
//
// Replaces, for single-precision plans whose grid sides have a static line plan (fft_inplace.cuh:
// 1600, 3200, 6400, 12800), the per-plane  cuFFT C2C (four passes over the padded plane for
// 12800^2)  +  w-screen kernel  +  cropping  that sit between gridding and the dirty image inside
// what the reference runs as ducc0.wgridder.ms2dirty / dirty2ms
// (/root/reference/src/ska_sdp_func_python/imaging/ng.py:240-256, :99-112).  Two passes per plane,
// every line transformed once, in place, in shared memory:
//
//   invert (grid plane -> image):
//     rows : persistent blocks; TMA bulk loads (cp.async.bulk + mbarrier) bring row u+1 into the
//            second line buffer while row u is transformed along v; only the npix_y outputs that
//            fall on the image are kept and leave as 16-byte pieces (two adjacent image columns,
//            built by a lane-pair exchange) into T[npix_y/2][nu][2];
//     cols : one TMA bulk load per PAIR of image columns (its nu 16-byte pieces are contiguous in
//            T); the first stage splits the pieces into the two line buffers, each column is
//            transformed along u, the npix_x kept outputs are multiplied by the w-screen
//            exp(-2 pi i w (n-1)) in registers and their real parts accumulated into the
//            transposed image accumulator accT[npix_y][npix_x]; the next pair's load is issued as
//            soon as the last stage of the second column has read its buffer.
//   predict (image -> grid plane) is the adjoint:
//     cols : pre-corrected image column * exp(+2 pi i w (n-1)), zero-padded to nu (the screen and
//            the padding happen in the first stage's loads), both columns of a pair transformed,
//            each stored into T as the 8-byte halves of the pieces;
//     rows : TMA tensor loads gather row u+1's pieces while row u is transformed (the rows between
//            the two kept ranges read as zero in the first stage); whole rows written to the plane.
//
// Only the part of the grid that a footprint can reach takes part (PlaneGeo::ukeep, vkeep, from the
// plan's max |u|, |v|): the rows passes run over the rows u < ukeep or u >= nu - ukeep and load /
// store their two ends in v, the columns passes read / write those rows of T only, and only the
// four corners of a ring slot are cleared.  All other cells are identically zero on the way to the
// image and are never read by the degridder, so this is exact; a benchmark configuration
// (cellsize = 1 / (4 longest baseline)) keeps about half of each axis.
//
// Per plane this moves (2 ukeep)(2 vkeep)*8 + 2*(2 ukeep)*npix*8 + 2*npix^2*4 bytes instead of
// cuFFT's 8 passes over nu*nv*8 plus the screen kernel's cropped read.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "fft_inplace.cuh"
#include "kernels.h"

namespace b2g {

namespace {

// ---- PTX helpers (TMA, mbarrier) ----------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tma_load_3d(void *smem, const CUtensorMap *map, uint64_t *bar, int c0,
                                            int c1, int c2)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
                 " [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0),
                 "r"(c1), "r"(c2)
                 : "memory");
}
// contiguous global -> shared bulk copy (bytes: multiple of 16, both addresses 16-byte aligned)
__device__ __forceinline__ void bulk_load(void *smem, const void *gmem, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(gmem)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// A transfer that never completes (a bad descriptor) must not hang the device: trap after ~2 s.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity))
        if (clock64() - t0 > 4000000000LL) __trap();
}

// ---- w-screen -----------------------------------------------------------------------------
// (sin, cos) of 2 pi * sgn * w * (n - 1) at r2 = x^2 + y^2; false beyond the horizon.  n - 1 =
// -r2 / (1 + sqrt(1 - r2)) in fp64 from fp32 seeds of rsqrt and 1/x with one Newton step each
// (relative error ~1e-14: no fp64 sqrt or division on the per-plane path).
__device__ __forceinline__ bool screen_sincos(double r2, double w_signed, float &s, float &c)
{
    if (!(r2 < 1.0)) return false;
    const double z = 1.0 - r2;
    double y = (double)rsqrtf((float)z);
    y = y * fma(-0.5 * z, y * y, 1.5);
    const double d = fma(z, y, 1.0);              // 1 + sqrt(z)
    double r = (double)__frcp_rn((float)d);
    r = r * fma(-d, r, 2.0);
    double ph = w_signed * (-r2 * r);
    ph -= rint(ph);
    sincospif(2.0f * (float)ph, &s, &c);
    return true;
}

// Per-plane screen constants.  phase [cycles] = w (n - 1) = -w r2 / 2 - w q^2 / 2 with r2 = x^2 + y^2
// and q = r2 / (1 + sqrt(1 - r2)) = -(n - 1).  The first term is tens of cycles and linear in the
// INTEGERS (a - hx)^2, (b - hy)^2: its fractional part comes exactly out of a 64-bit fixed-point
// product; the second is a fraction of a cycle and fp32 is enough for it.  `fast` is set per plane
// when that holds (|w| q_max^2 / 2 < 1/2 cycle); otherwise the fp64 evaluation above is used.
struct Screen {
    double w;                      // signed: -w for invert, +w for predict
    unsigned long long cx, cy;     // frac(-w psx^2 / 2), frac(-w psy^2 / 2) as 0.64 fixed point
    float k3;                      // -w / 2
    float psx, psy;
    int fast;
};

__device__ __forceinline__ float frac_fixed(unsigned long long c, int d)
{
    // frac(c * d^2) as a signed fraction in [-1/2, 1/2)
    const unsigned long long p = c * (unsigned long long)((unsigned)(d * d));
    return (float)(int)(unsigned)(p >> 32) * 2.3283064365386963e-10f;
}

// (sin, cos) of 2 pi * phase at pixel offset ax = a - hx of the line with fy = frac_fixed(cy, b - hy),
// y2f = y^2 (fp32), y2d = y^2 (fp64); false beyond the horizon
// FAST: 0 = fp64, 1 = fixed point + fp32, 2 = the same for small fields (r2 <= 0.01 everywhere: no
// horizon test, and q = r2 / (1 + sqrt(1 - r2)) from its series -- relative error 7 r2^4 / 128 < 1e-9
// -- instead of a reciprocal square root and a division: the screen is as many instructions of the
// columns passes as the transform itself)
template <int FAST>
__device__ __forceinline__ bool screen_eval(const Screen &S, double psx, int ax, float fy, float y2f,
                                            double y2d, float &s, float &c)
{
    if (FAST) {
        const float xf = (float)ax * S.psx;
        const float r2 = fmaf(xf, xf, y2f);
        float q;
        if (FAST == 2) {
            q = r2 * fmaf(r2, fmaf(r2, fmaf(r2, 0.0390625f, 0.0625f), 0.125f), 0.5f);
        } else {
            if (!(r2 < 1.f)) return false;
            const float z = 1.f - r2;
            q = __fdividef(r2, fmaf(z, rsqrtf(z), 1.f));      // r2 / (1 + sqrt(1 - r2))
        }
        float ph = frac_fixed(S.cx, ax) + fy + S.k3 * q * q;
        ph -= rintf(ph);
        __sincosf(6.283185307179586f * ph, &s, &c);
        return true;
    }
    const double x = (double)ax * psx;
    return screen_sincos(fma(x, x, y2d), S.w, s, c);
}

struct PlaneGeo {
    int nu, nv, npix_x, npix_y, hx, hy;
    double psx, psy;
    int wstack;
    // rows (u) of the grid that can hold samples: u < ukeep or u >= nu - ukeep (nu / 2: all).  The
    // others are identically zero when a plane is transformed to the image and are never read by
    // the degridder, so the rows passes leave them out
    int ukeep;
    // ... and the same along v: columns v < vkeep or v >= nv - vkeep (even; nv / 2: all).  The rows
    // passes neither load nor store the others
    int vkeep;
    __device__ __forceinline__ bool vkept(int v) const { return v < vkeep || v >= nv - vkeep; }
    __device__ __forceinline__ int nlines() const { return 2 * ukeep; }
    __device__ __forceinline__ int row_of(int line) const { return line < ukeep ? line : nu - 2 * ukeep + line; }
    __device__ __forceinline__ bool kept(int u) const { return u < ukeep || u >= nu - ukeep; }
};

__device__ __forceinline__ int image_index(int k, int n, int npix, int h)
{
    // FFT index k in [0, n) -> image pixel, or -1 for the lines the image does not hold
    return k < npix - h ? k + h : (k >= n - h ? k - (n - h) : -1);
}

// ---- shared-memory carve-up ---------------------------------------------------------------------
// [X0 | X1 | twiddle tables | two mbarriers]; line buffers rounded to 128 B so that both are valid
// TMA bases
template <class P> struct SmemLayout {
    static constexpr int NPADA = (P::NPAD + 15) & ~15;
    static constexpr size_t X_BYTES = size_t(NPADA) * 8;
    static constexpr size_t TW_OFF = 2 * X_BYTES;
    static constexpr size_t BAR_OFF = TW_OFF + ((size_t(P::TW_ENTRIES) * 8 + 15) & ~size_t(15));
    static constexpr size_t BYTES = BAR_OFF + 16;
    static_assert(2 * X_BYTES >= size_t(P::N) * 16, "a pair block of 16-byte pieces must fit the two buffers");
};

template <class P> constexpr int min_blocks() { return P::T >= 320 ? 1 : (P::T >= 160 ? 2 : 4); }

template <class P>
__device__ __forceinline__ void load_tables(float2 *tws, const float2 *__restrict__ tw, int tid)
{
    for (int i = tid; i < P::TW_ENTRIES; i += P::THREADS) tws[i] = tw[i];
}

// barrier between the phases of the line transform, with a hook after a chosen call
template <class Hook> struct HookSync {
    int calls, at;
    Hook hook;
    __device__ __forceinline__ void operator()()
    {
        fence_proxy_async();     // generic accesses of the buffers ordered before later TMA writes
        __syncthreads();
        if (++calls == at) hook();
    }
};
struct NoHook {
    __device__ __forceinline__ void operator()() const {}
};

// ---- invert, rows: plane[u][:] -> T[:, u, :] (kept columns) ----------------------------------
// tw: tables of the +1 transform along v.  The last stage's outputs stay in registers: lanes
// 2m / 2m+1 hold adjacent columns, one exchange builds the 16-byte piece (two adjacent image
// columns of row u) that goes to T[pair][u] with one vector store -- no staging buffer, so the line
// buffer is free (and refilled by TMA) as soon as the last stage has read it.
template <class P>
__global__ void __launch_bounds__(P::THREADS, min_blocks<P>())
hfft_rows_inv_kernel(PlaneGeo g, const float2 *__restrict__ tw, const float2 *plane, float4 *__restrict__ T4,
                     float2 *clear_plane)
{
    extern __shared__ __align__(128) unsigned char raw[];
    using L = SmemLayout<P>;
    using LS = hfft::LastStage<P>;
    static_assert(LS::C == 1 && LS::NS % 2 == 0 && LS::R % 2 == 0, "pair exchange layout");
    float2 *X0 = reinterpret_cast<float2 *>(raw);
    float2 *tws = reinterpret_cast<float2 *>(raw + L::TW_OFF);
    uint64_t *bar = reinterpret_cast<uint64_t *>(raw + L::BAR_OFF);
    const int tid = threadIdx.x;
    load_tables<P>(tws, tw, tid);
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_init(bar + 1, 1);
        mbar_fence_init();
    }
    // clear_plane (= plane, or null): every row is zeroed in global memory once it has arrived in
    // shared memory, so the ring slot is ready for the gridder's next slab without a separate
    // memset pass over it (the stores ride in this kernel's unused DRAM bandwidth)
    static_assert(P::N % (2 * P::T) == 0, "whole 16-byte stores per thread");
    __syncthreads();
    constexpr int nv = P::N;
    const int hy = g.hy, half_pairs = hy / 2;
    const bool active = tid < P::T;
    const int nlines = g.nlines();
    // a row arrives as its two ends (the columns that can hold samples); stage 0 reads the rest as 0
    const int vkeep = g.vkeep;
    const uint32_t end_bytes = (uint32_t)vkeep * 8u;
    auto fetch = [&](int line, float2 *buf, uint64_t *mb) {
        const float2 *src = plane + (size_t)g.row_of(line) * nv;
        mbar_expect_tx(mb, 2 * end_bytes);
        bulk_load(buf, src, end_bytes, mb);
        bulk_load(buf + (nv - vkeep), src + (nv - vkeep), end_bytes, mb);
    };
    if (tid == 0) {
        for (int b = 0; b < 2; ++b) {
            const int line = blockIdx.x + b * gridDim.x;
            if (line < nlines) fetch(line, X0 + b * L::NPADA, bar + b);
        }
    }
    int it = 0;
    for (int line = blockIdx.x; line < nlines; line += gridDim.x, ++it) {
        const int u = g.row_of(line);
        const int b = it & 1;
        float2 *X = X0 + b * L::NPADA;
        mbar_wait(bar + b, (uint32_t)(it >> 1) & 1u);
        if (clear_plane && active) {
            // measured: bulk (TMA) stores from a block of zeros here cost 1.2 ms per plane -- they
            // queue in front of the next rows' bulk loads; plain vector stores do not
            float4 *zrow = reinterpret_cast<float4 *>(clear_plane + (size_t)u * nv);
#pragma unroll
            for (int i = 0; i < P::N / (2 * P::T); ++i) zrow[tid + i * P::T] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        // linear, as the bulk loads left it
        auto rd0 = [X, vkeep](int i) { return (i < vkeep || i >= nv - vkeep) ? X[i] : make_float2(0.f, 0.f); };
        const int ln = line + 2 * gridDim.x;
        auto hook = [&]() {
            // the last stage has read X: refill it with the row after next
            if (tid == 0 && ln < nlines) fetch(ln, X, bar + b);
        };
        HookSync<NoHook> sync0{0, -1, NoHook()};
        hfft::line_stage0<P, 1>(tid, X, tws, rd0, sync0);
        HookSync<decltype(hook)> sync{0, 2 * P::NSTAGE - 3, hook};
        float2 v[1][LS::R];
        hfft::line_rest_open<P, 1>(tid, X, tws, v, sync);
        if (active) hfft::stage_compute<P, 1, LS::R, 1, LS::NS>(tid, v, tws + LS::TW_OFF);
        // v[0][r] is column k = tid + r * NS of row u
        const bool odd = tid & 1;
        // this lane's pieces start at columns kb = kbase + r NS (r = 0, 2, ...).  Image columns
        // b >= hy come from k = b - hy in [0, hy), b < hy from k = nv - hy + b, so the piece goes to
        // pair (kb + hy) / 2 of the low range or (kb - (nv - hy)) / 2 of the high one: two running
        // pointers that advance by NS pairs per step (the 20 pair indices are loop invariants that
        // the compiler otherwise keeps in registers across the line -- and spills: beside 227 KB of
        // shared memory a spill reload comes from L2)
        const int kbase = (tid & ~1) + (odd ? LS::NS : 0);
        const ptrdiff_t step = (ptrdiff_t)LS::NS * g.nu;
        float4 *plo = T4 + (ptrdiff_t)(half_pairs + (kbase >> 1)) * g.nu + u;
        float4 *phi = T4 + (ptrdiff_t)((kbase - (nv - hy)) >> 1) * g.nu + u;
#pragma unroll
        for (int r = 0; r < LS::R; r += 2) {
            const float2 send = odd ? v[0][r] : v[0][r + 1];
            float2 recv;
            recv.x = __shfl_xor_sync(0xffffffffu, send.x, 1);
            recv.y = __shfl_xor_sync(0xffffffffu, send.y, 1);
            const int kb = kbase + r * LS::NS;                    // first column of this lane's piece
            const float4 piece = odd ? make_float4(recv.x, recv.y, v[0][r + 1].x, v[0][r + 1].y)
                                     : make_float4(v[0][r].x, v[0][r].y, recv.x, recv.y);
            if (active && kb < hy) *plo = piece;
            else if (active && kb >= nv - hy) *phi = piece;
            plo += step;
            phi += step;
        }
    }
}

__device__ __forceinline__ void red_add_v4(float *addr, float a, float b, float c, float d)
{
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

// ---- invert, columns: T[b/2][:][b&1] -> accT[b][:] += Re(F * screen) ----------------------------
// The kept outputs are multiplied by the w-screen in registers; a 4 x 4 lane transpose gives every
// lane four consecutive image pixels, which are added to the accumulator with one vector reduction
// (no read-modify-write round trip; every pixel receives exactly one addition per plane, so the
// result does not depend on any ordering).
template <class P, int FAST>
__global__ void __launch_bounds__(P::THREADS, min_blocks<P>())
hfft_cols_inv_kernel(PlaneGeo g, const float2 *__restrict__ tw, const float2 *__restrict__ T, Screen S,
                     float *__restrict__ accT)
{
    extern __shared__ __align__(128) unsigned char raw[];
    using L = SmemLayout<P>;
    using LS = hfft::LastStage<P>;
    static_assert(LS::C == 1 && LS::NS % 4 == 0 && LS::R % 4 == 0, "quad exchange layout");
    float2 *X0 = reinterpret_cast<float2 *>(raw);
    float2 *X1 = X0 + L::NPADA;
    const float4 *P4 = reinterpret_cast<const float4 *>(raw);      // the pair block as loaded
    float2 *tws = reinterpret_cast<float2 *>(raw + L::TW_OFF);
    uint64_t *bar = reinterpret_cast<uint64_t *>(raw + L::BAR_OFF);
    const int tid = threadIdx.x;
    load_tables<P>(tws, tw, tid);
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    constexpr int nu = P::N;
    const int npx = g.npix_x, hx = g.hx, npairs = g.npix_y / 2;
    const bool active = tid < P::T;
    // only the rows of T that the rows pass wrote (the kept rows of the grid) are loaded: the two
    // ends of the pair block; the first stage reads the others as zero
    const int ukeep = g.ukeep;
    const uint32_t end_bytes = (uint32_t)ukeep * 16u;
    auto fetch = [&](int pr) {
        const float2 *src = T + (size_t)pr * nu * 2;
        mbar_expect_tx(bar, 2 * end_bytes);
        bulk_load(raw, src, end_bytes, bar);
        bulk_load(raw + (size_t)(nu - ukeep) * 16, src + (size_t)(nu - ukeep) * 2, end_bytes, bar);
    };
    if (tid == 0 && (int)blockIdx.x < npairs) fetch(blockIdx.x);
    int it = 0;
    for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
        mbar_wait(bar, (uint32_t)it & 1u);
        {
            // stage 0 of both columns: the 16-byte pieces split into the two line buffers
            float2 v0[P::C0][P::R0], v1[P::C0][P::R0];
            constexpr int q = P::N / P::R0;
            if (active) {
#pragma unroll
                for (int c = 0; c < P::C0; ++c)
#pragma unroll
                    for (int t = 0; t < P::R0; ++t) {
                        const int idx = tid + c * P::T + t * q;
                        const float4 p = g.kept(idx) ? P4[idx] : make_float4(0.f, 0.f, 0.f, 0.f);
                        v0[c][t] = make_float2(p.x, p.y);
                        v1[c][t] = make_float2(p.z, p.w);
                    }
            }
            __syncthreads();
            if (active) {
                hfft::stage0_write_padded<P, 1>(tid, v0, tws, X0);
                hfft::stage0_write_padded<P, 1>(tid, v1, tws, X1);
            }
            __syncthreads();
        }
        const int pn = pair + gridDim.x;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
            const int b = 2 * pair + c;
            float2 *X = c ? X1 : X0;
            // after the last stage of the SECOND column has read its buffer both are free: the
            // next pair's load runs under that stage's arithmetic and the epilogue
            auto hook = [=]() {
                if (c == 1 && tid == 0 && pn < npairs) fetch(pn);
            };
            HookSync<decltype(hook)> sync{0, 2 * P::NSTAGE - 3, hook};
            float2 v[1][LS::R];
            hfft::line_rest_open<P, 1>(tid, X, tws, v, sync);
            if (active) hfft::stage_compute<P, 1, LS::R, 1, LS::NS>(tid, v, tws + LS::TW_OFF);
            // v[0][r] is image row index k = tid + r * NS of column b: screen, keep the real part
            const double y = (double)(b - g.hy) * g.psy;
            const double y2 = y * y;
            const float y2f = (float)y2, fy = frac_fixed(S.cy, b - g.hy);
            float val[LS::R];
#pragma unroll
            for (int r = 0; r < LS::R; ++r) {
                val[r] = 0.f;
                const int a = image_index(tid + r * LS::NS, nu, npx, hx);
                if (active && a >= 0) {
                    if (g.wstack) {
                        float s, co;
                        if (screen_eval<FAST>(S, g.psx, a - hx, fy, y2f, y2, s, co)) val[r] = v[0][r].x * co - v[0][r].y * s;
                    } else {
                        val[r] = v[0][r].x;
                    }
                }
            }
            float *dst = accT + (size_t)b * npx;
            const int l4 = tid & 3;
            const bool hi2 = l4 & 2, hi1 = l4 & 1;
#pragma unroll
            for (int r = 0; r < LS::R; r += 4) {
                // 4 x 4 transpose over lanes (l4) x outputs (r .. r+3)
                const float s0 = hi2 ? val[r] : val[r + 2], s1 = hi2 ? val[r + 1] : val[r + 3];
                const float t0 = __shfl_xor_sync(0xffffffffu, s0, 2), t1 = __shfl_xor_sync(0xffffffffu, s1, 2);
                const float a0 = hi2 ? t0 : val[r], a1 = hi2 ? t1 : val[r + 1];
                const float b0 = hi2 ? val[r + 2] : t0, b1 = hi2 ? val[r + 3] : t1;
                const float u0 = hi1 ? a0 : a1, u1 = hi1 ? b0 : b1;
                const float w0 = __shfl_xor_sync(0xffffffffu, u0, 1), w1 = __shfl_xor_sync(0xffffffffu, u1, 1);
                // this lane now holds output r + l4 of the four lanes 4m .. 4m+3
                const float f0 = hi1 ? w0 : a0, f1 = hi1 ? a1 : w0, f2 = hi1 ? w1 : b0, f3 = hi1 ? b1 : w1;
                const int kb = (tid & ~3) + (r + l4) * LS::NS;
                const int a = image_index(kb, nu, npx, hx);
                if (active && a >= 0) red_add_v4(dst + a, f0, f1, f2, f3);
            }
        }
    }
}

// ---- predict, columns: preT[b][:] * screen, zero-padded -> T[b/2][:][b&1] -------------------------
template <class P, int FAST>
__global__ void __launch_bounds__(P::THREADS, min_blocks<P>())
hfft_cols_fwd_kernel(PlaneGeo g, const float2 *__restrict__ tw, const float *__restrict__ preT, Screen S,
                     float2 *__restrict__ T)
{
    extern __shared__ __align__(128) unsigned char raw[];
    using L = SmemLayout<P>;
    float2 *X0 = reinterpret_cast<float2 *>(raw);
    float2 *X1 = X0 + L::NPADA;
    float2 *tws = reinterpret_cast<float2 *>(raw + L::TW_OFF);
    const int tid = threadIdx.x;
    load_tables<P>(tws, tw, tid);
    __syncthreads();
    constexpr int nu = P::N;
    const int npx = g.npix_x, hx = g.hx, npairs = g.npix_y / 2;
    const bool active = tid < P::T;
    for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x) {
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
            const int b = 2 * pair + c;
            float2 *X = c ? X1 : X0;
            const float *src = preT + (size_t)b * npx;
            const double y = (double)(b - g.hy) * g.psy;
            const double y2 = y * y;
            const float y2f = (float)y2, fy = frac_fixed(S.cy, b - g.hy);
            // the last stage stores its outputs straight into T: 8-byte halves of the 16-byte
            // pieces, consecutive lanes consecutive rows (the other half follows with the pair's
            // second column and meets it in L2); only the rows the rows pass goes on to transform
            float2 *tcol = T + (size_t)pair * nu * 2 + c;
            auto wr = [tcol, &g](int k, float2 v) {
                if (g.kept(k)) tcol[2 * k] = v;
            };
            HookSync<NoHook> sync{0, -1, NoHook()};
            {
                // stage 0: all of the thread's image pixels are requested before the first screen
                // is evaluated (40 independent loads in flight), zero-padding by predication
                float2 v[P::C0][P::R0];
                constexpr int q = P::N / P::R0;
                if (active) {
                    float in[P::C0][P::R0];
#pragma unroll
                    for (int cc = 0; cc < P::C0; ++cc)
#pragma unroll
                        for (int t = 0; t < P::R0; ++t) {
                            const int a = image_index(tid + cc * P::T + t * q, nu, npx, hx);
                            in[cc][t] = a >= 0 ? __ldg(src + a) : 0.f;
                        }
#pragma unroll
                    for (int cc = 0; cc < P::C0; ++cc)
#pragma unroll
                        for (int t = 0; t < P::R0; ++t) {
                            const float val = in[cc][t];
                            float2 z = make_float2(val, 0.f);
                            if (g.wstack && val != 0.f) {
                                const int a = image_index(tid + cc * P::T + t * q, nu, npx, hx);
                                float s, co;
                                z = screen_eval<FAST>(S, g.psx, a - hx, fy, y2f, y2, s, co) ? make_float2(val * co, val * s)
                                                                                       : make_float2(0.f, 0.f);
                            }
                            v[cc][t] = z;
                        }
                }
                sync();
                if (active) hfft::stage0_write_padded<P, -1>(tid, v, tws, X);
                sync();
            }
            hfft::line_rest<P, -1>(tid, X, tws, wr, sync);
        }
        __syncthreads();     // the second column's last stage has read X1 before the next pair's stage 0 writes
    }
}

// ---- predict, rows: T[:, u, :] (kept columns, zero elsewhere) -> plane[u][:] ---------------------
template <class P>
__global__ void __launch_bounds__(P::THREADS, min_blocks<P>())
hfft_rows_fwd_kernel(PlaneGeo g, const float2 *__restrict__ tw, const __grid_constant__ CUtensorMap tmap,
                     int pb, float2 *__restrict__ plane)
{
    extern __shared__ __align__(128) unsigned char raw[];
    using L = SmemLayout<P>;
    float2 *X0 = reinterpret_cast<float2 *>(raw);
    float2 *tws = reinterpret_cast<float2 *>(raw + L::TW_OFF);
    uint64_t *bar = reinterpret_cast<uint64_t *>(raw + L::BAR_OFF);
    const int tid = threadIdx.x;
    load_tables<P>(tws, tw, tid);
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_init(bar + 1, 1);
        mbar_fence_init();
    }
    __syncthreads();
    constexpr int nv = P::N;
    const int hy = g.hy, half_pairs = hy / 2;
    const uint32_t row_bytes = (uint32_t)g.npix_y * 8u;
    // gather the kept columns of row u (16-byte pieces at a stride of nu * 16 bytes) into buf
    auto fetch = [&](int u, float2 *buf, uint64_t *mb) {
        mbar_expect_tx(mb, row_bytes);
        for (int q = 0; q < half_pairs; q += pb) {
            tma_load_3d(buf + 2 * q, &tmap, mb, 0, u, half_pairs + q);
            tma_load_3d(buf + (nv - hy) + 2 * q, &tmap, mb, 0, u, q);
        }
    };
    const int nlines = g.nlines();
    if (tid == 0) {
        for (int b = 0; b < 2; ++b) {
            const int line = blockIdx.x + b * gridDim.x;
            if (line < nlines) fetch(g.row_of(line), X0 + b * L::NPADA, bar + b);
        }
    }
    int it = 0;
    for (int line = blockIdx.x; line < nlines; line += gridDim.x, ++it) {
        const int u = g.row_of(line);
        const int b = it & 1;
        float2 *X = X0 + b * L::NPADA;
        mbar_wait(bar + b, (uint32_t)(it >> 1) & 1u);
        float2 *row = plane + (size_t)u * nv;
        auto rd0 = [X, hy](int i) {
            return (i < hy || i >= nv - hy) ? X[i] : make_float2(0.f, 0.f);
        };
        const int vkeep = g.vkeep;
        auto wr = [row, vkeep](int k, float2 v) {
            if (k < vkeep || k >= nv - vkeep) row[k] = v;     // only where the degridder can read
        };
        const int ln = line + 2 * gridDim.x;
        // X is free once the last stage has read it: prefetch the row after next under that
        // stage's arithmetic and stores
        auto hook = [&]() {
            if (tid == 0 && ln < nlines) fetch(g.row_of(ln), X, bar + b);
        };
        HookSync<decltype(hook)> sync{0, 2 * P::NSTAGE - 1, hook};
        hfft::line_inplace<P, -1>(tid, X, tws, rd0, wr, sync);
    }
}
// ---- image <-> transposed accumulator ----------------------------------------------------------
__device__ __forceinline__ double nm1_exact(double r2) { return -r2 / (sqrt(1.0 - r2) + 1.0); }

__device__ __forceinline__ double corr_at(const ImgGeo &g, int a, int b)
{
    double f = __ldg(g.cu + a) * __ldg(g.cv + b);
    if (g.wstack) {
        const double x = (double)(a - g.npix_x / 2) * g.pixsize_x;
        const double y = (double)(b - g.npix_y / 2) * g.pixsize_y;
        const double r2 = x * x + y * y;
        if (r2 >= 1.0) return 0.0;
        const double nm1 = nm1_exact(r2);
        const double z = (g.nm1min < 0.0) ? fma(-2.0 / g.nm1min, nm1, 1.0) : 0.0;
        double b1 = 0.0, b2 = 0.0;   // Clenshaw, as image_kernels.cu
        for (int j = g.ncheb - 1; j >= 1; --j) {
            const double b0 = fma(2.0 * z, b1, __ldg(g.cheb + j) - b2);
            b2 = b1;
            b1 = b0;
        }
        f *= fma(z, b1, __ldg(g.cheb) - b2) / (nm1 + 1.0);
    }
    return f;
}

// corrT[b][a] = corr(a, b): the gridding correction of the plan, tabulated once (fp64 square root,
// division and a ~20-term Chebyshev sum per pixel cost 4 ms per 8192^2 image; the table makes
// correct / precorrect plain HBM-bound multiplies)
__global__ void __launch_bounds__(256)
corr_table_kernel(ImgGeo g, float *__restrict__ corrT)
{
    const int a = blockIdx.x * 256 + threadIdx.x, b = blockIdx.y;
    if (a < g.npix_x) corrT[(size_t)b * g.npix_x + a] = (float)corr_at(g, a, b);
}

// out[a][b] = accT[b][a] * corr(a, b)   (32 x 32 tiles through shared memory)
template <typename Tout>
__global__ void __launch_bounds__(256)
correct_t_kernel(ImgGeo g, const float *__restrict__ accT, const float *__restrict__ corrT,
                 Tout *__restrict__ out)
{
    __shared__ float tile[32][33];
    const int a0 = blockIdx.y * 32, b0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) {
        const int b = b0 + r, a = a0 + tx;
        const size_t i = (size_t)b * g.npix_x + a;
        tile[r][tx] = (a < g.npix_x && b < g.npix_y) ? accT[i] * __ldg(corrT + i) : 0.f;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int a = a0 + r, b = b0 + tx;
        if (a < g.npix_x && b < g.npix_y) out[(size_t)a * g.npix_y + b] = (Tout)tile[tx][r];
    }
}

// preT[b][a] = dirty[a][b] * corr(a, b)
template <typename Tin>
__global__ void __launch_bounds__(256)
precorrect_t_kernel(ImgGeo g, const Tin *__restrict__ dirty, const float *__restrict__ corrT,
                    float *__restrict__ preT)
{
    __shared__ float tile[32][33];
    const int a0 = blockIdx.y * 32, b0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) {
        const int a = a0 + r, b = b0 + tx;
        tile[r][tx] = (a < g.npix_x && b < g.npix_y) ? (float)dirty[(size_t)a * g.npix_y + b] : 0.f;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int b = b0 + r, a = a0 + tx;
        const size_t i = (size_t)b * g.npix_x + a;
        if (a < g.npix_x && b < g.npix_y) preT[i] = tile[tx][r] * __ldg(corrT + i);
    }
}


typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                  const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn()
{
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// ---- per-plan launch tables ----------------------------------------------------------------------
struct LineOps {
    int n = 0, threads = 0, tw_entries = 0;
    size_t smem = 0;
    void (*build_tw)(int sign, float2 *tw) = nullptr;
    const void *rows_inv = nullptr, *cols_inv = nullptr, *cols_fwd = nullptr, *rows_fwd = nullptr;
    const void *cols_inv_fast = nullptr, *cols_fwd_fast = nullptr;   // fixed-point / fp32 screen
    const void *cols_inv_small = nullptr, *cols_fwd_small = nullptr; // ... of a small field
};

template <class P> LineOps make_ops()
{
    LineOps o;
    o.n = P::N;
    o.threads = P::THREADS;
    o.tw_entries = P::TW_ENTRIES;
    o.smem = SmemLayout<P>::BYTES;
    o.build_tw = &hfft::build_twiddles<P>;
    o.rows_inv = reinterpret_cast<const void *>(&hfft_rows_inv_kernel<P>);
    o.cols_inv = reinterpret_cast<const void *>(&hfft_cols_inv_kernel<P, 0>);
    o.cols_fwd = reinterpret_cast<const void *>(&hfft_cols_fwd_kernel<P, 0>);
    o.cols_inv_fast = reinterpret_cast<const void *>(&hfft_cols_inv_kernel<P, 1>);
    o.cols_fwd_fast = reinterpret_cast<const void *>(&hfft_cols_fwd_kernel<P, 1>);
    o.cols_inv_small = reinterpret_cast<const void *>(&hfft_cols_inv_kernel<P, 2>);
    o.cols_fwd_small = reinterpret_cast<const void *>(&hfft_cols_fwd_kernel<P, 2>);
    o.rows_fwd = reinterpret_cast<const void *>(&hfft_rows_fwd_kernel<P>);
    return o;
}

bool ops_for(int n, LineOps *o)
{
    switch (n) {
    case 1600: *o = make_ops<hfft::Plan1600>(); return true;
    case 3200: *o = make_ops<hfft::Plan3200>(); return true;
    case 6400: *o = make_ops<hfft::Plan6400>(); return true;
    case 12800: *o = make_ops<hfft::Plan12800>(); return true;
    default: return false;
    }
}

}  // namespace

struct HandFft {
    PlaneGeo g;
    LineOps ou, ov;               // lines along u (columns pass), along v (rows pass)
    float2 *d_tw = nullptr;       // [u+ | u- | v+ | v-]
    size_t off_up = 0, off_um = 0, off_vp = 0, off_vm = 0;
    float2 *d_T = nullptr;        // [npix_y/2][nu][2]
    float *d_corr = nullptr;      // [npix_y][npix_x] gridding correction, built by the first use
    CUtensorMap tmap;
    int pb = 0;
    int grid_rows_inv = 0, grid_rows_fwd = 0, grid_cols_inv = 0, grid_cols_fwd = 0;
};

int hfft_create(HandFft **out, const ImgGeo &ig, int ukeep, int vkeep, cudaStream_t s)
{
    *out = nullptr;
    const int nu = ig.nu, nv = ig.nv, npx = ig.npix_x, npy = ig.npix_y;
    const int hx = npx / 2, hy = npy / 2;
    if (std::getenv("B2G_NO_HFFT")) return -1;
    if (npy % 4 != 0 || npx % 2 != 0 || npx > nu || npy > nv) return -1;
    if ((nv - hy) % 16 != 0 || hy % 16 != 0) return -1;   // TMA box bases: 128-byte aligned
    LineOps ou, ov;
    if (!ops_for(nu, &ou) || !ops_for(nv, &ov)) return -1;
    int pb = 0;
    for (int d = 256; d >= 8; d -= 8)
        if ((hy / 2) % d == 0) {
            pb = d;
            break;
        }
    if (pb == 0) return -1;
    EncodeTiledFn enc = encode_fn();
    if (!enc) return -1;
    int dev = 0, max_smem = 0, nsm = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
    if ((int)ou.smem > max_smem || (int)ov.smem > max_smem) return -1;

    HandFft *h = new HandFft();
    if (ukeep < 1 || 2 * ukeep > nu || std::getenv("B2G_NO_ROW_PRUNING")) ukeep = nu / 2;
    vkeep = (vkeep + 1) & ~1;     // whole 16-byte pieces at both ends of a row
    if (vkeep < 2 || 2 * vkeep > nv || std::getenv("B2G_NO_ROW_PRUNING")) vkeep = nv / 2;
    h->g = PlaneGeo{nu, nv, npx, npy, hx, hy, ig.pixsize_x, ig.pixsize_y, ig.wstack, ukeep, vkeep};
    h->ou = ou;
    h->ov = ov;
    h->pb = pb;
    h->off_up = 0;
    h->off_um = ou.tw_entries;
    h->off_vp = 2 * size_t(ou.tw_entries);
    h->off_vm = 2 * size_t(ou.tw_entries) + ov.tw_entries;
    std::vector<float2> tw(2 * size_t(ou.tw_entries) + 2 * size_t(ov.tw_entries));
    ou.build_tw(+1, tw.data() + h->off_up);
    ou.build_tw(-1, tw.data() + h->off_um);
    ov.build_tw(+1, tw.data() + h->off_vp);
    ov.build_tw(-1, tw.data() + h->off_vm);
    cudaError_t e = cudaMallocAsync(&h->d_tw, tw.size() * sizeof(float2), s);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(h->d_tw, tw.data(), tw.size() * sizeof(float2), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMallocAsync(&h->d_T, size_t(npy / 2) * size_t(nu) * 16, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);   // tw (host vector) consumed
    if (e != cudaSuccess) {
        hfft_destroy(h, s);
        return (int)e;
    }
    // T as a rank-3 fp32 tensor: [npix_y/2][nu][4 floats]; boxes of pb column pairs of one row
    const cuuint64_t dims[3] = {4, (cuuint64_t)nu, (cuuint64_t)(npy / 2)};
    const cuuint64_t strides[2] = {16, (cuuint64_t)nu * 16};
    const cuuint32_t box[3] = {4, 1, (cuuint32_t)pb};
    const cuuint32_t estr[3] = {1, 1, 1};
    if (enc(&h->tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, h->d_T, dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        hfft_destroy(h, s);
        return -1;
    }
    // persistent grids: as many blocks as fit the device at once
    struct K {
        const void *fn;
        const LineOps *o;
        int *grid;
        int lines;
    } ks[8] = {{ov.rows_inv, &h->ov, &h->grid_rows_inv, 2 * ukeep},
               {ov.rows_fwd, &h->ov, &h->grid_rows_fwd, 2 * ukeep},
               {ou.cols_inv, &h->ou, &h->grid_cols_inv, npy / 2},
               {ou.cols_fwd, &h->ou, &h->grid_cols_fwd, npy / 2},
               {ou.cols_inv_fast, &h->ou, &h->grid_cols_inv, npy / 2},
               {ou.cols_fwd_fast, &h->ou, &h->grid_cols_fwd, npy / 2},
               {ou.cols_inv_small, &h->ou, &h->grid_cols_inv, npy / 2},
               {ou.cols_fwd_small, &h->ou, &h->grid_cols_fwd, npy / 2}};
    for (auto &k : ks) {
        int per_sm = 0;
        if (cudaFuncSetAttribute(k.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)k.o->smem) != cudaSuccess ||
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k.fn, k.o->threads, k.o->smem) != cudaSuccess ||
            per_sm < 1) {
            cudaGetLastError();
            hfft_destroy(h, s);
            return -1;
        }
        const long resident = (long)per_sm * nsm;
        *k.grid = (int)(k.lines < resident ? k.lines : resident);
    }
    *out = h;
    return 0;
}

void hfft_destroy(HandFft *h, cudaStream_t s)
{
    if (!h) return;
    if (h->d_tw) cudaFreeAsync(h->d_tw, s);
    if (h->d_T) cudaFreeAsync(h->d_T, s);
    if (h->d_corr) cudaFreeAsync(h->d_corr, s);
    delete h;
}

int hfft_ukeep(const HandFft *h) { return h ? h->g.ukeep : 0; }
int hfft_vkeep(const HandFft *h) { return h ? h->g.vkeep : 0; }

void hfft_describe(const HandFft *h, char *buf, int n)
{
    snprintf(buf, n, "u:%d (%d threads, %zu B smem, grids %d/%d) v:%d (%d threads, %zu B smem, grids %d/%d) box=%d pairs",
             h->ou.n, h->ou.threads, h->ou.smem, h->grid_cols_inv, h->grid_cols_fwd, h->ov.n, h->ov.threads,
             h->ov.smem, h->grid_rows_inv, h->grid_rows_fwd, h->pb);
}

namespace {
// the per-plane screen constants for signed w (see struct Screen)
Screen make_screen(const HandFft *h, double w_signed)
{
    const PlaneGeo &g = h->g;
    Screen S;
    S.w = w_signed;
    auto fixed = [](double v) {
        double f = v - std::floor(v);                      // [0, 1)
        if (!(f >= 0.0) || f >= 1.0) f = 0.0;
        return (unsigned long long)std::ldexp(f, 64 - 11) << 11;   // 53 significant bits
    };
    S.cx = fixed(-0.5 * w_signed * g.psx * g.psx);
    S.cy = fixed(-0.5 * w_signed * g.psy * g.psy);
    S.k3 = (float)(-0.5 * w_signed);
    S.psx = (float)g.psx;
    S.psy = (float)g.psy;
    // the fp32 remainder term must stay a fraction of a cycle (its relative error is ~1e-6), and
    // the squared pixel offsets must fit 32 bits
    const double xm = g.psx * g.hx, ym = g.psy * g.hy, r2 = xm * xm + ym * ym;
    const double q = r2 < 1.0 ? r2 / (1.0 + std::sqrt(1.0 - r2)) : 1.0;
    S.fast = (r2 < 0.5 && 0.5 * std::fabs(w_signed) * q * q < 0.5 && g.hx < 46000 && g.hy < 46000 &&
              !std::getenv("B2G_SCREEN_FP64")) ? 1 : 0;
    // small field: the series of q holds to 1e-9 (B2G_SCREEN_GENERAL=1 keeps the general fp32 path)
    if (S.fast && r2 <= 0.01 && !std::getenv("B2G_SCREEN_GENERAL")) S.fast = 2;
    return S;
}
}  // namespace

int hfft_invert_plane(HandFft *h, cplx<float> *plane, double w, float *accT, bool clear, cudaStream_t s)
{
    PlaneGeo g = h->g;
    Screen S = make_screen(h, -w);
    const float2 *twv = h->d_tw + h->off_vp, *twu = h->d_tw + h->off_up;
    const float2 *pl = reinterpret_cast<const float2 *>(plane);
    const float2 *T = h->d_T;
    float4 *T4 = reinterpret_cast<float4 *>(h->d_T);
    float2 *clr = clear ? reinterpret_cast<float2 *>(plane) : nullptr;
    void *a_rows[] = {&g, &twv, &pl, &T4, &clr};
    cudaError_t e = cudaLaunchKernel(h->ov.rows_inv, dim3(h->grid_rows_inv), dim3(h->ov.threads), a_rows,
                                     h->ov.smem, s);
    if (e != cudaSuccess) return (int)e;
    void *a_cols[] = {&g, &twu, &T, &S, &accT};
    e = cudaLaunchKernel(S.fast == 2 ? h->ou.cols_inv_small : (S.fast ? h->ou.cols_inv_fast : h->ou.cols_inv),
                         dim3(h->grid_cols_inv),
                         dim3(h->ou.threads), a_cols, h->ou.smem, s);
    return (int)e;
}

int hfft_predict_plane(HandFft *h, const float *preT, double w, cplx<float> *plane, cudaStream_t s)
{
    PlaneGeo g = h->g;
    Screen S = make_screen(h, w);
    const float2 *twv = h->d_tw + h->off_vm, *twu = h->d_tw + h->off_um;
    float2 *pl = reinterpret_cast<float2 *>(plane);
    float2 *T = h->d_T;
    int pb = h->pb;
    void *a_cols[] = {&g, &twu, &preT, &S, &T};
    cudaError_t e = cudaLaunchKernel(S.fast == 2 ? h->ou.cols_fwd_small : (S.fast ? h->ou.cols_fwd_fast : h->ou.cols_fwd),
                                     dim3(h->grid_cols_fwd),
                                     dim3(h->ou.threads), a_cols, h->ou.smem, s);
    if (e != cudaSuccess) return (int)e;
    void *a_rows[] = {&g, &twv, &h->tmap, &pb, &pl};
    e = cudaLaunchKernel(h->ov.rows_fwd, dim3(h->grid_rows_fwd), dim3(h->ov.threads), a_rows, h->ov.smem, s);
    return (int)e;
}

namespace {
int ensure_corr_table(HandFft *h, const ImgGeo &g, cudaStream_t s)
{
    if (h->d_corr) return 0;
    cudaError_t e = cudaMallocAsync(&h->d_corr, size_t(g.npix_x) * size_t(g.npix_y) * sizeof(float), s);
    if (e != cudaSuccess) return (int)e;
    dim3 grid((g.npix_x + 255) / 256, g.npix_y);
    corr_table_kernel<<<grid, 256, 0, s>>>(g, h->d_corr);
    return (int)cudaGetLastError();
}
}  // namespace

template <typename Tout>
int hfft_correct(HandFft *h, const ImgGeo &g, const float *accT, Tout *out, cudaStream_t s)
{
    if (int rc = ensure_corr_table(h, g, s)) return rc;
    dim3 grid((g.npix_y + 31) / 32, (g.npix_x + 31) / 32);
    correct_t_kernel<Tout><<<grid, 256, 0, s>>>(g, accT, h->d_corr, out);
    return (int)cudaGetLastError();
}
template <typename Tin>
int hfft_precorrect(HandFft *h, const ImgGeo &g, const Tin *dirty, float *preT, cudaStream_t s)
{
    if (int rc = ensure_corr_table(h, g, s)) return rc;
    dim3 grid((g.npix_y + 31) / 32, (g.npix_x + 31) / 32);
    precorrect_t_kernel<Tin><<<grid, 256, 0, s>>>(g, dirty, h->d_corr, preT);
    return (int)cudaGetLastError();
}
template int hfft_correct<float>(HandFft *, const ImgGeo &, const float *, float *, cudaStream_t);
template int hfft_correct<double>(HandFft *, const ImgGeo &, const float *, double *, cudaStream_t);
template int hfft_precorrect<float>(HandFft *, const ImgGeo &, const float *, float *, cudaStream_t);
template int hfft_precorrect<double>(HandFft *, const ImgGeo &, const double *, float *, cudaStream_t);

}  // namespace b2g
