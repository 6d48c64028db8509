// fft_planes.cu -- hand-written pruned 2-D FFT of one fp32 w-plane, fused with the w-screen.
//
// Replaces, for single-precision plans whose grid lines fit shared memory (nu, nv <= 14000), the
// per-plane  cuFFT C2C (four passes over the padded plane for 12800^2)  +  w-screen kernel  +
// cropping  that sit between gridding and the dirty image inside what the reference runs as
// ducc0.wgridder.ms2dirty / dirty2ms (/root/reference/src/ska_sdp_func_python/imaging/ng.py:240-256,
// :99-112).  Two passes per plane, each line transformed once in shared memory (fft_lines.cuh):
//
//   invert (grid plane -> image):
//     rows : FFT along v of every grid row u; only the npix_y outputs that fall on the image are
//            kept and go to the intermediate T[npix_y/2][nu][2] (pairs of image columns, so that a
//            row's outputs are 16-byte pieces at a stride of nu*16 bytes) by TMA tensor stores
//            (cp.async.bulk.tensor.3d) straight out of the line buffer;
//     cols : FFT along u of every kept column; each of the npix_x kept outputs is multiplied by the
//            w-screen exp(-2 pi i w (n-1)) in registers and its real part is accumulated into the
//            transposed image accumulator accT[npix_y][npix_x].
//   predict (image -> grid plane) is the adjoint:
//     cols : pre-corrected image column * exp(+2 pi i w (n-1)), zero-padded to nu, FFT along u,
//            written to T;
//     rows : TMA tensor loads gather a row's 16-byte pieces into the (zero-padded) line buffer,
//            FFT along v, whole row written to the plane.
//
// Per plane this moves nu*nv*8 + 2*nu*npix*8 + 2*npix^2*4 bytes instead of cuFFT's 8 passes over
// nu*nv*8 plus the screen kernel's cropped read.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "fft_lines.cuh"
#include "kernels.h"

namespace b2g {

using hfft::LinePlan;

namespace {

constexpr int kThreads = 512;

// ---- PTX helpers (TMA, mbarrier) ----------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap *map, const void *smem, int c0, int c1,
                                             int c2)
{
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(smem)), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void tma_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void tma_load_3d(void *smem, const CUtensorMap *map, uint64_t *bar, int c0,
                                            int c1, int c2)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
                 " [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0),
                 "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
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

struct PlaneGeo {
    int nu, nv, npix_x, npix_y, hx, hy;
    double psx, psy;
    int wstack;
};

__device__ __forceinline__ int image_index(int k, int n, int npix, int h)
{
    // FFT index k in [0, n) -> image pixel, or -1 for the lines the image does not hold
    return k < npix - h ? k + h : (k >= n - h ? k - (n - h) : -1);
}

struct Smem {
    float2 *a, *b, *t0, *t1;
    uint64_t *bar;
};
__device__ __forceinline__ Smem carve(unsigned char *raw, int nline, int n_hi)
{
    Smem s;
    s.a = reinterpret_cast<float2 *>(raw);
    s.b = s.a + nline;
    s.t0 = s.b + nline;
    s.t1 = s.t0 + hfft::kTwLow;
    s.bar = reinterpret_cast<uint64_t *>(s.t1 + ((n_hi + 1) & ~1));
    return s;
}
inline size_t smem_bytes(const LinePlan &p)
{
    // line buffers padded to 128 B so that both are valid TMA box bases
    const size_t nline = (size_t(p.n) + 15) & ~size_t(15);
    return 2 * nline * 8 + hfft::kTwLow * 8 + ((p.n_hi + 1) & ~1) * 8 + 16;
}

struct SyncThreads {
    __device__ __forceinline__ void operator()() const { __syncthreads(); }
};

// ---- invert, rows: plane[u][:] -> T[:, u, :] (kept columns) ----------------------------------
// tw: {t0, t1} of the +1 transform along v.  Boxes of `pb` column pairs.
__global__ void __launch_bounds__(kThreads, 1)
hfft_rows_inv_kernel(PlaneGeo g, LinePlan lp, const float2 *__restrict__ tw,
                     const float2 *__restrict__ plane, const __grid_constant__ CUtensorMap tmap, int pb)
{
    extern __shared__ __align__(128) unsigned char raw[];
    const int nline = (lp.n + 15) & ~15;
    Smem sm = carve(raw, nline, lp.n_hi);
    const int tid = threadIdx.x;
    for (int i = tid; i < hfft::kTwLow; i += kThreads) sm.t0[i] = tw[i];
    for (int i = tid; i < lp.n_hi; i += kThreads) sm.t1[i] = tw[hfft::kTwLow + i];
    __syncthreads();
    float2 *first = sm.a, *second = sm.b;
    const int nv = g.nv, hy = g.hy, half_pairs = hy / 2;
    bool pending = false;      // a TMA store may still be reading the buffer it was issued from
    for (int u = blockIdx.x; u < g.nu; u += gridDim.x) {
        const float2 *row = plane + (size_t)u * nv;
        // the last stage writes buffer `outb` (stage s writes `first` if s is even)
        float2 *outb = ((lp.nstage - 1) & 1) ? second : first;
        auto load = [row](int i) { return __ldg(row + i); };
        auto store = [outb](int k, float2 v) { outb[k] = v; };
        int calls = 0;
        auto sync = [&]() {
            // stage 0 wrote `first`, which the previous row's store never reads (see the swap
            // below); every later stage may write the buffer that store is still reading
            if (calls++ == 0 && pending) {
                if (tid == 0) tma_wait_read();
                pending = false;
            }
            __syncthreads();
        };
        if (lp.nstage == 1 && pending) {
            if (tid == 0) tma_wait_read();
            pending = false;
            __syncthreads();
        }
        hfft::fft_line<1>(lp, first, second, sm.t0, sm.t1, tid, kThreads, load, store, sync);
        fence_proxy_async();         // the line written by generic stores -> visible to the TMA
        __syncthreads();
        if (tid == 0) {
            // image columns b >= hy come from k = b - hy in [0, hy); b < hy from k = nv - hy + b
            for (int q = 0; q < half_pairs; q += pb) {
                tma_store_3d(&tmap, outb + 2 * q, 0, u, half_pairs + q);
                tma_store_3d(&tmap, outb + (nv - hy) + 2 * q, 0, u, q);
            }
            tma_commit();
        }
        pending = true;
        // next row: its stage 0 must not write `outb`
        if (outb == first) {
            float2 *t = first;
            first = second;
            second = t;
        }
    }
    if (tid == 0) tma_wait_all();
}

// ---- invert, columns: T[b/2][:][b&1] -> accT[b][:] (+)= Re(F * screen) --------------------------
__global__ void __launch_bounds__(kThreads, 1)
hfft_cols_inv_kernel(PlaneGeo g, LinePlan lp, const float2 *__restrict__ tw,
                     const float2 *__restrict__ T, double w, float *__restrict__ accT)
{
    extern __shared__ __align__(128) unsigned char raw[];
    const int nline = (lp.n + 15) & ~15;
    Smem sm = carve(raw, nline, lp.n_hi);
    const int tid = threadIdx.x;
    for (int i = tid; i < hfft::kTwLow; i += kThreads) sm.t0[i] = tw[i];
    for (int i = tid; i < lp.n_hi; i += kThreads) sm.t1[i] = tw[hfft::kTwLow + i];
    __syncthreads();
    const int nu = g.nu, npx = g.npix_x, hx = g.hx;
    for (int pair = blockIdx.x; pair < g.npix_y / 2; pair += gridDim.x) {
        for (int c = 0; c < 2; ++c) {
            const int b = 2 * pair + c;
            const float2 *src = T + (size_t)pair * nu * 2 + c;
            const double y = (double)(b - g.hy) * g.psy;
            const double y2 = y * y;
            float *dst = accT + (size_t)b * npx;
            auto load = [src](int i) { return __ldg(src + 2 * i); };
            auto store = [=](int k, float2 v) {
                const int a = image_index(k, nu, npx, hx);
                if (a < 0) return;
                float val = v.x;
                if (g.wstack) {
                    const double x = (double)(a - hx) * g.psx;
                    float s, co;
                    if (!screen_sincos(fma(x, x, y2), -w, s, co)) return;
                    val = v.x * co - v.y * s;
                }
                dst[a] += val;
            };
            hfft::fft_line<1>(lp, sm.a, sm.b, sm.t0, sm.t1, tid, kThreads, load, store, SyncThreads());
            __syncthreads();
        }
    }
}

// ---- predict, columns: preT[b][:] * screen, zero-padded -> T[b/2][:][b&1] -------------------------
__global__ void __launch_bounds__(kThreads, 1)
hfft_cols_fwd_kernel(PlaneGeo g, LinePlan lp, const float2 *__restrict__ tw,
                     const float *__restrict__ preT, double w, float2 *__restrict__ T)
{
    extern __shared__ __align__(128) unsigned char raw[];
    const int nline = (lp.n + 15) & ~15;
    Smem sm = carve(raw, nline, lp.n_hi);
    const int tid = threadIdx.x;
    for (int i = tid; i < hfft::kTwLow; i += kThreads) sm.t0[i] = tw[i];
    for (int i = tid; i < lp.n_hi; i += kThreads) sm.t1[i] = tw[hfft::kTwLow + i];
    __syncthreads();
    const int nu = g.nu, npx = g.npix_x, hx = g.hx;
    for (int pair = blockIdx.x; pair < g.npix_y / 2; pair += gridDim.x) {
        for (int c = 0; c < 2; ++c) {
            const int b = 2 * pair + c;
            const float *src = preT + (size_t)b * npx;
            float2 *dst = T + (size_t)pair * nu * 2 + c;
            const double y = (double)(b - g.hy) * g.psy;
            const double y2 = y * y;
            auto load = [=](int i) {
                const int a = image_index(i, nu, npx, hx);
                if (a < 0) return make_float2(0.f, 0.f);
                const float val = __ldg(src + a);
                if (val == 0.f) return make_float2(0.f, 0.f);
                if (!g.wstack) return make_float2(val, 0.f);
                const double x = (double)(a - hx) * g.psx;
                float s, co;
                if (!screen_sincos(fma(x, x, y2), w, s, co)) return make_float2(0.f, 0.f);
                return make_float2(val * co, val * s);
            };
            auto store = [dst](int k, float2 v) { dst[2 * k] = v; };
            hfft::fft_line<-1>(lp, sm.a, sm.b, sm.t0, sm.t1, tid, kThreads, load, store, SyncThreads());
            __syncthreads();
        }
    }
}

// ---- predict, rows: T[:, u, :] (kept columns, zero elsewhere) -> plane[u][:] ---------------------
__global__ void __launch_bounds__(kThreads, 1)
hfft_rows_fwd_kernel(PlaneGeo g, LinePlan lp, const float2 *__restrict__ tw,
                     const __grid_constant__ CUtensorMap tmap, int pb, float2 *__restrict__ plane)
{
    extern __shared__ __align__(128) unsigned char raw[];
    const int nline = (lp.n + 15) & ~15;
    Smem sm = carve(raw, nline, lp.n_hi);
    const int tid = threadIdx.x;
    for (int i = tid; i < hfft::kTwLow; i += kThreads) sm.t0[i] = tw[i];
    for (int i = tid; i < lp.n_hi; i += kThreads) sm.t1[i] = tw[hfft::kTwLow + i];
    if (tid == 0) mbar_init(sm.bar, 1);
    __syncthreads();
    const int nv = g.nv, hy = g.hy, half_pairs = hy / 2;
    const uint32_t row_bytes = (uint32_t)g.npix_y * 8u;
    // `inb` receives the row (TMA + zero fill); stage 0 reads it and writes `first` = the other one
    float2 *inb = sm.b, *first = sm.a;
    uint32_t parity = 0;
    auto fetch = [&](int u, float2 *buf) {
        // zero the lines between the two kept ranges (generic stores, ordered by the barriers of
        // the FFT stages / the one below), then gather the kept ones
        for (int i = hy + tid; i < nv - hy; i += kThreads) buf[i] = make_float2(0.f, 0.f);
        if (tid == 0) {
            mbar_expect_tx(sm.bar, row_bytes);
            for (int q = 0; q < half_pairs; q += pb) {
                tma_load_3d(buf + 2 * q, &tmap, sm.bar, 0, u, half_pairs + q);
                tma_load_3d(buf + (nv - hy) + 2 * q, &tmap, sm.bar, 0, u, q);
            }
        }
    };
    int u = blockIdx.x;
    if (u < g.nu) fetch(u, inb);
    for (; u < g.nu; u += gridDim.x) {
        mbar_wait(sm.bar, parity);
        parity ^= 1u;
        __syncthreads();             // zero fill by all threads complete
        float2 *row = plane + (size_t)u * nv;
        const float2 *in = inb;
        float2 *second = inb;
        auto load = [in](int i) { return in[i]; };
        auto store = [row](int k, float2 v) { row[k] = v; };
        // the buffer the last stage does not read is free from the barrier before it: prefetch
        // the next row into it while the last stage runs
        const int un = u + gridDim.x;
        float2 *lastread = lp.nstage == 1 ? inb : (((lp.nstage - 2) & 1) ? second : first);
        float2 *target = lastread == first ? second : first;
        int calls = 0;
        auto sync = [&]() {
            fence_proxy_async();     // generic accesses to `target` ordered before the TMA writes
            __syncthreads();
            if (++calls == lp.nstage - 1 && un < g.nu) fetch(un, target);
        };
        if (lp.nstage == 1 && un < g.nu) fetch(un, target);
        hfft::fft_line<-1>(lp, first, second, sm.t0, sm.t1, tid, kThreads, load, store, sync);
        __syncthreads();
        // next row: input sits in `target`
        inb = target;
        first = (target == sm.a) ? sm.b : sm.a;
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

// out[a][b] = accT[b][a] * corr(a, b)   (32 x 32 tiles through shared memory)
template <typename Tout>
__global__ void __launch_bounds__(256)
correct_t_kernel(ImgGeo g, const float *__restrict__ accT, Tout *__restrict__ out)
{
    __shared__ float tile[32][33];
    const int a0 = blockIdx.y * 32, b0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) {
        const int b = b0 + r, a = a0 + tx;
        tile[r][tx] = (a < g.npix_x && b < g.npix_y) ? accT[(size_t)b * g.npix_x + a] : 0.f;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int a = a0 + r, b = b0 + tx;
        if (a < g.npix_x && b < g.npix_y)
            out[(size_t)a * g.npix_y + b] = (Tout)((double)tile[tx][r] * corr_at(g, a, b));
    }
}

// preT[b][a] = dirty[a][b] * corr(a, b)
template <typename Tin>
__global__ void __launch_bounds__(256)
precorrect_t_kernel(ImgGeo g, const Tin *__restrict__ dirty, float *__restrict__ preT)
{
    __shared__ float tile[32][33];
    const int a0 = blockIdx.y * 32, b0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) {
        const int a = a0 + r, b = b0 + tx;
        tile[r][tx] = (a < g.npix_x && b < g.npix_y)
                          ? (float)((double)dirty[(size_t)a * g.npix_y + b] * corr_at(g, a, b))
                          : 0.f;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int b = b0 + r, a = a0 + tx;
        if (a < g.npix_x && b < g.npix_y) preT[(size_t)b * g.npix_x + a] = tile[tx][r];
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

}  // namespace

struct HandFft {
    PlaneGeo g;
    LinePlan lu, lv;
    float2 *d_tw = nullptr;       // [u+ | u- | v+ | v-], each kTwLow + n_hi entries
    size_t off_up = 0, off_um = 0, off_vp = 0, off_vm = 0;
    float2 *d_T = nullptr;        // [npix_y/2][nu][2]
    CUtensorMap tmap;
    int pb = 0;
    int nsm = 148;
    size_t smem_u = 0, smem_v = 0;
};

int hfft_create(HandFft **out, const ImgGeo &ig, cudaStream_t s)
{
    *out = nullptr;
    const int nu = ig.nu, nv = ig.nv, npx = ig.npix_x, npy = ig.npix_y;
    const int hx = npx / 2, hy = npy / 2;
    if (std::getenv("B2G_NO_HFFT")) return -1;
    if (npy % 4 != 0 || npx % 2 != 0 || npx > nu || npy > nv) return -1;
    if ((nv - hy) % 16 != 0 || hy % 16 != 0) return -1;   // TMA box bases: 128-byte aligned
    LinePlan lu, lv;
    if (hfft::plan_line(nu, &lu) || hfft::plan_line(nv, &lv)) return -1;
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
    const size_t su = smem_bytes(lu), sv = smem_bytes(lv);
    if ((int)su > max_smem || (int)sv > max_smem) return -1;

    HandFft *h = new HandFft();
    h->g = PlaneGeo{nu, nv, npx, npy, hx, hy, ig.pixsize_x, ig.pixsize_y, ig.wstack};
    h->lu = lu;
    h->lv = lv;
    h->pb = pb;
    h->nsm = nsm;
    h->smem_u = su;
    h->smem_v = sv;
    const size_t nu_tw = hfft::kTwLow + lu.n_hi, nv_tw = hfft::kTwLow + lv.n_hi;
    h->off_up = 0;
    h->off_um = nu_tw;
    h->off_vp = 2 * nu_tw;
    h->off_vm = 2 * nu_tw + nv_tw;
    std::vector<float2> tw(2 * nu_tw + 2 * nv_tw);
    hfft::build_twiddles(lu, +1, tw.data() + h->off_up, tw.data() + h->off_up + hfft::kTwLow);
    hfft::build_twiddles(lu, -1, tw.data() + h->off_um, tw.data() + h->off_um + hfft::kTwLow);
    hfft::build_twiddles(lv, +1, tw.data() + h->off_vp, tw.data() + h->off_vp + hfft::kTwLow);
    hfft::build_twiddles(lv, -1, tw.data() + h->off_vm, tw.data() + h->off_vm + hfft::kTwLow);
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
    cudaFuncSetAttribute(hfft_rows_inv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sv);
    cudaFuncSetAttribute(hfft_rows_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sv);
    cudaFuncSetAttribute(hfft_cols_inv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)su);
    cudaFuncSetAttribute(hfft_cols_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)su);
    if (cudaGetLastError() != cudaSuccess) {
        hfft_destroy(h, s);
        return -1;
    }
    *out = h;
    return 0;
}

void hfft_destroy(HandFft *h, cudaStream_t s)
{
    if (!h) return;
    if (h->d_tw) cudaFreeAsync(h->d_tw, s);
    if (h->d_T) cudaFreeAsync(h->d_T, s);
    delete h;
}

void hfft_describe(const HandFft *h, char *buf, int n)
{
    int o = snprintf(buf, n, "u:%d=", h->lu.n);
    for (int i = 0; i < h->lu.nstage; ++i) o += snprintf(buf + o, n - o, "%s%d", i ? "x" : "", h->lu.radix[i]);
    o += snprintf(buf + o, n - o, " v:%d=", h->lv.n);
    for (int i = 0; i < h->lv.nstage; ++i) o += snprintf(buf + o, n - o, "%s%d", i ? "x" : "", h->lv.radix[i]);
    snprintf(buf + o, n - o, " box=%d pairs", h->pb);
}

int hfft_invert_plane(HandFft *h, const cplx<float> *plane, double w, float *accT, cudaStream_t s)
{
    const PlaneGeo &g = h->g;
    const int grid_r = g.nu < h->nsm ? g.nu : h->nsm;
    const int grid_c = g.npix_y / 2 < h->nsm ? g.npix_y / 2 : h->nsm;
    hfft_rows_inv_kernel<<<grid_r, kThreads, h->smem_v, s>>>(
        g, h->lv, h->d_tw + h->off_vp, reinterpret_cast<const float2 *>(plane), h->tmap, h->pb);
    hfft_cols_inv_kernel<<<grid_c, kThreads, h->smem_u, s>>>(g, h->lu, h->d_tw + h->off_up, h->d_T, w,
                                                            accT);
    return (int)cudaGetLastError();
}

int hfft_predict_plane(HandFft *h, const float *preT, double w, cplx<float> *plane, cudaStream_t s)
{
    const PlaneGeo &g = h->g;
    const int grid_r = g.nu < h->nsm ? g.nu : h->nsm;
    const int grid_c = g.npix_y / 2 < h->nsm ? g.npix_y / 2 : h->nsm;
    hfft_cols_fwd_kernel<<<grid_c, kThreads, h->smem_u, s>>>(g, h->lu, h->d_tw + h->off_um, preT, w,
                                                            h->d_T);
    hfft_rows_fwd_kernel<<<grid_r, kThreads, h->smem_v, s>>>(g, h->lv, h->d_tw + h->off_vm, h->tmap,
                                                            h->pb, reinterpret_cast<float2 *>(plane));
    return (int)cudaGetLastError();
}

template <typename Tout>
int hfft_correct(const ImgGeo &g, const float *accT, Tout *out, cudaStream_t s)
{
    dim3 grid((g.npix_y + 31) / 32, (g.npix_x + 31) / 32);
    correct_t_kernel<Tout><<<grid, 256, 0, s>>>(g, accT, out);
    return (int)cudaGetLastError();
}
template <typename Tin>
int hfft_precorrect(const ImgGeo &g, const Tin *dirty, float *preT, cudaStream_t s)
{
    dim3 grid((g.npix_y + 31) / 32, (g.npix_x + 31) / 32);
    precorrect_t_kernel<Tin><<<grid, 256, 0, s>>>(g, dirty, preT);
    return (int)cudaGetLastError();
}
template int hfft_correct<float>(const ImgGeo &, const float *, float *, cudaStream_t);
template int hfft_correct<double>(const ImgGeo &, const float *, double *, cudaStream_t);
template int hfft_precorrect<float>(const ImgGeo &, const float *, float *, cudaStream_t);
template int hfft_precorrect<double>(const ImgGeo &, const double *, float *, cudaStream_t);

}  // namespace b2g
