// image_kernels.cu -- image-side kernels of the w-stacking gridder:
//   * w-plane phase screens (multiply-accumulate for invert, multiply-scatter for predict)
//   * gridding correction 1/(psihat_u * psihat_v * psihat_w * n)
//   * dtype conversion and the deterministic sum used for sumwt
// These replace the per-plane "apply w-screen", "grid correction" and cropping stages
// inside ducc0.wgridder.ms2dirty / dirty2ms (call sites
// /root/reference/src/ska_sdp_func_python/imaging/ng.py:240-256, :99-112).
//
// All of them are streaming kernels bound by HBM: threads run along y (the contiguous
// axis of both the image [npix_x][npix_y] and the grid planes [nu][nv]); each thread
// computes n-1 once in fp64 and loops over the planes of the slab, so the image is read
// and written once per slab instead of once per plane.
#include <math_constants.h>

#include "kernels.h"

namespace b2g {

namespace {

__device__ __forceinline__ double nm1_of(double r2) { return -r2 / (sqrt(1.0 - r2) + 1.0); }

__device__ __forceinline__ double cheb_eval(const double *__restrict__ c, int n, double z)
{
    // Clenshaw
    double b1 = 0.0, b2 = 0.0;
    for (int j = n - 1; j >= 1; --j) {
        double b0 = fma(2.0 * z, b1, __ldg(c + j) - b2);
        b2 = b1;
        b1 = b0;
    }
    return fma(z, b1, __ldg(c) - b2);
}

// gridding-correction factor of pixel (a,b); 0 beyond the horizon
__device__ __forceinline__ double corr_factor(const ImgGeo &g, int a, int b)
{
    double f = __ldg(g.cu + a) * __ldg(g.cv + b);
    if (g.wstack) {
        const double x = (double)(a - g.npix_x / 2) * g.pixsize_x;
        const double y = (double)(b - g.npix_y / 2) * g.pixsize_y;
        const double r2 = x * x + y * y;
        if (r2 >= 1.0) return 0.0;
        const double nm1 = nm1_of(r2);
        // nm1 in [nm1min, 0] -> z in [-1, 1]  (same map as cheb_fit(f, nm1min, 0) in es_kernel.h)
        const double z = (g.nm1min < 0.0) ? fma(-2.0 / g.nm1min, nm1, 1.0) : 0.0;
        f *= cheb_eval(g.cheb, g.ncheb, z) / (nm1 + 1.0);  // 1/(psi-hat_w * n)
    }
    return f;
}

__device__ __forceinline__ void phase_sincos(double ph, float &s, float &c)
{
    ph -= rint(ph);
    sincospif(2.0f * (float)ph, &s, &c);
}
__device__ __forceinline__ void phase_sincos(double ph, double &s, double &c)
{
    ph -= rint(ph);
    sincospi(2.0 * ph, &s, &c);
}

template <typename T>
__global__ void __launch_bounds__(256)
screen_accumulate_kernel(ImgGeo g, const cplx<T> *__restrict__ ring, int pfirst, int np,
                         T *__restrict__ acc, bool first)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    const int a = blockIdx.y;
    if (b >= g.npix_y) return;
    int ia = a - g.npix_x / 2;
    if (ia < 0) ia += g.nu;
    int ib = b - g.npix_y / 2;
    if (ib < 0) ib += g.nv;
    const long long cell = (long long)ia * g.nv + ib;
    const long long stride = (long long)g.nu * g.nv;
    T sum = T(0);
    if (!g.wstack) {
        sum = ring[cell].re;
    } else {
        const double x = (double)(a - g.npix_x / 2) * g.pixsize_x;
        const double y = (double)(b - g.npix_y / 2) * g.pixsize_y;
        const double r2 = x * x + y * y;
        if (r2 < 1.0) {
            const double nm1 = nm1_of(r2);
            int slot = pfirst % g.nslots;
            for (int p = 0; p < np; ++p) {
                const double w = fma((double)(pfirst + p), g.dw, g.w0);
                T s, c;
                phase_sincos(-w * nm1, s, c);
                const cplx<T> v = ring[(long long)slot * stride + cell];
                sum += v.re * c - v.im * s;
                slot = (slot + 1 == g.nslots) ? 0 : slot + 1;
            }
        }
    }
    const long long o = (long long)a * g.npix_y + b;
    acc[o] = first ? sum : acc[o] + sum;
}

template <typename T, typename Tout>
__global__ void __launch_bounds__(256)
correct_kernel(ImgGeo g, const T *__restrict__ acc, Tout *__restrict__ out)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    const int a = blockIdx.y;
    if (b >= g.npix_y) return;
    const long long o = (long long)a * g.npix_y + b;
    out[o] = (Tout)((double)acc[o] * corr_factor(g, a, b));
}

template <typename T, typename Tin>
__global__ void __launch_bounds__(256)
precorrect_kernel(ImgGeo g, const Tin *__restrict__ dirty, T *__restrict__ pre)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    const int a = blockIdx.y;
    if (b >= g.npix_y) return;
    const long long o = (long long)a * g.npix_y + b;
    pre[o] = (T)((double)dirty[o] * corr_factor(g, a, b));
}

// One thread per grid cell (iu, iv); writes that cell of every plane of the slab.
template <typename T>
__global__ void __launch_bounds__(256)
screen_apply_kernel(ImgGeo g, const T *__restrict__ pre, cplx<T> *__restrict__ ring, int pfirst,
                    int np)
{
    const int iv = blockIdx.x * blockDim.x + threadIdx.x;
    const int iu = blockIdx.y;
    if (iv >= g.nv) return;
    const long long cell = (long long)iu * g.nv + iv;
    const long long stride = (long long)g.nu * g.nv;
    // grid index -> image pixel (or none)
    const int hx = g.npix_x / 2, hy = g.npix_y / 2;
    int a = -1, b = -1;
    if (iu < g.npix_x - hx) a = iu + hx;
    else if (iu >= g.nu - hx) a = iu - g.nu + hx;
    if (iv < g.npix_y - hy) b = iv + hy;
    else if (iv >= g.nv - hy) b = iv - g.nv + hy;
    T val = T(0);
    double nm1 = 0.0;
    bool live = (a >= 0 && b >= 0);
    if (live) {
        val = pre[(long long)a * g.npix_y + b];
        if (g.wstack) {
            const double x = (double)(a - hx) * g.pixsize_x;
            const double y = (double)(b - hy) * g.pixsize_y;
            const double r2 = x * x + y * y;
            if (r2 < 1.0) nm1 = nm1_of(r2);
            else val = T(0);
        }
        live = (val != T(0));
    }
    int slot = pfirst % g.nslots;
    if (!live) {
        for (int p = 0; p < np; ++p) {
            ring[(long long)slot * stride + cell] = cplx<T>{T(0), T(0)};
            slot = (slot + 1 == g.nslots) ? 0 : slot + 1;
        }
        return;
    }
    if (!g.wstack) {
        ring[cell] = cplx<T>{val, T(0)};
        return;
    }
    for (int p = 0; p < np; ++p) {
        const double w = fma((double)(pfirst + p), g.dw, g.w0);
        T s, c;
        phase_sincos(w * nm1, s, c);
        ring[(long long)slot * stride + cell] = cplx<T>{val * c, val * s};
        slot = (slot + 1 == g.nslots) ? 0 : slot + 1;
    }
}

template <typename Tin, typename Tout>
__global__ void convert_kernel(const Tin *__restrict__ in, Tout *__restrict__ out, long long n)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long step = (long long)gridDim.x * blockDim.x;
    for (; i < n; i += step) out[i] = (Tout)in[i];
}

// fixed-shape two-stage reduction: block b sums elements [b*per, (b+1)*per) in a fixed order
template <typename Tin>
__global__ void __launch_bounds__(256)
sum_stage1(const Tin *__restrict__ in, long long n, long long per, double *__restrict__ part)
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
__global__ void __launch_bounds__(256) sum_stage2(const double *__restrict__ part, int nb,
                                                 double *__restrict__ out)
{
    __shared__ double sh[256];
    double s = 0.0;
    for (int i = threadIdx.x; i < nb; i += 256) s += part[i];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if ((int)threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = sh[0];
}

// min|w|, max|w|, max|u|, max|v| over rows; any non-finite coordinate makes max|w| = +inf
__global__ void __launch_bounds__(256)
uvw_stats_kernel(const double *__restrict__ uvw, long long nrow, double *__restrict__ out4)
{
    __shared__ double sh[4][256];
    double wmin = 1e300, wmax = 0.0, umax = 0.0, vmax = 0.0;
    for (long long r = (long long)blockIdx.x * 256 + threadIdx.x; r < nrow;
         r += (long long)gridDim.x * 256) {
        const double u = fabs(uvw[3 * r]), v = fabs(uvw[3 * r + 1]), w = fabs(uvw[3 * r + 2]);
        if (!(isfinite(u) && isfinite(v) && isfinite(w))) {  // reported as max|w| = inf
            wmax = CUDART_INF;
            continue;
        }
        wmin = fmin(wmin, w);
        wmax = fmax(wmax, w);
        umax = fmax(umax, u);
        vmax = fmax(vmax, v);
    }
    sh[0][threadIdx.x] = wmin;
    sh[1][threadIdx.x] = wmax;
    sh[2][threadIdx.x] = umax;
    sh[3][threadIdx.x] = vmax;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if ((int)threadIdx.x < w) {
            sh[0][threadIdx.x] = fmin(sh[0][threadIdx.x], sh[0][threadIdx.x + w]);
            sh[1][threadIdx.x] = fmax(sh[1][threadIdx.x], sh[1][threadIdx.x + w]);
            sh[2][threadIdx.x] = fmax(sh[2][threadIdx.x], sh[2][threadIdx.x + w]);
            sh[3][threadIdx.x] = fmax(sh[3][threadIdx.x], sh[3][threadIdx.x + w]);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        // doubles are non-negative here: integer compare on the bit pattern is order-preserving
        atomicMin((unsigned long long *)&out4[0], (unsigned long long)__double_as_longlong(sh[0][0]));
        atomicMax((unsigned long long *)&out4[1], (unsigned long long)__double_as_longlong(sh[1][0]));
        atomicMax((unsigned long long *)&out4[2], (unsigned long long)__double_as_longlong(sh[2][0]));
        atomicMax((unsigned long long *)&out4[3], (unsigned long long)__double_as_longlong(sh[3][0]));
    }
}

inline dim3 img_grid(int nfast, int nslow) { return dim3((unsigned)((nfast + 255) / 256), (unsigned)nslow); }

}  // namespace

template <typename T>
int launch_screen_accumulate(const ImgGeo &g, const cplx<T> *ring, int pfirst, int np, T *acc,
                             bool first, cudaStream_t s)
{
    screen_accumulate_kernel<T><<<img_grid(g.npix_y, g.npix_x), 256, 0, s>>>(g, ring, pfirst, np,
                                                                            acc, first);
    return (int)cudaGetLastError();
}
template <typename T, typename Tout>
int launch_correct(const ImgGeo &g, const T *acc, Tout *out, cudaStream_t s)
{
    correct_kernel<T, Tout><<<img_grid(g.npix_y, g.npix_x), 256, 0, s>>>(g, acc, out);
    return (int)cudaGetLastError();
}
template <typename T, typename Tin>
int launch_precorrect(const ImgGeo &g, const Tin *dirty, T *pre, cudaStream_t s)
{
    precorrect_kernel<T, Tin><<<img_grid(g.npix_y, g.npix_x), 256, 0, s>>>(g, dirty, pre);
    return (int)cudaGetLastError();
}
template <typename T>
int launch_screen_apply(const ImgGeo &g, const T *pre, cplx<T> *ring, int pfirst, int np,
                        cudaStream_t s)
{
    screen_apply_kernel<T><<<img_grid(g.nv, g.nu), 256, 0, s>>>(g, pre, ring, pfirst, np);
    return (int)cudaGetLastError();
}
template <typename Tin, typename Tout>
int launch_convert(const Tin *in, Tout *out, long long n, cudaStream_t s)
{
    if (n <= 0) return 0;
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    convert_kernel<Tin, Tout><<<(unsigned)blocks, 256, 0, s>>>(in, out, n);
    return (int)cudaGetLastError();
}
template <typename Tin>
int launch_sum(const Tin *in, long long n, double *out_dev, double *scratch, int nscratch,
               cudaStream_t s)
{
    int nb = nscratch;
    long long per = (n + nb - 1) / nb;
    if (per < 1) per = 1;
    sum_stage1<Tin><<<nb, 256, 0, s>>>(in, n, per, scratch);
    sum_stage2<<<1, 256, 0, s>>>(scratch, nb, out_dev);
    return (int)cudaGetLastError();
}
int launch_uvw_stats(const double *uvw, long long nrow, double *out4_dev, cudaStream_t stream)
{
    // out4 = {+big, 0, 0, 0}
    const double init[4] = {1e300, 0.0, 0.0, 0.0};
    cudaMemcpyAsync(out4_dev, init, sizeof(init), cudaMemcpyHostToDevice, stream);
    long long blocks = (nrow + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    uvw_stats_kernel<<<(unsigned)blocks, 256, 0, stream>>>(uvw, nrow, out4_dev);
    return (int)cudaGetLastError();
}

#define INST_T(T)                                                                              \
    template int launch_screen_accumulate<T>(const ImgGeo &, const cplx<T> *, int, int, T *,   \
                                             bool, cudaStream_t);                              \
    template int launch_correct<T, float>(const ImgGeo &, const T *, float *, cudaStream_t);    \
    template int launch_correct<T, double>(const ImgGeo &, const T *, double *, cudaStream_t);  \
    template int launch_precorrect<T, float>(const ImgGeo &, const float *, T *, cudaStream_t); \
    template int launch_precorrect<T, double>(const ImgGeo &, const double *, T *, cudaStream_t); \
    template int launch_screen_apply<T>(const ImgGeo &, const T *, cplx<T> *, int, int,         \
                                        cudaStream_t);
INST_T(float)
INST_T(double)
template int launch_convert<float, double>(const float *, double *, long long, cudaStream_t);
template int launch_convert<double, float>(const double *, float *, long long, cudaStream_t);
template int launch_sum<float>(const float *, long long, double *, double *, int, cudaStream_t);
template int launch_sum<double>(const double *, long long, double *, double *, int, cudaStream_t);

}  // namespace b2g
