// dft_kernels.cu -- direct Fourier transform of sky components to visibilities (SURVEY.md
// section 8f row N3), the other half of model prediction next to predict_ng.
//
// Replaces dft_kernel / dft_cpu_looped and the naive cupy kernel string of the reference
// (/root/reference/src/ska_sdp_func_python/imaging/dft.py:121-183, :265-285, :185-262):
//
//   vis[row, chan, pol] = sum_comp flux[comp, chan, pol] * exp(-2 pi i (u l + v m + w n'))
//
// with (u, v, w) in wavelengths (= uvw * f/c, the product the reference forms) and
// (l, m, n') = (l, m, sqrt(1 - l^2 - m^2) - 1) per component.
//
// Design: thread = one (row, chan) sample, lanes along chan (coalesced output, coalesced flux
// rows, uvw broadcast); every thread walks the component list, keeping the NPOL complex sums in
// registers.  Phases are formed in fp64 and reduced to [-1/2, 1/2] turns EXACTLY (phase -
// rint(phase)) before the trigonometric evaluation, so accuracy does not degrade with |uvw| (the
// reference multiplies by 2 pi first and loses ~|phase| * 1e-16).  The sine / cosine are
// sincospi in fp64 (default) or fp32 (opt-in, |error| ~ 1e-7 per term).  The kernel is bound by
// fp64 issue (about 60 DFMA-equivalents per sample and component), not by HBM: per sample it
// reads 24 B of uvw (shared by the row's channels) and writes 16 * NPOL bytes once.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200_wgridder.h"

namespace b2g {
int wrapper_fail(const char *msg);  // plan.cu
}

namespace {

struct c128 {
    double re, im;
};

template <bool F32TRIG> __device__ __forceinline__ void unit_phasor(double turns, double &c, double &s)
{
    // exp(-2 pi i turns), turns already reduced to [-1/2, 1/2]
    if (F32TRIG) {
        float sf, cf;
        sincospif((float)(-2.0 * turns), &sf, &cf);
        c = (double)cf;
        s = (double)sf;
    } else {
        sincospi(-2.0 * turns, &s, &c);
    }
}

// per_sample_uvw: uvw is [nrow*nchan][3] already in wavelengths (the reference's uvw_lambda
// array); otherwise uvw is [nrow][3] metres and kscale [nchan] = f/c.
template <int NPOL, bool F32TRIG>
__global__ void __launch_bounds__(256)
dft_predict_kernel(long long nrow, int nchan, long long ncomp, const double *__restrict__ uvw,
                   const double *__restrict__ kscale, const double *__restrict__ dircos,
                   const c128 *__restrict__ flux, int flux_nchan, c128 *__restrict__ out,
                   int accumulate)
{
    const long long n = nrow * nchan;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    long long row;
    int chan;
    if (i < 0x100000000ll) {
        const unsigned r = (unsigned)i / (unsigned)nchan;
        row = r;
        chan = (int)((unsigned)i - r * (unsigned)nchan);
    } else {
        row = i / nchan;
        chan = (int)(i - row * nchan);
    }
    double u, v, w;
    if (kscale) {
        const double k = kscale[chan];
        u = __dmul_rn(uvw[3 * row + 0], k);
        v = __dmul_rn(uvw[3 * row + 1], k);
        w = __dmul_rn(uvw[3 * row + 2], k);
    } else {
        u = uvw[3 * i + 0];
        v = uvw[3 * i + 1];
        w = uvw[3 * i + 2];
    }
    double are[NPOL], aim[NPOL];
#pragma unroll
    for (int p = 0; p < NPOL; ++p) are[p] = aim[p] = 0.0;
    const int fchan = flux_nchan == 1 ? 0 : chan;
    const c128 *fl = flux + (size_t)fchan * NPOL;
    const size_t fstride = (size_t)flux_nchan * NPOL;
#pragma unroll 2
    for (long long c = 0; c < ncomp; ++c) {
        const double l = __ldg(dircos + 3 * c), m = __ldg(dircos + 3 * c + 1),
                     nn = __ldg(dircos + 3 * c + 2);
        // the reference's numpy.sum(uvw_lambda * dc, axis=-1): products rounded, added in order
        const double phase = __dadd_rn(__dadd_rn(__dmul_rn(u, l), __dmul_rn(v, m)), __dmul_rn(w, nn));
        double cs, sn;
        unit_phasor<F32TRIG>(phase - rint(phase), cs, sn);
#pragma unroll
        for (int p = 0; p < NPOL; ++p) {
            const c128 f = fl[c * fstride + p];
            are[p] = fma(f.re, cs, fma(-f.im, sn, are[p]));
            aim[p] = fma(f.re, sn, fma(f.im, cs, aim[p]));
        }
    }
#pragma unroll
    for (int p = 0; p < NPOL; ++p) {
        c128 r{are[p], aim[p]};
        if (accumulate) {
            const c128 o = out[i * NPOL + p];
            r.re += o.re;
            r.im += o.im;
        }
        out[i * NPOL + p] = r;
    }
}

template <int NPOL>
void launch(bool f32, unsigned blocks, cudaStream_t s, long long nrow, int nchan, long long ncomp,
            const double *uvw, const double *kscale, const double *dircos, const c128 *flux,
            int flux_nchan, c128 *out, int accumulate)
{
    if (f32)
        dft_predict_kernel<NPOL, true><<<blocks, 256, 0, s>>>(nrow, nchan, ncomp, uvw, kscale, dircos,
                                                             flux, flux_nchan, out, accumulate);
    else
        dft_predict_kernel<NPOL, false><<<blocks, 256, 0, s>>>(nrow, nchan, ncomp, uvw, kscale, dircos,
                                                              flux, flux_nchan, out, accumulate);
}

}  // namespace

extern "C" int b2g_dft_predict(int64_t nrow, int64_t nchan, int npol, int64_t ncomp,
                               const double *uvw, const double *kscale, const double *dircos,
                               const void *flux, int64_t flux_nchan, void *vis_out,
                               int trig_precision, int accumulate, void *stream)
{
    if (npol < 1 || npol > 4) return b2g::wrapper_fail("b2g_dft_predict: need 1 <= npol <= 4");
    if (nrow < 0 || nchan < 1 || ncomp < 0) return b2g::wrapper_fail("b2g_dft_predict: bad shape");
    if (flux_nchan != 1 && flux_nchan != nchan)
        return b2g::wrapper_fail("b2g_dft_predict: flux must have 1 or nchan channels");
    if (!uvw || !vis_out || (ncomp > 0 && (!dircos || !flux)))
        return b2g::wrapper_fail("b2g_dft_predict: NULL pointer");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return b2g::wrapper_fail("b2g_dft_predict: no CUDA device available (this library has no CPU path)");
    const long long n = nrow * nchan;
    if (n <= 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    const bool f32 = trig_precision == B2G_F32;
    const c128 *f = static_cast<const c128 *>(flux);
    c128 *o = static_cast<c128 *>(vis_out);
    switch (npol) {
    case 1: launch<1>(f32, blocks, s, nrow, (int)nchan, ncomp, uvw, kscale, dircos, f, (int)flux_nchan, o, accumulate); break;
    case 2: launch<2>(f32, blocks, s, nrow, (int)nchan, ncomp, uvw, kscale, dircos, f, (int)flux_nchan, o, accumulate); break;
    case 3: launch<3>(f32, blocks, s, nrow, (int)nchan, ncomp, uvw, kscale, dircos, f, (int)flux_nchan, o, accumulate); break;
    default: launch<4>(f32, blocks, s, nrow, (int)nchan, ncomp, uvw, kscale, dircos, f, (int)flux_nchan, o, accumulate); break;
    }
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : b2g::wrapper_fail(cudaGetErrorString(e));
}
