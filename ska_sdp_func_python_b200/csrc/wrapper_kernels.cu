// wrapper_kernels.cu -- device side of the invert_ng / predict_ng marshalling.
//
// The reference wrapper prepares the gridder input with several full-size numpy passes
// (/root/reference/src/ska_sdp_func_python/imaging/ng.py:182-204: deep copy, phase rotation to
// the image centre via shift_vis_to_image -> phaserotate_visibility
// (imaging/base.py:48-92, visibility/base.py:27-45, :60-96), flagged_vis = vis*(1-flags),
// convert_pol_frame to the image polarisation frame, astype(c16), transposes, per-iteration
// ascontiguousarray at :237-239) and post-processes the degridder output the same way
// (ng.py:131-143).  Here each direction is ONE streaming kernel over the visibility cube:
//
//   prepare:  ms[p][row][chan]  = sum_q M[p][q] * vis[row][chan][q]*(1-flags[row][chan][q])
//                                 * exp(+2 pi i f/c (u l + v m + w (n-1)))        (conj phasor)
//             wgt[p][row][chan] = weight[row][chan][p]*(1-flags[row][chan][p])
//             (dopsf: ms[0] = 1, ms[p>0] = 0, ng.py:231-233)
//   finish:   vis[row][chan][q] = (sum_p M'[q][p] * ms[p][row][chan])
//                                 * exp(-2 pi i f/c (u l + v m + w (n-1)))
//
// Both are HBM-bound: every input element is read once, every output element written once.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200_wgridder.h"
#include "device_common.cuh"

namespace b2g {
int wrapper_fail(const char *msg);  // defined in plan.cu (sets the thread-local error)
}

namespace {

using b2g::cplx;

struct PolMat {
    double re[4][4], im[4][4];
};

struct Shift {
    int on;            // 0: no phase rotation
    double l, m, nm1;  // direction cosines of the image centre w.r.t. the vis phase centre,
                       // already signed for the (flipped) uvw passed alongside
};

__device__ __forceinline__ double flag_factor(const void *flags, int type, long long i)
{
    switch (type) {
    case B2G_FLAG_U8: return 1.0 - (double)static_cast<const uint8_t *>(flags)[i];
    case B2G_FLAG_I32: return 1.0 - (double)static_cast<const int32_t *>(flags)[i];
    case B2G_FLAG_I64: return 1.0 - (double)static_cast<const long long *>(flags)[i];
    case B2G_FLAG_F64: return 1.0 - static_cast<const double *>(flags)[i];
    default: return 1.0;
    }
}

template <typename Tv> __device__ __forceinline__ void load_c(const void *p, long long i, double &re, double &im)
{
    const cplx<Tv> v = static_cast<const cplx<Tv> *>(p)[i];
    re = (double)v.re;
    im = (double)v.im;
}

__device__ __forceinline__ void phasor(const Shift &sh, const double *__restrict__ uvw,
                                       const double *__restrict__ fscale, long long row, int chan,
                                       double sign, double &c, double &s)
{
    const double ph = (uvw[3 * row] * sh.l + uvw[3 * row + 1] * sh.m + uvw[3 * row + 2] * sh.nm1) *
                      fscale[chan];
    const double fr = ph - rint(ph);
    sincospi(2.0 * sign * fr, &s, &c);
}

template <typename Tv, typename Tw, typename T>
__global__ void __launch_bounds__(256)
vis_prepare_kernel(long long nrow, int nchan, int npol_in, int npol_out, const void *__restrict__ vis,
                   const void *__restrict__ flags, int flag_type, const Tw *__restrict__ weight,
                   PolMat M, Shift sh, const double *__restrict__ uvw,
                   const double *__restrict__ fscale, long long row0, long long nrow_total,
                   int dopsf, cplx<T> *__restrict__ ms_out, T *__restrict__ wgt_out)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // sample in the chunk
    if (i >= nrow * nchan) return;
    const long long r = i / nchan;
    const int c = (int)(i - r * nchan);
    double vr[4], vi[4], ff[4];
    for (int q = 0; q < npol_in; ++q) {
        const long long e = i * npol_in + q;
        ff[q] = flag_factor(flags, flag_type, e);
        load_c<Tv>(vis, e, vr[q], vi[q]);
        vr[q] *= ff[q];
        vi[q] *= ff[q];
    }
    double pc = 1.0, ps = 0.0;
    if (sh.on) phasor(sh, uvw, fscale, row0 + r, c, +1.0, pc, ps);
    const long long o = (row0 + r) * nchan + c;
    const long long pstride = nrow_total * nchan;
    for (int p = 0; p < npol_out; ++p) {
        double ar = 0.0, ai = 0.0;
        for (int q = 0; q < npol_in; ++q) {
            ar += M.re[p][q] * vr[q] - M.im[p][q] * vi[q];
            ai += M.re[p][q] * vi[q] + M.im[p][q] * vr[q];
        }
        double orr = ar * pc - ai * ps, oi = ar * ps + ai * pc;
        if (dopsf) {
            orr = p == 0 ? 1.0 : 0.0;
            oi = 0.0;
        }
        ms_out[p * pstride + o] = cplx<T>{(T)orr, (T)oi};
        if (wgt_out) {
            const double w = weight ? (double)weight[i * npol_in + p] * ff[p] : ff[p];
            wgt_out[p * pstride + o] = (T)w;
        }
    }
}

template <typename T, typename Tv>
__global__ void __launch_bounds__(256)
vis_finish_kernel(long long nrow, int nchan, int npol_img, int npol_vis,
                  const cplx<T> *__restrict__ ms, PolMat M, Shift sh,
                  const double *__restrict__ uvw, const double *__restrict__ fscale, long long row0,
                  long long nrow_total, cplx<Tv> *__restrict__ vis_out)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nrow * nchan) return;
    const long long r = i / nchan;
    const int c = (int)(i - r * nchan);
    const long long o = (row0 + r) * nchan + c;
    const long long pstride = nrow_total * nchan;
    double vr[4], vi[4];
    for (int p = 0; p < npol_img; ++p) {
        const cplx<T> v = ms[p * pstride + o];
        vr[p] = (double)v.re;
        vi[p] = (double)v.im;
    }
    double pc = 1.0, ps = 0.0;
    if (sh.on) phasor(sh, uvw, fscale, row0 + r, c, -1.0, pc, ps);
    for (int q = 0; q < npol_vis; ++q) {
        double ar = 0.0, ai = 0.0;
        for (int p = 0; p < npol_img; ++p) {
            ar += M.re[q][p] * vr[p] - M.im[q][p] * vi[p];
            ai += M.re[q][p] * vi[p] + M.im[q][p] * vr[p];
        }
        vis_out[i * npol_vis + q] = cplx<Tv>{(Tv)(ar * pc - ai * ps), (Tv)(ar * ps + ai * pc)};
    }
}

PolMat make_mat(const double *polmat, int nout, int nin)
{
    PolMat M;
    for (int a = 0; a < 4; ++a)
        for (int b = 0; b < 4; ++b) M.re[a][b] = M.im[a][b] = 0.0;
    for (int a = 0; a < nout; ++a)
        for (int b = 0; b < nin; ++b) {
            M.re[a][b] = polmat[2 * (a * nin + b)];
            M.im[a][b] = polmat[2 * (a * nin + b) + 1];
        }
    return M;
}

Shift make_shift(const double *lmn)
{
    Shift s{0, 0.0, 0.0, 0.0};
    if (lmn) {
        s.on = 1;
        s.l = lmn[0];
        s.m = lmn[1];
        s.nm1 = lmn[2];
    }
    return s;
}

}  // namespace

extern "C" int b2g_vis_prepare(int64_t nrow, int64_t nchan, int npol_in, int npol_out,
                               const void *vis, int vis_dtype, const void *flags, int flag_dtype,
                               const void *weight, int wgt_dtype, const double *polmat,
                               const double *lmn, const double *uvw, const double *fscale,
                               int64_t row0, int64_t nrow_total, int dopsf, void *ms_out,
                               void *wgt_out, int out_dtype, void *stream)
{
    if (npol_in < 1 || npol_in > 4 || npol_out < 1 || npol_out > npol_in)
        return b2g::wrapper_fail("b2g_vis_prepare: need 1 <= npol_out <= npol_in <= 4");
    if (!vis || !ms_out || !polmat) return b2g::wrapper_fail("b2g_vis_prepare: NULL pointer");
    if (lmn && (!uvw || !fscale)) return b2g::wrapper_fail("b2g_vis_prepare: phase shift needs uvw and fscale");
    const long long n = nrow * nchan;
    if (n <= 0) return 0;
    const PolMat M = make_mat(polmat, npol_out, npol_in);
    const Shift sh = make_shift(lmn);
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned blocks = (unsigned)((n + 255) / 256);
#define LAUNCH(Tv, Tw, T)                                                                         \
    vis_prepare_kernel<Tv, Tw, T><<<blocks, 256, 0, s>>>(                                         \
        nrow, (int)nchan, npol_in, npol_out, vis, flags, flags ? flag_dtype : B2G_FLAG_NONE,      \
        static_cast<const Tw *>(weight), M, sh, uvw, fscale, row0, nrow_total, dopsf,             \
        static_cast<cplx<T> *>(ms_out), static_cast<T *>(wgt_out))
    const int key = (vis_dtype == B2G_F64 ? 4 : 0) | (wgt_dtype == B2G_F64 ? 2 : 0) | (out_dtype == B2G_F64 ? 1 : 0);
    switch (key) {
    case 0: LAUNCH(float, float, float); break;
    case 1: LAUNCH(float, float, double); break;
    case 2: LAUNCH(float, double, float); break;
    case 3: LAUNCH(float, double, double); break;
    case 4: LAUNCH(double, float, float); break;
    case 5: LAUNCH(double, float, double); break;
    case 6: LAUNCH(double, double, float); break;
    default: LAUNCH(double, double, double); break;
    }
#undef LAUNCH
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : b2g::wrapper_fail(cudaGetErrorString(e));
}

extern "C" int b2g_vis_finish(int64_t nrow, int64_t nchan, int npol_img, int npol_vis,
                              const void *ms, int ms_dtype, const double *polmat,
                              const double *lmn, const double *uvw, const double *fscale,
                              int64_t row0, int64_t nrow_total, void *vis_out, int vis_dtype,
                              void *stream)
{
    if (npol_img < 1 || npol_img > 4 || npol_vis < 1 || npol_vis > 4)
        return b2g::wrapper_fail("b2g_vis_finish: need 1 <= npol <= 4");
    if (!ms || !vis_out || !polmat) return b2g::wrapper_fail("b2g_vis_finish: NULL pointer");
    if (lmn && (!uvw || !fscale)) return b2g::wrapper_fail("b2g_vis_finish: phase shift needs uvw and fscale");
    const long long n = nrow * nchan;
    if (n <= 0) return 0;
    const PolMat M = make_mat(polmat, npol_vis, npol_img);
    const Shift sh = make_shift(lmn);
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned blocks = (unsigned)((n + 255) / 256);
#define LAUNCH(T, Tv)                                                                             \
    vis_finish_kernel<T, Tv><<<blocks, 256, 0, s>>>(nrow, (int)nchan, npol_img, npol_vis,         \
                                                    static_cast<const cplx<T> *>(ms), M, sh, uvw,  \
                                                    fscale, row0, nrow_total,                     \
                                                    static_cast<cplx<Tv> *>(vis_out))
    if (ms_dtype == B2G_F64) {
        if (vis_dtype == B2G_F64) LAUNCH(double, double); else LAUNCH(double, float);
    } else {
        if (vis_dtype == B2G_F64) LAUNCH(float, double); else LAUNCH(float, float);
    }
#undef LAUNCH
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : b2g::wrapper_fail(cudaGetErrorString(e));
}
