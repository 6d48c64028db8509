// kernels.h -- internal launcher interface between plan.cu and the kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_common.cuh"

namespace b2g {

// One gridding / degridding launch over the sorted visibilities [begin, end) of a slab
// (= the samples whose first w-plane p0 lies in [plane0, plane0 + gs)).
template <typename T> struct VisLaunch {
    VisGeo geo;
    int W;
    bool wstack;
    const double *coef;     // host, [W][W+4], highest power first
    const uint32_t *sidx;   // device, sorted visibility indices
    long long begin, end;   // range in sidx
    int plane0;             // first origin plane of the slab
    int slot0;              // ring slot of plane0 (= plane0 mod nslots)
    const cplx<T> *ms;      // gridding input  [nrow][nchan]
    const T *wgt;           // optional weights [nrow][nchan]
    cplx<T> *ms_out;        // degridding output [nrow][nchan]
    const cplx<T> *sub_from = nullptr;  // degridding: ms_out = sub_from - model (residual), may alias ms_out
    cplx<T> *ring;          // [nslots][nu][nv]
    cudaStream_t stream;
    // degridding: class of every chunk of 1024 sorted samples of [begin, end) (vis_tc.cu:
    // tc_classify_kernel) -- 0: warp-run kernel, 1: tensor-core kernel (long runs), 2: one thread
    // per sample (hardly any two samples share a footprint).  NULL: the launched kernel takes all.
    const unsigned char *chunk_class = nullptr;
};

// host table [W][W+4] (highest power first, exactly mirror-symmetric: es_kernel.h) -> the
// even / odd half-tap table the device evaluates
template <typename T, int W> inline KernelTab<T, W> make_tab(const double *coef)
{
    using K = KernelTab<T, W>;
    K t;
    for (int h = 0; h < K::NH2; ++h) {
        for (int k = 0; k < K::NE; ++k) {
            const int power = 2 * (K::NE - 1 - k);
            t.e[k][h] = h < K::NH ? (T)coef[h * (K::D + 1) + (K::D - power)] : T(0);
        }
        for (int k = 0; k < K::NO; ++k) {
            const int power = 2 * (K::NO - 1 - k) + 1;
            t.o[k][h] = h < K::NH ? (T)coef[h * (K::D + 1) + (K::D - power)] : T(0);
        }
    }
    return t;
}

template <typename T> int launch_grid_simple(const VisLaunch<T> &L);
template <typename T> int launch_degrid_simple(const VisLaunch<T> &L);

// tensor-core degridder (fp32, w-stacking, W <= 7): degrids the chunks of 1024 sorted samples whose
// runs of equal footprint origin are long (at most max_runs runs per chunk: class 1) and records
// the class of every chunk in chunk_class (one byte per chunk; class 2 = at least sparse_runs runs)
// for the other launches (max_runs < 0: the table is already filled); chunk_class == NULL: all
// chunks go through this kernel
bool tc_supported(int W, bool wstack);
long long tc_chunks(long long nvis);
int launch_degrid_tc(const VisLaunch<float> &L, unsigned char *chunk_class, int max_runs, int sparse_runs);

template <typename T> bool warprun_supported(int W, bool wstack);
template <typename T> int launch_grid_warprun(const VisLaunch<T> &L);
template <typename T> int launch_degrid_warprun(const VisLaunch<T> &L);

// Histogram of the first w-plane p0 of every sample (plane_counts[nplanes]): slab boundaries.
int launch_plane_hist(const VisGeo &geo, int W, long long nvis, unsigned long long *plane_counts,
                      cudaStream_t stream);

// Radix sort of the sample indices 0..n-1 by key = ((chunk*nplanes + p0)*nu + iu0)*nv + iv0
// (chunk = row / rows_per_chunk, 0 if rows_per_chunk == 0) on its lowest key_bits bits; the
// key is recomputed from the index in every pass.  Result in idx_out, idx_tmp is scratch.
// Returns 0 or a cudaError_t.
int sort_samples(const VisGeo &geo, int W, uint32_t rows_per_chunk, uint32_t *idx_tmp,
                 uint32_t *idx_out, long long n, int key_bits, cudaStream_t stream);

// min / max |w| in wavelengths over the selected samples (geo.sel, nsel): out2 = {min, max}
int launch_sel_wstats(const VisGeo &geo, long long nsel, double *out2_dev, cudaStream_t stream);

// uvw statistics over rows: out = {min|w|, max|w|, max|u|, max|v|} (metres)
int launch_uvw_stats(const double *uvw, long long nrow, double *out4_dev, cudaStream_t stream);

// ---- image-side kernels -------------------------------------------------------------
struct ImgGeo {
    int npix_x, npix_y, nu, nv;
    double pixsize_x, pixsize_y;
    int wstack;
    int nslots;
    double w0, dw;          // plane p sits at w = w0 + p*dw (wavelengths)
    const double *cu;       // [npix_x] 1/psi-hat_u
    const double *cv;       // [npix_y] 1/psi-hat_v
    const double *cheb;     // Chebyshev coefficients of 1/psi-hat_w(nm1*dw) in nm1 on [nm1min,0]
    int ncheb;
    double nm1min;          // most negative n-1 over the image (<= 0)
};

// acc[a][b] (+)= sum_{p<np} Re( ring[slot(pfirst+p)][a'][b'] * exp(-2 pi i (w0+(pfirst+p)dw) nm1) )
template <typename T>
int launch_screen_accumulate(const ImgGeo &g, const cplx<T> *ring, int pfirst, int np, T *acc,
                             bool first, cudaStream_t s);
// out[a][b] = acc[a][b] * cu[a]*cv[b]*gw(nm1)          (Tout = float or double)
template <typename T, typename Tout>
int launch_correct(const ImgGeo &g, const T *acc, Tout *out, cudaStream_t s);
// pre[a][b] = dirty[a][b] * cu[a]*cv[b]*gw(nm1)
template <typename T, typename Tin>
int launch_precorrect(const ImgGeo &g, const Tin *dirty, T *pre, cudaStream_t s);
// ring[slot(pfirst+p)][iu][iv] = pre[a][b]*exp(+2 pi i (w0+(pfirst+p)dw) nm1) on the image
// footprint, 0 elsewhere
template <typename T>
int launch_screen_apply(const ImgGeo &g, const T *pre, cplx<T> *ring, int pfirst, int np,
                        cudaStream_t s);

// ---- hand-written pruned 2-D FFT of fp32 w-planes, fused with the w-screen (fft_planes.cu) -----
// With it the image accumulator is kept TRANSPOSED: accT[npix_y][npix_x].
struct HandFft;
// 0: ready; < 0: geometry not supported (the caller keeps cuFFT); > 0: a cudaError_t
// ukeep: the rows u < ukeep and u >= nu - ukeep of the grid can hold samples, the others are left
// out of the rows passes (<= 0 or >= nu / 2: all rows)
// vkeep: the same along v (the rows passes load / store only the two ends of a row)
int hfft_create(HandFft **out, const ImgGeo &g, int ukeep, int vkeep, cudaStream_t s);
int hfft_ukeep(const HandFft *h);
int hfft_vkeep(const HandFft *h);
void hfft_destroy(HandFft *h, cudaStream_t s);
void hfft_describe(const HandFft *h, char *buf, int n);
// accT += Re( FFT^-1(plane) on the image * exp(-2 pi i w (n-1)) ); clear: the plane is left zeroed
// (every row is cleared by the rows pass once it has loaded it: no separate memset of the ring slot)
int hfft_invert_plane(HandFft *h, cplx<float> *plane, double w, float *accT, bool clear, cudaStream_t s);
// plane = FFT( preT * exp(+2 pi i w (n-1)), zero-padded )
int hfft_predict_plane(HandFft *h, const float *preT, double w, cplx<float> *plane, cudaStream_t s);
// out[a][b] = accT[b][a] * correction(a, b);  preT[b][a] = dirty[a][b] * correction(a, b)
template <typename Tout>
int hfft_correct(HandFft *h, const ImgGeo &g, const float *accT, Tout *out, cudaStream_t s);
template <typename Tin>
int hfft_precorrect(HandFft *h, const ImgGeo &g, const Tin *dirty, float *preT, cudaStream_t s);

// dtype conversion helpers (n real elements)
template <typename Tin, typename Tout>
int launch_convert(const Tin *in, Tout *out, long long n, cudaStream_t s);
// deterministic fp64 sum; scratch holds nscratch doubles
template <typename Tin>
int launch_sum(const Tin *in, long long n, double *out_dev, double *scratch, int nscratch,
               cudaStream_t s);

}  // namespace b2g
