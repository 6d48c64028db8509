/*
 * b200_wgridder.h -- C ABI of the B200-native visibility gridder / degridder.
 *
 * This is the drop-in boundary for the hot path of
 *   ska_sdp_func_python.imaging.ng.invert_ng / predict_ng
 * (/root/reference/src/ska_sdp_func_python/imaging/ng.py:146-294, :38-143).  In the
 * reference all arithmetic of that path happens inside two calls into the third-party
 * C++ module ducc0.wgridder:
 *
 *   ducc0.wgridder.ms2dirty(uvw, freq, ms, wgt, npix_x, npix_y, pixsize_x, pixsize_y,
 *                           nu, nv, epsilon, do_wstacking, nthreads=, verbosity=,
 *                           double_precision_accumulation=)     ng.py:240-256, :271-287
 *   ducc0.wgridder.dirty2ms(uvw, freq, dirty, wgt, pixsize_x, pixsize_y, nu, nv,
 *                           epsilon, do_wstacking, nthreads, verbosity)
 *                                                               ng.py:99-112, :117-129
 *
 * The entry points below are what an FFI binding for those two calls binds instead
 * (see INTEGRATION.md for the ctypes stub).  All functions return 0 on success and a
 * non-zero status otherwise; b2g_last_error() returns a thread-local message.
 *
 * Conventions (identical to the ducc0 array-level convention, restated in
 * oracle/dft_oracle.c):
 *   uvw    [nrow][3]      float64, metres
 *   freq   [nchan]        float64, Hz
 *   ms     [nrow][nchan]  complex (interleaved re,im), float32 or float64 components
 *   wgt    [nrow][nchan]  real, same component type as ms; may be NULL (all ones)
 *   dirty  [npix_x][npix_y] real, float32 or float64
 *   dirty[ix,iy] = (1/n) Re sum wgt*ms*exp(+2 pi i f/c (u x + v y - w (n-1)))
 *   ms[row,chan] = wgt * sum dirty/n * exp(-2 pi i f/c (u x + v y - w (n-1)))
 *   with x=(ix-npix_x/2) pixsize_x, y=(iy-npix_y/2) pixsize_y, n=sqrt(1-x^2-y^2);
 *   do_wstacking=0 drops the w term and the 1/n factor.
 *
 * Pointers may be host or CUDA-device pointers, declared by the *_space arguments.
 * There is no CPU fallback: every call needs a CUDA device and fails otherwise.
 */
#ifndef B200_WGRIDDER_H
#define B200_WGRIDDER_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* component type codes */
#define B2G_F32 0
#define B2G_F64 1
/* memory spaces */
#define B2G_HOST 0
#define B2G_DEVICE 1
/* integer / real types accepted for the flags cube of b2g_vis_prepare */
#define B2G_FLAG_NONE 0
#define B2G_FLAG_U8 1
#define B2G_FLAG_I32 2
#define B2G_FLAG_I64 3
#define B2G_FLAG_F64 4
/* gridding kernel families */
#define B2G_ALGO_AUTO (-1)
#define B2G_ALGO_SIMPLE 0   /* one thread per visibility, global atomics (validation path) */
#define B2G_ALGO_WARPRUN 1  /* sorted runs, register window, warp per run (fast path)      */

typedef struct b2g_plan b2g_plan;

/* Thread-local description of the last failure in this thread ("" if none). */
const char *b2g_last_error(void);

/* Library version string. */
const char *b2g_version(void);

/*
 * Plan: everything that depends only on (uvw, freq, image geometry, epsilon):
 * kernel/oversampling choice, w-plane layout, gridding-correction tables and the
 * (uv-cell, w-plane)-sorted visibility index.  Replaces the per-call scan/bucket pass
 * ducc0 repeats on every ms2dirty/dirty2ms call (the reference wrapper calls it once per
 * polarisation and channel with identical uvw: ng.py:235-289).
 *
 *   precision : B2G_F32 / B2G_F64 arithmetic of the kernels
 *   uvw_space : B2G_HOST / B2G_DEVICE for uvw and freq
 *   nu, nv    : padded grid size, 0 = choose
 *   algo      : B2G_ALGO_*
 *   stream    : cudaStream_t (NULL = default stream); all work of the plan is ordered on it
 */
int b2g_plan_create(b2g_plan **plan, int64_t nrow, int64_t nchan, const double *uvw,
                    const double *freq, int uvw_space, int64_t npix_x, int64_t npix_y,
                    double pixsize_x, double pixsize_y, int64_t nu, int64_t nv,
                    double epsilon, int do_wstacking, int precision, int algo,
                    int verbosity, void *stream);

/* Plan over a SUBSET of the observation's samples (w-plane sharding across GPUs, SURVEY.md
 * section 8e): sel[i] = row*nchan + chan of plan sample i (uint32, same memory space as uvw,
 * entries < nrow*nchan, no duplicates).  The plan's w-planes cover the selected samples only, and
 * the ms / wgt arrays of b2g_plan_invert / b2g_plan_predict are then COMPACT: nsel entries in
 * the order of sel.  A device sel array is referenced, not copied, and must outlive the plan.
 * sel == NULL gives b2g_plan_create. */
int b2g_plan_create_sel(b2g_plan **plan, int64_t nrow, int64_t nchan, const double *uvw,
                        const double *freq, int uvw_space, const uint32_t *sel, int64_t nsel,
                        int64_t npix_x, int64_t npix_y, double pixsize_x, double pixsize_y,
                        int64_t nu, int64_t nv, double epsilon, int do_wstacking, int precision,
                        int algo, int verbosity, void *stream);

void b2g_plan_destroy(b2g_plan *plan);

/* ms2dirty on a plan.  ms/wgt: component type vis_dtype in vis_space;
 * dirty: dirty_dtype in dirty_space (written, not accumulated). */
int b2g_plan_invert(b2g_plan *plan, const void *ms, const void *wgt, int vis_dtype,
                    int vis_space, void *dirty, int dirty_dtype, int dirty_space);

/* dirty2ms on a plan.  ms_out: component type vis_dtype in vis_space (overwritten). */
int b2g_plan_predict(b2g_plan *plan, const void *dirty, int dirty_dtype, int dirty_space,
                     const void *wgt, void *ms_out, int vis_dtype, int vis_space);

/* Major-cycle residual on a plan: dirty_out = ms2dirty(ms - dirty2ms(model), wgt) in one call.
 * The degridder subtracts the model from the data as it writes each sample and the residual
 * samples feed the gridder without leaving HBM (the reference predicts, subtracts
 * [visibility/operations.py:86 subtract_visibility] and inverts in three passes over host
 * arrays: sky_model/skymodel_imaging.py:70-151, :154-235).  The model is degridded WITHOUT
 * weights (predict_ng passes wgt=None, ng.py:103), wgt weights the gridding.  model / dirty_out:
 * dirty_dtype in dirty_space; ms / wgt / ms_resid: vis_dtype in vis_space; ms_resid (the residual
 * visibilities) may be NULL, or equal to ms for an in-place update. */
int b2g_plan_residual(b2g_plan *plan, const void *model, int dirty_dtype, int dirty_space,
                      const void *ms, const void *wgt, int vis_dtype, int vis_space,
                      void *ms_resid, void *dirty_out);

/* Plan introspection: fills up to n entries of
 *   iinfo = {W, nu, nv, nplanes, nslabs, planes_per_slab, algo_used, poly_degree,
 *            nvis_sorted, key_bits, launches_last_call, hand_fft (1: the per-plane transform is
 *            the hand-written pruned FFT fused with the w-screen, 0: cuFFT + screen kernels)}
 *   dinfo = {sigma, eps_kernel, dw, wmin, wmax, beta, e0, t_plan_ms, t_last_ms} */
int b2g_plan_info(const b2g_plan *plan, int64_t *iinfo, int ni, double *dinfo, int nd);

/* Per-stage device times (ms, CUDA events on the plan's stream) of the last invert or
 * predict call: {h2d, prep, grid_or_degrid, fft, screen, correct, d2h, total}. */
int b2g_plan_timings(const b2g_plan *plan, double *ms_out, int n);

/* One-shot equivalents of the two ducc0 calls (plan + execute + destroy). */
int b2g_ms2dirty(int64_t nrow, int64_t nchan, const double *uvw, const double *freq,
                 const void *ms, const void *wgt, int vis_dtype, int space,
                 int64_t npix_x, int64_t npix_y, double pixsize_x, double pixsize_y,
                 int64_t nu, int64_t nv, double epsilon, int do_wstacking, int precision,
                 void *dirty, int dirty_dtype, void *stream);

int b2g_dirty2ms(int64_t nrow, int64_t nchan, const double *uvw, const double *freq,
                 const void *dirty, int dirty_dtype, const void *wgt, int vis_dtype,
                 int space, int64_t npix_x, int64_t npix_y, double pixsize_x,
                 double pixsize_y, int64_t nu, int64_t nv, double epsilon,
                 int do_wstacking, int precision, void *ms_out, void *stream);

/* Deterministic fp64 sum of n weights (fixed reduction tree; exact for 0/1 weights).
 * Mirrors numpy.sum(wgt) at ng.py:258, :289. */
int b2g_sumwt(const void *wgt, int dtype, int space, int64_t n, double *out, void *stream);

/* b2g_sumwt for callers that may not allocate device memory (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The 1024 block partials of the reduction live in a file-scope
 * __device__ array of the library (8 KiB, held by every GPU that loads the library until the
 * process exits; csrc/sum_static.cu) instead of coming from cudaMallocAsync.  The call allocates
 * nothing, copies nothing and does not synchronise: wgt (dtype B2G_F32 / B2G_F64, n entries) and
 * out (one double) are DEVICE pointers, and the two kernels are only enqueued on `stream`, on the
 * current device.  Same reduction tree as b2g_sumwt: the results agree bit for bit.
 * n > B2G_SUMWT_STATIC_MAX_N is rejected (non-zero return, message in b2g_last_error()).
 * All calls on one device SHARE the scratch array: calls that overlap on the GPU (different
 * streams) race; calls on one stream are ordered and safe.  Nothing needs clearing between
 * calls (every partial and *out are overwritten by each call). */
#define B2G_SUMWT_STATIC_MAX_N 4294967296ll /* 2^32 weights */
int b2g_sumwt_static(const void *wgt, int dtype, int64_t n, double *out, void *stream);

/* Plans created from HOST uvw stream the visibility arrays of b2g_plan_invert /
 * b2g_plan_predict over PCIe in row chunks of about `nsamples` samples, overlapped with
 * gridding / degridding (what a caller of ducc0.wgridder with host arrays gets, ng.py:240-256).
 * Default 2^24; 0 disables chunking.  Process-wide setting, applies to plans created later. */
int b2g_set_host_chunk_samples(int64_t nsamples);

/* Upper limit on the number of origin w-planes a plan grids per slab (0 = default: sized from
 * free device memory, at most 8 once the planes no longer fit at once).  With a small limit a
 * small problem runs the multi-slab ring-buffer path (slabs of `nplanes` origin planes in a
 * wrapping ring of nplanes + W - 1 slots) that the 8192^2 / 16384^2 configurations take; the
 * parity tests use it.  Process-wide setting, applies to plans created later. */
int b2g_set_max_slab_planes(int nplanes);

/* Tensor-core degridder (tcgen05 / TMEM, fp32 w-stacking plans with W <= 7): 0 = off, 1 = the
 * chunks of 1024 sorted samples whose runs of equal footprint origin are long (the sparse rest
 * stays with the warp-run kernel; the default), 2 = every chunk (parity tests).  Same results within fp32
 * rounding (3 x TF32 split).  Process-wide setting, applies to plans created later. */
int b2g_set_tc_mode(int mode);

/* Host-side kernel mathematics, exported so that tests can check them without a GPU.
 * b2g_kernel_select: choose (W, sigma) for epsilon; returns 0 and fills
 *   out = {W, sigma, beta, e0, eps_kernel}.
 * b2g_kernel_eval:  evaluate the piecewise-polynomial kernel of (W, beta, e0) at n
 *   points t in [-1,1].
 * b2g_kernel_corr:  Fourier transform psi-hat of that kernel at n points x (cycles/cell). */
int b2g_kernel_select(double epsilon, int do_wstacking, int precision, int64_t npix_x,
                      int64_t npix_y, double nvis, double *out5);
int b2g_kernel_eval(int W, double beta, double e0, const double *t, int n, double *out);
int b2g_kernel_corr(int W, double beta, double e0, const double *x, int n, double *out);

/*
 * Marshalling of the Visibility cube around the gridder, one streaming kernel per direction.
 * ALL pointers except polmat and lmn are DEVICE pointers; polmat / lmn are small host arrays.
 *
 * b2g_vis_prepare replaces the numpy passes of invert_ng before the ducc0 call
 * (ng.py:182-204, :231-233: phase rotation to the image centre [shift_vis_to_image,
 * imaging/base.py:48-92 -> visibility/base.py:27-45, :60-96], flagged_vis, flagged_imaging_weight,
 * convert_pol_frame, astype, transposes):
 *   vis     [nrow][nchan][npol_in] complex (vis_dtype), a chunk of rows starting at row0
 *   flags   same shape, type flag_dtype (B2G_FLAG_*), may be NULL;  factor = 1 - flags
 *   weight  same shape, real (wgt_dtype), may be NULL (all ones)
 *   polmat  [npol_out][npol_in] complex double (host): visibility frame -> image frame
 *   lmn     NULL (no rotation) or host {l, m, n-1} of the image centre relative to the
 *           visibility phase centre, signed for the uvw array passed here
 *   uvw     [nrow_total][3] metres, fscale [nchan] = f/c  (only read if lmn != NULL)
 *   ms_out  [npol_out][nrow_total][nchan] complex (out_dtype);  wgt_out same, real, may be NULL
 *   dopsf   pol 0 = 1, other pols = 0 (ng.py:231-233)
 *
 * b2g_vis_finish is the tail of predict_ng (ng.py:131-143): convert_pol_frame image frame ->
 * visibility frame, interleave to [nrow][nchan][npol_vis], inverse phase rotation.
 */
int b2g_vis_prepare(int64_t nrow, int64_t nchan, int npol_in, int npol_out, const void *vis,
                    int vis_dtype, const void *flags, int flag_dtype, const void *weight,
                    int wgt_dtype, const double *polmat, const double *lmn, const double *uvw,
                    const double *fscale, int64_t row0, int64_t nrow_total, int dopsf,
                    void *ms_out, void *wgt_out, int out_dtype, void *stream);

int b2g_vis_finish(int64_t nrow, int64_t nchan, int npol_img, int npol_vis, const void *ms,
                   int ms_dtype, const double *polmat, const double *lmn, const double *uvw,
                   const double *fscale, int64_t row0, int64_t nrow_total, void *vis_out,
                   int vis_dtype, void *stream);

/*
 * Density weighting and tapering of the imaging weights (the step before invert_ng).  ALL array
 * pointers are DEVICE pointers except sumwt_host / sumsq_host (host outputs).  Cubes are
 * [nrow][nchan][npol] float64 (flags: B2G_FLAG_* type, may be NULL), uvw [nrow][3] metres,
 * kscale [nchan] = freq / c, vis_to_im [nchan] int32 = image channel of each visibility channel.
 * The weight grid [nchan_img][npol][ny][nx] float64 has the image's shape and the linear UU / VV
 * axes of the reference's GridData: cell = round(u_lambda / cdelt + crpix - 1), crpix 1-based.
 *
 * b2g_weight_grid   replaces grid_visibility_weight_to_griddata
 *                   (/root/reference/src/ska_sdp_func_python/grid_data/gridding.py:258-334): zeroes
 *                   grid, adds weight*(1-flags) of every sample to its cell and to the cell of the
 *                   conjugate sample (both must be on the grid); sumwt_host [nchan_img][npol] =
 *                   2 * in-grid weight sum; sumsq_host (1 value, may be NULL) = sum(grid^2).
 * b2g_weight_apply  replaces griddata_visibility_reweight (gridding.py:362-499):
 *                   mode 0 uniform  imaging_weight = w / g;  mode 1 robust  w / (1 + f2 * g);
 *                   0 where g <= 0 or the sample is off the grid.  w = weight*(1-flags).
 * b2g_weight_taper  replaces taper_visibility_gaussian (mode 0, param = pi^2 beam^2 / (4 ln 2))
 *                   and taper_visibility_tukey (mode 1, param = tukey) of
 *                   imaging/weighting.py:68-140; wave [nchan] = c / freq; imaging_weight is
 *                   multiplied in place by (1-flags) * taper.
 */
int b2g_weight_grid(int64_t nrow, int64_t nchan, int npol, const double *uvw,
                    const double *kscale, const double *weight, const void *flags,
                    int flag_dtype, const int32_t *vis_to_im, int nchan_img, int64_t nx,
                    int64_t ny, double cdelt_u, double cdelt_v, double crpix_u, double crpix_v,
                    double *grid, double *sumwt_host, double *sumsq_host, void *stream);

int b2g_weight_apply(int64_t nrow, int64_t nchan, int npol, const double *uvw,
                     const double *kscale, const double *weight, const void *flags,
                     int flag_dtype, const int32_t *vis_to_im, int nchan_img, int64_t nx,
                     int64_t ny, double cdelt_u, double cdelt_v, double crpix_u, double crpix_v,
                     const double *grid, int mode, double f2, double *imaging_weight,
                     void *stream);

int b2g_weight_taper(int64_t nrow, int64_t nchan, int npol, const double *uvw,
                     const double *wave, const void *flags, int flag_dtype, int mode,
                     double param, double *imaging_weight, void *stream);

/*
 * Direct Fourier transform of sky components to visibilities: replaces dft_kernel /
 * dft_cpu_looped and the cupy kernel string of
 * /root/reference/src/ska_sdp_func_python/imaging/dft.py:121-183, :185-262, :265-285.
 *   vis[row][chan][pol] (+)= sum_comp flux[comp][fchan][pol] * exp(-2 pi i (u l + v m + w n'))
 * ALL pointers are DEVICE pointers.  uvw [nrow][3] metres with kscale [nchan] = freq / c, or --
 * kscale == NULL -- uvw [nrow*nchan][3] already in wavelengths (the reference's uvw_lambda);
 * dircos [ncomp][3] = (l, m, sqrt(1-l^2-m^2)-1); flux [ncomp][flux_nchan][npol] complex double
 * (flux_nchan = nchan, or 1 = same flux in every channel); vis_out [nrow][nchan][npol] complex
 * double.  trig_precision B2G_F64 (default) or B2G_F32 (fp32 sine / cosine after the exact fp64
 * range reduction); accumulate != 0 adds to vis_out instead of overwriting it.
 */
int b2g_dft_predict(int64_t nrow, int64_t nchan, int npol, int64_t ncomp, const double *uvw,
                    const double *kscale, const double *dircos, const void *flux,
                    int64_t flux_nchan, void *vis_out, int trig_precision, int accumulate,
                    void *stream);

/*
 * Hogbom CLEAN minor cycle on device-resident images: replaces hogbom()
 * (/root/reference/src/ska_sdp_func_python/image/cleaners.py:23-133).  ALL image pointers are
 * DEVICE pointers to C-contiguous float64 arrays: dirty / comps / res / window [nx][ny], psf
 * [px][py]; window may be NULL; res may equal dirty (in place).  pmax = psf.max() (the caller
 * checks that it is 1 as the reference asserts).  Every step is the reference's fp64 operation in
 * the reference's order, so comps and res are bit-exact; iters_host receives the number of
 * iterations performed.
 *
 * b2g_image_peak: mode 0: value and flat index of max |img * window| (first occurrence, as
 * numpy.argmax); mode 1: max img.  Host outputs.
 */
int b2g_hogbom(const double *dirty, const double *psf, const double *window, int64_t nx,
               int64_t ny, int64_t px, int64_t py, double gain, double pmax, double thresh,
               double fracthresh, int64_t niter, double *comps, double *res, int64_t *iters_host,
               void *stream);

int b2g_image_peak(const double *img, const double *window, int64_t nx, int64_t ny, int mode,
                   double *val_host, int64_t *idx_host, void *stream);

/*
 * Complex Hogbom CLEAN of a polarised image Q + iU: replaces hogbom_complex()
 * (/root/reference/src/ska_sdp_func_python/image/cleaners.py:136-232).  DEVICE pointers,
 * C-contiguous float64 [nx][ny] (psf [px][py], shared by Q and U as the reference asserts); window
 * may be NULL; res_* may equal dirty_* (in place).  pmax = psf.max().  numpy's complex arithmetic
 * is followed operation by operation (|z| = hypot, division of a complex by a real = product with
 * 1 / pmax), so components and residuals match the reference bit for bit.
 */
int b2g_hogbom_complex(const double *dirty_q, const double *dirty_u, const double *psf,
                       const double *window, int64_t nx, int64_t ny, int64_t px, int64_t py,
                       double gain, double pmax, double thresh, double fracthresh, int64_t niter,
                       double *comps_q, double *comps_u, double *res_q, double *res_u,
                       int64_t *iters_host, void *stream);

/*
 * Multi-scale CLEAN minor cycle: replaces the loop of msclean()
 * (/root/reference/src/ska_sdp_func_python/image/cleaners.py:398-451) with its peak search
 * find_max_abs_stack() (:565-610).  DEVICE pointers, C-contiguous float64:
 *   res_stack      [nscales][nx][ny]          scale-convolved residuals, updated in place
 *   psf_scalescale [nscales][nscales][px][py] doubly convolved PSFs
 *   psf_scale      [nscales][px][py]          the scale shapes on the PSF's raster
 *   window_stack   [nscales][nx][ny] or NULL, sensitivity [nx][ny] or NULL
 *   comps          [nx][ny]                   component image (zeroed here)
 * coupling_diag is a HOST array [nscales] (= max of psf_scalescale[s][s]); absolutethresh =
 * max(thresh, fracthresh * max|res_stack[0]|) (:380).  The arithmetic and its order are the
 * reference's, so for the same stacks the result is bit-exact; iters_host receives the number of
 * components subtracted.  nscales <= 16.
 */
int b2g_msclean(double *res_stack, const double *psf_scalescale, const double *psf_scale,
                const double *window_stack, const double *sensitivity, const double *coupling_diag,
                int64_t nscales, int64_t nx, int64_t ny, int64_t px, int64_t py, double gain,
                double absolutethresh, int64_t niter, double *comps, int64_t *iters_host,
                void *stream);

/*
 * Multi-scale multi-frequency CLEAN minor cycle: replaces the loop of msmfsclean()
 * (/root/reference/src/ska_sdp_func_python/image/cleaners.py:826-880) with find_global_optimum()
 * (:901-975, the "RASCIL" / "Algorithm1" peak finder), update_moment_model() (:1004-1031) and
 * update_scale_moment_residual() (:977-1001).  DEVICE pointers, C-contiguous float64:
 *   smres        [nscales][nmoment][nx][ny]                scale-convolved moment residuals, in place
 *   psf_moments  [nscales][nscales][2 nmoment - 1][px][py] [a][b][k] = PSF moment k convolved with
 *                                                          scale b and correlated with scale a
 *                                                          (= the reference's ssmmpsf[a][b][t][q], k = t + q)
 *   psf_scale    [nscales][px][py]                         the scale shapes on the PSF's raster
 *   window_stack [nscales][nx][ny] or NULL, sensitivity [nx][ny] or NULL
 *   model        [nmoment][nx][ny]                         moment model (zeroed here)
 * inv_hessian and hessian are HOST arrays [nscales][nmoment][nmoment] (:1081-1104; hessian may be
 * NULL unless findpeak == 1).  findpeak: 0 = "RASCIL" / "Algorithm1" (peak of the moment-0 principal
 * solution), 1 = "CASA" (peak of the chi^2 decrease, :936-960).  nscales <= 8, nmoment <= 4.
 * iters_host receives the number of components subtracted.
 */
int b2g_msmfsclean(double *smres, const double *psf_moments, const double *psf_scale,
                   const double *window_stack, const double *sensitivity, const double *inv_hessian,
                   const double *hessian, int findpeak, int64_t nscales, int64_t nmoment, int64_t nx,
                   int64_t ny, int64_t px, int64_t py, double gain, double absolutethresh,
                   int64_t niter, double *model, int64_t *iters_host, void *stream);

/*
 * Adjoint of b2g_dft_predict: replaces the sums of idft_visibility_skycomponent
 * (/root/reference/src/ska_sdp_func_python/imaging/dft.py:334-384).  DEVICE pointers.
 *   flux_out[comp][chan][pol] = sum_rows weight*(1-flags) * vis*(1-flags) * exp(+2 pi i (u l + v m + w n'))
 *   wsum_out[chan][pol]       = sum_rows weight*(1-flags)
 * vis [nrow][nchan][npol] complex double, weight / flags same shape (weight may be NULL = ones,
 * flags may be NULL), npol 1, 2 or 4.  The caller divides flux by wsum where wsum > 0.
 */
int b2g_dft_invert(int64_t nrow, int64_t nchan, int npol, int64_t ncomp, const double *uvw,
                   const double *kscale, const double *dircos, const void *vis,
                   const double *weight, const void *flags, int flag_dtype, void *flux_out,
                   double *wsum_out, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_WGRIDDER_H */
