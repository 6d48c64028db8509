// weight_kernels.cu -- density weighting and tapering of the imaging weights (SURVEY.md
// section 8f row N1: the step immediately before invert_ng).
//
// Replaces the interpreted per-row loops of
//   grid_visibility_weight_to_griddata   /root/reference/src/ska_sdp_func_python/grid_data/gridding.py:258-334
//   griddata_visibility_reweight         gridding.py:362-499   (uniform / robust, Briggs 3.15-3.16)
//   taper_visibility_gaussian / _tukey   imaging/weighting.py:68-140
// reached through weight_visibility (imaging/weighting.py:35-65).
//
// Cell mapping (spatial_mapping, gridding.py:146-157): the weight grid has the image's shape and
// a linear UU/VV WCS; a sample at (u, v) wavelengths lands in
//     pu = round(u / cdelt_u + crpix_u - 1),  pv = round(v / cdelt_v + crpix_v - 1)
// (numpy.round = round-half-even = rint) and its conjugate in the cell of (-u, -v).  A sample
// counts only if BOTH cells are inside the grid.  u, v in wavelengths are uvw * (freq / c), the
// product the reference forms (visibility_acc.uvw_lambda); NaN coordinates become 0
// (numpy.nan_to_num, gridding.py:53-55).
//
// All three passes are HBM / L2-atomic bound streaming kernels over the weight cube
// [nrow][nchan][npol] (fp64, as the reference): algorithmic bytes per sample = 8 (weight) +
// flag bytes for the histogram; 8 + flag bytes + 8 (imaging weight written) for the re-weighting.
#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>

#include "../../include/b200_wgridder.h"

namespace b2g {
int wrapper_fail(const char *msg);  // plan.cu
}

namespace {

constexpr unsigned kFull = 0xffffffffu;

struct WGeo {
    int nx, ny;               // grid (= image) size along u (fastest) and v
    double cdelt_u, cdelt_v;  // wavelengths per cell (cdelt_u < 0 for RA-decreasing images)
    double rcdelt_u, rcdelt_v;  // 1 / cdelt (fast path of cell_pair)
    double crpix_u, crpix_v;  // 1-based reference pixel (nx//2 + 1, ny//2 + 1)
    int nchan_img, npol;
};

__device__ __forceinline__ double flagf(const void *flags, int type, long long i)
{
    switch (type) {
    case B2G_FLAG_U8: return 1.0 - (double)static_cast<const uint8_t *>(flags)[i];
    case B2G_FLAG_I32: return 1.0 - (double)static_cast<const int32_t *>(flags)[i];
    case B2G_FLAG_I64: return 1.0 - (double)static_cast<const long long *>(flags)[i];
    case B2G_FLAG_F64: return 1.0 - static_cast<const double *>(flags)[i];
    default: return 1.0;
    }
}

// numpy.nan_to_num for one coordinate
__device__ __forceinline__ double nan_to_num(double x)
{
    if (x != x) return 0.0;
    if (x > DBL_MAX) return DBL_MAX;
    if (x < -DBL_MAX) return -DBL_MAX;
    return x;
}

// Cells of x and of -x along one axis: round(+-x / cdelt + crpix - 1) as the reference computes
// it (correctly rounded division, then the two additions, then round-half-even), -1 if outside
// [0, n).  (-x) / cdelt == -(x / cdelt) exactly, so one quotient serves both.  The quotient is
// first approximated by a multiplication with 1/cdelt (relative error < 4e-16, i.e. < 4e-7 cells
// for |q| < 1e9); only when the result lies within 1e-6 of a rounding boundary (where the
// approximation could change the cell) is the exact division evaluated -- the cell indices are
// therefore always those of the exact expression.  rint() is the add-and-subtract-2^52*1.5
// idiom (round-half-even, exact for |t| < 2^51).
__device__ __forceinline__ void cell_pair(double x, double cdelt, double rcdelt, double crpix, int n,
                                          int &p, int &pc)
{
    const double kMagic = 6755399441055744.0;
    const double c1 = crpix - 1.0;
    double q = __dmul_rn(x, rcdelt);
    const double t = __dadd_rn(q, c1), tc = __dadd_rn(c1, -q);
    double r = __dadd_rn(__dadd_rn(t, kMagic), -kMagic), rc = __dadd_rn(__dadd_rn(tc, kMagic), -kMagic);
    const double lim = 0.5 - 1e-6;
    if (!(fabs(t - r) < lim && fabs(tc - rc) < lim && fabs(q) < 1e9)) {
        q = __ddiv_rn(x, cdelt);
        r = rint(__dadd_rn(q, crpix) - 1.0);
        rc = rint(__dadd_rn(-q, crpix) - 1.0);
    }
    p = (r >= 0.0 && r < (double)n) ? (int)r : -1;
    pc = (rc >= 0.0 && rc < (double)n) ? (int)rc : -1;
}

// cell index pv*nx + pu of the sample and of its conjugate; false if either is off the grid
__device__ __forceinline__ bool map_sample(const WGeo &g, const double *uvw, const double *kscale,
                                           long long row, int chan, int &cell, int &cellc)
{
    const double k = kscale[chan];
    const double u = nan_to_num(__dmul_rn(uvw[3 * row + 0], k));
    const double v = nan_to_num(__dmul_rn(uvw[3 * row + 1], k));
    int pu, pv, puc, pvc;
    cell_pair(u, g.cdelt_u, g.rcdelt_u, g.crpix_u, g.nx, pu, puc);
    cell_pair(v, g.cdelt_v, g.rcdelt_v, g.crpix_v, g.ny, pv, pvc);
    if ((pu | pv | puc | pvc) < 0) return false;
    cell = pv * g.nx + pu;
    cellc = pvc * g.nx + puc;
    return true;
}

// flattened sample index -> (row, chan): exact division by the run-time constant nchan with a
// multiply-high (Granlund & Montgomery) whenever the index fits 32 bits
struct ChanDiv {
    unsigned d, mul;
    int sh1, sh2;
};

inline ChanDiv make_chandiv(unsigned d)
{
    ChanDiv c;
    c.d = d;
    int l = 0;
    while ((1ull << l) < d) ++l;
    c.mul = (unsigned)(((1ull << 32) * ((1ull << l) - d)) / d + 1);
    c.sh1 = l < 1 ? l : 1;
    c.sh2 = l > 1 ? l - 1 : 0;
    return c;
}

__device__ __forceinline__ void split_index(long long i, const ChanDiv &c, long long &row, int &chan)
{
    if (i < 0x100000000ll) {
        const unsigned x = (unsigned)i;
        const unsigned t = __umulhi(c.mul, x);
        const unsigned r = (t + ((x - t) >> c.sh1)) >> c.sh2;
        row = r;
        chan = (int)(x - r * c.d);
    } else {
        row = i / c.d;
        chan = (int)(i - row * c.d);
    }
}

// Histogram of the flagged weights.  Thread = (row, chan), lanes run along chan, so neighbouring
// lanes mostly hit the same cell: a full-warp segmented scan over runs of equal cells leaves one
// fp64 atomic per run instead of one per sample.
__global__ void __launch_bounds__(256)
weight_grid_kernel(WGeo g, long long nrow, int nchan, ChanDiv cd, const double *__restrict__ uvw,
                   const double *__restrict__ kscale, const double *__restrict__ weight,
                   const void *__restrict__ flags, int flag_type,
                   const int *__restrict__ vis_to_im, double *__restrict__ grid)
{
    const long long n = nrow * nchan;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    int cell = -1, cellc = -1, imchan = 0;
    bool ok = false;
    if (i < n) {
        long long row;
        int chan;
        split_index(i, cd, row, chan);
        imchan = vis_to_im[chan];
        ok = imchan >= 0 && imchan < g.nchan_img && map_sample(g, uvw, kscale, row, chan, cell, cellc);
    }
    // run key: plane and cell (the conjugate cell is a function of the same (u, v), but two
    // samples can share a cell without sharing the conjugate cell at the rounding boundary)
    const long long key = ok ? (((long long)imchan * g.ny * g.nx + cell) << 1) | 1 : -(long long)lane - 2;
    const long long keyc = ok ? cellc : -1;
    const long long kp = __shfl_up_sync(kFull, key, 1), kcp = __shfl_up_sync(kFull, keyc, 1);
    const bool head = lane == 0 || kp != key || kcp != keyc;
    const unsigned heads = __ballot_sync(kFull, head);
    // first lane of my run / am I its last lane
    const unsigned below = heads & (lane == 31 ? kFull : ((2u << lane) - 1u));
    const int first = 31 - __clz(below);
    const unsigned above = lane == 31 ? 0u : heads >> (lane + 1);
    const bool tail = lane == 31 || (above & 1u);
    const size_t plane = (size_t)g.ny * g.nx;
    for (int pol = 0; pol < g.npol; ++pol) {
        double w = 0.0;
        if (ok) {
            const long long e = i * g.npol + pol;
            w = weight[e] * flagf(flags, flag_type, e);
        }
        // inclusive segmented scan (runs are contiguous in lane order)
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const double t = __shfl_up_sync(kFull, w, d);
            if (lane - d >= first) w += t;
        }
        if (ok && tail && w != 0.0) {
            double *p = grid + ((size_t)imchan * g.npol + pol) * plane;
            atomicAdd(p + cell, w);
            atomicAdd(p + cellc, w);
        }
    }
}

// imaging_weight = f(weight * (1 - flags), gridded weight at the sample's cell)
//   mode 0 uniform: w / g          mode 1 robust: w / (1 + f2 * g)        (g > 0, in grid)
//   0 where g <= 0 or the sample (or its conjugate) is off the grid   gridding.py:449-493
__global__ void __launch_bounds__(256)
weight_apply_kernel(WGeo g, long long nrow, int nchan, ChanDiv cd, const double *__restrict__ uvw,
                    const double *__restrict__ kscale, const double *__restrict__ weight,
                    const void *__restrict__ flags, int flag_type,
                    const int *__restrict__ vis_to_im, const double *__restrict__ grid, int mode,
                    double f2, double *__restrict__ imaging_weight)
{
    const long long n = nrow * nchan;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    long long row;
    int chan;
    split_index(i, cd, row, chan);
    const int imchan = vis_to_im[chan];
    int cell = -1, cellc = -1;
    const bool ok = imchan >= 0 && imchan < g.nchan_img && map_sample(g, uvw, kscale, row, chan, cell, cellc);
    const size_t plane = (size_t)g.ny * g.nx;
    for (int pol = 0; pol < g.npol; ++pol) {
        const long long e = i * g.npol + pol;
        double out = 0.0;
        if (ok) {
            const double gd = grid[((size_t)imchan * g.npol + pol) * plane + cell];
            if (gd > 0.0) {
                const double w = weight[e] * flagf(flags, flag_type, e);
                out = mode == 0 ? __ddiv_rn(w, gd) : __ddiv_rn(w, __dadd_rn(1.0, __dmul_rn(f2, gd)));
            }
        }
        imaging_weight[e] = out;
    }
}

// mode 0 Gaussian: imw = imw*(1-flags) * exp(-scale * (u^2 + v^2) / wave^2)   weighting.py:88-99
// mode 1 Tukey:    imw = imw*(1-flags) * tukey((d / wave) / (dmax / wave), r)  weighting.py:121-138
__global__ void __launch_bounds__(256)
weight_taper_kernel(long long nrow, int nchan, ChanDiv cd, int npol, const double *__restrict__ uvw,
                    const double *__restrict__ wave, const void *__restrict__ flags, int flag_type,
                    int mode, double param, double dmax, double *__restrict__ imaging_weight)
{
    const long long n = nrow * nchan;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    long long row;
    int chan;
    split_index(i, cd, row, chan);
    const double u = uvw[3 * row], v = uvw[3 * row + 1];
    const double wv = wave[chan];
    const double d2 = __dadd_rn(__dmul_rn(u, u), __dmul_rn(v, v));
    double wt;
    if (mode == 0) {
        const double q = __ddiv_rn(d2, __dmul_rn(wv, wv));
        wt = exp(__dmul_rn(-param, q));
    } else {
        const double x = __ddiv_rn(__ddiv_rn(sqrt(d2), wv), __ddiv_rn(dmax, wv));
        const double r = param;
        const double twopi = 2.0 * 3.14159265358979323846;
        if (x >= 0.0 && x < r / 2.0)
            wt = 0.5 * (1.0 + cos(twopi * (x - r / 2.0) / r));
        else if (x >= 1.0 - r / 2.0 && x <= 1.0)
            wt = 0.5 * (1.0 + cos(twopi * (x - 1.0 + r / 2.0) / r));
        else
            wt = 1.0;
    }
    for (int pol = 0; pol < npol; ++pol) {
        const long long e = i * npol + pol;
        imaging_weight[e] = __dmul_rn(__dmul_rn(imaging_weight[e], flagf(flags, flag_type, e)), wt);
    }
}

// ---- deterministic reductions (fixed tree: 1024 blocks x 256 threads, then one block) ---------
// op 0: sum x      op 1: sum x^2      op 2: max sqrt(x[3i]^2 + x[3i+1]^2) over rows
constexpr int kRedBlocks = 1024;

__device__ __forceinline__ double red_combine(int op, double a, double b)
{
    if (op != 2) return a + b;
    return (a != a || b != b) ? a + b : fmax(a, b);  // numpy.max propagates NaN
}

__device__ double block_reduce(int op, double v)
{
    __shared__ double sm[8];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = red_combine(op, v, __shfl_xor_sync(kFull, v, d));
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    double r = sm[0];
    for (int k = 1; k < (int)(blockDim.x >> 5); ++k) r = red_combine(op, r, sm[k]);
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(256)
reduce_stage1(int op, const double *__restrict__ x, long long n, long long seg_len,
              double *__restrict__ partial)
{
    // segment blockIdx.y (e.g. one grid plane) of length seg_len, kRedBlocks partials per segment
    const double *xs = x + (size_t)blockIdx.y * (size_t)(op == 2 ? 0 : seg_len);
    const long long len = op == 2 ? n : seg_len;
    double v = 0.0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < len;
         i += (long long)gridDim.x * blockDim.x) {
        if (op == 0) v += xs[i];
        else if (op == 1) v += xs[i] * xs[i];
        else v = red_combine(2, v, sqrt(xs[3 * i] * xs[3 * i] + xs[3 * i + 1] * xs[3 * i + 1]));
    }
    v = block_reduce(op, v);
    if (threadIdx.x == 0) partial[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = v;
}

__global__ void __launch_bounds__(256)
reduce_stage2(int op, const double *__restrict__ partial, int nper, double *__restrict__ out)
{
    const double *p = partial + (size_t)blockIdx.x * nper;
    double v = 0.0;
    for (int i = threadIdx.x; i < nper; i += blockDim.x) v = red_combine(op, v, p[i]);
    v = block_reduce(op, v);
    if (threadIdx.x == 0) out[blockIdx.x] = v;
}

int check_launch()
{
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : b2g::wrapper_fail(cudaGetErrorString(e));
}

WGeo make_geo(int64_t nx, int64_t ny, double cdelt_u, double cdelt_v, double crpix_u, double crpix_v,
              int nchan_img, int npol)
{
    WGeo g;
    g.nx = (int)nx;
    g.ny = (int)ny;
    g.cdelt_u = cdelt_u;
    g.cdelt_v = cdelt_v;
    g.rcdelt_u = 1.0 / cdelt_u;
    g.rcdelt_v = 1.0 / cdelt_v;
    g.crpix_u = crpix_u;
    g.crpix_v = crpix_v;
    g.nchan_img = nchan_img;
    g.npol = npol;
    return g;
}

}  // namespace

// Device-side reduction of `nseg` segments; result copied to the host array `out` (nseg doubles).
static int reduce_to_host(int op, const double *x, long long n, long long seg_len, int nseg,
                          double *out_host, cudaStream_t s)
{
    double *scratch = nullptr;
    cudaError_t e = cudaMallocAsync(&scratch, sizeof(double) * ((size_t)nseg * kRedBlocks + nseg), s);
    if (e != cudaSuccess) return b2g::wrapper_fail(cudaGetErrorString(e));
    double *res = scratch + (size_t)nseg * kRedBlocks;
    reduce_stage1<<<dim3(kRedBlocks, nseg), 256, 0, s>>>(op, x, n, seg_len, scratch);
    reduce_stage2<<<nseg, 256, 0, s>>>(op, scratch, kRedBlocks, res);
    e = cudaMemcpyAsync(out_host, res, sizeof(double) * nseg, cudaMemcpyDeviceToHost, s);
    cudaFreeAsync(scratch, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e == cudaSuccess) e = cudaGetLastError();
    return e == cudaSuccess ? 0 : b2g::wrapper_fail(cudaGetErrorString(e));
}

extern "C" int b2g_weight_grid(int64_t nrow, int64_t nchan, int npol, const double *uvw,
                               const double *kscale, const double *weight, const void *flags,
                               int flag_dtype, const int32_t *vis_to_im, int nchan_img, int64_t nx,
                               int64_t ny, double cdelt_u, double cdelt_v, double crpix_u,
                               double crpix_v, double *grid, double *sumwt_host,
                               double *sumsq_host, void *stream)
{
    if (npol < 1 || npol > 4 || nchan_img < 1 || nx < 1 || ny < 1 || nx * ny > 0x3fffffff)
        return b2g::wrapper_fail("b2g_weight_grid: bad shape");
    if (!grid || !kscale || !vis_to_im || (nrow * nchan > 0 && (!uvw || !weight)))
        return b2g::wrapper_fail("b2g_weight_grid: NULL pointer");
    if (cdelt_u == 0.0 || cdelt_v == 0.0) return b2g::wrapper_fail("b2g_weight_grid: zero cell size");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return b2g::wrapper_fail("b2g_weight_grid: no CUDA device available (this library has no CPU path)");
    cudaStream_t s = (cudaStream_t)stream;
    const WGeo g = make_geo(nx, ny, cdelt_u, cdelt_v, crpix_u, crpix_v, nchan_img, npol);
    const size_t cells = (size_t)nchan_img * npol * nx * ny;
    cudaError_t e = cudaMemsetAsync(grid, 0, cells * sizeof(double), s);
    if (e != cudaSuccess) return b2g::wrapper_fail(cudaGetErrorString(e));
    const long long n = nrow * nchan;
    if (n > 0) {
        const unsigned blocks = (unsigned)((n + 255) / 256);
        weight_grid_kernel<<<blocks, 256, 0, s>>>(g, nrow, (int)nchan, make_chandiv((unsigned)nchan), uvw, kscale, weight, flags,
                                                  flags ? flag_dtype : B2G_FLAG_NONE, vis_to_im, grid);
        if (int rc = check_launch()) return rc;
    }
    // sumwt[imchan][pol] = 2 * (sum of the in-grid weights) = sum of the plane (every sample adds
    // its weight to two cells): gridding.py:318-324
    if (sumwt_host)
        if (int rc = reduce_to_host(0, grid, 0, nx * ny, nchan_img * npol, sumwt_host, s)) return rc;
    if (sumsq_host)
        if (int rc = reduce_to_host(1, grid, 0, (long long)cells, 1, sumsq_host, s)) return rc;
    return 0;
}

extern "C" int b2g_weight_apply(int64_t nrow, int64_t nchan, int npol, const double *uvw,
                                const double *kscale, const double *weight, const void *flags,
                                int flag_dtype, const int32_t *vis_to_im, int nchan_img, int64_t nx,
                                int64_t ny, double cdelt_u, double cdelt_v, double crpix_u,
                                double crpix_v, const double *grid, int mode, double f2,
                                double *imaging_weight, void *stream)
{
    if (npol < 1 || npol > 4 || nchan_img < 1 || nx < 1 || ny < 1 || nx * ny > 0x3fffffff)
        return b2g::wrapper_fail("b2g_weight_apply: bad shape");
    if (mode != 0 && mode != 1) return b2g::wrapper_fail("b2g_weight_apply: mode must be 0 (uniform) or 1 (robust)");
    if (nrow * nchan > 0 && (!uvw || !kscale || !weight || !vis_to_im || !grid || !imaging_weight))
        return b2g::wrapper_fail("b2g_weight_apply: NULL pointer");
    if (cdelt_u == 0.0 || cdelt_v == 0.0) return b2g::wrapper_fail("b2g_weight_apply: zero cell size");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return b2g::wrapper_fail("b2g_weight_apply: no CUDA device available (this library has no CPU path)");
    const long long n = nrow * nchan;
    if (n <= 0) return 0;
    const WGeo g = make_geo(nx, ny, cdelt_u, cdelt_v, crpix_u, crpix_v, nchan_img, npol);
    const unsigned blocks = (unsigned)((n + 255) / 256);
    weight_apply_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(
        g, nrow, (int)nchan, make_chandiv((unsigned)nchan), uvw, kscale, weight, flags,
        flags ? flag_dtype : B2G_FLAG_NONE, vis_to_im, grid, mode, f2, imaging_weight);
    return check_launch();
}

extern "C" int b2g_weight_taper(int64_t nrow, int64_t nchan, int npol, const double *uvw,
                                const double *wave, const void *flags, int flag_dtype, int mode,
                                double param, double *imaging_weight, void *stream)
{
    if (npol < 1 || npol > 4) return b2g::wrapper_fail("b2g_weight_taper: bad npol");
    if (mode != 0 && mode != 1) return b2g::wrapper_fail("b2g_weight_taper: mode must be 0 (gaussian) or 1 (tukey)");
    if (nrow * nchan > 0 && (!uvw || !wave || !imaging_weight))
        return b2g::wrapper_fail("b2g_weight_taper: NULL pointer");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return b2g::wrapper_fail("b2g_weight_taper: no CUDA device available (this library has no CPU path)");
    const long long n = nrow * nchan;
    if (n <= 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    double dmax = 1.0;
    if (mode == 1)
        if (int rc = reduce_to_host(2, uvw, nrow, 0, 1, &dmax, s)) return rc;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    weight_taper_kernel<<<blocks, 256, 0, s>>>(nrow, (int)nchan, make_chandiv((unsigned)nchan), npol, uvw, wave, flags,
                                               flags ? flag_dtype : B2G_FLAG_NONE, mode, param, dmax,
                                               imaging_weight);
    return check_launch();
}
