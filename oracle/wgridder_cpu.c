/*
 * oracle/wgridder_cpu.c -- TEST INFRASTRUCTURE / CPU BASELINE ONLY (never linked into the
 * product; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline and
 * --impl reference legs may load it).
 *
 * Multi-threaded (OpenMP) fp64 CPU restatement of the algorithm the reference runs for this
 * path: the w-stacking ES-kernel gridder / degridder of ducc0.wgridder.ms2dirty / dirty2ms
 * (third-party, pinned ducc0==0.27.0: /root/reference/pyproject.toml:18, poetry.lock:305-306;
 * call sites /root/reference/src/ska_sdp_func_python/imaging/ng.py:99-112, :117-129,
 * :240-256, :271-287).  ducc0's source is not under /root/reference and not installable
 * here, so this follows the PUBLISHED algorithm (Arras, Reinecke, Westermann & Ensslin 2021,
 * A&A 646 A58, sect. 3; ES kernel of Barnett, Magland & af Klinteberg 2019):
 *
 *   1. scan: u,v,w -> wavelengths per channel, (u,v,w) -> -(u,v,w) + conjugate for w < 0,
 *      plane spacing dw = 1/(2 sigma max|n-1|), nplanes = ceil((wmax-wmin)/dw) + W;
 *      bucket the samples by (first w-plane, u-strip)                    [wg_bucket]
 *   2. per w-plane: grid every sample whose w-support touches the plane with the separable
 *      kernel phi(u) phi(v) phi(w) (piecewise polynomial, Horner)        [wg_grid_plane]
 *      -> 2-D FFT (done by the Python driver with scipy.fft = pocketfft, the same FFT
 *      family ducc0 uses) -> crop, multiply by the w-screen exp(-2 pi i w_p (n-1)),
 *      accumulate                                                        [wg_screen_accumulate]
 *   3. divide by the kernel's Fourier transform along u, v, w and by n   (Python driver)
 *   dirty2ms is the adjoint with the stages reversed        [wg_screen_apply, wg_degrid_plane]
 *
 * Parity status: checked against the direct-DFT oracle (oracle/dft_oracle.c) in
 * tests/test_oracle.py; "parity unpinned" against ducc0 output itself.
 * Threads: strips of grid rows are processed in two colours (even / odd strips), so that
 * concurrent threads never touch the same grid row and no locks or atomics are needed.
 */
#define _GNU_SOURCE
#include <math.h>
#include <omp.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define WG_MAXW 16
#define WG_MAXDEG 20
#define SPEED_OF_LIGHT 299792458.0

typedef struct {
    int64_t nrow, nchan;
    const double *uvw;   /* [nrow][3] metres */
    const double *freq;  /* [nchan] Hz */
    int64_t nu, nv;      /* padded grid */
    double psx, psy;     /* pixel sizes (rad) */
    int W, deg, wstack, nplanes;
    double w0, dw;       /* plane p at w = w0 + p dw (wavelengths) */
    double coef[WG_MAXW][WG_MAXDEG + 1]; /* piece s, highest power first, variable xi in [-1,1] */
    int nstrips, strip_h;
    /* bucketed sample index: bucket b = p0 * nstrips + strip */
    uint32_t *sidx;
    int64_t *bstart;     /* [nplanes*nstrips + 1] */
} wg_ctx;

static inline void kernel_row(const wg_ctx *c, double delta, double *out)
{
    const double xi = 2.0 * delta - 1.0;
    for (int s = 0; s < c->W; ++s) {
        const double *p = c->coef[s];
        double r = p[0];
        for (int d = 1; d <= c->deg; ++d) r = r * xi + p[d];
        out[s] = r;
    }
}
static inline double kernel_one(const wg_ctx *c, int s, double delta)
{
    const double xi = 2.0 * delta - 1.0;
    const double *p = c->coef[s];
    double r = p[0];
    for (int d = 1; d <= c->deg; ++d) r = r * xi + p[d];
    return r;
}

typedef struct { int iu0, iv0, p0, flip; double du, dv, dwf; } wg_loc;

static inline void locate_axis(double cyc, int64_t n, double halfW, int *i0, double *delta)
{
    double f = cyc - rint(cyc);
    double s = f * (double)n - halfW;
    double c = ceil(s);
    *delta = c - s;
    int64_t i = (int64_t)c;
    if (i < 0) i += n;
    if (i >= n) i -= n;
    *i0 = (int)i;
}

static inline wg_loc locate(const wg_ctx *c, int64_t row, int64_t chan)
{
    wg_loc L;
    const double halfW = 0.5 * c->W;
    const double fs = c->freq[chan] / SPEED_OF_LIGHT;
    double u = c->uvw[3 * row] * fs, v = c->uvw[3 * row + 1] * fs, w = c->uvw[3 * row + 2] * fs;
    L.flip = 0;
    L.p0 = 0;
    L.dwf = 0.0;
    if (c->wstack) {
        if (w < 0.0) { u = -u; v = -v; w = -w; L.flip = 1; }
        double s = (w - c->w0) / c->dw - halfW;
        double cs = ceil(s);
        int p0 = (int)cs;
        int pmax = c->nplanes - c->W;
        if (p0 < 0) p0 = 0;
        if (p0 > pmax) p0 = pmax;
        L.p0 = p0;
        L.dwf = (double)p0 - s;
    }
    locate_axis(u * c->psx, c->nu, halfW, &L.iu0, &L.du);
    locate_axis(v * c->psy, c->nv, halfW, &L.iv0, &L.dv);
    return L;
}

wg_ctx *wg_create(int64_t nrow, int64_t nchan, const double *uvw, const double *freq,
                  int64_t nu, int64_t nv, double psx, double psy, int W, int deg,
                  const double *coef /* [W][deg+1] */, int wstack, int nplanes, double w0,
                  double dw, int nthreads)
{
    if (W < 2 || W > WG_MAXW || deg > WG_MAXDEG) return NULL;
    wg_ctx *c = (wg_ctx *)calloc(1, sizeof(wg_ctx));
    c->nrow = nrow; c->nchan = nchan; c->uvw = uvw; c->freq = freq;
    c->nu = nu; c->nv = nv; c->psx = psx; c->psy = psy;
    c->W = W; c->deg = deg; c->wstack = wstack; c->nplanes = wstack ? nplanes : 1;
    c->w0 = w0; c->dw = dw;
    for (int s = 0; s < W; ++s)
        for (int d = 0; d <= deg; ++d) c->coef[s][d] = coef[s * (deg + 1) + d];
    /* strips: height >= max(W, 8), an even count, about 8 per thread */
    int minh = W > 8 ? W : 8;
    int ns = (int)(nu / minh);
    if (ns > 8 * nthreads) ns = 8 * nthreads;
    ns &= ~1;
    if (ns < 2) ns = 1;  /* tiny grid: single strip, processed serially */
    c->nstrips = ns;
    c->strip_h = (int)(nu / ns);
    return c;
}

void wg_destroy(wg_ctx *c)
{
    if (!c) return;
    free(c->sidx);
    free(c->bstart);
    free(c);
}

static inline int strip_of(const wg_ctx *c, int iu0)
{
    int s = iu0 / c->strip_h;
    return s >= c->nstrips ? c->nstrips - 1 : s;
}

/* Counting sort of all nrow*nchan samples into (p0, strip) buckets; samples with
 * keep[i]==0 (zero weight / zero visibility) are dropped.  keep may be NULL. */
int wg_bucket(wg_ctx *c, const uint8_t *keep, int nthreads)
{
    const int64_t n = c->nrow * c->nchan;
    const int64_t nb = (int64_t)c->nplanes * c->nstrips;
    if (n >= ((int64_t)1 << 32)) return 1;
    free(c->sidx);
    free(c->bstart);
    c->bstart = (int64_t *)calloc((size_t)nb + 1, sizeof(int64_t));
    int64_t *cnt = (int64_t *)calloc((size_t)nb * nthreads, sizeof(int64_t));
    uint32_t *bk = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(n > 0 ? n : 1));
    if (!c->bstart || !cnt || !bk) return 2;
#pragma omp parallel num_threads(nthreads)
    {
        const int t = omp_get_thread_num();
        const int64_t lo = n * t / nthreads, hi = n * (t + 1) / nthreads;
        int64_t *my = cnt + (size_t)t * nb;
        for (int64_t i = lo; i < hi; ++i) {
            if (keep && !keep[i]) { bk[i] = UINT32_MAX; continue; }
            wg_loc L = locate(c, i / c->nchan, i % c->nchan);
            uint32_t b = (uint32_t)(L.p0 * c->nstrips + strip_of(c, L.iu0));
            bk[i] = b;
            my[b]++;
        }
    }
    /* exclusive scan in (bucket, thread) order -> stable */
    int64_t run = 0;
    for (int64_t b = 0; b < nb; ++b) {
        c->bstart[b] = run;
        for (int t = 0; t < nthreads; ++t) {
            int64_t k = cnt[(size_t)t * nb + b];
            cnt[(size_t)t * nb + b] = run;
            run += k;
        }
    }
    c->bstart[nb] = run;
    c->sidx = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(run > 0 ? run : 1));
    if (!c->sidx) return 2;
#pragma omp parallel num_threads(nthreads)
    {
        const int t = omp_get_thread_num();
        const int64_t lo = n * t / nthreads, hi = n * (t + 1) / nthreads;
        int64_t *my = cnt + (size_t)t * nb;
        for (int64_t i = lo; i < hi; ++i) {
            uint32_t b = bk[i];
            if (b == UINT32_MAX) continue;
            c->sidx[my[b]++] = (uint32_t)i;
        }
    }
    free(cnt);
    free(bk);
    return 0;
}

int64_t wg_nkept(const wg_ctx *c) { return c->bstart ? c->bstart[(int64_t)c->nplanes * c->nstrips] : 0; }

/* grid[nu][nv] (interleaved complex, zeroed here) += all samples touching plane p.
 * ms interleaved complex [nrow][nchan]; wgt [nrow][nchan] or NULL. */
void wg_grid_plane(const wg_ctx *c, int p, const double *ms, const double *wgt, double *grid,
                   int nthreads)
{
    const int W = c->W;
    const int64_t nu = c->nu, nv = c->nv;
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int64_t i = 0; i < nu; ++i) memset(grid + 2 * i * nv, 0, sizeof(double) * 2 * nv);
    const int ncol = c->nstrips > 1 ? 2 : 1;
    for (int colour = 0; colour < ncol; ++colour) {
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 1)
        for (int st = colour; st < c->nstrips; st += ncol) {
            double ku[WG_MAXW], kv[WG_MAXW];
            for (int q = 0; q < (c->wstack ? W : 1); ++q) {
                const int p0 = p - q;
                if (p0 < 0 || p0 > c->nplanes - (c->wstack ? W : 1)) continue;
                const int64_t b = (int64_t)p0 * c->nstrips + st;
                for (int64_t k = c->bstart[b]; k < c->bstart[b + 1]; ++k) {
                    const int64_t i = c->sidx[k];
                    const wg_loc L = locate(c, i / c->nchan, i % c->nchan);
                    double re = ms[2 * i], im = ms[2 * i + 1];
                    if (wgt) { re *= wgt[i]; im *= wgt[i]; }
                    if (L.flip) im = -im;
                    if (c->wstack) {
                        const double kw = kernel_one(c, q, L.dwf);
                        re *= kw; im *= kw;
                    }
                    kernel_row(c, L.du, ku);
                    kernel_row(c, L.dv, kv);
                    for (int a = 0; a < W; ++a) {
                        int64_t iu = L.iu0 + a;
                        if (iu >= nu) iu -= nu;
                        double *rowp = grid + 2 * iu * nv;
                        const double fr = re * ku[a], fi = im * ku[a];
                        if (L.iv0 + W <= nv) {
                            double *cp = rowp + 2 * L.iv0;
                            for (int bb = 0; bb < W; ++bb) {
                                cp[2 * bb] += fr * kv[bb];
                                cp[2 * bb + 1] += fi * kv[bb];
                            }
                        } else {
                            for (int bb = 0; bb < W; ++bb) {
                                int64_t iv = L.iv0 + bb;
                                if (iv >= nv) iv -= nv;
                                rowp[2 * iv] += fr * kv[bb];
                                rowp[2 * iv + 1] += fi * kv[bb];
                            }
                        }
                    }
                }
            }
        }
    }
}

/* ms_out[i] += contribution of plane p (ms_out must be zeroed by the caller before the first
 * plane); the weight and the conjugation for flipped samples are applied by wg_finish_ms. */
void wg_degrid_plane(const wg_ctx *c, int p, const double *grid, double *ms_out, int nthreads)
{
    const int W = c->W;
    const int64_t nu = c->nu, nv = c->nv;
    /* reads only: all strips in parallel; each sample is written by exactly one (q, strip) */
    for (int q = 0; q < (c->wstack ? W : 1); ++q) {
        const int p0 = p - q;
        if (p0 < 0 || p0 > c->nplanes - (c->wstack ? W : 1)) continue;
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 1)
        for (int st = 0; st < c->nstrips; ++st) {
            double ku[WG_MAXW], kv[WG_MAXW];
            const int64_t b = (int64_t)p0 * c->nstrips + st;
            for (int64_t k = c->bstart[b]; k < c->bstart[b + 1]; ++k) {
                const int64_t i = c->sidx[k];
                const wg_loc L = locate(c, i / c->nchan, i % c->nchan);
                kernel_row(c, L.du, ku);
                kernel_row(c, L.dv, kv);
                double sr = 0.0, si = 0.0;
                for (int a = 0; a < W; ++a) {
                    int64_t iu = L.iu0 + a;
                    if (iu >= nu) iu -= nu;
                    const double *rowp = grid + 2 * iu * nv;
                    double rr = 0.0, ri = 0.0;
                    for (int bb = 0; bb < W; ++bb) {
                        int64_t iv = L.iv0 + bb;
                        if (iv >= nv) iv -= nv;
                        rr += rowp[2 * iv] * kv[bb];
                        ri += rowp[2 * iv + 1] * kv[bb];
                    }
                    sr += rr * ku[a];
                    si += ri * ku[a];
                }
                const double kw = c->wstack ? kernel_one(c, q, L.dwf) : 1.0;
                ms_out[2 * i] += sr * kw;
                ms_out[2 * i + 1] += si * kw;
            }
        }
    }
}

/* conjugate flipped samples, apply weights */
void wg_finish_ms(const wg_ctx *c, const double *wgt, double *ms_out, int nthreads)
{
    const int64_t n = c->nrow * c->nchan;
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        if (c->wstack && c->uvw[3 * (i / c->nchan) + 2] < 0.0) ms_out[2 * i + 1] = -ms_out[2 * i + 1];
        if (wgt) { ms_out[2 * i] *= wgt[i]; ms_out[2 * i + 1] *= wgt[i]; }
    }
}

static inline double nm1_of(double r2) { return -r2 / (sqrt(1.0 - r2) + 1.0); }

/* acc[a][b] += Re( fftgrid[a'][b'] * exp(-2 pi i w (n-1)) ) over the image footprint */
void wg_screen_accumulate(const wg_ctx *c, const double *fftgrid, int64_t npix_x, int64_t npix_y,
                          double w, double *acc, int nthreads)
{
    const int64_t nu = c->nu, nv = c->nv;
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int64_t a = 0; a < npix_x; ++a) {
        int64_t ia = a - npix_x / 2;
        if (ia < 0) ia += nu;
        const double x = (double)(a - npix_x / 2) * c->psx;
        for (int64_t b = 0; b < npix_y; ++b) {
            int64_t ib = b - npix_y / 2;
            if (ib < 0) ib += nv;
            const double re = fftgrid[2 * (ia * nv + ib)], im = fftgrid[2 * (ia * nv + ib) + 1];
            if (!c->wstack) { acc[a * npix_y + b] += re; continue; }
            const double y = (double)(b - npix_y / 2) * c->psy;
            const double r2 = x * x + y * y;
            if (r2 >= 1.0) continue;
            const double ph = -2.0 * M_PI * w * nm1_of(r2);
            acc[a * npix_y + b] += re * cos(ph) - im * sin(ph);
        }
    }
}

/* grid = zero-padded pre[a][b] * exp(+2 pi i w (n-1)) */
void wg_screen_apply(const wg_ctx *c, const double *pre, int64_t npix_x, int64_t npix_y, double w,
                     double *grid, int nthreads)
{
    const int64_t nu = c->nu, nv = c->nv;
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int64_t i = 0; i < nu; ++i) memset(grid + 2 * i * nv, 0, sizeof(double) * 2 * nv);
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int64_t a = 0; a < npix_x; ++a) {
        int64_t ia = a - npix_x / 2;
        if (ia < 0) ia += nu;
        const double x = (double)(a - npix_x / 2) * c->psx;
        for (int64_t b = 0; b < npix_y; ++b) {
            int64_t ib = b - npix_y / 2;
            if (ib < 0) ib += nv;
            const double val = pre[a * npix_y + b];
            double re = val, im = 0.0;
            if (c->wstack) {
                const double y = (double)(b - npix_y / 2) * c->psy;
                const double r2 = x * x + y * y;
                if (r2 >= 1.0) { re = 0.0; }
                else {
                    const double ph = 2.0 * M_PI * w * nm1_of(r2);
                    re = val * cos(ph);
                    im = val * sin(ph);
                }
            }
            grid[2 * (ia * nv + ib)] = re;
            grid[2 * (ia * nv + ib) + 1] = im;
        }
    }
}
