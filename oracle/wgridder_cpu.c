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
 * Threads: samples are bucketed by (first plane, 16-row strip, 32-cell v-tile); buckets are cut
 * into chunks; a thread grids one chunk into a private tile and adds the tile to the shared
 * grid under per-strip locks (the scheme ducc0 uses, with its tile size 16).
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
    double coef[WG_MAXDEG + 1][WG_MAXW] __attribute__((aligned(64))); /* [power (highest first)][piece] */
    int nstrips, strip_h;
    int nvt, vt_w;       /* v-tiles per strip and their width (cells) */
    /* bucketed sample index: bucket b = (p0 * nstrips + strip) * nvt + vtile */
    uint32_t *sidx;
    int64_t *bstart;     /* [nbuckets + 1] */
    int64_t *chunk_lo;   /* [nchunks + 1] first sorted sample of each chunk */
    int32_t *chunk_bucket;
    int64_t *bucket_chunk0; /* [nbuckets + 1] first chunk of each bucket */
    omp_lock_t *locks;   /* one per strip */
} wg_ctx;

/* all W pieces at once: the inner loop runs over pieces and vectorises (AVX2) */
static inline void kernel_row(const wg_ctx *c, double delta, double *restrict out)
{
    const double xi = 2.0 * delta - 1.0;
    double r[WG_MAXW] __attribute__((aligned(64)));
    for (int s = 0; s < WG_MAXW; ++s) r[s] = c->coef[0][s];
    for (int d = 1; d <= c->deg; ++d)
        for (int s = 0; s < WG_MAXW; ++s) r[s] = r[s] * xi + c->coef[d][s];
    for (int s = 0; s < c->W; ++s) out[s] = r[s];
}
static inline double kernel_one(const wg_ctx *c, int s, double delta)
{
    const double xi = 2.0 * delta - 1.0;
    double r = c->coef[0][s];
    for (int d = 1; d <= c->deg; ++d) r = r * xi + c->coef[d][s];
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
    wg_ctx *c = (wg_ctx *)aligned_alloc(64, (sizeof(wg_ctx) + 63) / 64 * 64);
    if (!c) return NULL;
    memset(c, 0, sizeof(wg_ctx));
    c->nrow = nrow; c->nchan = nchan; c->uvw = uvw; c->freq = freq;
    c->nu = nu; c->nv = nv; c->psx = psx; c->psy = psy;
    c->W = W; c->deg = deg; c->wstack = wstack; c->nplanes = wstack ? nplanes : 1;
    c->w0 = w0; c->dw = dw;
    for (int s = 0; s < W; ++s)
        for (int d = 0; d <= deg; ++d) c->coef[d][s] = coef[s * (deg + 1) + d];
    /* strips of 16 grid rows (the last one takes the remainder, < 32 rows) and v-tiles of 32
     * cells: one bucket's footprint (strip + W) x (32 + W) cells is the thread-local tile */
    (void)nthreads;
    int ns = (int)(nu / 16);
    if (ns < 1) ns = 1;
    c->nstrips = ns;
    c->strip_h = ns > 1 ? 16 : (int)nu;
    c->vt_w = 32;
    c->nvt = (int)((nv + c->vt_w - 1) / c->vt_w);
    c->locks = (omp_lock_t *)malloc(sizeof(omp_lock_t) * (size_t)c->nstrips);
    for (int i = 0; i < c->nstrips; ++i) omp_init_lock(&c->locks[i]);
    return c;
}

void wg_destroy(wg_ctx *c)
{
    if (!c) return;
    free(c->sidx);
    free(c->bstart);
    free(c->chunk_lo);
    free(c->chunk_bucket);
    free(c->bucket_chunk0);
    for (int i = 0; i < c->nstrips; ++i) omp_destroy_lock(&c->locks[i]);
    free(c->locks);
    free(c);
}

static void build_chunks(wg_ctx *c);
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
    const int64_t nb = (int64_t)c->nplanes * c->nstrips * c->nvt;
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
            uint32_t b = (uint32_t)((L.p0 * c->nstrips + strip_of(c, L.iu0)) * c->nvt + L.iv0 / c->vt_w);
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
    build_chunks(c);
    return 0;
}


int64_t wg_nkept(const wg_ctx *c) { return c->bstart ? c->bstart[(int64_t)c->nplanes * c->nstrips * c->nvt] : 0; }

/* Work units: every bucket is cut into chunks of at most WG_CHUNK samples so that the few
 * very full buckets at the centre of the uv-plane are shared between threads. */
#define WG_CHUNK 2048
#define WG_LANES 61
#define WG_TILE_MAX 64  /* (max strip height 2*16-1) + WG_MAXW, (32 + WG_MAXW) rounded up */

static void build_chunks(wg_ctx *c)
{
    const int64_t nb = (int64_t)c->nplanes * c->nstrips * c->nvt;
    free(c->chunk_lo);
    free(c->chunk_bucket);
    free(c->bucket_chunk0);
    c->bucket_chunk0 = (int64_t *)malloc(sizeof(int64_t) * (size_t)(nb + 1));
    int64_t nc = 0;
    for (int64_t b = 0; b < nb; ++b) {
        c->bucket_chunk0[b] = nc;
        nc += (c->bstart[b + 1] - c->bstart[b] + WG_CHUNK - 1) / WG_CHUNK;
    }
    c->bucket_chunk0[nb] = nc;
    c->chunk_lo = (int64_t *)malloc(sizeof(int64_t) * (size_t)(nc + 1));
    c->chunk_bucket = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nc + 1));
    int64_t k = 0;
    for (int64_t b = 0; b < nb; ++b)
        for (int64_t lo = c->bstart[b]; lo < c->bstart[b + 1]; lo += WG_CHUNK) {
            c->chunk_lo[k] = lo;
            c->chunk_bucket[k++] = (int32_t)b;
        }
    c->chunk_lo[nc] = c->bstart[nb];
}

/* grid[nu][nv] (interleaved complex, zeroed here) += all samples touching plane p.
 * ms interleaved complex [nrow][nchan]; wgt [nrow][nchan] or NULL.
 * Each thread accumulates one chunk into a private tile buffer and adds the tile to the
 * shared grid under the locks of the (one or two) strips it covers. */
void wg_grid_plane(wg_ctx *c, int p, const double *ms, const double *wgt, double *grid,
                   int nthreads)
{
    const int W = c->W, PW = c->wstack ? W : 1;
    const int64_t nu = c->nu, nv = c->nv;
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int64_t i = 0; i < nu; ++i) memset(grid + 2 * i * nv, 0, sizeof(double) * 2 * nv);
    const int64_t per_plane = (int64_t)c->nstrips * c->nvt;
    int plo = p - PW + 1, phi = p;
    if (plo < 0) plo = 0;
    if (phi > c->nplanes - PW) phi = c->nplanes - PW;
    if (phi < plo) return;
    const int64_t c0 = c->bucket_chunk0[plo * per_plane], c1 = c->bucket_chunk0[(phi + 1) * per_plane];
#pragma omp parallel num_threads(nthreads)
    {
        double (*buf)[WG_TILE_MAX][2] = malloc(sizeof(double) * WG_TILE_MAX * WG_TILE_MAX * 2);
        double ku[WG_MAXW], kv[WG_MAXW];
        /* neighbouring iterations are WG_LANES-th parts of the chunk list apart, so that
         * concurrently running threads rarely want the same strip lock */
        const int64_t nch = c1 - c0, span = (nch + WG_LANES - 1) / WG_LANES;
#pragma omp for schedule(dynamic, 1)
        for (int64_t it = 0; it < span * WG_LANES; ++it) {
            const int64_t ch = c0 + (it % WG_LANES) * span + it / WG_LANES;
            if (ch >= c1) continue;
            const int64_t b = c->chunk_bucket[ch];
            const int64_t khi = (ch + 1 < c->bucket_chunk0[b + 1]) ? c->chunk_lo[ch + 1] : c->bstart[b + 1];
            const int p0 = (int)(b / per_plane), q = p - p0;
            const int st = (int)((b % per_plane) / c->nvt), vt = (int)(b % c->nvt);
            const int ubase = st * c->strip_h, vbase = vt * c->vt_w;
            const int sh = (st == c->nstrips - 1 ? (int)nu - ubase : c->strip_h) + W;
            const int sw = c->vt_w + W;
            /* bounding box of the footprint origins of this chunk */
            int umin = sh, umax = -1, vmin = sw, vmax = -1;
            for (int64_t k = c->chunk_lo[ch]; k < khi; ++k) {
                const int64_t i = c->sidx[k];
                const wg_loc L = locate(c, i / c->nchan, i % c->nchan);
                const int lu = L.iu0 - ubase, lv = L.iv0 - vbase;
                if (lu < umin) umin = lu;
                if (lu > umax) umax = lu;
                if (lv < vmin) vmin = lv;
                if (lv > vmax) vmax = lv;
            }
            umax += W; vmax += W;   /* exclusive upper bounds of the touched cells */
            for (int a = umin; a < umax; ++a) memset(&buf[a][vmin], 0, sizeof(double) * 2 * (vmax - vmin));
            for (int64_t k = c->chunk_lo[ch]; k < khi; ++k) {
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
                const int lu = L.iu0 - ubase, lv = L.iv0 - vbase;
                for (int a = 0; a < W; ++a) {
                    double *cp = buf[lu + a][lv];
                    const double fr = re * ku[a], fi = im * ku[a];
                    for (int bb = 0; bb < W; ++bb) {
                        cp[2 * bb] += fr * kv[bb];
                        cp[2 * bb + 1] += fi * kv[bb];
                    }
                }
            }
            /* flush: rows of the own strip under its lock, overhanging rows under the next */
            const int own = sh - W;
            for (int part = 0; part < 2; ++part) {
                int a0 = part ? own : 0, a1 = part ? sh : own;
                if (a0 < umin) a0 = umin;
                if (a1 > umax) a1 = umax;
                if (a0 >= a1) continue;
                const int lk = part ? (st + 1) % c->nstrips : st;
                omp_set_lock(&c->locks[lk]);
                for (int a = a0; a < a1; ++a) {
                    int64_t iu = ubase + a;
                    if (iu >= nu) iu -= nu;
                    double *rowp = grid + 2 * iu * nv;
                    for (int bb = vmin; bb < vmax; ++bb) {
                        int64_t iv = vbase + bb;
                        if (iv >= nv) iv -= nv;
                        rowp[2 * iv] += buf[a][bb][0];
                        rowp[2 * iv + 1] += buf[a][bb][1];
                    }
                }
                omp_unset_lock(&c->locks[lk]);
            }
        }
        free(buf);
    }
}

/* ms_out[i] += contribution of plane p (ms_out must be zeroed by the caller before the first
 * plane); the weight and the conjugation for flipped samples are applied by wg_finish_ms. */
void wg_degrid_plane(const wg_ctx *c, int p, const double *grid, double *ms_out, int nthreads)
{
    const int W = c->W, PW = c->wstack ? W : 1;
    const int64_t nu = c->nu, nv = c->nv;
    const int64_t per_plane = (int64_t)c->nstrips * c->nvt;
    int plo = p - PW + 1, phi = p;
    if (plo < 0) plo = 0;
    if (phi > c->nplanes - PW) phi = c->nplanes - PW;
    if (phi < plo) return;
    const int64_t c0 = c->bucket_chunk0[plo * per_plane], c1 = c->bucket_chunk0[(phi + 1) * per_plane];
    /* reads of the grid only; every sample belongs to one bucket, hence one chunk */
#pragma omp parallel num_threads(nthreads)
    {
        double ku[WG_MAXW], kv[WG_MAXW];
#pragma omp for schedule(dynamic, 1)
        for (int64_t ch = c0; ch < c1; ++ch) {
            const int64_t b = c->chunk_bucket[ch];
            const int64_t khi = (ch + 1 < c->bucket_chunk0[b + 1]) ? c->chunk_lo[ch + 1] : c->bstart[b + 1];
            const int q = p - (int)(b / per_plane);
            for (int64_t k = c->chunk_lo[ch]; k < khi; ++k) {
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
                    if (L.iv0 + W <= nv) {
                        const double *cp = rowp + 2 * L.iv0;
                        for (int bb = 0; bb < W; ++bb) {
                            rr += cp[2 * bb] * kv[bb];
                            ri += cp[2 * bb + 1] * kv[bb];
                        }
                    } else {
                        for (int bb = 0; bb < W; ++bb) {
                            int64_t iv = L.iv0 + bb;
                            if (iv >= nv) iv -= nv;
                            rr += rowp[2 * iv] * kv[bb];
                            ri += rowp[2 * iv + 1] * kv[bb];
                        }
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

/* out[a][b] = cu[a] cv[b] * (wstack ? g(nm1) / n : 1), g = Chebyshev series of
 * 1/psihat_w(nm1 dw) on [nm1min, 0] (coefficients with c0 NOT halved, numpy convention);
 * 0 beyond the horizon. */
void wg_correction(const wg_ctx *c, int64_t npix_x, int64_t npix_y, const double *cu,
                   const double *cv, const double *cheb, int ncheb, double nm1min, double *out,
                   int nthreads)
{
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int64_t a = 0; a < npix_x; ++a) {
        const double x = (double)(a - npix_x / 2) * c->psx;
        for (int64_t b = 0; b < npix_y; ++b) {
            double f = cu[a] * cv[b];
            if (c->wstack) {
                const double y = (double)(b - npix_y / 2) * c->psy;
                const double r2 = x * x + y * y;
                if (r2 >= 1.0) { out[a * npix_y + b] = 0.0; continue; }
                const double nm1 = nm1_of(r2);
                const double z = nm1min < 0.0 ? 1.0 - 2.0 * nm1 / nm1min : 0.0;
                double b1 = 0.0, b2 = 0.0;
                for (int j = ncheb - 1; j >= 1; --j) {
                    const double b0 = 2.0 * z * b1 - b2 + cheb[j];
                    b2 = b1;
                    b1 = b0;
                }
                f *= (z * b1 - b2 + cheb[0]) / (nm1 + 1.0);
            }
            out[a * npix_y + b] = f;
        }
    }
}
