/*
 * oracle/dft_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Direct evaluation (O(Nvis*Npix), fp64) of the measurement equation that the
 * reference's hot path approximates.  The reference computes it through the
 * third-party module ducc0.wgridder (pinned ducc0==0.27.0, /root/reference/pyproject.toml:18,
 * poetry.lock:305-306; call sites /root/reference/src/ska_sdp_func_python/imaging/ng.py:99-112,
 * :117-129, :240-256, :271-287), whose source is NOT in /root/reference and which is not
 * installable here.  This file restates the published definition of ms2dirty / dirty2ms
 * (Arras et al. 2021, A&A 646 A58, eq. 1-3; array-level convention of ducc0):
 *
 *   ms2dirty:  dirty[ix,iy] = (1/n) * Re sum_{row,chan} wgt*ms *
 *                  exp(+2 pi i (f/c) (u x + v y - w (n-1)))
 *   dirty2ms:  ms[row,chan] = wgt * sum_{ix,iy} dirty[ix,iy]/n *
 *                  exp(-2 pi i (f/c) (u x + v y - w (n-1)))
 *   x = (ix - npix_x/2) pixsize_x,  y = (iy - npix_y/2) pixsize_y,  n = sqrt(1-x^2-y^2);
 *   do_wstacking = 0  =>  w term and 1/n dropped.
 *
 * The sign / axis convention is pinned against the reference's own direct DFT
 * (/root/reference/src/ska_sdp_func_python/imaging/dft.py:265-285, dft_cpu_looped) through
 * tests/golden/ (see tests/golden/make_golden.py); it is NOT pinned against ducc0 output
 * ("parity unpinned" w.r.t. ducc0 itself -- see DESIGN.md).
 *
 * Both entry points work on an explicit list of output pixels / samples so that large
 * configurations can be spot-checked.
 */
#define _GNU_SOURCE
#include <math.h>
#include <stddef.h>
#include <stdint.h>

#define SPEED_OF_LIGHT 299792458.0

/* out[k] = dirty[px[k], py[k]] for k < npts.  ms is interleaved complex (re,im). */
void oracle_ms2dirty_points(int64_t nrow, int64_t nchan, const double *uvw,
                            const double *freq, const double *ms, const double *wgt,
                            int64_t npix_x, int64_t npix_y, double pixsize_x,
                            double pixsize_y, int do_wstacking, int64_t npts,
                            const int64_t *px, const int64_t *py, double *out)
{
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t k = 0; k < npts; ++k) {
        double x = ((double)px[k] - (double)(npix_x / 2)) * pixsize_x;
        double y = ((double)py[k] - (double)(npix_y / 2)) * pixsize_y;
        double r2 = x * x + y * y;
        double nm1 = 0.0, n = 1.0;
        if (do_wstacking) {
            if (r2 > 1.0) { out[k] = 0.0; continue; }
            nm1 = -r2 / (sqrt(1.0 - r2) + 1.0);
            n = nm1 + 1.0;
        }
        double acc = 0.0, comp = 0.0; /* Kahan */
        for (int64_t r = 0; r < nrow; ++r) {
            double pu = uvw[3 * r] * x + uvw[3 * r + 1] * y - uvw[3 * r + 2] * nm1;
            for (int64_t c = 0; c < nchan; ++c) {
                size_t i = (size_t)r * nchan + c;
                double w = wgt ? wgt[i] : 1.0;
                double re = ms[2 * i] * w, im = ms[2 * i + 1] * w;
                if (re == 0.0 && im == 0.0) continue;
                double ph = freq[c] / SPEED_OF_LIGHT * pu;
                ph -= rint(ph);
                double s, co;
                sincos(2.0 * M_PI * ph, &s, &co);
                double term = re * co - im * s - comp;
                double t = acc + term;
                comp = (t - acc) - term;
                acc = t;
            }
        }
        out[k] = acc / n;
    }
}

/* out[2k..2k+1] = ms[rows[k], chans[k]] for k < nsel.  dirty is [npix_x][npix_y]. */
void oracle_dirty2ms_samples(int64_t nchan, const double *uvw, const double *freq,
                             const double *dirty, const double *wgt, int64_t npix_x,
                             int64_t npix_y, double pixsize_x, double pixsize_y,
                             int do_wstacking, int64_t nsel, const int64_t *rows,
                             const int64_t *chans, double *out)
{
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t k = 0; k < nsel; ++k) {
        int64_t r = rows[k], c = chans[k];
        double fl = freq[c] / SPEED_OF_LIGHT;
        double u = uvw[3 * r] * fl, v = uvw[3 * r + 1] * fl, w = uvw[3 * r + 2] * fl;
        double are = 0.0, aim = 0.0;
        for (int64_t ix = 0; ix < npix_x; ++ix) {
            double x = ((double)ix - (double)(npix_x / 2)) * pixsize_x;
            for (int64_t iy = 0; iy < npix_y; ++iy) {
                double d = dirty[ix * npix_y + iy];
                if (d == 0.0) continue;
                double y = ((double)iy - (double)(npix_y / 2)) * pixsize_y;
                double r2 = x * x + y * y;
                double nm1 = 0.0, n = 1.0;
                if (do_wstacking) {
                    if (r2 > 1.0) continue;
                    nm1 = -r2 / (sqrt(1.0 - r2) + 1.0);
                    n = nm1 + 1.0;
                }
                double ph = u * x + v * y - w * nm1;
                ph -= rint(ph);
                double s, co;
                sincos(-2.0 * M_PI * ph, &s, &co);
                are += d / n * co;
                aim += d / n * s;
            }
        }
        double wt = wgt ? wgt[(size_t)r * nchan + c] : 1.0;
        out[2 * k] = are * wt;
        out[2 * k + 1] = aim * wt;
    }
}
