// es_kernel.h -- host-side mathematics of the gridding kernel (header only, plain C++).
//
// Kernel family: generalised exponential-of-semicircle
//     phi(t) = exp(beta * W * ((1 - t^2)^e0 - 1)),  |t| <= 1,  support W grid cells
// (Barnett, Magland & af Klinteberg 2019; Arras et al. 2021, A&A 646 A58 sect. 3.2).
// This replaces, from first principles, the kernel machinery inside ducc0.wgridder that
// the reference calls at /root/reference/src/ska_sdp_func_python/imaging/ng.py:240-256.
//
// The device code never evaluates phi directly: it evaluates a piecewise polynomial
// (one polynomial of degree D per grid cell of the support) by Horner's rule, and all
// correction factors are computed from that same piecewise polynomial, so that the
// polynomial IS the kernel (its fit error only perturbs the kernel shape).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

namespace b2g {

struct KernelRow {
    int W;
    double sigma, beta, e0, eps, eps_max;
    double amp;  // psihat(0) / psihat(1/(2 sigma)): rim amplification of the gridding correction
};

static const KernelRow kKernelDB[] = {
#include "kernel_db.inc"
};
static const int kKernelDBSize = int(sizeof(kKernelDB) / sizeof(kKernelDB[0]));

constexpr int kMaxW = 16;
constexpr int kMinW = 4;
inline int poly_degree(int W) { return W + 3; }

inline double es_phi(double t, int W, double beta, double e0)
{
    double a = 1.0 - t * t;
    if (a < 0.0) a = 0.0;
    return std::exp(beta * W * (std::pow(a, e0) - 1.0));
}

// Gauss-Legendre nodes/weights on [-1,1] (Newton on P_n).
inline void gauss_legendre(int n, std::vector<double> &x, std::vector<double> &w)
{
    x.assign(n, 0.0);
    w.assign(n, 0.0);
    const long double pi = 3.14159265358979323846264338327950288L;
    for (int i = 0; i < (n + 1) / 2; ++i) {
        long double z = std::cos(pi * (i + 0.75L) / (n + 0.5L));
        long double pp = 1.0L;
        for (int it = 0; it < 100; ++it) {
            long double p1 = 1.0L, p2 = 0.0L;
            for (int j = 0; j < n; ++j) {
                long double p3 = p2;
                p2 = p1;
                p1 = ((2.0L * j + 1.0L) * z * p2 - j * p3) / (j + 1.0L);
            }
            pp = n * (z * p1 - p2) / (z * z - 1.0L);
            long double dz = p1 / pp;
            z -= dz;
            if (std::fabs((double)dz) < 1e-18) break;
        }
        x[i] = (double)(-z);
        x[n - 1 - i] = (double)z;
        w[i] = w[n - 1 - i] = (double)(2.0L / ((1.0L - z * z) * pp * pp));
    }
}

// Piecewise-polynomial representation of the kernel.
//   piece s (0 <= s < W) covers t in [-1 + 2s/W, -1 + 2(s+1)/W];
//   local variable xi in [-1,1]:  t = -1 + (2s + 1 + xi)/W;
//   phi_s(xi) = sum_{d=0..D} c[s][d] * xi^(D-d)            (highest power first: Horner)
struct PolyKernel {
    int W = 0, D = 0;
    double beta = 0, e0 = 0;
    std::vector<double> c;  // [W][D+1]

    void build(int W_, double beta_, double e0_)
    {
        W = W_;
        beta = beta_;
        e0 = e0_;
        D = poly_degree(W);
        const int n = D + 1;
        c.assign(size_t(W) * n, 0.0);
        const long double pi = 3.14159265358979323846264338327950288L;
        std::vector<long double> f(n), a(n);
        for (int s = 0; s < W; ++s) {
            // Chebyshev interpolation at the n first-kind nodes
            for (int k = 0; k < n; ++k) {
                long double xi = std::cos(pi * (k + 0.5L) / n);
                double t = -1.0 + (2.0 * s + 1.0 + (double)xi) / W;
                f[k] = es_phi(t, W, beta, e0);
            }
            for (int j = 0; j < n; ++j) {
                long double sum = 0.0L;
                for (int k = 0; k < n; ++k) sum += f[k] * std::cos(pi * j * (k + 0.5L) / n);
                a[j] = sum * 2.0L / n;
            }
            a[0] *= 0.5L;
            // Chebyshev series -> monomials, via T_{j+1} = 2 x T_j - T_{j-1}
            std::vector<long double> mono(n, 0.0L), tprev(n, 0.0L), tcur(n, 0.0L), tnext(n);
            tprev[0] = 1.0L;  // T0
            mono[0] += a[0];
            if (n > 1) {
                tcur[1] = 1.0L;  // T1
                mono[1] += a[1];
            }
            for (int j = 2; j < n; ++j) {
                std::fill(tnext.begin(), tnext.end(), 0.0L);
                for (int d = 0; d + 1 < n; ++d) tnext[d + 1] += 2.0L * tcur[d];
                for (int d = 0; d < n; ++d) tnext[d] -= tprev[d];
                for (int d = 0; d < n; ++d) mono[d] += a[j] * tnext[d];
                tprev = tcur;
                tcur = tnext;
            }
            for (int d = 0; d < n; ++d) c[size_t(s) * n + d] = (double)mono[D - d];
        }
        // The kernel is even: piece W-1-s is piece s mirrored (xi -> -xi).  Make the table
        // exactly so, because the device evaluates mirrored pairs from one set of even / odd
        // coefficients (device_common.cuh: KernelTab) and the correction factors below must
        // describe the very same polynomial.
        for (int s = 0; s <= W - 1 - s; ++s) {
            const int m = W - 1 - s;
            for (int d = 0; d < n; ++d) {
                const double sign = ((D - d) & 1) ? -1.0 : 1.0;
                const double a = c[size_t(s) * n + d], b = sign * c[size_t(m) * n + d];
                const double mean = (s == m && sign < 0) ? 0.0 : 0.5 * (a + b);
                c[size_t(s) * n + d] = mean;
                c[size_t(m) * n + d] = sign * mean;
            }
        }
    }

    double eval_piece(int s, double xi) const
    {
        const double *p = &c[size_t(s) * (D + 1)];
        double r = p[0];
        for (int d = 1; d <= D; ++d) r = r * xi + p[d];
        return r;
    }

    // kernel value at t in [-1,1] (0 outside)
    double eval(double t) const
    {
        if (!(t >= -1.0 && t <= 1.0)) return 0.0;
        double g = (t + 1.0) * 0.5 * W;  // [0, W]
        int s = std::min(W - 1, std::max(0, (int)std::floor(g)));
        double xi = 2.0 * (g - s) - 1.0;
        return eval_piece(s, xi);
    }

    // psi-hat(x): Fourier transform of psi(g) = phi(2g/W), g in cells, x in cycles/cell.
    double corr(double x) const
    {
        static thread_local std::vector<double> gx, gw;
        // 24 nodes integrate (degree <= 19 polynomial) x cos(<= pi/2 t + phase) to < 1e-17
        if (gx.empty()) gauss_legendre(24, gx, gw);
        const double twopi = 6.283185307179586476925286766559;
        double sum = 0.0;
        for (int s = 0; s < W; ++s) {
            for (size_t k = 0; k < gx.size(); ++k) {
                double g = -0.5 * W + s + 0.5 * (1.0 + gx[k]);
                sum += 0.5 * gw[k] * eval_piece(s, gx[k]) * std::cos(twopi * g * x);
            }
        }
        return sum;
    }
};

// FFT-friendly even size >= n: 2^a * 3^b * 5^c with a >= 3, b <= 2, c <= 2.  Measured on B200
// (cuFFT 12.9, 2-D c2c in place): such sizes run at 3.5-3.7 TB/s of plane traffic (4 plane
// passes), sizes with a factor 7 or many odd factors (16800 = 2^5*3*5^2*7) at 1.4 TB/s.
inline int64_t good_size_even(int64_t n)
{
    if (n < 8) n = 8;
    int64_t best = -1;
    for (int64_t p5 = 1, c = 0; c <= 2; ++c, p5 *= 5)
        for (int64_t p3 = 1, b = 0; b <= 2; ++b, p3 *= 3) {
            int64_t m = 8 * p3 * p5;
            while (m < n) m *= 2;
            if (best < 0 || m < best) best = m;
        }
    return best;
}

inline int cols_per_lane(int W) { return (W * W + 31) / 32; }

// Estimated run time (seconds, B200) used to pick (W, sigma): the visibility-side kernels
// are issue-bound (packed FMAs on register-resident cells plus ~W*pw cell flushes per
// visibility where the uv-plane is sparsely filled), the image side (FFT passes, screens)
// streams the padded grid through HBM about six times per w-plane.
//   planes_per_sigma = (wmax-wmin)*2*|nm1min|  (so that nplanes = that*sigma + W)
//   csize            = bytes per complex grid cell (8 fp32 / 16 fp64)
inline double cost_model(int W, double sigma, bool wstack, double nvis, int64_t npix_x,
                         int64_t npix_y, double planes_per_sigma, int csize)
{
    const int64_t nu = good_size_even((int64_t)std::ceil(sigma * npix_x));
    const int64_t nv = good_size_even((int64_t)std::ceil(sigma * npix_y));
    const int pw = wstack ? W : 1;
    const double nplanes = wstack ? std::ceil(planes_per_sigma * sigma) + W : 1.0;
    // visibility side: warp instructions per sample of the warp-run kernels (measured on C2:
    // 40 for W = 7 gridding, ~1.3x that for degridding), issued at ~62 % of 148 SMs x 4
    // schedulers x 1.965 GHz; fp64 FMAs issue at half rate and wide kernels re-run the
    // preparation once per group of planes
    const int jl = W <= 5 ? W : (W <= 8 ? 8 : 16);
    const int ipl = (W + 32 / jl - 1) / (32 / jl);
    const double fp = csize == 8 ? 1.0 : 2.0;
    const double prep = 16.0 * (csize == 8 ? 1.0 : W / 8.0 + 1.0);
    const double passes = (wstack && W > 8) ? std::ceil(W / (W <= 12 ? 4.0 : 3.0)) : 1.0;
    const double instr = passes * prep + 0.5 * pw + ipl * (2.0 + pw) * fp;
    double t_vis = 1.15 * nvis * instr / 7.2e11;
    // sparse uv coverage: every sample that opens a new footprint flushes / reloads its cells
    const double density = nvis / (double(nu) * double(nv) * std::max(1.0, nplanes - W + 1));
    const double moved = std::min(1.0, 1.0 / std::max(density, 1e-30));
    t_vis += nvis * moved * (W * pw * 1.2 + 60.0) * fp / 7.2e11 / 4.0;
    // image side per w-plane: cuFFT = 4 passes over the padded plane at ~3.3 TB/s, one more
    // plane pass for the memset (invert) or the screen write (predict) and the cropped read at
    // ~6 TB/s
    const double plane = double(nu) * double(nv) * csize;
    const double t_img = nplanes * (plane * (4.0 / 3.3e12 + 1.5 / 6.0e12) +
                                    double(npix_x) * npix_y * csize / 6.0e12);
    return t_vis + t_img;
}

// Pick the cheapest table row whose error meets eps_dim (per-dimension target).
// Returns index into kKernelDB or -1.
inline int select_kernel(double eps_dim, double amp_max, bool wstack, int maxW, double nvis,
                         int64_t npix_x, int64_t npix_y, double planes_per_sigma, int csize)
{
    int best = -1;
    double bestc = 0;
    for (int i = 0; i < kKernelDBSize; ++i) {
        const KernelRow &r = kKernelDB[i];
        if (r.W > maxW || r.eps > eps_dim || r.amp > amp_max) continue;
        double c = cost_model(r.W, r.sigma, wstack, nvis, npix_x, npix_y, planes_per_sigma, csize);
        if (best < 0 || c < bestc) {
            best = i;
            bestc = c;
        }
    }
    return best;
}

// Chebyshev fit of a smooth function on [a,b]; returns coefficients (c0 already halved)
// so that f(x) ~ sum_j c_j T_j(z), z = (2x - a - b)/(b - a).  Degree is grown until the
// trailing coefficients fall below tol * max|c|.
template <class F>
inline std::vector<double> cheb_fit(F f, double a, double b, double tol, int maxdeg = 96)
{
    const double pi = 3.14159265358979323846;
    for (int n = 16;; n += 16) {
        std::vector<double> fv(n), c(n);
        for (int k = 0; k < n; ++k) {
            double z = std::cos(pi * (k + 0.5) / n);
            fv[k] = f(0.5 * (a + b) + 0.5 * (b - a) * z);
        }
        double cmax = 0;
        for (int j = 0; j < n; ++j) {
            double s = 0;
            for (int k = 0; k < n; ++k) s += fv[k] * std::cos(pi * j * (k + 0.5) / n);
            c[j] = 2.0 * s / n;
            cmax = std::max(cmax, std::fabs(c[j]));
        }
        c[0] *= 0.5;
        double tail = 0;
        for (int j = n - 4; j < n; ++j) tail = std::max(tail, std::fabs(c[j]));
        if (tail <= tol * cmax || n >= maxdeg) {
            int m = n;
            while (m > 1 && std::fabs(c[m - 1]) <= 0.1 * tol * cmax) --m;
            c.resize(m);
            return c;
        }
    }
}

}  // namespace b2g
