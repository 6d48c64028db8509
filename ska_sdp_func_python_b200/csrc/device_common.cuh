// device_common.cuh -- structures and per-visibility device functions shared by all
// visibility-side kernels (key generation, gridding, degridding).
//
// Layout in HBM (per plan):
//   uvw      [nrow][3]   f64   metres (as given by the caller; read-only)
//   fscale   [nchan]     f64   f/c
//   ms       [nrow][nchan] complex<T> ; wgt [nrow][nchan] T
//   sidx     [nvis]      u32   visibility index (row*nchan+chan) sorted by
//                              (p0, iu0, iv0) = the w-plane / cell of the footprint origin
//   ring     [nslots][nu][nv] complex<T>   ring buffer of w-planes of the uv-grid;
//                              plane P lives in slot P mod nslots, nslots = gs + W - 1
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2g {

template <typename T> struct cplx { T re, im; };

// Piecewise-polynomial kernel table, passed BY VALUE as a kernel argument so that the
// Horner FMAs take their coefficient operand straight from the constant bank.
template <typename T, int W> struct KernelTab {
    static constexpr int D = W + 3;
    T c[W][W + 4];  // highest power first
};

struct VisGeo {
    const double *uvw;     // [nrow][3]
    const double *fscale;  // [nchan]
    uint32_t nchan;
    int nu, nv;           // padded grid
    int nplanes;          // total w-planes (1 without w-stacking)
    int nslots;           // ring slots
    int wstack;
    double su, sv;        // pixsize_x, pixsize_y  (cycles per pixel per wavelength)
    double sw, w0s;       // plane coordinate of a sample: |w_lambda| * sw - w0s
    long long plane_stride;  // nu*nv
};

// Footprint origin + sub-cell offsets of one visibility.
struct Located {
    int iu0, iv0;        // wrapped to [0,nu), [0,nv)
    int p0;              // first w-plane touched
    double du, dv, dw;   // in [0,1): distance (cells) from the footprint's lower edge
    bool flip;           // (u,v,w) -> -(u,v,w), vis -> conj(vis)
};

__device__ __forceinline__ void locate_axis(double cyc, int n, double halfW, int &i0w,
                                            double &delta)
{
    // cyc: coordinate in cycles per pixel; the grid is periodic with period n cells
    double f = cyc - rint(cyc);               // [-0.5, 0.5]
    double s = fma(f, (double)n, -halfW);     // lower edge of the support, in cells
    double c = ceil(s);
    delta = c - s;
    int i0 = (int)c;
    if (i0 < 0) i0 += n;
    if (i0 >= n) i0 -= n;
    i0w = i0;
}

// The same function is used by key generation and by the gridding kernels, so the
// sort key always describes the origin the kernels recompute.
__device__ __forceinline__ Located locate(const VisGeo &g, int W, uint32_t row, uint32_t chan)
{
    Located L;
    const double halfW = 0.5 * (double)W;
    const double fs = __ldg(g.fscale + chan);
    double u = __ldg(g.uvw + 3ull * row + 0) * fs;
    double v = __ldg(g.uvw + 3ull * row + 1) * fs;
    double w = __ldg(g.uvw + 3ull * row + 2) * fs;
    L.flip = false;
    L.p0 = 0;
    L.dw = 0.0;
    if (g.wstack) {
        if (w < 0.0) {
            u = -u; v = -v; w = -w;
            L.flip = true;
        }
        double s = fma(w, g.sw, -g.w0s) - halfW;
        double c = ceil(s);
        int p0 = (int)c;
        int pmax = g.nplanes - W;
        p0 = p0 < 0 ? 0 : (p0 > pmax ? pmax : p0);
        L.p0 = p0;
        L.dw = (double)p0 - s;
    }
    locate_axis(u * g.su, g.nu, halfW, L.iu0, L.du);
    locate_axis(v * g.sv, g.nv, halfW, L.iv0, L.dv);
    return L;
}

// All W kernel values along one axis for offset delta (Horner per piece).
template <typename T, int W>
__device__ __forceinline__ void kernel_values(const KernelTab<T, W> &tab, T delta, T (&out)[W])
{
    const T xi = T(2) * delta - T(1);
#pragma unroll
    for (int s = 0; s < W; ++s) {
        T r = tab.c[s][0];
#pragma unroll
        for (int d = 1; d <= KernelTab<T, W>::D; ++d) r = fma(r, xi, tab.c[s][d]);
        out[s] = r;
    }
}

__device__ __forceinline__ void red_add(cplx<float> *p, float re, float im)
{
    asm volatile("red.global.add.v2.f32 [%0], {%1,%2};" ::"l"(p), "f"(re), "f"(im) : "memory");
}
__device__ __forceinline__ void red_add(cplx<double> *p, double re, double im)
{
    atomicAdd(&p->re, re);
    atomicAdd(&p->im, im);
}

__device__ __forceinline__ int wrap_inc(int i, int n) { return i >= n ? i - n : i; }

}  // namespace b2g
