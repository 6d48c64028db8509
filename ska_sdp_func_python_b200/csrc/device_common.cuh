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

template <typename T> struct V2;
template <> struct V2<float> { using type = float2; };
template <> struct V2<double> { using type = double2; };

// packed (x, y) arithmetic: one FFMA2 / FMUL2 / FADD2 on sm_100 for float, two scalar
// operations for double
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ double2 fma2(double2 a, double2 b, double2 c)
{
    return make_double2(fma(a.x, b.x, c.x), fma(a.y, b.y, c.y));
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ double2 mul2(double2 a, double2 b) { return make_double2(a.x * b.x, a.y * b.y); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ double2 add2(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 mk2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ double2 mk2(double a, double b) { return make_double2(a, b); }

// Piecewise-polynomial kernel table, passed BY VALUE as a kernel argument so that the
// Horner FMAs take their coefficient operand straight from the constant bank.
//
// The kernel is even, so piece W-1-s is piece s mirrored: P_{W-1-s}(xi) = P_s(-xi).  With
// P_s(xi) = E_s(xi^2) + xi * O_s(xi^2) only the (W+1)/2 "half taps" are stored, split into
// their even- and odd-power coefficients; one pair of taps then costs about D/2 + D/2 + 2
// FMAs instead of 2 D, and two half taps are evaluated per packed FMA.
template <typename T, int W> struct KernelTab {
    static constexpr int D = W + 3;              // polynomial degree per piece
    static constexpr int NH = (W + 1) / 2;       // half taps: W/2 mirrored pairs (+ the middle)
    static constexpr int NH2 = (NH + 1) & ~1;    // padded to even for packed evaluation
    static constexpr int NE = D / 2 + 1;         // even powers 0, 2, ..
    static constexpr int NO = (D + 1) / 2;       // odd powers 1, 3, ..
    T e[NE][NH2];  // coefficient of xi^(2*(NE-1-k)) of half tap h at e[k][h]: highest first
    T o[NO][NH2];  // coefficient of xi^(2*(NO-1-k)+1)
};

struct VisGeo {
    const double *uvw;     // [nrow][3]
    const double *fscale;  // [nchan]
    const uint32_t *sel;   // optional sample selection: plan sample i is observation sample
                           // sel[i] = row*nchan + chan (w-plane-sharded plans); NULL = all
    uint32_t nchan;
    uint32_t nchan_mul;   // idx / nchan == (t + ((idx - t) >> nchan_sh1)) >> nchan_sh2,
    int nchan_sh1, nchan_sh2;  //   t = umulhi(nchan_mul, idx)   (set by set_nchan)
    int nu, nv;           // padded grid
    int nplanes;          // total w-planes (1 without w-stacking)
    int nslots;           // ring slots
    int wstack;
    double su, sv;        // pixsize_x, pixsize_y  (cycles per pixel per wavelength)
    double sw, w0s;       // plane coordinate of a sample: |w_lambda| * sw - w0s
    long long plane_stride;  // nu*nv

    // exact unsigned 32-bit division by the run-time constant nchan (Granlund & Montgomery)
    __host__ void set_nchan(uint32_t d)
    {
        nchan = d;
        int l = 0;
        while ((1ull << l) < d) ++l;
        nchan_mul = (uint32_t)(((1ull << 32) * ((1ull << l) - d)) / d + 1);
        nchan_sh1 = l < 1 ? l : 1;
        nchan_sh2 = l > 1 ? l - 1 : 0;
    }
    __device__ __forceinline__ void split(uint32_t idx, uint32_t &row, uint32_t &chan) const
    {
        if (sel) idx = __ldg(sel + idx);
        const uint32_t t = __umulhi(nchan_mul, idx);
        row = (t + ((idx - t) >> nchan_sh1)) >> nchan_sh2;
        chan = idx - row * nchan;
    }
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
// (u, v, w) in wavelengths -> footprint origin and offsets
__device__ __forceinline__ Located locate_uvw(const VisGeo &g, int W, double u, double v, double w)
{
    Located L;
    const double halfW = 0.5 * (double)W;
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

__device__ __forceinline__ Located locate(const VisGeo &g, int W, uint32_t row, uint32_t chan)
{
    const double fs = __ldg(g.fscale + chan);
    const double u = __ldg(g.uvw + 3ull * row + 0) * fs;
    const double v = __ldg(g.uvw + 3ull * row + 1) * fs;
    const double w = __ldg(g.uvw + 3ull * row + 2) * fs;
    return locate_uvw(g, W, u, v, w);
}

// All W kernel values along one axis for offset delta in [0,1): out[s] = P_s(2*delta - 1).
template <typename T, int W>
__device__ __forceinline__ void kernel_values(const KernelTab<T, W> &tab, T delta, T (&out)[W])
{
    using K = KernelTab<T, W>;
    using P = typename V2<T>::type;
    const T xi = T(2) * delta - T(1);
    const T x2 = xi * xi;
    const P x22 = mk2(x2, x2);
#pragma unroll
    for (int hp = 0; hp < K::NH2 / 2; ++hp) {
        P ev = mk2(tab.e[0][2 * hp], tab.e[0][2 * hp + 1]);
#pragma unroll
        for (int k = 1; k < K::NE; ++k) ev = fma2(ev, x22, mk2(tab.e[k][2 * hp], tab.e[k][2 * hp + 1]));
        P od = mk2(tab.o[0][2 * hp], tab.o[0][2 * hp + 1]);
#pragma unroll
        for (int k = 1; k < K::NO; ++k) od = fma2(od, x22, mk2(tab.o[k][2 * hp], tab.o[k][2 * hp + 1]));
        const T e_[2] = {ev.x, ev.y}, o_[2] = {od.x, od.y};
#pragma unroll
        for (int t = 0; t < 2; ++t) {
            const int h = 2 * hp + t;
            if (h >= K::NH) continue;
            if (2 * h + 1 == W) {
                out[h] = e_[t];  // the middle tap is even
            } else {
                const P r = fma2(mk2(xi, -xi), mk2(o_[t], o_[t]), mk2(e_[t], e_[t]));
                out[h] = r.x;
                out[W - 1 - h] = r.y;
            }
        }
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
