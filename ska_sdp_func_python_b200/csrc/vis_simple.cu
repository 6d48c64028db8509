// vis_simple.cu -- validation-path gridding / degridding kernels ("B2G_ALGO_SIMPLE").
//
// One thread per visibility, W*W*W global reductions (gridding) or loads (degridding).
// Works for every (precision, W, w-stacking) combination and is the on-device
// cross-check for the fast warp-run kernels in vis_warprun.cu.  It replaces the
// gridding / degridding stage inside ducc0.wgridder.ms2dirty / dirty2ms
// (call sites /root/reference/src/ska_sdp_func_python/imaging/ng.py:240-256, :99-112).
#include "kernels.h"

namespace b2g {

template <typename T, int W, bool WS>
__global__ void __launch_bounds__(128)
grid_simple_kernel(VisGeo g, KernelTab<T, W> tab, const uint32_t *__restrict__ sidx,
                   long long begin, long long end, int plane0, int slot0,
                   const cplx<T> *__restrict__ ms, const T *__restrict__ wgt,
                   cplx<T> *__restrict__ ring)
{
    constexpr int PW = WS ? W : 1;
    long long i = begin + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= end) return;
    const uint32_t idx = sidx[i];
    uint32_t row, chan;
    g.split(idx, row, chan);
    const Located L = locate(g, W, row, chan);
    cplx<T> val = ms[idx];
    if (wgt) {
        T w = wgt[idx];
        val.re *= w;
        val.im *= w;
    }
    if (L.flip) val.im = -val.im;
    if (val.re == T(0) && val.im == T(0)) return;
    T ku[W], kv[W], kw[W];
    kernel_values<T, W>(tab, (T)L.du, ku);
    kernel_values<T, W>(tab, (T)L.dv, kv);
    if (WS) kernel_values<T, W>(tab, (T)L.dw, kw);
    const int sbase = L.p0 - plane0 + slot0;
#pragma unroll 1
    for (int q = 0; q < PW; ++q) {
        cplx<T> *plane = ring + (long long)wrap_inc(sbase + q, g.nslots) * g.plane_stride;
        const T fq = WS ? kw[q] : T(1);
#pragma unroll 1
        for (int a = 0; a < W; ++a) {
            const int iu = wrap_inc(L.iu0 + a, g.nu);
            const T fa = fq * ku[a];
#pragma unroll 1
            for (int b = 0; b < W; ++b) {
                const int iv = wrap_inc(L.iv0 + b, g.nv);
                const T f = fa * kv[b];
                red_add(plane + (long long)iu * g.nv + iv, val.re * f, val.im * f);
            }
        }
    }
}

template <typename T, int W, bool WS>
__global__ void __launch_bounds__(128)
degrid_simple_kernel(VisGeo g, KernelTab<T, W> tab, const uint32_t *__restrict__ sidx,
                     long long begin, long long end, int plane0, int slot0,
                     const T *__restrict__ wgt, cplx<T> *out, const cplx<T> *sub_from,
                     const cplx<T> *__restrict__ ring, const unsigned char *__restrict__ chunk_class)
{
    constexpr int PW = WS ? W : 1;
    // with a chunk table: only the chunks of 1024 samples of class 2 (hardly any shared
    // footprints: one thread per sample beats the warp-run kernel there); blocks of 128 samples
    // never straddle a chunk
    if (chunk_class && chunk_class[blockIdx.x >> 3] != 2) return;
    long long i = begin + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= end) return;
    const uint32_t idx = sidx[i];
    uint32_t row, chan;
    g.split(idx, row, chan);
    T wv = T(1);
    if (wgt) {
        wv = wgt[idx];
        if (wv == T(0)) {
            out[idx] = sub_from ? sub_from[idx] : cplx<T>{T(0), T(0)};
            return;
        }
    }
    const Located L = locate(g, W, row, chan);
    T ku[W], kv[W], kw[W];
    kernel_values<T, W>(tab, (T)L.du, ku);
    kernel_values<T, W>(tab, (T)L.dv, kv);
    if (WS) kernel_values<T, W>(tab, (T)L.dw, kw);
    const int sbase = L.p0 - plane0 + slot0;
    T sre = T(0), sim = T(0);
#pragma unroll 1
    for (int q = 0; q < PW; ++q) {
        const cplx<T> *plane = ring + (long long)wrap_inc(sbase + q, g.nslots) * g.plane_stride;
        T pre = T(0), pim = T(0);
#pragma unroll 1
        for (int a = 0; a < W; ++a) {
            const int iu = wrap_inc(L.iu0 + a, g.nu);
            T rre = T(0), rim = T(0);
#pragma unroll 1
            for (int b = 0; b < W; ++b) {
                const int iv = wrap_inc(L.iv0 + b, g.nv);
                const cplx<T> c = plane[(long long)iu * g.nv + iv];
                rre = fma(c.re, kv[b], rre);
                rim = fma(c.im, kv[b], rim);
            }
            pre = fma(rre, ku[a], pre);
            pim = fma(rim, ku[a], pim);
        }
        const T fq = WS ? kw[q] : T(1);
        sre = fma(pre, fq, sre);
        sim = fma(pim, fq, sim);
    }
    if (L.flip) sim = -sim;
    cplx<T> r{sre * wv, sim * wv};
    if (sub_from) {  // residual mode: data - model
        const cplx<T> d = sub_from[idx];
        r.re = d.re - r.re;
        r.im = d.im - r.im;
    }
    out[idx] = r;
}


template <typename T, int W, bool WS> static int grid_simple_inst(const VisLaunch<T> &L)
{
    long long n = L.end - L.begin;
    if (n <= 0) return 0;
    unsigned blocks = (unsigned)((n + 127) / 128);
    grid_simple_kernel<T, W, WS><<<blocks, 128, 0, L.stream>>>(
        L.geo, make_tab<T, W>(L.coef), L.sidx, L.begin, L.end, L.plane0, L.slot0, L.ms, L.wgt, L.ring);
    return (int)cudaGetLastError();
}
template <typename T, int W, bool WS> static int degrid_simple_inst(const VisLaunch<T> &L)
{
    long long n = L.end - L.begin;
    if (n <= 0) return 0;
    unsigned blocks = (unsigned)((n + 127) / 128);
    degrid_simple_kernel<T, W, WS><<<blocks, 128, 0, L.stream>>>(
        L.geo, make_tab<T, W>(L.coef), L.sidx, L.begin, L.end, L.plane0, L.slot0, L.wgt, L.ms_out,
        L.sub_from, L.ring, L.chunk_class);
    return (int)cudaGetLastError();
}

#define B2G_W_CASES(FN)                                                             \
    switch (L.W) {                                                                  \
    case 4: return L.wstack ? FN<T, 4, true>(L) : FN<T, 4, false>(L);               \
    case 5: return L.wstack ? FN<T, 5, true>(L) : FN<T, 5, false>(L);               \
    case 6: return L.wstack ? FN<T, 6, true>(L) : FN<T, 6, false>(L);               \
    case 7: return L.wstack ? FN<T, 7, true>(L) : FN<T, 7, false>(L);               \
    case 8: return L.wstack ? FN<T, 8, true>(L) : FN<T, 8, false>(L);               \
    case 9: return L.wstack ? FN<T, 9, true>(L) : FN<T, 9, false>(L);               \
    case 10: return L.wstack ? FN<T, 10, true>(L) : FN<T, 10, false>(L);            \
    case 11: return L.wstack ? FN<T, 11, true>(L) : FN<T, 11, false>(L);            \
    case 12: return L.wstack ? FN<T, 12, true>(L) : FN<T, 12, false>(L);            \
    case 13: return L.wstack ? FN<T, 13, true>(L) : FN<T, 13, false>(L);            \
    case 14: return L.wstack ? FN<T, 14, true>(L) : FN<T, 14, false>(L);            \
    case 15: return L.wstack ? FN<T, 15, true>(L) : FN<T, 15, false>(L);            \
    case 16: return L.wstack ? FN<T, 16, true>(L) : FN<T, 16, false>(L);            \
    default: return -1;                                                             \
    }

template <typename T> int launch_grid_simple(const VisLaunch<T> &L) { B2G_W_CASES(grid_simple_inst) }
template <typename T> int launch_degrid_simple(const VisLaunch<T> &L) { B2G_W_CASES(degrid_simple_inst) }

template int launch_grid_simple<float>(const VisLaunch<float> &);
template int launch_grid_simple<double>(const VisLaunch<double> &);
template int launch_degrid_simple<float>(const VisLaunch<float> &);
template int launch_degrid_simple<double>(const VisLaunch<double> &);

}  // namespace b2g
