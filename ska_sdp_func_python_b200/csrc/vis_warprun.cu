// vis_warprun.cu -- fast gridding / degridding kernels ("B2G_ALGO_WARPRUN").
//
// Replaces the ES-kernel gridding / degridding hot loops of ducc0.wgridder.ms2dirty /
// dirty2ms (call sites /root/reference/src/ska_sdp_func_python/imaging/ng.py:240-256,
// :99-112) with a B200-first design:
//
//  * Visibilities arrive sorted by (p0, iu0, iv0) = the first w-plane and the first grid
//    cell of their W x W x W footprint, so consecutive visibilities touch the same planes
//    and overlapping cells.
//  * One warp owns a contiguous chunk of the sorted list.  Its lanes own the W*W residue
//    classes (i, j) of uv-cell indices modulo W (CPL = ceil(W*W/32) classes per lane).
//    Any W x W footprint contains exactly one cell of every class, so for every
//    visibility each lane has exactly one cell per class to update, and it keeps the W
//    w-planes of that cell in REGISTERS.  When the footprint moves by one cell only the
//    W lanes whose cell left the footprint flush (gridding) or reload (degridding) their
//    registers; everything else stays put.  Dense uv regions therefore cost ~2W packed
//    FMAs per lane and visibility and almost no memory traffic, sparse regions cost about
//    W*W_w cell flushes per visibility instead of W*W*W_w.
//  * Per 32 visibilities: a lane-parallel preparation phase (one visibility per lane:
//    fp64 coordinates -> cell/plane origin + sub-cell offsets, 3*W Horner evaluations of
//    the kernel polynomial with coefficients taken from the constant bank, staged in
//    shared memory), then a visibility-serial accumulation phase on the register cells.
//  * Gridding flushes with vectorised red.global.add.v2.f32 (one complex cell per
//    lane-op, no return value, no CAS loop); degridding reduces the 32 lane partials of
//    32 visibilities through a padded shared-memory transpose.
//  * fp32 arithmetic uses the packed FFMA2 (fma.rn.f32x2) of sm_100: (re, im) of a cell
//    is one 64-bit register pair.
#include "kernels.h"

namespace b2g {

namespace {

constexpr int kWarps = 4;        // warps per block
constexpr int kChunk = 256;      // visibilities per warp (multiple of 32)

template <typename T> struct V2;
template <> struct V2<float> { using type = float2; };
template <> struct V2<double> { using type = double2; };

__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ double2 fma2(double2 a, double2 b, double2 c)
{
    return make_double2(fma(a.x, b.x, c.x), fma(a.y, b.y, c.y));
}
__device__ __forceinline__ float2 mk2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ double2 mk2(double a, double b) { return make_double2(a, b); }

// row stride of the duplicated w-kernel table, in (k,k) pairs: even (16-byte rows for
// float) and == 2 mod 4 (spreads the lane-parallel stores over the banks)
constexpr int kw_stride(int pw) { return ((pw + 1) / 4) * 4 + 2; }

template <typename T, int W, bool WS> struct Stage {
    using P = typename V2<T>::type;
    static constexpr int PW = WS ? W : 1;
    static constexpr int KS = W | 1;                 // odd row stride: conflict-free STS
    static constexpr int KWS = kw_stride(PW);
    alignas(16) P kwv[32 * KWS];                     // gridding: (kw*re, kw*im) of plane p0+q at [q]
                                                     // degridding: (kw, kw)
    alignas(16) int4 org[32];                        // iu0, iv0, p0, ru | rv<<8 | active<<16
    T ku[32 * KS];
    T kv[32 * KS];
};

template <typename T> struct Partials {
    typename V2<T>::type p[32 * 33];
};

// Lane-parallel preparation of one visibility (lane `lane` of the warp).  `val` is the
// (weighted, conjugated if flipped) visibility for gridding; for degridding pass (1, 1) so that
// kwv holds the duplicated w-kernel values.
template <typename T, int W, bool WS, bool GRID>
__device__ __forceinline__ void prepare(const VisGeo &g, const KernelTab<T, W> &tab,
                                        uint32_t idx, int lane, Stage<T, W, WS> &st,
                                        cplx<T> val, bool &flip)
{
    using S = Stage<T, W, WS>;
    const uint32_t row = idx / g.nchan, chan = idx - row * g.nchan;
    const Located L = locate(g, W, row, chan);
    flip = L.flip;
    st.org[lane] = make_int4(L.iu0, L.iv0, L.p0, (L.iu0 % W) | ((L.iv0 % W) << 8) | (1 << 16));
    T k[W];
    kernel_values<T, W>(tab, (T)L.du, k);
#pragma unroll
    for (int s = 0; s < W; ++s) st.ku[lane * S::KS + s] = k[s];
    kernel_values<T, W>(tab, (T)L.dv, k);
#pragma unroll
    for (int s = 0; s < W; ++s) st.kv[lane * S::KS + s] = k[s];
    if (GRID && flip) val.im = -val.im;
    if (WS) {
        kernel_values<T, W>(tab, (T)L.dw, k);
#pragma unroll
        for (int s = 0; s < W; ++s)
            st.kwv[lane * S::KWS + s] = GRID ? mk2(k[s] * val.re, k[s] * val.im) : mk2(k[s], k[s]);
    } else if (GRID) {
        st.kwv[lane * S::KWS] = mk2(val.re, val.im);
    }
}

}  // namespace

// ------------------------------------------------------------------------------------
// Gridding
// ------------------------------------------------------------------------------------
template <typename T, int W, bool WS>
__global__ void __launch_bounds__(32 * kWarps)
grid_warprun_kernel(VisGeo g, KernelTab<T, W> tab, const uint32_t *__restrict__ sidx,
                    long long begin, long long end, int plane0, int slot0,
                    const cplx<T> *__restrict__ ms, const T *__restrict__ wgt,
                    cplx<T> *__restrict__ ring)
{
    using S = Stage<T, W, WS>;
    using P = typename V2<T>::type;
    constexpr int PW = S::PW;
    constexpr int CPL = (W * W + 31) / 32;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    S *stage_all = reinterpret_cast<S *>(smem_raw);

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    S &st = stage_all[warp];
    const long long cbeg = begin + ((long long)blockIdx.x * kWarps + warp) * kChunk;
    if (cbeg >= end) return;
    const long long cend = (cbeg + kChunk < end) ? cbeg + kChunk : end;

    // residue classes owned by this lane: c = lane + 32k -> (i, j) = (c / W, c % W)
    int ri[CPL], rj[CPL];
    bool valid[CPL];
#pragma unroll
    for (int k = 0; k < CPL; ++k) {
        int c = lane + 32 * k;
        valid[k] = c < W * W;
        if (!valid[k]) c = 0;
        ri[k] = c / W;
        rj[k] = c - ri[k] * W;
    }
    P acc[CPL][PW];
    int coff[CPL];       // grid cell (iu*nv + iv) the accumulators of class k belong to, -1 = none
    int aoff[CPL], boff[CPL];   // position of that cell inside the current footprint
#pragma unroll
    for (int k = 0; k < CPL; ++k) {
        coff[k] = -1;
        aoff[k] = boff[k] = 0;
#pragma unroll
        for (int q = 0; q < PW; ++q) acc[k][q] = mk2(T(0), T(0));
    }
    int cur_p = 0;
    int pox = -1, poy = -1, poz = -1;   // footprint origin of the previous sample

    auto flush_k = [&](int k) {
        const int sbase = cur_p - plane0 + slot0;
#pragma unroll
        for (int q = 0; q < PW; ++q) {
            cplx<T> *base = ring + (long long)wrap_inc(sbase + q, g.nslots) * g.plane_stride;
            red_add(base + coff[k], acc[k][q].x, acc[k][q].y);
            acc[k][q] = mk2(T(0), T(0));
        }
    };

    for (long long b0 = cbeg; b0 < cend; b0 += 32) {
        const int n = (int)((cend - b0 < 32) ? cend - b0 : 32);
        if (lane < n) {
            const uint32_t idx = sidx[b0 + lane];
            cplx<T> val = ms[idx];
            if (wgt) {
                const T w = wgt[idx];
                val.re *= w;
                val.im *= w;
            }
            if (val.re == T(0) && val.im == T(0)) {
                st.org[lane] = make_int4(0, 0, 0, 0);  // inactive: skipped below
            } else {
                bool flip;
                prepare<T, W, WS, true>(g, tab, idx, lane, st, val, flip);
            }
        }
        __syncwarp();
        for (int v = 0; v < n; ++v) {
            const int4 o = st.org[v];
            if ((o.w >> 16) == 0) continue;
            if (o.x != pox || o.y != poy || o.z != poz) {
                // the footprint moved: re-derive which cell of it every class owns
                if (WS && o.z != cur_p) {
#pragma unroll
                    for (int k = 0; k < CPL; ++k)
                        if (valid[k] && coff[k] >= 0) {
                            flush_k(k);
                            coff[k] = -1;
                        }
                    cur_p = o.z;
                }
                const int ru = o.w & 0xff, rv = (o.w >> 8) & 0xff;
#pragma unroll
                for (int k = 0; k < CPL; ++k) {
                    if (!valid[k]) continue;
                    int a = ri[k] - ru;
                    a += (a < 0) ? W : 0;
                    int b = rj[k] - rv;
                    b += (b < 0) ? W : 0;
                    const int cu = wrap_inc(o.x + a, g.nu), cv = wrap_inc(o.y + b, g.nv);
                    const int off = cu * g.nv + cv;
                    if (off != coff[k]) {
                        if (coff[k] >= 0) flush_k(k);
                        coff[k] = off;
                    }
                    aoff[k] = a;
                    boff[k] = b;
                }
                pox = o.x;
                poy = o.y;
                poz = o.z;
            }
            P kwv[PW];
#pragma unroll
            for (int q = 0; q < PW; ++q) kwv[q] = st.kwv[v * S::KWS + q];
            const T *kur = st.ku + v * S::KS, *kvr = st.kv + v * S::KS;
#pragma unroll
            for (int k = 0; k < CPL; ++k) {
                // classes beyond W*W accumulate into registers that are never flushed
                const T kk = kur[aoff[k]] * kvr[boff[k]];
                const P kk2 = mk2(kk, kk);
#pragma unroll
                for (int q = 0; q < PW; ++q) acc[k][q] = fma2(kk2, kwv[q], acc[k][q]);
            }
        }
        __syncwarp();
    }
#pragma unroll
    for (int k = 0; k < CPL; ++k)
        if (valid[k] && coff[k] >= 0) flush_k(k);
}

// ------------------------------------------------------------------------------------
// Degridding
// ------------------------------------------------------------------------------------
template <typename T, int W, bool WS>
__global__ void __launch_bounds__(32 * kWarps)
degrid_warprun_kernel(VisGeo g, KernelTab<T, W> tab, const uint32_t *__restrict__ sidx,
                      long long begin, long long end, int plane0, int slot0,
                      const T *__restrict__ wgt, cplx<T> *__restrict__ out,
                      const cplx<T> *__restrict__ ring)
{
    using S = Stage<T, W, WS>;
    using P = typename V2<T>::type;
    constexpr int PW = S::PW;
    constexpr int CPL = (W * W + 31) / 32;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    S *stage_all = reinterpret_cast<S *>(smem_raw);
    Partials<T> *part_all = reinterpret_cast<Partials<T> *>(smem_raw + kWarps * sizeof(S));

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    S &st = stage_all[warp];
    P *part = part_all[warp].p;
    const long long cbeg = begin + ((long long)blockIdx.x * kWarps + warp) * kChunk;
    if (cbeg >= end) return;
    const long long cend = (cbeg + kChunk < end) ? cbeg + kChunk : end;

    int ri[CPL], rj[CPL];
    bool valid[CPL];
#pragma unroll
    for (int k = 0; k < CPL; ++k) {
        int c = lane + 32 * k;
        valid[k] = c < W * W;
        if (!valid[k]) c = 0;
        ri[k] = c / W;
        rj[k] = c - ri[k] * W;
    }
    P gv[CPL][PW];       // grid values of the cell of class k, planes cur_p .. cur_p+PW-1
    int coff[CPL], aoff[CPL], boff[CPL];
#pragma unroll
    for (int k = 0; k < CPL; ++k) {
        coff[k] = -1;
        aoff[k] = boff[k] = 0;
#pragma unroll
        for (int q = 0; q < PW; ++q) gv[k][q] = mk2(T(0), T(0));
    }
    int cur_p = -1;
    int pox = -1, poy = -1, poz = -1;

    auto load_k = [&](int k) {
        const int sbase = cur_p - plane0 + slot0;
#pragma unroll
        for (int q = 0; q < PW; ++q) {
            const cplx<T> *base = ring + (long long)wrap_inc(sbase + q, g.nslots) * g.plane_stride;
            const cplx<T> c = base[coff[k]];
            gv[k][q] = mk2(c.re, c.im);
        }
    };

    for (long long b0 = cbeg; b0 < cend; b0 += 32) {
        const int n = (int)((cend - b0 < 32) ? cend - b0 : 32);
        uint32_t my_idx = 0;
        T my_w = T(1);
        bool my_flip = false;
        if (lane < n) {
            my_idx = sidx[b0 + lane];
            if (wgt) my_w = wgt[my_idx];
            if (my_w == T(0)) {
                st.org[lane] = make_int4(0, 0, 0, 0);
            } else {
                prepare<T, W, WS, false>(g, tab, my_idx, lane, st, cplx<T>{T(1), T(1)}, my_flip);
            }
        }
        __syncwarp();
        for (int v = 0; v < n; ++v) {
            const int4 o = st.org[v];
            P s = mk2(T(0), T(0));
            if ((o.w >> 16) != 0) {
                if (o.x != pox || o.y != poy || o.z != poz) {
                    const bool newplane = (o.z != cur_p);
                    cur_p = o.z;
                    const int ru = o.w & 0xff, rv = (o.w >> 8) & 0xff;
#pragma unroll
                    for (int k = 0; k < CPL; ++k) {
                        if (!valid[k]) continue;
                        int a = ri[k] - ru;
                        a += (a < 0) ? W : 0;
                        int b = rj[k] - rv;
                        b += (b < 0) ? W : 0;
                        const int cu = wrap_inc(o.x + a, g.nu), cv = wrap_inc(o.y + b, g.nv);
                        const int off = cu * g.nv + cv;
                        if (newplane || off != coff[k]) {
                            coff[k] = off;
                            load_k(k);
                        }
                        aoff[k] = a;
                        boff[k] = b;
                    }
                    pox = o.x;
                    poy = o.y;
                    poz = o.z;
                }
                P kw2[PW];
                if (WS) {
#pragma unroll
                    for (int q = 0; q < PW; ++q) kw2[q] = st.kwv[v * S::KWS + q];
                }
                const T *kur = st.ku + v * S::KS, *kvr = st.kv + v * S::KS;
#pragma unroll
                for (int k = 0; k < CPL; ++k) {
                    // classes beyond W*W hold zeros in gv and contribute nothing
                    const T kk = kur[aoff[k]] * kvr[boff[k]];
                    P t;
                    if (WS) {
                        t = mk2(gv[k][0].x * kw2[0].x, gv[k][0].y * kw2[0].y);
#pragma unroll
                        for (int q = 1; q < PW; ++q) t = fma2(gv[k][q], kw2[q], t);
                    } else {
                        t = gv[k][0];
                    }
                    s = fma2(mk2(kk, kk), t, s);
                }
            }
            part[v * 33 + lane] = s;
        }
        __syncwarp();
        if (lane < n) {
            T sre = T(0), sim = T(0);
#pragma unroll 8
            for (int l = 0; l < 32; ++l) {
                const P c = part[lane * 33 + l];
                sre += c.x;
                sim += c.y;
            }
            if (my_flip) sim = -sim;
            out[my_idx] = cplx<T>{sre * my_w, sim * my_w};
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------
// Launchers
// ------------------------------------------------------------------------------------
template <typename T, int W> static KernelTab<T, W> make_tab(const double *coef)
{
    KernelTab<T, W> t;
    for (int s = 0; s < W; ++s)
        for (int d = 0; d < W + 4; ++d) t.c[s][d] = (T)coef[s * (W + 4) + d];
    return t;
}

static unsigned nblocks_for(long long n)
{
    const long long per_block = (long long)kWarps * kChunk;
    return (unsigned)((n + per_block - 1) / per_block);
}

template <typename T, int W, bool WS> static int grid_inst(const VisLaunch<T> &L)
{
    const long long n = L.end - L.begin;
    if (n <= 0) return 0;
    const size_t smem = kWarps * sizeof(Stage<T, W, WS>);
    cudaError_t e = cudaFuncSetAttribute(grid_warprun_kernel<T, W, WS>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    grid_warprun_kernel<T, W, WS><<<nblocks_for(n), 32 * kWarps, smem, L.stream>>>(
        L.geo, make_tab<T, W>(L.coef), L.sidx, L.begin, L.end, L.plane0, L.slot0, L.ms, L.wgt,
        L.ring);
    return (int)cudaGetLastError();
}
template <typename T, int W, bool WS> static int degrid_inst(const VisLaunch<T> &L)
{
    const long long n = L.end - L.begin;
    if (n <= 0) return 0;
    const size_t smem = kWarps * (sizeof(Stage<T, W, WS>) + sizeof(Partials<T>));
    cudaError_t e = cudaFuncSetAttribute(degrid_warprun_kernel<T, W, WS>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    degrid_warprun_kernel<T, W, WS><<<nblocks_for(n), 32 * kWarps, smem, L.stream>>>(
        L.geo, make_tab<T, W>(L.coef), L.sidx, L.begin, L.end, L.plane0, L.slot0, L.wgt,
        L.ms_out, L.ring);
    return (int)cudaGetLastError();
}

template <> bool warprun_supported<float>(int W, bool) { return W >= 4 && W <= 8; }
template <> bool warprun_supported<double>(int W, bool wstack)
{
    return W >= 4 && (wstack ? W <= 8 : W <= 16);
}

#define B2G_CASE(FN, WW) \
    case WW: return L.wstack ? FN<T, WW, true>(L) : FN<T, WW, false>(L);
#define B2G_CASE2D(FN, WW) \
    case WW: return L.wstack ? -1 : FN<T, WW, false>(L);

template <> int launch_grid_warprun<float>(const VisLaunch<float> &L)
{
    using T = float;
    switch (L.W) {
        B2G_CASE(grid_inst, 4) B2G_CASE(grid_inst, 5) B2G_CASE(grid_inst, 6)
        B2G_CASE(grid_inst, 7) B2G_CASE(grid_inst, 8)
    default: return -1;
    }
}
template <> int launch_degrid_warprun<float>(const VisLaunch<float> &L)
{
    using T = float;
    switch (L.W) {
        B2G_CASE(degrid_inst, 4) B2G_CASE(degrid_inst, 5) B2G_CASE(degrid_inst, 6)
        B2G_CASE(degrid_inst, 7) B2G_CASE(degrid_inst, 8)
    default: return -1;
    }
}
template <> int launch_grid_warprun<double>(const VisLaunch<double> &L)
{
    using T = double;
    switch (L.W) {
        B2G_CASE(grid_inst, 4) B2G_CASE(grid_inst, 5) B2G_CASE(grid_inst, 6)
        B2G_CASE(grid_inst, 7) B2G_CASE(grid_inst, 8)
        B2G_CASE2D(grid_inst, 9) B2G_CASE2D(grid_inst, 10) B2G_CASE2D(grid_inst, 11)
        B2G_CASE2D(grid_inst, 12) B2G_CASE2D(grid_inst, 13) B2G_CASE2D(grid_inst, 14)
        B2G_CASE2D(grid_inst, 15) B2G_CASE2D(grid_inst, 16)
    default: return -1;
    }
}
template <> int launch_degrid_warprun<double>(const VisLaunch<double> &L)
{
    using T = double;
    switch (L.W) {
        B2G_CASE(degrid_inst, 4) B2G_CASE(degrid_inst, 5) B2G_CASE(degrid_inst, 6)
        B2G_CASE(degrid_inst, 7) B2G_CASE(degrid_inst, 8)
        B2G_CASE2D(degrid_inst, 9) B2G_CASE2D(degrid_inst, 10) B2G_CASE2D(degrid_inst, 11)
        B2G_CASE2D(degrid_inst, 12) B2G_CASE2D(degrid_inst, 13) B2G_CASE2D(degrid_inst, 14)
        B2G_CASE2D(degrid_inst, 15) B2G_CASE2D(degrid_inst, 16)
    default: return -1;
    }
}

}  // namespace b2g
