// vis_warprun.cu -- fast gridding / degridding kernels ("B2G_ALGO_WARPRUN").
//
// Replaces the ES-kernel gridding / degridding hot loops of ducc0.wgridder.ms2dirty /
// dirty2ms (call sites /root/reference/src/ska_sdp_func_python/imaging/ng.py:240-256,
// :99-112) with a B200-first design:
//
//  * Visibilities arrive sorted by (iu0, iv0, p0) = the grid cell and w-plane of the
//    lower corner of their W x W x W footprint, so consecutive visibilities touch the
//    same cells.
//  * One warp owns a contiguous chunk of the sorted list.  Its 32 lanes own the W*W
//    uv-cells of the current footprint (CPL = ceil(W*W/32) cells per lane, consecutive
//    lanes = consecutive v cells, so every global access is a run of 8-byte cells), and
//    every lane keeps all W planes of its cells in REGISTERS ("register window").
//    Plane P lives in register slot P mod W, so when the footprint origin moves up in w
//    the window rolls: only the planes that fell out of the window are flushed
//    (gridding) or fetched (degridding).
//  * Per 32 visibilities: a lane-parallel preparation phase (one visibility per lane:
//    fp64 coordinates -> cell/plane origin + sub-cell offsets, 3*W Horner evaluations of
//    the kernel polynomial with coefficients taken from the constant bank, staging in
//    shared memory), then a visibility-serial accumulation phase in which each lane does
//    CPL*(3 + 2W) FMAs per visibility on its register window and touches no memory but
//    a few broadcast shared-memory loads.
//  * Gridding flushes with vectorised red.global.add.v2.f32 (one complex cell per
//    instruction-lane, no return value, no CAS loop); degridding reduces the 32 lane
//    partials of 32 visibilities through a padded shared-memory transpose.
//
// Per visibility this is ~60 issued warp-instructions for W=8 (vs 512 read-modify-writes
// for a scatter gridder); the flush traffic is amortised over all visibilities sharing a
// (iu0, iv0) origin.
#include "kernels.h"

namespace b2g {

namespace {

constexpr int kWarps = 4;        // warps per block
constexpr int kChunk = 512;      // visibilities per warp (multiple of 32)

template <typename T, int W, bool WS> struct Stage {
    static constexpr int PW = WS ? W : 1;
    static constexpr int KS = W | 1;                 // odd row stride: conflict-free STS
    static constexpr int KWS = (PW + 3) / 4 * 4;     // 16-byte aligned rows
    alignas(16) T kw[32 * KWS];
    alignas(16) cplx<T> val[32];
    alignas(16) int4 org[32];                        // iu0, iv0, prel, active
    T ku[32 * KS];
    T kv[32 * KS];
};

template <typename T> struct Partials {
    cplx<T> p[32 * 33];
};

// Lane-parallel preparation of one visibility (lane `lane` of the warp).
template <typename T, int W, bool WS>
__device__ __forceinline__ void prepare(const VisGeo &g, const KernelTab<T, W> &tab, int slab,
                                        uint32_t idx, int lane, Stage<T, W, WS> &st,
                                        bool &flip)
{
    using S = Stage<T, W, WS>;
    const uint32_t row = idx / g.nchan, chan = idx - row * g.nchan;
    const Located L = locate(g, W, row, chan);
    flip = L.flip;
    const int prel = L.p0 - slab * g.gs;
    st.org[lane] = make_int4(L.iu0, L.iv0, prel, 1);
    T k[W];
    kernel_values<T, W>(tab, (T)L.du, k);
#pragma unroll
    for (int s = 0; s < W; ++s) st.ku[lane * S::KS + s] = k[s];
    kernel_values<T, W>(tab, (T)L.dv, k);
#pragma unroll
    for (int s = 0; s < W; ++s) st.kv[lane * S::KS + s] = k[s];
    if (WS) {
        kernel_values<T, W>(tab, (T)L.dw, k);
        int slot = prel % W;  // plane P = prel+s lives in slot P mod W
#pragma unroll
        for (int s = 0; s < W; ++s) {
            st.kw[lane * S::KWS + slot] = k[s];
            slot = (slot + 1 == W) ? 0 : slot + 1;
        }
    }
}

}  // namespace

// ------------------------------------------------------------------------------------
// Gridding
// ------------------------------------------------------------------------------------
template <typename T, int W, bool WS>
__global__ void __launch_bounds__(32 * kWarps)
grid_warprun_kernel(VisGeo g, KernelTab<T, W> tab, const uint32_t *__restrict__ sidx,
                    long long begin, long long end, int slab,
                    const cplx<T> *__restrict__ ms, const T *__restrict__ wgt,
                    cplx<T> *__restrict__ slabbuf)
{
    using S = Stage<T, W, WS>;
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

    // cells owned by this lane: c = lane + 32k -> (a, b) = (c / W, c % W), a along u
    int ca[CPL], cb[CPL];
    bool valid[CPL];
#pragma unroll
    for (int k = 0; k < CPL; ++k) {
        int c = lane + 32 * k;
        valid[k] = c < W * W;
        if (!valid[k]) c = 0;
        ca[k] = c / W;
        cb[k] = c - ca[k] * W;
    }
    T are[CPL][PW], aim[CPL][PW];
#pragma unroll
    for (int k = 0; k < CPL; ++k)
#pragma unroll
        for (int q = 0; q < PW; ++q) are[k][q] = aim[k][q] = T(0);
    int coff[CPL];
#pragma unroll
    for (int k = 0; k < CPL; ++k) coff[k] = 0;
    int cur_iu = -1, cur_iv = -1, cur_p = 0, r0 = 0;

    // plane held in slot q of a window whose lowest plane is cur_p (r0 = cur_p % W)
    auto plane_of = [&](int q) { return cur_p + q - r0 + (q < r0 ? W : 0); };

    auto flush_slot = [&](int q, int P) {
        cplx<T> *base = slabbuf + (long long)P * g.plane_stride;
#pragma unroll
        for (int k = 0; k < CPL; ++k) {
            if (valid[k]) red_add(base + coff[k], are[k][q], aim[k][q]);
            are[k][q] = T(0);
            aim[k][q] = T(0);
        }
    };
    auto flush_all = [&]() {
#pragma unroll
        for (int q = 0; q < PW; ++q) flush_slot(q, WS ? plane_of(q) : 0);
    };
    auto flush_below = [&](int pnew) {  // planes [cur_p, pnew) leave the window
#pragma unroll
        for (int q = 0; q < PW; ++q) {
            const int P = plane_of(q);
            if (P < pnew) flush_slot(q, P);
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
                prepare<T, W, WS>(g, tab, slab, idx, lane, st, flip);
                if (flip) val.im = -val.im;
                st.val[lane] = val;
            }
        }
        __syncwarp();
        for (int v = 0; v < n; ++v) {
            const int4 o = st.org[v];
            if (o.w == 0) continue;
            if (o.x != cur_iu || o.y != cur_iv) {
                if (cur_iu >= 0) flush_all();
                cur_iu = o.x;
                cur_iv = o.y;
                cur_p = o.z;
                r0 = WS ? cur_p % W : 0;
#pragma unroll
                for (int k = 0; k < CPL; ++k)
                    coff[k] = wrap_inc(cur_iu + ca[k], g.nu) * g.nv + wrap_inc(cur_iv + cb[k], g.nv);
            } else if (WS && o.z != cur_p) {
                if (o.z < cur_p || o.z >= cur_p + W) flush_all();
                else flush_below(o.z);
                cur_p = o.z;
                r0 = cur_p % W;
            }
            const cplx<T> val = st.val[v];
            T kwv[PW];
            if (WS) {
#pragma unroll
                for (int q = 0; q < PW; ++q) kwv[q] = st.kw[v * S::KWS + q];
            }
#pragma unroll
            for (int k = 0; k < CPL; ++k) {
                const T kk = st.ku[v * S::KS + ca[k]] * st.kv[v * S::KS + cb[k]];
                const T pr = kk * val.re, pi = kk * val.im;
                if (WS) {
#pragma unroll
                    for (int q = 0; q < PW; ++q) {
                        are[k][q] = fma(pr, kwv[q], are[k][q]);
                        aim[k][q] = fma(pi, kwv[q], aim[k][q]);
                    }
                } else {
                    are[k][0] += pr;
                    aim[k][0] += pi;
                }
            }
        }
        __syncwarp();
    }
    if (cur_iu >= 0) flush_all();
}

// ------------------------------------------------------------------------------------
// Degridding
// ------------------------------------------------------------------------------------
template <typename T, int W, bool WS>
__global__ void __launch_bounds__(32 * kWarps)
degrid_warprun_kernel(VisGeo g, KernelTab<T, W> tab, const uint32_t *__restrict__ sidx,
                      long long begin, long long end, int slab, const T *__restrict__ wgt,
                      cplx<T> *__restrict__ out, const cplx<T> *__restrict__ slabbuf)
{
    using S = Stage<T, W, WS>;
    constexpr int PW = S::PW;
    constexpr int CPL = (W * W + 31) / 32;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    S *stage_all = reinterpret_cast<S *>(smem_raw);
    Partials<T> *part_all = reinterpret_cast<Partials<T> *>(smem_raw + kWarps * sizeof(S));

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    S &st = stage_all[warp];
    cplx<T> *part = part_all[warp].p;
    const long long cbeg = begin + ((long long)blockIdx.x * kWarps + warp) * kChunk;
    if (cbeg >= end) return;
    const long long cend = (cbeg + kChunk < end) ? cbeg + kChunk : end;

    int ca[CPL], cb[CPL];
    bool valid[CPL];
#pragma unroll
    for (int k = 0; k < CPL; ++k) {
        int c = lane + 32 * k;
        valid[k] = c < W * W;
        if (!valid[k]) c = 0;
        ca[k] = c / W;
        cb[k] = c - ca[k] * W;
    }
    T gre[CPL][PW], gim[CPL][PW];
#pragma unroll
    for (int k = 0; k < CPL; ++k)
#pragma unroll
        for (int q = 0; q < PW; ++q) gre[k][q] = gim[k][q] = T(0);
    int coff[CPL];
#pragma unroll
    for (int k = 0; k < CPL; ++k) coff[k] = 0;
    int cur_iu = -1, cur_iv = -1, cur_p = 0, r0 = 0;

    auto plane_of = [&](int q) { return cur_p + q - r0 + (q < r0 ? W : 0); };
    auto load_slot = [&](int q, int P) {
        const cplx<T> *base = slabbuf + (long long)P * g.plane_stride;
#pragma unroll
        for (int k = 0; k < CPL; ++k) {
            cplx<T> c{T(0), T(0)};
            if (valid[k]) c = base[coff[k]];
            gre[k][q] = c.re;
            gim[k][q] = c.im;
        }
    };
    auto load_all = [&]() {
#pragma unroll
        for (int q = 0; q < PW; ++q) load_slot(q, WS ? plane_of(q) : 0);
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
                prepare<T, W, WS>(g, tab, slab, my_idx, lane, st, my_flip);
            }
        }
        __syncwarp();
        for (int v = 0; v < n; ++v) {
            const int4 o = st.org[v];
            cplx<T> s{T(0), T(0)};
            if (o.w != 0) {
                if (o.x != cur_iu || o.y != cur_iv) {
                    cur_iu = o.x;
                    cur_iv = o.y;
                    cur_p = o.z;
                    r0 = WS ? cur_p % W : 0;
#pragma unroll
                    for (int k = 0; k < CPL; ++k)
                        coff[k] = wrap_inc(cur_iu + ca[k], g.nu) * g.nv +
                                  wrap_inc(cur_iv + cb[k], g.nv);
                    load_all();
                } else if (WS && o.z != cur_p) {
                    if (o.z < cur_p || o.z >= cur_p + W) {
                        cur_p = o.z;
                        r0 = cur_p % W;
                        load_all();
                    } else {
                        // slots whose plane dropped below the new origin take the plane W above
                        const int pnew = o.z;
                        bool need[PW];
#pragma unroll
                        for (int q = 0; q < PW; ++q) need[q] = plane_of(q) < pnew;
                        cur_p = pnew;
                        r0 = cur_p % W;
#pragma unroll
                        for (int q = 0; q < PW; ++q)
                            if (need[q]) load_slot(q, plane_of(q));
                    }
                }
                T kwv[PW];
                if (WS) {
#pragma unroll
                    for (int q = 0; q < PW; ++q) kwv[q] = st.kw[v * S::KWS + q];
                }
#pragma unroll
                for (int k = 0; k < CPL; ++k) {
                    const T kk = st.ku[v * S::KS + ca[k]] * st.kv[v * S::KS + cb[k]];
                    T tre, tim;
                    if (WS) {
                        tre = gre[k][0] * kwv[0];
                        tim = gim[k][0] * kwv[0];
#pragma unroll
                        for (int q = 1; q < PW; ++q) {
                            tre = fma(gre[k][q], kwv[q], tre);
                            tim = fma(gim[k][q], kwv[q], tim);
                        }
                    } else {
                        tre = gre[k][0];
                        tim = gim[k][0];
                    }
                    s.re = fma(kk, tre, s.re);
                    s.im = fma(kk, tim, s.im);
                }
            }
            part[v * 33 + lane] = s;
        }
        __syncwarp();
        if (lane < n) {
            T sre = T(0), sim = T(0);
#pragma unroll 8
            for (int l = 0; l < 32; ++l) {
                const cplx<T> c = part[lane * 33 + l];
                sre += c.re;
                sim += c.im;
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
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(grid_warprun_kernel<T, W, WS>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attr_set = true;
    }
    grid_warprun_kernel<T, W, WS><<<nblocks_for(n), 32 * kWarps, smem, L.stream>>>(
        L.geo, make_tab<T, W>(L.coef), L.sidx, L.begin, L.end, L.slab, L.ms, L.wgt, L.slabbuf);
    return (int)cudaGetLastError();
}
template <typename T, int W, bool WS> static int degrid_inst(const VisLaunch<T> &L)
{
    const long long n = L.end - L.begin;
    if (n <= 0) return 0;
    const size_t smem = kWarps * (sizeof(Stage<T, W, WS>) + sizeof(Partials<T>));
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(degrid_warprun_kernel<T, W, WS>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attr_set = true;
    }
    degrid_warprun_kernel<T, W, WS><<<nblocks_for(n), 32 * kWarps, smem, L.stream>>>(
        L.geo, make_tab<T, W>(L.coef), L.sidx, L.begin, L.end, L.slab, L.wgt, L.ms_out,
        L.slabbuf);
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
