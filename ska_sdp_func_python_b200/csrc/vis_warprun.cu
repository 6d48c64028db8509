// vis_warprun.cu -- fast gridding / degridding kernels ("B2G_ALGO_WARPRUN").
//
// Replaces the ES-kernel gridding / degridding hot loops of ducc0.wgridder.ms2dirty /
// dirty2ms (call sites /root/reference/src/ska_sdp_func_python/imaging/ng.py:240-256,
// :99-112) with a B200-first design:
//
//  * Samples arrive sorted by (p0, iu0, iv0) = the first w-plane and the first grid cell of
//    their W x W x W footprint, so consecutive samples touch the same planes and mostly the
//    same cells (on the C2 benchmark 99.3 % of the samples have the origin of their
//    predecessor).
//  * One warp owns a contiguous run of the sorted list.  Its lanes own the W*W residue
//    classes (i, j) of uv-cell indices modulo W: lane = (ig, j) owns column class j and the
//    IPL row classes ig*IPL .. ig*IPL+IPL-1.  Any W x W footprint contains exactly one cell
//    of every class, so for every sample each lane has exactly one cell per class to update,
//    and it keeps the W w-planes of that cell in REGISTERS.  When the footprint moves only
//    the lanes whose cell left it flush (gridding) or reload (degridding).
//  * Per 32 samples a lane-parallel preparation phase (one sample per lane: fp64 coordinates
//    -> origin + sub-cell offsets, the 3 W kernel values from the even/odd half-tap
//    polynomials with packed FMAs) stages the kernel values in shared memory ROTATED by the
//    origin's residue, so that in the sample-serial accumulation phase every lane reads its
//    operands at lane-constant addresses: per sample and lane 4 LDS.128 + 2 LDS + IPL FMUL +
//    IPL*W FFMA2, no index arithmetic.  "Did the footprint move" is one bit per sample in a
//    warp-uniform mask computed during preparation.
//  * Gridding flushes with vectorised red.global.add.v2.f32 (one complex cell per lane-op);
//    degridding reduces the 32 lane partials of 16 samples at a time through a padded
//    shared-memory transpose.
//  * fp32 arithmetic uses the packed FFMA2 (fma.rn.f32x2) of sm_100: (re, im) of a cell is
//    one 64-bit register pair.
#include "kernels.h"

#ifndef B2G_MINB_GRID
#define B2G_MINB_GRID 6
#endif
#ifndef B2G_MINB_DEGRID
#define B2G_MINB_DEGRID 6
#endif
#ifndef B2G_CHUNK
#define B2G_CHUNK 256
#endif

namespace b2g {

namespace {

constexpr int kWarps = 4;        // warps per block
constexpr int kChunk = B2G_CHUNK;  // samples per warp (multiple of 32)
constexpr unsigned kFull = 0xffffffffu;

// resident blocks per SM the register allocation is asked to allow (fp32 kernels only; the
// fp64 kernels need their registers for the accumulators)
template <typename T> constexpr int min_blocks(bool grid) { return 1; }
template <> constexpr int min_blocks<float>(bool grid) { return grid ? B2G_MINB_GRID : B2G_MINB_DEGRID; }

template <typename T> struct Vec16;
template <> struct Vec16<float> { using type = float4; };
template <> struct Vec16<double> { using type = double2; };

// PWB: w-planes of a sample's support handled per pass (0 = no w-stacking, W = all in one pass)
template <typename T, int W, int PWB, bool GRID> struct Cfg {
    static constexpr bool WS = PWB > 0;
    static constexpr int PW = WS ? PWB : 1;
    static constexpr int JL = W <= 5 ? W : (W <= 8 ? 8 : 16);   // lanes along v (column classes)
    static constexpr int IG = 32 / JL;                          // complete lane groups along u
    static constexpr int NG = 31 / JL + 1;                      // incl. a partial last group
    static constexpr int IPL = (W + IG - 1) / IG;               // row classes per lane
    static constexpr int KU_N = NG * IPL;                       // ku entries a row must hold
    static constexpr bool KU_VEC = IPL == 2;                    // read as one (T, T) vector
    static constexpr int KUS = KU_VEC ? KU_N + 2 : (KU_N | 1);  // row strides, in T
    static constexpr int KVS = (JL + 1) | 1;
    // w-kernel row: gridding (kw*re, kw*im) per plane, degridding kw per plane
    static constexpr int KW_T = GRID ? 2 * PW : PW;
    static constexpr int KW_V = 16 / (int)sizeof(T);            // T per 16-byte vector
    static constexpr int KW_N16 = (KW_T + KW_V - 1) / KW_V;     // vectors read per row
    static constexpr int KW_S16 = KW_N16 | 1;                   // odd stride: conflict-free STS.128
};

template <typename T, int W, int PWB, bool GRID> struct Stage {
    using C = Cfg<T, W, PWB, GRID>;
    alignas(16) T kw[32 * C::KW_S16 * C::KW_V];
    alignas(16) int4 org[32];     // iu0, iv0, p0, ru | rv << 8   (read when the footprint moved)
    alignas(16) T ku[32 * C::KUS];  // ku[(s + ru) mod W] = kernel value of footprint row s
    T kv[32 * C::KVS];
};

template <typename T> struct Partials {
    typename V2<T>::type p[16 * 33];
};

// (u, v, w) of sample idx in wavelengths: loaded one batch ahead of their use (the loads depend on
// the sorted index, so they cannot be issued any earlier than that)
__device__ __forceinline__ void fetch_uvw(const VisGeo &g, uint32_t idx, double &ul, double &vl, double &wl)
{
    uint32_t row, chan;
    g.split(idx, row, chan);
    const double fs = __ldg(g.fscale + chan);
    ul = __ldg(g.uvw + 3ull * row + 0) * fs;
    vl = __ldg(g.uvw + 3ull * row + 1) * fs;
    wl = __ldg(g.uvw + 3ull * row + 2) * fs;
}

// Lane-parallel preparation of one sample (lane `lane` of the warp), located from its (u, v, w) in
// wavelengths.  `val` is the (weighted) visibility for gridding.  Returns the packed origin
// (iu0 << 16 | iv0, p0).
template <typename T, int W, int PWB, bool GRID>
__device__ __forceinline__ void prepare(const VisGeo &g, const KernelTab<T, W> &tab, double ul, double vl,
                                        double wl, int lane, Stage<T, W, PWB, GRID> &st, cplx<T> val,
                                        int q0, bool &flip, int &oxy, int &oz)
{
    using C = Cfg<T, W, PWB, GRID>;
    constexpr bool WS = C::WS;
    const Located L = locate_uvw(g, W, ul, vl, wl);
    flip = L.flip;
    oxy = (L.iu0 << 16) | L.iv0;
    oz = L.p0;
    const int ru = L.iu0 % W, rv = L.iv0 % W;
    st.org[lane] = make_int4(L.iu0, L.iv0, L.p0, ru | (rv << 8));
    T k[W];
    kernel_values<T, W>(tab, (T)L.du, k);
#pragma unroll
    for (int s = 0; s < W; ++s) {
        int pos = s + ru;
        pos -= pos >= W ? W : 0;
        st.ku[lane * C::KUS + pos] = k[s];
    }
    kernel_values<T, W>(tab, (T)L.dv, k);
#pragma unroll
    for (int s = 0; s < W; ++s) {
        int pos = s + rv;
        pos -= pos >= W ? W : 0;
        st.kv[lane * C::KVS + pos] = k[s];
    }
    if (GRID && flip) val.im = -val.im;
    if (WS || GRID) {
        T row_[C::KW_N16 * C::KW_V];
#pragma unroll
        for (int i = 0; i < C::KW_N16 * C::KW_V; ++i) row_[i] = T(0);
        if (WS) {
            kernel_values<T, W>(tab, (T)L.dw, k);
            // taps q0 .. q0+PW-1 of the w-kernel (0 beyond the support); q0 is a multiple of
            // PW, one statically indexed copy per pass keeps k[] in registers
#pragma unroll
            for (int g0 = 0; g0 < W; g0 += C::PW) {
                if (q0 != g0) continue;
#pragma unroll
                for (int s = 0; s < C::PW; ++s) {
                    const T kq = g0 + s < W ? k[g0 + s < W ? g0 + s : 0] : T(0);
                    if (GRID) {
                        row_[2 * s] = kq * val.re;
                        row_[2 * s + 1] = kq * val.im;
                    } else {
                        row_[s] = kq;
                    }
                }
            }
        } else {
            row_[0] = val.re;
            row_[1] = val.im;
        }
        using V = typename Vec16<T>::type;
        V *dst = reinterpret_cast<V *>(st.kw) + lane * C::KW_S16;
#pragma unroll
        for (int n = 0; n < C::KW_N16; ++n) dst[n] = *reinterpret_cast<const V *>(row_ + n * C::KW_V);
    }
}

// the w-kernel row of sample v, as C::KW_T values of T
template <typename T, int W, int PWB, bool GRID>
__device__ __forceinline__ void load_kw(const Stage<T, W, PWB, GRID> &st, int v,
                                        T (&out)[Cfg<T, W, PWB, GRID>::KW_N16 * Cfg<T, W, PWB, GRID>::KW_V])
{
    using C = Cfg<T, W, PWB, GRID>;
    using V = typename Vec16<T>::type;
    const V *src = reinterpret_cast<const V *>(st.kw) + v * C::KW_S16;
    // whole 16-byte vectors, then the remainder in the widest aligned pieces (a broadcast
    // LDS.128 costs two shared-memory wavefronts, LDS.64 / LDS.32 one)
    constexpr int NF = C::KW_T / C::KW_V, REM = C::KW_T % C::KW_V;
#pragma unroll
    for (int n = 0; n < NF; ++n) *reinterpret_cast<V *>(out + n * C::KW_V) = src[n];
    const T *tail = reinterpret_cast<const T *>(src + NF);
    if (REM >= 2) {
        using P = typename V2<T>::type;
        const P t = *reinterpret_cast<const P *>(tail);
        out[NF * C::KW_V] = t.x;
        out[NF * C::KW_V + 1] = t.y;
    }
    if (REM & 1) out[NF * C::KW_V + (REM & ~1)] = tail[REM & ~1];
}

// ku values of this lane's IPL row classes (lane-constant offset i0) and its kv value
template <typename T, int W, int PWB, bool GRID>
__device__ __forceinline__ void load_kuv(const Stage<T, W, PWB, GRID> &st, int v, int i0, int j,
                                         T (&ku)[Cfg<T, W, PWB, GRID>::IPL], T &kv)
{
    using C = Cfg<T, W, PWB, GRID>;
    using P = typename V2<T>::type;
    if (C::KU_VEC) {
        const P t = *reinterpret_cast<const P *>(st.ku + v * C::KUS + i0);
        ku[0] = t.x;
        ku[C::IPL - 1] = t.y;
    } else {
#pragma unroll
        for (int k = 0; k < C::IPL; ++k) ku[k] = st.ku[v * C::KUS + i0 + k];
    }
    kv = st.kv[v * C::KVS + j];
}

// previous active sample's origin for every lane -> "moved" flags; updates the carry
__device__ __forceinline__ unsigned moved_mask(bool active, int oxy, int oz, int lane, unsigned act,
                                               int &carry_xy, int &carry_z)
{
    const unsigned below = act & ((1u << lane) - 1u);
    const int src = below ? 31 - __clz(below) : 0;
    int pxy = __shfl_sync(kFull, oxy, src), pz = __shfl_sync(kFull, oz, src);
    if (!below) {
        pxy = carry_xy;
        pz = carry_z;
    }
    const unsigned mov = __ballot_sync(kFull, active && (oxy != pxy || oz != pz));
    if (act) {
        const int last = 31 - __clz(act);
        carry_xy = __shfl_sync(kFull, oxy, last);
        carry_z = __shfl_sync(kFull, oz, last);
    }
    return mov;
}

}  // namespace

// ------------------------------------------------------------------------------------
// Gridding
// ------------------------------------------------------------------------------------
template <typename T, int W, int PWB>
__global__ void __launch_bounds__(32 * kWarps, min_blocks<T>(true))
grid_warprun_kernel(VisGeo g, KernelTab<T, W> tab, const uint32_t *__restrict__ sidx,
                    long long begin, long long end, int plane0, int slot0, int q0,
                    const cplx<T> *__restrict__ ms, const T *__restrict__ wgt,
                    cplx<T> *__restrict__ ring)
{
    using C = Cfg<T, W, PWB, true>;
    using S = Stage<T, W, PWB, true>;
    using P = typename V2<T>::type;
    constexpr int PW = C::PW, IPL = C::IPL;
    constexpr bool WS = C::WS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    S *stage_all = reinterpret_cast<S *>(smem_raw);

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    S &st = stage_all[warp];
    const long long cbeg = begin + ((long long)blockIdx.x * kWarps + warp) * kChunk;
    if (cbeg >= end) return;
    const long long cend = (cbeg + kChunk < end) ? cbeg + kChunk : end;

    const int j = lane % C::JL, i0 = (lane / C::JL) * IPL;
    bool valid[IPL];
#pragma unroll
    for (int k = 0; k < IPL; ++k) valid[k] = j < W && i0 + k < W;
    P acc[IPL][PW];
    int coff[IPL];       // grid cell (iu*nv + iv) the accumulators of class k belong to, -1 = none
#pragma unroll
    for (int k = 0; k < IPL; ++k) {
        coff[k] = -1;
#pragma unroll
        for (int q = 0; q < PW; ++q) acc[k][q] = mk2(T(0), T(0));
    }
    int cur_p = 0;
    int carry_xy = -1, carry_z = -1;

    auto flush_k = [&](int k) {
        const int sbase = cur_p - plane0 + slot0 + q0;
#pragma unroll
        for (int q = 0; q < PW; ++q) {
            if (WS && W % PW != 0 && q0 + q >= W) continue;  // ragged last pass
            cplx<T> *base = ring + (long long)wrap_inc(sbase + q, g.nslots) * g.plane_stride;
            red_add(base + coff[k], acc[k][q].x, acc[k][q].y);
            acc[k][q] = mk2(T(0), T(0));
        }
    };

    // software pipeline, two deep: the sorted index is loaded two batches ahead, the sample and
    // its weight (whose addresses depend on the index) one batch ahead, and nothing is done with a
    // loaded value before the batch that uses it (the weighting multiply included), so neither
    // load latency sits in front of the arithmetic
    // (the (u, v, w) loads are NOT issued a batch ahead here, unlike in the degridder: measured,
    // the six extra live registers cost this kernel more than the hidden latency gains)
    uint32_t idx1 = 0, idx2 = 0;
    cplx<T> val1{T(0), T(0)};
    T w1 = T(1);
    if (cbeg + lane < cend) {
        idx1 = sidx[cbeg + lane];
        val1 = ms[idx1];
        if (wgt) w1 = wgt[idx1];
    }
    if (cbeg + 32 + lane < cend) idx2 = sidx[cbeg + 32 + lane];

    for (long long b0 = cbeg; b0 < cend; b0 += 32) {
        const int n = (int)((cend - b0 < 32) ? cend - b0 : 32);
        const uint32_t idx = idx1;
        const cplx<T> val{val1.re * w1, val1.im * w1};
        idx1 = idx2;
        if (b0 + 32 + lane < cend) {
            val1 = ms[idx1];
            if (wgt) w1 = wgt[idx1];
        }
        if (b0 + 64 + lane < cend) idx2 = sidx[b0 + 64 + lane];
        const bool active = lane < n && (val.re != T(0) || val.im != T(0));
        int oxy = 0, oz = 0;
        if (active) {
            bool flip;
            double ul, vl, wl;
            fetch_uvw(g, idx, ul, vl, wl);
            prepare<T, W, PWB, true>(g, tab, ul, vl, wl, lane, st, val, q0, flip, oxy, oz);
        }
        const unsigned act = __ballot_sync(kFull, active);
        const unsigned mov = moved_mask(active, oxy, oz, lane, act, carry_xy, carry_z);
        __syncwarp();
        // samples that need more than the plain accumulation: footprint moved, or inactive
        const unsigned stop = (mov | ~act) & (n == 32 ? kFull : ((1u << n) - 1u));
        int v = 0;
        while (v < n) {
            if ((stop >> v) & 1u) {
                if (!((act >> v) & 1u)) {
                    ++v;
                    continue;
                }
                // the footprint moved: re-derive which cell of it every class owns
                const int4 o = st.org[v];
                if (WS && o.z != cur_p) {
#pragma unroll
                    for (int k = 0; k < IPL; ++k)
                        if (valid[k] && coff[k] >= 0) {
                            flush_k(k);
                            coff[k] = -1;
                        }
                    cur_p = o.z;
                }
                const int ru = o.w & 0xff, rv = o.w >> 8;
                int b = j - rv;
                b += (b < 0) ? W : 0;
                const int cv = wrap_inc(o.y + b, g.nv);
#pragma unroll
                for (int k = 0; k < IPL; ++k) {
                    if (!valid[k]) continue;
                    int a = i0 + k - ru;
                    a += (a < 0) ? W : 0;
                    const int off = wrap_inc(o.x + a, g.nu) * g.nv + cv;
                    if (off != coff[k]) {
                        if (coff[k] >= 0) flush_k(k);
                        coff[k] = off;
                    }
                }
            }
            // this sample and the run of plain samples behind it
            const unsigned rest = v < 31 ? stop >> (v + 1) : 0u;
            const int vend = rest ? v + __ffs(rest) : n;
#pragma unroll 4
            for (; v < vend; ++v) {
                T kwr[C::KW_N16 * C::KW_V], kur[IPL], kvr;
                load_kw<T, W, PWB, true>(st, v, kwr);
                load_kuv<T, W, PWB, true>(st, v, i0, j, kur, kvr);
#pragma unroll
                for (int k = 0; k < IPL; ++k) {
                    // classes outside the footprint accumulate into registers never flushed
                    const T kk = kur[k] * kvr;
                    const P kk2 = mk2(kk, kk);
#pragma unroll
                    for (int q = 0; q < PW; ++q)
                        acc[k][q] = fma2(kk2, mk2(kwr[2 * q], kwr[2 * q + 1]), acc[k][q]);
                }
            }
        }
        __syncwarp();
    }
#pragma unroll
    for (int k = 0; k < IPL; ++k)
        if (valid[k] && coff[k] >= 0) flush_k(k);
}

// ------------------------------------------------------------------------------------
// Degridding
// ------------------------------------------------------------------------------------
template <typename T, int W, int PWB>
__global__ void __launch_bounds__(32 * kWarps, min_blocks<T>(false))
degrid_warprun_kernel(VisGeo g, KernelTab<T, W> tab, const uint32_t *__restrict__ sidx,
                      long long begin, long long end, int plane0, int slot0, int q0,
                      const T *__restrict__ wgt, cplx<T> *out, const cplx<T> *sub_from,
                      const cplx<T> *__restrict__ ring, const unsigned char *__restrict__ skip)
{
    using C = Cfg<T, W, PWB, false>;
    static_assert(kWarps * kChunk == 1024, "skip flags are per chunk of 1024 samples (vis_tc.cu)");
    if (skip && skip[blockIdx.x]) return;   // another kernel's chunk (vis_tc.cu classes 1, 2)
    using S = Stage<T, W, PWB, false>;
    using P = typename V2<T>::type;
    constexpr int PW = C::PW, IPL = C::IPL;
    constexpr bool WS = C::WS;
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

    // rows / columns of ku, kv that no sample writes must read as zero: the classes outside
    // the footprint multiply them with their (zero) grid registers
    for (int i = lane; i < 32 * C::KUS; i += 32) st.ku[i] = T(0);
    for (int i = lane; i < 32 * C::KVS; i += 32) st.kv[i] = T(0);
    for (int i = lane; i < 16 * 33; i += 32) part[i] = mk2(T(0), T(0));
    __syncwarp();

    const int j = lane % C::JL, i0 = (lane / C::JL) * IPL;
    bool valid[IPL];
#pragma unroll
    for (int k = 0; k < IPL; ++k) valid[k] = j < W && i0 + k < W;
    P gv[IPL][PW];       // grid values of the cell of class k, planes cur_p .. cur_p+PW-1
    int coff[IPL];
#pragma unroll
    for (int k = 0; k < IPL; ++k) {
        coff[k] = -1;
#pragma unroll
        for (int q = 0; q < PW; ++q) gv[k][q] = mk2(T(0), T(0));
    }
    int cur_p = -1;
    int carry_xy = -1, carry_z = -1;

    auto load_k = [&](int k) {
        const int sbase = cur_p - plane0 + slot0 + q0;
#pragma unroll
        for (int q = 0; q < PW; ++q) {
            if (WS && W % PW != 0 && q0 + q >= W) continue;  // ragged last pass: stays zero
            const cplx<T> *base = ring + (long long)wrap_inc(sbase + q, g.nslots) * g.plane_stride;
            const cplx<T> c = base[coff[k]];
            gv[k][q] = mk2(c.re, c.im);
        }
    };

    uint32_t nidx = 0;
    T nw = T(1);
    double nul = 0.0, nvl = 0.0, nwl = 0.0;
    if (cbeg + lane < cend) {
        nidx = sidx[cbeg + lane];
        if (wgt) nw = wgt[nidx];
        fetch_uvw(g, nidx, nul, nvl, nwl);
    }

    for (long long b0 = cbeg; b0 < cend; b0 += 32) {
        const int n = (int)((cend - b0 < 32) ? cend - b0 : 32);
        const uint32_t my_idx = nidx;
        const T my_w = nw;
        const double ul = nul, vl = nvl, wl = nwl;
        if (b0 + 32 + lane < cend) {
            nidx = sidx[b0 + 32 + lane];
            if (wgt) nw = wgt[nidx];
            fetch_uvw(g, nidx, nul, nvl, nwl);
        }
        const bool active = lane < n && my_w != T(0);
        bool my_flip = false;
        int oxy = 0, oz = 0;
        if (active)
            prepare<T, W, PWB, false>(g, tab, ul, vl, wl, lane, st, cplx<T>{T(1), T(1)}, q0, my_flip, oxy, oz);
        const unsigned act = __ballot_sync(kFull, active);
        const unsigned mov = moved_mask(active, oxy, oz, lane, act, carry_xy, carry_z);
        __syncwarp();
        P res = mk2(T(0), T(0));
        const unsigned stop = (mov | ~act) & (n == 32 ? kFull : ((1u << n) - 1u));
        for (int h0 = 0; h0 < n; h0 += 16) {
            const int hn = n < h0 + 16 ? n : h0 + 16;
            int v = h0;
            while (v < hn) {
                if ((stop >> v) & 1u) {
                    if (!((act >> v) & 1u)) {
                        part[(v & 15) * 33 + lane] = mk2(T(0), T(0));
                        ++v;
                        continue;
                    }
                    // the footprint moved: reload the cells that entered it
                    const int4 o = st.org[v];
                    const bool newplane = (o.z != cur_p);
                    cur_p = o.z;
                    const int ru = o.w & 0xff, rv = o.w >> 8;
                    int b = j - rv;
                    b += (b < 0) ? W : 0;
                    const int cv = wrap_inc(o.y + b, g.nv);
#pragma unroll
                    for (int k = 0; k < IPL; ++k) {
                        if (!valid[k]) continue;
                        int a = i0 + k - ru;
                        a += (a < 0) ? W : 0;
                        const int off = wrap_inc(o.x + a, g.nu) * g.nv + cv;
                        if (newplane || off != coff[k]) {
                            coff[k] = off;
                            load_k(k);
                        }
                    }
                }
                const unsigned rest = v < 31 ? stop >> (v + 1) : 0u;
                int vend = rest ? v + __ffs(rest) : hn;
                vend = vend < hn ? vend : hn;
#pragma unroll 4
                for (; v < vend; ++v) {
                    T kwr[C::KW_N16 * C::KW_V], kur[IPL], kvr;
                    if (WS) load_kw<T, W, PWB, false>(st, v, kwr);
                    load_kuv<T, W, PWB, false>(st, v, i0, j, kur, kvr);
                    P s = mk2(T(0), T(0));
#pragma unroll
                    for (int k = 0; k < IPL; ++k) {
                        // classes outside the footprint hold zeros in gv and contribute nothing
                        const T kk = kur[k] * kvr;
                        P t;
                        if (WS) {
                            t = mul2(gv[k][0], mk2(kwr[0], kwr[0]));
#pragma unroll
                            for (int q = 1; q < PW; ++q) t = fma2(gv[k][q], mk2(kwr[q], kwr[q]), t);
                        } else {
                            t = gv[k][0];
                        }
                        s = fma2(mk2(kk, kk), t, s);
                    }
                    part[(v & 15) * 33 + lane] = s;
                }
            }
            // the 32 lane partials of (up to) 16 samples: lane = (sample r, half h)
            __syncwarp();
            const P *src = part + (lane & 15) * 33 + (lane >> 4) * 16;
            P t = src[0];
#pragma unroll
            for (int i = 1; i < 16; ++i) t = add2(t, src[i]);
            t.x += __shfl_xor_sync(kFull, t.x, 16);
            t.y += __shfl_xor_sync(kFull, t.y, 16);
            if ((lane >> 4) == (h0 >> 4)) res = t;
            __syncwarp();
        }
        if (lane < n) {
            if (my_flip) res.y = -res.y;
            cplx<T> r{res.x * my_w, res.y * my_w};
            if (sub_from) {  // residual mode: data - model (later passes subtract their share)
                const cplx<T> prev = q0 > 0 ? out[my_idx] : sub_from[my_idx];
                r.re = prev.re - r.re;
                r.im = prev.im - r.im;
            } else if (q0 > 0) {  // later passes add their planes' share
                const cplx<T> prev = out[my_idx];
                r.re += prev.re;
                r.im += prev.im;
            }
            out[my_idx] = r;
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------
// Launchers
// ------------------------------------------------------------------------------------
static unsigned nblocks_for(long long n)
{
    const long long per_block = (long long)kWarps * kChunk;
    return (unsigned)((n + per_block - 1) / per_block);
}

template <typename T, int W, int PWB> static int grid_inst(const VisLaunch<T> &L)
{
    const long long n = L.end - L.begin;
    if (n <= 0) return 0;
    const size_t smem = kWarps * sizeof(Stage<T, W, PWB, true>);
    cudaError_t e = cudaFuncSetAttribute(grid_warprun_kernel<T, W, PWB>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    // one pass per group of PWB w-planes of the samples' support
    for (int q0 = 0; q0 < (PWB > 0 ? W : 1); q0 += (PWB > 0 ? PWB : 1)) {
        grid_warprun_kernel<T, W, PWB><<<nblocks_for(n), 32 * kWarps, smem, L.stream>>>(
            L.geo, make_tab<T, W>(L.coef), L.sidx, L.begin, L.end, L.plane0, L.slot0, q0, L.ms,
            L.wgt, L.ring);
        if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
    }
    return 0;
}
template <typename T, int W, int PWB> static int degrid_inst(const VisLaunch<T> &L)
{
    const long long n = L.end - L.begin;
    if (n <= 0) return 0;
    const size_t smem = kWarps * (sizeof(Stage<T, W, PWB, false>) + sizeof(Partials<T>));
    cudaError_t e = cudaFuncSetAttribute(degrid_warprun_kernel<T, W, PWB>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    for (int q0 = 0; q0 < (PWB > 0 ? W : 1); q0 += (PWB > 0 ? PWB : 1)) {
        degrid_warprun_kernel<T, W, PWB><<<nblocks_for(n), 32 * kWarps, smem, L.stream>>>(
            L.geo, make_tab<T, W>(L.coef), L.sidx, L.begin, L.end, L.plane0, L.slot0, q0, L.wgt,
            L.ms_out, L.sub_from, L.ring, L.chunk_class);
        if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
    }
    return 0;
}

template <> bool warprun_supported<float>(int W, bool) { return W >= 4 && W <= 8; }
template <> bool warprun_supported<double>(int W, bool) { return W >= 4 && W <= 16; }

// w-stacking: all W planes in one pass up to W = 8; wider (fp64-only) kernels keep 4 (W <= 12)
// or 3 planes per pass in registers and run ceil(W/4) or ceil(W/3) passes
#define B2G_CASE(FN, WW) \
    case WW: return L.wstack ? FN<T, WW, WW>(L) : FN<T, WW, 0>(L);
#define B2G_CASE_WIDE(FN, WW) \
    case WW: return L.wstack ? FN<T, WW, (WW <= 12 ? 4 : 3)>(L) : FN<T, WW, 0>(L);

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
        B2G_CASE_WIDE(grid_inst, 9) B2G_CASE_WIDE(grid_inst, 10) B2G_CASE_WIDE(grid_inst, 11)
        B2G_CASE_WIDE(grid_inst, 12) B2G_CASE_WIDE(grid_inst, 13) B2G_CASE_WIDE(grid_inst, 14)
        B2G_CASE_WIDE(grid_inst, 15) B2G_CASE_WIDE(grid_inst, 16)
    default: return -1;
    }
}
template <> int launch_degrid_warprun<double>(const VisLaunch<double> &L)
{
    using T = double;
    switch (L.W) {
        B2G_CASE(degrid_inst, 4) B2G_CASE(degrid_inst, 5) B2G_CASE(degrid_inst, 6)
        B2G_CASE(degrid_inst, 7) B2G_CASE(degrid_inst, 8)
        B2G_CASE_WIDE(degrid_inst, 9) B2G_CASE_WIDE(degrid_inst, 10) B2G_CASE_WIDE(degrid_inst, 11)
        B2G_CASE_WIDE(degrid_inst, 12) B2G_CASE_WIDE(degrid_inst, 13) B2G_CASE_WIDE(degrid_inst, 14)
        B2G_CASE_WIDE(degrid_inst, 15) B2G_CASE_WIDE(degrid_inst, 16)
    default: return -1;
    }
}

}  // namespace b2g
