// fft_lines.cuh -- one complex fp32 line FFT per thread block, entirely in shared memory.
//
// Building block of the hand-written pruned 2-D FFT of the w-planes (csrc/fft_planes.cu), which
// replaces the per-plane cuFFT call + w-screen kernel + cropping inside what the reference runs as
// ducc0.wgridder.ms2dirty / dirty2ms (call sites
// /root/reference/src/ska_sdp_func_python/imaging/ng.py:240-256, :99-112).
//
// Algorithm: Stockham autosort, mixed radix (2^a 3^b 5^c), natural order in and out.  The first
// stage takes its inputs from a caller-supplied LOADER (global memory, or the shared-memory
// buffer a TMA tensor load filled) and the last stage hands its outputs to a caller-supplied
// STORER (global memory, an epilogue in registers, or the shared-memory buffer a TMA tensor store
// reads), so a line of N points makes nstage - 1 round trips through shared memory, ping-ponging
// between two N-point buffers.  Twiddles W_N^e come from a two-level table in shared memory
// (W^e = T1[e >> 7] * T0[e & 127], both tabulated in fp64 and rounded once), one complex product
// each.  The butterflies are generated straight-line codelets (fft_codelets.inc).
//
// Everything here is __host__ __device__ so that tests/test_fft_host.cpp (compiled with g++ via
// tests/test_fft_lines.py) checks the index arithmetic, the codelets and the twiddle scheme against
// a direct O(N^2) DFT without a GPU.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#define B2G_HD __host__ __device__ __forceinline__
#else
#define B2G_HD inline
#ifndef B2G_HOST_FLOAT2
#define B2G_HOST_FLOAT2
struct float2 { float x, y; };
static inline float2 make_float2(float x, float y) { float2 r; r.x = x; r.y = y; return r; }
#endif
#endif

namespace b2g {
namespace hfft {

#include "fft_codelets.inc"

constexpr int kTwLowBits = 7;             // T0 holds 2^7 entries
constexpr int kTwLow = 1 << kTwLowBits;
constexpr int kMaxStages = 6;
constexpr int kMaxLine = 14000;           // two N-point fp32 complex buffers must fit 227 KB

struct LinePlan {
    int n;                    // line length
    int nstage;
    int radix[kMaxStages];
    int n_hi;                 // entries of T1 = ceil(n / 128)
};

// Radix schedule for n (0 on success).  Preference: few stages, an odd first radix (its
// shared-memory writes, at a stride of r1 elements, are then free of bank conflicts), no radix
// above 32 (register pressure), later radices balanced.
inline int plan_line(int n, LinePlan *out)
{
    static const int allowed[] = {32, 25, 16, 15, 9, 8, 5, 4, 3, 2};
    if (n < 2 || n > kMaxLine) return 1;
    int best[kMaxStages], bestn = 0;
    long bestcost = -1;
    int cur[kMaxStages];
    // depth-first search over ordered factorizations
    struct Rec {
        static void go(int rem, int depth, int *cur, int *best, int &bestn, long &bestcost)
        {
            if (rem == 1) {
                if (depth == 0) return;
                long cost = 1000L * depth;
                if (cur[0] % 2 == 0) cost += 400;           // even first radix: bank conflicts
                int mx = 0;
                for (int i = 0; i < depth; ++i) mx = cur[i] > mx ? cur[i] : mx;
                cost += 4 * mx;                               // registers / imbalance
                if (depth >= 2 && cur[depth - 1] < 4) cost += 50;   // tiny last radix: many stores
                if (bestcost < 0 || cost < bestcost) {
                    bestcost = cost;
                    bestn = depth;
                    for (int i = 0; i < depth; ++i) best[i] = cur[i];
                }
                return;
            }
            if (depth == kMaxStages) return;
            for (int r : allowed) {
                if (rem % r) continue;
                if (depth > 0 && (r == 25 || r == 15 || r == 9) && cur[0] % 2 == 0) continue;
                cur[depth] = r;
                go(rem / r, depth + 1, cur, best, bestn, bestcost);
            }
        }
    };
    Rec::go(n, 0, cur, best, bestn, bestcost);
    if (bestcost < 0) return 1;
    out->n = n;
    out->nstage = bestn;
    for (int i = 0; i < kMaxStages; ++i) out->radix[i] = i < bestn ? best[i] : 1;
    out->n_hi = (n + kTwLow - 1) / kTwLow;
    return 0;
}

// Twiddle tables for sign s: t0[m] = exp(s 2 pi i m / n), m < 128; t1[h] = exp(s 2 pi i 128 h / n)
inline void build_twiddles(const LinePlan &p, int sign, float2 *t0, float2 *t1)
{
    const double w = (sign >= 0 ? 2.0 : -2.0) * 3.14159265358979323846 / double(p.n);
    for (int m = 0; m < kTwLow; ++m) t0[m] = make_float2((float)cos(w * m), (float)sin(w * m));
    for (int h = 0; h < p.n_hi; ++h) {
        const double a = w * double(h) * double(kTwLow);
        t1[h] = make_float2((float)cos(a), (float)sin(a));
    }
}

B2G_HD float2 cmul(float2 a, float2 b)
{
    return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}

// W_N^e, 0 <= e < N
B2G_HD float2 twiddle(const float2 *t0, const float2 *t1, int e)
{
    return cmul(t1[e >> kTwLowBits], t0[e & (kTwLow - 1)]);
}

// One butterfly of a Stockham stage: j in [0, n/R); ns = product of the earlier radices.
//   in(j + t n/R) * W^(t k m), k = j mod ns, m = n / (ns R)  ->  DFT_R  ->  out(j0 + u ns),
//   j0 = (j / ns) ns R + k
template <int R, int S, class In, class Out>
B2G_HD void butterfly(int n, int ns, int j, const float2 *t0, const float2 *t1, In in, Out out)
{
    float2 v[R];
    const int q = n / R;
#pragma unroll
    for (int t = 0; t < R; ++t) v[t] = in(j + t * q);
    int k = 0, j0 = j * R;
    if (ns > 1) {
        const int blk = j / ns;
        k = j - blk * ns;
        j0 = blk * ns * R + k;
        const int m = q / ns;             // n / (ns R)
        const int e1 = k * m;
#pragma unroll
        for (int t = 1; t < R; ++t) v[t] = cmul(v[t], twiddle(t0, t1, t * e1));
    }
    SmallDft<R, S>::run(v);
#pragma unroll
    for (int u = 0; u < R; ++u) out(j0 + u * ns, v[u]);
}

// All butterflies of one stage assigned to this thread (tid of nthr)
template <int R, int S, class In, class Out>
B2G_HD void stage_r(int n, int ns, int tid, int nthr, const float2 *t0, const float2 *t1, In in,
                    Out out)
{
    for (int j = tid; j < n / R; j += nthr) butterfly<R, S>(n, ns, j, t0, t1, in, out);
}

template <int S, class In, class Out>
B2G_HD void stage(int radix, int n, int ns, int tid, int nthr, const float2 *t0, const float2 *t1,
                  In in, Out out)
{
    switch (radix) {
    case 2: stage_r<2, S>(n, ns, tid, nthr, t0, t1, in, out); break;
    case 3: stage_r<3, S>(n, ns, tid, nthr, t0, t1, in, out); break;
    case 4: stage_r<4, S>(n, ns, tid, nthr, t0, t1, in, out); break;
    case 5: stage_r<5, S>(n, ns, tid, nthr, t0, t1, in, out); break;
    case 8: stage_r<8, S>(n, ns, tid, nthr, t0, t1, in, out); break;
    case 9: stage_r<9, S>(n, ns, tid, nthr, t0, t1, in, out); break;
    case 15: stage_r<15, S>(n, ns, tid, nthr, t0, t1, in, out); break;
    case 16: stage_r<16, S>(n, ns, tid, nthr, t0, t1, in, out); break;
    case 25: stage_r<25, S>(n, ns, tid, nthr, t0, t1, in, out); break;
    case 32: stage_r<32, S>(n, ns, tid, nthr, t0, t1, in, out); break;
    default: break;
    }
}

struct NoSync {
    B2G_HD void operator()() const {}
};

// A whole line: loader -> [bufA <-> bufB] -> storer.  `sync` is called between stages
// (__syncthreads on the device).  Which buffer holds what: stage 0 writes `first`, then the
// buffers alternate.  Returns nothing; the caller knows from nstage where the data went (the
// storer gets every output).
template <int S, class Load, class Store, class Sync>
B2G_HD void fft_line(const LinePlan &p, float2 *first, float2 *second, const float2 *t0,
                     const float2 *t1, int tid, int nthr, Load load, Store store, Sync sync)
{
    const int n = p.n;
    float2 *src = nullptr, *dst = first;
    int ns = 1;
    for (int s = 0; s < p.nstage; ++s) {
        const int r = p.radix[s];
        const bool last = (s == p.nstage - 1);
        float2 *rd = src, *wr = dst;
        auto in_smem = [rd](int i) { return rd[i]; };
        auto out_smem = [wr](int i, float2 val) { wr[i] = val; };
        if (s == 0 && last) stage<S>(r, n, ns, tid, nthr, t0, t1, load, store);
        else if (s == 0) stage<S>(r, n, ns, tid, nthr, t0, t1, load, out_smem);
        else if (last) stage<S>(r, n, ns, tid, nthr, t0, t1, in_smem, store);
        else stage<S>(r, n, ns, tid, nthr, t0, t1, in_smem, out_smem);
        ns *= r;
        src = dst;
        dst = (dst == first) ? second : first;
        if (!last) sync();
    }
}

}  // namespace hfft
}  // namespace b2g
