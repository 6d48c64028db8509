// fft_inplace.cuh -- one complex fp32 line FFT per thread block, IN PLACE in one shared-memory
// buffer, every stage staged through registers.
//
// Building block of the hand-written pruned 2-D FFT of the w-planes (csrc/fft_planes.cu), which
// replaces the per-plane cuFFT call + w-screen kernel + cropping inside what the reference runs as
// ducc0.wgridder.ms2dirty / dirty2ms (call sites
// /root/reference/src/ska_sdp_func_python/imaging/ng.py:240-256, :99-112).
//
// Why in place: a 12800-point line is 102 KB, so an out-of-place (ping-pong) transform fills the
// 227 KB of shared memory of an SM with ONE line and nothing can be loaded while it is transformed.
// In place, the second buffer receives the next line (TMA) during the whole transform.
//
// Algorithm: Stockham autosort, three (or two) stages with compile-time radices, natural order in
// and out.  A stage is  read (every thread pulls the CNT x R inputs of its CNT butterflies into
// registers) -> barrier -> twiddle, DFT_R, write -> barrier,  so the buffer may change layout
// between stages: linear where TMA touches it, padded (i + (i >> PS)) where a stage's writes would
// otherwise hit one bank.  Plans are static: thread count T divides every stage's butterfly count,
// so all threads carry the same R*CNT = 40 points through every stage.
//
// Twiddles: butterfly k of a stage needs w^t, t < R, w = exp(S 2 pi i k / L).  Three shared-memory
// tables hold w, w^4 and w^16 per k (fp64-computed, rounded once); the other powers are products of
// at most three of { w^c, w^(4b), w^(16a) } with c, b < 4 -- at most five roundings, no table of
// n entries (there is no shared memory left for one) and no bank conflicts (lanes read
// consecutive k).
//
// Everything here is __host__ __device__: tests/native/test_fft_host.cpp runs the same templates on
// one host thread (read phase of all "threads", then write phase) against a direct DFT.
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

// complex add / subtract: one packed FADD2 / FFMA2 on the device (sm_100 f32x2 arithmetic; the
// subtraction is b * (-1, -1) + a, exact like a - b), two scalar operations on the host
B2G_HD float2 cadd(float2 a, float2 b)
{
#if defined(__CUDA_ARCH__)
    return __fadd2_rn(a, b);
#else
    return make_float2(a.x + b.x, a.y + b.y);
#endif
}
B2G_HD float2 csub(float2 a, float2 b)
{
#if defined(__CUDA_ARCH__)
    return __ffma2_rn(b, make_float2(-1.f, -1.f), a);
#else
    return make_float2(a.x - b.x, a.y - b.y);
#endif
}

#include "fft_codelets.inc"

B2G_HD float2 cmul(float2 a, float2 b)
{
    return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}

// ---- static plans ----------------------------------------------------------------------------
// N = R0 * R1 * R2 (R2 = 1: two stages); T threads; stage s runs C_s = N / (R_s T) butterflies per
// thread.  PS: padding shift of the intermediate layouts.
template <int N_, int T_, int PS_, int R0_, int R1_, int R2_>
struct StaticPlan {
    static constexpr int N = N_, T = T_, PS = PS_;
    static constexpr int R0 = R0_, R1 = R1_, R2 = R2_;
    static constexpr int NSTAGE = R2_ > 1 ? 3 : 2;
    static constexpr int C0 = N_ / (R0_ * T_), C1 = N_ / (R1_ * T_), C2 = R2_ > 1 ? N_ / (R2_ * T_) : 0;
    static_assert(R0_ * R1_ * R2_ == N_, "radices must multiply to N");
    static_assert(C0 * R0_ * T_ == N_ && C1 * R1_ * T_ == N_, "T must divide every butterfly count");
    static_assert(R2_ == 1 || C2 * R2_ * T_ == N_, "T must divide every butterfly count");
    static constexpr int NPAD = N_ + (N_ >> PS_);        // elements of one padded line buffer
    // twiddle tables: stage 1 has ns = R0 entries, stage 2 ns = R0 R1; three tables each
    static constexpr int NS1 = R0_, NS2 = R2_ > 1 ? R0_ * R1_ : 0;
    static constexpr int TW_ENTRIES = 3 * (NS1 + NS2);
    static constexpr int THREADS = (T_ + 31) / 32 * 32;  // launched block size
};

// the grids of the benchmark configurations and their test-sized relatives: N = 40 T
typedef StaticPlan<12800, 320, 4, 8, 40, 40> Plan12800;
typedef StaticPlan<6400, 160, 4, 8, 20, 40> Plan6400;
typedef StaticPlan<3200, 80, 4, 8, 10, 40> Plan3200;
typedef StaticPlan<1600, 40, 4, 8, 5, 40> Plan1600;

template <class P> B2G_HD int padded(int i) { return i + (i >> P::PS); }

// tables for sign S: for stage 1 (L = R0 R1, k < R0) then stage 2 (L = N, k < R0 R1):
// [w^1 | w^4 | w^16] each of ns entries
template <class P> inline void build_twiddles(int sign, float2 *tw)
{
    const double two_pi = (sign >= 0 ? 2.0 : -2.0) * 3.14159265358979323846;
    int off = 0;
    const int ns_of[2] = {P::NS1, P::NS2};
    const int len_of[2] = {P::R0 * P::R1, P::N};
    for (int s = 0; s < 2; ++s) {
        const int ns = ns_of[s];
        for (int pw = 0; pw < 3; ++pw) {
            const int mult = pw == 0 ? 1 : (pw == 1 ? 4 : 16);
            for (int k = 0; k < ns; ++k) {
                const long e = (long)k * mult % len_of[s];
                const double a = two_pi * double(e) / double(len_of[s]);
                tw[off + pw * ns + k] = make_float2((float)cos(a), (float)sin(a));
            }
        }
        off += 3 * ns;
    }
}

// v[t] *= w^t, t = 1 .. R-1, from w, w^4, w^16
template <int R> B2G_HD void apply_twiddles(float2 (&v)[R], float2 w1, float2 w4, float2 w16)
{
    float2 p1[4], p4[4], p16[4];
    p1[1] = w1;
    p1[2] = cmul(w1, w1);
    p1[3] = cmul(p1[2], w1);
    p4[1] = w4;
    p4[2] = cmul(w4, w4);
    p4[3] = cmul(p4[2], w4);
    p16[1] = w16;
    p16[2] = cmul(w16, w16);
    p16[3] = cmul(p16[2], w16);
#pragma unroll
    for (int a = 0; a < 4; ++a) {
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            if (16 * a + 4 * b >= R) continue;
            float2 rb = make_float2(1.f, 0.f);
            if (a > 0 && b > 0) rb = cmul(p16[a], p4[b]);
            else if (a > 0) rb = p16[a];
            else if (b > 0) rb = p4[b];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int t = 16 * a + 4 * b + c;
                if (t == 0 || t >= R) continue;
                float2 w = rb;
                if (c > 0) w = (a == 0 && b == 0) ? p1[c] : cmul(rb, p1[c]);
                v[t] = cmul(v[t], w);
            }
        }
    }
}

// ---- one stage, split in its two phases --------------------------------------------------------
// read phase: v[c][t] = rd(j + t q), j = tid + c T, q = N / R
template <class P, int R, int CNT, class Rd>
B2G_HD void stage_read(int tid, float2 (&v)[CNT][R], Rd rd)
{
    constexpr int q = P::N / R;
#pragma unroll
    for (int c = 0; c < CNT; ++c) {
        const int j = tid + c * P::T;
#pragma unroll
        for (int t = 0; t < R; ++t) v[c][t] = rd(j + t * q);
    }
}

// compute phase: twiddle (ns > 1) and DFT_R of the thread's CNT butterflies, in registers.
// tw: this stage's [w | w^4 | w^16] tables (ns entries each); unused for NS == 1
template <class P, int S, int R, int CNT, int NS>
B2G_HD void stage_compute(int tid, float2 (&v)[CNT][R], const float2 *tw)
{
#pragma unroll
    for (int c = 0; c < CNT; ++c) {
        if (NS > 1) {
            const int k = (tid + c * P::T) % NS;
            apply_twiddles<R>(v[c], tw[k], tw[NS + k], tw[2 * NS + k]);
        }
        SmallDft<R, S>::run(v[c]);
    }
}

// output index of v[c][u]: j0 + u ns, j0 = (j / ns) ns R + (j mod ns), j = tid + c T
template <class P, int R, int NS> B2G_HD int out_index(int tid, int c, int u)
{
    const int j = tid + c * P::T;
    return (j / NS) * (NS * R) + (j % NS) + u * NS;
}

// compute + write phase
template <class P, int S, int R, int CNT, int NS, class Wr>
B2G_HD void stage_write(int tid, float2 (&v)[CNT][R], const float2 *tw, Wr wr)
{
    stage_compute<P, S, R, CNT, NS>(tid, v, tw);
#pragma unroll
    for (int c = 0; c < CNT; ++c) {
        const int j0 = out_index<P, R, NS>(tid, c, 0);
#pragma unroll
        for (int u = 0; u < R; ++u) wr(j0 + u * NS, v[c][u]);
    }
}

// Stage 0 (ns = 1: no twiddles) written into the padded layout with the address algebra spelled
// out.  Butterfly j writes padded(j R0 + u) = padded(j R0) + u (its R0 outputs do not straddle a
// padding step when R0 divides 2^PS), and padded((tid + c T) R0) = padded(tid R0) + c padded(T R0)
// when T R0 is a multiple of 2^PS: one address register and immediates.  Through the generic
// stage_write the compiler sees C0 R0 = 40 unrelated loop-invariant addresses, keeps them in
// registers across the line loop and spills them (measured: 35 of them in the predict rows pass,
// and beside 227 KB of shared memory a spill reload comes from L2).
template <class P, int S>
B2G_HD void stage0_write_padded(int tid, float2 (&v)[P::C0][P::R0], const float2 *tw, float2 *X)
{
    static_assert((1 << P::PS) % P::R0 == 0, "a butterfly's outputs share one padding step");
    static_assert((P::T * P::R0) % (1 << P::PS) == 0, "the per-c offset is a constant");
    stage_compute<P, S, P::R0, P::C0, 1>(tid, v, tw);
    float2 *base = X + padded<P>(tid * P::R0);
    constexpr int CST = P::T * P::R0 + ((P::T * P::R0) >> P::PS);
#pragma unroll
    for (int c = 0; c < P::C0; ++c)
#pragma unroll
        for (int u = 0; u < P::R0; ++u) base[c * CST + u] = v[c][u];
}

// The later stages' accesses to the padded buffer, same idea.  Reads: padded(tid + c T + t q) =
// padded(tid) + c padded(T) + t padded(q) when T and q are multiples of 2^PS.  Writes of a stage
// with ns | 2^PS: element (j / ns) ns R + j mod ns + u ns lands at
// [padded((j / ns) ns R) + j mod ns] + [u ns + (u ns >> PS)] when ns R is a multiple of 2^PS.
// Plans that do not meet the conditions (1600 points: T = 40) take the generic path.
template <class P, int R, int CNT>
B2G_HD void stage_read_padded(int tid, float2 (&v)[CNT][R], const float2 *X)
{
    constexpr int q = P::N / R, M = 1 << P::PS;
    if constexpr (q % M == 0 && P::T % M == 0) {
        const float2 *base = X + padded<P>(tid);
        constexpr int CT = P::T + (P::T >> P::PS), CQ = q + (q >> P::PS);
#pragma unroll
        for (int c = 0; c < CNT; ++c)
#pragma unroll
            for (int t = 0; t < R; ++t) v[c][t] = base[c * CT + t * CQ];
    } else {
        stage_read<P, R, CNT>(tid, v, [X](int i) { return X[padded<P>(i)]; });
    }
}

template <class P, int S, int R, int CNT, int NS>
B2G_HD void stage_write_padded(int tid, float2 (&v)[CNT][R], const float2 *tw, float2 *X)
{
    constexpr int M = 1 << P::PS;
    if constexpr (M % NS == 0 && (NS * R) % M == 0 && P::T % NS == 0 && (P::T * R) % M == 0) {
        stage_compute<P, S, R, CNT, NS>(tid, v, tw);
        const int a0 = (tid / NS) * (NS * R);
        float2 *base = X + (a0 + (a0 >> P::PS) + tid % NS);
        constexpr int CC = P::T * R + ((P::T * R) >> P::PS);
#pragma unroll
        for (int c = 0; c < CNT; ++c)
#pragma unroll
            for (int u = 0; u < R; ++u) base[c * CC + u * NS + ((u * NS) >> P::PS)] = v[c][u];
    } else {
        stage_write<P, S, R, CNT, NS>(tid, v, tw, [X](int i, float2 val) { X[padded<P>(i)] = val; });
    }
}

// ---- a whole line ------------------------------------------------------------------------------
// Stages 1 .. of a line whose stage 0 has left it in X in the padded layout.  `sync` is the block
// barrier; it is called 2 * NSTAGE - 3 times, the last call being the barrier after the final
// stage's reads (from then on X is free).  Threads with tid >= T only take part in the barriers.
template <class P, int S, class WrLast, class Sync>
B2G_HD void line_rest(int tid, float2 *X, const float2 *tw, WrLast wr_last, Sync &sync)
{
    const bool active = tid < P::T;
    if (P::NSTAGE == 2) {
        float2 v[P::C1][P::R1];
        if (active) stage_read_padded<P, P::R1, P::C1>(tid, v, X);
        sync();
        if (active) stage_write<P, S, P::R1, P::C1, P::NS1>(tid, v, tw, wr_last);
        return;
    }
    {
        float2 v[P::C1][P::R1];
        if (active) stage_read_padded<P, P::R1, P::C1>(tid, v, X);
        sync();
        if (active) stage_write_padded<P, S, P::R1, P::C1, P::NS1>(tid, v, tw, X);
        sync();
    }
    {
        constexpr int R2 = P::R2 > 1 ? P::R2 : 2, C2 = P::C2 > 0 ? P::C2 : 1, NS2 = P::NS2 > 0 ? P::NS2 : 1;
        float2 v[C2][R2];
        if (active) stage_read_padded<P, R2, C2>(tid, v, X);
        sync();
        if (active) stage_write<P, S, R2, C2, NS2>(tid, v, tw + 3 * P::NS1, wr_last);
    }
}

// Same, but the final stage is left to the caller: runs stages 1 .. NSTAGE-2 and the READ phase of
// the last stage (its barrier included), returning the last stage's inputs in v.  The caller then
// runs stage_compute<P, S, LAST_R, LAST_C, LAST_NS> and disposes of the outputs itself
// (out_index gives their positions).  Used where the outputs leave through lane exchanges.
template <class P> struct LastStage {
    static constexpr int R = P::NSTAGE == 3 ? P::R2 : P::R1;
    static constexpr int C = P::NSTAGE == 3 ? P::C2 : P::C1;
    static constexpr int NS = P::NSTAGE == 3 ? P::NS2 : P::NS1;
    static constexpr int TW_OFF = P::NSTAGE == 3 ? 3 * P::NS1 : 0;
};

template <class P, int S, class Sync>
B2G_HD void line_rest_open(int tid, float2 *X, const float2 *tw,
                           float2 (&v)[LastStage<P>::C][LastStage<P>::R], Sync &sync)
{
    const bool active = tid < P::T;
    if (P::NSTAGE == 3) {
        float2 m[P::C1][P::R1];
        if (active) stage_read_padded<P, P::R1, P::C1>(tid, m, X);
        sync();
        if (active) stage_write_padded<P, S, P::R1, P::C1, P::NS1>(tid, m, tw, X);
        sync();
    }
    if (active) stage_read_padded<P, LastStage<P>::R, LastStage<P>::C>(tid, v, X);
    sync();
}

// rd0(i): input element i (any source); X: the line's shared-memory buffer (NPAD elements), used
// in the padded layout between stages; wr_last(k, v): output element k.  rd0 may read X itself
// (in its own layout) and wr_last may write it: every stage's reads are separated from its writes
// by a barrier.  `sync` is called 2 * NSTAGE - 1 times.
template <class P, int S, class Rd0, class Sync>
B2G_HD void line_stage0(int tid, float2 *X, const float2 *tw, Rd0 rd0, Sync &sync)
{
    const bool active = tid < P::T;
    float2 v[P::C0][P::R0];
    if (active) stage_read<P, P::R0, P::C0>(tid, v, rd0);
    sync();
    if (active) stage0_write_padded<P, S>(tid, v, tw, X);
    sync();
}

template <class P, int S, class Rd0, class WrLast, class Sync>
B2G_HD void line_inplace(int tid, float2 *X, const float2 *tw, Rd0 rd0, WrLast wr_last, Sync &sync)
{
    line_stage0<P, S>(tid, X, tw, rd0, sync);
    line_rest<P, S>(tid, X, tw, wr_last, sync);
}

}  // namespace hfft
}  // namespace b2g
