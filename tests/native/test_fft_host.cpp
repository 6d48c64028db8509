// Host check of csrc/fft_lines.cuh (index arithmetic, codelets, twiddle tables) against a direct
// O(N^2) DFT in double precision.  Built and run by tests/test_fft_lines.py (g++, no GPU).
//   usage: test_fft_host N [N ...]   prints "N nstage radices... err_fwd err_inv" per size
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "../../ska_sdp_func_python_b200/csrc/fft_lines.cuh"

using namespace b2g::hfft;

template <int S> static double run(const LinePlan &p, const std::vector<float2> &x)
{
    const int n = p.n;
    std::vector<float2> a(n), b(n), out(n), t0(kTwLow), t1(p.n_hi);
    build_twiddles(p, S, t0.data(), t1.data());
    auto load = [&](int i) { return x[i]; };
    auto store = [&](int i, float2 v) { out[i] = v; };
    fft_line<S>(p, a.data(), b.data(), t0.data(), t1.data(), 0, 1, load, store, NoSync());
    // reference: direct DFT on a subset of outputs (all if n small)
    double num = 0, den = 0;
    const int step = n > 2048 ? n / 512 : 1;
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

int main(int argc, char **argv)
{
    int bad = 0;
    std::mt19937 rng(12345);
    std::normal_distribution<float> g(0.f, 1.f);
    for (int a = 1; a < argc; ++a) {
        const int n = std::atoi(argv[a]);
        LinePlan p;
        if (plan_line(n, &p)) {
            printf("%d unsupported\n", n);
            continue;
        }
        std::vector<float2> x(n);
        for (auto &v : x) v = make_float2(g(rng), g(rng));
        const double ef = run<-1>(p, x), ei = run<1>(p, x);
        printf("%d %d", n, p.nstage);
        for (int s = 0; s < p.nstage; ++s) printf(" %d", p.radix[s]);
        printf(" %.3e %.3e\n", ef, ei);
        if (!(ef < 2e-6) || !(ei < 2e-6)) bad = 1;
    }
    return bad;
}
