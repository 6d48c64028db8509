// plan.cu -- host side of the B200 gridder: plan construction (kernel choice, w-plane
// layout, correction tables, visibility sort), the invert / predict drivers and the C ABI
// declared in include/b200_wgridder.h.
//
// Replaces what happens inside ducc0.wgridder.ms2dirty / dirty2ms, which the reference
// calls at /root/reference/src/ska_sdp_func_python/imaging/ng.py:240-256, :271-287
// (ms2dirty) and :99-112, :117-129 (dirty2ms).  There is no CPU fallback in this file:
// every entry point needs a CUDA device.
#include <cuda_runtime.h>
#include <cufft.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

#include "../../include/b200_wgridder.h"
#include "es_kernel.h"
#include "kernels.h"

using namespace b2g;

namespace {

constexpr double kSpeedOfLight = 299792458.0;

thread_local std::string g_err;

// Plans built from HOST uvw stream the visibilities of invert / predict over PCIe in row chunks
// of about this many samples, overlapped with gridding / degridding (0 disables chunking).
std::atomic<long long> g_host_chunk_samples{1ll << 24};
constexpr int kMaxChunks = 32;
// Upper limit on the origin w-planes per slab (0 = sized from free memory): lets a small
// geometry run the multi-slab ring-buffer path that the large configurations take.
std::atomic<int> g_max_slab_planes{0};
// Tensor-core degridder (vis_tc.cu): 0 = off, 1 = chunks of the sorted sample list with long runs
// (the rest stays with the warp-run kernel), 2 = every chunk (tests).
std::atomic<int> g_tc_mode{1};
constexpr int kTcMaxRuns = 16;     // a chunk of 1024 samples is "dense" with at most this many runs
constexpr int kSparseRuns = 384;   // ... and "sparse" (one thread per sample) with at least this many

int fail(const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return 1;
}

#define CUDA_OK(expr)                                                                     \
    do {                                                                                  \
        cudaError_t e_ = (expr);                                                          \
        if (e_ != cudaSuccess)                                                            \
            return fail("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_)); \
    } while (0)
#define FFT_OK(expr)                                                                      \
    do {                                                                                  \
        cufftResult r_ = (expr);                                                          \
        if (r_ != CUFFT_SUCCESS)                                                          \
            return fail("%s:%d: %s -> cufft error %d", __FILE__, __LINE__, #expr, (int)r_); \
    } while (0)
#define KERNEL_OK(expr)                                                                   \
    do {                                                                                  \
        int k_ = (expr);                                                                  \
        if (k_ != 0)                                                                      \
            return fail("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,                     \
                        k_ < 0 ? "unsupported kernel configuration"                       \
                               : cudaGetErrorString((cudaError_t)k_));                    \
    } while (0)
#define B2G_OK(expr)               \
    do {                           \
        int b_ = (expr);           \
        if (b_ != 0) return b_;    \
    } while (0)

enum Stage { S_H2D = 0, S_PREP, S_VIS, S_FFT, S_SCREEN, S_CORRECT, S_D2H, S_TOTAL, S_COUNT };

// CUDA-event stage timer: begin/end pairs are summed per stage after one synchronise.
struct StageTimer {
    cudaStream_t stream = nullptr;
    bool enabled = true;
    struct Span { int stage; cudaEvent_t a, b; };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> pool;
    cudaEvent_t get()
    {
        if (!pool.empty()) {
            cudaEvent_t e = pool.back();
            pool.pop_back();
            return e;
        }
        cudaEvent_t e;
        cudaEventCreate(&e);
        return e;
    }
    void begin(int stage)
    {
        if (!enabled) return;
        Span s{stage, get(), get()};
        cudaEventRecord(s.a, stream);
        spans.push_back(s);
    }
    void end()
    {
        if (!enabled || spans.empty()) return;
        cudaEventRecord(spans.back().b, stream);
    }
    // drop spans left behind by a call that failed half-way
    void reset()
    {
        for (auto &s : spans) {
            pool.push_back(s.a);
            pool.push_back(s.b);
        }
        spans.clear();
        (void)cudaGetLastError();
    }
    // call after the stream has been synchronised
    void collect(double *out)
    {
        for (int i = 0; i < S_COUNT; ++i) out[i] = 0.0;
        for (auto &s : spans) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, s.a, s.b) == cudaSuccess) out[s.stage] += ms;
            pool.push_back(s.a);
            pool.push_back(s.b);
        }
        spans.clear();
    }
    ~StageTimer()
    {
        for (auto &s : spans) {
            cudaEventDestroy(s.a);
            cudaEventDestroy(s.b);
        }
        for (auto e : pool) cudaEventDestroy(e);
    }
};

size_t real_size(int dtype) { return dtype == B2G_F64 ? 8 : 4; }

}  // namespace

struct b2g_plan {
    int64_t nrow = 0, nchan = 0, nvis = 0, npix_x = 0, npix_y = 0, nu = 0, nv = 0;
    double psx = 0, psy = 0, epsilon = 0;
    int wstack = 0, precision = B2G_F32, algo = B2G_ALGO_AUTO, verbosity = 0;
    cudaStream_t stream = nullptr;
    int device = 0;
    // kernel
    int W = 0;
    double sigma = 0, beta = 0, e0 = 0, eps_kernel = 0;
    PolyKernel pk;
    // w-planes
    int nplanes = 1, gs = 1, nslabs = 1, nslots = 1;
    double dw = 0, w0 = 0, wmin = 0, wmax = 0, nm1min = 0;
    double umax_l = 0, vmax_l = 0;   // max |u|, |v| over all rows and channels, wavelengths
    // device buffers
    double *d_uvw = nullptr;
    bool own_uvw = false;
    double *d_fscale = nullptr;
    double *d_cu = nullptr, *d_cv = nullptr, *d_cheb = nullptr;
    int ncheb = 0;
    uint32_t *d_sidx = nullptr;
    const uint32_t *d_sel = nullptr;  // sample selection of a w-plane-sharded plan (NULL = all)
    bool own_sel = false;
    std::vector<long long> slab_begin;
    // host-streaming plans (nslabs == 1): sorted by (row chunk, p0, iu0, iv0); chunk c holds
    // rows [c*rows_per_chunk, (c+1)*rows_per_chunk)
    int nchunks = 1;
    long long rows_per_chunk = 0;
    cudaStream_t copy_stream = nullptr;
    std::vector<cudaEvent_t> chunk_ev;
    cudaEvent_t ev_a = nullptr, ev_b = nullptr;
    cudaEvent_t ev_plan0 = nullptr, ev_plan1 = nullptr;  // plan creation on the stream
    void *d_ring = nullptr;
    size_t plane_bytes = 0;
    void *d_acc = nullptr;
    double *d_scratch = nullptr;  // small scratch (sums)
    cufftHandle fft = 0;
    bool fft_ok = false;
    // fp32 plans whose grid lines fit shared memory: hand-written two-pass pruned FFT fused with
    // the w-screen (fft_planes.cu); d_acc then holds the TRANSPOSED accumulator [npix_y][npix_x]
    HandFft *hfft = nullptr;
    void *d_fftwork = nullptr;
    // chunk classification of the tensor-core degridder: one byte per chunk of 1024 samples of
    // every launch range seen so far (ranges are fixed per plan: slabs x host chunks), filled by
    // the first predict that degrids the range
    unsigned char *d_dense = nullptr;
    long long dense_cap = 0, dense_used = 0;
    std::vector<std::pair<std::pair<long long, long long>, long long>> dense_ranges;   // (begin, end) -> offset
    int tc_mode = 0;
    int key_bits = 0;
    double t_plan_ms = -1, t_last_ms = 0;
    // device time of plan creation; the sort of a single-slab plan is left running on the
    // stream when b2g_plan_create returns, so this may have to wait for it
    double plan_ms()
    {
        if (t_plan_ms < 0 && ev_plan0 && ev_plan1) {
            float ms = 0.f;
            if (cudaEventSynchronize(ev_plan1) == cudaSuccess &&
                cudaEventElapsedTime(&ms, ev_plan0, ev_plan1) == cudaSuccess)
                t_plan_ms = ms;
        }
        return t_plan_ms < 0 ? 0.0 : t_plan_ms;
    }
    double timings[S_COUNT] = {0};
    long long launches_last = 0;
    StageTimer timer;

    VisGeo vis_geo() const
    {
        VisGeo g;
        g.uvw = d_uvw;
        g.sel = d_sel;
        g.fscale = d_fscale;
        g.set_nchan((uint32_t)nchan);
        g.nu = (int)nu;
        g.nv = (int)nv;
        g.nplanes = nplanes;
        g.nslots = nslots;
        g.wstack = wstack;
        g.su = psx;
        g.sv = psy;
        g.sw = wstack ? 1.0 / dw : 0.0;
        g.w0s = wstack ? w0 / dw : 0.0;
        g.plane_stride = (long long)nu * nv;
        return g;
    }
    ImgGeo img_geo() const
    {
        ImgGeo g;
        g.npix_x = (int)npix_x;
        g.npix_y = (int)npix_y;
        g.nu = (int)nu;
        g.nv = (int)nv;
        g.pixsize_x = psx;
        g.pixsize_y = psy;
        g.wstack = wstack;
        g.nslots = nslots;
        g.w0 = w0;
        g.dw = dw;
        g.cu = d_cu;
        g.cv = d_cv;
        g.cheb = d_cheb;
        g.ncheb = ncheb;
        g.nm1min = nm1min;
        return g;
    }
    void release()
    {
        // all plan buffers come from the stream-ordered pool (see pool_setup)
        if (own_uvw && d_uvw) cudaFreeAsync(d_uvw, stream);
        if (d_fscale) cudaFreeAsync(d_fscale, stream);
        if (d_cu) cudaFreeAsync(d_cu, stream);
        if (d_cv) cudaFreeAsync(d_cv, stream);
        if (d_cheb) cudaFreeAsync(d_cheb, stream);
        if (d_sidx) cudaFreeAsync(d_sidx, stream);
        if (own_sel && d_sel) cudaFreeAsync(const_cast<uint32_t *>(d_sel), stream);
        d_sel = nullptr;
        if (d_ring) cudaFreeAsync(d_ring, stream);
        if (d_acc) cudaFreeAsync(d_acc, stream);
        if (d_scratch) cudaFreeAsync(d_scratch, stream);
        if (d_fftwork) cudaFreeAsync(d_fftwork, stream);
        if (d_dense) cudaFreeAsync(d_dense, stream);
        d_dense = nullptr;
        if (hfft) hfft_destroy(hfft, stream);
        hfft = nullptr;
        if (fft_ok) cufftDestroy(fft);
        for (auto e : chunk_ev) cudaEventDestroy(e);
        chunk_ev.clear();
        if (ev_a) cudaEventDestroy(ev_a);
        if (ev_b) cudaEventDestroy(ev_b);
        if (ev_plan0) cudaEventDestroy(ev_plan0);
        if (ev_plan1) cudaEventDestroy(ev_plan1);
        ev_plan0 = ev_plan1 = nullptr;
        if (copy_stream) cudaStreamDestroy(copy_stream);
        ev_a = ev_b = nullptr;
        copy_stream = nullptr;
        d_uvw = d_fscale = d_cu = d_cv = d_cheb = d_scratch = nullptr;
        d_sidx = nullptr;
        d_ring = d_acc = d_fftwork = nullptr;
        fft_ok = false;
    }
};

namespace {

// (W, sigma, ...) for the request; honours an explicit padded grid (nu, nv) if given.
int choose_kernel(double epsilon, bool wstack, int precision, int64_t npix_x, int64_t npix_y,
                  int64_t nu_req, int64_t nv_req, double nvis, double planes_per_sigma,
                  KernelRow *out)
{
    if (!(epsilon > 0.0)) return fail("epsilon must be positive (got %g)", epsilon);
    // Error budget: the aliasing errors of the 2 (3) separable dimensions add in quadrature;
    // 1.75x head-room covers the polynomial fit and the difference between the tabulated
    // mean over |x| <= 1/(2 sigma) and the actual image; `floor` is the rounding noise of
    // the arithmetic type, which no kernel can remove.
    const double floor = precision == B2G_F32 ? 6e-7 : 3e-13;
    const double eps_dim = (epsilon - floor) / (1.75 * std::sqrt(wstack ? 3.0 : 2.0));
    // Small oversampling with a wide kernel makes psi-hat tiny at the image rim: the correction
    // 1/psi-hat(x) psi-hat(y) then amplifies a corner pixel (and the rounding noise around it) by
    // amp^2.  Keep machine-epsilon * amp^2 below a quarter of the requested accuracy.
    const double amp_max = std::sqrt(0.25 * epsilon / (precision == B2G_F32 ? 6e-8 : 1.1e-16));
    const int maxW = precision == B2G_F32 ? 8 : kMaxW;
    const int csize = precision == B2G_F32 ? 8 : 16;
    if (precision == B2G_F32 && epsilon < 2e-6)
        return fail("epsilon=%g is below what single precision supports (>= 2e-6); use "
                    "double precision", epsilon);
    if (epsilon < 4e-13)
        return fail("epsilon=%g is below what double precision supports (>= 4e-13)", epsilon);
    int best = -1;
    // tuning aid: B2G_KERNEL="W,sigma" forces a table row that meets the error budget
    if (const char *force = std::getenv("B2G_KERNEL")) {
        int fw = 0;
        double fsig = 0;
        if (std::sscanf(force, "%d,%lf", &fw, &fsig) == 2)
            for (int i = 0; i < kKernelDBSize; ++i)
                if (kKernelDB[i].W == fw && std::fabs(kKernelDB[i].sigma - fsig) < 1e-6 &&
                    kKernelDB[i].eps <= eps_dim && kKernelDB[i].amp <= amp_max && fw <= maxW)
                    best = i;
    }
    if (best >= 0) {
    } else if (nu_req > 0 && nv_req > 0) {
        const double smax = std::min(double(nu_req) / npix_x, double(nv_req) / npix_y);
        if (smax < 1.15)
            return fail("requested grid %lldx%lld is less than 1.15x the image", (long long)nu_req,
                        (long long)nv_req);
        for (int i = 0; i < kKernelDBSize; ++i) {
            const KernelRow &r = kKernelDB[i];
            if (r.W > maxW || r.eps > eps_dim || r.amp > amp_max || r.sigma > smax + 1e-9) continue;
            if (best < 0 || r.W < kKernelDB[best].W ||
                (r.W == kKernelDB[best].W && r.sigma > kKernelDB[best].sigma))
                best = i;
        }
    } else {
        best = select_kernel(eps_dim, amp_max, wstack, maxW, nvis, npix_x, npix_y, planes_per_sigma, csize);
    }
    if (best < 0)
        return fail("no gridding kernel reaches epsilon=%g in %s precision", epsilon,
                    precision == B2G_F32 ? "single" : "double");
    *out = kKernelDB[best];
    return 0;
}

int build_tables(b2g_plan *p)
{
    // 1/psi-hat along u and v at the image pixels (x in cycles per grid cell)
    // psi-hat is even: evaluate one half, mirror the other; reuse u for v on square grids
    auto table = [&](int64_t npix, int64_t n, std::vector<double> &out) {
        out.resize(npix);
        const int64_t h = npix / 2;
        for (int64_t k = 0; k <= h; ++k) {
            const double c = 1.0 / p->pk.corr(double(k) / double(n));
            out[h - k] = c;                    // pixel h-k sits at -k
            if (h + k < npix) out[h + k] = c;  // pixel h+k at +k
        }
    };
    std::vector<double> cu, cv;
    table(p->npix_x, p->nu, cu);
    if (p->npix_y == p->npix_x && p->nv == p->nu) cv = cu;
    else table(p->npix_y, p->nv, cv);
    CUDA_OK(cudaMallocAsync(&p->d_cu, sizeof(double) * p->npix_x, p->stream));
    CUDA_OK(cudaMallocAsync(&p->d_cv, sizeof(double) * p->npix_y, p->stream));
    CUDA_OK(cudaMemcpyAsync(p->d_cu, cu.data(), sizeof(double) * p->npix_x,
                            cudaMemcpyHostToDevice, p->stream));
    CUDA_OK(cudaMemcpyAsync(p->d_cv, cv.data(), sizeof(double) * p->npix_y,
                            cudaMemcpyHostToDevice, p->stream));
    // 1/psi-hat_w(nm1*dw) as a Chebyshev series in nm1 on [nm1min, 0]
    std::vector<double> cheb(1, 1.0);
    if (p->wstack) {
        const PolyKernel &pk = p->pk;
        const double dw = p->dw;
        cheb = cheb_fit([&](double nm1) { return 1.0 / pk.corr(nm1 * dw); }, p->nm1min, 0.0,
                        1e-16);
    }
    p->ncheb = (int)cheb.size();
    CUDA_OK(cudaMallocAsync(&p->d_cheb, sizeof(double) * cheb.size(), p->stream));
    CUDA_OK(cudaMemcpyAsync(p->d_cheb, cheb.data(), sizeof(double) * cheb.size(),
                            cudaMemcpyHostToDevice, p->stream));
    CUDA_OK(cudaStreamSynchronize(p->stream));  // host vectors go out of scope
    return 0;
}

int build_sort(b2g_plan *p)
{
    const long long n = p->nvis;
    p->slab_begin.assign(p->nslabs + 1, 0);
    if (n == 0) return 0;
    if (n >= (1ll << 32)) return fail("more than 2^32 samples per plan (%lld)", n);
    cudaStream_t s = p->stream;
    uint32_t *idx_tmp = nullptr;
    unsigned long long *counts = nullptr;
    CUDA_OK(cudaMallocAsync(&p->d_sidx, sizeof(uint32_t) * n, s));
    CUDA_OK(cudaMallocAsync(&idx_tmp, sizeof(uint32_t) * n, s));
    const VisGeo g = p->vis_geo();
    const double kmax = double(p->nchunks) * double(p->nplanes) * double(p->nu) * double(p->nv);
    p->key_bits = std::max(1, (int)std::ceil(std::log2(kmax + 1.0)));
    int rc = sort_samples(g, p->W, p->nchunks > 1 ? (uint32_t)p->rows_per_chunk : 0u, idx_tmp,
                          p->d_sidx, n, p->key_bits, s);
    cudaFreeAsync(idx_tmp, s);
    KERNEL_OK(rc);
    if (p->nslabs == 1) {
        // one slab holds every sample: no histogram needed, the sort stays in flight
        p->slab_begin[1] = n;
        return 0;
    }
    CUDA_OK(cudaMallocAsync(&counts, sizeof(unsigned long long) * p->nplanes, s));
    rc = launch_plane_hist(g, p->W, n, counts, s);
    std::vector<unsigned long long> h(p->nplanes, 0);
    cudaError_t e = cudaMemcpyAsync(h.data(), counts, sizeof(unsigned long long) * p->nplanes,
                                    cudaMemcpyDeviceToHost, s);
    cudaFreeAsync(counts, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    KERNEL_OK(rc);
    CUDA_OK(e);
    // slab sl = origin planes [sl*gs, (sl+1)*gs)
    long long run = 0;
    for (int sl = 0; sl < p->nslabs; ++sl) {
        p->slab_begin[sl] = run;
        for (int q = sl * p->gs; q < std::min((sl + 1) * p->gs, p->nplanes); ++q) run += (long long)h[q];
    }
    p->slab_begin[p->nslabs] = run;
    if (run != n) {
        // origin planes beyond nplanes-W cannot occur (locate() clamps); count them anyway
        for (int q = p->nslabs * p->gs; q < p->nplanes; ++q) run += (long long)h[q];
        if (run != n) return fail("internal: plane histogram covers %lld of %lld samples", run, n);
        p->slab_begin[p->nslabs] = run;
    }
    return 0;
}

// The stream-ordered pool keeps freed blocks instead of returning them to the driver at every
// synchronisation: plan creation / destruction per call (the one-shot ms2dirty / dirty2ms
// pattern of the reference) then costs no cudaMalloc / cudaFree round trips after the first.
int pool_setup(int device)
{
    static std::atomic<unsigned> done{0};
    if (device < 0 || device >= 32 || (done.load() & (1u << device))) return 0;
    cudaMemPool_t pool;
    CUDA_OK(cudaDeviceGetDefaultMemPool(&pool, device));
    size_t free_b = 0, total_b = 0;
    CUDA_OK(cudaMemGetInfo(&free_b, &total_b));
    unsigned long long keep = (unsigned long long)(total_b / 4);
    CUDA_OK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    done.fetch_or(1u << device);
    return 0;
}

}  // namespace

namespace b2g {
int wrapper_fail(const char *msg) { return fail("%s", msg); }
}  // namespace b2g

extern "C" const char *b2g_last_error(void) { return g_err.c_str(); }
extern "C" const char *b2g_version(void) { return "b200-wgridder 0.1.0 (sm_100a)"; }

extern "C" int b2g_plan_create(b2g_plan **plan, int64_t nrow, int64_t nchan, const double *uvw,
                               const double *freq, int uvw_space, int64_t npix_x, int64_t npix_y,
                               double pixsize_x, double pixsize_y, int64_t nu, int64_t nv,
                               double epsilon, int do_wstacking, int precision, int algo,
                               int verbosity, void *stream)
{
    return b2g_plan_create_sel(plan, nrow, nchan, uvw, freq, uvw_space, nullptr, 0, npix_x, npix_y,
                               pixsize_x, pixsize_y, nu, nv, epsilon, do_wstacking, precision, algo,
                               verbosity, stream);
}

extern "C" int b2g_plan_create_sel(b2g_plan **plan, int64_t nrow, int64_t nchan, const double *uvw,
                                   const double *freq, int uvw_space, const uint32_t *sel,
                                   int64_t nsel, int64_t npix_x, int64_t npix_y, double pixsize_x,
                                   double pixsize_y, int64_t nu, int64_t nv, double epsilon,
                                   int do_wstacking, int precision, int algo, int verbosity,
                                   void *stream)
{
    g_err.clear();
    if (!plan) return fail("plan pointer is NULL");
    *plan = nullptr;
    if (nrow < 0 || nchan < 1) return fail("bad shape nrow=%lld nchan=%lld", (long long)nrow, (long long)nchan);
    if (nsel < 0 || (nsel > 0 && !sel)) return fail("bad sample selection (nsel=%lld)", (long long)nsel);
    if (sel && nrow * nchan >= (1ll << 32)) return fail("selection needs nrow*nchan < 2^32");
    if (npix_x < 2 || npix_y < 2 || (npix_x & 1) || (npix_y & 1))
        return fail("image dimensions must be even and >= 2 (got %lld x %lld)", (long long)npix_x,
                    (long long)npix_y);
    if (!(pixsize_x > 0.0) || !(pixsize_y > 0.0)) return fail("pixel sizes must be positive");
    if (precision != B2G_F32 && precision != B2G_F64) return fail("bad precision code %d", precision);
    if (nrow > 0 && !uvw) return fail("uvw is NULL");
    if (!freq) return fail("freq is NULL");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail("no CUDA device available (this library has no CPU path)");

    auto wall0 = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (verbosity < 2) return;
        auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[b2g] plan_create %-14s %8.3f ms\n", what,
                std::chrono::duration<double, std::milli>(now - wall0).count());
        wall0 = now;
    };
    b2g_plan *p = new b2g_plan();
    struct Guard {
        b2g_plan *p;
        ~Guard() { if (p) { p->release(); delete p; } }
    } guard{p};
    p->stream = (cudaStream_t)stream;
    p->timer.stream = p->stream;
    CUDA_OK(cudaGetDevice(&p->device));
    B2G_OK(pool_setup(p->device));
    CUDA_OK(cudaEventCreate(&p->ev_plan0));
    CUDA_OK(cudaEventCreate(&p->ev_plan1));
    CUDA_OK(cudaEventRecord(p->ev_plan0, p->stream));
    p->nrow = nrow;
    p->nchan = nchan;
    p->nvis = sel ? nsel : nrow * nchan;
    p->npix_x = npix_x;
    p->npix_y = npix_y;
    p->psx = pixsize_x;
    p->psy = pixsize_y;
    p->epsilon = epsilon;
    p->wstack = do_wstacking ? 1 : 0;
    p->precision = precision;
    p->verbosity = verbosity;

    // frequencies -> f/c on the device
    std::vector<double> fs(nchan);
    if (uvw_space == B2G_DEVICE) {
        CUDA_OK(cudaMemcpy(fs.data(), freq, sizeof(double) * nchan, cudaMemcpyDeviceToHost));
    } else {
        std::memcpy(fs.data(), freq, sizeof(double) * nchan);
    }
    double fsmin = 1e300, fsmax = 0.0;
    for (auto &f : fs) {
        if (!(f > 0.0)) return fail("frequencies must be positive");
        f /= kSpeedOfLight;
        fsmin = std::min(fsmin, f);
        fsmax = std::max(fsmax, f);
    }
    CUDA_OK(cudaMallocAsync(&p->d_fscale, sizeof(double) * nchan, p->stream));
    CUDA_OK(cudaMemcpyAsync(p->d_fscale, fs.data(), sizeof(double) * nchan, cudaMemcpyHostToDevice,
                            p->stream));
    // uvw
    if (uvw_space == B2G_DEVICE) {
        p->d_uvw = const_cast<double *>(uvw);
        p->own_uvw = false;
    } else {
        CUDA_OK(cudaMallocAsync(&p->d_uvw, sizeof(double) * 3 * std::max<int64_t>(nrow, 1), p->stream));
        p->own_uvw = true;
        if (nrow > 0)
            CUDA_OK(cudaMemcpyAsync(p->d_uvw, uvw, sizeof(double) * 3 * nrow,
                                    cudaMemcpyHostToDevice, p->stream));
    }
    CUDA_OK(cudaMallocAsync(&p->d_scratch, sizeof(double) * 2048, p->stream));
    double st[4] = {0, 0, 0, 0};
    if (nrow > 0) {
        KERNEL_OK(launch_uvw_stats(p->d_uvw, nrow, p->d_scratch, p->stream));
        CUDA_OK(cudaMemcpyAsync(st, p->d_scratch, sizeof(st), cudaMemcpyDeviceToHost, p->stream));
    }
    CUDA_OK(cudaStreamSynchronize(p->stream));  // also: fs (host vector) has been consumed
    lap("uvw+stats");
    if (!(st[1] == st[1]) || !(st[2] == st[2]) || !(st[3] == st[3]) || st[1] > 1e300)
        return fail("uvw contains NaN or Inf");
    p->wmin = st[0] * fsmin;
    p->wmax = st[1] * fsmax;
    p->umax_l = st[2] * fsmax;
    p->vmax_l = st[3] * fsmax;
    if (sel) {
        // w-plane-sharded plan: the w-range (hence the planes) of the selected samples only
        if (uvw_space == B2G_DEVICE) {
            p->d_sel = sel;
        } else {
            uint32_t *d = nullptr;
            CUDA_OK(cudaMallocAsync(&d, sizeof(uint32_t) * std::max<int64_t>(nsel, 1), p->stream));
            p->d_sel = d;
            p->own_sel = true;
            if (nsel > 0)
                CUDA_OK(cudaMemcpyAsync(d, sel, sizeof(uint32_t) * nsel, cudaMemcpyHostToDevice, p->stream));
        }
        VisGeo g0;
        std::memset(&g0, 0, sizeof(g0));
        g0.uvw = p->d_uvw;
        g0.fscale = p->d_fscale;
        g0.sel = p->d_sel;
        g0.set_nchan((uint32_t)nchan);
        double ws[2] = {0, 0};
        KERNEL_OK(launch_sel_wstats(g0, nsel, p->d_scratch + 8, p->stream));
        CUDA_OK(cudaMemcpyAsync(ws, p->d_scratch + 8, sizeof(ws), cudaMemcpyDeviceToHost, p->stream));
        CUDA_OK(cudaStreamSynchronize(p->stream));
        p->wmin = ws[0];
        p->wmax = ws[1];
        if (nsel == 0) p->wmin = p->wmax = 0.0;
    }
    if (nrow == 0) p->wmin = p->wmax = 0.0;

    // field of view -> most negative n-1
    {
        const double xm = pixsize_x * double(npix_x / 2), ym = pixsize_y * double(npix_y / 2);
        const double r2 = xm * xm + ym * ym;
        p->nm1min = r2 >= 1.0 ? -1.0 : -r2 / (std::sqrt(1.0 - r2) + 1.0);
    }
    const double planes_per_sigma = p->wstack ? (p->wmax - p->wmin) * 2.0 * std::fabs(p->nm1min) : 0.0;
    KernelRow kr;
    B2G_OK(choose_kernel(epsilon, p->wstack, precision, npix_x, npix_y, nu, nv, double(p->nvis),
                         planes_per_sigma, &kr));
    p->W = kr.W;
    p->sigma = kr.sigma;
    p->beta = kr.beta;
    p->e0 = kr.e0;
    p->eps_kernel = kr.eps;
    p->pk.build(kr.W, kr.beta, kr.e0);
    if (nu > 0 && nv > 0) {
        if ((nu & 1) || (nv & 1)) return fail("nu and nv must be even");
        p->nu = nu;
        p->nv = nv;
    } else {
        p->nu = std::max<int64_t>(good_size_even((int64_t)std::ceil(kr.sigma * npix_x)), 32);
        p->nv = std::max<int64_t>(good_size_even((int64_t)std::ceil(kr.sigma * npix_y)), 32);
    }
    if (p->nu > 32768 || p->nv > 32768) return fail("padded grid %lldx%lld too large", (long long)p->nu, (long long)p->nv);

    // w-planes
    const size_t csize = 2 * real_size(precision);
    p->plane_bytes = size_t(p->nu) * size_t(p->nv) * csize;
    if (p->wstack) {
        p->dw = 0.5 / (kr.sigma * std::max(std::fabs(p->nm1min), 1e-12));
        const double span = (p->wmax - p->wmin) / p->dw;
        if (span > 60000.0) return fail("w range needs %g planes", span);
        p->nplanes = (int)std::ceil(span) + p->W;
        p->w0 = 0.5 * (p->wmin + p->wmax) - 0.5 * double(p->nplanes - 1) * p->dw;
        size_t free_b = 0, total_b = 0;
        CUDA_OK(cudaMemGetInfo(&free_b, &total_b));
        // working set besides the ring: sorted index (4 B/sample) + sort scratch
        // (20 B/sample, released after the sort) + caller's visibilities
        const double avail = 0.35 * double(free_b);
        int gs = (int)(avail / double(p->plane_bytes)) - (p->W - 1);
        // all origin planes in one slab when that is cheap (enables host streaming), else <= 8
        const int all = p->nplanes - p->W + 1;
        gs = (all <= 24 && gs >= all) ? all : std::max(1, std::min(gs, 8));
        gs = std::min(gs, all);
        if (const int cap = g_max_slab_planes.load(); cap > 0) gs = std::min(gs, cap);
        p->gs = gs;
        p->nslots = gs + p->W - 1;
        p->nslabs = (p->nplanes - p->W + 1 + gs - 1) / gs;
    } else {
        p->dw = 0.0;
        p->w0 = 0.0;
        p->nplanes = 1;
        p->gs = 1;
        p->nslots = 1;
        p->nslabs = 1;
    }

    // host-streaming plans: row chunks as the outermost sort component (all planes resident)
    {
        const long long per = g_host_chunk_samples.load();
        if (uvw_space == B2G_HOST && !sel && p->nslabs == 1 && per > 0 && p->nvis >= 2 * per) {
            int nc = (int)std::min<long long>(kMaxChunks, p->nvis / per);
            p->rows_per_chunk = (nrow + nc - 1) / nc;
            p->nchunks = (int)((nrow + p->rows_per_chunk - 1) / p->rows_per_chunk);
        }
        if (p->nchunks > 1) {
            CUDA_OK(cudaStreamCreateWithFlags(&p->copy_stream, cudaStreamNonBlocking));
            p->chunk_ev.resize(p->nchunks);
            for (auto &e : p->chunk_ev) CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            CUDA_OK(cudaEventCreateWithFlags(&p->ev_a, cudaEventDisableTiming));
            CUDA_OK(cudaEventCreateWithFlags(&p->ev_b, cudaEventDisableTiming));
        }
    }

    // gridding kernel family
    bool fast_ok = precision == B2G_F32 ? warprun_supported<float>(p->W, p->wstack)
                                        : warprun_supported<double>(p->W, p->wstack);
    if (algo == B2G_ALGO_AUTO) algo = fast_ok ? B2G_ALGO_WARPRUN : B2G_ALGO_SIMPLE;
    if (algo == B2G_ALGO_WARPRUN && !fast_ok)
        return fail("warp-run kernels do not support W=%d (%s, wstacking=%d)", p->W,
                    precision == B2G_F32 ? "f32" : "f64", p->wstack);
    if (algo != B2G_ALGO_SIMPLE && algo != B2G_ALGO_WARPRUN) return fail("bad algo code %d", algo);
    p->algo = algo;
    p->tc_mode = g_tc_mode.load();

    lap("setup");
    B2G_OK(build_tables(p));
    lap("tables");

    // per-plane transform: the hand-written pruned FFT where it applies, else cuFFT
    if (precision == B2G_F32) {
        // rows of the grid that a footprint can reach (locate_axis: cells f nu - W/2 .. + W, |f| <=
        // max |u| pixsize_x, wrapped): beyond them a plane is identically zero / never degridded
        const double fu = p->umax_l * p->psx;
        int ukeep = 0;
        if (fu < 0.49) ukeep = (int)std::ceil(fu * double(p->nu) + 0.5 * p->W) + 3;
        const double fv = p->vmax_l * p->psy;
        int vkeep = 0;
        if (fv < 0.49) vkeep = (int)std::ceil(fv * double(p->nv) + 0.5 * p->W) + 3;
        const int hrc = hfft_create(&p->hfft, p->img_geo(), ukeep, vkeep, p->stream);
        if (hrc > 0) return fail("hand FFT set-up failed: %s", cudaGetErrorString((cudaError_t)hrc));
    }
    size_t work = 0;
    if (!p->hfft) {
    // FFT plan (own work area so that memory use is explicit)
    FFT_OK(cufftCreate(&p->fft));
    p->fft_ok = true;
    FFT_OK(cufftSetAutoAllocation(p->fft, 0));
    FFT_OK(cufftMakePlan2d(p->fft, (int)p->nu, (int)p->nv,
                           precision == B2G_F32 ? CUFFT_C2C : CUFFT_Z2Z, &work));
    FFT_OK(cufftSetStream(p->fft, p->stream));
    }
    lap("cufft plan");

    // the sort goes last: for a single-slab plan it is still running when this call returns,
    // so the caller's next call (e.g. the H2D copies of b2g_plan_invert) overlaps with it
    B2G_OK(build_sort(p));
    if (work > 0) CUDA_OK(cudaMallocAsync(&p->d_fftwork, work, p->stream));
    if (p->fft_ok) FFT_OK(cufftSetWorkArea(p->fft, p->d_fftwork));
    CUDA_OK(cudaMallocAsync(&p->d_ring, p->plane_bytes * size_t(p->nslots), p->stream));
    CUDA_OK(cudaMallocAsync(&p->d_acc, size_t(npix_x) * size_t(npix_y) * real_size(precision), p->stream));
    CUDA_OK(cudaEventRecord(p->ev_plan1, p->stream));
    lap("sort+alloc");
    if (verbosity > 0)
        fprintf(stderr,
                "[b2g] plan: nvis=%lld W=%d sigma=%.2f grid=%lldx%lld planes=%d (gs=%d, slabs=%d) "
                "dw=%g w=[%g,%g] algo=%d key_bits=%d host_chunks=%d work=%.1f MB plan=%.2f ms\n",
                (long long)p->nvis, p->W, p->sigma, (long long)p->nu, (long long)p->nv, p->nplanes,
                p->gs, p->nslabs, p->dw, p->wmin, p->wmax, p->algo, p->key_bits, p->nchunks,
                work / 1048576.0, p->plan_ms());
    if (verbosity > 0 && p->hfft) {
        char buf[256];
        hfft_describe(p->hfft, buf, sizeof(buf));
        fprintf(stderr, "[b2g] plane transform: hand-written pruned FFT + fused w-screen (%s)\n", buf);
    }
    guard.p = nullptr;
    *plan = p;
    return 0;
}

extern "C" void b2g_plan_destroy(b2g_plan *plan)
{
    if (!plan) return;
    cudaStreamSynchronize(plan->stream);
    plan->release();
    delete plan;
}

namespace {

// A device view of caller data in the plan's precision; owns a staging copy if the caller's
// data is on the host or has the other component type.
template <typename T> struct Staged {
    T *ptr = nullptr;
    bool owned = false;
    cudaStream_t stream = nullptr;
    cudaStream_t filler = nullptr;  // second stream that writes into the buffer, if any
    ~Staged()
    {
        if (owned && ptr) {
            if (filler) cudaStreamSynchronize(filler);
            cudaFreeAsync(ptr, stream);
        }
    }
};

constexpr long long kChunkReals = 1ll << 26;  // staging chunk (reals) for converting copies

// src (n reals of `dtype` in `space`) -> device array of T
template <typename T>
int stage_in(const void *src, int dtype, int space, long long n, cudaStream_t s, Staged<T> *out)
{
    out->stream = s;
    const int want = sizeof(T) == 8 ? B2G_F64 : B2G_F32;
    if (n == 0 || src == nullptr) return 0;
    if (space == B2G_DEVICE && dtype == want) {
        out->ptr = const_cast<T *>(static_cast<const T *>(src));
        return 0;
    }
    CUDA_OK(cudaMallocAsync(&out->ptr, sizeof(T) * n, s));
    out->owned = true;
    if (dtype == want) {
        CUDA_OK(cudaMemcpyAsync(out->ptr, src, sizeof(T) * n, cudaMemcpyHostToDevice, s));
        return 0;
    }
    // other component type: convert on the device, chunk by chunk
    if (dtype == B2G_F64) {
        const double *sp = static_cast<const double *>(src);
        if (space == B2G_DEVICE) {
            KERNEL_OK((launch_convert<double, float>(sp, reinterpret_cast<float *>(out->ptr), n, s)));
            return 0;
        }
        const long long cn = std::min(n, kChunkReals);
        double *tmp = nullptr;
        CUDA_OK(cudaMallocAsync(&tmp, sizeof(double) * cn, s));
        for (long long o = 0; o < n; o += cn) {
            const long long m = std::min(cn, n - o);
            CUDA_OK(cudaMemcpyAsync(tmp, sp + o, sizeof(double) * m, cudaMemcpyHostToDevice, s));
            KERNEL_OK((launch_convert<double, float>(tmp, reinterpret_cast<float *>(out->ptr) + o, m, s)));
        }
        CUDA_OK(cudaFreeAsync(tmp, s));
    } else {
        const float *sp = static_cast<const float *>(src);
        if (space == B2G_DEVICE) {
            KERNEL_OK((launch_convert<float, double>(sp, reinterpret_cast<double *>(out->ptr), n, s)));
            return 0;
        }
        const long long cn = std::min(n, kChunkReals);
        float *tmp = nullptr;
        CUDA_OK(cudaMallocAsync(&tmp, sizeof(float) * cn, s));
        for (long long o = 0; o < n; o += cn) {
            const long long m = std::min(cn, n - o);
            CUDA_OK(cudaMemcpyAsync(tmp, sp + o, sizeof(float) * m, cudaMemcpyHostToDevice, s));
            KERNEL_OK((launch_convert<float, double>(tmp, reinterpret_cast<double *>(out->ptr) + o, m, s)));
        }
        CUDA_OK(cudaFreeAsync(tmp, s));
    }
    return 0;
}

// device array of T (n reals) -> dst (`dtype` in `space`); no-op if dst aliases src
template <typename T>
int stage_out(const T *src, void *dst, int dtype, int space, long long n, cudaStream_t s)
{
    const int have = sizeof(T) == 8 ? B2G_F64 : B2G_F32;
    if (n == 0 || static_cast<const void *>(src) == dst) return 0;
    if (dtype == have) {
        CUDA_OK(cudaMemcpyAsync(dst, src, sizeof(T) * n,
                                space == B2G_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, s));
        return 0;
    }
    if (dtype == B2G_F64) {  // T == float
        double *dp = static_cast<double *>(dst);
        const float *sp = reinterpret_cast<const float *>(src);
        if (space == B2G_DEVICE) {
            KERNEL_OK((launch_convert<float, double>(sp, dp, n, s)));
            return 0;
        }
        const long long cn = std::min(n, kChunkReals);
        double *tmp = nullptr;
        CUDA_OK(cudaMallocAsync(&tmp, sizeof(double) * cn, s));
        for (long long o = 0; o < n; o += cn) {
            const long long m = std::min(cn, n - o);
            KERNEL_OK((launch_convert<float, double>(sp + o, tmp, m, s)));
            CUDA_OK(cudaMemcpyAsync(dp + o, tmp, sizeof(double) * m, cudaMemcpyDeviceToHost, s));
        }
        CUDA_OK(cudaFreeAsync(tmp, s));
    } else {  // T == double, dst float
        float *dp = static_cast<float *>(dst);
        const double *sp = reinterpret_cast<const double *>(src);
        if (space == B2G_DEVICE) {
            KERNEL_OK((launch_convert<double, float>(sp, dp, n, s)));
            return 0;
        }
        const long long cn = std::min(n, kChunkReals);
        float *tmp = nullptr;
        CUDA_OK(cudaMallocAsync(&tmp, sizeof(float) * cn, s));
        for (long long o = 0; o < n; o += cn) {
            const long long m = std::min(cn, n - o);
            KERNEL_OK((launch_convert<double, float>(sp + o, tmp, m, s)));
            CUDA_OK(cudaMemcpyAsync(dp + o, tmp, sizeof(float) * m, cudaMemcpyDeviceToHost, s));
        }
        CUDA_OK(cudaFreeAsync(tmp, s));
    }
    return 0;
}

// degridding of one launch range with the plan's kernel family
template <typename T> int degrid_dispatch(b2g_plan *p, VisLaunch<T> &L)
{
    if (p->algo == B2G_ALGO_WARPRUN) KERNEL_OK(launch_degrid_warprun<T>(L));
    else KERNEL_OK(launch_degrid_simple<T>(L));
    p->launches_last++;
    return 0;
}
// fp32: chunks with long runs go to the tensor-core kernel, the rest to the warp-run kernel
template <> int degrid_dispatch<float>(b2g_plan *p, VisLaunch<float> &L)
{
    if (p->algo == B2G_ALGO_WARPRUN && p->tc_mode > 0 && tc_supported(p->W, p->wstack != 0)) {
        unsigned char *dense = nullptr;
        bool classify = false;
        if (p->tc_mode == 1) {
            if (!p->d_dense) {
                p->dense_cap = tc_chunks(p->nvis) + 4 * (long long)(p->nslabs + 1) * std::max(1, p->nchunks) + 64;
                CUDA_OK(cudaMallocAsync(&p->d_dense, (size_t)p->dense_cap, p->stream));
            }
            long long off = -1;
            for (const auto &r : p->dense_ranges)
                if (r.first.first == L.begin && r.first.second == L.end) off = r.second;
            if (off < 0) {
                const long long need = tc_chunks(L.end - L.begin);
                if (p->dense_used + need > p->dense_cap) return fail("tensor-core chunk table overflow");
                off = p->dense_used;
                p->dense_used += need;
                p->dense_ranges.push_back({{L.begin, L.end}, off});
                classify = true;
            }
            dense = p->d_dense + off;
        }
        static const int max_runs = [] {
            const char *e = std::getenv("B2G_TC_MAX_RUNS");
            return e ? std::atoi(e) : kTcMaxRuns;
        }();
        static const int sparse_runs = [] {
            const char *e = std::getenv("B2G_SPARSE_RUNS");
            return e ? std::atoi(e) : kSparseRuns;
        }();
        KERNEL_OK(launch_degrid_tc(L, dense, classify ? max_runs : -1, sparse_runs));
        p->launches_last += classify ? 2 : 1;
        if (!dense) return 0;
        // the chunks the tensor-core kernel left: hardly any shared footprints -> one thread per
        // sample; the rest -> warp-run
        L.chunk_class = dense;
        KERNEL_OK(launch_degrid_warprun<float>(L));
        KERNEL_OK(launch_degrid_simple<float>(L));
        L.chunk_class = nullptr;
        p->launches_last += 2;
        return 0;
    }
    if (p->algo == B2G_ALGO_WARPRUN) KERNEL_OK(launch_degrid_warprun<float>(L));
    else KERNEL_OK(launch_degrid_simple<float>(L));
    p->launches_last++;
    return 0;
}

template <typename T> VisLaunch<T> make_launch(b2g_plan *p, int slab)
{
    VisLaunch<T> L;
    L.geo = p->vis_geo();
    L.W = p->W;
    L.wstack = p->wstack != 0;
    L.coef = p->pk.c.data();
    L.sidx = p->d_sidx;
    L.begin = p->slab_begin.empty() ? 0 : p->slab_begin[slab];
    L.end = p->slab_begin.empty() ? 0 : p->slab_begin[slab + 1];
    L.plane0 = slab * p->gs;
    L.slot0 = L.plane0 % p->nslots;
    L.ms = nullptr;
    L.wgt = nullptr;
    L.ms_out = nullptr;
    L.ring = static_cast<cplx<T> *>(p->d_ring);
    L.stream = p->stream;
    return L;
}

template <typename T> cplx<T> *slot_ptr(b2g_plan *p, int plane)
{
    return static_cast<cplx<T> *>(p->d_ring) + (size_t)(plane % p->nslots) * size_t(p->nu) * size_t(p->nv);
}

// Zero the ring slot of `plane` for gridding.  With the hand-written transform only the rows of
// the grid that a footprint can reach are ever read or written (fft_planes.cu: ukeep), so only
// those are cleared: the two ends of the plane.
template <typename T> int clear_slot(b2g_plan *p, int plane)
{
    cplx<T> *d = slot_ptr<T>(p, plane);
    const int ukeep = p->hfft ? hfft_ukeep(p->hfft) : 0, vkeep = p->hfft ? hfft_vkeep(p->hfft) : 0;
    if (ukeep > 0 && 2 * (int64_t)ukeep < p->nu && vkeep > 0 && 2 * (int64_t)vkeep < p->nv) {
        // the four corners of the plane (rows and columns both wrap around 0)
        const size_t pitch = size_t(p->nv) * sizeof(cplx<T>), width = size_t(vkeep) * sizeof(cplx<T>);
        cplx<T> *hi = d + size_t(p->nu - ukeep) * size_t(p->nv);
        CUDA_OK(cudaMemset2DAsync(d, pitch, 0, width, ukeep, p->stream));
        CUDA_OK(cudaMemset2DAsync(d + (p->nv - vkeep), pitch, 0, width, ukeep, p->stream));
        CUDA_OK(cudaMemset2DAsync(hi, pitch, 0, width, ukeep, p->stream));
        CUDA_OK(cudaMemset2DAsync(hi + (p->nv - vkeep), pitch, 0, width, ukeep, p->stream));
    } else if (ukeep > 0 && 2 * (int64_t)ukeep < p->nu) {
        const size_t rows = size_t(ukeep) * size_t(p->nv);
        CUDA_OK(cudaMemsetAsync(d, 0, rows * sizeof(cplx<T>), p->stream));
        CUDA_OK(cudaMemsetAsync(d + size_t(p->nu - ukeep) * size_t(p->nv), 0, rows * sizeof(cplx<T>), p->stream));
    } else {
        CUDA_OK(cudaMemsetAsync(d, 0, p->plane_bytes, p->stream));
    }
    return 0;
}

template <typename T> int fft_plane(b2g_plan *p, int plane, int direction)
{
    cplx<T> *d = slot_ptr<T>(p, plane);
    if (sizeof(T) == 4)
        FFT_OK(cufftExecC2C(p->fft, reinterpret_cast<cufftComplex *>(d),
                            reinterpret_cast<cufftComplex *>(d), direction));
    else
        FFT_OK(cufftExecZ2Z(p->fft, reinterpret_cast<cufftDoubleComplex *>(d),
                            reinterpret_cast<cufftDoubleComplex *>(d), direction));
    return 0;
}

// fp32 plans with the hand-written FFT: per-plane fused transform + w-screen
template <typename T> int hfft_planes_invert(b2g_plan *, int, int, bool) { return -1; }
template <> int hfft_planes_invert<float>(b2g_plan *p, int first, int count, bool clear)
{
    for (int q = 0; q < count; ++q) {
        const double w = p->w0 + double(first + q) * p->dw;
        KERNEL_OK(hfft_invert_plane(p->hfft, slot_ptr<float>(p, first + q), w,
                                    static_cast<float *>(p->d_acc), clear, p->stream));
        p->launches_last += 2;
    }
    return 0;
}
template <typename T> int hfft_planes_predict(b2g_plan *, int, int) { return -1; }
template <> int hfft_planes_predict<float>(b2g_plan *p, int first, int count)
{
    for (int q = 0; q < count; ++q) {
        const double w = p->w0 + double(first + q) * p->dw;
        KERNEL_OK(hfft_predict_plane(p->hfft, static_cast<const float *>(p->d_acc), w,
                                     slot_ptr<float>(p, first + q), p->stream));
        p->launches_last += 2;
    }
    return 0;
}
template <typename T, typename Tout> int hfft_correct_any(b2g_plan *, const ImgGeo &, Tout *) { return -1; }
template <> int hfft_correct_any<float, float>(b2g_plan *p, const ImgGeo &g, float *out)
{
    return hfft_correct<float>(p->hfft, g, static_cast<const float *>(p->d_acc), out, p->stream);
}
template <> int hfft_correct_any<float, double>(b2g_plan *p, const ImgGeo &g, double *out)
{
    return hfft_correct<double>(p->hfft, g, static_cast<const float *>(p->d_acc), out, p->stream);
}
template <typename T, typename Tin> int hfft_precorrect_any(b2g_plan *, const ImgGeo &, const Tin *) { return -1; }
template <> int hfft_precorrect_any<float, float>(b2g_plan *p, const ImgGeo &g, const float *in)
{
    return hfft_precorrect<float>(p->hfft, g, in, static_cast<float *>(p->d_acc), p->stream);
}
template <> int hfft_precorrect_any<float, double>(b2g_plan *p, const ImgGeo &g, const double *in)
{
    return hfft_precorrect<double>(p->hfft, g, in, static_cast<float *>(p->d_acc), p->stream);
}

// planes finished by slab s (invert) : [first, first+count)
void slab_planes_done(const b2g_plan *p, int s, int *first, int *count)
{
    *first = s * p->gs;
    *count = (s == p->nslabs - 1) ? p->nplanes - *first : p->gs;
}
// planes that must be produced before slab s can be degridded (predict)
void slab_planes_new(const b2g_plan *p, int s, int *first, int *count)
{
    if (s == 0) {
        *first = 0;
        *count = std::min(p->nslots, p->nplanes);
    } else {
        *first = s * p->gs + p->W - 1;
        *count = std::min(p->gs, p->nplanes - *first);
        if (*count < 0) *count = 0;
    }
}

template <typename T>
int invert_impl(b2g_plan *p, const void *ms, const void *wgt, int vis_dtype, int vis_space,
                void *dirty, int dirty_dtype, int dirty_space)
{
    StageTimer &tm = p->timer;
    cudaStream_t s = p->stream;
    const long long n = p->nvis;
    p->launches_last = 0;
    tm.reset();
    tm.begin(S_TOTAL);
    const size_t tot_idx = tm.spans.size() - 1;

    const int want = sizeof(T) == 8 ? B2G_F64 : B2G_F32;
    // host arrays of the plan's own type are streamed chunk by chunk, overlapped with gridding
    const bool stream_in = p->nchunks > 1 && vis_space == B2G_HOST && vis_dtype == want && n > 0;
    Staged<T> d_ms, d_wgt;
    if (stream_in) {
        d_ms.stream = d_wgt.stream = s;
        d_ms.filler = d_wgt.filler = p->copy_stream;
        // allocated on the copy stream: the copies must not wait for what is still queued on
        // the compute stream (the plan's sort); the gridding launches order themselves behind
        // the copies through the chunk events
        CUDA_OK(cudaMallocAsync(&d_ms.ptr, sizeof(T) * 2 * n, p->copy_stream));
        d_ms.owned = true;
        if (wgt) {
            CUDA_OK(cudaMallocAsync(&d_wgt.ptr, sizeof(T) * n, p->copy_stream));
            d_wgt.owned = true;
        }
        for (int c = 0; c < p->nchunks; ++c) {
            const long long r0 = std::min<long long>(p->nrow, c * p->rows_per_chunk);
            const long long r1 = std::min<long long>(p->nrow, (c + 1) * p->rows_per_chunk);
            const long long o = r0 * p->nchan, m = (r1 - r0) * p->nchan;
            CUDA_OK(cudaMemcpyAsync(d_ms.ptr + 2 * o, static_cast<const T *>(ms) + 2 * o,
                                    sizeof(T) * 2 * m, cudaMemcpyHostToDevice, p->copy_stream));
            if (wgt)
                CUDA_OK(cudaMemcpyAsync(d_wgt.ptr + o, static_cast<const T *>(wgt) + o, sizeof(T) * m,
                                        cudaMemcpyHostToDevice, p->copy_stream));
            CUDA_OK(cudaEventRecord(p->chunk_ev[c], p->copy_stream));
        }
    } else {
        tm.begin(S_H2D);
        B2G_OK(stage_in<T>(ms, vis_dtype, vis_space, 2 * n, s, &d_ms));
        B2G_OK(stage_in<T>(wgt, vis_dtype, vis_space, n, s, &d_wgt));
        tm.end();
    }

    tm.begin(S_PREP);
    for (int q = 0; q < p->nslots; ++q) B2G_OK(clear_slot<T>(p, q));
    if (p->hfft)   // the fused column pass accumulates into the (transposed) image from plane 0 on
        CUDA_OK(cudaMemsetAsync(p->d_acc, 0, size_t(p->npix_x) * size_t(p->npix_y) * sizeof(T), s));
    tm.end();

    const ImgGeo ig = p->img_geo();
    for (int sl = 0; sl < p->nslabs; ++sl) {
        VisLaunch<T> L = make_launch<T>(p, sl);
        L.ms = reinterpret_cast<const cplx<T> *>(d_ms.ptr);
        L.wgt = d_wgt.ptr;
        tm.begin(S_VIS);
        // one launch per row chunk (nslabs == 1), each behind its own H2D copy when streaming
        for (int c = 0; c < p->nchunks; ++c) {
            if (p->nchunks > 1) {
                L.begin = std::min<long long>(p->nrow, c * p->rows_per_chunk) * p->nchan;
                L.end = std::min<long long>(p->nrow, (c + 1) * p->rows_per_chunk) * p->nchan;
                if (stream_in) CUDA_OK(cudaStreamWaitEvent(s, p->chunk_ev[c], 0));
            }
            if (L.end > L.begin) {
                if (p->algo == B2G_ALGO_WARPRUN) KERNEL_OK(launch_grid_warprun<T>(L));
                else KERNEL_OK(launch_grid_simple<T>(L));
                p->launches_last++;
            }
        }
        tm.end();
        int first, count;
        slab_planes_done(p, sl, &first, &count);
        if (p->hfft) {
            tm.begin(S_FFT);
            // B2G_FUSED_CLEAR=1: the slots a later slab reuses are cleared by the transform itself
            // (its rows pass zeroes every row it has loaded) instead of by the memsets below.  Off
            // by default: on most boxes it is 2.5 ms per step faster (plane transforms + 19 ms,
            // memsets - 22 ms), but the transform's time then varies from box to box (429 ... 496 ms
            // for the same build, against 407 ... 410 ms without)
            static const bool fused_clear = [] {
                const char *e = std::getenv("B2G_FUSED_CLEAR");
                return e && e[0] == '1';
            }();
            B2G_OK(hfft_planes_invert<T>(p, first, count, fused_clear && sl + 1 < p->nslabs));
            tm.end();
            if (!fused_clear && sl + 1 < p->nslabs) {
                tm.begin(S_PREP);
                for (int q = 0; q < count; ++q)
                    B2G_OK(clear_slot<T>(p, first + q));
                tm.end();
            }
        } else {
            tm.begin(S_FFT);
            for (int q = 0; q < count; ++q) B2G_OK(fft_plane<T>(p, first + q, CUFFT_INVERSE));
            tm.end();
            tm.begin(S_SCREEN);
            KERNEL_OK(launch_screen_accumulate<T>(ig, static_cast<const cplx<T> *>(p->d_ring), first,
                                                  count, static_cast<T *>(p->d_acc), sl == 0, s));
            p->launches_last++;
            tm.end();
        }
        if (sl + 1 < p->nslabs && !p->hfft) {
            tm.begin(S_PREP);
            for (int q = 0; q < count; ++q)
                B2G_OK(clear_slot<T>(p, first + q));
            tm.end();
        }
    }

    tm.begin(S_CORRECT);
    void *d_out = dirty;
    const long long npx = p->npix_x * p->npix_y;
    if (dirty_space != B2G_DEVICE) CUDA_OK(cudaMallocAsync(&d_out, real_size(dirty_dtype) * npx, s));
    if (p->hfft) {
        if (dirty_dtype == B2G_F64) KERNEL_OK((hfft_correct_any<T, double>(p, ig, static_cast<double *>(d_out))));
        else KERNEL_OK((hfft_correct_any<T, float>(p, ig, static_cast<float *>(d_out))));
    } else if (dirty_dtype == B2G_F64)
        KERNEL_OK((launch_correct<T, double>(ig, static_cast<const T *>(p->d_acc), static_cast<double *>(d_out), s)));
    else
        KERNEL_OK((launch_correct<T, float>(ig, static_cast<const T *>(p->d_acc), static_cast<float *>(d_out), s)));
    p->launches_last++;
    tm.end();
    if (dirty_space != B2G_DEVICE) {
        tm.begin(S_D2H);
        CUDA_OK(cudaMemcpyAsync(dirty, d_out, real_size(dirty_dtype) * npx, cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaFreeAsync(d_out, s));
        tm.end();
    }
    if (tm.enabled) cudaEventRecord(tm.spans[tot_idx].b, s);
    CUDA_OK(cudaStreamSynchronize(s));
    tm.collect(p->timings);
    p->t_last_ms = p->timings[S_TOTAL];
    return 0;
}

template <typename T>
int predict_impl(b2g_plan *p, const void *dirty, int dirty_dtype, int dirty_space,
                 const void *wgt, void *ms_out, int vis_dtype, int vis_space,
                 const void *sub_from = nullptr, int sub_dtype = 0, int sub_space = 0)
{
    StageTimer &tm = p->timer;
    cudaStream_t s = p->stream;
    const long long n = p->nvis;
    const int want = sizeof(T) == 8 ? B2G_F64 : B2G_F32;
    p->launches_last = 0;
    tm.reset();
    tm.begin(S_TOTAL);
    const size_t tot_idx = tm.spans.size() - 1;
    const long long npx = p->npix_x * p->npix_y;
    const ImgGeo ig = p->img_geo();

    tm.begin(S_H2D);
    const void *d_dirty = dirty;
    void *d_dirty_tmp = nullptr;
    if (dirty_space != B2G_DEVICE) {
        CUDA_OK(cudaMallocAsync(&d_dirty_tmp, real_size(dirty_dtype) * npx, s));
        CUDA_OK(cudaMemcpyAsync(d_dirty_tmp, dirty, real_size(dirty_dtype) * npx, cudaMemcpyHostToDevice, s));
        d_dirty = d_dirty_tmp;
    }
    Staged<T> d_wgt, d_sub;
    B2G_OK(stage_in<T>(wgt, vis_dtype, vis_space, n, s, &d_wgt));
    // residual mode (b2g_plan_residual): ms_out = sub_from - model
    B2G_OK(stage_in<T>(sub_from, sub_dtype, sub_space, 2 * n, s, &d_sub));
    tm.end();

    tm.begin(S_CORRECT);
    if (p->hfft) {
        if (dirty_dtype == B2G_F64) KERNEL_OK((hfft_precorrect_any<T, double>(p, ig, static_cast<const double *>(d_dirty))));
        else KERNEL_OK((hfft_precorrect_any<T, float>(p, ig, static_cast<const float *>(d_dirty))));
    } else if (dirty_dtype == B2G_F64)
        KERNEL_OK((launch_precorrect<T, double>(ig, static_cast<const double *>(d_dirty), static_cast<T *>(p->d_acc), s)));
    else
        KERNEL_OK((launch_precorrect<T, float>(ig, static_cast<const float *>(d_dirty), static_cast<T *>(p->d_acc), s)));
    p->launches_last++;
    if (d_dirty_tmp) CUDA_OK(cudaFreeAsync(d_dirty_tmp, s));
    tm.end();

    cplx<T> *d_out = static_cast<cplx<T> *>(ms_out);
    bool own_out = false;
    if (n > 0 && !(vis_space == B2G_DEVICE && vis_dtype == want)) {
        CUDA_OK(cudaMallocAsync(&d_out, sizeof(cplx<T>) * n, s));
        own_out = true;
    }
    // host output of the plan's own type: each row chunk goes back while the next is degridded
    const bool stream_out = own_out && p->nchunks > 1 && vis_space == B2G_HOST && vis_dtype == want;

    for (int sl = 0; sl < p->nslabs; ++sl) {
        int first, count;
        slab_planes_new(p, sl, &first, &count);
        if (count > 0 && p->hfft) {
            tm.begin(S_FFT);
            B2G_OK(hfft_planes_predict<T>(p, first, count));
            tm.end();
        } else if (count > 0) {
            tm.begin(S_SCREEN);
            KERNEL_OK(launch_screen_apply<T>(ig, static_cast<const T *>(p->d_acc),
                                             static_cast<cplx<T> *>(p->d_ring), first, count, s));
            p->launches_last++;
            tm.end();
            tm.begin(S_FFT);
            for (int q = 0; q < count; ++q) B2G_OK(fft_plane<T>(p, first + q, CUFFT_FORWARD));
            tm.end();
        }
        VisLaunch<T> L = make_launch<T>(p, sl);
        L.wgt = d_wgt.ptr;
        L.ms_out = d_out;
        L.sub_from = reinterpret_cast<const cplx<T> *>(d_sub.ptr);
        tm.begin(S_VIS);
        for (int c = 0; c < p->nchunks; ++c) {
            if (p->nchunks > 1) {
                L.begin = std::min<long long>(p->nrow, c * p->rows_per_chunk) * p->nchan;
                L.end = std::min<long long>(p->nrow, (c + 1) * p->rows_per_chunk) * p->nchan;
            }
            if (L.end > L.begin) B2G_OK(degrid_dispatch<T>(p, L));
            if (stream_out && L.end > L.begin) {
                CUDA_OK(cudaEventRecord(p->chunk_ev[c], s));
                CUDA_OK(cudaStreamWaitEvent(p->copy_stream, p->chunk_ev[c], 0));
                CUDA_OK(cudaMemcpyAsync(static_cast<cplx<T> *>(ms_out) + L.begin, d_out + L.begin,
                                        sizeof(cplx<T>) * (L.end - L.begin), cudaMemcpyDeviceToHost,
                                        p->copy_stream));
            }
        }
        tm.end();
    }
    if (stream_out) {
        CUDA_OK(cudaEventRecord(p->ev_b, p->copy_stream));
        CUDA_OK(cudaStreamWaitEvent(s, p->ev_b, 0));
        CUDA_OK(cudaFreeAsync(d_out, s));
    } else if (own_out) {
        tm.begin(S_D2H);
        B2G_OK(stage_out<T>(reinterpret_cast<const T *>(d_out), ms_out, vis_dtype, vis_space, 2 * n, s));
        CUDA_OK(cudaFreeAsync(d_out, s));
        tm.end();
    }
    if (tm.enabled) cudaEventRecord(tm.spans[tot_idx].b, s);
    CUDA_OK(cudaStreamSynchronize(s));
    tm.collect(p->timings);
    p->t_last_ms = p->timings[S_TOTAL];
    return 0;
}

}  // namespace

static int check_codes(int dtype, int space, const char *what)
{
    if (dtype != B2G_F32 && dtype != B2G_F64) return fail("bad %s dtype code %d", what, dtype);
    if (space != B2G_HOST && space != B2G_DEVICE) return fail("bad %s memory-space code %d", what, space);
    return 0;
}

extern "C" int b2g_plan_invert(b2g_plan *plan, const void *ms, const void *wgt, int vis_dtype,
                               int vis_space, void *dirty, int dirty_dtype, int dirty_space)
{
    g_err.clear();
    if (!plan) return fail("plan is NULL");
    if (!dirty) return fail("dirty is NULL");
    if (plan->nvis > 0 && !ms) return fail("ms is NULL");
    B2G_OK(check_codes(vis_dtype, vis_space, "visibility"));
    B2G_OK(check_codes(dirty_dtype, dirty_space, "image"));
    CUDA_OK(cudaSetDevice(plan->device));
    return plan->precision == B2G_F32
               ? invert_impl<float>(plan, ms, wgt, vis_dtype, vis_space, dirty, dirty_dtype, dirty_space)
               : invert_impl<double>(plan, ms, wgt, vis_dtype, vis_space, dirty, dirty_dtype, dirty_space);
}

extern "C" int b2g_plan_predict(b2g_plan *plan, const void *dirty, int dirty_dtype, int dirty_space,
                                const void *wgt, void *ms_out, int vis_dtype, int vis_space)
{
    g_err.clear();
    if (!plan) return fail("plan is NULL");
    if (!dirty) return fail("dirty is NULL");
    if (plan->nvis > 0 && !ms_out) return fail("ms_out is NULL");
    B2G_OK(check_codes(vis_dtype, vis_space, "visibility"));
    B2G_OK(check_codes(dirty_dtype, dirty_space, "image"));
    CUDA_OK(cudaSetDevice(plan->device));
    return plan->precision == B2G_F32
               ? predict_impl<float>(plan, dirty, dirty_dtype, dirty_space, wgt, ms_out, vis_dtype, vis_space)
               : predict_impl<double>(plan, dirty, dirty_dtype, dirty_space, wgt, ms_out, vis_dtype, vis_space);
}

namespace {

// residual image = invert(ms - predict(model)): the degridder subtracts the model from the data
// as it writes each sample, and the residual samples go straight into the gridder without
// leaving HBM
template <typename T>
int residual_impl(b2g_plan *p, const void *model, int dirty_dtype, int dirty_space, const void *ms,
                  const void *wgt, int vis_dtype, int vis_space, void *ms_resid, void *dirty_out)
{
    const long long n = p->nvis;
    const int want = sizeof(T) == 8 ? B2G_F64 : B2G_F32;
    cudaStream_t s = p->stream;
    // residual samples in the plan's own type, on the device; the caller's array if it is one
    T *d_res = nullptr;
    bool own = false;
    if (ms_resid && vis_space == B2G_DEVICE && vis_dtype == want) {
        d_res = static_cast<T *>(ms_resid);
    } else if (n > 0) {
        CUDA_OK(cudaMallocAsync(&d_res, sizeof(T) * 2 * n, s));
        own = true;
    }
    int rc = predict_impl<T>(p, model, dirty_dtype, dirty_space, nullptr, d_res, want, B2G_DEVICE, ms,
                             vis_dtype, vis_space);
    double t_pred[8];
    for (int i = 0; i < 8; ++i) t_pred[i] = p->timings[i];
    const int l_pred = p->launches_last;
    // the weights in the plan's type on the device (the residual samples already are)
    Staged<T> d_wgt;
    if (rc == 0) rc = stage_in<T>(wgt, vis_dtype, vis_space, n, s, &d_wgt);
    if (rc == 0)
        rc = invert_impl<T>(p, d_res, d_wgt.ptr, want, B2G_DEVICE, dirty_out, dirty_dtype, dirty_space);
    if (rc == 0 && ms_resid && own) rc = stage_out<T>(d_res, ms_resid, vis_dtype, vis_space, 2 * n, s);
    if (own) {
        cudaStreamSynchronize(s);
        cudaFreeAsync(d_res, s);
    }
    if (rc) return rc;
    for (int i = 0; i < 8; ++i) p->timings[i] += t_pred[i];
    p->launches_last += l_pred;
    p->t_last_ms = p->timings[S_TOTAL];
    return 0;
}

}  // namespace

extern "C" int b2g_plan_residual(b2g_plan *plan, const void *model, int dirty_dtype, int dirty_space,
                                 const void *ms, const void *wgt, int vis_dtype, int vis_space,
                                 void *ms_resid, void *dirty_out)
{
    g_err.clear();
    if (!plan) return fail("plan is NULL");
    if (!model || !dirty_out) return fail("model / dirty_out is NULL");
    if (plan->nvis > 0 && !ms) return fail("ms is NULL");
    B2G_OK(check_codes(vis_dtype, vis_space, "visibility"));
    B2G_OK(check_codes(dirty_dtype, dirty_space, "image"));
    CUDA_OK(cudaSetDevice(plan->device));
    return plan->precision == B2G_F32
               ? residual_impl<float>(plan, model, dirty_dtype, dirty_space, ms, wgt, vis_dtype,
                                      vis_space, ms_resid, dirty_out)
               : residual_impl<double>(plan, model, dirty_dtype, dirty_space, ms, wgt, vis_dtype,
                                       vis_space, ms_resid, dirty_out);
}

extern "C" int b2g_plan_info(const b2g_plan *p, int64_t *iinfo, int ni, double *dinfo, int nd)
{
    g_err.clear();
    if (!p) return fail("plan is NULL");
    const_cast<b2g_plan *>(p)->plan_ms();
    const int64_t iv[12] = {p->W, p->nu, p->nv, p->nplanes, p->nslabs, p->gs, p->algo,
                            poly_degree(p->W), p->nvis, p->key_bits, p->launches_last,
                            p->hfft ? 1 : 0};
    const double dv[9] = {p->sigma, p->eps_kernel, p->dw, p->wmin, p->wmax, p->beta, p->e0,
                          p->t_plan_ms, p->t_last_ms};
    for (int i = 0; i < ni && i < 12; ++i) iinfo[i] = iv[i];
    for (int i = 0; i < nd && i < 9; ++i) dinfo[i] = dv[i];
    return 0;
}

extern "C" int b2g_plan_timings(const b2g_plan *p, double *ms_out, int n)
{
    g_err.clear();
    if (!p) return fail("plan is NULL");
    for (int i = 0; i < n && i < S_COUNT; ++i) ms_out[i] = p->timings[i];
    return 0;
}

extern "C" int b2g_ms2dirty(int64_t nrow, int64_t nchan, const double *uvw, const double *freq,
                            const void *ms, const void *wgt, int vis_dtype, int space,
                            int64_t npix_x, int64_t npix_y, double pixsize_x, double pixsize_y,
                            int64_t nu, int64_t nv, double epsilon, int do_wstacking, int precision,
                            void *dirty, int dirty_dtype, void *stream)
{
    b2g_plan *p = nullptr;
    B2G_OK(b2g_plan_create(&p, nrow, nchan, uvw, freq, space, npix_x, npix_y, pixsize_x, pixsize_y,
                           nu, nv, epsilon, do_wstacking, precision, B2G_ALGO_AUTO, 0, stream));
    int rc = b2g_plan_invert(p, ms, wgt, vis_dtype, space, dirty, dirty_dtype, space);
    b2g_plan_destroy(p);
    return rc;
}

extern "C" int b2g_dirty2ms(int64_t nrow, int64_t nchan, const double *uvw, const double *freq,
                            const void *dirty, int dirty_dtype, const void *wgt, int vis_dtype,
                            int space, int64_t npix_x, int64_t npix_y, double pixsize_x,
                            double pixsize_y, int64_t nu, int64_t nv, double epsilon,
                            int do_wstacking, int precision, void *ms_out, void *stream)
{
    b2g_plan *p = nullptr;
    B2G_OK(b2g_plan_create(&p, nrow, nchan, uvw, freq, space, npix_x, npix_y, pixsize_x, pixsize_y,
                           nu, nv, epsilon, do_wstacking, precision, B2G_ALGO_AUTO, 0, stream));
    int rc = b2g_plan_predict(p, dirty, dirty_dtype, space, wgt, ms_out, vis_dtype, space);
    b2g_plan_destroy(p);
    return rc;
}

extern "C" int b2g_sumwt(const void *wgt, int dtype, int space, int64_t n, double *out, void *stream)
{
    g_err.clear();
    if (!out) return fail("out is NULL");
    B2G_OK(check_codes(dtype, space, "weight"));
    *out = 0.0;
    if (n <= 0) return 0;
    if (!wgt) return fail("wgt is NULL");
    cudaStream_t s = (cudaStream_t)stream;
    constexpr int kParts = 1024;
    double *scratch = nullptr;
    CUDA_OK(cudaMallocAsync(&scratch, sizeof(double) * (kParts + 1), s));
    int rc = 0;
    if (dtype == B2G_F64) {
        Staged<double> w;
        rc = stage_in<double>(wgt, dtype, space, n, s, &w);
        if (!rc) rc = launch_sum<double>(w.ptr, n, scratch + kParts, scratch, kParts, s) ? fail("sum kernel launch failed") : 0;
        if (!rc && cudaMemcpyAsync(out, scratch + kParts, sizeof(double), cudaMemcpyDeviceToHost, s) != cudaSuccess)
            rc = fail("sumwt copy failed");
        if (!rc && cudaStreamSynchronize(s) != cudaSuccess) rc = fail("sumwt sync failed");
    } else {
        Staged<float> w;
        rc = stage_in<float>(wgt, dtype, space, n, s, &w);
        if (!rc) rc = launch_sum<float>(w.ptr, n, scratch + kParts, scratch, kParts, s) ? fail("sum kernel launch failed") : 0;
        if (!rc && cudaMemcpyAsync(out, scratch + kParts, sizeof(double), cudaMemcpyDeviceToHost, s) != cudaSuccess)
            rc = fail("sumwt copy failed");
        if (!rc && cudaStreamSynchronize(s) != cudaSuccess) rc = fail("sumwt sync failed");
    }
    cudaFreeAsync(scratch, s);
    return rc;
}

extern "C" int b2g_set_host_chunk_samples(int64_t nsamples)
{
    g_err.clear();
    g_host_chunk_samples.store(nsamples > 0 ? (long long)nsamples : 0ll);
    return 0;
}

extern "C" int b2g_set_max_slab_planes(int nplanes)
{
    g_err.clear();
    g_max_slab_planes.store(nplanes > 0 ? nplanes : 0);
    return 0;
}

extern "C" int b2g_set_tc_mode(int mode)
{
    g_err.clear();
    if (mode < 0 || mode > 2) return fail("tensor-core mode must be 0 (off), 1 (dense chunks) or 2 (all chunks)");
    g_tc_mode.store(mode);
    return 0;
}

extern "C" int b2g_kernel_select(double epsilon, int do_wstacking, int precision, int64_t npix_x,
                                 int64_t npix_y, double nvis, double *out5)
{
    g_err.clear();
    KernelRow kr;
    B2G_OK(choose_kernel(epsilon, do_wstacking != 0, precision, npix_x, npix_y, 0, 0, nvis, 0.0, &kr));
    out5[0] = kr.W;
    out5[1] = kr.sigma;
    out5[2] = kr.beta;
    out5[3] = kr.e0;
    out5[4] = kr.eps;
    return 0;
}

extern "C" int b2g_kernel_eval(int W, double beta, double e0, const double *t, int n, double *out)
{
    g_err.clear();
    if (W < kMinW || W > kMaxW) return fail("W=%d out of range", W);
    PolyKernel pk;
    pk.build(W, beta, e0);
    for (int i = 0; i < n; ++i) out[i] = pk.eval(t[i]);
    return 0;
}

extern "C" int b2g_kernel_corr(int W, double beta, double e0, const double *x, int n, double *out)
{
    g_err.clear();
    if (W < kMinW || W > kMaxW) return fail("W=%d out of range", W);
    PolyKernel pk;
    pk.build(W, beta, e0);
    for (int i = 0; i < n; ++i) out[i] = pk.corr(x[i]);
    return 0;
}
