// clean_kernels.cu -- Hogbom CLEAN minor cycle on a device-resident residual (SURVEY.md section
// 8f row N4: the consumer of invert_ng's dirty image and PSF).
//
// Replaces the interpreted loop of hogbom()
// (/root/reference/src/ska_sdp_func_python/image/cleaners.py:23-133, overlapIndices :235-267):
//   repeat niter times
//     (mx, my) = argmax |res * window|            (first occurrence in C order, like numpy.argmax)
//     mval     = res[mx, my] * gain / pmax
//     comps[mx, my] += mval
//     res[overlap] -= psf[overlap] * mval         (PSF centred on the peak, clipped to the image)
//     stop if |res[mx, my]| < 0.9 * absolutethresh
//
// Every arithmetic step is the reference's own fp64 operation in the reference's order (separate
// multiply and subtract, no fused multiply-add), so components and residual are BIT-EXACT.
//
// Design: the residual, PSF and window of a 2048^2 image (3 x 33 MB) live in the 126 MB L2, so one
// iteration is one pass over L2-resident data: the kernel that subtracts the shifted PSF also
// reduces |res * window| of the UPDATED residual to one (value, index) candidate per block, and a
// one-block kernel picks the next peak, updates the component and evaluates the stop test.  The
// host replays batches of 64 iterations from a CUDA graph (iterate_until_stop) and looks at the
// stop flag after each batch; kernels of iterations enqueued after the stop exit immediately.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200_wgridder.h"

namespace b2g {
int wrapper_fail(const char *msg);  // plan.cu
}

namespace {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kBlocks = 1184;  // 148 SMs x 8
constexpr int kThreads = 256;

struct Cand {
    double val;
    long long idx;
};

__device__ __forceinline__ double neg_inf() { return __longlong_as_double(0xfff0000000000000ll); }

// numpy.argmax order: larger value wins, equal values keep the smaller flat index
__device__ __forceinline__ bool better(double v, long long i, double bv, long long bi)
{
    return v > bv || (v == bv && i < bi);
}

__device__ __forceinline__ Cand warp_best(Cand c)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        const double ov = __shfl_xor_sync(kFull, c.val, d);
        const long long oi = __shfl_xor_sync(kFull, c.idx, d);
        if (better(ov, oi, c.val, c.idx)) {
            c.val = ov;
            c.idx = oi;
        }
    }
    return c;
}

__device__ Cand block_best(Cand c)
{
    __shared__ double sv[32];
    __shared__ long long si[32];
    c = warp_best(c);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
        sv[warp] = c.val;
        si[warp] = c.idx;
    }
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    Cand r{lane < nw ? sv[lane] : neg_inf(), lane < nw ? si[lane] : (long long)0x7fffffffffffffffll};
    r = warp_best(r);
    __syncthreads();
    return r;
}

// The minor cycles are launch bound on small images (two short kernels per iteration), so the
// iterations are replayed from a CUDA graph of kBatch iterations; the state block tells the host
// after every batch whether the stop test fired (iterations enqueued after it are no-ops, and the
// kernels themselves enforce niter).  The legacy default stream cannot be captured: the whole call
// then runs on a private stream ordered after the caller's work (LoopStream).
constexpr int kBatch = 64;

struct LoopStream {
    cudaStream_t s;
    bool own = false;
    explicit LoopStream(cudaStream_t user) : s(user)
    {
        if (user != nullptr && user != cudaStreamLegacy) return;
        cudaStream_t ls = nullptr;
        cudaEvent_t ev = nullptr;
        if (cudaStreamCreateWithFlags(&ls, cudaStreamNonBlocking) != cudaSuccess) return;
        if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) {
            cudaStreamDestroy(ls);
            return;
        }
        cudaEventRecord(ev, user);
        cudaStreamWaitEvent(ls, ev, 0);
        cudaEventDestroy(ev);
        s = ls;
        own = true;
    }
    ~LoopStream()
    {
        if (own) {
            cudaStreamSynchronize(s);
            cudaStreamDestroy(s);
        }
    }
    LoopStream(const LoopStream &) = delete;
    LoopStream &operator=(const LoopStream &) = delete;
};

// one(): enqueue one iteration on s.  Returns 0 or the wrapper's error code; *h holds the final state.
template <typename State, typename F>
int iterate_until_stop(cudaStream_t s, long long niter, F one, const State *d_st, State *h)
{
    cudaGraphExec_t exec = nullptr;
    if (niter >= 2 * kBatch && cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
        for (int k = 0; k < kBatch; ++k) one();
        cudaGraph_t g = nullptr;
        if (cudaStreamEndCapture(s, &g) == cudaSuccess && g != nullptr &&
            cudaGraphInstantiate(&exec, g, 0) != cudaSuccess)
            exec = nullptr;
        if (g) cudaGraphDestroy(g);
    }
    (void)cudaGetLastError();  // not capturable: plain launches below
    int rc = 0;
    h->stop = 0;
    // short runs: no more launches per batch than iterations asked for (+ 1: the multi-scale loops
    // need the select after the last subtraction to see the limit)
    const int per_batch = exec ? kBatch : (int)(niter + 1 < kBatch ? niter + 1 : kBatch);
    for (long long batch = 0; batch <= niter / per_batch + 1 && !h->stop; ++batch) {
        cudaError_t e;
        if (exec) {
            e = cudaGraphLaunch(exec, s);
        } else {
            for (int k = 0; k < per_batch; ++k) one();
            e = cudaGetLastError();
        }
        if (e == cudaSuccess) e = cudaMemcpyAsync(h, d_st, sizeof(State), cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) {
            rc = b2g::wrapper_fail(cudaGetErrorString(e));
            break;
        }
    }
    if (exec) cudaGraphExecDestroy(exec);
    return rc;
}

struct CleanState {
    unsigned int ticket;  // blocks of the running iteration that have delivered their candidate
    int mx, my;        // current peak
    double mval;       // component just added at the peak
    int stop;          // set once the stop test fires (or niter is reached)
    int iters;         // iterations completed
    double peak0;      // max |dirty| (no window), for absolutethresh
};

// mode 0: candidate = |a[i] * w[i]| (w may be NULL), mode 1: candidate = a[i]
__global__ void __launch_bounds__(kThreads)
peak_partial_kernel(const double *__restrict__ a, const double *__restrict__ w, long long n, int mode,
                    Cand *__restrict__ part)
{
    Cand best{neg_inf(), 0x7fffffffffffffffll};  // -inf
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const double v = mode == 1 ? a[i] : fabs(w ? __dmul_rn(a[i], w[i]) : a[i]);
        if (better(v, i, best.val, best.idx)) {
            best.val = v;
            best.idx = i;
        }
    }
    best = block_best(best);
    if (threadIdx.x == 0) part[blockIdx.x] = best;
}

__global__ void __launch_bounds__(kThreads)
peak_final_kernel(const Cand *__restrict__ part, int npart, Cand *__restrict__ out)
{
    Cand best{neg_inf(), 0x7fffffffffffffffll};
    for (int i = threadIdx.x; i < npart; i += blockDim.x)
        if (better(part[i].val, part[i].idx, best.val, best.idx)) best = part[i];
    best = block_best(best);
    if (threadIdx.x == 0) *out = best;
}

// Coherent (L2) loads of values written by other blocks of the same launch.
__device__ __forceinline__ Cand load_cand(const Cand *p)
{
    return Cand{__ldcg(&p->val), __ldcg(&p->idx)};
}

// Executed by ONE whole block: stop test of the iteration just subtracted (first == 0), then the
// next peak, the component value and the component image update.
__device__ void clean_select_block(const double *res, double *__restrict__ comps, int ny, const Cand *part,
                                   int npart, double gain, double pmax, double stop_level, int niter,
                                   int first, CleanState *st)
{
    __shared__ int s_stop;
    if (threadIdx.x == 0) {
        s_stop = 0;
        if (!first) {
            st->iters += 1;
            const double r = __ldcg(&res[(long long)st->mx * ny + st->my]);
            if (fabs(r) < stop_level || st->iters >= niter) {
                st->stop = 1;
                s_stop = 1;
            }
        }
    }
    __syncthreads();
    if (s_stop) return;
    Cand best{neg_inf(), 0x7fffffffffffffffll};
    for (int i = threadIdx.x; i < npart; i += blockDim.x) {
        const Cand c = load_cand(part + i);
        if (better(c.val, c.idx, best.val, best.idx)) best = c;
    }
    best = block_best(best);
    if (threadIdx.x == 0) {
        const int mx = (int)(best.idx / ny), my = (int)(best.idx - (long long)mx * ny);
        const double mval = __ddiv_rn(__dmul_rn(__ldcg(&res[best.idx]), gain), pmax);
        st->mx = mx;
        st->my = my;
        st->mval = mval;
        comps[best.idx] = __dadd_rn(comps[best.idx], mval);
    }
}

// The first peak (no subtraction yet).
__global__ void __launch_bounds__(kThreads)
clean_select_kernel(const double *res, double *__restrict__ comps, int ny, const Cand *part, int npart,
                    double gain, double pmax, double stop_level, int niter, int first, CleanState *st)
{
    if (st->stop) return;
    clean_select_block(res, comps, ny, part, npart, gain, pmax, stop_level, niter, first, st);
}

// One iteration in one launch: res -= psf(shifted to the peak) * mval on the overlap, candidates of
// the updated residual per block, and -- in the block that finishes last (ticket counter behind a
// __threadfence) -- the stop test, the next peak and the component update.  Every block reads the
// state at its start; the last block rewrites it only after all others have taken their ticket.
// The pointers stay __restrict__ for the streaming loop (without it the unrolled loads are not
// batched: measured 2 x slower); the last block reads other blocks' results with explicit L2 loads
// (__ldcg) behind the fences.
__global__ void __launch_bounds__(kThreads)
clean_iter_kernel(double *__restrict__ res, const double *__restrict__ psf,
                  const double *__restrict__ window, double *__restrict__ comps, int nx, int ny, int px,
                  int py, double gain, double pmax, double stop_level, int niter,
                  CleanState *__restrict__ st, Cand *__restrict__ part)
{
    if (st->stop) return;
    const int mx = st->mx, my = st->my;
    const double mval = st->mval;
    const int pwx = px / 2, pwy = py / 2;
    const int xlo = max(0, mx - pwx), xhi = min(nx, mx + pwx);
    const int ylo = max(0, my - pwy), yhi = min(ny, my + pwy);
    Cand best{neg_inf(), 0x7fffffffffffffffll};
    // rows over blocks, columns over threads: no index division in the loop
    for (int x = blockIdx.x; x < nx; x += gridDim.x) {
        const bool xin = x >= xlo && x < xhi;
        const double *prow = psf + (long long)(x - mx + pwx) * py + (pwy - my);
#pragma unroll 4
        for (int y = threadIdx.x; y < ny; y += blockDim.x) {
            const long long i = (long long)x * ny + y;
            double r = res[i];
            if (xin && y >= ylo && y < yhi) {
                r = __dadd_rn(r, -__dmul_rn(prow[y], mval));
                res[i] = r;
            }
            const double v = fabs(window ? __dmul_rn(r, window[i]) : r);
            if (better(v, i, best.val, best.idx)) {
                best.val = v;
                best.idx = i;
            }
        }
    }
    best = block_best(best);
    __shared__ int s_last;
    if (threadIdx.x == 0) {
        part[blockIdx.x] = best;
        __threadfence();
        s_last = atomicAdd(&st->ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (threadIdx.x == 0) st->ticket = 0;
    clean_select_block(res, comps, ny, part, (int)gridDim.x, gain, pmax, stop_level, niter, 0, st);
}

// ---- complex Hogbom CLEAN of Q + iU (hogbom_complex(), cleaners.py:136-232) --------------------
//
// Same loop as hogbom() on the complex image res = Q + iU, kept as two real planes.  numpy's
// arithmetic is mirrored operation by operation: |z| is hypot(re, im); complex x real products are
// component-wise; mval = res * gain / pmax divides a complex by a real through Smith's method,
// i.e. multiplies both parts by 1.0 / pmax; the stop test is |res[peak]| < absolutethresh (no 0.9).

struct CplxCleanState {
    int mx, my;
    double mre, mim;
    int stop;
    int iters;
};

__device__ __forceinline__ double cabs_win(double re, double im, const double *window, long long i)
{
    if (window) {
        const double w = window[i];
        re = __dmul_rn(re, w);
        im = __dmul_rn(im, w);
    }
    return hypot(re, im);
}

// first != 0: candidates of the initial residual only
__global__ void __launch_bounds__(kThreads)
cclean_subtract_kernel(double *__restrict__ res_re, double *__restrict__ res_im,
                       const double *__restrict__ psf, const double *__restrict__ window, int nx,
                       int ny, int px, int py, int first, const CplxCleanState *__restrict__ st,
                       Cand *__restrict__ part)
{
    if (st->stop) return;
    const int mx = st->mx, my = st->my;
    const double mre = st->mre, mim = st->mim;
    const int pwx = px / 2, pwy = py / 2;
    const int xlo = first ? 0 : max(0, mx - pwx), xhi = first ? 0 : min(nx, mx + pwx);
    const int ylo = max(0, my - pwy), yhi = min(ny, my + pwy);
    Cand best{neg_inf(), 0x7fffffffffffffffll};
    for (int x = blockIdx.x; x < nx; x += gridDim.x) {
        const bool xin = x >= xlo && x < xhi;
        const double *prow = psf + (long long)(x - mx + pwx) * py + (pwy - my);
        for (int y = threadIdx.x; y < ny; y += blockDim.x) {
            const long long i = (long long)x * ny + y;
            double re = res_re[i], im = res_im[i];
            if (xin && y >= ylo && y < yhi) {
                const double p = prow[y];
                re = __dadd_rn(re, -__dmul_rn(p, mre));
                im = __dadd_rn(im, -__dmul_rn(p, mim));
                res_re[i] = re;
                res_im[i] = im;
            }
            const double v = cabs_win(re, im, window, i);
            if (better(v, i, best.val, best.idx)) {
                best.val = v;
                best.idx = i;
            }
        }
    }
    best = block_best(best);
    if (threadIdx.x == 0) part[blockIdx.x] = best;
}

__global__ void __launch_bounds__(kThreads)
cclean_select_kernel(const double *__restrict__ res_re, const double *__restrict__ res_im,
                     double *__restrict__ comps_re, double *__restrict__ comps_im, int ny,
                     const Cand *__restrict__ part, int npart, double gain, double pmax,
                     double absolutethresh, int niter, int first, CplxCleanState *__restrict__ st)
{
    if (st->stop) return;
    __shared__ int s_stop;
    if (threadIdx.x == 0) {
        s_stop = 0;
        if (!first) {
            st->iters += 1;
            const long long i = (long long)st->mx * ny + st->my;
            if (hypot(res_re[i], res_im[i]) < absolutethresh || st->iters >= niter) {
                st->stop = 1;
                s_stop = 1;
            }
        }
    }
    __syncthreads();
    if (s_stop) return;
    Cand best{neg_inf(), 0x7fffffffffffffffll};
    for (int i = threadIdx.x; i < npart; i += blockDim.x)
        if (better(part[i].val, part[i].idx, best.val, best.idx)) best = part[i];
    best = block_best(best);
    if (threadIdx.x == 0) {
        const int mx = (int)(best.idx / ny), my = (int)(best.idx - (long long)mx * ny);
        const double scl = __ddiv_rn(1.0, pmax);
        const double mre = __dmul_rn(__dmul_rn(res_re[best.idx], gain), scl);
        const double mim = __dmul_rn(__dmul_rn(res_im[best.idx], gain), scl);
        st->mx = mx;
        st->my = my;
        st->mre = mre;
        st->mim = mim;
        comps_re[best.idx] = __dadd_rn(comps_re[best.idx], mre);
        comps_im[best.idx] = __dadd_rn(comps_im[best.idx], mim);
    }
}

// ---- multi-scale CLEAN (msclean(), cleaners.py:279-470; find_max_abs_stack :565-610) -----------
//
// State on the device: the stack of scale-convolved residuals res[ns][nx][ny], the doubly
// convolved PSFs pss[ns][ns][px][py], the scale shapes pscale[ns][px][py], an optional window
// stack and sensitivity image, and the coupling diagonal c[s] = max pss[s][s].  One iteration =
// one launch that subtracts gain * mval * pss[s][mscale] (shifted to the peak) from every scale,
// adds the scale shape to the component image and reduces the search metric of the UPDATED stack
// to one candidate per (block, scale), plus a one-block launch that applies the reference's
// across-scale rule and stop tests.  Arithmetic and its order are the reference's.

constexpr int kMsBlocks = 592;  // upper bound of blocks per scale plane (see plane_blocks)
constexpr int kMaxScales = 16;

struct MsState {
    int mx, my, mscale;
    double mval;
    int stop;
    int iters;  // subtractions done
};

struct MsCoupling {
    double c[kMaxScales];
};

// value the search compares across scales (sensitivity applied once) ...
__device__ __forceinline__ double ms_scaled(double r, const double *wrow, const double *sens,
                                            long long i, double c)
{
    double v = wrow ? __dmul_rn(r, wrow[i]) : r;
    v = __ddiv_rn(v, c);
    if (sens) v = __dmul_rn(v, sens[i]);
    return v;
}
// ... and the metric whose argmax picks the pixel within one scale (sensitivity applied twice
// when there is one, cleaners.py:591-596)
__device__ __forceinline__ double ms_metric(double v, const double *sens, long long i)
{
    return fabs(sens ? __dmul_rn(v, sens[i]) : v);
}

// Running per-thread argmax of the metric over pixels visited in ascending index order.  Without
// a sensitivity image and with c > 0 the metric |r w / c| is a non-decreasing function of
// |r w|, so a pixel whose |r w| is strictly below the one behind the current best cannot beat it
// (an equal quotient would lose the tie to the earlier index): the division is only done for new
// running maxima.
struct MsBest {
    Cand best;
    double pre;  // |r w| behind best (fast path only)
};

__device__ __forceinline__ void ms_visit(MsBest &b, double r, const double *wrow, const double *sens,
                                         long long i, double c, bool fast)
{
    if (fast) {
        const double a = fabs(wrow ? __dmul_rn(r, wrow[i]) : r);
        if (a < b.pre) return;
        const double m = fabs(__ddiv_rn(wrow ? __dmul_rn(r, wrow[i]) : r, c));
        if (better(m, i, b.best.val, b.best.idx)) {
            b.best.val = m;
            b.best.idx = i;
            b.pre = a;
        }
    } else {
        const double m = ms_metric(ms_scaled(r, wrow, sens, i, c), sens, i);
        if (better(m, i, b.best.val, b.best.idx)) {
            b.best.val = m;
            b.best.idx = i;
        }
    }
}

// grid (plane_blocks, ns + 1): y < ns updates scale y and reduces its metric; y == ns adds the
// component.  first != 0: no subtraction, candidates of the initial stack only.
__global__ void __launch_bounds__(kThreads)
ms_subtract_kernel(double *__restrict__ res, const double *__restrict__ pss,
                   const double *__restrict__ pscale, const double *__restrict__ wstack,
                   const double *__restrict__ sens, double *__restrict__ comps, MsCoupling cp, int ns,
                   int nx, int ny, int px, int py, double gain, int first,
                   const MsState *__restrict__ st, Cand *__restrict__ part)
{
    if (st->stop) return;
    const int s = blockIdx.y;
    const int mx = st->mx, my = st->my, ms = st->mscale;
    const double mval = st->mval;
    const int pwx = px / 2, pwy = py / 2;
    const int xlo = first ? 0 : max(0, mx - pwx), xhi = first ? 0 : min(nx, mx + pwx);
    const int ylo = max(0, my - pwy), yhi = min(ny, my + pwy);
    const long long n = (long long)nx * ny, pn = (long long)px * py;
    if (s == ns) {
        if (first) return;
        const double *shape = pscale + (long long)ms * pn;
        for (int x = xlo + blockIdx.x; x < xhi; x += gridDim.x) {
            const double *prow = shape + (long long)(x - mx + pwx) * py + (pwy - my);
            for (int y = ylo + threadIdx.x; y < yhi; y += blockDim.x) {
                const long long i = (long long)x * ny + y;
                comps[i] = __dadd_rn(comps[i], __dmul_rn(__dmul_rn(prow[y], gain), mval));
            }
        }
        return;
    }
    double *r_s = res + (long long)s * n;
    const double *w_s = wstack ? wstack + (long long)s * n : nullptr;
    const double *p_s = pss + ((long long)s * ns + ms) * pn;
    const double c = cp.c[s];
    const bool fast = sens == nullptr && c > 0.0;
    MsBest b{{neg_inf(), 0x7fffffffffffffffll}, -1.0};
    for (int x = blockIdx.x; x < nx; x += gridDim.x) {
        const bool xin = x >= xlo && x < xhi;
        const double *prow = p_s + (long long)(x - mx + pwx) * py + (pwy - my);
        for (int y = threadIdx.x; y < ny; y += blockDim.x) {
            const long long i = (long long)x * ny + y;
            double r = r_s[i];
            if (xin && y >= ylo && y < yhi) {
                r = __dadd_rn(r, -__dmul_rn(__dmul_rn(prow[y], gain), mval));
                r_s[i] = r;
            }
            ms_visit(b, r, w_s, sens, i, c, fast);
        }
    }
    const Cand best = block_best(b.best);
    if (threadIdx.x == 0) part[(long long)s * gridDim.x + blockIdx.x] = best;
}

// One block: the peak over all scales (strict '>' in scale order, starting from 0 at pixel
// (0, 0) of scale 0), the component value and the stop tests of the iteration about to start.
__global__ void __launch_bounds__(kThreads)
ms_select_kernel(const double *__restrict__ res, const double *__restrict__ wstack,
                 const double *__restrict__ sens, MsCoupling cp, int ns, int nx, int ny,
                 const Cand *__restrict__ part, int npart, double stop_level, int niter,
                 MsState *__restrict__ st)
{
    if (st->stop) return;
    const long long n = (long long)nx * ny;
    double pabsmax = 0.0;
    long long pidx = 0;
    int pscale = 0;
    for (int s = 0; s < ns; ++s) {
        Cand best{neg_inf(), 0x7fffffffffffffffll};
        for (int i = threadIdx.x; i < npart; i += blockDim.x) {
            const Cand c = part[(long long)s * npart + i];
            if (better(c.val, c.idx, best.val, best.idx)) best = c;
        }
        best = block_best(best);
        // every thread of warp 0 holds the block's best; thread 0 decides
        if (threadIdx.x == 0) {
            const double v = ms_scaled(res[(long long)s * n + best.idx], wstack ? wstack + (long long)s * n : nullptr,
                                       sens, best.idx, cp.c[s]);
            if (fabs(v) > pabsmax) {
                pabsmax = fabs(v);
                pidx = best.idx;
                pscale = s;
            }
        }
    }
    if (threadIdx.x == 0) {
        const double r = res[(long long)pscale * n + pidx];
        const double mval = __ddiv_rn(r, cp.c[pscale]);
        if (st->iters >= niter || fabs(r) < stop_level || !(fabs(mval) > 0.0)) {
            st->stop = 1;
        } else {
            st->mx = (int)(pidx / ny);
            st->my = (int)(pidx - (long long)st->mx * ny);
            st->mscale = pscale;
            st->mval = mval;
            st->iters += 1;
        }
    }
}

// ---- multi-scale multi-frequency CLEAN (msmfsclean(), cleaners.py:686-898) ---------------------
//
// State: scale-convolved moment residuals smres[ns][nm][nx][ny]; coupled PSFs indexed by the SUM of
// the two moment indices, psfm[ns (peak scale)][ns][2 nm - 1][px][py] (the reference stores all
// nm x nm pairs, :1055-1078, which only depend on t + q); scale shapes pscale[ns][px][py]; inverse
// moment Hessians ih[ns][nm][nm].  One iteration = one launch that, for every scale, subtracts
// gain * sum_q psfm[mscale][s][t + q] * mval[q] from moment t on the clipped overlap, forms the
// moment-0 principal solution sum_m ih[s][m][0] * smres[s][m] of the UPDATED residuals and reduces
// (a) max |solution * window * sensitivity| and (b) argmax |solution| per block -- the reference
// picks the scale on (a) and the pixel on (b), :1124-1157 -- plus one launch that applies the
// across-scale rule, evaluates all moments of the principal solution at the peak and the stop test.

constexpr int kMaxMoments = 4;

struct MfState {
    int mx, my, mscale;
    int stop;
    int iters;
    double mval[kMaxMoments];
};

constexpr int kMaxMfScales = 8;

struct MfHess {
    double ih[kMaxMfScales][kMaxMoments][kMaxMoments];  // inverse moment Hessians
    double h[kMaxMfScales][kMaxMoments][kMaxMoments];   // the Hessians themselves (findpeak "CASA")
};

template <int NM, bool CASA>
__global__ void __launch_bounds__(kThreads)
mf_subtract_kernel(double *__restrict__ smres, const double *__restrict__ psfm,
                   const double *__restrict__ pscale, const double *__restrict__ wstack,
                   const double *__restrict__ sens, double *__restrict__ model, MfHess hs, int ns,
                   int nx, int ny, int px, int py, double gain, int first,
                   const MfState *__restrict__ st, Cand *__restrict__ part_scale,
                   Cand *__restrict__ part_pixel)
{
    if (st->stop) return;
    constexpr int NK = 2 * NM - 1;
    const int s = blockIdx.y;
    const int mx = st->mx, my = st->my, ms = st->mscale;
    double mval[NM];
#pragma unroll
    for (int t = 0; t < NM; ++t) mval[t] = st->mval[t];
    const int pwx = px / 2, pwy = py / 2;
    const int xlo = first ? 0 : max(0, mx - pwx), xhi = first ? 0 : min(nx, mx + pwx);
    const int ylo = max(0, my - pwy), yhi = min(ny, my + pwy);
    const long long n = (long long)nx * ny, pn = (long long)px * py;
    if (s == ns) {
        if (first) return;
        const double *shape = pscale + (long long)ms * pn;
        for (int x = xlo + blockIdx.x; x < xhi; x += gridDim.x) {
            const double *prow = shape + (long long)(x - mx + pwx) * py + (pwy - my);
            for (int y = ylo + threadIdx.x; y < yhi; y += blockDim.x) {
                const long long i = (long long)x * ny + y;
                const double pg = __dmul_rn(prow[y], gain);
#pragma unroll
                for (int t = 0; t < NM; ++t)
                    model[t * n + i] = __dadd_rn(model[t * n + i], __dmul_rn(pg, mval[t]));
            }
        }
        return;
    }
    double *r_s = smres + (long long)s * NM * n;
    const double *w_s = wstack ? wstack + (long long)s * n : nullptr;
    const double *p_s = psfm + ((long long)ms * ns + s) * NK * pn;
    double ih0[NM];
#pragma unroll
    for (int m = 0; m < NM; ++m) ih0[m] = hs.ih[s][m][0];
    Cand best_scale{neg_inf(), 0x7fffffffffffffffll}, best_pixel{neg_inf(), 0x7fffffffffffffffll};
    for (int x = blockIdx.x; x < nx; x += gridDim.x) {
        const bool xin = x >= xlo && x < xhi;
        const long long poff = (long long)(x - mx + pwx) * py + (pwy - my);
#pragma unroll 4
        for (int y = threadIdx.x; y < ny; y += blockDim.x) {
            const long long i = (long long)x * ny + y;
            double r[NM];
#pragma unroll
            for (int t = 0; t < NM; ++t) r[t] = r_s[t * n + i];
            if (xin && y >= ylo && y < yhi) {
                double pk[NK];
#pragma unroll
                for (int k = 0; k < NK; ++k) pk[k] = p_s[k * pn + poff + y];
#pragma unroll
                for (int t = 0; t < NM; ++t) {
                    double acc = __dmul_rn(pk[t], mval[0]);
#pragma unroll
                    for (int q = 1; q < NM; ++q) acc = __dadd_rn(acc, __dmul_rn(pk[t + q], mval[q]));
                    r[t] = __dadd_rn(r[t], -__dmul_rn(gain, acc));
                    r_s[t * n + i] = r[t];
                }
            }
            double sol = __dmul_rn(ih0[0], r[0]);
#pragma unroll
            for (int m = 1; m < NM; ++m) sol = __dadd_rn(sol, __dmul_rn(ih0[m], r[m]));
            if (CASA) {
                // CASA 4.7 peak finder (cleaners.py:936-960): the search image is the decrease of
                // chi^2, sum_m 2 sol[m] res[m] - sum_{m,q} H[m][q] sol[m] sol[q], accumulated in the
                // reference's order
                double sl[NM];
                sl[0] = sol;
#pragma unroll
                for (int nn = 1; nn < NM; ++nn) {
                    double acc = __dmul_rn(hs.ih[s][0][nn], r[0]);
#pragma unroll
                    for (int m = 1; m < NM; ++m) acc = __dadd_rn(acc, __dmul_rn(hs.ih[s][m][nn], r[m]));
                    sl[nn] = acc;
                }
                double dch = 0.0;
#pragma unroll
                for (int m1 = 0; m1 < NM; ++m1) {
                    dch = __dadd_rn(dch, __dmul_rn(__dmul_rn(2.0, sl[m1]), r[m1]));
#pragma unroll
                    for (int m2 = 0; m2 < NM; ++m2)
                        dch = __dadd_rn(dch, -__dmul_rn(__dmul_rn(hs.h[s][m1][m2], sl[m1]), sl[m2]));
                }
                sol = dch;
            }
            double v = w_s ? __dmul_rn(sol, w_s[i]) : sol;
            if (sens) v = __dmul_rn(sens[i], v);
            v = fabs(v);
            if (v > best_scale.val) best_scale.val = v;
            const double a = fabs(sol);
            if (better(a, i, best_pixel.val, best_pixel.idx)) {
                best_pixel.val = a;
                best_pixel.idx = i;
            }
        }
    }
    best_scale.idx = 0;
    best_scale = block_best(best_scale);
    best_pixel = block_best(best_pixel);
    if (threadIdx.x == 0) {
        part_scale[(long long)s * gridDim.x + blockIdx.x] = best_scale;
        part_pixel[(long long)s * gridDim.x + blockIdx.x] = best_pixel;
    }
}

template <int NM>
__global__ void __launch_bounds__(kThreads)
mf_select_kernel(const double *__restrict__ smres, MfHess hs, int ns, int nx, int ny,
                 const Cand *__restrict__ part_scale, const Cand *__restrict__ part_pixel, int npart,
                 double absolutethresh, int niter, MfState *__restrict__ st)
{
    if (st->stop) return;
    const long long n = (long long)nx * ny;
    double optimum = 0.0;
    long long pidx = 0;
    int pscale = 0;
    for (int s = 0; s < ns; ++s) {
        Cand top{neg_inf(), 0}, pix{neg_inf(), 0x7fffffffffffffffll};
        for (int i = threadIdx.x; i < npart; i += blockDim.x) {
            const Cand a = part_scale[(long long)s * npart + i];
            if (a.val > top.val) top.val = a.val;
            const Cand b = part_pixel[(long long)s * npart + i];
            if (better(b.val, b.idx, pix.val, pix.idx)) pix = b;
        }
        top = block_best(top);
        pix = block_best(pix);
        if (top.val > optimum) {
            optimum = top.val;
            pscale = s;
            pidx = pix.idx;
        }
    }
    if (threadIdx.x == 0) {
        double mval[NM];
#pragma unroll
        for (int nn = 0; nn < NM; ++nn) {
            double acc = __dmul_rn(hs.ih[pscale][0][nn], smres[((long long)pscale * NM) * n + pidx]);
#pragma unroll
            for (int m = 1; m < NM; ++m)
                acc = __dadd_rn(acc, __dmul_rn(hs.ih[pscale][m][nn], smres[((long long)pscale * NM + m) * n + pidx]));
            mval[nn] = acc;
        }
        if (st->iters >= niter || fabs(mval[0]) < absolutethresh) {
            st->stop = 1;
        } else {
            st->mx = (int)(pidx / ny);
            st->my = (int)(pidx - (long long)st->mx * ny);
            st->mscale = pscale;
#pragma unroll
            for (int nn = 0; nn < NM; ++nn) st->mval[nn] = mval[nn];
            st->iters += 1;
        }
    }
}

template <int NM>
int mf_run(double *smres, const double *psfm, const double *pscale, const double *wstack,
           const double *sens, const MfHess &hs, int ns, int nx, int ny, int px, int py, double gain,
           double absolutethresh, long long niter, int casa, double *model, int64_t *iters_host,
           cudaStream_t s);

int cuda_fail(cudaError_t e) { return b2g::wrapper_fail(cudaGetErrorString(e)); }

// Blocks per scale plane so that the whole (planes + 1)-deep grid is exactly one resident wave:
// more blocks than fit leave a partial second wave (measured: 1480 blocks at 6 resident per SM ran
// 1.67 waves).
template <typename K> int plane_blocks(K kernel, int planes)
{
    int occ = 0, dev = 0, nsm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kThreads, 0) != cudaSuccess || occ < 1) occ = 4;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || nsm < 1) nsm = 148;
    const int bx = occ * nsm / planes;
    return bx < 1 ? 1 : (bx > kMsBlocks ? kMsBlocks : bx);
}


template <int NM>
int mf_run(double *smres, const double *psfm, const double *pscale, const double *wstack,
           const double *sens, const MfHess &hs, int ns, int nx, int ny, int px, int py, double gain,
           double absolutethresh, long long niter, int casa, double *model, int64_t *iters_host,
           cudaStream_t s)
{
    Cand *d_part = nullptr;
    const int bx = casa ? plane_blocks(mf_subtract_kernel<NM, true>, ns + 1)
                        : plane_blocks(mf_subtract_kernel<NM, false>, ns + 1);
    const size_t npart = (size_t)bx * ns;
    cudaError_t e = cudaMallocAsync(&d_part, sizeof(Cand) * 2 * npart + sizeof(MfState), s);
    if (e != cudaSuccess) return cuda_fail(e);
    Cand *d_scale = d_part, *d_pixel = d_part + npart;
    MfState *d_st = reinterpret_cast<MfState *>(d_part + 2 * npart);
    auto done = [&](int code) {
        cudaStreamSynchronize(s);
        cudaFreeAsync(d_part, s);
        return code;
    };
#define CK(call)                                   \
    do {                                           \
        cudaError_t e_ = (call);                   \
        if (e_ != cudaSuccess) return done(cuda_fail(e_)); \
    } while (0)
    CK(cudaMemsetAsync(d_st, 0, sizeof(MfState), s));
    CK(cudaMemsetAsync(model, 0, sizeof(double) * NM * nx * ny, s));
    const dim3 grid(bx, ns + 1);
    auto subtract = [&](int first) {
        if (casa)
            mf_subtract_kernel<NM, true><<<grid, kThreads, 0, s>>>(smres, psfm, pscale, wstack, sens, model, hs,
                                                                   ns, nx, ny, px, py, gain, first, d_st,
                                                                   d_scale, d_pixel);
        else
            mf_subtract_kernel<NM, false><<<grid, kThreads, 0, s>>>(smres, psfm, pscale, wstack, sens, model,
                                                                    hs, ns, nx, ny, px, py, gain, first, d_st,
                                                                    d_scale, d_pixel);
    };
    subtract(1);
    MfState h;
    h.iters = 0;
    const int rc = iterate_until_stop(s, niter, [&] {
        mf_select_kernel<NM><<<1, kThreads, 0, s>>>(smres, hs, ns, nx, ny, d_scale, d_pixel, bx,
                                                    absolutethresh, (int)niter, d_st);
        subtract(0);
    }, d_st, &h);
    if (rc) return done(rc);
#undef CK
    if (iters_host) *iters_host = h.iters;
    return done(0);
}

}  // namespace

static int run_peak(const double *img, const double *window, long long n, int mode, Cand *d_part,
                    Cand *d_out, cudaStream_t s)
{
    peak_partial_kernel<<<kBlocks, kThreads, 0, s>>>(img, window, n, mode, d_part);
    peak_final_kernel<<<1, kThreads, 0, s>>>(d_part, kBlocks, d_out);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : cuda_fail(e);
}

extern "C" int b2g_image_peak(const double *img, const double *window, int64_t nx, int64_t ny,
                              int mode, double *val_host, int64_t *idx_host, void *stream)
{
    if (!img || nx < 1 || ny < 1) return b2g::wrapper_fail("b2g_image_peak: bad arguments");
    if (mode != 0 && mode != 1) return b2g::wrapper_fail("b2g_image_peak: mode must be 0 (argmax |a*w|) or 1 (max a)");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return b2g::wrapper_fail("b2g_image_peak: no CUDA device available (this library has no CPU path)");
    cudaStream_t s = (cudaStream_t)stream;
    Cand *d = nullptr;
    cudaError_t e = cudaMallocAsync(&d, sizeof(Cand) * (kBlocks + 1), s);
    if (e != cudaSuccess) return cuda_fail(e);
    int rc = run_peak(img, window, nx * ny, mode, d, d + kBlocks, s);
    Cand h{0.0, 0};
    if (rc == 0) {
        e = cudaMemcpyAsync(&h, d + kBlocks, sizeof(Cand), cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) rc = cuda_fail(e);
    }
    cudaFreeAsync(d, s);
    if (rc) return rc;
    if (val_host) *val_host = h.val;
    if (idx_host) *idx_host = h.idx;
    return 0;
}

extern "C" int b2g_hogbom(const double *dirty, const double *psf, const double *window, int64_t nx,
                          int64_t ny, int64_t px, int64_t py, double gain, double pmax,
                          double thresh, double fracthresh, int64_t niter, double *comps,
                          double *res, int64_t *iters_host, void *stream)
{
    if (!dirty || !psf || !comps || !res) return b2g::wrapper_fail("b2g_hogbom: NULL pointer");
    if (nx < 1 || ny < 1 || px < 1 || py < 1 || nx * ny > 0x7fffffffll)
        return b2g::wrapper_fail("b2g_hogbom: bad image shape");
    if (!(gain > 0.0 && gain < 2.0)) return b2g::wrapper_fail("b2g_hogbom: need 0 < gain < 2");
    if (niter < 1) return b2g::wrapper_fail("b2g_hogbom: need niter > 0");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return b2g::wrapper_fail("b2g_hogbom: no CUDA device available (this library has no CPU path)");
    LoopStream loop_stream((cudaStream_t)stream);
    cudaStream_t s = loop_stream.s;
    const long long n = nx * ny;
    Cand *d_part = nullptr;
    cudaError_t e = cudaMallocAsync(&d_part, sizeof(Cand) * (kBlocks + 1) + sizeof(CleanState), s);
    if (e != cudaSuccess) return cuda_fail(e);
    Cand *d_peak = d_part + kBlocks;
    CleanState *d_st = reinterpret_cast<CleanState *>(d_part + kBlocks + 1);
    int rc = 0;
    auto done = [&](int code) {
        cudaStreamSynchronize(s);
        cudaFreeAsync(d_part, s);
        return code;
    };
#define CK(call)                                   \
    do {                                           \
        cudaError_t e_ = (call);                   \
        if (e_ != cudaSuccess) return done(cuda_fail(e_)); \
    } while (0)
    // absolutethresh = max(thresh, fracthresh * max|dirty|)          cleaners.py:57
    if ((rc = run_peak(dirty, nullptr, n, 0, d_part, d_peak, s))) return done(rc);
    Cand h{0.0, 0};
    CK(cudaMemcpyAsync(&h, d_peak, sizeof(Cand), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    const double absolutethresh = thresh > fracthresh * h.val ? thresh : fracthresh * h.val;
    const double stop_level = 0.9 * absolutethresh;
    CK(cudaMemsetAsync(d_st, 0, sizeof(CleanState), s));
    CK(cudaMemsetAsync(comps, 0, sizeof(double) * n, s));
    if (res != dirty) CK(cudaMemcpyAsync(res, dirty, sizeof(double) * n, cudaMemcpyDeviceToDevice, s));
    // first peak
    peak_partial_kernel<<<kBlocks, kThreads, 0, s>>>(res, window, n, 0, d_part);
    clean_select_kernel<<<1, kThreads, 0, s>>>(res, comps, (int)ny, d_part, kBlocks, gain, pmax,
                                               stop_level, (int)niter, 1, d_st);
    CleanState hs;
    hs.iters = 0;
    rc = iterate_until_stop(s, niter, [&] {
        clean_iter_kernel<<<kBlocks, kThreads, 0, s>>>(res, psf, window, comps, (int)nx, (int)ny, (int)px,
                                                       (int)py, gain, pmax, stop_level, (int)niter, d_st,
                                                       d_part);
    }, d_st, &hs);
    if (rc) return done(rc);
#undef CK
    if (iters_host) *iters_host = hs.iters;
    return done(0);
}

extern "C" int b2g_msclean(double *res_stack, const double *psf_scalescale, const double *psf_scale,
                           const double *window_stack, const double *sensitivity,
                           const double *coupling_diag, int64_t nscales, int64_t nx, int64_t ny,
                           int64_t px, int64_t py, double gain, double absolutethresh, int64_t niter,
                           double *comps, int64_t *iters_host, void *stream)
{
    if (!res_stack || !psf_scalescale || !psf_scale || !coupling_diag || !comps)
        return b2g::wrapper_fail("b2g_msclean: NULL pointer");
    if (nscales < 1 || nscales > kMaxScales) return b2g::wrapper_fail("b2g_msclean: need 1..16 scales");
    if (nx < 1 || ny < 1 || px < 1 || py < 1 || nx * ny > 0x7fffffffll)
        return b2g::wrapper_fail("b2g_msclean: bad image shape");
    if (!(gain > 0.0 && gain < 2.0)) return b2g::wrapper_fail("b2g_msclean: need 0 < gain < 2");
    if (niter < 1) return b2g::wrapper_fail("b2g_msclean: need niter > 0");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return b2g::wrapper_fail("b2g_msclean: no CUDA device available (this library has no CPU path)");
    LoopStream loop_stream((cudaStream_t)stream);
    cudaStream_t s = loop_stream.s;
    const int ns = (int)nscales;
    MsCoupling cp;
    for (int i = 0; i < kMaxScales; ++i) cp.c[i] = i < ns ? coupling_diag[i] : 1.0;
    Cand *d_part = nullptr;
    const int bx = plane_blocks(ms_subtract_kernel, ns + 1);
    cudaError_t e = cudaMallocAsync(&d_part, sizeof(Cand) * bx * ns + sizeof(MsState), s);
    if (e != cudaSuccess) return cuda_fail(e);
    MsState *d_st = reinterpret_cast<MsState *>(d_part + (size_t)bx * ns);
    auto done = [&](int code) {
        cudaStreamSynchronize(s);
        cudaFreeAsync(d_part, s);
        return code;
    };
#define CK(call)                                   \
    do {                                           \
        cudaError_t e_ = (call);                   \
        if (e_ != cudaSuccess) return done(cuda_fail(e_)); \
    } while (0)
    const double stop_level = 0.9 * absolutethresh;
    CK(cudaMemsetAsync(d_st, 0, sizeof(MsState), s));
    CK(cudaMemsetAsync(comps, 0, sizeof(double) * nx * ny, s));
    const dim3 grid(bx, ns + 1);
    ms_subtract_kernel<<<grid, kThreads, 0, s>>>(res_stack, psf_scalescale, psf_scale, window_stack,
                                                 sensitivity, comps, cp, ns, (int)nx, (int)ny, (int)px,
                                                 (int)py, gain, 1, d_st, d_part);
    MsState hs;
    hs.iters = 0;
    // iteration: select (stop tests, peak) then subtract; the select after the last subtraction
    // sees iters == niter and stops
    const int rc = iterate_until_stop(s, niter, [&] {
        ms_select_kernel<<<1, kThreads, 0, s>>>(res_stack, window_stack, sensitivity, cp, ns, (int)nx,
                                                (int)ny, d_part, bx, stop_level, (int)niter, d_st);
        ms_subtract_kernel<<<grid, kThreads, 0, s>>>(res_stack, psf_scalescale, psf_scale, window_stack,
                                                     sensitivity, comps, cp, ns, (int)nx, (int)ny, (int)px,
                                                     (int)py, gain, 0, d_st, d_part);
    }, d_st, &hs);
    if (rc) return done(rc);
#undef CK
    if (iters_host) *iters_host = hs.iters;
    return done(0);
}

extern "C" int b2g_msmfsclean(double *smres, const double *psf_moments, const double *psf_scale,
                              const double *window_stack, const double *sensitivity,
                              const double *inv_hessian, const double *hessian, int findpeak,
                              int64_t nscales, int64_t nmoment, int64_t nx, int64_t ny, int64_t px,
                              int64_t py, double gain, double absolutethresh, int64_t niter,
                              double *model, int64_t *iters_host, void *stream)
{
    if (!smres || !psf_moments || !psf_scale || !inv_hessian || !model)
        return b2g::wrapper_fail("b2g_msmfsclean: NULL pointer");
    if (findpeak != 0 && findpeak != 1) return b2g::wrapper_fail("b2g_msmfsclean: findpeak must be 0 (RASCIL / Algorithm1) or 1 (CASA)");
    if (findpeak == 1 && !hessian) return b2g::wrapper_fail("b2g_msmfsclean: findpeak CASA needs the Hessians");
    if (nscales < 1 || nscales > kMaxMfScales) return b2g::wrapper_fail("b2g_msmfsclean: need 1..8 scales");
    if (nmoment < 1 || nmoment > kMaxMoments) return b2g::wrapper_fail("b2g_msmfsclean: need 1..4 moments");
    if (nx < 1 || ny < 1 || px < 1 || py < 1 || nx * ny > 0x7fffffffll)
        return b2g::wrapper_fail("b2g_msmfsclean: bad image shape");
    if (!(gain > 0.0 && gain < 2.0)) return b2g::wrapper_fail("b2g_msmfsclean: need 0 < gain < 2");
    if (niter < 1) return b2g::wrapper_fail("b2g_msmfsclean: need niter > 0");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return b2g::wrapper_fail("b2g_msmfsclean: no CUDA device available (this library has no CPU path)");
    MfHess hs;
    for (int s = 0; s < kMaxMfScales; ++s)
        for (int m = 0; m < kMaxMoments; ++m)
            for (int q = 0; q < kMaxMoments; ++q) {
                const bool in = s < nscales && m < nmoment && q < nmoment;
                hs.ih[s][m][q] = in ? inv_hessian[(s * nmoment + m) * nmoment + q] : 0.0;
                hs.h[s][m][q] = in && hessian ? hessian[(s * nmoment + m) * nmoment + q] : 0.0;
            }
    LoopStream loop_stream((cudaStream_t)stream);
    cudaStream_t st = loop_stream.s;
#define RUN(NM)                                                                                     \
    return mf_run<NM>(smres, psf_moments, psf_scale, window_stack, sensitivity, hs, (int)nscales,   \
                      (int)nx, (int)ny, (int)px, (int)py, gain, absolutethresh, niter, findpeak,    \
                      model, iters_host, st)
    switch (nmoment) {
    case 1: RUN(1);
    case 2: RUN(2);
    case 3: RUN(3);
    default: RUN(4);
    }
#undef RUN
}

extern "C" int b2g_hogbom_complex(const double *dirty_q, const double *dirty_u, const double *psf,
                                  const double *window, int64_t nx, int64_t ny, int64_t px, int64_t py,
                                  double gain, double pmax, double thresh, double fracthresh,
                                  int64_t niter, double *comps_q, double *comps_u, double *res_q,
                                  double *res_u, int64_t *iters_host, void *stream)
{
    if (!dirty_q || !dirty_u || !psf || !comps_q || !comps_u || !res_q || !res_u)
        return b2g::wrapper_fail("b2g_hogbom_complex: NULL pointer");
    if (nx < 1 || ny < 1 || px < 1 || py < 1 || nx * ny > 0x7fffffffll)
        return b2g::wrapper_fail("b2g_hogbom_complex: bad image shape");
    if (!(gain > 0.0 && gain < 2.0)) return b2g::wrapper_fail("b2g_hogbom_complex: need 0 < gain < 2");
    if (niter < 1) return b2g::wrapper_fail("b2g_hogbom_complex: need niter > 0");
    if (!(pmax > 0.0)) return b2g::wrapper_fail("b2g_hogbom_complex: need psf.max() > 0");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return b2g::wrapper_fail("b2g_hogbom_complex: no CUDA device available (this library has no CPU path)");
    LoopStream loop_stream((cudaStream_t)stream);
    cudaStream_t s = loop_stream.s;
    const long long n = nx * ny;
    Cand *d_part = nullptr;
    cudaError_t e = cudaMallocAsync(&d_part, sizeof(Cand) * (kBlocks + 1) + sizeof(CplxCleanState), s);
    if (e != cudaSuccess) return cuda_fail(e);
    Cand *d_peak = d_part + kBlocks;
    CplxCleanState *d_st = reinterpret_cast<CplxCleanState *>(d_part + kBlocks + 1);
    auto done = [&](int code) {
        cudaStreamSynchronize(s);
        cudaFreeAsync(d_part, s);
        return code;
    };
#define CK(call)                                   \
    do {                                           \
        cudaError_t e_ = (call);                   \
        if (e_ != cudaSuccess) return done(cuda_fail(e_)); \
    } while (0)
    CK(cudaMemsetAsync(d_st, 0, sizeof(CplxCleanState), s));
    CK(cudaMemsetAsync(comps_q, 0, sizeof(double) * n, s));
    CK(cudaMemsetAsync(comps_u, 0, sizeof(double) * n, s));
    if (res_q != dirty_q) CK(cudaMemcpyAsync(res_q, dirty_q, sizeof(double) * n, cudaMemcpyDeviceToDevice, s));
    if (res_u != dirty_u) CK(cudaMemcpyAsync(res_u, dirty_u, sizeof(double) * n, cudaMemcpyDeviceToDevice, s));
    // absolutethresh = max(thresh, fracthresh * max |Q + iU|) (no window)          cleaners.py:176-178
    cclean_subtract_kernel<<<kBlocks, kThreads, 0, s>>>(res_q, res_u, psf, nullptr, (int)nx, (int)ny,
                                                        (int)px, (int)py, 1, d_st, d_part);
    peak_final_kernel<<<1, kThreads, 0, s>>>(d_part, kBlocks, d_peak);
    Cand h{0.0, 0};
    CK(cudaMemcpyAsync(&h, d_peak, sizeof(Cand), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    const double absolutethresh = thresh > fracthresh * h.val ? thresh : fracthresh * h.val;
    // first peak
    cclean_subtract_kernel<<<kBlocks, kThreads, 0, s>>>(res_q, res_u, psf, window, (int)nx, (int)ny,
                                                        (int)px, (int)py, 1, d_st, d_part);
    cclean_select_kernel<<<1, kThreads, 0, s>>>(res_q, res_u, comps_q, comps_u, (int)ny, d_part, kBlocks,
                                                gain, pmax, absolutethresh, (int)niter, 1, d_st);
    CplxCleanState hs;
    hs.iters = 0;
    const int rc = iterate_until_stop(s, niter, [&] {
        cclean_subtract_kernel<<<kBlocks, kThreads, 0, s>>>(res_q, res_u, psf, window, (int)nx, (int)ny,
                                                            (int)px, (int)py, 0, d_st, d_part);
        cclean_select_kernel<<<1, kThreads, 0, s>>>(res_q, res_u, comps_q, comps_u, (int)ny, d_part,
                                                    kBlocks, gain, pmax, absolutethresh, (int)niter, 0, d_st);
    }, d_st, &hs);
    if (rc) return done(rc);
#undef CK
    if (iters_host) *iters_host = hs.iters;
    return done(0);
}
