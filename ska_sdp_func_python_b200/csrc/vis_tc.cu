// vis_tc.cu -- tensor-core degridder (tcgen05 / TMEM) for dense runs of the sorted sample list.
//
// Replaces, where consecutive sorted samples share their footprint origin, the degridding hot
// loop of ducc0.wgridder.dirty2ms (call site
// /root/reference/src/ska_sdp_func_python/imaging/ng.py:99-112).  For a run of samples with one
// origin (iu0, iv0, p0) degridding is a dense contraction over the W x W cells of the footprint,
//
//     P[s, (c, re/im)] = sum_(a,b)  ku_s[a] kv_s[b]  *  G[(a,b), (c, re/im)] ,
//     vis_s            = sum_c kw_s[c] (P[s, c, re] + i P[s, c, im]) ,
//
// i.e. a GEMM with M = samples, K = W^2 (49 -> 56) and N = 2 W (14 -> 16), accumulated in tensor
// memory.
//
//   * one thread = one sample = one row of A: it evaluates its 3 W kernel values, forms the W^2
//     products ku[a] kv[b] and writes them straight from registers into its TMEM lane
//     (tcgen05.st) -- no shared-memory staging and no cross-lane reduction; the w-kernel stays in
//     the thread's registers for the W-term epilogue on its own accumulator row (tcgen05.ld);
//   * a tile = up to 128 consecutive sorted samples of one run = the 128 TMEM lanes of the block;
//     thread 0 issues the MMAs, completion comes back through an mbarrier (tcgen05.commit);
//   * B = the W^2 x 2W block of grid cells under the run's footprint, loaded once per run into
//     shared memory in the K-major core-matrix layout (the next run's block is loaded while the
//     current tile's MMAs run);
//   * fp32 accuracy from TF32 tensor cores: A = A_hi + A_lo, B = B_hi + B_lo (11 + 11 mantissa
//     bits each), three MMAs per K-step (hi*hi, lo*hi, hi*lo), fp32 accumulation: measured
//     1.1e-6 of the row maximum against fp64 (tools/probe/tc_probe.cu), against 8e-4 for plain
//     TF32;
//   * persistent blocks work through chunks of 1024 sorted samples: a lane-parallel first phase
//     locates all of them (fp64, as the sort key did) into shared memory, then the tiles follow
//     each other.  Chunks whose runs are short (sparse outer uv-plane) are left to the warp-run
//     kernel (vis_warprun.cu), which costs less per run.
//
// Measured alternative (same file, earlier revision): every warp running its own 32-row tiles in
// its TMEM lane quadrant with the other quadrants masked off (disable-output-lane) needs no block
// barrier, but the 21 MMA issues per tile (about 11 instructions each: operands moved to uniform
// registers) are then paid per 27 samples instead of per 73, and it measured 13 % slower.
#include <cuda_runtime.h>

#include "kernels.h"

namespace b2g {

namespace {

constexpr int kTcThreads = 128;
constexpr int kTcChunk = 1024;                       // samples per block (== 4 warps x B2G_CHUNK)
constexpr int kTcPer = kTcChunk / kTcThreads;
constexpr int kTmCols = 128;                         // TMEM columns per block: A_hi | A_lo | D
constexpr int kN = 16;                               // accumulator columns: (plane, re/im), padded

template <int W> struct TcCfg {
    static constexpr int K = W * W;
    static constexpr int KP = (K + 7) & ~7;          // padded to whole K-steps of 8
    static constexpr int KS = KP / 8;
    static constexpr int COL_ALO = KP, COL_D = 2 * KP;
    static_assert(2 * KP + kN <= kTmCols, "A_hi, A_lo and D must fit the TMEM allocation");
    static_assert(2 * W <= kN, "one accumulator column per (plane, re/im)");
    static constexpr int B_FLOATS = KS * 128;        // one operand copy (hi or lo) of one buffer
};

template <int W> struct TcSmem {
    uint32_t idx[kTcChunk];
    float wt[kTcChunk];
    int oxy[kTcChunk];        // iu0 << 16 | iv0
    int oz[kTcChunk];         // p0 | flip << 31
    float du[kTcChunk], dv[kTcChunk], dw[kTcChunk];
    alignas(128) float B[2][2][TcCfg<W>::B_FLOATS];   // [buffer][hi, lo]
    alignas(8) unsigned long long bar;
    uint32_t tm_base;
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void tm_st8(uint32_t taddr, const float (&v)[8])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr),
                 "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
                 "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
                 "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7]))
                 : "memory");
}
__device__ __forceinline__ void tm_ld16(uint32_t taddr, float (&v)[16])
{
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, "
                 "%13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]),
                   "=r"(r[15])
                 : "r"(taddr)
                 : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
// D[tmem] (+)= A[tmem] * B[smem]^T, kind::tf32, issued by one thread; mask bit = 1: that TMEM lane
// of D is not written (bit i of m_j <-> lane 32 j + i), which lets every warp of the block run its
// own 32-row tiles in its own lane quadrant of the shared columns
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t bdesc, uint32_t idesc, uint32_t acc,
                                       uint32_t m0, uint32_t m1, uint32_t m2, uint32_t m3)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, {%5, %6, %7, %8}, p;\n\t"
        "}\n" ::"r"(d), "r"(a), "l"(bdesc), "r"(idesc), "r"(acc), "r"(m0), "r"(m1), "r"(m2), "r"(m3)
        : "memory");
}
__device__ __forceinline__ bool mbar_try(unsigned long long *bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(ok)
                 : "r"(smem_u32(bar)), "r"(parity)
                 : "memory");
    return ok != 0;
}
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

// B element (n, k), in floats from the start of one operand copy: K-steps of 512 B; inside a
// step the two 8-row groups are 256 B apart (stride byte offset), the two 16-byte K chunks 128 B
// apart (leading byte offset), rows 16 B apart -- the no-swizzle K-major core-matrix layout
__device__ __forceinline__ int b_off(int n, int k)
{
    return (k >> 3) * 128 + (n >> 3) * 64 + ((k >> 2) & 1) * 32 + (n & 7) * 4 + (k & 3);
}

// The W x W x W block of grid cells under the footprint with origin (oxy, p0), split in hi / lo
// TF32 parts, into one B buffer, by the whole block.  Lane = (plane c = lane / 4, cell k mod 4),
// the four warps take every fourth group of four cells; the pad rows (k >= W^2) and columns
// (n >= 2 W) stay zero from the initial clear.
template <int W>
__device__ __forceinline__ void load_B(const VisGeo &g, const cplx<float> *__restrict__ ring, int plane0,
                                       int slot0, int oxy, int p0, float *Bhi, float *Blo, int tid)
{
    using C = TcCfg<W>;
    constexpr int ROUNDS = (C::K + 15) / 16;
    const int lane = tid & 31, warp = tid >> 5;
    const int c = lane >> 2;
    if (c >= W) return;
    const int iu0 = oxy >> 16, iv0 = oxy & 0xffff;
    const cplx<float> *base = ring + (long long)wrap_inc(p0 - plane0 + slot0 + c, g.nslots) * g.plane_stride;
    cplx<float> v[ROUNDS];
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
        const int k = 4 * (warp + 4 * r) + (lane & 3);
        v[r] = cplx<float>{0.f, 0.f};
        if (k < C::K) {
            const int a = k / W, b = k - a * W;
            v[r] = base[(long long)wrap_inc(iu0 + a, g.nu) * g.nv + wrap_inc(iv0 + b, g.nv)];
        }
    }
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
        const int k = 4 * (warp + 4 * r) + (lane & 3);
        if (k < C::K) {
            const int o = b_off(2 * c, k);
            const float hr = tf32_hi(v[r].re), hi = tf32_hi(v[r].im);
            Bhi[o] = hr;
            Bhi[o + 4] = hi;
            Blo[o] = v[r].re - hr;
            Blo[o + 4] = v[r].im - hi;
        }
    }
}

template <int W>
__global__ void __launch_bounds__(kTcThreads, 4)
degrid_tc_kernel(VisGeo g, KernelTab<float, W> tab, const uint32_t *__restrict__ sidx, long long begin,
                 long long end, int plane0, int slot0, const float *__restrict__ wgt, cplx<float> *out,
                 const cplx<float> *sub_from, const cplx<float> *__restrict__ ring,
                 const unsigned char *__restrict__ dense)
{
    using C = TcCfg<W>;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    TcSmem<W> &sm = *reinterpret_cast<TcSmem<W> *>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tm_base)),
                     "r"(kTmCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0) {
        // four arrivals per tile: lane 0 of every warp issues a share of the MMAs and commits
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 4;" ::"r"(smem_u32(&sm.bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < 4 * C::B_FLOATS; i += kTcThreads) (&sm.B[0][0][0])[i] = 0.f;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = sm.tm_base;
    const uint32_t lane_base = tm + ((uint32_t)(warp * 32) << 16);
    // instruction descriptor: D fp32, A / B tf32, both K-major, N = 16, M = 128
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kN >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t dbase = (uint64_t(128 >> 4) << 16) | (uint64_t(256 >> 4) << 32) | (1ull << 46);

    int bb = 0;                           // B buffer of the current tile
    int cur_xy = -1, cur_z = -1;          // origin whose cells buffer bb holds
    int oth_xy = -1, oth_z = -1;          // ... and buffer bb ^ 1
    uint32_t parity = 0;
    // persistent blocks: chunk = 1024 sorted samples, dealt out round-robin
    const long long nchunk = (end - begin + kTcChunk - 1) / kTcChunk;
    for (long long ch = blockIdx.x; ch < nchunk; ch += gridDim.x) {
        if (dense && dense[ch] != 1) continue;        // not a long-run chunk: another kernel's
        const long long cbeg = begin + ch * kTcChunk;
        const int n = (int)((end - cbeg < kTcChunk) ? end - cbeg : kTcChunk);
        __syncthreads();                              // the previous chunk's arrays are no longer read

        // ---- phase 1: locate every sample of the chunk (one sample per thread and round); the
        // loads of four rounds are in flight together ------------------------------------------
#pragma unroll
        for (int r0 = 0; r0 < kTcPer; r0 += 4) {
            uint32_t idx[4];
            double uu[4], vv[4], ww[4], fs[4];
            float wt[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int s = tid + kTcThreads * (r0 + r);
                idx[r] = s < n ? __ldg(sidx + cbeg + s) : 0u;
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int s = tid + kTcThreads * (r0 + r);
                uint32_t row = 0, chan = 0;
                if (s < n) g.split(idx[r], row, chan);
                fs[r] = __ldg(g.fscale + chan);
                uu[r] = __ldg(g.uvw + 3ull * row + 0);
                vv[r] = __ldg(g.uvw + 3ull * row + 1);
                ww[r] = __ldg(g.uvw + 3ull * row + 2);
                wt[r] = (wgt && s < n) ? __ldg(wgt + idx[r]) : 1.f;
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int s = tid + kTcThreads * (r0 + r);
                if (s < n) {
                    const Located L = locate_uvw(g, W, uu[r] * fs[r], vv[r] * fs[r], ww[r] * fs[r]);
                    sm.idx[s] = idx[r];
                    sm.wt[s] = wt[r];
                    sm.oxy[s] = (L.iu0 << 16) | L.iv0;
                    sm.oz[s] = L.p0 | (L.flip ? (int)0x80000000 : 0);
                    sm.du[s] = (float)L.du;
                    sm.dv[s] = (float)L.dv;
                    sm.dw[s] = (float)L.dw;
                }
            }
        }
        __syncthreads();

        int pos = 0;
        while (pos < n) {                 // block-uniform
            const int s = pos + tid;
            const int oxy0 = sm.oxy[pos], oz0 = sm.oz[pos] & 0x7fffffff;
            const bool match = s < n && sm.oxy[s] == oxy0 && (sm.oz[s] & 0x7fffffff) == oz0;
            // the samples sharing the first one's origin are a prefix of the sorted window
            const int tile_len = __syncthreads_count(match);
            const float my_w = match ? sm.wt[s] : 0.f;
            const bool active = match && my_w != 0.f;
            const bool warp_on = __any_sync(0xffffffffu, match);

            // ---- B: the grid cells of this run (usually prefetched during the previous tile) ----
            if (cur_xy != oxy0 || cur_z != oz0) {
                if (oth_xy == oxy0 && oth_z == oz0) {
                    bb ^= 1;
                    oth_xy = cur_xy;
                    oth_z = cur_z;
                } else {
                    load_B<W>(g, ring, plane0, slot0, oxy0, oz0, sm.B[bb][0], sm.B[bb][1], tid);
                }
                cur_xy = oxy0;
                cur_z = oz0;
            }

            // ---- A: this thread's row = ku (x) kv, hi and lo parts, registers -> TMEM ------------
            float kw[W];
            if (warp_on) {
                float ku[W], kv[W];
                if (active) {
                    kernel_values<float, W>(tab, sm.du[s], ku);
                    kernel_values<float, W>(tab, sm.dv[s], kv);
                    kernel_values<float, W>(tab, sm.dw[s], kw);
                } else {
#pragma unroll
                    for (int i = 0; i < W; ++i) ku[i] = kv[i] = kw[i] = 0.f;
                }
#pragma unroll
                for (int c0 = 0; c0 < C::KP; c0 += 8) {
                    float hi[8], lo[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int k = c0 + i;
                        float p = 0.f;
                        if (k < C::K) p = ku[k / W] * kv[k % W];
                        hi[i] = p;                      // the tensor core reads the upper 19 bits
                        lo[i] = p - tf32_hi(p);
                    }
                    tm_st8(lane_base + c0, hi);
                    tm_st8(lane_base + C::COL_ALO + c0, lo);
                }
                // every MMA of the tile accumulates (they are issued by four threads, in no
                // particular order): the accumulator rows start from zero
                const float zero[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                tm_st8(lane_base + C::COL_D, zero);
                tm_st8(lane_base + C::COL_D + 8, zero);
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // B stores -> tensor core reads
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncthreads();
            if (lane == 0) {
                // K-steps warp, warp + 4, ...: the ~11 instructions of operand set-up per MMA are the
                // serial part of a tile, so they are spread over the four warps (measured safe:
                // tools/probe/tc_probe.cu mode 3)
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t bh_desc = dbase | (uint64_t)((smem_u32(sm.B[bb][0]) >> 4) & 0x3fff);
                const uint64_t bl_desc = dbase | (uint64_t)((smem_u32(sm.B[bb][1]) >> 4) & 0x3fff);
#pragma unroll
                for (int k0 = 0; k0 < C::KS; k0 += 4) {
                    const int ks = k0 + warp;
                    if (ks < C::KS) {
                        const uint64_t bh = bh_desc + (uint64_t)(ks * 32), bl = bl_desc + (uint64_t)(ks * 32);
                        mma_ts(tm + C::COL_D, tm + ks * 8, bh, idesc, 1, 0u, 0u, 0u, 0u);
                        mma_ts(tm + C::COL_D, tm + C::COL_ALO + ks * 8, bh, idesc, 1, 0u, 0u, 0u, 0u);
                        mma_ts(tm + C::COL_D, tm + ks * 8, bl, idesc, 1, 0u, 0u, 0u, 0u);
                    }
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                                 smem_u32(&sm.bar))
                             : "memory");
            }
            // ---- while the MMAs run: the next run's cells into the other buffer ------------------
            const int npos = pos + tile_len;
            if (npos < n) {
                const int nxy = sm.oxy[npos], nz = sm.oz[npos] & 0x7fffffff;
                if ((nxy != oxy0 || nz != oz0) && (oth_xy != nxy || oth_z != nz)) {
                    load_B<W>(g, ring, plane0, slot0, nxy, nz, sm.B[bb ^ 1][0], sm.B[bb ^ 1][1], tid);
                    oth_xy = nxy;
                    oth_z = nz;
                }
            }
            if (lane == 0) {              // one poller per warp
                const long long t0 = clock64();
                while (!mbar_try(&sm.bar, parity))
                    if (clock64() - t0 > 4000000000LL) __trap();    // never hang the device
            }
            __syncwarp();
            parity ^= 1u;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

            // ---- epilogue: this thread's accumulator row, W-term w-kernel sum, store -------------
            if (warp_on) {
                float d[16];
                tm_ld16(lane_base + C::COL_D, d);
                if (match) {
                    float re = 0.f, im = 0.f;
                    if (active) {
#pragma unroll
                        for (int c = 0; c < W; ++c) {
                            re = fmaf(kw[c], d[2 * c], re);
                            im = fmaf(kw[c], d[2 * c + 1], im);
                        }
                        if (sm.oz[s] < 0) im = -im;   // flipped sample: conjugate
                    }
                    const uint32_t idx = sm.idx[s];
                    cplx<float> r{re * my_w, im * my_w};
                    if (sub_from) {                   // residual mode: data - model
                        const cplx<float> prev = sub_from[idx];
                        r.re = prev.re - r.re;
                        r.im = prev.im - r.im;
                    }
                    out[idx] = r;
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            pos = npos;
        }
    }
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(kTmCols) : "memory");
}

// class of chunk c of [begin, end): 1 = at most `max_runs` runs of equal footprint origin, 2 = at
// least `sparse_runs` runs, else 0 (and 0 for the ragged last chunk)
template <int W>
__global__ void __launch_bounds__(256)
tc_classify_kernel(VisGeo g, const uint32_t *__restrict__ sidx, long long begin, long long end, int max_runs,
                   int sparse_runs, unsigned char *__restrict__ dense)
{
    __shared__ int s_count;
    const long long cbeg = begin + (long long)blockIdx.x * kTcChunk;
    const int n = (int)((end - cbeg < kTcChunk) ? end - cbeg : kTcChunk);
    if (threadIdx.x == 0) s_count = 0;
    __syncthreads();
    int changes = 0;
    // thread t looks at samples 4t .. 4t+3 and their predecessor
    long long key_prev = -1;
    for (int i = -1; i < 4; ++i) {
        const int s = 4 * (int)threadIdx.x + i;
        if (s < 0 || s >= n) {
            if (i == -1) key_prev = -1;
            continue;
        }
        uint32_t row, chan;
        g.split(sidx[cbeg + s], row, chan);
        const Located L = locate(g, W, row, chan);
        const long long key = ((long long)L.p0 << 32) | ((long long)L.iu0 << 16) | L.iv0;
        if (i >= 0 && key != key_prev) ++changes;
        key_prev = key;
    }
    atomicAdd(&s_count, changes);
    __syncthreads();
    if (threadIdx.x == 0)
        dense[blockIdx.x] = n != kTcChunk ? 0 : (s_count <= max_runs ? 1 : (s_count >= sparse_runs ? 2 : 0));
}

template <int W> int degrid_tc_inst(const VisLaunch<float> &L, unsigned char *dense_scratch, int max_runs,
                                   int sparse_runs)
{
    const long long nvis = L.end - L.begin;
    if (nvis <= 0) return 0;
    const unsigned nblk = (unsigned)((nvis + kTcChunk - 1) / kTcChunk);
    const size_t smem = sizeof(TcSmem<W>);
    cudaError_t e = cudaFuncSetAttribute(degrid_tc_kernel<W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    if (dense_scratch && max_runs >= 0) {
        tc_classify_kernel<W><<<nblk, 256, 0, L.stream>>>(L.geo, L.sidx, L.begin, L.end, max_runs, sparse_runs,
                                                          dense_scratch);
        if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
    }
    // persistent warps: as many blocks as are resident at once (4 per SM: 128 TMEM columns each)
    static int nsm = 0;
    if (nsm == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        if (cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || nsm <= 0) nsm = 148;
    }
    const unsigned ngrid = nblk < (unsigned)(4 * nsm) ? nblk : (unsigned)(4 * nsm);
    degrid_tc_kernel<W><<<ngrid, kTcThreads, smem, L.stream>>>(L.geo, make_tab<float, W>(L.coef), L.sidx, L.begin,
                                                              L.end, L.plane0, L.slot0, L.wgt, L.ms_out, L.sub_from,
                                                              L.ring, dense_scratch);
    return (int)cudaGetLastError();
}

}  // namespace

bool tc_supported(int W, bool wstack) { return wstack && W >= 4 && W <= 7; }

long long tc_chunks(long long nvis) { return (nvis + kTcChunk - 1) / kTcChunk; }

// Degrids the dense chunks of [begin, end) (all chunks if dense_scratch is null).  dense_scratch
// holds the classification (one byte per chunk of 1024 samples of the range), which the warp-run
// launch then uses to skip those chunks; it is (re)computed first unless max_runs < 0 (the
// caller has kept it from an earlier call on the same range of the same plan).
int launch_degrid_tc(const VisLaunch<float> &L, unsigned char *dense_scratch, int max_runs, int sparse_runs)
{
    switch (L.W) {
    case 4: return degrid_tc_inst<4>(L, dense_scratch, max_runs, sparse_runs);
    case 5: return degrid_tc_inst<5>(L, dense_scratch, max_runs, sparse_runs);
    case 6: return degrid_tc_inst<6>(L, dense_scratch, max_runs, sparse_runs);
    case 7: return degrid_tc_inst<7>(L, dense_scratch, max_runs, sparse_runs);
    default: return -1;
    }
}

}  // namespace b2g
