// vis_sort.cu -- the uv-cell / w-plane sort ("bucket") pass.
//
// Every visibility has the key  ((chunk*nplanes + p0)*nu + iu0)*nv + iv0  where chunk is the
// row chunk of a host-streaming plan (0 otherwise) and (iu0, iv0, p0) is the
// lower corner of its kernel footprint (first cell in u and v, first w-plane), and the
// (key, index) pairs are radix-sorted; a histogram of p0 gives the slab boundaries.  This is the B200 counterpart of the scan/bucket pass at the start
// of ducc0.wgridder.ms2dirty / dirty2ms (call sites
// /root/reference/src/ska_sdp_func_python/imaging/ng.py:240-256, :99-112): ducc0 buckets
// (row, channel-range) runs per 16x16 tile and w-plane on the host; here the sort is
// per visibility and exact, because the gridding kernels accumulate runs of identical
// origins in registers.
//
// The sort moves 4-byte sample indices only: every pass recomputes the key of a sample from
// its index (locate() is ~60 fp64/integer instructions, cheaper than carrying 8-byte keys
// through HBM twice per pass), so the scratch is one index array (4 B/sample) next to the
// result instead of 20 B/sample -- what lets the 4e9-sample C4 plan fit one GPU.
//
// LSD radix sort, 8 bits per pass, three kernels per pass:
//   1. per-tile digit histogram (shared-memory counters, one tile = 4096 pairs per block)
//   2. exclusive scan of the [digit][tile] table (single cooperative-free two-level scan)
//   3. stable scatter: each warp ranks its 128-pair strip with match.any, strips are
//      combined through a shared-memory digit table.
#include <algorithm>

#include "kernels.h"

namespace b2g {

namespace {

// key of sample idx: ((chunk*nplanes + p0)*nu + iu0)*nv + iv0
__device__ __forceinline__ uint64_t sample_key(const VisGeo &g, int W, uint32_t rows_per_chunk,
                                               uint32_t idx, int *p0_out = nullptr)
{
    uint32_t row, chan;
    g.split(idx, row, chan);
    const Located L = locate(g, W, row, chan);
    if (p0_out) *p0_out = L.p0;
    // row chunks (host-streaming plans) are the most significant key component
    uint64_t key = rows_per_chunk ? (uint64_t)(row / rows_per_chunk) * (uint64_t)g.nplanes : 0ull;
    key = (key + (uint64_t)L.p0) * (uint64_t)g.nu + (uint64_t)L.iu0;
    return key * (uint64_t)g.nv + (uint64_t)L.iv0;
}

__global__ void __launch_bounds__(256)
plane_hist_kernel(VisGeo g, int W, long long nvis, unsigned long long *__restrict__ plane_counts)
{
    extern __shared__ unsigned int sh_cnt[];  // [nplanes] when nplanes <= 4096
    const int np = g.nplanes;
    const bool use_sh = np <= 4096;
    if (use_sh) {
        for (int i = threadIdx.x; i < np; i += blockDim.x) sh_cnt[i] = 0;
        __syncthreads();
    }
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvis;
         i += (long long)gridDim.x * blockDim.x) {
        int p0;
        sample_key(g, W, 0u, (uint32_t)i, &p0);
        if (use_sh) atomicAdd(&sh_cnt[p0], 1u);
        else atomicAdd(&plane_counts[p0], 1ull);
    }
    if (use_sh) {
        __syncthreads();
        for (int s = threadIdx.x; s < np; s += blockDim.x)
            if (sh_cnt[s]) atomicAdd(&plane_counts[s], (unsigned long long)sh_cnt[s]);
    }
}

// min / max of |w| in wavelengths over the selected samples of a plan: out2 = {min, max}
__global__ void __launch_bounds__(256)
sel_wstats_kernel(VisGeo g, long long nsel, double *__restrict__ out2)
{
    __shared__ double sh[2][256];
    double wmin = 1e300, wmax = 0.0;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < nsel;
         i += (long long)gridDim.x * 256) {
        uint32_t row, chan;
        g.split((uint32_t)i, row, chan);
        const double w = fabs(__ldg(g.uvw + 3ull * row + 2) * __ldg(g.fscale + chan));
        wmin = fmin(wmin, w);
        wmax = fmax(wmax, w);
    }
    sh[0][threadIdx.x] = wmin;
    sh[1][threadIdx.x] = wmax;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) {
            sh[0][threadIdx.x] = fmin(sh[0][threadIdx.x], sh[0][threadIdx.x + s]);
            sh[1][threadIdx.x] = fmax(sh[1][threadIdx.x], sh[1][threadIdx.x + s]);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        // non-negative doubles: integer compare on the bit pattern is order-preserving
        atomicMin((unsigned long long *)&out2[0], (unsigned long long)__double_as_longlong(sh[0][0]));
        atomicMax((unsigned long long *)&out2[1], (unsigned long long)__double_as_longlong(sh[1][0]));
    }
}

constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;
constexpr int kSortThreads = 256;
constexpr int kItems = 16;                          // pairs per thread
constexpr int kTile = kSortThreads * kItems;        // 4096 pairs per block

__global__ void __launch_bounds__(kSortThreads)
radix_hist_kernel(VisGeo g, int W, uint32_t rows_per_chunk, const uint32_t *__restrict__ idx_in,
                  long long n, int shift, int ntiles,
                  uint32_t *__restrict__ table /* [kRadix][ntiles] */)
{
    __shared__ unsigned int cnt[kRadix];
    cnt[threadIdx.x] = 0;  // kSortThreads == kRadix
    __syncthreads();
    const long long base = (long long)blockIdx.x * kTile;
#pragma unroll
    for (int it = 0; it < kItems; ++it) {
        const long long i = base + (long long)it * kSortThreads + threadIdx.x;
        const bool live = i < n;
        const unsigned livemask = __ballot_sync(0xffffffffu, live);
        if (live) {
            const uint32_t id = idx_in ? idx_in[i] : (uint32_t)i;  // first pass: identity
            const unsigned d = (unsigned)(sample_key(g, W, rows_per_chunk, id) >> shift) & (kRadix - 1);
            // warp-aggregated increment: consecutive visibilities mostly share the digit
            const unsigned peers = __match_any_sync(livemask, d);
            if ((int)(threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&cnt[d], __popc(peers));
        }
    }
    __syncthreads();
    table[(size_t)threadIdx.x * ntiles + blockIdx.x] = cnt[threadIdx.x];
}

// Exclusive scan over the table in (digit-major, tile-minor) order.  Two levels:
// scan_partial sums blocks of 1024 entries, scan_finish scans the block sums (single
// block) and scan_apply writes the exclusive prefix back.
__global__ void __launch_bounds__(256)
scan_partial_kernel(const uint32_t *__restrict__ in, long long n, uint64_t *__restrict__ bsum)
{
    __shared__ unsigned long long sh[256];
    const long long base = (long long)blockIdx.x * 1024;
    unsigned long long s = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const long long i = base + threadIdx.x * 4 + k;
        if (i < n) s += in[i];
    }
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if ((int)threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) bsum[blockIdx.x] = sh[0];
}
__global__ void __launch_bounds__(1024)
scan_finish_kernel(uint64_t *__restrict__ bsum, long long nb)
{
    // single block: sequential over chunks of 1024 with a running carry
    __shared__ unsigned long long sh[1024];
    __shared__ unsigned long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (long long base = 0; base < nb; base += 1024) {
        const long long i = base + threadIdx.x;
        const unsigned long long v = (i < nb) ? bsum[i] : 0ull;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int off = 1; off < 1024; off <<= 1) {
            unsigned long long t = 0;
            if ((int)threadIdx.x >= off) t = sh[threadIdx.x - off];
            __syncthreads();
            sh[threadIdx.x] += t;
            __syncthreads();
        }
        const unsigned long long incl = sh[threadIdx.x];
        if (i < nb) bsum[i] = carry + incl - v;  // exclusive
        __syncthreads();
        if (threadIdx.x == 1023) carry += incl;
        __syncthreads();
    }
}
__global__ void __launch_bounds__(256)
scan_apply_kernel(const uint32_t *__restrict__ in, long long n, const uint64_t *__restrict__ bsum,
                  uint64_t *__restrict__ out)
{
    __shared__ unsigned long long sh[256];
    const long long base = (long long)blockIdx.x * 1024;
    unsigned int v[4];
    unsigned long long s = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const long long i = base + threadIdx.x * 4 + k;
        v[k] = (i < n) ? in[i] : 0u;
        s += v[k];
    }
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int off = 1; off < 256; off <<= 1) {
        unsigned long long t = 0;
        if ((int)threadIdx.x >= off) t = sh[threadIdx.x - off];
        __syncthreads();
        sh[threadIdx.x] += t;
        __syncthreads();
    }
    unsigned long long run = bsum[blockIdx.x] + sh[threadIdx.x] - s;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const long long i = base + threadIdx.x * 4 + k;
        if (i < n) out[i] = run;
        run += v[k];
    }
}

// Stable scatter.  Block = one tile of kTile pairs, processed in kItems rounds of 256
// pairs in input order.  Within a round: warp-level ranks by match.any, cross-warp
// offsets through warp_cnt[warp][digit]; running per-digit offsets carry across rounds.
__global__ void __launch_bounds__(kSortThreads)
radix_scatter_kernel(VisGeo g, int W, uint32_t rows_per_chunk, const uint32_t *__restrict__ idx_in,
                     uint32_t *__restrict__ idx_out, long long n, int shift, int ntiles,
                     const uint64_t *__restrict__ offs /* [kRadix][ntiles] */)
{
    constexpr int NW = kSortThreads / 32;
    __shared__ unsigned long long dig_base[kRadix];   // global position of next pair of digit d
    __shared__ unsigned int warp_cnt[NW][kRadix];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    dig_base[threadIdx.x] = offs[(size_t)threadIdx.x * ntiles + blockIdx.x];
    const long long base = (long long)blockIdx.x * kTile;
    for (int it = 0; it < kItems; ++it) {
        for (int w = 0; w < NW; ++w) warp_cnt[w][threadIdx.x] = 0;
        __syncthreads();
        const long long i = base + (long long)it * kSortThreads + threadIdx.x;
        const bool live = i < n;
        uint32_t id = 0;
        unsigned d = 0, rank_in_warp = 0;
        if (live) {
            id = idx_in ? idx_in[i] : (uint32_t)i;
            d = (unsigned)(sample_key(g, W, rows_per_chunk, id) >> shift) & (kRadix - 1);
        }
        const unsigned livemask = __ballot_sync(0xffffffffu, live);
        if (live) {
            const unsigned peers = __match_any_sync(livemask, d);
            rank_in_warp = __popc(peers & ((1u << lane) - 1u));
            if (rank_in_warp == 0) warp_cnt[warp][d] = __popc(peers);
        }
        __syncthreads();
        unsigned long long pos = 0;
        if (live) {
            unsigned before = 0;
            for (int w = 0; w < warp; ++w) before += warp_cnt[w][d];
            pos = dig_base[d] + before + rank_in_warp;
        }
        __syncthreads();
        {   // advance the running offsets (thread t handles digit t)
            unsigned tot = 0;
#pragma unroll
            for (int w = 0; w < NW; ++w) tot += warp_cnt[w][threadIdx.x];
            dig_base[threadIdx.x] += tot;
        }
        if (live) idx_out[pos] = id;
        __syncthreads();
    }
}

}  // namespace

int launch_sel_wstats(const VisGeo &geo, long long nsel, double *out2_dev, cudaStream_t stream)
{
    const double init[2] = {1e300, 0.0};
    cudaMemcpyAsync(out2_dev, init, sizeof(init), cudaMemcpyHostToDevice, stream);
    if (nsel <= 0) return 0;
    const unsigned blocks = (unsigned)std::min<long long>((nsel + 255) / 256, 148 * 16);
    sel_wstats_kernel<<<blocks, 256, 0, stream>>>(geo, nsel, out2_dev);
    return (int)cudaGetLastError();
}

int launch_plane_hist(const VisGeo &geo, int W, long long nvis, unsigned long long *plane_counts,
                      cudaStream_t stream)
{
    cudaMemsetAsync(plane_counts, 0, sizeof(unsigned long long) * geo.nplanes, stream);
    if (nvis <= 0) return 0;
    const unsigned blocks = (unsigned)std::min<long long>((nvis + 255) / 256, 148 * 32);
    const size_t sh = geo.nplanes <= 4096 ? sizeof(unsigned int) * geo.nplanes : 0;
    plane_hist_kernel<<<blocks, 256, sh, stream>>>(geo, W, nvis, plane_counts);
    return (int)cudaGetLastError();
}

// Sorted sample indices into idx_out; idx_tmp is the one scratch array (n entries each).
// Table scratch is allocated per call with the stream-ordered allocator.
int sort_samples(const VisGeo &geo, int W, uint32_t rows_per_chunk, uint32_t *idx_tmp,
                 uint32_t *idx_out, long long n, int key_bits, cudaStream_t stream)
{
    if (n <= 0) return 0;
    const int ntiles = (int)((n + kTile - 1) / kTile);
    const long long tab_n = (long long)kRadix * ntiles;
    const long long nb = (tab_n + 1023) / 1024;
    uint32_t *table = nullptr;
    uint64_t *offs = nullptr, *bsum = nullptr;
    cudaError_t e;
    if ((e = cudaMallocAsync(&table, sizeof(uint32_t) * tab_n, stream)) != cudaSuccess) return (int)e;
    if ((e = cudaMallocAsync(&offs, sizeof(uint64_t) * tab_n, stream)) != cudaSuccess) return (int)e;
    if ((e = cudaMallocAsync(&bsum, sizeof(uint64_t) * nb, stream)) != cudaSuccess) return (int)e;
    int npass = (key_bits + kRadixBits - 1) / kRadixBits;
    if (npass < 1) npass = 1;
    // ping-pong so that the last pass writes idx_out; the first pass reads the identity
    const uint32_t *src = nullptr;
    uint32_t *dst = (npass & 1) ? idx_out : idx_tmp;
    for (int p = 0; p < npass; ++p) {
        const int shift = p * kRadixBits;
        radix_hist_kernel<<<ntiles, kSortThreads, 0, stream>>>(geo, W, rows_per_chunk, src, n, shift,
                                                              ntiles, table);
        scan_partial_kernel<<<(unsigned)nb, 256, 0, stream>>>(table, tab_n, bsum);
        scan_finish_kernel<<<1, 1024, 0, stream>>>(bsum, nb);
        scan_apply_kernel<<<(unsigned)nb, 256, 0, stream>>>(table, tab_n, bsum, offs);
        radix_scatter_kernel<<<ntiles, kSortThreads, 0, stream>>>(geo, W, rows_per_chunk, src, dst, n,
                                                                 shift, ntiles, offs);
        src = dst;
        dst = (dst == idx_out) ? idx_tmp : idx_out;
    }
    cudaFreeAsync(table, stream);
    cudaFreeAsync(offs, stream);
    cudaFreeAsync(bsum, stream);
    return (int)cudaGetLastError();
}

}  // namespace b2g
