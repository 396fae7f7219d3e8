// Stage 4b: binning for all frames of a call, without a global key sort and without a host sync.
//
// Upstream (InclusiveSum over Gaussians -> duplicateWithKeys -> 64-bit SortPairs ->
// identifyTileRanges) sorts D (tile<<32 | depth) keys with ~6 radix passes
// over HBM.  Here the tile is known when an entry is emitted, so:
//   1. preprocess already counted entries per tile (tile_count);
//   2. exclusive scan of tile_count (warp-shuffle scan, 3 small launches) -> tile_start = the
//      upstream `ranges`, plus D at tile_start[last];
//   3. scatter: every (Gaussian, tile) pair drops (depth_bits<<32 | gaussian) into its tile's bucket;
//   4. the buckets are sorted in shared memory on the 64-bit entry (depth bits, then Gaussian index: unique keys) and
//      the Gaussian indices written out: groups of consecutive tiles of up to 4096 entries per pass of an O(n)
//      depth-bin sort (sort_groups), larger buckets one at a time (sort_tiles<1024, 16384>).
// The result equals the stable sort of upstream keys: within a tile ascending depth bits, ties in
// ascending Gaussian index (upstream emits in index order and its radix sort is stable).
// HBM traffic: D*(8 w + 8 r + 4 w) instead of D*12*2*passes.
#include "common.cuh"

#ifndef AMAV_SCAT_CTAS
#define AMAV_SCAT_CTAS 4            // resident scatter CTAs per SM (measured: 2, 3 and 4 time the same)
#endif
#ifndef AMAV_SCAT_CHUNK
// Gaussians per scatter CTA.  512 keeps the frames in flight (resident CTAs / CTAs per frame: 2.3 at 100k Gaussians)
// and their ~12 MB each of half-written bucket sectors inside the L2; at 2048 nine frames were in flight, partially
// written sectors were evicted and the DRAM paid read-modify-writes: bin 2.20 -> 1.82 ms per 120 frames (1024: 2.05,
// 256: 1.86), 6.30 -> 5.47 ms at 500k Gaussians / 2048^2.
#define AMAV_SCAT_CHUNK 512
#endif
#ifndef AMAV_SCAT_LDCS
#define AMAV_SCAT_LDCS 0            // 1: rectangle / depth loads with the evict-first hint
#endif
#ifndef AMAV_SORT_GROUPS
#define AMAV_SORT_GROUPS 1          // 1: grouped small-bucket sort (sort_groups); 0: one tile at a time (sort_tiles<256, 4096>)
#endif

namespace {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 4;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;   // 1024 counts per block

__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) >= o) v += n;
    }
    return v;
}

// Exclusive scan of one value per thread across a 256-thread block; returns the block total in `total`.
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t &total)
{
    __shared__ uint32_t warp_sums[SCAN_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t inc = warp_inclusive_scan(v);
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; ++w) {
        const uint32_t s = warp_sums[w];
        if (w < warp) base += s;
        tot += s;
    }
    __syncthreads();
    total = tot;
    return base + inc - v;
}

// block_sums[b] = entries of block b's tiles, ne_sums[b] = how many of them are non-empty
__global__ void __launch_bounds__(SCAN_THREADS) scan_block_sums(const uint32_t *__restrict__ count, int64_t n,
                                                                 uint32_t *__restrict__ block_sums, uint32_t *__restrict__ ne_sums)
{
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t s = 0, z = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k)
        if (base + k < n) { const uint32_t c = count[base + k]; s += c; z += c != 0u; }
    uint32_t total, ztotal;
    block_exclusive_scan(s, total);
    block_exclusive_scan(z, ztotal);
    // did the block's entry count leave 32 bits?  (a 16-bit digest of the high part is enough: counts are < 2^31 each)
    uint32_t hi = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k)
        if (base + k < n) hi += count[base + k] >> 16;
    uint32_t hitotal;
    block_exclusive_scan(hi, hitotal);
    if (threadIdx.x == 0) {
        block_sums[blockIdx.x] = total;
        // the low halves of 1024 counts add up to < 2^26, so a total >= 2^32 needs the high halves to reach 2^16 - 2^10
        ne_sums[blockIdx.x] = ztotal | (hitotal >= 0xfc00u ? 0x80000000u : 0u);
    }
}

// Single block: in-place exclusive scan of block_sums[0..nb); block_sums[nb] = grand total.
__global__ void __launch_bounds__(SCAN_THREADS) scan_spine(uint32_t *block_sums, uint32_t *ne_sums, uint32_t *ne_count, int nb)
{
    if (threadIdx.x == 0) { block_sums[nb + 1] = 0u; block_sums[nb + 2] = 0u; }      // deep-tile list counters
    uint32_t carry = 0, zcarry = 0;
    bool wrapped = false;                                 // the 32-bit entry count of the call overflowed
    for (int base = 0; base < nb; base += SCAN_THREADS) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < nb ? block_sums[i] : 0u, zraw = i < nb ? ne_sums[i] : 0u;
        const uint32_t z = zraw & 0x7fffffffu;
        wrapped |= __syncthreads_or((int)(zraw >> 31)) != 0;
        uint32_t total, ztotal;
        const uint32_t ex = block_exclusive_scan(v, total);
        const uint32_t zex = block_exclusive_scan(z, ztotal);
        if (i < nb) { block_sums[i] = carry + ex; ne_sums[i] = zcarry + zex; }
        wrapped |= carry + total < carry;
        carry += total;
        zcarry += ztotal;
    }
    // D = 0xffffffff is the "too many entries for one call" sentinel: the host raises on it, the scatter kernel sets the
    // overflow flag (a wrapped scan would otherwise scatter into the wrong buckets without any sign)
    if (threadIdx.x == 0) { block_sums[nb] = (wrapped || carry == 0xffffffffu) ? 0xffffffffu : carry; *ne_count = zcarry; }
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_apply(const uint32_t *__restrict__ count, int64_t n,
                                                            const uint32_t *__restrict__ block_sums, int nb,
                                                            uint32_t *__restrict__ start, uint32_t *__restrict__ order,
                                                            const uint32_t *__restrict__ ne_sums)
{
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t c[SCAN_ITEMS];
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        c[k] = base + k < n ? count[base + k] : 0u;
        s += c[k];
    }
    uint32_t total;
    uint32_t ex = block_exclusive_scan(s, total) + block_sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        if (base + k < n) start[base + k] = ex;
        ex += c[k];
        // deep tiles, for the blend kernels' queues: >= AMAV_DEEP0 entries listed from the front, >= AMAV_DEEP1 from the back
        if (c[k] >= (uint32_t)AMAV_DEEP0) order[2 + atomicAdd(order, 1u)] = (uint32_t)(base + k);
        else if (c[k] >= (uint32_t)AMAV_DEEP1) order[2 + (n - 1 - atomicAdd(order + 1, 1u))] = (uint32_t)(base + k);
    }
    // the non-empty tiles (four in five are empty at 1024^2), compacted IN ASCENDING ORDER (a second scan, no atomics:
    // the blend walks this list front to back, and frame order is what keeps a frame's records in L2).  The blend
    // kernel's work queue holds only these; empty tiles just get the background.
    uint32_t *ne = order + 2 + n;                        // [0] = count (written by the spine kernel), list from [1]
    uint32_t z = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) z += c[k] != 0u;
    uint32_t ztotal;
    uint32_t zex = block_exclusive_scan(z, ztotal) + ne_sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k)
        if (c[k] != 0u) ne[1 + zex++] = (uint32_t)(base + k);
    if (blockIdx.x == 0 && threadIdx.x == 0) start[n] = block_sums[nb];
}

// CTA = SCAT_CHUNK consecutive Gaussians of one frame.  Three passes over the chunk's tile rectangles:
//   A. count entries per tile in a shared-memory histogram;
//   B. one global atomicAdd per non-empty (CTA, tile) reserves the chunk's range in the tile's bucket; the histogram
//      slot is overwritten with the absolute position of that range;
//   C. every entry takes its slot with a shared-memory atomicAdd on that position and is written out.
// The order inside a bucket is arbitrary (the per-tile sort fixes it).  ~20x fewer global atomics than one cursor
// bump per entry.  use_smem == 0: direct global cursor per entry (more tiles than fit in shared memory).
constexpr int SCAT_CHUNK = AMAV_SCAT_CHUNK;
constexpr int SCAT_PER = SCAT_CHUNK / 256;

__global__ void __launch_bounds__(256, AMAV_SCAT_CTAS) scatter_entries(int N, int gx, int tiles, const float4 *__restrict__ geom,
                                                        const ushort4 *__restrict__ rect,
                                                        const uint32_t *__restrict__ tile_start,
                                                        uint32_t *__restrict__ tile_cursor,
                                                        uint64_t *__restrict__ entries, int64_t capacity,
                                                        int32_t *overflow, const int use_smem)
{
    extern __shared__ uint32_t s_pos[];
    const int f = blockIdx.y;
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0 && tile_start[(int64_t)gridDim.y * tiles] == 0xffffffffu) *overflow = 1;
    const int64_t tbase = (int64_t)f * tiles;
    ushort4 rc[SCAT_PER];
    float depth[SCAT_PER];                                   // fetched up front: the loads overlap the histogram phases
#pragma unroll
    for (int k = 0; k < SCAT_PER; ++k) {
        const int i = blockIdx.x * SCAT_CHUNK + k * 256 + threadIdx.x;
        #if AMAV_SCAT_LDCS
        rc[k] = make_ushort4(0, 0, 0, 0);
        if (i < N) {
            const uint2 r2 = __ldcs(reinterpret_cast<const uint2 *>(rect) + (int64_t)f * N + i);
            rc[k] = make_ushort4((unsigned short)(r2.x & 0xffffu), (unsigned short)(r2.x >> 16), (unsigned short)(r2.y & 0xffffu),
                                 (unsigned short)(r2.y >> 16));
        }
#else
        rc[k] = i < N ? rect[(int64_t)f * N + i] : make_ushort4(0, 0, 0, 0);
#endif
    }
#pragma unroll
    for (int k = 0; k < SCAT_PER; ++k) {
        const int i = blockIdx.x * SCAT_CHUNK + k * 256 + threadIdx.x;
        #if AMAV_SCAT_LDCS
        depth[k] = (rc[k].z > rc[k].x && rc[k].w > rc[k].y) ? __ldcs(reinterpret_cast<const float *>(geom + 3 * ((int64_t)f * N + i) + 2) + 2) : 0.0f;
#else
        depth[k] = (rc[k].z > rc[k].x && rc[k].w > rc[k].y) ? geom[3 * ((int64_t)f * N + i) + 2].z : 0.0f;
#endif
    }
    if (use_smem) {
        for (int t = threadIdx.x; t < tiles; t += 256) s_pos[t] = 0u;
        __syncthreads();
        // one flat loop per rectangle (w * h trips): lanes whose rectangles differ in shape but not in area stay together
#pragma unroll
        for (int k = 0; k < SCAT_PER; ++k) {
            const int w = (int)rc[k].z - (int)rc[k].x, h = (int)rc[k].w - (int)rc[k].y;
            const int cnt = (w > 0 && h > 0) ? w * h : 0;
            int t = rc[k].y * gx + rc[k].x, x = 0;
            for (int e = 0; e < cnt; ++e) {
                atomicAdd(&s_pos[t], 1u);
                ++t;
                if (++x == w) { x = 0; t += gx - w; }
            }
        }
        __syncthreads();
        // four tiles per thread and step: the tile_start loads and the cursor atomics of a step are in flight together
        for (int t0 = threadIdx.x; t0 < tiles; t0 += 4 * 256) {
            uint32_t c[4], st[4], old[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int t = t0 + u * 256;
                c[u] = t < tiles ? s_pos[t] : 0u;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) st[u] = c[u] ? tile_start[tbase + t0 + u * 256] : 0u;
#pragma unroll
            for (int u = 0; u < 4; ++u) old[u] = c[u] ? atomicAdd(tile_cursor + tbase + t0 + u * 256, c[u]) : 0u;
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (c[u]) s_pos[t0 + u * 256] = st[u] + old[u];
        }
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < SCAT_PER; ++k) {
        if (rc[k].z <= rc[k].x || rc[k].w <= rc[k].y) continue;
        const uint32_t i = (uint32_t)(blockIdx.x * SCAT_CHUNK + k * 256 + threadIdx.x);
        const uint64_t e = ((uint64_t)__float_as_uint(depth[k]) << 32) | (uint64_t)i;
        const int w = (int)rc[k].z - (int)rc[k].x, cnt = w * ((int)rc[k].w - (int)rc[k].y);
        int t = rc[k].y * gx + rc[k].x, x = 0;
        for (int q = 0; q < cnt; ++q) {
            const int64_t pos = use_smem ? (int64_t)atomicAdd(&s_pos[t], 1u)
                                         : (int64_t)tile_start[tbase + t] + atomicAdd(tile_cursor + tbase + t, 1u);
            if (pos < capacity) entries[pos] = e;
            else *overflow = 1;
            ++t;
            if (++x == w) { x = 0; t += gx - w; }
        }
    }
}

// Ascending bitonic sort of n entries (virtual +inf padding to P = 2^logP): every compare-exchange moves the
// minimum to the lower index, so pairs with hi >= n are no-ops.  All index math is shifts and masks.
template <typename Ptr>
__device__ __forceinline__ void bitonic_sort(Ptr a, int n, int logP)
{
    const int half = 1 << (logP - 1);
    for (int lk = 1; lk <= logP; ++lk) {                 // merge size k = 2^lk
        const int lhk = lk - 1, hk = 1 << lhk;
        for (int i = threadIdx.x; i < half; i += blockDim.x) {            // flip step
            const int blk = i >> lhk, r = i & (hk - 1);
            const int lo = (blk << lk) + r, hi = (blk << lk) + (2 * hk - 1) - r;
            if (hi < n) {
                const uint64_t x = a[lo], y = a[hi];
                if (y < x) { a[lo] = y; a[hi] = x; }
            }
        }
        __syncthreads();
        for (int lj = lhk - 1; lj >= 0; --lj) {          // half-cleaners, stride j = 2^lj
            const int j = 1 << lj;
            for (int i = threadIdx.x; i < half; i += blockDim.x) {
                const int lo = ((i >> lj) << (lj + 1)) + (i & (j - 1)), hi = lo + j;
                if (hi < n) {
                    const uint64_t x = a[lo], y = a[hi];
                    if (y < x) { a[lo] = y; a[hi] = x; }
                }
            }
            __syncthreads();
        }
    }
}

// Bitonic sort of a bucket that does not fit the shared-memory tile (n > CH = 2^LC entries, in place in global memory
// g): the same network as bitonic_sort, but every compare-exchange step whose stride is below CH runs on CH-entry
// chunks staged in shared memory (padded with +inf keys, which the network leaves in place); only the
// O(log^2(n / CH)) steps with a stride of CH or more touch global memory.  A 32768-entry bucket makes 3 passes over
// global memory instead of 120.
template <int LC>
__device__ void bitonic_sort_large(uint64_t *__restrict__ g, uint64_t *s, int n, int logP)
{
    constexpr int CH = 1 << LC;
    const int nch = (n + CH - 1) >> LC;
    const int half = 1 << (logP - 1);
    auto load = [&](int c) {
        const int base = c << LC;
        for (int i = threadIdx.x; i < CH; i += blockDim.x) s[i] = base + i < n ? g[base + i] : ~0ull;
        __syncthreads();
    };
    auto store = [&](int c) {
        const int base = c << LC;
        for (int i = threadIdx.x; i < CH; i += blockDim.x)
            if (base + i < n) g[base + i] = s[i];
        __syncthreads();
    };
    for (int c = 0; c < nch; ++c) {                      // merges up to size CH: each chunk sorted on its own
        load(c);
        bitonic_sort(s, CH, LC);
        store(c);
    }
    for (int lk = LC + 1; lk <= logP; ++lk) {            // merge size k = 2^lk > CH
        const int lhk = lk - 1, hk = 1 << lhk;
        for (int i = threadIdx.x; i < half; i += blockDim.x) {            // flip step, across chunks
            const int blk = i >> lhk, r = i & (hk - 1);
            const int lo = (blk << lk) + r, hi = (blk << lk) + (2 * hk - 1) - r;
            if (hi < n) {
                const uint64_t x = g[lo], y = g[hi];
                if (y < x) { g[lo] = y; g[hi] = x; }
            }
        }
        __syncthreads();
        for (int lj = lhk - 1; lj >= LC; --lj) {         // half-cleaners whose stride spans chunks
            const int j = 1 << lj;
            for (int i = threadIdx.x; i < half; i += blockDim.x) {
                const int lo = ((i >> lj) << (lj + 1)) + (i & (j - 1)), hi = lo + j;
                if (hi < n) {
                    const uint64_t x = g[lo], y = g[hi];
                    if (y < x) { g[lo] = y; g[hi] = x; }
                }
            }
            __syncthreads();
        }
        for (int c = 0; c < nch; ++c) {                  // the remaining half-cleaners (strides CH/2 .. 1) per chunk
            load(c);
            for (int lj = LC - 1; lj >= 0; --lj) {
                const int j = 1 << lj;
                for (int i = threadIdx.x; i < CH / 2; i += blockDim.x) {
                    const int lo = ((i >> lj) << (lj + 1)) + (i & (j - 1)), hi = lo + j;
                    const uint64_t x = s[lo], y = s[hi];
                    if (y < x) { s[lo] = y; s[hi] = x; }
                }
                __syncthreads();
            }
            store(c);
        }
    }
}

// Two instantiations of the per-tile sort share one body:
//   sort_tiles<256, 4096, 16>    buckets of up to 4096 entries: 48 KB of smem, 4 CTAs per SM;
//   sort_tiles<1024, 16384, 64>  buckets of 4097..16384 entries (512x512 training views, 500k-Gaussian close-ups):
//                                192 KB of smem, one CTA per SM; larger buckets still: bitonic network in place in
//                                global memory with every stride below 16384 done on chunks in shared memory.
constexpr int SORT_SMALL_CAP = 4096;
#if !AMAV_SORT_GROUPS
constexpr int SORT_SMALL_T = 256, SORT_SMALL_POP = 16;
#endif
constexpr int SORT_LARGE_T = 1024, SORT_LARGE_CAP = 16384, SORT_LARGE_POP = 64;
constexpr int SORT_MAX_BIN = 48;                     // a depth bin fuller than this sends the tile to the bitonic network

// One tile's bucket (n <= T * PER entries in s_ent) sorted in O(n): entries are spread over B >= n depth bins
// (bin = monotone linear map of the depth bits between the bucket's min and max), a scan of the bin counts gives
// each bin its output range, and every entry's final position is its bin's start plus its rank among the (about
// one) entries of the bin on the full 64-bit (depth bits, Gaussian) key.  Returns false (nothing guaranteed about s_ent's order) when the depths
// are too clustered for that to be cheap; the caller then runs the bitonic network.
template <int T, int PER>
__device__ bool bin_sort_tile(uint64_t *s_ent, uint32_t *s_bin, int n, int logB, int32_t *__restrict__ out)
{
    __shared__ uint32_t s_red[2 * (T / 32)];
    __shared__ uint32_t s_wsum[T / 32];
    __shared__ int s_bad;
    const int B = 1 << logB;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t lo = 0xffffffffu, hi = 0u;
    for (int i = threadIdx.x; i < n; i += T) {
        const uint32_t d = (uint32_t)(s_ent[i] >> 32);
        lo = min(lo, d); hi = max(hi, d);
    }
    lo = __reduce_min_sync(0xffffffffu, lo);
    hi = __reduce_max_sync(0xffffffffu, hi);
    if (lane == 0) { s_red[warp] = lo; s_red[T / 32 + warp] = hi; }
    for (int i = threadIdx.x; i < B; i += T) s_bin[i] = 0u;
    if (threadIdx.x == 0) s_bad = 0;
    __syncthreads();
#pragma unroll
    for (int w = 0; w < T / 32; ++w) { lo = min(lo, s_red[w]); hi = max(hi, s_red[T / 32 + w]); }
    const float scale = (float)B / ((float)(hi - lo) + 1.0f);
    // pass 1: bin of every entry (kept in registers) + bin counts
    uint64_t ent[PER];
    int bin[PER];
    const int slots = (n + T - 1) / T;                    // slots in use (uniform): the unrolled loops stop there
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        if (k >= slots) break;
        const int i = threadIdx.x + k * T;
        if (i < n) {
            ent[k] = s_ent[i];
            bin[k] = min(B - 1, (int)((float)((uint32_t)(ent[k] >> 32) - lo) * scale));
            atomicAdd(&s_bin[bin[k]], 1u);
        }
    }
    __syncthreads();
    // exclusive scan of the B bin counts (B / T consecutive bins per thread)
    const int per = B / T;                                // B >= T is guaranteed by the caller
    uint32_t sum = 0, mx = 0;
    for (int q = 0; q < per; ++q) { const uint32_t c = s_bin[threadIdx.x * per + q]; sum += c; mx = max(mx, c); }
    if (mx > SORT_MAX_BIN) s_bad = 1;
    const uint32_t inc = warp_inclusive_scan(sum);
    if (lane == 31) s_wsum[warp] = inc;
    __syncthreads();
    if (s_bad) return false;
    uint32_t base = inc - sum;
#pragma unroll
    for (int w = 0; w < T / 32; ++w)
        if (w < warp) base += s_wsum[w];
    for (int q = 0; q < per; ++q) {
        const uint32_t c = s_bin[threadIdx.x * per + q];
        s_bin[threadIdx.x * per + q] = base;
        base += c;
    }
    __syncthreads();
    // pass 2: scatter (slot inside the bin from an atomic cursor: any order, the bin is sorted next)
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        if (k >= slots) break;
        const int i = threadIdx.x + k * T;
        if (i < n) s_ent[atomicAdd(&s_bin[bin[k]], 1u)] = ent[k];
    }
    __syncthreads();
    // pass 3: s_bin[b] is now the END of bin b.  Every entry finds its final position = bin start + its rank among
    // the (about one, at most SORT_MAX_BIN) entries of its bin, and its Gaussian index goes straight to the output.
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        if (k >= slots) break;
        const int i = threadIdx.x + k * T;
        if (i < n) {
            const int e0 = bin[k] == 0 ? 0 : (int)s_bin[bin[k] - 1], e1 = (int)s_bin[bin[k]];
            int rank = e0;
            for (int q = e0; q < e1; ++q) rank += s_ent[q] < ent[k];
            out[rank] = (int32_t)(uint32_t)ent[k];
        }
    }
    return true;
}

// CTAs pull POP tiles at a time from a device counter (four tiles in five are empty: the queue pop, its two barriers
// and the tile_start loads are amortised over the batch).  A bucket of MINN < n <= CAP entries is sorted in shared
// memory: rank sort for n <= 64, the O(n) depth-bin sort otherwise, the bitonic network when the depths are too
// clustered for the bins; with GLOBAL_FALLBACK buckets above CAP are sorted in place in global memory.
template <int T, int CAP, int POP, int MINN, bool GLOBAL_FALLBACK>
__global__ void __launch_bounds__(T, T == 256 ? 4 : 1) sort_tiles(unsigned int num_tiles_total,
                                                                  const uint32_t *__restrict__ tile_start,
                                                                  uint64_t *__restrict__ entries,
                                                                  int32_t *__restrict__ point_list, int64_t capacity,
                                                                  unsigned int *work_counter)
{
    extern __shared__ uint64_t s_sort[];
    uint64_t *s_ent = s_sort;                                            // [CAP]
    uint32_t *s_bin = reinterpret_cast<uint32_t *>(s_sort + CAP);        // [CAP]
    __shared__ unsigned int s_next;
    __shared__ uint32_t s_ts[POP + 1];
    __shared__ unsigned int s_mask[2];
    static_assert(POP <= 64, "one mask word per 32 popped tiles");
    const int tid = threadIdx.x;
    for (;;) {
        __syncthreads();
        if (tid == 0) s_next = atomicAdd(work_counter, (unsigned int)POP);
        __syncthreads();
        const unsigned int t0 = s_next;
        if (t0 >= num_tiles_total) break;
        const unsigned int nt = min((unsigned int)POP, num_tiles_total - t0);
        if (tid <= (int)nt) s_ts[tid] = tile_start[t0 + tid];
        __syncthreads();
        // which of the popped tiles this instantiation sorts (four in five are empty; the large-bucket instantiation
        // finds none in most batches): one lane per tile, then the CTA walks the set bits only
        if (tid < 64) {
            bool mine = false;
            if (tid < (int)nt) {
                int64_t e = s_ts[tid + 1];
                if (e > capacity) e = capacity;
                const int64_t n = e - (int64_t)s_ts[tid];
                mine = n > MINN && n >= 1 && (GLOBAL_FALLBACK || n <= CAP);
            }
            const unsigned int m = __ballot_sync(0xffffffffu, mine);
            if ((tid & 31) == 0) s_mask[tid >> 5] = m;
        }
        __syncthreads();
        for (int w = 0; w < (POP + 31) / 32; ++w)
        for (unsigned int todo = s_mask[w]; todo != 0u; todo &= todo - 1u) {
            const unsigned int q = 32u * w + (unsigned int)(__ffs(todo) - 1);
            const int64_t s = s_ts[q];
            int64_t e = s_ts[q + 1];
            if (e > capacity) e = capacity;
            const int n = (int)(e - s);
            if (n == 1) {
                if (tid == 0) point_list[s] = (int32_t)(uint32_t)entries[s];
                continue;
            }
            const int logP = max(1, 32 - __clz(n - 1));            // smallest power of two >= n
            __syncthreads();                                       // the previous tile is done with shared memory
            if (n > CAP) {
                uint64_t *g = entries + s;
                static_assert((CAP & (CAP - 1)) == 0, "CAP must be a power of two");
                bitonic_sort_large<31 - __builtin_clz(CAP)>(g, s_ent, n, logP);
                for (int i = tid; i < n; i += T) point_list[s + i] = (int32_t)(uint32_t)g[i];
                continue;
            }
            for (int i = tid; i < n; i += T) s_ent[i] = entries[s + i];
            __syncthreads();
            if (n <= 64) {                                         // rank sort (keys are distinct)
                if (tid < n) {
                    const uint64_t x = s_ent[tid];
                    int rank = 0;
                    for (int u = 0; u < n; ++u) rank += s_ent[u] < x;
                    point_list[s + rank] = (int32_t)(uint32_t)x;
                }
                continue;
            }
            const int logB = max(logP, 31 - __clz(T));
            if (!bin_sort_tile<T, CAP / T>(s_ent, s_bin, n, logB, point_list + s)) {
                __syncthreads();
                bitonic_sort(s_ent, n, logP);
                for (int i = tid; i < n; i += T) point_list[s + i] = (int32_t)(uint32_t)s_ent[i];
            }
        }
    }
}


// ---------------------------------------------------------------------------------------------------------------------
// Grouped small-bucket sort.  The buckets of consecutive tiles are adjacent in `entries` (tile_start is a scan), so a
// CTA sorts a GROUP of consecutive tiles whose buckets add up to <= CAP entries in ONE pass of the depth-bin sort: the
// barriers, the bin scan and the queue pop are paid once per ~CAP entries instead of once per tile (mean bucket: 665
// entries at 1024^2, a fifth of the tiles non-empty), every thread has CAP / T entries in flight per phase, and tiny
// buckets ride along with their neighbours.  Tile j of the group owns the bins [start_j - start_group, +n_j) -- as many
// bins as entries -- and maps its depths linearly between its own min and max; one scan over the group's bins then
// gives every entry's bin its final range in the group, and the rank among the (about one) entries of the bin on the
// full 64-bit (depth bits, Gaussian) key finishes the order.  Entries stay in registers from the global load to the
// scatter; the slot inside the bin is the return value of the counting atomic (no second atomic pass).
// A group with a bin fuller than SORT_MAX_BIN (depths too clustered) is redone tile by tile on the bitonic network.
#ifndef AMAV_GRP_POP
#define AMAV_GRP_POP 32
#endif
constexpr int GRP_T = 256, GRP_CAP = 4096, GRP_POP = AMAV_GRP_POP, GRP_PER = GRP_CAP / GRP_T;

__global__ void __launch_bounds__(GRP_T, 4) sort_groups(unsigned int num_tiles_total, const uint32_t *__restrict__ tile_start,
                                                        const uint64_t *__restrict__ entries,
                                                        int32_t *__restrict__ point_list, int64_t capacity,
                                                        unsigned int *work_counter)
{
    constexpr int T = GRP_T, CAP = GRP_CAP, POP = GRP_POP, PER = GRP_PER;
    extern __shared__ uint64_t s_sort[];
    uint64_t *s_ent = s_sort;                                            // [CAP]
    uint32_t *s_bin = reinterpret_cast<uint32_t *>(s_sort + CAP);        // [CAP + T + 1]: counts, then exclusive starts
    __shared__ unsigned int s_next;
    __shared__ uint32_t s_ts[POP + 1];
    __shared__ uint32_t s_lo[POP], s_hi[POP];
    __shared__ float s_scale[POP];
    __shared__ uint32_t s_wsum[T / 32];
    __shared__ int s_bad;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (;;) {
        __syncthreads();
        if (tid == 0) s_next = atomicAdd(work_counter, (unsigned int)POP);
        __syncthreads();
        const unsigned int t0 = s_next;
        if (t0 >= num_tiles_total) break;
        const int nt = (int)min((unsigned int)POP, num_tiles_total - t0);
        if (tid <= nt) {
            const int64_t v = tile_start[t0 + tid];
            s_ts[tid] = (uint32_t)(v > capacity ? capacity : v);
        }
        __syncthreads();
        int a = 0;
        while (a < nt) {                                          // uniform: every thread forms the same groups
            const uint32_t sa = s_ts[a];
            int b = a + 1;
            const uint32_t n0 = s_ts[b] - sa;
            if (n0 == 0u || n0 > (uint32_t)CAP) { a = b; continue; }   // empty, or the large-bucket kernel's
            while (b < nt && s_ts[b + 1] - sa <= (uint32_t)CAP) ++b;
            const int n = (int)(s_ts[b] - sa);
            const int ga = a;
            a = b;
            if (n == 1) {
                if (tid == 0) point_list[sa] = (int32_t)(uint32_t)entries[sa];
                continue;
            }
            const int slots = (n + T - 1) / T;                   // uniform
            const int per = ((n + 1 + T - 1) / T) | 1;           // bins per thread in the scan (odd: conflict-free stride)
            __syncthreads();                                     // the previous group is done with shared memory
            uint64_t ent[PER];
            uint32_t meta[PER];
#pragma unroll
            for (int k = 0; k < PER; ++k) {
                if (k >= slots) break;
                const int i = tid + k * T;
                ent[k] = i < n ? entries[(int64_t)sa + i] : 0ull;
            }
            for (int i = tid; i < T * per; i += T) s_bin[i] = 0u;
            if (tid < b - ga) { s_lo[ga + tid] = 0xffffffffu; s_hi[ga + tid] = 0u; }
            if (tid == 0) s_bad = 0;
            __syncthreads();
            // tile of every entry (a forward walk: a thread's entries are T apart) and the tiles' depth ranges
            {
                int j = ga;
#pragma unroll
                for (int k = 0; k < PER; ++k) {
                    if (k >= slots) break;
                    const int i = tid + k * T;
                    const bool valid = i < n;
                    if (valid) {
                        const uint32_t pos = sa + (uint32_t)i;
                        while (pos >= s_ts[j + 1]) ++j;
                    }
                    const uint32_t d = (uint32_t)(ent[k] >> 32);
                    const int j0 = __shfl_sync(0xffffffffu, j, 0);
                    const bool v0 = __shfl_sync(0xffffffffu, (int)valid, 0) != 0;      // lane 0 holds the warp's lowest i
                    if (v0) {
                        if (__all_sync(0xffffffffu, !valid || j == j0)) {
                            const uint32_t lo = __reduce_min_sync(0xffffffffu, valid ? d : 0xffffffffu);
                            const uint32_t hi = __reduce_max_sync(0xffffffffu, valid ? d : 0u);
                            if (lane == 0) { atomicMin(&s_lo[j0], lo); atomicMax(&s_hi[j0], hi); }
                        } else if (valid) {
                            atomicMin(&s_lo[j], d); atomicMax(&s_hi[j], d);
                        }
                    }
                    meta[k] = (uint32_t)j;
                }
            }
            __syncthreads();
            if (tid < b - ga) {
                const int j = ga + tid;
                s_scale[j] = (float)(s_ts[j + 1] - s_ts[j]) / ((float)(s_hi[j] - s_lo[j]) + 1.0f);
            }
            __syncthreads();
            // pass 1: bin of every entry; the counting atomic's return value is the entry's slot inside the bin
#pragma unroll
            for (int k = 0; k < PER; ++k) {
                if (k >= slots) break;
                const int i = tid + k * T;
                if (i < n) {
                    const int j = (int)meta[k];
                    const uint32_t tsj = s_ts[j];
                    const int nj = (int)(s_ts[j + 1] - tsj);
                    const int bl = min(nj - 1, (int)((float)((uint32_t)(ent[k] >> 32) - s_lo[j]) * s_scale[j]));
                    const uint32_t bin = (tsj - sa) + (uint32_t)bl;
                    const uint32_t slot = atomicAdd(&s_bin[bin], 1u);
                    meta[k] = bin | (slot << 13);
                }
            }
            __syncthreads();
            // exclusive scan of the n + 1 bin counts (`per` consecutive bins per thread; bins past n hold zero)
            {
                uint32_t sum = 0, mx = 0;
                for (int q = 0; q < per; ++q) { const uint32_t c = s_bin[tid * per + q]; sum += c; mx = max(mx, c); }
                if (mx > SORT_MAX_BIN) s_bad = 1;
                const uint32_t inc = warp_inclusive_scan(sum);
                if (lane == 31) s_wsum[warp] = inc;
                __syncthreads();
                uint32_t base = inc - sum;
#pragma unroll
                for (int w = 0; w < T / 32; ++w)
                    if (w < warp) base += s_wsum[w];
                for (int q = 0; q < per; ++q) {
                    const uint32_t c = s_bin[tid * per + q];
                    s_bin[tid * per + q] = base;
                    base += c;
                }
            }
            __syncthreads();
            if (s_bad) {                                         // clustered depths: tile by tile on the bitonic network
                for (int j = ga; j < b; ++j) {
                    const uint32_t sj = s_ts[j];
                    const int nj = (int)(s_ts[j + 1] - sj);
                    if (nj == 0) continue;
                    __syncthreads();
                    for (int i = tid; i < nj; i += T) s_ent[i] = entries[(int64_t)sj + i];
                    __syncthreads();
                    if (nj > 1) bitonic_sort(s_ent, nj, max(1, 32 - __clz(nj - 1)));
                    for (int i = tid; i < nj; i += T) point_list[(int64_t)sj + i] = (int32_t)(uint32_t)s_ent[i];
                }
                continue;
            }
            // pass 2: scatter into the bins
#pragma unroll
            for (int k = 0; k < PER; ++k) {
                if (k >= slots) break;
                const int i = tid + k * T;
                if (i < n) s_ent[s_bin[meta[k] & 0x1fffu] + (meta[k] >> 13)] = ent[k];
            }
            __syncthreads();
            // pass 3: final position = bin start + rank among the entries of the bin; the Gaussian index goes out
            int32_t *out = point_list + (int64_t)sa;
#pragma unroll
            for (int k = 0; k < PER; ++k) {
                if (k >= slots) break;
                const int i = tid + k * T;
                if (i < n) {
                    const uint32_t bin = meta[k] & 0x1fffu;
                    const int e0 = (int)s_bin[bin], e1 = (int)s_bin[bin + 1];
                    int rank = e0;
                    if (e1 - e0 > 1)
                        for (int q = e0; q < e1; ++q) rank += s_ent[q] < ent[k];
                    out[rank] = (int32_t)(uint32_t)ent[k];
                }
            }
        }
    }
}

}  // namespace

extern "C" int64_t amav_splat_scan_tmp_elems(int64_t num_tiles_total)
{
    const int64_t nb = (num_tiles_total + SCAN_TILE - 1) / SCAN_TILE;
    return nb + 1 + 2 + num_tiles_total + 1 + num_tiles_total + nb;
}

extern "C" int64_t amav_splat_scan_order_offset(int64_t num_tiles_total)
{
    return (num_tiles_total + SCAN_TILE - 1) / SCAN_TILE + 1;
}

extern "C" int amav_splat_scan(const AmavSplatDims *d, const uint32_t *tile_count, uint32_t *tile_start,
                               uint32_t *scan_tmp, void *stream_)
{
    AMAV_CHECK_ARG(d && d->frames >= 0 && d->width > 0 && d->height > 0, "bad dims");
    AMAV_CHECK_ARG(tile_count && tile_start && scan_tmp, "null buffer");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int gx = (d->width + 15) / 16, gy = (d->height + 15) / 16;
    const int64_t nt = (int64_t)d->frames * gx * gy;
    AMAV_CHECK_ARG(nt < (int64_t)1 << 31, "too many tiles");
    if (nt == 0) return AMAV_OK;
    const int nb = amav_div_up(nt, SCAN_TILE);
    uint32_t *order = scan_tmp + nb + 1;                 // [2 counters][nt deep-tile slots][1 count][nt non-empty tiles]
    uint32_t *ne_sums = order + 2 + nt + 1 + nt;         // [nb]
    scan_block_sums<<<nb, SCAN_THREADS, 0, stream>>>(tile_count, nt, scan_tmp, ne_sums);
    scan_spine<<<1, SCAN_THREADS, 0, stream>>>(scan_tmp, ne_sums, order + 2 + nt, nb);
    scan_apply<<<nb, SCAN_THREADS, 0, stream>>>(tile_count, nt, scan_tmp, nb, tile_start, order, ne_sums);
    amav_count_launches(2);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}

extern "C" int amav_splat_bin(const AmavSplatDims *d, const float *geom, const uint16_t *rect,
                              const uint32_t *tile_start, uint32_t *tile_cursor, uint64_t *entries,
                              int32_t *point_list, int64_t capacity, int32_t *overflow, void *stream_)
{
    AMAV_CHECK_ARG(d && d->frames >= 0 && d->num_gaussians >= 0, "bad dims");
    AMAV_CHECK_ARG(tile_start && tile_cursor && overflow, "null buffer");
    AMAV_CHECK_ARG((int64_t)d->frames * ((d->width + 15) / 16) * ((d->height + 15) / 16) < ((int64_t)1 << 31), "too many tiles");
    AMAV_CHECK_ARG(capacity >= 0 && (capacity == 0 || (entries && point_list)), "null entries");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int gx = (d->width + 15) / 16, gy = (d->height + 15) / 16;
    const int tiles = gx * gy;
    const int64_t nt = (int64_t)d->frames * tiles;
    const int64_t total = (int64_t)d->frames * d->num_gaussians;
    if (nt == 0 || total == 0) return AMAV_OK;
    AMAV_CHECK_ARG(geom && rect, "null geometry");
    AMAV_CHECK_ARG(d->frames <= 65535, "too many frames for one call");
    const size_t hist = (size_t)tiles * sizeof(uint32_t);
    const int use_smem = hist <= 160 * 1024;
    if (use_smem) amav_ensure_dyn_smem((const void *)scatter_entries, (int)hist);
    dim3 sgrid(amav_div_up(d->num_gaussians, SCAT_CHUNK), d->frames);
    scatter_entries<<<sgrid, 256, use_smem ? hist : 0, stream>>>(
        d->num_gaussians, gx, tiles, reinterpret_cast<const float4 *>(geom), reinterpret_cast<const ushort4 *>(rect),
        tile_start, tile_cursor, entries, capacity, overflow, use_smem);
    AMAV_CHECK_LAUNCH();
    // small buckets: queue head = overflow[1]; large buckets: queue head = tile_cursor[0], free once the scatter is done
    auto large = sort_tiles<SORT_LARGE_T, SORT_LARGE_CAP, SORT_LARGE_POP, SORT_SMALL_CAP, true>;
    const int large_smem = SORT_LARGE_CAP * 12;
    amav_ensure_dyn_smem((const void *)large, large_smem);
    unsigned int *small_queue = reinterpret_cast<unsigned int *>(overflow) + 1;
#if AMAV_SORT_GROUPS
    static_assert(GRP_CAP == SORT_SMALL_CAP, "the large-bucket kernel takes every bucket above the grouped sort's capacity");
    const int64_t want = (nt + GRP_POP - 1) / GRP_POP;
    const int small_smem = GRP_CAP * 8 + (GRP_CAP + GRP_T + 1) * 4;
    amav_ensure_dyn_smem((const void *)sort_groups, small_smem);
    sort_groups<<<(int)(want < AMAV_SM_COUNT * 4 ? want : AMAV_SM_COUNT * 4), GRP_T, small_smem, stream>>>(
        (unsigned int)nt, tile_start, entries, point_list, capacity, small_queue);
#else
    auto small = sort_tiles<SORT_SMALL_T, SORT_SMALL_CAP, SORT_SMALL_POP, 0, false>;
    const int64_t want = (nt + SORT_SMALL_POP - 1) / SORT_SMALL_POP;
    const int small_smem = SORT_SMALL_CAP * 12;
    amav_ensure_dyn_smem((const void *)small, small_smem);
    small<<<(int)(want < AMAV_SM_COUNT * 4 ? want : AMAV_SM_COUNT * 4), SORT_SMALL_T, small_smem, stream>>>(
        (unsigned int)nt, tile_start, entries, point_list, capacity, small_queue);
#endif
    AMAV_CHECK_LAUNCH();
    if (cudaMemsetAsync(tile_cursor, 0, sizeof(uint32_t), stream) != cudaSuccess) {
        amav_set_error("%s: cudaMemsetAsync failed", __func__);
        return AMAV_ECUDA;
    }
    const int64_t wantl = (nt + SORT_LARGE_POP - 1) / SORT_LARGE_POP;
    large<<<(int)(wantl < AMAV_SM_COUNT ? wantl : AMAV_SM_COUNT), SORT_LARGE_T, large_smem, stream>>>(
        (unsigned int)nt, tile_start, entries, point_list, capacity, tile_cursor);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
