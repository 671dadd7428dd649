// gut_binning.cu -- tile binning: bucket scan, key emission, per-tile radix sort, record expansion.
// Built with --fmad=false (the emit pass must repeat K1's culling decisions bit-for-bit).
//
// What the reference does (src/gutRenderer.cu:297-373): inclusive scan over particles (CUB) -> host
// read-back of M -> expandTileProjections writes M (tile<<32|depth, particle) pairs -> CUB
// DeviceRadixSort over 32+log2(T) key bits (6 passes over M x 12 B in HBM) -> a kernel finds the tile
// ranges.  Here the same ordering is produced as a hybrid MSD/LSD radix sort:
//   MSD digit = tile id.  K1 already histogrammed the tiles, so a T-element scan gives every tile's bucket
//                and the emit pass drops (depth, particle) pairs straight into their bucket (one 8-byte
//                store per intersection; no per-particle scan, no key array in HBM);
//   LSD       = one CTA per tile sorts its bucket on (depth bits, particle index) in shared memory with
//                8-bit stable radix passes, skipping digits that are constant over the bucket.
// The particle index is part of the sort key because buckets are filled by atomics in arbitrary order;
// the reference's stable sort leaves equal (tile, depth) keys in emission order = ascending particle
// index, which (depth, index) ordering reproduces exactly.  Tile ranges fall out of the bucket scan.
#include "gut_internal.h"

namespace {

constexpr int SCAN_THREADS = 1024;

// ---------------------------------------------------------------------------------------------------
// Exclusive scan of the tile histogram (T is a few thousand: one CTA, chunks of 1024 tiles with a carry).
// hist holds GUT_TILE_SUB sub-counts per tile; start[t] = first slot of tile t's bucket, cursor[GUT_SUB_SLOT(t,k,T)] = first
// slot of sub-bucket k (absolute positions: the emit pass appends with one atomicAdd and no further look-up).
// start has 2T+1 entries: start[0..T] the bucket offsets, start[T+1 .. 2T] the heavy-first tile order (see the end).
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(SCAN_THREADS) gut_tile_scan_kernel(uint32_t T, const uint32_t* __restrict__ hist,
                                                                    uint32_t* __restrict__ start, uint32_t* __restrict__ cursor,
                                                                    uint32_t* __restrict__ stats) {
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry_s;
    __shared__ uint32_t max_s;
    const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) { carry_s = 0; max_s = 0; }
    __syncthreads();
    uint32_t local_max = 0;
    for (uint32_t base = 0; base < T; base += SCAN_THREADS) {
        const uint32_t i = base + tid;
        uint32_t c[GUT_TILE_SUB];
#pragma unroll
        for (int k = 0; k < GUT_TILE_SUB; ++k) c[k] = 0;
        if (i < T) {
#pragma unroll
            for (int k = 0; k < GUT_TILE_SUB; ++k) c[k] = hist[GUT_SUB_SLOT(i, k, T)];
        }
        uint32_t v = 0;
#pragma unroll
        for (int k = 0; k < GUT_TILE_SUB; ++k) v += c[k];
        local_max = max(local_max, v);
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (uint32_t)o) incl += n;
        }
        if (lane == 31) warp_sums[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            uint32_t ws = warp_sums[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t n = __shfl_up_sync(0xffffffffu, ws, o);
                if (lane >= (uint32_t)o) ws += n;
            }
            warp_sums[lane] = ws;  // inclusive over warps
        }
        __syncthreads();
        const uint32_t carry = carry_s;
        const uint32_t excl = carry + (wid ? warp_sums[wid - 1] : 0u) + incl - v;
        if (i < T) {
            start[i] = excl;
            uint32_t run = excl;
            uint32_t o[GUT_TILE_SUB];
#pragma unroll
            for (int k = 0; k < GUT_TILE_SUB; ++k) { o[k] = run; run += c[k]; }
#pragma unroll
            for (int k = 0; k < GUT_TILE_SUB; ++k) cursor[GUT_SUB_SLOT(i, k, T)] = o[k];
        }
        __syncthreads();
        if (tid == SCAN_THREADS - 1) carry_s = carry + warp_sums[31];
        __syncthreads();
    }
    atomicMax(&max_s, local_max);
    __syncthreads();
    if (tid == 0) { start[T] = carry_s; stats[0] = carry_s; stats[1] = max_s; stats[2] = 0u; /* buckets handed to the radix sort */ }

    // Heavy-first launch order for the compositing kernels: order[T] (stored behind start[T+1]) lists the tiles by descending
    // list length, as a counting sort on SCAN_THREADS length classes.  One-warp CTAs of the compositing grids live for up to
    // a millisecond; dispatched in image order the long ones of the image centre start mid-way and the grid drains with the
    // SMs a fifth empty on average (ncu r02a: 17.9 of 21 resident warps in K5).  Longest-first lets the short ones fill the end.
    __shared__ uint32_t cls_off[SCAN_THREADS];
    uint32_t* order = start + T + 1;
    const uint32_t mx = max(max_s, 1u);
    cls_off[tid] = 0;
    __syncthreads();
    for (uint32_t i = tid; i < T; i += SCAN_THREADS) {
        const uint32_t len = start[i + 1] - start[i];
        atomicAdd(&cls_off[(SCAN_THREADS - 1) - (uint32_t)(((uint64_t)len * (SCAN_THREADS - 1)) / mx)], 1u);
    }
    __syncthreads();
    {   // exclusive scan over the classes (class 0 = longest)
        const uint32_t c = cls_off[tid];
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (uint32_t)o) incl += n;
        }
        if (lane == 31) warp_sums[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            uint32_t ws = warp_sums[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t n = __shfl_up_sync(0xffffffffu, ws, o);
                if (lane >= (uint32_t)o) ws += n;
            }
            warp_sums[lane] = ws;
        }
        __syncthreads();
        cls_off[tid] = (wid ? warp_sums[wid - 1] : 0u) + incl - c;
    }
    __syncthreads();
    for (uint32_t i = tid; i < T; i += SCAN_THREADS) {
        const uint32_t len = start[i + 1] - start[i];
        const uint32_t pos = atomicAdd(&cls_off[(SCAN_THREADS - 1) - (uint32_t)(((uint64_t)len * (SCAN_THREADS - 1)) / mx)], 1u);
        order[pos] = i;
    }
}

// ---------------------------------------------------------------------------------------------------
// K2: emit.  One thread per particle appends (depth bits, particle) to the bucket of every tile that survived K1's culling
// test.  reference: GUTProjector::expand, gutProjector.cuh:324-388 (which re-walks the box and repeats the tests).
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, 8) gut_emit_kernel(GutConfig cfg, GutFrame fr, uint32_t N, GutProjBuffers proj,
                                                       uint32_t* __restrict__ tile_cursor, uint2* __restrict__ entries, uint32_t capacity) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    // K1 left the culled tile box as a bit mask (gut_math.cuh, GUT_EMIT_*): one 16-byte load, no tests to repeat.
    const uint4 rec = __ldg(proj.emit_rec + i);
    const uint32_t depth_bits = rec.w;
    const uint32_t sub = i & (GUT_TILE_SUB - 1);
    // The bucket slot comes back from an L2 atomic (~1 us under contention): EMIT_ILP independent atomics are kept in
    // flight per thread instead of one dependent round trip per tile.
    constexpr int EMIT_ILP = 4;
    if (!(rec.z & GUT_EMIT_BIG)) {
        uint64_t mask = ((uint64_t)rec.y << 32) | rec.x;
        if (mask == 0ull) return;
        const uint32_t x0 = rec.z & 0xfffu, y0 = (rec.z >> 12) & 0xfffu, bw = (rec.z >> 24) & 0x7fu;
        const uint32_t inv = 65536u / bw + 1u;   // row = (bit * inv) >> 16, exact for bit < 64, bw <= 64
        while (mask) {
            uint32_t t[EMIT_ILP], slot[EMIT_ILP];
            bool ok[EMIT_ILP];
#pragma unroll
            for (int k = 0; k < EMIT_ILP; ++k) {
                ok[k] = mask != 0ull;
                const uint32_t bit = ok[k] ? (uint32_t)__ffsll((long long)mask) - 1u : 0u;
                mask &= mask - 1ull;
                const uint32_t row = (bit * inv) >> 16;
                t[k] = (y0 + row) * fr.grid_x + x0 + (bit - row * bw);
            }
#pragma unroll
            for (int k = 0; k < EMIT_ILP; ++k) slot[k] = ok[k] ? atomicAdd(&tile_cursor[GUT_SUB_SLOT(t[k], sub, fr.num_tiles)], 1u) : 0u;
#pragma unroll
            for (int k = 0; k < EMIT_ILP; ++k)
                if (ok[k] && slot[k] < capacity) entries[slot[k]] = make_uint2(depth_bits, i);
        }
        return;
    }
    // box above 64 tiles: walk it with the same culling test as K1 (reference: GUTProjector::expand, gutProjector.cuh:324-388)
    const float2 ext = proj.extent[i];
    const float2 centre = proj.proj_pos[i];
    const float4 co = proj.conic_opacity[i];
    const float max_power = gut_logf(co.w / cfg.min_alpha);
    const GutTileBox bb = gut_tile_bbox(fr.grid_x, fr.grid_y, centre, ext);
    const uint32_t bw = bb.x1 - bb.x0;
    const uint32_t total = bw * (bb.y1 - bb.y0);
    uint32_t x = bb.x0, y = bb.y0;
    for (uint32_t base = 0; base < total; base += EMIT_ILP) {
        uint32_t t[EMIT_ILP], slot[EMIT_ILP];
        bool ok[EMIT_ILP];
#pragma unroll
        for (int k = 0; k < EMIT_ILP; ++k) {
            ok[k] = (base + k) < total;
            t[k] = y * fr.grid_x + x;
            if (ok[k] && cfg.tile_based_culling) ok[k] = gut_tile_min_power((float)x, (float)y, co, centre) < max_power;
            if (++x == bb.x1) { x = bb.x0; ++y; }
        }
#pragma unroll
        for (int k = 0; k < EMIT_ILP; ++k) slot[k] = ok[k] ? atomicAdd(&tile_cursor[GUT_SUB_SLOT(t[k], sub, fr.num_tiles)], 1u) : 0u;
#pragma unroll
        for (int k = 0; k < EMIT_ILP; ++k) {
            if (ok[k] && slot[k] < capacity) entries[slot[k]] = make_uint2(depth_bits, i);
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// Per-tile sort + record expansion.
//
// One CTA per tile.  The bucket (depth bits, particle) is sorted on the DEPTH bits only, with stable 8-bit LSD
// radix passes over just the bit range that varies inside the bucket (3-4 passes for typical depth ranges);
// runs of bit-identical depths (rare) are then put in ascending particle order, which is the order the
// reference's stable sort leaves them in (emission order).  Pass structure: every warp owns a contiguous
// segment of the bucket, counts its digits into warp-private counters, one scan over (digit, warp) gives each
// warp its output cursors, and the warp scatters its segment 32 keys at a time ranking equal digits with
// match.any -- three CTA barriers per pass, no per-round barriers.
// The output is the sorted particle list only (4 B per intersection); the compositing kernels gather and expand
// the particle parameters themselves, one batch ahead of their walk, where that latency hides behind FP32 work.
// ---------------------------------------------------------------------------------------------------
constexpr uint32_t SORT_RADIX_BITS = 8;
constexpr uint32_t SORT_BINS = 1u << SORT_RADIX_BITS;
template <int WARPS>
struct SortShared {
    uint32_t warp_cnt[WARPS][SORT_BINS];  // per-warp digit counts, then per-warp output cursors
    uint32_t wsum[32];
    uint32_t or_bits, and_bits, ties;
};

// One stable counting pass on `bits` key bits starting at `shift`.  src/dst may be shared or global memory.
template <int THREADS>
__device__ void radix_pass(const uint32_t* src_k, const uint32_t* src_v, uint32_t* dst_k, uint32_t* dst_v, uint32_t n, uint32_t shift,
                           uint32_t mask, SortShared<THREADS / 32>& sh) {
    constexpr int WARPS = THREADS / 32;
    const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (uint32_t i = tid; i < WARPS * SORT_BINS; i += THREADS) (&sh.warp_cnt[0][0])[i] = 0;
    __syncthreads();
    const uint32_t seg = (((n + WARPS - 1) / WARPS) + 31u) & ~31u;   // keys per warp, multiple of 32
    const uint32_t s0 = min(n, wid * seg), s1 = min(n, s0 + seg);
    for (uint32_t i = s0 + lane; i < s1; i += 32) atomicAdd(&sh.warp_cnt[wid][(src_k[i] >> shift) & mask], 1u);
    __syncthreads();
    // cursors: cnt[w][d] <- sum_{d' < d} total(d') + sum_{w' < w} cnt[w'][d]
    for (uint32_t d0 = 0; d0 < SORT_BINS; d0 += THREADS) {   // one iteration when THREADS >= SORT_BINS
        const uint32_t d = d0 + tid;
        uint32_t c[WARPS], tot = 0;
        if (d < SORT_BINS) {
#pragma unroll
            for (int w = 0; w < WARPS; ++w) { c[w] = sh.warp_cnt[w][d]; tot += c[w]; }
        }
        uint32_t incl = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (uint32_t)o) incl += v;
        }
        if (lane == 31) sh.wsum[wid] = incl;
        __syncthreads();
        uint32_t off = 0;
        for (uint32_t w = 0; w < wid; ++w) off += sh.wsum[w];
        uint32_t run = off + incl - tot;
        if (d < SORT_BINS) {
#pragma unroll
            for (int w = 0; w < WARPS; ++w) { sh.warp_cnt[w][d] = run; run += c[w]; }
        }
        __syncthreads();
    }
    // stable scatter of the warp's own segment, 32 keys at a time
    for (uint32_t i0 = s0; i0 < s1; i0 += 32) {
        const uint32_t i = i0 + lane;
        const bool act = i < s1;
        uint32_t key = 0, val = 0, d = 0xffffffffu;
        if (act) { key = src_k[i]; val = src_v[i]; d = (key >> shift) & mask; }
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
        uint32_t pos = 0;
        if (act) pos = sh.warp_cnt[wid][d] + rank;
        __syncwarp();
        if (act) {
            dst_k[pos] = key; dst_v[pos] = val;
            if (rank == 0) sh.warp_cnt[wid][d] += __popc(peers);
        }
        __syncwarp();
    }
    __syncthreads();
}

template <int THREADS, uint32_t CAP>
__global__ void __launch_bounds__(THREADS) gut_tile_radix_sort_kernel(GutFrame fr, const uint32_t* __restrict__ tile_start,
                                                                      uint2* __restrict__ entries, uint2* __restrict__ scratch,
                                                                      const uint32_t* __restrict__ tile_flags,
                                                                      const uint32_t* __restrict__ flagged_count, bool oversize,
                                                                      uint32_t* __restrict__ sorted_idx) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ SortShared<THREADS / 32> sh;
    const uint32_t tid = threadIdx.x, lane = tid & 31;
    if (!oversize && *flagged_count == 0u) return;   // the ordinary frame: nothing was handed on

    // persistent: the launch is idle for ordinary frames (no flagged bucket, none above CAP), so it must cost next to nothing
    for (uint32_t tile = blockIdx.x; tile < fr.num_tiles; tile += gridDim.x) {
    const uint32_t b0 = tile_start[tile], b1 = tile_start[tile + 1];
    const uint32_t n = b1 - b0;
    if (n == 0 || !(n > CAP || tile_flags[tile])) continue;   // block-uniform

    // shared-memory layout: [key A | val A | key B | val B], CAP entries each; buckets above CAP sort in global memory
    // (keys and values de-interleaved in place over `entries` / `scratch`, which hold 2 x n words each).
    const bool in_smem = n <= CAP;
    uint32_t* ka = in_smem ? reinterpret_cast<uint32_t*>(smem_raw) : reinterpret_cast<uint32_t*>(scratch + b0);
    uint32_t* va = ka + (in_smem ? CAP : n);
    uint32_t* kb = in_smem ? va + CAP : reinterpret_cast<uint32_t*>(entries + b0);
    uint32_t* vb = kb + (in_smem ? CAP : n);

    __syncthreads();   // previous tile of this CTA is done with the shared buffers
    if (tid == 0) { sh.or_bits = 0; sh.and_bits = 0xffffffffu; sh.ties = 0; }
    __syncthreads();
    {
        uint32_t o = 0, a = 0xffffffffu;
        for (uint32_t i = tid; i < n; i += THREADS) {
            const uint2 e = entries[b0 + i];
            ka[i] = e.x; va[i] = e.y;
            o |= e.x; a &= e.x;
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) { o |= __shfl_xor_sync(0xffffffffu, o, s); a &= __shfl_xor_sync(0xffffffffu, a, s); }
        if (lane == 0) { atomicOr(&sh.or_bits, o); atomicAnd(&sh.and_bits, a); }
    }
    __syncthreads();
    const uint32_t varying = sh.or_bits ^ sh.and_bits;
    if (varying) {
        const uint32_t lo = __ffs(varying) - 1u, hi = 32u - __clz(varying);   // bits [lo, hi) differ somewhere in the bucket
        const uint32_t npass = (hi - lo + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
        for (uint32_t p = 0; p < npass; ++p) {
            const uint32_t shift = lo + p * SORT_RADIX_BITS;
            const uint32_t bits = min(SORT_RADIX_BITS, hi - shift);
            radix_pass<THREADS>(ka, va, kb, vb, n, shift, (1u << bits) - 1u, sh);
            uint32_t* t = ka; ka = kb; kb = t;
            t = va; va = vb; vb = t;
        }
    }
    // runs of identical depth: ascending particle index (what the reference's stable sort leaves, gutRenderer.cu:356-365)
    for (uint32_t i = tid; i + 1 < n; i += THREADS) {
        const uint32_t k = ka[i];
        if (k == ka[i + 1] && (i == 0 || ka[i - 1] != k)) {     // first element of a run of equal keys
            uint32_t e = i + 2;
            while (e < n && ka[e] == k) ++e;
            for (uint32_t a = i + 1; a < e; ++a) {              // insertion sort, runs are tiny
                const uint32_t v = va[a];
                uint32_t b = a;
                while (b > i && va[b - 1] > v) { va[b] = va[b - 1]; --b; }
                va[b] = v;
            }
        }
    }
    __syncthreads();

    // sorted particle list of the tile (the compositing kernels gather the parameters themselves)
    for (uint32_t i = tid; i < n; i += THREADS) sorted_idx[b0 + i] = va[i];
    }
}

// ---------------------------------------------------------------------------------------------------
// Per-tile sort, common path: ONE counting pass instead of 3-4 radix passes.
//
// The bucket's depth keys are binned monotonically on their value range, bin = (key - min) >> shift with BINS >= the
// bucket size, so the typical bin holds zero to a few entries.  Every thread keeps its EPT entries in registers:
//   load (+ min/max)  ->  bin count with shared atomics (the returned old value is the entry's slot in its bin)
//   ->  exclusive scan of the bins  ->  scatter into shared memory (bins in order, entries of a bin in arrival order)
//   ->  every entry ranks itself inside its bin by counting the bin's smaller (depth, particle) pairs and stores its
//       particle index at bin start + rank of the sorted list in global memory (neighbouring threads write neighbouring
//       words).
// (depth, particle) is a strict total order, so the result does not depend on the order the emit atomics filled the
// bucket in, and equal depths come out in ascending particle order like the reference's stable sort leaves them.
// Ranking costs sum(bin size^2): a bucket with a bin above BIN_LIMIT (depths piled up on a few values with outliers
// stretching the range) is flagged and left to the radix kernel below, whose cost does not depend on the distribution.
// ---------------------------------------------------------------------------------------------------
constexpr uint32_t BIN_LIMIT = 256;

template <int THREADS, uint32_t CAP, uint32_t LOG_BINS, int MIN_BLOCKS>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS) gut_tile_bin_sort_kernel(const uint32_t* __restrict__ tile_start, const uint2* __restrict__ entries,
                                                                                uint32_t n_min, uint32_t n_max, uint32_t* __restrict__ sorted_idx,
                                                                                uint32_t* __restrict__ tile_flags, uint32_t* __restrict__ flagged_count) {
    constexpr uint32_t BINS = 1u << LOG_BINS;
    constexpr int EPT = CAP / THREADS;        // entries per thread
    constexpr int BPT = BINS / THREADS;       // bins per thread in the scan
    constexpr int WARPS = THREADS / 32;
    static_assert(CAP % THREADS == 0 && BINS % THREADS == 0 && BPT % 4 == 0 && EPT % 4 == 0 && BIN_LIMIT <= 256, "tiling");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint2* sorted = reinterpret_cast<uint2*>(smem_raw);                    // [CAP]
    uint32_t* bins = reinterpret_cast<uint32_t*>(sorted + CAP);            // [BINS + 4]: counts, then exclusive starts (+ total)
    __shared__ uint32_t s_min, s_max, s_big, s_wsum[WARPS];

    const uint32_t tile = blockIdx.x;
    const uint32_t b0 = tile_start[tile], n = tile_start[tile + 1] - b0;
    if (n < n_min || n > n_max) return;       // this launch handles buckets of n_min..n_max entries (n_min >= 1)
    const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;

    uint2 ent[EPT];
    uint32_t kmin = 0xffffffffu, kmax = 0u;
#pragma unroll
    for (int e = 0; e < EPT; ++e) {
        const uint32_t i = tid + e * THREADS;
        ent[e] = make_uint2(0u, 0u);
        if (i < n) {
            ent[e] = __ldg(entries + b0 + i);
            kmin = min(kmin, ent[e].x); kmax = max(kmax, ent[e].x);
        }
    }
    if (tid == 0) { s_min = 0xffffffffu; s_max = 0u; s_big = 0u; }
#pragma unroll
    for (int k = 0; k < BPT / 4; ++k) reinterpret_cast<uint4*>(bins)[tid + k * THREADS] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    kmin = __reduce_min_sync(0xffffffffu, kmin); kmax = __reduce_max_sync(0xffffffffu, kmax);
    if (lane == 0) { atomicMin(&s_min, kmin); atomicMax(&s_max, kmax); }
    __syncthreads();
    kmin = s_min;
    const uint32_t range_bits = 32u - __clz(s_max - kmin);                 // 0 when every key is the same
    const uint32_t shift = range_bits > LOG_BINS ? range_bits - LOG_BINS : 0u;

    // slot of the entry inside its bin, 8 bits each: a bin that overflows them is above BIN_LIMIT and the bucket is handed on
    uint32_t slot4[EPT / 4];
#pragma unroll
    for (int e = 0; e < EPT; ++e) {
        if ((e & 3) == 0) slot4[e >> 2] = 0;
        if (tid + e * THREADS < n) slot4[e >> 2] |= (atomicAdd(&bins[(ent[e].x - kmin) >> shift], 1u) & 0xffu) << (8 * (e & 3));
    }
    __syncthreads();

    // exclusive scan of the bin counts, BPT consecutive bins per thread
    {
        uint32_t c[BPT], sum = 0, big = 0;
#pragma unroll
        for (int k = 0; k < BPT / 4; ++k) {
            const uint4 q = reinterpret_cast<const uint4*>(bins)[tid * (BPT / 4) + k];
            c[4 * k] = q.x; c[4 * k + 1] = q.y; c[4 * k + 2] = q.z; c[4 * k + 3] = q.w;
        }
#pragma unroll
        for (int k = 0; k < BPT; ++k) { sum += c[k]; big = max(big, c[k]); }
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (uint32_t)o) incl += v;
        }
        if (lane == 31) s_wsum[wid] = incl;
        if (big > BIN_LIMIT) s_big = 1u;
        __syncthreads();
        uint32_t run = incl - sum;
        for (uint32_t w = 0; w < wid; ++w) run += s_wsum[w];
#pragma unroll
        for (int k = 0; k < BPT; ++k) { const uint32_t t = c[k]; c[k] = run; run += t; }
#pragma unroll
        for (int k = 0; k < BPT / 4; ++k)
            reinterpret_cast<uint4*>(bins)[tid * (BPT / 4) + k] = make_uint4(c[4 * k], c[4 * k + 1], c[4 * k + 2], c[4 * k + 3]);
        if (tid == THREADS - 1) bins[BINS] = run;                          // = n
    }
    const bool big = s_big != 0u;
    if (tid == 0) {
        tile_flags[tile] = big ? 1u : 0u;
        if (big) atomicAdd(flagged_count, 1u);
    }
    if (big) return;                                                       // block-uniform: the radix kernel sorts this bucket
    __syncthreads();

#pragma unroll
    for (int e = 0; e < EPT; ++e)
        if (tid + e * THREADS < n) sorted[bins[(ent[e].x - kmin) >> shift] + ((slot4[e >> 2] >> (8 * (e & 3))) & 0xffu)] = ent[e];
    __syncthreads();

    for (uint32_t i = tid; i < n; i += THREADS) {
        const uint2 me = sorted[i];
        const uint32_t b = (me.x - kmin) >> shift;
        const uint32_t s0 = bins[b], s1 = bins[b + 1];
        uint32_t rank = 0;
        for (uint32_t j = s0; j < s1; ++j) {
            const uint2 o = sorted[j];
            rank += (o.x < me.x || (o.x == me.x && o.y < me.y)) ? 1u : 0u;
        }
        sorted_idx[b0 + s0 + rank] = me.y;
    }
}

// reference-shaped debug outputs: sorted u64 keys (tile<<32 | depth bits) and uvec2 tile ranges ((0,0) for empty tiles,
// as left by the memset at src/gutRenderer.cu:161)
__global__ void gut_export_keys_kernel(GutFrame fr, const uint32_t* __restrict__ tile_start, const uint32_t* __restrict__ sorted_idx,
                                       const float* __restrict__ depth, uint64_t* __restrict__ keys, uint32_t* __restrict__ ranges) {
    const uint32_t tile = blockIdx.x;
    const uint32_t b0 = tile_start[tile], b1 = tile_start[tile + 1];
    if (threadIdx.x == 0 && ranges) {
        ranges[2 * tile] = (b1 > b0) ? b0 : 0u;
        ranges[2 * tile + 1] = (b1 > b0) ? b1 : 0u;
    }
    if (keys)
        for (uint32_t i = b0 + threadIdx.x; i < b1; i += blockDim.x)
            keys[i] = ((uint64_t)tile << 32) | (uint64_t)__float_as_uint(depth[sorted_idx[i]]);
}

__global__ void gut_fill_outputs_kernel(uint32_t P, float4* __restrict__ rgba, float* __restrict__ dist, float* __restrict__ hits) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    rgba[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    dist[i] = 1e06f;
    hits[i] = 0.0f;
}

}  // namespace

cudaError_t gut_launch_tile_scan(uint32_t T, const uint32_t* tile_hist, uint32_t* tile_start, uint32_t* tile_cursor,
                                 uint32_t* stats, cudaStream_t stream) {
    gut_tile_scan_kernel<<<1, SCAN_THREADS, 0, stream>>>(T, tile_hist, tile_start, tile_cursor, stats);
    return cudaGetLastError();
}

cudaError_t gut_launch_emit(const GutConfig& cfg, const GutFrame& frame, uint32_t N, GutProjBuffers proj,
                            uint32_t* tile_cursor, uint2* entries, uint32_t capacity, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    gut_emit_kernel<<<(N + 255) / 256, 256, 0, stream>>>(cfg, frame, N, proj, tile_cursor, entries, capacity);
    return cudaGetLastError();
}

namespace {
constexpr uint32_t SORT_CAP_M = 8192, SORT_CAP_L = 12288, SORT_LOG_BINS_M = 13, SORT_LOG_BINS_L = 14;
constexpr size_t SORT_SMEM_M = SORT_CAP_M * sizeof(uint2) + ((1u << SORT_LOG_BINS_M) + 4) * sizeof(uint32_t);
constexpr size_t SORT_SMEM_L = SORT_CAP_L * sizeof(uint2) + ((1u << SORT_LOG_BINS_L) + 4) * sizeof(uint32_t);
constexpr size_t SORT_SMEM_R = 4 * SORT_CAP_L * sizeof(uint32_t);
}  // namespace

// Function attributes belong to the device (its primary context): gut_create calls this under its DeviceGuard for every
// handle, so a second handle on another GPU of the same process gets the dynamic shared-memory opt-in too.
cudaError_t gut_configure_tile_sort(int* sm_count) {
    cudaError_t e = cudaFuncSetAttribute(gut_tile_bin_sort_kernel<512, SORT_CAP_M, SORT_LOG_BINS_M, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SORT_SMEM_M);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(gut_tile_bin_sort_kernel<512, SORT_CAP_L, SORT_LOG_BINS_L, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SORT_SMEM_L);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(gut_tile_radix_sort_kernel<512, SORT_CAP_L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SORT_SMEM_R);
    if (e != cudaSuccess) return e;
    int dev = 0;
    e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    return cudaDeviceGetAttribute(sm_count, cudaDevAttrMultiProcessorCount, dev);
}

cudaError_t gut_launch_tile_sort(const GutFrame& frame, uint32_t N, uint32_t max_list, const uint32_t* tile_start, uint2* entries,
                                 uint2* scratch, uint32_t* tile_flags, uint32_t* flagged_count, uint32_t* sorted_idx, int sm_count,
                                 cudaStream_t stream) {
    // buckets up to 8192 entries: 512 threads, 8192 bins (96 KB, two CTAs per SM); up to 12288: 16384 bins (160 KB);
    // flagged buckets and buckets above 12288 entries: radix passes (in 192 KB of shared memory / over global memory).
    (void)N;
    gut_tile_bin_sort_kernel<512, SORT_CAP_M, SORT_LOG_BINS_M, 2><<<frame.num_tiles, 512, SORT_SMEM_M, stream>>>(
        tile_start, entries, 1u, SORT_CAP_M, sorted_idx, tile_flags, flagged_count);
    if (max_list > SORT_CAP_M)
        gut_tile_bin_sort_kernel<512, SORT_CAP_L, SORT_LOG_BINS_L, 1><<<frame.num_tiles, 512, SORT_SMEM_L, stream>>>(
            tile_start, entries, SORT_CAP_M + 1u, SORT_CAP_L, sorted_idx, tile_flags, flagged_count);
    const uint32_t grid = frame.num_tiles < (uint32_t)sm_count ? frame.num_tiles : (uint32_t)sm_count;
    gut_tile_radix_sort_kernel<512, SORT_CAP_L><<<grid, 512, SORT_SMEM_R, stream>>>(frame, tile_start, entries, scratch, tile_flags, flagged_count,
                                                                                max_list > SORT_CAP_L, sorted_idx);
    return cudaGetLastError();
}

cudaError_t gut_launch_export_keys(const GutFrame& frame, const uint32_t* tile_start, const uint32_t* sorted_idx, const float* depth,
                                   uint64_t* sorted_keys, uint32_t* ranges, cudaStream_t stream) {
    gut_export_keys_kernel<<<frame.num_tiles, 128, 0, stream>>>(frame, tile_start, sorted_idx, depth, sorted_keys, ranges);
    return cudaGetLastError();
}

cudaError_t gut_launch_fill_outputs(uint32_t num_pixels, float* rgba, float* dist, float* hits, cudaStream_t stream) {
    if (num_pixels == 0) return cudaSuccess;
    gut_fill_outputs_kernel<<<(num_pixels + 255) / 256, 256, 0, stream>>>(num_pixels, reinterpret_cast<float4*>(rgba), dist, hits);
    return cudaGetLastError();
}
