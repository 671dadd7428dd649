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
// Exclusive scan of the tile histogram (T is a few thousand: one CTA, chunks of 1024 with a carry).
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
        const uint32_t v = i < T ? hist[i] : 0u;
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
        if (i < T) { start[i] = excl; cursor[i] = 0u; }
        __syncthreads();
        if (tid == SCAN_THREADS - 1) carry_s = carry + warp_sums[31];
        __syncthreads();
    }
    atomicMax(&max_s, local_max);
    __syncthreads();
    if (tid == 0) { start[T] = carry_s; stats[0] = carry_s; stats[1] = max_s; }
}

// ---------------------------------------------------------------------------------------------------
// K2: emit.  One thread per particle re-walks its tile box with the same culling test as K1 and appends
// (depth bits, particle) to each surviving tile's bucket.  reference: GUTProjector::expand, gutProjector.cuh:324-388.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) gut_emit_kernel(GutConfig cfg, GutFrame fr, uint32_t N, GutProjBuffers proj,
                                                       const uint32_t* __restrict__ tile_start, uint32_t* __restrict__ tile_cursor,
                                                       uint2* __restrict__ entries) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    if (proj.tiles_count[i] == 0) return;
    const float2 ext = proj.extent[i];
    if (ext.x <= 1e-06f) return;
    const uint32_t depth_bits = __float_as_uint(proj.depth[i]);
    const float2 centre = proj.proj_pos[i];
    const float4 co = proj.conic_opacity[i];
    const float max_power = gut_logf(co.w / cfg.min_alpha);
    const GutTileBox bb = gut_tile_bbox(fr.grid_x, fr.grid_y, centre, ext);
    for (uint32_t y = bb.y0; y < bb.y1; ++y)
        for (uint32_t x = bb.x0; x < bb.x1; ++x) {
            if (!cfg.tile_based_culling || gut_tile_min_power((float)x, (float)y, co, centre) < max_power) {
                const uint32_t t = y * fr.grid_x + x;
                const uint32_t slot = atomicAdd(&tile_cursor[t], 1u);
                const uint32_t b0 = tile_start[t], b1 = tile_start[t + 1];
                if (b0 + slot < b1) entries[b0 + slot] = make_uint2(depth_bits, i);
            }
        }
}

// ---------------------------------------------------------------------------------------------------
// Per-tile sort + record expansion.
// ---------------------------------------------------------------------------------------------------
constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr uint32_t SORT_SMEM_CAP = 4096;  // entries per buffer kept in shared memory (2 x 32 KB)

struct SortShared {
    uint32_t hist[256];
    uint32_t base[256];
    uint32_t warp_cnt[SORT_WARPS][256];
    uint32_t or_lo, or_hi, and_lo, and_hi;
};

// One stable 8-bit counting pass over n keys, src -> dst (generic pointers: shared or global memory).
// Returns false (and leaves dst untouched) when the digit is constant over the bucket.
__device__ bool radix_pass(const uint64_t* src, uint64_t* dst, uint32_t n, int shift, SortShared& sh) {
    const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    sh.hist[tid] = 0;
    __syncthreads();
    for (uint32_t i = tid; i < n; i += SORT_THREADS) atomicAdd(&sh.hist[(uint32_t)(src[i] >> shift) & 255u], 1u);
    __syncthreads();
    // exclusive scan of 256 bins by warp 0..7 cooperatively (each thread owns one bin)
    const uint32_t h = sh.hist[tid];
    const int uniform = __syncthreads_or(h == n);
    if (uniform) return false;
    uint32_t incl = h;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (uint32_t)o) incl += v;
    }
    __shared__ uint32_t wsum[SORT_WARPS];
    if (lane == 31) wsum[wid] = incl;
    __syncthreads();
    uint32_t off = 0;
    for (uint32_t w = 0; w < wid; ++w) off += wsum[w];
    sh.base[tid] = off + incl - h;
    for (int w = 0; w < SORT_WARPS; ++w) sh.warp_cnt[w][tid] = 0;
    __syncthreads();

    // stable scatter, SORT_THREADS keys per round, rounds in order
    for (uint32_t r0 = 0; r0 < n; r0 += SORT_THREADS) {
        const uint32_t i = r0 + tid;
        const bool act = i < n;
        uint64_t key = 0;
        uint32_t d = 0xffffffffu;
        if (act) { key = src[i]; d = (uint32_t)(key >> shift) & 255u; }
        // rank among the lanes of this warp that hold the same digit (lower lanes first)
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
        if (act && rank == 0) sh.warp_cnt[wid][d] = __popc(peers);
        __syncthreads();
        if (act) {
            uint32_t before = 0;
            for (uint32_t w = 0; w < wid; ++w) before += sh.warp_cnt[w][d];
            dst[sh.base[d] + before + rank] = key;
        }
        __syncthreads();
        // advance the bases by this round's totals and clear the per-warp counters
        {
            uint32_t tot = 0;
#pragma unroll
            for (int w = 0; w < SORT_WARPS; ++w) { tot += sh.warp_cnt[w][tid]; sh.warp_cnt[w][tid] = 0; }
            sh.base[tid] += tot;
        }
        __syncthreads();
    }
    return true;
}

__global__ void __launch_bounds__(SORT_THREADS) gut_tile_sort_kernel(GutFrame fr, uint32_t N, const uint32_t* __restrict__ tile_start,
                                                                     uint2* __restrict__ entries, uint2* __restrict__ scratch,
                                                                     const float4* __restrict__ density12,
                                                                     const float4* __restrict__ features,
                                                                     uint32_t* __restrict__ sorted_idx, GutRecord* __restrict__ records,
                                                                     GutRecordBwd* __restrict__ records_bwd) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* buf_a = reinterpret_cast<uint64_t*>(smem_raw);
    uint64_t* buf_b = buf_a + SORT_SMEM_CAP;
    __shared__ SortShared sh;

    const uint32_t tile = blockIdx.x;
    const uint32_t b0 = tile_start[tile], b1 = tile_start[tile + 1];
    const uint32_t n = b1 - b0;
    if (n == 0) return;
    const uint32_t tid = threadIdx.x;

    // keys are (depth_bits << 32) | particle; the bucket in global memory is uint2{depth_bits, particle} = little-endian
    // {lo = depth, hi = particle}, so swap halves while loading.
    const bool in_smem = n <= SORT_SMEM_CAP;
    uint64_t* src = in_smem ? buf_a : reinterpret_cast<uint64_t*>(entries + b0);
    uint64_t* dst = in_smem ? buf_b : reinterpret_cast<uint64_t*>(scratch + b0);
    if (tid == 0) { sh.or_lo = 0; sh.or_hi = 0; sh.and_lo = 0xffffffffu; sh.and_hi = 0xffffffffu; }
    __syncthreads();
    {
        uint32_t olo = 0, ohi = 0, alo = 0xffffffffu, ahi = 0xffffffffu;
        for (uint32_t i = tid; i < n; i += SORT_THREADS) {
            const uint2 e = entries[b0 + i];
            const uint64_t key = ((uint64_t)e.x << 32) | e.y;
            src[i] = key;  // in the global case this rewrites the bucket in key order
            olo |= e.y; alo &= e.y; ohi |= e.x; ahi &= e.x;
        }
        atomicOr(&sh.or_lo, olo); atomicOr(&sh.or_hi, ohi); atomicAnd(&sh.and_lo, alo); atomicAnd(&sh.and_hi, ahi);
    }
    __syncthreads();
    const uint64_t varying = (((uint64_t)(sh.or_hi ^ sh.and_hi)) << 32) | (uint64_t)(sh.or_lo ^ sh.and_lo);
    for (int shift = 0; shift < 64; shift += 8) {
        if (((varying >> shift) & 255ull) == 0) continue;  // digit identical for every key of the bucket
        if (radix_pass(src, dst, n, shift, sh)) { uint64_t* t = src; src = dst; dst = t; }
        __syncthreads();
    }

    // ---- epilogue: sorted particle list + record stream (contiguous per tile) ----
    for (uint32_t i = tid; i < n; i += SORT_THREADS) {
        const uint32_t p = (uint32_t)(src[i] & 0xffffffffull);
        sorted_idx[b0 + i] = p;
        const float4 r0 = __ldg(density12 + 3 * (size_t)p);
        const float4 q = __ldg(density12 + 3 * (size_t)p + 1);
        const float4 r2 = __ldg(density12 + 3 * (size_t)p + 2);
        const float4 ft = __ldg(features + p);
        float3 c0, c1, c2;
        quat_to_columns(q.x, q.y, q.z, q.w, c0, c1, c2);  // columns of R == rows of R^T
        const float ix = 1.0f / r2.x, iy = 1.0f / r2.y, iz = 1.0f / r2.z;
        const float3 m0 = c0 * ix, m1 = c1 * iy, m2 = c2 * iz;  // rows of ms = diag(1/s) R^T
        const float3 v = fr.cam_centre - f3(r0.x, r0.y, r0.z);  // frame origin - position
        float4* out = reinterpret_cast<float4*>(records + b0 + i);
        out[0] = make_float4(m0.x, m0.y, m0.z, dot(m0, v));
        out[1] = make_float4(m1.x, m1.y, m1.z, dot(m1, v));
        out[2] = make_float4(m2.x, m2.y, m2.z, dot(m2, v));
        out[3] = make_float4(r2.x, r2.y, r2.z, r0.w);
        out[4] = make_float4(fmaxf(ft.x, 0.0f), fmaxf(ft.y, 0.0f), fmaxf(ft.z, 0.0f), __uint_as_float(p));
        float4* outb = reinterpret_cast<float4*>(records_bwd + b0 + i);
        outb[0] = q;
        outb[1] = make_float4(ix, iy, iz, 0.0f);
        outb[2] = make_float4(r0.x, r0.y, r0.z, 0.0f);
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
                            const uint32_t* tile_start, uint32_t* tile_cursor, uint2* entries, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    gut_emit_kernel<<<(N + 255) / 256, 256, 0, stream>>>(cfg, frame, N, proj, tile_start, tile_cursor, entries);
    return cudaGetLastError();
}

cudaError_t gut_launch_tile_sort(const GutFrame& frame, uint32_t N, const uint32_t* tile_start, uint2* entries, uint2* scratch,
                                 const float* density12, const float4* features, uint32_t* sorted_idx, GutRecord* records,
                                 GutRecordBwd* records_bwd, cudaStream_t stream) {
    const size_t smem = 2 * SORT_SMEM_CAP * sizeof(uint64_t);
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(gut_tile_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    gut_tile_sort_kernel<<<frame.num_tiles, SORT_THREADS, smem, stream>>>(frame, N, tile_start, entries, scratch,
                                                                         reinterpret_cast<const float4*>(density12), features,
                                                                         sorted_idx, records, records_bwd);
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
