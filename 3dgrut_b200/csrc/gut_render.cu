// gut_render.cu -- K4 (per-tile forward compositing) and K5 (backward) of the 3DGUT rasteriser.
//
// reference: render / renderBackward kernels (include/3dgut/kernels/cuda/renderers/gutRenderer.cuh:83-165),
//   GUTKBufferRenderer::evalKBuffer / evalBackwardNoKBuffer (renderers/gutKBufferRenderer.cuh:217-386),
//   ray set-up initializeRay / initializeBackwardRay (common/rayPayload.cuh:75-129, rayPayloadBackward.cuh:30-58),
//   hit math gaussianParticle.hit / integrateHit (kernels/slang/models/gaussianParticles.slang:96-254),
//   per-hit gradient processHitBwd (kernels/cuda/models/gaussianParticles.cuh:480-738).
//
// B200 design (DESIGN.md "K4/K5").  Both kernels are bound by FP32 instruction issue / the FMA pipe (ncu: DRAM < 2 %),
// so the design minimises issue slots per ray x particle test and keeps the FMA pipe fed:
//  * one WARP per CTA, owning an 8x8 pixel block (four blocks per 16x16 tile share the tile's sorted list).  A warp
//    stages, walks and retires on its own: no CTA barrier, no waiting for the slowest warp of a tile, registers and
//    shared memory are released the moment a block saturates;
//  * every lane carries TWO rays (vertically adjacent pixels) in packed fp32 registers.  sm_100a issues FFMA2 / FMUL2 /
//    FADD2 (fma.rn.f32x2 ...) in one slot for two fp32 lanes; per-particle record values enter as scalar broadcast
//    operands.  The test, the compositing update and the whole per-hit gradient math run once for both rays; a ray
//    that does not take the hit is masked (alpha = 0), not branched around;
//  * staging runs one batch (32 particles, one per lane) ahead of the walk: the lane's entry of the tile's sorted list
//    is a coalesced load two batches ahead, its parameters (3 x 16 B + its K1 colour) are gathered by cp.async (LDGSTS,
//    no registers held across the walk) and expanded once into an 80-byte shared-memory record that all 64 rays of
//    the block read by broadcast -- quaternion->matrix math happens once per (block, particle), not per ray.  DRAM
//    sits at ~2 % during compositing, so the gathers are free;
//  * records carry ms = diag(1/s) R^T and go = ms (frame origin - position).  When every ray of the block starts
//    at the frame origin (true for pinhole and fisheye bundles; checked per warp, bitwise) the canonical ray origin
//    is go itself and a test is ONE mat-vec, a cross product, one MUFU.RCP and one MUFU.EX2; otherwise go is
//    corrected by ms (ray origin - frame origin).  Normalisation and the hit distance are computed for accepted
//    hits only, from the unnormalised direction: t = |u.go| sqrt(|s*u|^2) / |u|^2;
//  * the translation unit is built with -use_fast_math like the reference (setup_3dgut.py:83): FTZ, MUFU
//    division / sqrt / exp2;
//  * backward: the 15 per-particle gradient values of the block's 64 rays are summed with a packed 16-shuffle
//    transpose-reduce (not 14 x 5 butterflies) and ONE red.global.add.f32 instruction per (warp, particle) -- and
//    nothing at all when no ray of the block hit (the reference reduces and does 14 atomics regardless,
//    shRadiativeGaussianParticles.cuh:409-482 called outside the isAlive branch at gutKBufferRenderer.cuh:379-383).
#include <stdlib.h>

#include "gut_render_common.cuh"

namespace {

constexpr int BLOCKS_PER_TILE = 4;    // 8x8 pixel blocks (one warp, two rays per lane) of a 16x16 tile
#ifndef GUT_FWD_MIN_BLOCKS
#define GUT_FWD_MIN_BLOCKS 24         // resident one-warp CTAs per SM the forward is compiled for (80 registers); measured 20: 1.282, 24: 1.286, 28: 1.311 ms
#endif
#ifndef GUT_BWD_MIN_BLOCKS
#define GUT_BWD_MIN_BLOCKS 20         // resident one-warp CTAs per SM the backward kernel is compiled for (96 registers);
#endif                                // measured: 24 (80 regs, spills) 4.76 ms, 20 4.05 ms, 18 (112 regs) 4.21 ms per view

// ---- packed FP32 (sm_100a FFMA2 / FMUL2 / FADD2: two fp32 lanes per instruction, one issue slot) -----------
// Both compositing kernels are issue bound, and a packed instruction does the work of two scalar ones for the same
// issue slot (scripts/microbench/ffma2.cu: 4 FFMA2 + 8 IADD take 0.75x the time of 8 FFMA + 8 IADD).  Every lane
// therefore carries TWO rays (vertically adjacent pixels): all per-ray quantities are F2 pairs, per-particle
// quantities enter as scalar broadcast operands (ptxas folds mov.b64 {x,x} into the R.F32 operand form) and the
// whole test, the state update and the per-hit gradient math run once for both rays.
typedef unsigned long long u64;
struct F2 { u64 v; };
GUT_D F2 mk2(float lo, float hi) { F2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi)); return r; }
GUT_D F2 bc(float x) { return mk2(x, x); }
GUT_D float lo(F2 a) { float l, h; asm("mov.b64 {%0, %1}, %2;" : "=f"(l), "=f"(h) : "l"(a.v)); return l; }
GUT_D float hi(F2 a) { float l, h; asm("mov.b64 {%0, %1}, %2;" : "=f"(l), "=f"(h) : "l"(a.v)); return h; }
GUT_D F2 fma(F2 a, F2 b, F2 c) { F2 r; asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v)); return r; }
GUT_D F2 operator*(F2 a, F2 b) { F2 r; asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
GUT_D F2 operator+(F2 a, F2 b) { F2 r; asm("add.rn.ftz.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
GUT_D F2 operator-(F2 a, F2 b) { F2 r; asm("sub.rn.ftz.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
// pure sign flip (neg.f32 without .ftz), so that ptxas folds it into the consumer's operand modifier
GUT_D F2 operator-(F2 a) {
    float nl, nh;
    asm("neg.f32 %0, %1;" : "=f"(nl) : "f"(lo(a)));
    asm("neg.f32 %0, %1;" : "=f"(nh) : "f"(hi(a)));
    return mk2(nl, nh);
}
struct F2x3 { F2 x, y, z; };
GUT_D F2x3 bc3(float3 a) { F2x3 r; r.x = bc(a.x); r.y = bc(a.y); r.z = bc(a.z); return r; }
GUT_D F2 dot(const F2x3& a, const F2x3& b) { return fma(a.z, b.z, fma(a.y, b.y, a.x * b.x)); }
GUT_D F2x3 cross(const F2x3& a, const F2x3& b) {
    F2x3 r;
    r.x = fma(a.y, b.z, -(a.z * b.y));
    r.y = fma(a.z, b.x, -(a.x * b.z));
    r.z = fma(a.x, b.y, -(a.y * b.x));
    return r;
}
GUT_D F2x3 operator*(const F2x3& a, F2 s) { F2x3 r; r.x = a.x * s; r.y = a.y * s; r.z = a.z * s; return r; }

// pixels owned by this thread: warp w -> 8x8 block (w&1, w>>1), lane -> column lane&7, rows 2*(lane>>3) and +1
GUT_D void thread_pixels(const GutFrame& fr, uint32_t tile, uint32_t w, uint32_t& px, uint32_t& py) {
    const uint32_t tx = tile % fr.grid_x, ty = tile / fr.grid_x;
    const uint32_t l = threadIdx.x & 31;
    px = tx * GUT_TILE + (w & 1) * 8 + (l & 7);
    py = ty * GUT_TILE + (w >> 1) * 8 + (l >> 3) * 2;
}

// the two rays of a lane, packed
struct RayPair {
    F2x3 d, dlt;          // world direction; ray origin - frame origin
    float tmin[2], tmax[2];
    bool inside[2], valid[2];
    uint32_t pix[2];
};
GUT_D RayPair make_ray_pair(const GutFrame& fr, const float* __restrict__ ray_ori, const float* __restrict__ ray_dir, uint32_t tile,
                            uint32_t block, bool& at_origin) {
    uint32_t px, py;
    thread_pixels(fr, tile, block, px, py);
    RayPair rp;
    float3 d[2], e[2];
    at_origin = true;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        rp.inside[h] = px < fr.width && (py + h) < fr.height;
        rp.pix[h] = rp.inside[h] ? px + fr.width * (py + h) : 0u;
        Ray r;
        r.valid = false; r.o = fr.cam_centre; r.d = f3(0.f, 0.f, 1.f); r.tmin = 0.f; r.tmax = 0.f;
        if (rp.inside[h]) r = make_ray(fr, ray_ori, ray_dir, rp.pix[h]);
        d[h] = r.d; e[h] = r.o - fr.cam_centre;
        rp.tmin[h] = r.tmin; rp.tmax[h] = r.tmax; rp.valid[h] = r.valid;
        at_origin = at_origin && (!rp.inside[h] || ((e[h].x == 0.0f) && (e[h].y == 0.0f) && (e[h].z == 0.0f)));
    }
    rp.d.x = mk2(d[0].x, d[1].x); rp.d.y = mk2(d[0].y, d[1].y); rp.d.z = mk2(d[0].z, d[1].z);
    rp.dlt.x = mk2(e[0].x, e[1].x); rp.dlt.y = mk2(e[0].y, e[1].y); rp.dlt.z = mk2(e[0].z, e[1].z);
    return rp;
}


// The four compositing thresholds are -D defines in the reference (setup_3dgut.py:47-70) and run-time fields of GutConfig
// here.  STD = true instantiates the kernels with the reference's default values as immediates (no constant-bank loads,
// no registers); gut_launch_render* picks it when the handle's configuration holds exactly those values.
constexpr float STD_MIN_RESPONSE = 0.0113f, STD_MIN_ALPHA = (float)(1.0 / 255.0), STD_MAX_ALPHA = 0.99f, STD_MIN_T = 0.0001f;
template <bool STD> GUT_D float thr_min_response(const GutConfig& c) { return STD ? STD_MIN_RESPONSE : c.min_kernel_response; }
template <bool STD> GUT_D float thr_min_alpha(const GutConfig& c) { return STD ? STD_MIN_ALPHA : c.min_alpha; }
template <bool STD> GUT_D float thr_max_alpha(const GutConfig& c) { return STD ? STD_MAX_ALPHA : c.max_alpha; }
template <bool STD> GUT_D float thr_min_T(const GutConfig& c) { return STD ? STD_MIN_T : c.min_transmittance; }
inline bool thresholds_are_standard(const GutConfig& c) {
    return c.min_kernel_response == STD_MIN_RESPONSE && c.min_alpha == STD_MIN_ALPHA && c.max_alpha == STD_MAX_ALPHA &&
           c.min_transmittance == STD_MIN_T;
}

// canonical-frame test shared by forward and backward, for the two rays of a lane.  u = ms d (unnormalised canonical
// direction), go = canonical origin, c = u x go.
struct Canon2 {
    F2x3 go, u, c;
    F2 l2, il2, g2, resp, araw;
};
template <int DEG, bool UNIFORM, bool STD>
GUT_D void particle_test(const GutConfig& cfg, const float4 A, const float4 B, const float4 Cm, const float density, const RayPair& ray,
                         Canon2& k, bool& accept0, bool& accept1) {
    k.go.x = bc(A.w); k.go.y = bc(B.w); k.go.z = bc(Cm.w);
    if (!UNIFORM) {  // the rays do not start at the frame origin: go += ms (ray origin - frame origin)
        k.go.x = fma(bc(A.z), ray.dlt.z, fma(bc(A.y), ray.dlt.y, fma(bc(A.x), ray.dlt.x, k.go.x)));
        k.go.y = fma(bc(B.z), ray.dlt.z, fma(bc(B.y), ray.dlt.y, fma(bc(B.x), ray.dlt.x, k.go.y)));
        k.go.z = fma(bc(Cm.z), ray.dlt.z, fma(bc(Cm.y), ray.dlt.y, fma(bc(Cm.x), ray.dlt.x, k.go.z)));
    }
    k.u.x = fma(bc(A.z), ray.d.z, fma(bc(A.y), ray.d.y, bc(A.x) * ray.d.x));
    k.u.y = fma(bc(B.z), ray.d.z, fma(bc(B.y), ray.d.y, bc(B.x) * ray.d.x));
    k.u.z = fma(bc(Cm.z), ray.d.z, fma(bc(Cm.y), ray.d.y, bc(Cm.x) * ray.d.x));
    k.l2 = dot(k.u, k.u);
    k.c = cross(k.u, k.go);
    k.il2 = mk2(mufu_rcp(lo(k.l2)), mufu_rcp(hi(k.l2)));
    k.g2 = dot(k.c, k.c) * k.il2;  // |normalize(u) x go|^2
    float r0, r1;
    if (DEG == 2) {   // exp(-g/2) = 2^(-g log2(e) / 2): one packed FMUL + one MUFU.EX2 per ray
        const F2 e = k.g2 * bc(-0.72134752044448170368f);
        r0 = mufu_ex2(lo(e)); r1 = mufu_ex2(hi(e));
    } else {
        r0 = kernel_response<DEG>(lo(k.g2)); r1 = kernel_response<DEG>(hi(k.g2));
    }
    k.resp = mk2(r0, r1);
    k.araw = k.resp * bc(density);
    // alpha = min(max_alpha, araw) > min_alpha  <=>  araw > min_alpha  (max_alpha > min_alpha is checked at gut_create)
    accept0 = (r0 > thr_min_response<STD>(cfg)) && (lo(k.araw) > thr_min_alpha<STD>(cfg));
    accept1 = (r1 > thr_min_response<STD>(cfg)) && (hi(k.araw) > thr_min_alpha<STD>(cfg));
}

// ---------------------------------------------------------------------------------------------------
// K4 forward
// ---------------------------------------------------------------------------------------------------
struct FwdState {
    F2 T, D, cr, cg, cb;
    uint32_t nhits[2], ntests[2];
    bool alive[2];
};

// Returns, as a bit per record of the batch, which records one of this lane's two rays accepted while alive (before the
// tMinMax test): the union over the warp is what the backward has to revisit (hit_mask_slot below).
template <int DEG, bool UNIFORM, bool COUNT, bool STD>
GUT_D uint32_t forward_walk(const GutConfig& cfg, const float4* __restrict__ rec, uint32_t cnt, const RayPair& ray, FwdState& st) {
    uint32_t accepted = 0;
    const uint32_t cnt_pad = (cnt + WALK_UNROLL - 1) & ~(WALK_UNROLL - 1);
    for (uint32_t j0 = 0; j0 < cnt_pad; j0 += ALIVE_CHECK) {
        if (!__any_sync(0xffffffffu, st.alive[0] || st.alive[1])) break;
        const uint32_t je = min(cnt_pad, j0 + ALIVE_CHECK);
        for (uint32_t j = j0; j < je; j += WALK_UNROLL) {
#pragma unroll
            for (uint32_t q = 0; q < WALK_UNROLL; q += 2) {
                // two independent tests in flight (the walk is latency bound at 4-5 warps per scheduler), then their hits in order
                const float4* r = rec + 5 * (j + q);
                Canon2 k[2];
                bool acc[2][2];
                float4 S[2];
#pragma unroll
                for (int p = 0; p < 2; ++p) {
                    const float4 A = r[5 * p], B = r[5 * p + 1], Cm = r[5 * p + 2];
                    S[p] = r[5 * p + 3];
                    particle_test<DEG, UNIFORM, STD>(cfg, A, B, Cm, S[p].w, ray, k[p], acc[p][0], acc[p][1]);
                }
#pragma unroll
                for (int p = 0; p < 2; ++p) {
                    if (COUNT && (j + q + p) < cnt) { st.ntests[0] += st.alive[0]; st.ntests[1] += st.alive[1]; }
                    const bool h0 = acc[p][0] && st.alive[0], h1 = acc[p][1] && st.alive[1];
                    if (h0 || h1) {
                        accepted |= 1u << (j + q + p);
                        // hit distance |s * grd * dot(grd, -go)| with grd = u / |u|  (canonicalRayDistance, gaussianParticles.slang:165-171)
                        F2x3 su;
                        su.x = bc(S[p].x) * k[p].u.x; su.y = bc(S[p].y) * k[p].u.y; su.z = bc(S[p].z) * k[p].u.z;
                        const F2 ss = dot(su, su), ug = dot(k[p].u, k[p].go);
                        const float t0 = fabsf(lo(ug)) * lo(k[p].il2) * mufu_sqrt(lo(ss));
                        const float t1 = fabsf(hi(ug)) * hi(k[p].il2) * mufu_sqrt(hi(ss));
                        const bool in0 = h0 && t0 > ray.tmin[0] && t0 < ray.tmax[0];
                        const bool in1 = h1 && t1 > ray.tmin[1] && t1 < ray.tmax[1];
                        // a ray that does not take the hit composites alpha = 0: its state is left unchanged
                        const F2 alpha = mk2(in0 ? fminf(thr_max_alpha<STD>(cfg), lo(k[p].araw)) : 0.0f, in1 ? fminf(thr_max_alpha<STD>(cfg), hi(k[p].araw)) : 0.0f);
                        const float4 F = r[5 * p + 4];
                        const F2 w = alpha * st.T;
                        st.D = fma(mk2(in0 ? t0 : 0.0f, in1 ? t1 : 0.0f), w, st.D);
                        st.T = st.T * (bc(1.0f) - alpha);
                        st.cr = fma(bc(F.x), w, st.cr); st.cg = fma(bc(F.y), w, st.cg); st.cb = fma(bc(F.z), w, st.cb);
                        if (lo(w) > 0.0f) st.nhits[0]++;
                        if (hi(w) > 0.0f) st.nhits[1]++;
                        if (lo(st.T) < thr_min_T<STD>(cfg)) st.alive[0] = false;
                        if (hi(st.T) < thr_min_T<STD>(cfg)) st.alive[1] = false;
                    }
                }
            }
        }
    }
    return accepted;
}

// The forward leaves one 32-bit mask per (8x8 block, batch of 32 list entries): bit j set = some ray of the block accepted
// entry j while alive.  The backward walks only those entries (41 % of a block's tests find no hit at the bench
// workload).  The mask is a superset of what the backward accepts: its rays die no later than the forward's (it
// composites the forward's hits plus those the forward's tMinMax test dropped), and it re-tests every revisited entry.
// Slot of batch b of a tile whose list starts at b0: (b0 / 32 + tile + b) is unique across tiles.
GUT_D size_t hit_mask_slot(uint32_t b0, uint32_t tile, uint32_t b, uint32_t block) {
    return ((size_t)(b0 >> 5) + tile + b) * BLOCKS_PER_TILE + block;
}

// COUNT = true is the instrumented variant used by gut_measure_work(): it writes no image, it counts the ray x particle
// tests and accepted hits the frame performs (the units the FP32 roofline of K4/K5 is expressed in).
template <int DEG, bool COUNT, bool STD>
__global__ void __launch_bounds__(RENDER_THREADS, GUT_FWD_MIN_BLOCKS) gut_render_kernel(GutConfig cfg, GutFrame fr, const uint32_t* __restrict__ tile_start,
                                                                     GutParticleSource src,
                                                                     const float* __restrict__ ray_ori, const float* __restrict__ ray_dir,
                                                                     float4* __restrict__ rgba, float* __restrict__ dist,
                                                                     float* __restrict__ hits, unsigned long long* __restrict__ counters,
                                                                     uint32_t* __restrict__ hit_masks /* nullable: inference */, bool block_culling) {
    __shared__ __align__(16) unsigned char smem_raw[StageLayout<false>::TOTAL];

    // CTAs are dispatched in blockIdx order: tiles longest list first (gut_tile_scan_kernel leaves the order behind tile_start)
    const uint32_t tile = __ldg(tile_start + fr.num_tiles + 1 + blockIdx.x / BLOCKS_PER_TILE), block = blockIdx.x % BLOCKS_PER_TILE;
    bool at_origin;
    const RayPair ray = make_ray_pair(fr, ray_ori, ray_dir, tile, block, at_origin);
    const bool uniform = __all_sync(0xffffffffu, at_origin);

    const uint32_t b0 = tile_start[tile], b1 = tile_start[tile + 1];
    WarpStager<false> stg;
    stg.init(smem_raw, src, b0, b1 - b0, fr.cam_centre);

    FwdState st;
    st.T = bc(1.0f); st.D = bc(0.0f); st.cr = st.cg = st.cb = bc(0.0f);
#pragma unroll
    for (int h = 0; h < 2; ++h) { st.nhits[h] = 0; st.ntests[h] = 0; st.alive[h] = ray.inside[h] && ray.valid[h]; }

    // cone around the block's live rays for block_cull (kept in shared memory: the walk has no registers to spare)
    __shared__ float4 s_cone;
    {
        const uint32_t lane = threadIdx.x & 31;
        float3 dn[2], sum = f3(0.f, 0.f, 0.f);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const float3 d = f3(h ? hi(ray.d.x) : lo(ray.d.x), h ? hi(ray.d.y) : lo(ray.d.y), h ? hi(ray.d.z) : lo(ray.d.z));
            dn[h] = rsqrtf(dot(d, d)) * d;
            if (st.alive[h]) sum = sum + dn[h];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            sum.x += __shfl_xor_sync(0xffffffffu, sum.x, o); sum.y += __shfl_xor_sync(0xffffffffu, sum.y, o); sum.z += __shfl_xor_sync(0xffffffffu, sum.z, o);
        }
        const float n2 = dot(sum, sum);
        const float3 axis = n2 > 0.0f ? rsqrtf(n2) * sum : f3(0.f, 0.f, 1.f);
        float cmin = 1.0f;
#pragma unroll
        for (int h = 0; h < 2; ++h) if (st.alive[h]) cmin = fminf(cmin, dot(axis, dn[h]));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cmin = fminf(cmin, __shfl_xor_sync(0xffffffffu, cmin, o));
        // tan(half angle), widened; no cull (w < 0) for degenerate bundles, per-ray origins or another kernel degree
        float tan_h = -1.0f;
        if (DEG == 2 && block_culling && uniform && cmin > 0.5f) tan_h = sqrtf(fmaxf(1.0f - cmin * cmin, 0.0f)) / cmin * 1.001f + 1e-6f;
        if (lane == 0) s_cone = make_float4(axis.x, axis.y, axis.z, tan_h);
        __syncwarp();
    }
    for (uint32_t b = 0; b < stg.nb; ++b) {
        // records no ray of the block can accept are dropped while staging; the walk sees the kept ones, compacted
        uint32_t keep;
        const uint32_t kept = stg.stage_culled(b, cfg, s_cone, keep);
        const uint32_t accepted = uniform ? forward_walk<DEG, true, COUNT, STD>(cfg, stg.rec, kept, ray, st)
                                          : forward_walk<DEG, false, COUNT, STD>(cfg, stg.rec, kept, ray, st);
        if (!COUNT && hit_masks) {
            // bit k of `accepted` = k-th kept record; the backward indexes the batch by list position
            const uint32_t lane = threadIdx.x & 31;
            const uint32_t comp = __reduce_or_sync(0xffffffffu, accepted);
            const bool mine = ((keep >> lane) & 1u) && ((comp >> __popc(keep & ((1u << lane) - 1u))) & 1u);
            const uint32_t m = __ballot_sync(0xffffffffu, mine);
            if (lane == 0) hit_masks[hit_mask_slot(b0, tile, b, block)] = m;
        }
        if (!__any_sync(0xffffffffu, st.alive[0] || st.alive[1])) break;   // the block saturated early
        stg.rotate();
    }
    stg.drain();
    if (COUNT) {
        unsigned long long t = st.ntests[0] + st.ntests[1], hh = st.nhits[0] + st.nhits[1];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { t += __shfl_xor_sync(0xffffffffu, t, o); hh += __shfl_xor_sync(0xffffffffu, hh, o); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(&counters[0], t); atomicAdd(&counters[1], hh); }
        return;
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t pix = ray.pix[h];
        if (ray.inside[h] && ray.valid[h]) {
            rgba[pix] = make_float4(h ? hi(st.cr) : lo(st.cr), h ? hi(st.cg) : lo(st.cg), h ? hi(st.cb) : lo(st.cb),
                                    1.0f - (h ? hi(st.T) : lo(st.T)));
            dist[pix] = h ? hi(st.D) : lo(st.D);
            if (cfg.enable_hitcounts) hits[pix] = (float)st.nhits[h];
        } else if (ray.inside[h]) {
            rgba[pix] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            dist[pix] = 1e06f;
            hits[pix] = 0.0f;
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// K5 backward (k_buffer_size == 0)
// ---------------------------------------------------------------------------------------------------

// Sum 16 values x 2 rays per lane over the warp.  Stage 0 adds the two rays of a lane and leaves value k in the low
// and value k+8 in the high half of 8 packed registers; stages 1-3 (lane xor 16, 8, 4) exchange the packed half of
// the values the partner lane will own (selects and shuffles per 32-bit half, adds packed); stage 4 (xor 2) splits
// the last packed register; stage 5 (xor 1) finishes.  16 shuffles in all.  On return lane l holds the warp total of
// value reduce_slot(l); lanes l and l^1 hold the same value.
GUT_D uint32_t reduce_slot(uint32_t lane) { return ((lane >> 2) & 7u) | ((lane & 2u) << 2); }
GUT_D F2 sel2(bool c, F2 a, F2 b) { F2 r; r.v = c ? a.v : b.v; return r; }
GUT_D F2 shfl2(F2 a, int m) { F2 r; r.v = __shfl_xor_sync(0xffffffffu, a.v, m); return r; }
GUT_D float warp_transpose_reduce(const F2 (&V)[16], uint32_t lane) {
    F2 W[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) W[k] = mk2(lo(V[k]) + hi(V[k]), lo(V[k + 8]) + hi(V[k + 8]));
    {
        const bool up = lane & 16;
#pragma unroll
        for (int k = 0; k < 4; ++k) W[k] = sel2(up, W[k + 4], W[k]) + shfl2(sel2(up, W[k], W[k + 4]), 16);
    }
    {
        const bool up = lane & 8;
#pragma unroll
        for (int k = 0; k < 2; ++k) W[k] = sel2(up, W[k + 2], W[k]) + shfl2(sel2(up, W[k], W[k + 2]), 8);
    }
    {
        const bool up = lane & 4;
        W[0] = sel2(up, W[1], W[0]) + shfl2(sel2(up, W[0], W[1]), 4);
    }
    const bool up = lane & 2;
    float x = (up ? hi(W[0]) : lo(W[0])) + __shfl_xor_sync(0xffffffffu, up ? lo(W[0]) : hi(W[0]), 2);
    x += __shfl_xor_sync(0xffffffffu, x, 1);
    return x;
}

constexpr uint32_t REDUCE_PITCH = 20;                           // floats per row
constexpr uint32_t REDUCE_TILE_BYTES = 32 * REDUCE_PITCH * 4;

// The same sum through a warp-private shared-memory tile, software-pipelined (the default; -DGUT_BWD_SHUFFLE_REDUCE selects
// the shuffle version above): a hit's rows are stored at the end of its iteration and summed during the NEXT hit's
// arithmetic, so that the store -> load -> add -> shuffle chain of one hit overlaps the FMA chain of the next instead of
// following its own -- the walk is bound by the length of its dependent chains at ~4.6 warps per scheduler.  44
// instructions instead of 71, and off the critical path: 3.27 -> 3.12 ms per view.  (Not pipelined, the tile version was
// SLOWER than the shuffles, 3.955 vs 3.91 ms at the time: store -> __syncwarp -> load is the longer chain.)
//  reduce_store : the two rays of a lane are added and the lane stores its row of 16 floats with four 128-bit stores (row
//                 pitch 20 floats, so the stores of a quarter warp cover all 32 banks exactly once);
//  reduce_load  : lane l owns the value pair p = l & 7 and the row group g = l >> 3 (8 rows): eight 64-bit loads (odd groups
//                 start 4 rows in, which moves their banks by 16: conflict free within a half warp);
//  reduce_finish: packed adds, then the four row groups are combined with two 64-bit shuffle rounds (xor 8, xor 16): every
//                 lane holds the warp totals of values 2p (low) and 2p + 1 (high).  Value 15 is padding.
GUT_D void reduce_store(const F2 (&V)[16], uint32_t lane, float* __restrict__ tile) {
    float4* row = reinterpret_cast<float4*>(tile + lane * REDUCE_PITCH);
    row[0] = make_float4(lo(V[0]) + hi(V[0]), lo(V[1]) + hi(V[1]), lo(V[2]) + hi(V[2]), lo(V[3]) + hi(V[3]));
    row[1] = make_float4(lo(V[4]) + hi(V[4]), lo(V[5]) + hi(V[5]), lo(V[6]) + hi(V[6]), lo(V[7]) + hi(V[7]));
    row[2] = make_float4(lo(V[8]) + hi(V[8]), lo(V[9]) + hi(V[9]), lo(V[10]) + hi(V[10]), lo(V[11]) + hi(V[11]));
    row[3] = make_float4(lo(V[12]) + hi(V[12]), lo(V[13]) + hi(V[13]), lo(V[14]) + hi(V[14]), 0.0f);
}
GUT_D F2 reduce_load(uint32_t lane, const float* __restrict__ tile) {   // sum over the lane's 8 rows of its value pair
    const uint32_t g = lane >> 3, rot = (g & 1u) * 4u;
    const float* col = tile + g * 8 * REDUCE_PITCH + 2 * (lane & 7u);
    F2 x[8];
#pragma unroll
    for (uint32_t i = 0; i < 8; ++i) x[i].v = *reinterpret_cast<const u64*>(col + ((i + rot) & 7u) * REDUCE_PITCH);
    return ((x[0] + x[1]) + (x[2] + x[3])) + ((x[4] + x[5]) + (x[6] + x[7]));
}
GUT_D F2 reduce_finish(F2 t) {
    t = t + shfl2(t, 8);
    t = t + shfl2(t, 16);
    return t;
}
// lane p < 8 adds values 2p and 2p + 1 of particle `idx`: one 64-bit vector reduction (value 15 is padding)
// (value 14 is zero when every ray starts at the frame origin: with14 = false skips its reduction)
GUT_D void reduce_commit(F2 total, uint32_t lane, uint32_t idx, float* __restrict__ raw_grad, bool with14) {
    if (lane < 8u) {
        float* dst = raw_grad + 16 * (size_t)idx + 2 * lane;
        if (lane < 7u) atomicAdd(reinterpret_cast<float2*>(dst), make_float2(lo(total), hi(total)));
        else if (with14) atomicAdd(dst, lo(total));
    }
}

struct BwdRay {
    F2 T_f, dA, dD;          // final transmittance of the forward pass, upstream gradients
    F2x3 dC;
    F2 T;                    // running transmittance of the re-walk
    F2x3 Crem;               // forward colour still to be composited behind the current hit
    bool alive[2];
    bool pending;            // the reduce tile holds the rows of a hit that is not summed yet ...
    uint32_t pending_idx;    // ... of this particle (warp-uniform)
};

// Per accepted hit a ray produces 15 values that are LINEAR in the ray, so they can be summed over rays first (the two
// rays of the lane, then the warp transpose-reduce, then one RED per value into the per-particle accumulator `raw`,
// 16 floats per particle) and turned into parameter gradients once per particle by K6 (gut_project.cu):
//   [0..2]  A   = dL/d go            (canonical ray origin)        -> dL/dpos   = -ms^T A
//   [3]     dL/d density
//   [4..6]  tau' = body-frame torque x (s_j s_k), direction half     -> dL/dquat  = rho q + 2 q (x) (0, tau)
//   [7]     tr  = sum_i (dL/dR^T)_ii, direction half                -> rho (radial part of dL/dquat, R(q) is a polynomial)
//   [8..10] Z_i = u_i B_i - s_i * (hit-distance term)  [+ go_i A_i]  -> dL/dscale_i = -Z_i / s_i
//   [11..13] dL/d colour
//   [14]    sum_i P_ii = go . A  (the direction half u . B vanishes)  -> rho
// with B = dL/d u.  This is the chain rule of processHitBwd (gaussianParticles.cuh:635-738) with the per-ray
// quaternion / matrix products (matmul_bw_quat, matmul_bw_vec) factored out of the ray loop.  tau, tr and Z each have a
// half that is go (x) A: with all rays starting at the frame origin go is a per-particle constant, so that half is
// go_frame (x) sum_rays(A) and K6 adds it from [0..2] (the "origin fold", gut_project.cu) -- 12 packed FP32 instructions
// per hit that the 64 rays of a block do not execute.  Only the per-ray remainder of go (non-UNIFORM rays) is added here.
// A ray of the lane that does not take the hit runs the same packed instructions with alpha = 0 and a zeroed
// upstream sum: every one of its 15 values is then exactly 0 and its state is unchanged.
template <int DEG, bool UNIFORM, bool STD>
GUT_D void backward_walk(const GutConfig& cfg, const float4* __restrict__ rec, const float4* __restrict__ recb, uint32_t mask,
                         const RayPair& ray, BwdRay& st, uint32_t lane, float* __restrict__ raw_grad, float* __restrict__ tile) {
    const float minT = thr_min_T<STD>(cfg);
    const F2 one = bc(1.0f);
    while (mask) {
      {
        const uint32_t j = __ffs(mask) - 1;
        mask &= mask - 1;
#ifndef GUT_BWD_SHUFFLE_REDUCE
        // rows of the previous hit (stored before the last __syncwarp): summed while this hit's arithmetic runs.  Two
        // registers stay live, and the two shuffle rounds are placed further down between independent stretches of it.
        F2 prev = reduce_load(lane, tile);
#endif
        const float4* r = rec + 5 * j;
        const float4 A = r[0], B = r[1], Cm = r[2], S = r[3];
        Canon2 k;
        bool acc0, acc1;
        particle_test<DEG, UNIFORM, STD>(cfg, A, B, Cm, S.w, ray, k, acc0, acc1);
        const bool h0 = acc0 && st.alive[0], h1 = acc1 && st.alive[1];  // no tMinMax test here
#ifndef GUT_BWD_SHUFFLE_REDUCE
        prev = prev + shfl2(prev, 8);
#endif

        // every lane runs the hit math (rays that do not take the hit are masked, see above): no lane-level branch
        F2 v[16];
        const float4 F = r[4];
        {
            const float4 IS = recb[2 * j], S2 = recb[2 * j + 1];   // 1 / scale, scale^2
            const F2x3& go = k.go;
            const F2x3& u = k.u;
            const F2 il = mk2(mufu_rsq(h0 ? lo(k.l2) : 1.0f), mufu_rsq(h1 ? hi(k.l2) : 1.0f));   // 1 / |u|
            const F2x3 grd = u * il;                   // normalised canonical direction
            const F2 dgo = dot(grd, go);
            // hit distance gdist = |s * grd * dot(grd, -go)| = |dgo| * |sg|, sg = s * grd.  hv = s^2 grd, sg2 = sg^2 = hv grd
            F2x3 hv, sg2;
            hv.x = bc(S2.x) * grd.x; hv.y = bc(S2.y) * grd.y; hv.z = bc(S2.z) * grd.z;
            sg2.x = hv.x * grd.x; sg2.y = hv.y * grd.y; sg2.z = hv.z * grd.z;
            const F2 ssum = sg2.x + sg2.y + sg2.z;
            const F2 gdist = mk2(h0 ? fabsf(lo(dgo)) * mufu_sqrt(lo(ssum)) : 0.0f, h1 ? fabsf(hi(dgo)) * mufu_sqrt(hi(ssum)) : 0.0f);
            const F2 alpha = mk2(h0 ? fminf(thr_max_alpha<STD>(cfg), lo(k.araw)) : 0.0f, h1 ? fminf(thr_max_alpha<STD>(cfg), hi(k.araw)) : 0.0f);
            const F2 T = st.T;
            const F2 w = alpha * T;
            const F2 om = one - alpha;
            const F2 nextT = om * T;
            const F2 invT = mk2(lo(nextT) <= minT ? 0.0f : mufu_rcp(lo(nextT)), hi(nextT) <= minT ? 0.0f : mufu_rcp(hi(nextT)));

            // hit-distance terms (gaussianParticles.cuh:536-567; the Jacobians of grds wrt grd / gro are the reference's
            // element-wise ones).  grds = -dgo sg;  dL/dgrds = grds hs = kap sg  with  kap = -dgo hs
            // The alpha gradient's hit-distance part is T * gdist * dD: the reference's residual distance is identically 0.
            // Its wrapper hands processHitBwd (hitT, hitTBackward, hitTGradient) where the callee declares (integratedDepth,
            // depth&, depthGrad) (shRadiativeGaussianParticles.cuh:369-406, gaussianParticles.cuh:499-501), so the residual is
            // max((0 - (D_f + w gdist)) / nextT, 0) = 0 for the forward's non-negative D_f.  Pinned on the reference's own
            // kernels by tests/test_gpu_reference.py.
            const F2 wd = w * st.dD;
            // gdist is 0 for a ray that does not take the hit, so "> 0" also is the hit test (the reference tests gdist^2 > 0)
            const F2 hs = mk2(lo(gdist) > 0.0f ? lo(wd) * mufu_rcp(lo(gdist)) : 0.0f, hi(gdist) > 0.0f ? hi(wd) * mufu_rcp(hi(gdist)) : 0.0f);
            const F2 nkap = dgo * hs;                  // -kap
            const F2 nd = nkap * dgo;                  // -kd, with s_i * sclHit_i = -kd * sg2_i
            // groHit_i = -kap sg2_i grd_i ; grdHit_i = -kap (sg2_i go_i + dgo hv_i)
            F2x3 mk;
            mk.x = nkap * sg2.x; mk.y = nkap * sg2.y; mk.z = nkap * sg2.z;

            // opacity and colour terms (:569-633)
            const F2 resT = mk2(lo(alpha) < 0.999999f ? lo(st.T_f) * mufu_rcp(lo(om)) : lo(T),
                                hi(alpha) < 0.999999f ? hi(st.T_f) * mufu_rcp(hi(om)) : hi(T));
            const F2 aDns = resT * st.dA;
            st.Crem.x = fma(-w, bc(F.x), st.Crem.x); st.Crem.y = fma(-w, bc(F.y), st.Crem.y); st.Crem.z = fma(-w, bc(F.z), st.Crem.z);
            // residual colour max(Crem / nextT, 0) = invT max(Crem, 0) (invT >= 0):  sum_c (F_c - resC_c) dC_c = F.dC - invT (max(Crem, 0).dC)
            const F2 mcx = mk2(fmaxf(lo(st.Crem.x), 0.0f), fmaxf(hi(st.Crem.x), 0.0f));
            const F2 mcy = mk2(fmaxf(lo(st.Crem.y), 0.0f), fmaxf(hi(st.Crem.y), 0.0f));
            const F2 mcz = mk2(fmaxf(lo(st.Crem.z), 0.0f), fmaxf(hi(st.Crem.z), 0.0f));
            const F2 fdc = fma(bc(F.z), st.dC.z, fma(bc(F.y), st.dC.y, bc(F.x) * st.dC.x));
            const F2 mdc = fma(mcz, st.dC.z, fma(mcy, st.dC.y, mcx * st.dC.x));
            const F2 cs = fma(-invT, mdc, fdc);
#ifndef GUT_BWD_SHUFFLE_REDUCE
            prev = prev + shfl2(prev, 16);
#endif
            const F2 Sall = fma(T, fma(gdist, st.dD, cs), aDns);
            const F2 Ssum = mk2(h0 ? lo(Sall) : 0.0f, h1 ? hi(Sall) : 0.0f);
            const F2 dDens = k.resp * Ssum;
            // g2 = |grd x go|^2 : adjoint of the cross product (:671-689); grd x go = c / |u|, so dL/d(grd x go) = c (2 g2G / |u|).
            // Degree 2: g2G = -resp respG / 2 with respG = density Ssum: the factors 2 and -1/2 cancel and resp Ssum is dDens
            F2 gscale;
            if (DEG == 2) gscale = -(bc(S.w) * dDens) * il;
            else {
                const F2 respG = bc(S.w) * Ssum;
                gscale = bc(2.0f) * mk2(kernel_response_grad<DEG>(lo(k.g2), lo(k.resp), lo(respG)), kernel_response_grad<DEG>(hi(k.g2), hi(k.resp), hi(respG))) * il;
            }
            const F2x3 gcG = k.c * gscale;
            const F2x3 grdG = cross(go, gcG);
            const F2x3 groG = cross(gcG, grd);
            // A = dL/dgo (total), X = dL/dgrd (total)
            F2x3 Ag, X, Bg;
            Ag.x = fma(mk.x, grd.x, groG.x); Ag.y = fma(mk.y, grd.y, groG.y); Ag.z = fma(mk.z, grd.z, groG.z);
            X.x = fma(mk.x, go.x, fma(nd, hv.x, grdG.x)); X.y = fma(mk.y, go.y, fma(nd, hv.y, grdG.y)); X.z = fma(mk.z, go.z, fma(nd, hv.z, grdG.z));
            // normalisation adjoint (safe_normalize_bw, mathUtils.cuh:420-430): B = dL/du = Y / |u| with Y = X - grd (grd . X).
            // Every use of B below but one is a product u_i B_j = grd_i Y_j, so Y is used directly and only the trace term
            // is divided by |u|.
            const F2 gx = dot(grd, X);
            Bg.x = fma(-grd.x, gx, X.x); Bg.y = fma(-grd.y, gx, X.y); Bg.z = fma(-grd.z, gx, X.z);   // = Y

            // P_ij = go_i A_j + u_i B_j = go_i A_j + grd_i Y_j : diagonal -> scale, off-diagonal -> body-frame torque,
            // tau_x = (s_z/s_y) P_zy - (s_y/s_z) P_yz = [(Y x hv)_x + (A x s^2 go)_x] / (s_y s_z), and cyclic.
            // The go (x) A half is linear in A with a per-PARTICLE factor when every ray starts at the frame origin, and A is
            // summed over the rays anyway (v[0..2]): K6 adds go_frame (x) sum(A) once per particle (gut_project.cu), and so
            // for the trace and the scale terms; here only the part of go that differs per ray (dl = ms (ray origin - frame
            // origin), non-UNIFORM rays) is left.  K6 also applies the 1 / (s_j s_k) of the torque.
            const F2x3 tq = cross(Bg, hv);
            v[0] = (Ag.x); v[1] = (Ag.y); v[2] = (Ag.z);
            v[3] = (dDens);
            v[4] = tq.x; v[5] = tq.y; v[6] = tq.z;
            F2 tr = fma(bc(IS.z) * ray.d.z, Bg.z, fma(bc(IS.y) * ray.d.y, Bg.y, (bc(IS.x) * ray.d.x) * Bg.x)) * il;   // sum_i B_i d_i / s_i
            const F2 kd = -nd;
            v[8] = fma(kd, sg2.x, grd.x * Bg.x);
            v[9] = fma(kd, sg2.y, grd.y * Bg.y);
            v[10] = fma(kd, sg2.z, grd.z * Bg.z);
            F2 v14 = bc(0.0f);
            if (!UNIFORM) {
                // dl = ms (ray origin - frame origin) = go - go_frame (the rounding of the difference is that of go itself)
                F2x3 dl, gd;
                dl.x = go.x - bc(A.w); dl.y = go.y - bc(B.w); dl.z = go.z - bc(Cm.w);
                gd.x = bc(S2.x) * dl.x; gd.y = bc(S2.y) * dl.y; gd.z = bc(S2.z) * dl.z;
                const F2x3 ta = cross(Ag, gd);
                v[4] = v[4] + ta.x; v[5] = v[5] + ta.y; v[6] = v[6] + ta.z;
                tr = fma(bc(IS.z), Ag.z * ray.dlt.z, fma(bc(IS.y), Ag.y * ray.dlt.y, fma(bc(IS.x), Ag.x * ray.dlt.x, tr)));
                v[8] = fma(dl.x, Ag.x, v[8]); v[9] = fma(dl.y, Ag.y, v[9]); v[10] = fma(dl.z, Ag.z, v[10]);
                v14 = dot(dl, Ag);
            }
            v[7] = (tr);
            v[11] = (w * st.dC.x); v[12] = (w * st.dC.y); v[13] = (w * st.dC.z);
            // [14]: sum_i P_ii = go . A + grd . Y, and grd . Y = 0 (Y is X minus its component along the unit vector grd): only
            // the part of go . A that differs per ray is left; K6's origin fold adds go_frame . sum(A)
            v[14] = UNIFORM ? bc(0.0f) : v14;
            v[15] = bc(0.0f);

            st.T = nextT;
            if (lo(nextT) < minT) st.alive[0] = false;
            if (hi(nextT) < minT) st.alive[1] = false;
        }
#ifndef GUT_BWD_SHUFFLE_REDUCE
        if (st.pending) reduce_commit(prev, lane, st.pending_idx, raw_grad, !UNIFORM);
        __syncwarp();   // every lane has read the previous hit's rows
        reduce_store(v, lane, tile);
        __syncwarp();   // this hit's rows are visible to the next iteration's loads
        st.pending = true;
        st.pending_idx = __float_as_uint(F.w);
#else
        const float total = warp_transpose_reduce(v, lane);
        const uint32_t slot = reduce_slot(lane);
        if ((lane & 1u) == 0u && slot < 15u) atomicAdd(raw_grad + 16 * (size_t)__float_as_uint(F.w) + slot, total);
#endif
      }
    }
}

// UNIFORM (every ray of the block starts at the frame origin) is a launch-level template parameter: the kernel is launched
// once per variant and a block leaves at once when its rays are of the other kind.  Keeping the per-ray origin offsets
// out of the common variant's live set is worth 0.27 ms per view at the 96-register cap; the idle launch costs ~10 us.
template <int DEG, bool UNIFORM, bool STD>
__global__ void __launch_bounds__(RENDER_THREADS, GUT_BWD_MIN_BLOCKS) gut_render_backward_kernel(
    GutConfig cfg, GutFrame fr, const uint32_t* __restrict__ tile_start, GutParticleSource src,
    const float* __restrict__ ray_ori, const float* __restrict__ ray_dir,
    const float4* __restrict__ rgba, const float4* __restrict__ rgba_grad, const float* __restrict__ dist,
    const float* __restrict__ dist_grad, float* __restrict__ raw_grad, const uint32_t* __restrict__ hit_masks) {
    __shared__ __align__(16) unsigned char smem_raw[StageLayout<true>::TOTAL];
#ifndef GUT_BWD_SHUFFLE_REDUCE
    __shared__ __align__(16) float reduce_tile[REDUCE_TILE_BYTES / 4];
#else
    float* const reduce_tile = nullptr;
#endif

    // CTAs are dispatched in blockIdx order: tiles longest list first (gut_tile_scan_kernel leaves the order behind tile_start)
    const uint32_t tile = __ldg(tile_start + fr.num_tiles + 1 + blockIdx.x / BLOCKS_PER_TILE), block = blockIdx.x % BLOCKS_PER_TILE;
    const uint32_t lane = threadIdx.x & 31;
    bool at_origin;
    const RayPair ray = make_ray_pair(fr, ray_ori, ray_dir, tile, block, at_origin);
    if (__all_sync(0xffffffffu, at_origin) != UNIFORM) return;

    // initializeBackwardRay (rayPayloadBackward.cuh:46-55)
    BwdRay st;
    float Tf[2], dA[2], dD[2];
    float3 Cf[2], dC[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        st.alive[h] = ray.inside[h] && ray.valid[h];
        Tf[h] = 1.0f; dA[h] = 0.0f; dD[h] = 0.0f;
        Cf[h] = f3(0.f, 0.f, 0.f); dC[h] = f3(0.f, 0.f, 0.f);
        if (st.alive[h]) {
            const float4 fd = __ldg(rgba + ray.pix[h]), gd = __ldg(rgba_grad + ray.pix[h]);
            Tf[h] = 1.0f - fd.w;       // transmittanceBackward
            dA[h] = gd.w;              // transmittanceGradient = -dA; used as residualTrm * -(-dA)
            dD[h] = __ldg(dist_grad + ray.pix[h]);
            Cf[h] = f3(fd.x, fd.y, fd.z);
            dC[h] = f3(gd.x, gd.y, gd.z);
        }
    }
    st.T_f = mk2(Tf[0], Tf[1]); st.dA = mk2(dA[0], dA[1]); st.dD = mk2(dD[0], dD[1]);
    st.Crem.x = mk2(Cf[0].x, Cf[1].x); st.Crem.y = mk2(Cf[0].y, Cf[1].y); st.Crem.z = mk2(Cf[0].z, Cf[1].z);
    st.dC.x = mk2(dC[0].x, dC[1].x); st.dC.y = mk2(dC[0].y, dC[1].y); st.dC.z = mk2(dC[0].z, dC[1].z);
    st.T = bc(1.0f);
    st.pending = false; st.pending_idx = 0u;

    const uint32_t b0 = tile_start[tile], b1 = tile_start[tile + 1];
    WarpStager<true> stg;
    stg.init(smem_raw, src, b0, b1 - b0, fr.cam_centre);

    for (uint32_t b = 0; b < stg.nb; ++b) {
        stg.stage(b);
        const uint32_t mask = __ldg(hit_masks + hit_mask_slot(b0, tile, b, block));   // entries some ray of the block accepted in the forward
        backward_walk<DEG, UNIFORM, STD>(cfg, stg.rec, stg.recb, mask, ray, st, lane, raw_grad, reduce_tile);
        if (!__any_sync(0xffffffffu, st.alive[0] || st.alive[1])) break;
        stg.rotate();
    }
#ifndef GUT_BWD_SHUFFLE_REDUCE
    if (st.pending) reduce_commit(reduce_finish(reduce_load(lane, reduce_tile)), lane, st.pending_idx, raw_grad, !UNIFORM);   // the last hit
#endif
    stg.drain();
}

template <int DEG>
cudaError_t launch_render_deg(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start, GutParticleSource src,
                              const float* ray_ori, const float* ray_dir, float* rgba, float* dist, float* hits,
                              unsigned long long* counters, uint32_t* hit_masks, bool cull, cudaStream_t stream) {
    const uint32_t grid = frame.num_tiles * BLOCKS_PER_TILE;
    // cull == false (gut_set_block_cull(h, 0), a test control) walks every record: proves that the cull changes no output bit.
    // The kernel applies the cone bound only for DEG == 2 (block_cull's g2max is the degree-2 response bound).
    if (counters)
        gut_render_kernel<DEG, true, false><<<grid, RENDER_THREADS, 0, stream>>>(cfg, frame, tile_start, src, ray_ori, ray_dir, nullptr, nullptr,
                                                                              nullptr, counters, nullptr, cull);
    else if (DEG == 2 && thresholds_are_standard(cfg))   // the reference's defaults as immediates (default kernel degree only)
        gut_render_kernel<DEG, false, DEG == 2><<<grid, RENDER_THREADS, 0, stream>>>(cfg, frame, tile_start, src, ray_ori, ray_dir,
                                                                                  reinterpret_cast<float4*>(rgba), dist, hits, nullptr, hit_masks, cull);
    else
        gut_render_kernel<DEG, false, false><<<grid, RENDER_THREADS, 0, stream>>>(cfg, frame, tile_start, src, ray_ori, ray_dir,
                                                                               reinterpret_cast<float4*>(rgba), dist, hits, nullptr, hit_masks, cull);
    return cudaGetLastError();
}

template <int DEG>
cudaError_t launch_render_bwd_deg(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start, GutParticleSource src,
                                  const float* ray_ori, const float* ray_dir, const float* rgba,
                                  const float* rgba_grad, const float* dist, const float* dist_grad, float* raw_grad,
                                  const uint32_t* hit_masks, cudaStream_t stream) {
    const uint32_t grid = frame.num_tiles * BLOCKS_PER_TILE;
    if (DEG == 2 && thresholds_are_standard(cfg)) {
        gut_render_backward_kernel<DEG, true, DEG == 2><<<grid, RENDER_THREADS, 0, stream>>>(
            cfg, frame, tile_start, src, ray_ori, ray_dir, reinterpret_cast<const float4*>(rgba),
            reinterpret_cast<const float4*>(rgba_grad), dist, dist_grad, raw_grad, hit_masks);
        gut_render_backward_kernel<DEG, false, DEG == 2><<<grid, RENDER_THREADS, 0, stream>>>(
            cfg, frame, tile_start, src, ray_ori, ray_dir, reinterpret_cast<const float4*>(rgba),
            reinterpret_cast<const float4*>(rgba_grad), dist, dist_grad, raw_grad, hit_masks);
    } else {
        gut_render_backward_kernel<DEG, true, false><<<grid, RENDER_THREADS, 0, stream>>>(
            cfg, frame, tile_start, src, ray_ori, ray_dir, reinterpret_cast<const float4*>(rgba),
            reinterpret_cast<const float4*>(rgba_grad), dist, dist_grad, raw_grad, hit_masks);
        gut_render_backward_kernel<DEG, false, false><<<grid, RENDER_THREADS, 0, stream>>>(
            cfg, frame, tile_start, src, ray_ori, ray_dir, reinterpret_cast<const float4*>(rgba),
            reinterpret_cast<const float4*>(rgba_grad), dist, dist_grad, raw_grad, hit_masks);
    }
    return cudaGetLastError();
}

}  // namespace

#define GUT_DISPATCH_DEGREE(deg, CALL)                 \
    switch (deg) {                                     \
    case 0: return CALL(0);                            \
    case 1: return CALL(1);                            \
    case 2: return CALL(2);                            \
    case 3: return CALL(3);                            \
    case 4: return CALL(4);                            \
    case 5: return CALL(5);                            \
    case 8: return CALL(8);                            \
    default: return cudaErrorInvalidValue;             \
    }

cudaError_t gut_launch_render(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start, GutParticleSource src,
                              const float* ray_ori, const float* ray_dir, float* rgba, float* dist, float* hits,
                              unsigned long long* counters, uint32_t* hit_masks, bool block_cull, cudaStream_t stream) {
#define CALL_FWD(D) launch_render_deg<D>(cfg, frame, tile_start, src, ray_ori, ray_dir, rgba, dist, hits, counters, hit_masks, block_cull, stream)
    GUT_DISPATCH_DEGREE(cfg.kernel_degree, CALL_FWD)
#undef CALL_FWD
}

cudaError_t gut_launch_render_backward(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start,
                                       GutParticleSource src, const float* ray_ori, const float* ray_dir,
                                       const float* rgba, const float* rgba_grad, const float* dist, const float* dist_grad,
                                       float* raw_grad, const uint32_t* hit_masks, cudaStream_t stream) {
#define CALL_BWD(D) \
    launch_render_bwd_deg<D>(cfg, frame, tile_start, src, ray_ori, ray_dir, rgba, rgba_grad, dist, dist_grad, raw_grad, hit_masks, stream)
    GUT_DISPATCH_DEGREE(cfg.kernel_degree, CALL_BWD)
#undef CALL_BWD
}

size_t gut_hit_mask_words(uint64_t num_intersections, uint32_t num_tiles) {
    return ((size_t)(num_intersections >> 5) + num_tiles + 2) * BLOCKS_PER_TILE;
}
