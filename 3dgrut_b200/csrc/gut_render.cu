// gut_render.cu -- K4 (per-tile forward compositing) and K5 (backward) of the 3DGUT rasteriser.
//
// reference: render / renderBackward kernels (include/3dgut/kernels/cuda/renderers/gutRenderer.cuh:83-165),
//   GUTKBufferRenderer::evalKBuffer / evalBackwardNoKBuffer (renderers/gutKBufferRenderer.cuh:217-386),
//   ray set-up initializeRay / initializeBackwardRay (common/rayPayload.cuh:75-129, rayPayloadBackward.cuh:30-58),
//   hit math gaussianParticle.hit / integrateHit (kernels/slang/models/gaussianParticles.slang:96-254),
//   per-hit gradient processHitBwd (kernels/cuda/models/gaussianParticles.cuh:480-738).
//
// B200 design (DESIGN.md "K4/K5").  Both kernels are bound by instruction issue (ncu: sm__throughput > 80 %, DRAM
// < 2 %), so the design minimises SASS instructions per ray x particle test:
//  * one CTA per 16x16 tile, 8 warps, each warp owns a compact 8x4 pixel block (more lanes hit the same
//    particle than with the reference's 16x2 rows -> better SIMT efficiency in the divergent hit branch);
//  * staging runs one batch (256 particles) ahead of the walk and costs the walk nothing: the batch's slice of the
//    sorted particle list is contiguous and arrives by cp.async.bulk (TMA bulk copy, UBLKCP) on an mbarrier; every
//    thread then gathers ONE particle's parameters (3 x 16 B + its K1 colour) with cp.async (LDGSTS, no registers
//    held across the walk) and, when the batch is due, expands them once into an 80-byte shared-memory record that
//    all 256 rays of the tile read by broadcast -- quaternion->matrix math happens once per (tile, particle), not
//    per ray.  DRAM sits at ~2 % during compositing, so the gathers are free; writing the records to HBM from the
//    sort kernel instead (tried first) cost 1 ms per view;
//  * records carry ms = diag(1/s) R^T and go = ms (frame origin - position).  When every ray of the tile starts
//    at the frame origin (true for pinhole and fisheye bundles; checked per CTA, bitwise) the canonical ray origin
//    is go itself and a test is ONE mat-vec, a cross product, one MUFU.RCP and one MUFU.EX2; otherwise go is
//    corrected by ms (ray origin - frame origin).  Normalisation and the hit distance are computed for accepted
//    hits only, from the unnormalised direction: t = |u.go| sqrt(|s*u|^2) / |u|^2;
//  * the translation unit is built with -use_fast_math like the reference (setup_3dgut.py:83): FTZ, MUFU
//    division / sqrt / exp2;
//  * backward: lanes that hit compute the 14 per-particle gradient values; the warp sums them with a 16-shuffle
//    transpose-reduce (not 14 x 5 butterflies) and issues ONE red.global.add.f32 instruction per (warp, particle)
//    -- and nothing at all when no lane of the warp hit (the reference reduces and does 14 atomics regardless,
//    shRadiativeGaussianParticles.cuh:409-482 called outside the isAlive branch at gutKBufferRenderer.cuh:379-383).
#include "gut_internal.h"

namespace {

constexpr int RENDER_THREADS = 256;
constexpr int BATCH = 256;  // particles per staged batch = one per thread

// ---- mbarrier / bulk-copy PTX -------------------------------------------------------------------------
GUT_D uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
GUT_D void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
GUT_D void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
GUT_D void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
GUT_D void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
GUT_D void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// ---- generalized-Gaussian kernel response and its derivative wrt the squared canonical distance ---------
// reference: canonicalRayMaxKernelResponse (gaussianParticles.slang:119-163), particleResponseGrd (gaussianParticles.cuh:219-261)
template <int DEG>
GUT_D float kernel_response(float g) {
    if (DEG == 8) { const float g2 = g * g; return __expf(-0.000685871056241f * g2 * g2); }
    if (DEG == 5) return __expf(-0.0185185185185f * g * g * sqrtf(g));
    if (DEG == 4) return __expf(-0.0555555555556f * g * g);
    if (DEG == 3) return __expf(-0.166666666667f * g * sqrtf(g));
    if (DEG == 1) return __expf(-1.5f * sqrtf(g));
    if (DEG == 0) return fmaxf(1.0f + -0.329630334487f * sqrtf(g), 0.0f);
    return __expf(-0.5f * g);
}
template <int DEG>
GUT_D float kernel_response_grad(float g, float resp, float resp_grad) {
    if (DEG == 8) return (-0.000685871056241f * 4.0f) * g * g * g * resp * resp_grad;
    if (DEG == 5) return (-0.0185185185185f * 2.5f) * g * sqrtf(g) * resp * resp_grad;
    if (DEG == 4) return (-0.0555555555556f * 2.0f) * g * resp * resp_grad;
    if (DEG == 3) return (-0.166666666667f * 1.5f) * sqrtf(g) * resp * resp_grad;
    if (DEG == 1) return -0.75f * sqrtf(g) * resp * resp_grad;
    if (DEG == 0) return resp > 0.0f ? (0.5f * -0.329630334487f * rsqrtf(g)) * resp_grad : 0.0f;
    return -0.5f * resp * resp_grad;
}


// ---- single-instruction MUFU wrappers (flush-to-zero, approximate: what -use_fast_math gives the reference) ----
GUT_D float mufu_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
GUT_D float mufu_ex2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
GUT_D float mufu_rsq(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
GUT_D float mufu_sqrt(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

struct Ray {
    float3 o, d;
    float tmin, tmax;
    bool valid;
};

// slab test against the +-1e6 scene box (BoundingBox::ray_intersect, include/3dgut/utils/bounding_box.h:79-128)
GUT_D void box_intersect(float3 o, float3 d, float& tmin, float& tmax) {
    const float lo = -1e06f, hi = 1e06f;
    float t0 = (lo - o.x) / d.x, t1 = (hi - o.x) / d.x;
    if (t0 > t1) { const float t = t0; t0 = t1; t1 = t; }
    float y0 = (lo - o.y) / d.y, y1 = (hi - o.y) / d.y;
    if (y0 > y1) { const float t = y0; y0 = y1; y1 = t; }
    if (t0 > y1 || y0 > t1) { tmin = 3.402823466e+38f; tmax = 3.402823466e+38f; return; }
    if (y0 > t0) t0 = y0;
    if (y1 < t1) t1 = y1;
    float z0 = (lo - o.z) / d.z, z1 = (hi - o.z) / d.z;
    if (z0 > z1) { const float t = z0; z0 = z1; z1 = t; }
    if (t0 > z1 || z0 > t1) { tmin = 3.402823466e+38f; tmax = 3.402823466e+38f; return; }
    if (z0 > t0) t0 = z0;
    if (z1 < t1) t1 = z1;
    tmin = t0; tmax = t1;
}

GUT_D Ray make_ray(const GutFrame& fr, const float* __restrict__ ray_ori, const float* __restrict__ ray_dir, uint32_t pix) {
    Ray r;
    const float3 o = f3(__ldg(ray_ori + 3 * (size_t)pix), __ldg(ray_ori + 3 * (size_t)pix + 1), __ldg(ray_ori + 3 * (size_t)pix + 2));
    const float3 d = f3(__ldg(ray_dir + 3 * (size_t)pix), __ldg(ray_dir + 3 * (size_t)pix + 1), __ldg(ray_dir + 3 * (size_t)pix + 2));
    r.o = rotate_cols(fr.c2w.c[0], fr.c2w.c[1], fr.c2w.c[2], o) + fr.c2w.c[3];
    r.d = rotate_cols(fr.c2w.c[0], fr.c2w.c[1], fr.c2w.c[2], d);
    box_intersect(r.o, r.d, r.tmin, r.tmax);
    r.tmin = fmaxf(r.tmin, 0.0f);
    r.valid = r.tmax > r.tmin;
    return r;
}

// pixel owned by this thread: warp w -> 8x4 block (w&1, w>>1), lane -> (lane&7, lane>>3)
GUT_D void thread_pixel(const GutFrame& fr, uint32_t tile, uint32_t& px, uint32_t& py) {
    const uint32_t tx = tile % fr.grid_x, ty = tile / fr.grid_x;
    const uint32_t w = threadIdx.x >> 5, l = threadIdx.x & 31;
    px = tx * GUT_TILE + (w & 1) * 8 + (l & 7);
    py = ty * GUT_TILE + (w >> 1) * 4 + (l >> 3);
}

// The records of a batch are walked four at a time; the tail of a tile's last batch is padded in shared memory with
// records that can never be accepted (density 0), so the inner loop has no remainder handling.
constexpr uint32_t WALK_UNROLL = 4;
constexpr uint32_t ALIVE_CHECK = 16;  // records between two warp-wide "is any ray still alive" votes

// ---- packed FP32 (sm_100a FFMA2 / FMUL2: two fp32 lanes per instruction, one issue slot) --------------------
// Both compositing kernels are issue bound, and a packed instruction does the work of two scalar ones for the same
// issue slot (scripts/microbench/ffma2.cu: 4 FFMA2 + 8 IADD take 0.75x the time of 8 FFMA + 8 IADD).  The test
// therefore runs on TWO particles at once: records are stored pair-interleaved so that one LDS.128 yields two packed
// operands, and the ray direction enters as a scalar broadcast operand (ptxas folds mov.b64 {x,x} into R.F32).
typedef unsigned long long u64;
GUT_D u64 pk2(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
GUT_D void upk2(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
GUT_D float half_of(u64 v, int h) { float lo, hi; upk2(v, lo, hi); return h ? hi : lo; }
GUT_D u64 bc2(float x) { return pk2(x, x); }
GUT_D u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
GUT_D u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
// pure sign flip (neg.f32 without .ftz), so that ptxas folds it into the consumer's operand modifier
GUT_D u64 neg2(u64 a) {
    float lo, hi, nl, nh;
    upk2(a, lo, hi);
    asm("neg.f32 %0, %1;" : "=f"(nl) : "f"(lo));
    asm("neg.f32 %0, %1;" : "=f"(nh) : "f"(hi));
    return pk2(nl, nh);
}

// Pair-interleaved record: particles (2i, 2i+1) of a batch share 10 quads (160 B).  A vector X = (x, y, z, w) of the
// two particles is stored as (x0, x1, y0, y1) (z0, z1, w0, w1); quads 0-1 = row 0 of ms + go.x, 2-3 = row 1 + go.y,
// 4-5 = row 2 + go.z, 6-7 = scale + density, 8 / 9 = clamped colour + particle index of particle 0 / 1.
constexpr uint32_t PAIR_QUADS = 10;
GUT_D void store_pair_vec(float* pair, uint32_t v, uint32_t h, float x, float y, float z, float w) {
    float* q = pair + 8 * v + h;
    q[0] = x; q[2] = y; q[4] = z; q[6] = w;
}

// canonical-frame test shared by forward and backward, two particles at a time.  u = ms d (unnormalised canonical
// direction), go = canonical origin, c = u x go.
struct PairCanon {
    u64 gox, goy, goz, ux, uy, uz, cx, cy, cz, l2, il2, g2, resp, araw;
};
struct Canon {
    float3 go, u, c;
    float l2, il2, g2, resp, araw;
};
GUT_D Canon canon_half(const PairCanon& k, int h) {
    Canon c;
    c.go = f3(half_of(k.gox, h), half_of(k.goy, h), half_of(k.goz, h));
    c.u = f3(half_of(k.ux, h), half_of(k.uy, h), half_of(k.uz, h));
    c.c = f3(half_of(k.cx, h), half_of(k.cy, h), half_of(k.cz, h));
    c.l2 = half_of(k.l2, h); c.il2 = half_of(k.il2, h); c.g2 = half_of(k.g2, h);
    c.resp = half_of(k.resp, h); c.araw = half_of(k.araw, h);
    return c;
}
template <int DEG, bool UNIFORM>
GUT_D void pair_test(const GutConfig& cfg, const ulonglong2* __restrict__ r, const float3 d, const float3 dlt, PairCanon& k,
                     bool& accept0, bool& accept1) {
    const ulonglong2 a0 = r[0], a1 = r[1], b0 = r[2], b1 = r[3], c0 = r[4], c1 = r[5];
    const u64 dens = r[7].y;
    const u64 dx = bc2(d.x), dy = bc2(d.y), dz = bc2(d.z);
    k.gox = a1.y; k.goy = b1.y; k.goz = c1.y;
    if (!UNIFORM) {  // the ray does not start at the frame origin: go += ms (ray origin - frame origin)
        const u64 ex = bc2(dlt.x), ey = bc2(dlt.y), ez = bc2(dlt.z);
        k.gox = fma2(a1.x, ez, fma2(a0.y, ey, fma2(a0.x, ex, k.gox)));
        k.goy = fma2(b1.x, ez, fma2(b0.y, ey, fma2(b0.x, ex, k.goy)));
        k.goz = fma2(c1.x, ez, fma2(c0.y, ey, fma2(c0.x, ex, k.goz)));
    }
    k.ux = fma2(a1.x, dz, fma2(a0.y, dy, mul2(a0.x, dx)));
    k.uy = fma2(b1.x, dz, fma2(b0.y, dy, mul2(b0.x, dx)));
    k.uz = fma2(c1.x, dz, fma2(c0.y, dy, mul2(c0.x, dx)));
    k.l2 = fma2(k.uz, k.uz, fma2(k.uy, k.uy, mul2(k.ux, k.ux)));
    k.cx = fma2(k.uy, k.goz, neg2(mul2(k.uz, k.goy)));
    k.cy = fma2(k.uz, k.gox, neg2(mul2(k.ux, k.goz)));
    k.cz = fma2(k.ux, k.goy, neg2(mul2(k.uy, k.gox)));
    float l0, l1;
    upk2(k.l2, l0, l1);
    k.il2 = pk2(mufu_rcp(l0), mufu_rcp(l1));
    k.g2 = mul2(fma2(k.cz, k.cz, fma2(k.cy, k.cy, mul2(k.cx, k.cx))), k.il2);  // |normalize(u) x go|^2
    float r0, r1;
    if (DEG == 2) {
        float e0, e1;
        upk2(mul2(k.g2, bc2(-0.72134752044448170368f)), e0, e1);
        r0 = mufu_ex2(e0); r1 = mufu_ex2(e1);
    } else {
        float g0, g1;
        upk2(k.g2, g0, g1);
        r0 = kernel_response<DEG>(g0); r1 = kernel_response<DEG>(g1);
    }
    k.resp = pk2(r0, r1);
    k.araw = mul2(k.resp, dens);
    float w0, w1;
    upk2(k.araw, w0, w1);
    // alpha = min(max_alpha, araw) > min_alpha  <=>  araw > min_alpha  (max_alpha > min_alpha is checked at gut_create)
    accept0 = (r0 > cfg.min_kernel_response) && (w0 > cfg.min_alpha);
    accept1 = (r1 > cfg.min_kernel_response) && (w1 > cfg.min_alpha);
}

// ---------------------------------------------------------------------------------------------------
// K4 forward
// ---------------------------------------------------------------------------------------------------
struct FwdState {
    float T, D, cr, cg, cb;
    uint32_t nhits, ntests;
    bool alive;
};

template <int DEG, bool UNIFORM, bool COUNT>
GUT_D void forward_walk(const GutConfig& cfg, const float4* __restrict__ rec, uint32_t cnt, const Ray& ray, const float3 dlt, FwdState& st) {
    const uint32_t cnt_pad = (cnt + WALK_UNROLL - 1) & ~(WALK_UNROLL - 1);
    for (uint32_t j0 = 0; j0 < cnt_pad; j0 += ALIVE_CHECK) {
        if (!__any_sync(0xffffffffu, st.alive)) break;
        const uint32_t je = min(cnt_pad, j0 + ALIVE_CHECK);
        for (uint32_t j = j0; j < je; j += WALK_UNROLL) {
#pragma unroll
            for (uint32_t q = 0; q < WALK_UNROLL; q += 2) {
                const float4* r = rec + PAIR_QUADS * ((j + q) >> 1);
                PairCanon pk;
                bool acc[2];
                pair_test<DEG, UNIFORM>(cfg, reinterpret_cast<const ulonglong2*>(r), ray.d, dlt, pk, acc[0], acc[1]);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    if (COUNT && st.alive && (j + q + h) < cnt) st.ntests++;
                    if (acc[h] && st.alive) {
                        // hit distance |s * grd * dot(grd, -go)| with grd = u / |u|  (canonicalRayDistance, gaussianParticles.slang:165-171)
                        const Canon k = canon_half(pk, h);
                        const float4 Sxy = r[6], Szd = r[7];
                        const float3 su = f3((h ? Sxy.y : Sxy.x) * k.u.x, (h ? Sxy.w : Sxy.z) * k.u.y, (h ? Szd.y : Szd.x) * k.u.z);
                        const float t = fabsf(dot(k.u, k.go)) * k.il2 * mufu_sqrt(dot(su, su));
                        if (t > ray.tmin && t < ray.tmax) {
                            const float alpha = fminf(cfg.max_alpha, k.araw);
                            const float4 F = r[8 + h];
                            const float w = alpha * st.T;
                            st.D = fmaf(t, w, st.D);
                            st.T *= (1.0f - alpha);
                            if (w > 0.0f) { st.cr = fmaf(F.x, w, st.cr); st.cg = fmaf(F.y, w, st.cg); st.cb = fmaf(F.z, w, st.cb); st.nhits++; }
                            if (st.T < cfg.min_transmittance) st.alive = false;
                        }
                    }
                }
            }
        }
    }
}

// ---- batch staging shared by K4 and K5 --------------------------------------------------------------------
GUT_D void cp_async16(void* sdst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(sdst)), "l"(gsrc) : "memory");
}
GUT_D void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
GUT_D void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

constexpr uint32_t IDX_STAGE_WORDS = BATCH + 4;   // a batch's slice of the sorted list, widened to 16-byte alignment

template <bool BWD>
struct StageLayout {
    static constexpr uint32_t REC = BATCH * sizeof(GutRecord);
    static constexpr uint32_t RECB = BWD ? BATCH * sizeof(GutRecordBwd) : 0;
    static constexpr uint32_t RAW = BATCH * 4 * sizeof(float4);
    static constexpr uint32_t IDX = 2 * IDX_STAGE_WORDS * sizeof(uint32_t);
    static constexpr uint32_t TOTAL = REC + RECB + RAW + IDX;
};

// Per-CTA staging pipeline (see the header comment).  Batch bi of the tile is: idx slice by TMA (issued two batches
// ahead by thread 0), raw parameter gather by cp.async (issued one batch ahead by every thread), expansion into
// rec/recb right before the walk.
template <bool BWD>
struct TileStager {
    float4* rec; float4* recb; float4* raw; uint32_t* idx; uint64_t* bar;
    GutParticleSource src;
    uint32_t b0, n, nb;
    float3 origin;
    uint32_t p_next;   // particle this thread gathers for the batch in flight

    GUT_D void init(unsigned char* smem, uint64_t* bars, const GutParticleSource& s, uint32_t b0_, uint32_t n_, float3 origin_) {
        rec = reinterpret_cast<float4*>(smem);
        recb = reinterpret_cast<float4*>(smem + StageLayout<BWD>::REC);
        raw = reinterpret_cast<float4*>(smem + StageLayout<BWD>::REC + StageLayout<BWD>::RECB);
        idx = reinterpret_cast<uint32_t*>(smem + StageLayout<BWD>::REC + StageLayout<BWD>::RECB + StageLayout<BWD>::RAW);
        bar = bars; src = s; b0 = b0_; n = n_; nb = (n_ + BATCH - 1) / BATCH; origin = origin_; p_next = 0;
    }
    GUT_D uint32_t count(uint32_t bi) const { return min((uint32_t)BATCH, n - bi * BATCH); }
    // thread 0: bulk-copy the 16-byte aligned superset of sorted_idx[b0 + bi*BATCH, +count)
    GUT_D void issue_idx(uint32_t bi) {
        const uint32_t g = b0 + bi * BATCH, a = g & ~3u;
        const uint32_t bytes = (((g - a) + count(bi)) * 4u + 15u) & ~15u;
        mbar_expect_tx(&bar[bi & 1], bytes);
        bulk_g2s(idx + (bi & 1) * IDX_STAGE_WORDS, src.sorted_idx + a, bytes, &bar[bi & 1]);
    }
    // every thread: gather the parameters of its particle of batch bi into the raw planes
    GUT_D void issue_gather(uint32_t bi) {
        mbar_wait(&bar[bi & 1], (bi >> 1) & 1);
        const uint32_t t = threadIdx.x;
        if (t < count(bi)) {
            const uint32_t p = idx[(bi & 1) * IDX_STAGE_WORDS + ((b0 + bi * BATCH) & 3u) + t];
            p_next = p;
            cp_async16(raw + t, src.density12 + 3 * (size_t)p);
            cp_async16(raw + BATCH + t, src.density12 + 3 * (size_t)p + 1);
            cp_async16(raw + 2 * BATCH + t, src.density12 + 3 * (size_t)p + 2);
            cp_async16(raw + 3 * BATCH + t, src.features + p);
        }
        cp_async_commit();
    }
    // every thread: its gathered particle -> its half of a pair-interleaved record (the walk of the previous batch must
    // be over: rec is single buffered).  Threads past the end of the list pad it to a multiple of WALK_UNROLL with
    // records that can never be accepted (density 0).
    GUT_D void expand(uint32_t bi) {
        cp_async_wait_all();
        const uint32_t t = threadIdx.x, cnt = count(bi);
        const uint32_t cnt_pad = (cnt + WALK_UNROLL - 1) & ~(WALK_UNROLL - 1);
        float* pair = reinterpret_cast<float*>(rec + PAIR_QUADS * (t >> 1));
        const uint32_t h = t & 1u;
        if (t < cnt) {
            const float4 r0 = raw[t], q = raw[BATCH + t], r2 = raw[2 * BATCH + t], ft = raw[3 * BATCH + t];
            float3 c0, c1, c2;
            quat_to_columns(q.x, q.y, q.z, q.w, c0, c1, c2);   // columns of R == rows of R^T
            const float ix = mufu_rcp(r2.x), iy = mufu_rcp(r2.y), iz = mufu_rcp(r2.z);
            const float3 m0 = c0 * ix, m1 = c1 * iy, m2 = c2 * iz;   // rows of ms = diag(1/s) R^T
            const float3 v = origin - f3(r0.x, r0.y, r0.z);          // frame origin - position
            store_pair_vec(pair, 0, h, m0.x, m0.y, m0.z, dot(m0, v));
            store_pair_vec(pair, 1, h, m1.x, m1.y, m1.z, dot(m1, v));
            store_pair_vec(pair, 2, h, m2.x, m2.y, m2.z, dot(m2, v));
            store_pair_vec(pair, 3, h, r2.x, r2.y, r2.z, r0.w);
            rec[PAIR_QUADS * (t >> 1) + 8 + h] =
                make_float4(fmaxf(ft.x, 0.0f), fmaxf(ft.y, 0.0f), fmaxf(ft.z, 0.0f), __uint_as_float(p_next));
            if (BWD) {   // scale ratios s_i / s_j of the body-frame torque, and 1 / s
                float4* outb = recb + 3 * t;
                outb[0] = make_float4(r2.y * iz, r2.z * iy, r2.x * iz, r2.z * ix);
                outb[1] = make_float4(r2.y * ix, r2.x * iy, ix, iy);
                outb[2] = make_float4(iz, 0.0f, 0.0f, 0.0f);
            }
        } else if (t < cnt_pad) {
            store_pair_vec(pair, 0, h, 1.0f, 1.0f, 1.0f, 0.0f);
            store_pair_vec(pair, 1, h, 1.0f, 1.0f, 1.0f, 0.0f);
            store_pair_vec(pair, 2, h, 1.0f, 1.0f, 1.0f, 0.0f);
            store_pair_vec(pair, 3, h, 1.0f, 1.0f, 1.0f, 0.0f);
        }
    }
    // prologue: idx slices of batches 0 and 1, gather of batch 0 (call after the barriers are initialised and published)
    GUT_D void start() {
        if (nb == 0) return;
        if (threadIdx.x == 0) { issue_idx(0); if (nb > 1) issue_idx(1); }
        issue_gather(0);
    }
    // after the CTA barrier that publishes rec of batch bi: put batch bi+1 (gather) and bi+2 (idx slice) in flight
    GUT_D void advance(uint32_t bi) {
        if (bi + 1 < nb) issue_gather(bi + 1);
        if (threadIdx.x == 0 && bi + 2 < nb) issue_idx(bi + 2);   // stage bi&1 was last read by issue_gather(bi), one barrier ago
    }
    // early exit after batch bi: let everything in flight land before the CTA retires
    GUT_D void drain(uint32_t bi) {
        cp_async_wait_all();
        if (bi + 2 < nb) mbar_wait(&bar[bi & 1], ((bi + 2) >> 1) & 1);
    }
};

// COUNT = true is the instrumented variant used by gut_measure_work(): it writes no image, it counts the ray x particle
// tests and accepted hits the frame performs (the units the FP32 roofline of K4/K5 is expressed in).
template <int DEG, bool COUNT>
__global__ void __launch_bounds__(RENDER_THREADS) gut_render_kernel(GutConfig cfg, GutFrame fr, const uint32_t* __restrict__ tile_start,
                                                                    GutParticleSource src,
                                                                    const float* __restrict__ ray_ori, const float* __restrict__ ray_dir,
                                                                    float4* __restrict__ rgba, float* __restrict__ dist,
                                                                    float* __restrict__ hits, unsigned long long* __restrict__ counters) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t idx_bar[2];

    const uint32_t tile = blockIdx.x;
    uint32_t px, py;
    thread_pixel(fr, tile, px, py);
    const bool inside = px < fr.width && py < fr.height;
    const uint32_t pix = inside ? px + fr.width * py : 0u;
    Ray ray;
    ray.valid = false; ray.o = fr.cam_centre; ray.d = f3(0.f, 0.f, 1.f); ray.tmin = 0.f; ray.tmax = 0.f;
    if (inside) ray = make_ray(fr, ray_ori, ray_dir, pix);
    const float3 dlt = ray.o - fr.cam_centre;
    const bool at_origin = (dlt.x == 0.0f) && (dlt.y == 0.0f) && (dlt.z == 0.0f);

    const uint32_t b0 = tile_start[tile], b1 = tile_start[tile + 1];
    TileStager<false> stg;
    stg.init(smem_raw, idx_bar, src, b0, b1 - b0, fr.cam_centre);

    if (threadIdx.x == 0) {
        mbar_init(&idx_bar[0], 1);
        mbar_init(&idx_bar[1], 1);
        fence_mbar_init();
    }
    const bool uniform = __syncthreads_and(at_origin || !inside) != 0;  // also publishes the barrier initialisation
    stg.start();

    FwdState st;
    st.T = 1.0f; st.D = 0.0f; st.cr = st.cg = st.cb = 0.0f; st.nhits = 0; st.ntests = 0;
    st.alive = inside && ray.valid;
    for (uint32_t b = 0; b < stg.nb; ++b) {
        stg.expand(b);
        __syncthreads();
        stg.advance(b);
        if (uniform) forward_walk<DEG, true, COUNT>(cfg, stg.rec, stg.count(b), ray, dlt, st);
        else forward_walk<DEG, false, COUNT>(cfg, stg.rec, stg.count(b), ray, dlt, st);
        if (__syncthreads_and(!st.alive)) { stg.drain(b); break; }   // the tile saturated early
    }
    if (COUNT) {
        unsigned long long t = st.ntests, hh = st.nhits;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { t += __shfl_xor_sync(0xffffffffu, t, o); hh += __shfl_xor_sync(0xffffffffu, hh, o); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(&counters[0], t); atomicAdd(&counters[1], hh); }
        return;
    }
    if (inside && ray.valid) {
        rgba[pix] = make_float4(st.cr, st.cg, st.cb, 1.0f - st.T);
        dist[pix] = st.D;
        if (cfg.enable_hitcounts) hits[pix] = (float)st.nhits;
    } else if (inside) {
        rgba[pix] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        dist[pix] = 1e06f;
        hits[pix] = 0.0f;
    }
}


// ---------------------------------------------------------------------------------------------------
// K5 backward (k_buffer_size == 0)
// ---------------------------------------------------------------------------------------------------

// Sum 16 per-lane values over the warp with 16 shuffles: each stage exchanges the half of the values the
// partner lane will own.  On return lane l holds in v[0] the warp total of value
// ((l>>4)&1)*8 + ((l>>3)&1)*4 + ((l>>2)&1)*2 + ((l>>1)&1); lanes l and l^1 hold the same value.
GUT_D float warp_transpose_reduce16(float (&v)[16], uint32_t lane) {
    {
        const bool up = lane & 16;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float send = up ? v[k] : v[k + 8];
            const float keep = up ? v[k + 8] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
    }
    {
        const bool up = lane & 8;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float send = up ? v[k] : v[k + 4];
            const float keep = up ? v[k + 4] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
    }
    {
        const bool up = lane & 4;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const float send = up ? v[k] : v[k + 2];
            const float keep = up ? v[k + 2] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        }
    }
    {
        const bool up = lane & 2;
        const float send = up ? v[0] : v[1];
        const float keep = up ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
    return v[0];
}

struct BwdRay {
    float T_f, D_f, dA, dD;   // final transmittance / distance of the forward pass, upstream gradients
    float3 C_f, dC;
    float T, D;               // running state of the re-walk
    float3 C;
    bool alive;
};

// Per accepted hit a lane produces 15 values that are LINEAR in the ray, so they can be summed over rays first (warp
// transpose-reduce, then one RED per value into the per-particle accumulator `raw`, 16 floats per particle) and
// turned into parameter gradients once per particle by K6 (gut_project.cu, finish_density_gradient):
//   [0..2]  A   = dL/d go            (canonical ray origin)        -> dL/dpos   = -ms^T A
//   [3]     dL/d density
//   [4..6]  tau = body-frame torque                                 -> dL/dquat  = rho q + 2 q (x) (0, tau)
//   [7]     tr  = sum_i (dL/dR^T)_ii, direction part                -> rho (radial part of dL/dquat, R(q) is a polynomial)
//   [8..10] Z_i = go_i A_i + u_i B_i - s_i * (hit-distance term)    -> dL/dscale_i = -Z_i / s_i
//   [11..13] dL/d colour
//   [14]    W   = sum_i s_i * (hit-distance scale term)             -> rho
// with B = dL/d u.  This is the chain rule of processHitBwd (gaussianParticles.cuh:635-738) with the per-ray
// quaternion / matrix products (matmul_bw_quat, matmul_bw_vec) factored out of the ray loop.
template <int DEG, bool UNIFORM>
GUT_D void backward_walk(const GutConfig& cfg, const float4* __restrict__ rec, const float4* __restrict__ recb, uint32_t cnt, const Ray& ray,
                         const float3 dlt, BwdRay& st, uint32_t lane, float* __restrict__ raw_grad) {
    const float minT = cfg.min_transmittance;
    const uint32_t cnt_pad = (cnt + WALK_UNROLL - 1) & ~(WALK_UNROLL - 1);   // pad records are never accepted
    for (uint32_t jp = 0; jp < cnt_pad; jp += 2) {
        if ((jp & (ALIVE_CHECK - 1)) == 0u && !__any_sync(0xffffffffu, st.alive)) break;
        const float4* r = rec + PAIR_QUADS * (jp >> 1);
        PairCanon pk;
        bool acc[2];
        pair_test<DEG, UNIFORM>(cfg, reinterpret_cast<const ulonglong2*>(r), ray.d, dlt, pk, acc[0], acc[1]);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
        const bool hit = acc[h] && st.alive;  // no tMinMax test here
        if (!__any_sync(0xffffffffu, hit)) continue;
        const uint32_t j = jp + h;
        const Canon k = canon_half(pk, h);
        const float4 Sxy = r[6], Szd = r[7];
        const float4 S = make_float4(h ? Sxy.y : Sxy.x, h ? Sxy.w : Sxy.z, h ? Szd.y : Szd.x, h ? Szd.w : Szd.z);

        float v[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) v[q] = 0.0f;
        const float4 F = r[8 + h];
        if (hit) {
            const float4 R0 = recb[3 * j], R1 = recb[3 * j + 1];
            const float isz = recb[3 * j + 2].x;
            const float3 go = k.go, u = k.u;
            const float il = mufu_rsq(k.l2);           // 1 / |u|
            const float3 grd = u * il;                 // normalised canonical direction
            const float dgo = dot(grd, go);
            // hit distance gdist = |s * grd * dot(grd, -go)| = |dgo| * |sg|, sg = s * grd
            const float3 sg = f3(S.x * grd.x, S.y * grd.y, S.z * grd.z);
            const float3 sg2 = sg * sg;
            const float sum2 = sg2.x + sg2.y + sg2.z;
            const float gsq = dgo * dgo * sum2;
            const float gdist = fabsf(dgo) * mufu_sqrt(sum2);
            const float alpha = fminf(cfg.max_alpha, k.araw);
            const float T = st.T;
            const float w = alpha * T;
            const float om = 1.0f - alpha;
            const float nextT = om * T;
            const float invT = nextT <= minT ? 0.0f : mufu_rcp(nextT);

            // hit-distance terms (gaussianParticles.cuh:536-567; the Jacobians of grds wrt grd / gro are the reference's
            // element-wise ones).  grds = -dgo sg;  dL/dgrds = grds hs = kap sg  with  kap = -dgo hs
            st.D = fmaf(w, gdist, st.D);
            const float resD = fmaxf((st.D_f - st.D) * invT, 0.0f);
            const float aHit = (gdist - resD) * T * st.dD;
            const float hs = gsq > 0.0f ? w * st.dD * mufu_rcp(gdist) : 0.0f;
            const float kap = -dgo * hs;
            const float kd = kap * dgo;                // s_i * sclHit_i = -kd * sg2_i
            // grdHit_i = -kap (sg2_i go_i + dgo s_i sg_i) ; groHit_i = -kap sg2_i grd_i
            const float3 tH = f3(fmaf(sg2.x, go.x, dgo * S.x * sg.x), fmaf(sg2.y, go.y, dgo * S.y * sg.y), fmaf(sg2.z, go.z, dgo * S.z * sg.z));

            // opacity and colour terms (:569-633)
            const float resT = alpha < 0.999999f ? st.T_f * mufu_rcp(om) : T;
            const float aDns = resT * st.dA;
            st.C = f3(fmaf(w, F.x, st.C.x), fmaf(w, F.y, st.C.y), fmaf(w, F.z, st.C.z));
            const float3 resC = f3(fmaxf((st.C_f.x - st.C.x) * invT, 0.0f), fmaxf((st.C_f.y - st.C.y) * invT, 0.0f),
                                   fmaxf((st.C_f.z - st.C.z) * invT, 0.0f));
            const float Ssum = aHit + aDns + T * ((F.x - resC.x) * st.dC.x + (F.y - resC.y) * st.dC.y + (F.z - resC.z) * st.dC.z);
            const float dDens = k.resp * Ssum;
            const float respG = S.w * Ssum;
            const float g2G = kernel_response_grad<DEG>(k.g2, k.resp, respG);

            // g2 = |grd x go|^2 : adjoint of the cross product (:671-689); grd x go = c / |u|
            const float3 gcG = k.c * (2.0f * g2G * il);
            const float3 grdG = cross(go, gcG);
            const float3 groG = cross(gcG, grd);
            // A = dL/dgo (total), X = dL/dgrd (total)
            const float3 Ag = f3(fmaf(-kap, sg2.x * grd.x, groG.x), fmaf(-kap, sg2.y * grd.y, groG.y), fmaf(-kap, sg2.z * grd.z, groG.z));
            const float3 X = f3(fmaf(-kap, tH.x, grdG.x), fmaf(-kap, tH.y, grdG.y), fmaf(-kap, tH.z, grdG.z));
            // normalisation adjoint (safe_normalize_bw, mathUtils.cuh:420-430): B = dL/du = (X - grd (grd . X)) / |u|
            const float gx = dot(grd, X);
            const float3 Bg = f3(il * (X.x - grd.x * gx), il * (X.y - grd.y * gx), il * (X.z - grd.z * gx));

            // P_ij = go_i A_j + u_i B_j : diagonal -> scale, off-diagonal -> body-frame torque
            v[0] = Ag.x; v[1] = Ag.y; v[2] = Ag.z;
            v[3] = dDens;
            v[4] = R0.y * fmaf(go.z, Ag.y, u.z * Bg.y) - R0.x * fmaf(go.y, Ag.z, u.y * Bg.z);   // rho_zy P_zy - rho_yz P_yz
            v[5] = R0.z * fmaf(go.x, Ag.z, u.x * Bg.z) - R0.w * fmaf(go.z, Ag.x, u.z * Bg.x);   // rho_xz P_xz - rho_zx P_zx
            v[6] = R1.x * fmaf(go.y, Ag.x, u.y * Bg.x) - R1.y * fmaf(go.x, Ag.y, u.x * Bg.y);   // rho_yx P_yx - rho_xy P_xy
            float tr = R1.z * Bg.x * ray.d.x + R1.w * Bg.y * ray.d.y + isz * Bg.z * ray.d.z;
            if (!UNIFORM) tr += R1.z * Ag.x * dlt.x + R1.w * Ag.y * dlt.y + isz * Ag.z * dlt.z;
            v[7] = tr;
            v[8] = fmaf(go.x, Ag.x, fmaf(u.x, Bg.x, kd * sg2.x));
            v[9] = fmaf(go.y, Ag.y, fmaf(u.y, Bg.y, kd * sg2.y));
            v[10] = fmaf(go.z, Ag.z, fmaf(u.z, Bg.z, kd * sg2.z));
            v[11] = w * st.dC.x; v[12] = w * st.dC.y; v[13] = w * st.dC.z;
            v[14] = -kd * sum2;

            st.T = nextT;
            if (nextT < minT) st.alive = false;
        }
        const float total = warp_transpose_reduce16(v, lane);
        const uint32_t slot = ((lane >> 4) & 1u) * 8u + ((lane >> 3) & 1u) * 4u + ((lane >> 2) & 1u) * 2u + ((lane >> 1) & 1u);
        if ((lane & 1u) == 0u && slot < 15u) atomicAdd(raw_grad + 16 * (size_t)__float_as_uint(F.w) + slot, total);
        }
    }
}

template <int DEG>
__global__ void __launch_bounds__(RENDER_THREADS, 3) gut_render_backward_kernel(
    GutConfig cfg, GutFrame fr, const uint32_t* __restrict__ tile_start, GutParticleSource src,
    const float* __restrict__ ray_ori, const float* __restrict__ ray_dir,
    const float4* __restrict__ rgba, const float4* __restrict__ rgba_grad, const float* __restrict__ dist,
    const float* __restrict__ dist_grad, float* __restrict__ raw_grad) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t idx_bar[2];

    const uint32_t tile = blockIdx.x;
    const uint32_t lane = threadIdx.x & 31;
    uint32_t px, py;
    thread_pixel(fr, tile, px, py);
    const bool inside = px < fr.width && py < fr.height;
    const uint32_t pix = inside ? px + fr.width * py : 0u;
    Ray ray;
    ray.valid = false; ray.o = fr.cam_centre; ray.d = f3(0.f, 0.f, 1.f); ray.tmin = 0.f; ray.tmax = 0.f;
    if (inside) ray = make_ray(fr, ray_ori, ray_dir, pix);
    const float3 dlt = ray.o - fr.cam_centre;
    const bool at_origin = (dlt.x == 0.0f) && (dlt.y == 0.0f) && (dlt.z == 0.0f);

    // initializeBackwardRay (rayPayloadBackward.cuh:46-55)
    BwdRay st;
    st.alive = inside && ray.valid;
    st.T_f = 1.0f; st.D_f = 0.0f; st.dA = 0.0f; st.dD = 0.0f;
    st.C_f = f3(0.f, 0.f, 0.f); st.dC = f3(0.f, 0.f, 0.f);
    st.T = 1.0f; st.D = 0.0f; st.C = f3(0.f, 0.f, 0.f);
    if (st.alive) {
        const float4 fd = __ldg(rgba + pix), gd = __ldg(rgba_grad + pix);
        st.T_f = 1.0f - fd.w;      // transmittanceBackward
        st.dA = gd.w;              // transmittanceGradient = -dA; used as residualTrm * -(-dA)
        st.D_f = __ldg(dist + pix);
        st.dD = __ldg(dist_grad + pix);
        st.C_f = f3(fd.x, fd.y, fd.z);
        st.dC = f3(gd.x, gd.y, gd.z);
    }

    const uint32_t b0 = tile_start[tile], b1 = tile_start[tile + 1];
    TileStager<true> stg;
    stg.init(smem_raw, idx_bar, src, b0, b1 - b0, fr.cam_centre);

    if (threadIdx.x == 0) {
        mbar_init(&idx_bar[0], 1);
        mbar_init(&idx_bar[1], 1);
        fence_mbar_init();
    }
    const bool uniform = __syncthreads_and(at_origin || !inside) != 0;
    stg.start();

    for (uint32_t b = 0; b < stg.nb; ++b) {
        stg.expand(b);
        __syncthreads();
        stg.advance(b);
        if (uniform) backward_walk<DEG, true>(cfg, stg.rec, stg.recb, stg.count(b), ray, dlt, st, lane, raw_grad);
        else backward_walk<DEG, false>(cfg, stg.rec, stg.recb, stg.count(b), ray, dlt, st, lane, raw_grad);
        if (__syncthreads_and(!st.alive)) { stg.drain(b); break; }
    }
}

template <int DEG>
cudaError_t launch_render_deg(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start, GutParticleSource src,
                              const float* ray_ori, const float* ray_dir, float* rgba, float* dist, float* hits,
                              unsigned long long* counters, cudaStream_t stream) {
    const size_t smem = StageLayout<false>::TOTAL;
    cudaError_t e;
    if (counters) {
        e = cudaFuncSetAttribute(gut_render_kernel<DEG, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        gut_render_kernel<DEG, true><<<frame.num_tiles, RENDER_THREADS, smem, stream>>>(cfg, frame, tile_start, src, ray_ori, ray_dir,
                                                                                      nullptr, nullptr, nullptr, counters);
    } else {
        e = cudaFuncSetAttribute(gut_render_kernel<DEG, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        gut_render_kernel<DEG, false><<<frame.num_tiles, RENDER_THREADS, smem, stream>>>(cfg, frame, tile_start, src, ray_ori, ray_dir,
                                                                                       reinterpret_cast<float4*>(rgba), dist, hits, nullptr);
    }
    return cudaGetLastError();
}

template <int DEG>
cudaError_t launch_render_bwd_deg(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start, GutParticleSource src,
                                  const float* ray_ori, const float* ray_dir, const float* rgba,
                                  const float* rgba_grad, const float* dist, const float* dist_grad, float* raw_grad,
                                  cudaStream_t stream) {
    const size_t smem = StageLayout<true>::TOTAL;
    cudaError_t e = cudaFuncSetAttribute(gut_render_backward_kernel<DEG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    gut_render_backward_kernel<DEG><<<frame.num_tiles, RENDER_THREADS, smem, stream>>>(
        cfg, frame, tile_start, src, ray_ori, ray_dir, reinterpret_cast<const float4*>(rgba),
        reinterpret_cast<const float4*>(rgba_grad), dist, dist_grad, raw_grad);
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
                              unsigned long long* counters, cudaStream_t stream) {
#define CALL_FWD(D) launch_render_deg<D>(cfg, frame, tile_start, src, ray_ori, ray_dir, rgba, dist, hits, counters, stream)
    GUT_DISPATCH_DEGREE(cfg.kernel_degree, CALL_FWD)
#undef CALL_FWD
}

cudaError_t gut_launch_render_backward(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start,
                                       GutParticleSource src, const float* ray_ori, const float* ray_dir,
                                       const float* rgba, const float* rgba_grad, const float* dist, const float* dist_grad,
                                       float* raw_grad, cudaStream_t stream) {
#define CALL_BWD(D) \
    launch_render_bwd_deg<D>(cfg, frame, tile_start, src, ray_ori, ray_dir, rgba, rgba_grad, dist, dist_grad, raw_grad, stream)
    GUT_DISPATCH_DEGREE(cfg.kernel_degree, CALL_BWD)
#undef CALL_BWD
}
