// gut_render_kbuffer.cu -- K4 / K5 of the 3DGUT rasteriser for k_buffer_size > 0 (the "sorted" configurations,
// configs/paper/3dgut/base_sorted.yaml: k_buffer_size = 16).
//
// reference: HitParticleKBuffer<K> (include/3dgut/kernels/cuda/renderers/gutKBufferRenderer.cuh:28-85),
//   GUTKBufferRenderer::evalKBuffer (:217-292) and processHitParticle (:108-171), forward and backward branch;
//   per-hit backward: particleFeaturesIntegrateBwd (kernels/slang/models/shRadiativeParticles.slang:179-207) and
//   particleDensityProcessHitBwdToBuffer (kernels/slang/models/gaussianParticles.slang:394-451), i.e. Slang autodiff of
//   hit (:188-222) + integrateHit<true> (:224-252): true derivatives, alpha clamp masked, tMinMax test also in backward.
//
// Design.  A ray keeps the ORDER of its K closest pending hits in K registers -- one 32-bit key per hit, the distance bits
// with the payload slot number in the low 4 mantissa bits -- and their exact (distance, alpha, particle) in a
// conflict-free shared-memory column; insertion is a fully unrolled integer min/max chain over static indices, the
// drain walks the static indices too, so nothing spills to local memory.  One warp per CTA owns an 8x4 pixel block (eight blocks per 16x16 tile, one ray per
// lane) and stages the tile's sorted list through the same
// warp-private pipeline as the K = 0 kernels (gut_render_common.cuh).  A hit leaves the buffer long after its batch
// left shared memory, so processing gathers the particle by index: its K1 colour in the forward, its 48-byte
// parameters too in the backward.
// Backward: the reference adds 11 + 3 scalar atomics per hit and thread into [N,12] / [N,3].  Here a hit produces the same
// 15 ray-linear canonical-frame values as the K = 0 kernel (gut_render.cu), with the true hit-distance Jacobians and the
// masked clamp, and adds them to the [N,16] accumulators with four red.global.add.v4.f32 -- K6 is shared.
#include "gut_render_common.cuh"

namespace {

#ifndef GUT_KB_MIN_BLOCKS
#define GUT_KB_MIN_BLOCKS 20   // resident one-warp CTAs per SM the forward is compiled for (96 registers: 4.59 -> 4.38 ms; the backward is better uncapped, 8.72 vs 8.86 ms)
#endif
constexpr int KB_BLOCKS_PER_TILE = 8;   // 8x4 pixel blocks (one warp, one ray per lane) of a 16x16 tile

// The ordering state of a ray's buffer is ONE 32-bit key per entry: the hit distance's bit pattern (positive floats order
// like unsigned integers) with its low 4 mantissa bits replaced by the number of the payload slot that holds the hit's
// exact (distance, alpha, particle) in shared memory.  Insertion is then a min/max chain -- two integer instructions per
// level instead of a compare and six selects on three 32-bit fields -- and the buffer costs K registers, not 3K.
// Two hits whose distances agree in all but the last 4 mantissa bits (relative 1.9e-6) are ordered by slot number;
// the oracle marks rays with such near ties as borderline (another fp32 implementation may order them either way).
constexpr uint32_t KB_SLOT_MASK = 15u;
template <int K>
struct KBuffer {
    static_assert(K <= 16, "payload slot number is 4 bits");
    uint32_t key[K];     // ascending; 0 = empty (empties sit at the bottom, like t = -1 in HitParticleKBuffer)
    uint32_t n;
    GUT_D void clear() {
        n = 0;
#pragma unroll
        for (int i = 0; i < K; ++i) key[i] = 0u;
    }
    GUT_D bool full() const { return n == (uint32_t)K; }
    GUT_D uint32_t front_slot() const { return key[0] & KB_SLOT_MASK; }
    // insert (the caller has processed entry 0 if the buffer was full): HitParticleKBuffer::insert, :45-62.
    // Returns the payload slot the caller writes the new hit to (the popped entry's slot when the buffer was full).
    GUT_D uint32_t insert(float ht) {
        uint32_t slot;
        if (full()) { slot = front_slot(); key[0] = 0u; } else slot = n++;
        uint32_t carry = (__float_as_uint(ht) & ~KB_SLOT_MASK) | slot;
#pragma unroll
        for (int i = K - 1; i >= 0; --i) {
            const uint32_t small = min(key[i], carry);
            key[i] = max(key[i], carry);
            carry = small;
        }
        return slot;
    }
};

// exact payload of the buffered hits: [field][slot][lane] in shared memory, a lane only ever touches its own column
// (bank = lane: conflict free, no synchronisation)
template <int K>
struct KPayload {
    float* t; float* a; uint32_t* p;
    GUT_D void init(float* base) { t = base; a = base + K * 32; p = reinterpret_cast<uint32_t*>(base + 2 * K * 32); }
    GUT_D void put(uint32_t slot, uint32_t lane, float ht, float ha, uint32_t hp) {
        const uint32_t o = slot * 32u + lane;
        t[o] = ht; a[o] = ha; p[o] = hp;
    }
};

struct RayK {
    float3 o, d;
    float tmin, tmax;
    bool inside, valid;
    uint32_t pix;
};
GUT_D RayK make_ray_k(const GutFrame& fr, const float* __restrict__ ray_ori, const float* __restrict__ ray_dir, uint32_t tile, uint32_t block) {
    const uint32_t tx = tile % fr.grid_x, ty = tile / fr.grid_x, l = threadIdx.x & 31;
    const uint32_t px = tx * GUT_TILE + (block & 1) * 8 + (l & 7), py = ty * GUT_TILE + (block >> 1) * 4 + (l >> 3);
    RayK r;
    r.inside = px < fr.width && py < fr.height;
    r.pix = r.inside ? px + fr.width * py : 0u;
    r.o = fr.cam_centre; r.d = f3(0.f, 0.f, 1.f); r.tmin = 0.f; r.tmax = 0.f; r.valid = false;
    if (r.inside) {
        const Ray q = make_ray(fr, ray_ori, ray_dir, r.pix);
        r.o = q.o; r.d = q.d; r.tmin = q.tmin; r.tmax = q.tmax; r.valid = q.valid;
    }
    return r;
}

// Cone (axis, tan of the half angle) around the live rays of the warp's block for WarpStager::stage_culled; w < 0 = no
// cull (per-ray origins, degenerate bundles, or a kernel degree the bound is not derived for).  Same construction as in
// gut_render_kernel (gut_render.cu); an 8x4 block gives a narrower cone than the 8x8 block of the K = 0 kernels.
template <int DEG>
GUT_D float4 block_cone(const RayK& ray, const float3 dlt, bool alive) {
    const bool at_origin = !ray.inside || (dlt.x == 0.0f && dlt.y == 0.0f && dlt.z == 0.0f);
    const bool uniform = __all_sync(0xffffffffu, at_origin);
    const float3 dn = rsqrtf(dot(ray.d, ray.d)) * ray.d;
    float3 sum = alive ? dn : f3(0.f, 0.f, 0.f);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum.x += __shfl_xor_sync(0xffffffffu, sum.x, o); sum.y += __shfl_xor_sync(0xffffffffu, sum.y, o); sum.z += __shfl_xor_sync(0xffffffffu, sum.z, o);
    }
    const float n2 = dot(sum, sum);
    const float3 axis = n2 > 0.0f ? rsqrtf(n2) * sum : f3(0.f, 0.f, 1.f);
    float cmin = alive ? dot(axis, dn) : 1.0f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cmin = fminf(cmin, __shfl_xor_sync(0xffffffffu, cmin, o));
    float tan_h = -1.0f;
    if (DEG == 2 && uniform && cmin > 0.5f) tan_h = sqrtf(fmaxf(1.0f - cmin * cmin, 0.0f)) / cmin * 1.001f + 1e-6f;
    return make_float4(axis.x, axis.y, axis.z, tan_h);
}

// test of one staged record against one ray; on acceptance also the hit distance (canonicalRayDistance,
// gaussianParticles.slang:165-171).  dlt = ray origin - frame origin.
template <int DEG>
GUT_D bool record_hit(const GutConfig& cfg, const float4* __restrict__ r, const float3 d, const float3 dlt, float tmin, float tmax,
                      float& alpha, float& t) {
    const float4 A = r[0], B = r[1], Cm = r[2], S = r[3];
    const float3 go = f3(A.w + A.x * dlt.x + A.y * dlt.y + A.z * dlt.z, B.w + B.x * dlt.x + B.y * dlt.y + B.z * dlt.z,
                         Cm.w + Cm.x * dlt.x + Cm.y * dlt.y + Cm.z * dlt.z);
    const float3 u = f3(A.x * d.x + A.y * d.y + A.z * d.z, B.x * d.x + B.y * d.y + B.z * d.z, Cm.x * d.x + Cm.y * d.y + Cm.z * d.z);
    const float l2 = dot(u, u);
    const float3 c = cross(u, go);
    const float il2 = mufu_rcp(l2);
    const float g2 = dot(c, c) * il2;
    const float resp = (DEG == 2) ? mufu_ex2(g2 * -0.72134752044448170368f) : kernel_response<DEG>(g2);
    const float araw = resp * S.w;
    if (!((resp > cfg.min_kernel_response) && (araw > cfg.min_alpha))) return false;
    const float3 su = f3(S.x * u.x, S.y * u.y, S.z * u.z);
    t = fabsf(dot(u, go)) * il2 * mufu_sqrt(dot(su, su));
    alpha = fminf(cfg.max_alpha, araw);
    return t > tmin && t < tmax;
}

// ---------------------------------------------------------------------------------------------------
// K4 forward, K > 0
// ---------------------------------------------------------------------------------------------------
struct FwdStateK {
    float T, D, cr, cg, cb;
    uint32_t nhits;
    bool alive;
};
// processHitParticle, forward branch (:153-170): integrateHit<false> + integrateRadiance<false>
GUT_D void process_hit_fwd(const GutConfig& cfg, const float4* __restrict__ features, float t, float alpha, uint32_t p, FwdStateK& st) {
    const float w = alpha * st.T;
    st.D = fmaf(t, w, st.D);
    st.T *= (1.0f - alpha);
    if (w > 0.0f) {
        const float4 F = __ldg(features + p);
        st.cr = fmaf(fmaxf(F.x, 0.0f), w, st.cr); st.cg = fmaf(fmaxf(F.y, 0.0f), w, st.cg); st.cb = fmaf(fmaxf(F.z, 0.0f), w, st.cb);
        st.nhits++;
    }
    if (st.T < cfg.min_transmittance) st.alive = false;
}

template <int DEG, int K>
__global__ void __launch_bounds__(RENDER_THREADS, GUT_KB_MIN_BLOCKS) gut_render_kbuffer_kernel(GutConfig cfg, GutFrame fr, const uint32_t* __restrict__ tile_start,
                                                                            GutParticleSource src, const float* __restrict__ ray_ori,
                                                                            const float* __restrict__ ray_dir, float4* __restrict__ rgba,
                                                                            float* __restrict__ dist, float* __restrict__ hits) {
    __shared__ __align__(16) unsigned char smem_raw[StageLayout<false>::TOTAL];
    __shared__ float kb_payload[3 * K * 32];
    // CTAs are dispatched in blockIdx order: tiles longest list first (gut_tile_scan_kernel leaves the order behind tile_start)
    const uint32_t tile = __ldg(tile_start + fr.num_tiles + 1 + blockIdx.x / KB_BLOCKS_PER_TILE), block = blockIdx.x % KB_BLOCKS_PER_TILE;
    const RayK ray = make_ray_k(fr, ray_ori, ray_dir, tile, block);
    const float3 dlt = ray.o - fr.cam_centre;

    const uint32_t b0 = tile_start[tile], b1 = tile_start[tile + 1];
    WarpStager<false> stg;
    stg.init(smem_raw, src, b0, b1 - b0, fr.cam_centre);

    FwdStateK st;
    st.T = 1.0f; st.D = 0.0f; st.cr = st.cg = st.cb = 0.0f; st.nhits = 0;
    st.alive = ray.inside && ray.valid;
    const uint32_t lane = threadIdx.x & 31;
    KBuffer<K> kb;
    kb.clear();
    KPayload<K> pay;
    pay.init(kb_payload);
    const float4 cone = block_cone<DEG>(ray, dlt, st.alive);
    for (uint32_t b = 0; b < stg.nb; ++b) {
        uint32_t keep;
        const uint32_t cnt = stg.stage_culled(b, cfg, cone, keep);   // records no ray of the block can accept are dropped while staging
        for (uint32_t j = 0; j < cnt; ++j) {
            float alpha, t;
            if (st.alive && record_hit<DEG>(cfg, stg.rec + 5 * j, ray.d, dlt, ray.tmin, ray.tmax, alpha, t)) {
                if (kb.full()) {
                    const uint32_t o = kb.front_slot() * 32u + lane;
                    process_hit_fwd(cfg, src.features, pay.t[o], pay.a[o], pay.p[o], st);
                }
                const uint32_t slot = kb.insert(t);
                pay.put(slot, lane, t, alpha, __float_as_uint(stg.rec[5 * j + 4].w));
            }
        }
        if (!__any_sync(0xffffffffu, st.alive)) break;
        stg.rotate();
    }
    stg.drain();
    // drain the buffer closest first (:282-290): valid entries sit at the top, empty ones have key 0
#pragma unroll
    for (int i = 0; i < K; ++i)
        if (st.alive && kb.key[i] != 0u) {
            const uint32_t o = (kb.key[i] & KB_SLOT_MASK) * 32u + lane;
            process_hit_fwd(cfg, src.features, pay.t[o], pay.a[o], pay.p[o], st);
        }

    if (ray.inside && ray.valid) {
        rgba[ray.pix] = make_float4(st.cr, st.cg, st.cb, 1.0f - st.T);
        dist[ray.pix] = st.D;
        if (cfg.enable_hitcounts) hits[ray.pix] = (float)st.nhits;
    } else if (ray.inside) {
        rgba[ray.pix] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        dist[ray.pix] = 1e06f;
        hits[ray.pix] = 0.0f;
    }
}

// ---------------------------------------------------------------------------------------------------
// K5 backward, K > 0
// ---------------------------------------------------------------------------------------------------
struct BwdStateK {
    float Tb, Tg, Db, Dg;    // transmittance / distance still behind the current hit, and their gradients
    float3 Fb, Fg;           // colour still behind the current hit, and its gradient
    float T;
    bool alive;
};

// processHitParticle, backward branch (:116-152), for ONE ray (called by the lanes whose buffer releases a hit).  A hit
// yields the 15 canonical-frame values of gut_render.cu's backward_walk (true hit-distance Jacobians, masked clamp),
// added to raw_grad[16 p + .] -- which K6 turns into parameter gradients, as for K = 0 -- with four 128-bit vector
// reductions per ray (the reference: 14 scalar atomics per hit and thread).  No warp reduction: the rays of a block
// leave their buffers with different particles most of the time (5.7 distinct particles per pop event of a warp at
// the bench workload), and one reduction per distinct particle was 44 % of this kernel's instructions.
template <int DEG>
GUT_D void process_hit_bwd(const GutConfig& cfg, const GutParticleSource& src, const float3 centre, const float3 o, const float3 d,
                           float t, float alpha, uint32_t p, BwdStateK& st, float* __restrict__ raw_grad) {
    if (alpha > 0.0f) {
        float v[15];
        const float4 r0 = __ldg(src.density12 + 3 * (size_t)p), q = __ldg(src.density12 + 3 * (size_t)p + 1),
                     r2 = __ldg(src.density12 + 3 * (size_t)p + 2), F = __ldg(src.features + p);
        // particleFeaturesIntegrateBwd: un-blend the UNclamped feature, adjoint of lerp(I, f, alpha)
        const float wgt = mufu_rcp(1.0f - alpha);
        st.Fb = f3((st.Fb.x - F.x * alpha) * wgt, (st.Fb.y - F.y * alpha) * wgt, (st.Fb.z - F.z * alpha) * wgt);
        const float alpha_grad = st.Fg.x * (F.x - st.Fb.x) + st.Fg.y * (F.y - st.Fb.y) + st.Fg.z * (F.z - st.Fb.z);
        const float3 fgrad = st.Fg * alpha;
        st.Fg = st.Fg * (1.0f - alpha);
        // particleDensityProcessHitBwdToBuffer: un-blend transmittance and distance, adjoint of integrateHit<true>
        st.Tb *= wgt;
        st.Db = (st.Db - t * alpha) * wgt;
        const float d_alpha = alpha_grad + st.Dg * (t - st.Db) - st.Tg * st.Tb;
        const float d_depth = st.Dg * alpha;
        st.Dg *= (1.0f - alpha);
        st.Tg *= (1.0f - alpha);

        // the particle in its canonical frame (same quantities as the staged record, gut_render_common.cuh)
        float3 c0, c1, c2;
        quat_to_columns(q.x, q.y, q.z, q.w, c0, c1, c2);
        const float ix = mufu_rcp(r2.x), iy = mufu_rcp(r2.y), iz = mufu_rcp(r2.z);
        const float3 m0 = c0 * ix, m1 = c1 * iy, m2 = c2 * iz;
        const float3 rel = o - f3(r0.x, r0.y, r0.z);
        const float3 go = f3(dot(m0, rel), dot(m1, rel), dot(m2, rel));
        const float3 u = f3(dot(m0, d), dot(m1, d), dot(m2, d));
        const float il = mufu_rsq(dot(u, u));
        const float3 grd = u * il;
        const float dgo = dot(grd, go);
        const float3 cr = cross(grd, go);
        const float g2 = dot(cr, cr);
        const float resp = (DEG == 2) ? mufu_ex2(g2 * -0.72134752044448170368f) : kernel_response<DEG>(g2);
        // alpha = min(max_alpha, resp * density): no gradient when clamped
        const bool unclamped = !(cfg.max_alpha < resp * r0.w);
        const float dDens = unclamped ? d_alpha * resp : 0.0f;
        const float g2G = kernel_response_grad<DEG>(g2, resp, unclamped ? d_alpha * r0.w : 0.0f);
        // depth = |dgo| |sg|, sg = s * grd: true Jacobians wrt go, grd and the explicit scale
        const float3 sg = f3(r2.x * grd.x, r2.y * grd.y, r2.z * grd.z);
        const float3 sg2 = sg * sg;
        const float ssum = sg2.x + sg2.y + sg2.z;
        const float depth = fabsf(dgo) * mufu_sqrt(ssum);
        const float kap = depth > 0.0f ? -dgo * d_depth * mufu_rcp(depth) : 0.0f;   // dL/dgrds = kap sg
        const float kd = kap * dgo;
        const float3 gcG = cr * (2.0f * g2G);
        const float3 grdG = cross(go, gcG);
        const float3 groG = cross(gcG, grd);
        const float3 Ag = f3(fmaf(-kap * ssum, grd.x, groG.x), fmaf(-kap * ssum, grd.y, groG.y), fmaf(-kap * ssum, grd.z, groG.z));
        const float3 X = f3(grdG.x - kap * fmaf(dgo * r2.x, sg.x, ssum * go.x), grdG.y - kap * fmaf(dgo * r2.y, sg.y, ssum * go.y),
                            grdG.z - kap * fmaf(dgo * r2.z, sg.z, ssum * go.z));
        const float gx = dot(grd, X);
        const float3 Y = f3(X.x - grd.x * gx, X.y - grd.y * gx, X.z - grd.z * gx);   // |u| dL/du
        // The accumulators follow gut_render.cu's backward_walk: K6 adds the halves of the torque, the trace and the scale
        // terms that are go_frame (x) sum(A) (origin fold, gut_project.cu) and scales the torque by 1 / (s_j s_k); a ray adds
        // the direction halves and what its own origin contributes beyond the frame origin (dl = ms (o - frame origin)).
        const float3 dlt = o - centre;
        const float3 dl = f3(dot(m0, dlt), dot(m1, dlt), dot(m2, dlt));
        const float3 hv = f3(r2.x * sg.x, r2.y * sg.y, r2.z * sg.z);                       // s^2 grd
        const float3 gd = f3(r2.x * r2.x * dl.x, r2.y * r2.y * dl.y, r2.z * r2.z * dl.z);  // s^2 dl
        const float3 tq = cross(Y, hv) + cross(Ag, gd);
        v[0] = Ag.x; v[1] = Ag.y; v[2] = Ag.z;
        v[3] = dDens;
        v[4] = tq.x; v[5] = tq.y; v[6] = tq.z;
        v[7] = il * (ix * (Y.x * d.x) + iy * (Y.y * d.y) + iz * (Y.z * d.z)) + (ix * (Ag.x * dlt.x) + iy * (Ag.y * dlt.y) + iz * (Ag.z * dlt.z));
        v[8] = fmaf(dl.x, Ag.x, fmaf(grd.x, Y.x, kd * sg2.x));
        v[9] = fmaf(dl.y, Ag.y, fmaf(grd.y, Y.y, kd * sg2.y));
        v[10] = fmaf(dl.z, Ag.z, fmaf(grd.z, Y.z, kd * sg2.z));
        v[11] = fgrad.x; v[12] = fgrad.y; v[13] = fgrad.z;
        v[14] = dot(dl, Ag);   // per-ray part of go . A; the direction half of the trace vanishes (gut_render.cu)
    // Rays of a block leave their buffers with DIFFERENT particles most of the time (measured: 5.7 distinct particles per
    // pop event of a warp at the bench workload), so a warp reduction per distinct particle costs more than it saves
    // (it was 44 % of the kernel's instructions).  Each lane adds its 15 values itself with four 128-bit vector
    // reductions (red.global.add.v4.f32, sm_90+): 4 instructions per pop event instead of ~90 per distinct particle.
        float4* dst = reinterpret_cast<float4*>(raw_grad + 16 * (size_t)p);
        atomicAdd(dst, make_float4(v[0], v[1], v[2], v[3]));
        atomicAdd(dst + 1, make_float4(v[4], v[5], v[6], v[7]));
        atomicAdd(dst + 2, make_float4(v[8], v[9], v[10], v[11]));
        atomicAdd(dst + 3, make_float4(v[12], v[13], v[14], 0.0f));
    }
    st.T *= (1.0f - alpha);
    if (st.T < cfg.min_transmittance) st.alive = false;
}

template <int DEG, int K>
__global__ void __launch_bounds__(RENDER_THREADS) gut_render_kbuffer_backward_kernel(
    GutConfig cfg, GutFrame fr, const uint32_t* __restrict__ tile_start, GutParticleSource src,
    const float* __restrict__ ray_ori, const float* __restrict__ ray_dir,
    const float4* __restrict__ rgba, const float4* __restrict__ rgba_grad, const float* __restrict__ dist,
    const float* __restrict__ dist_grad, float* __restrict__ raw_grad) {
    __shared__ __align__(16) unsigned char smem_raw[StageLayout<false>::TOTAL];
    __shared__ float kb_payload[3 * K * 32];
    // CTAs are dispatched in blockIdx order: tiles longest list first (gut_tile_scan_kernel leaves the order behind tile_start)
    const uint32_t tile = __ldg(tile_start + fr.num_tiles + 1 + blockIdx.x / KB_BLOCKS_PER_TILE), block = blockIdx.x % KB_BLOCKS_PER_TILE;
    const RayK ray = make_ray_k(fr, ray_ori, ray_dir, tile, block);
    const float3 dlt = ray.o - fr.cam_centre;
    const uint32_t lane = threadIdx.x & 31;

    BwdStateK st;   // initializeBackwardRay (rayPayloadBackward.cuh:46-55)
    st.alive = ray.inside && ray.valid;
    st.T = 1.0f; st.Tb = 1.0f; st.Tg = 0.0f; st.Db = 0.0f; st.Dg = 0.0f;
    st.Fb = f3(0.f, 0.f, 0.f); st.Fg = f3(0.f, 0.f, 0.f);
    if (st.alive) {
        const float4 fd = __ldg(rgba + ray.pix), gd = __ldg(rgba_grad + ray.pix);
        st.Tb = 1.0f - fd.w; st.Tg = -gd.w;
        st.Db = __ldg(dist + ray.pix); st.Dg = __ldg(dist_grad + ray.pix);
        st.Fb = f3(fd.x, fd.y, fd.z); st.Fg = f3(gd.x, gd.y, gd.z);
    }

    const uint32_t b0 = tile_start[tile], b1 = tile_start[tile + 1];
    WarpStager<false> stg;
    stg.init(smem_raw, src, b0, b1 - b0, fr.cam_centre);
    KBuffer<K> kb;
    kb.clear();
    KPayload<K> pay;
    pay.init(kb_payload);
    const float4 cone = block_cone<DEG>(ray, dlt, st.alive);
    for (uint32_t b = 0; b < stg.nb; ++b) {
        uint32_t keep;
        const uint32_t cnt = stg.stage_culled(b, cfg, cone, keep);
        for (uint32_t j = 0; j < cnt; ++j) {
            float alpha, t;
            const bool hit = st.alive && record_hit<DEG>(cfg, stg.rec + 5 * j, ray.d, dlt, ray.tmin, ray.tmax, alpha, t);
            const bool need = hit && kb.full();
            if (need) {
                const uint32_t o = kb.front_slot() * 32u + lane;
                process_hit_bwd<DEG>(cfg, src, fr.cam_centre, ray.o, ray.d, pay.t[o], pay.a[o], pay.p[o], st, raw_grad);
            }
            if (hit) {
                const uint32_t slot = kb.insert(t);
                pay.put(slot, lane, t, alpha, __float_as_uint(stg.rec[5 * j + 4].w));
            }
        }
        if (!__any_sync(0xffffffffu, st.alive)) break;
        stg.rotate();
    }
    stg.drain();
#pragma unroll
    for (int i = 0; i < K; ++i) {
        if (st.alive && kb.key[i] != 0u) {
            const uint32_t o = (kb.key[i] & KB_SLOT_MASK) * 32u + lane;
            process_hit_bwd<DEG>(cfg, src, fr.cam_centre, ray.o, ray.d, pay.t[o], pay.a[o], pay.p[o], st, raw_grad);
        }
    }
}

template <int DEG, int K>
cudaError_t launch_fwd(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start, GutParticleSource src, const float* ray_ori,
                       const float* ray_dir, float* rgba, float* dist, float* hits, cudaStream_t stream) {
    gut_render_kbuffer_kernel<DEG, K><<<frame.num_tiles * KB_BLOCKS_PER_TILE, RENDER_THREADS, 0, stream>>>(
        cfg, frame, tile_start, src, ray_ori, ray_dir, reinterpret_cast<float4*>(rgba), dist, hits);
    return cudaGetLastError();
}
template <int DEG, int K>
cudaError_t launch_bwd(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start, GutParticleSource src, const float* ray_ori,
                       const float* ray_dir, const float* rgba, const float* rgba_grad, const float* dist, const float* dist_grad,
                       float* raw_grad, cudaStream_t stream) {
    gut_render_kbuffer_backward_kernel<DEG, K><<<frame.num_tiles * KB_BLOCKS_PER_TILE, RENDER_THREADS, 0, stream>>>(
        cfg, frame, tile_start, src, ray_ori, ray_dir, reinterpret_cast<const float4*>(rgba), reinterpret_cast<const float4*>(rgba_grad),
        dist, dist_grad, raw_grad);
    return cudaGetLastError();
}

}  // namespace

// every K the 4-bit payload slot allows that is a power of two, plus 12 (the reference: any compile-time K, setup_3dgut.py:57)
bool gut_kbuffer_size_supported(int k) { return k == 1 || k == 2 || k == 4 || k == 8 || k == 12 || k == 16; }

#define GUT_DISPATCH_DEG_K(deg, k, CALL)                                  \
    switch (deg) {                                                        \
    case 0: GUT_DISPATCH_K(0, k, CALL)                                    \
    case 1: GUT_DISPATCH_K(1, k, CALL)                                    \
    case 2: GUT_DISPATCH_K(2, k, CALL)                                    \
    case 3: GUT_DISPATCH_K(3, k, CALL)                                    \
    case 4: GUT_DISPATCH_K(4, k, CALL)                                    \
    case 5: GUT_DISPATCH_K(5, k, CALL)                                    \
    case 8: GUT_DISPATCH_K(8, k, CALL)                                    \
    default: return cudaErrorInvalidValue;                                \
    }
#define GUT_DISPATCH_K(D, k, CALL)                                        \
    switch (k) {                                                          \
    case 1: return CALL(D, 1);                                            \
    case 2: return CALL(D, 2);                                            \
    case 4: return CALL(D, 4);                                            \
    case 8: return CALL(D, 8);                                            \
    case 12: return CALL(D, 12);                                          \
    case 16: return CALL(D, 16);                                          \
    default: return cudaErrorInvalidValue;                                \
    }

cudaError_t gut_launch_render_kbuffer(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start, GutParticleSource src,
                                      const float* ray_ori, const float* ray_dir, float* rgba, float* dist, float* hits,
                                      cudaStream_t stream) {
#define CALL_FWD(D, KK) launch_fwd<D, KK>(cfg, frame, tile_start, src, ray_ori, ray_dir, rgba, dist, hits, stream)
    GUT_DISPATCH_DEG_K(cfg.kernel_degree, cfg.k_buffer_size, CALL_FWD)
#undef CALL_FWD
}

cudaError_t gut_launch_render_kbuffer_backward(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start,
                                               GutParticleSource src, const float* ray_ori, const float* ray_dir, const float* rgba,
                                               const float* rgba_grad, const float* dist, const float* dist_grad, float* raw_grad,
                                               cudaStream_t stream) {
#define CALL_BWD(D, KK) \
    launch_bwd<D, KK>(cfg, frame, tile_start, src, ray_ori, ray_dir, rgba, rgba_grad, dist, dist_grad, raw_grad, stream)
    GUT_DISPATCH_DEG_K(cfg.kernel_degree, cfg.k_buffer_size, CALL_BWD)
#undef CALL_BWD
}
