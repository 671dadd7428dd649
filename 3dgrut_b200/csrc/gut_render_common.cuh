// gut_render_common.cuh -- device helpers shared by the compositing kernels (gut_render.cu: k_buffer_size == 0,
// gut_render_kbuffer.cu: k_buffer_size > 0): kernel response, MUFU wrappers, ray set-up, warp-private record staging.
#pragma once
#include "gut_internal.h"

namespace {

constexpr int RENDER_THREADS = 32;    // one warp per CTA
constexpr int BATCH = 32;             // particles per staged batch = one per lane

// The records of a batch are walked four at a time; the tail of a tile's last batch is padded in shared memory with
// records that can never be accepted (density 0), so the inner loop has no remainder handling.
constexpr uint32_t WALK_UNROLL = 4;
constexpr uint32_t ALIVE_CHECK = 16;  // records between two warp-wide "is any ray still alive" votes

GUT_D uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

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

// ---- warp-private batch staging shared by K4 and K5 -----------------------------------------------------
GUT_D void cp_async16(void* sdst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(sdst)), "l"(gsrc) : "memory");
}
GUT_D void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
GUT_D void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <bool BWD>
struct StageLayout {
    static constexpr uint32_t REC = BATCH * sizeof(GutRecord);
    static constexpr uint32_t RECB = BWD ? BATCH * sizeof(GutRecordBwd) : 0;
    static constexpr uint32_t RAW = BATCH * 4 * sizeof(float4);
    static constexpr uint32_t TOTAL = REC + RECB + RAW;
};

// Staging pipeline of ONE warp (= one CTA, see the header comment).  A batch is 32 consecutive entries of the tile's
// sorted list, one per lane.  Batch bi: the lane's list entry is a coalesced load two batches ahead (kept in a
// register), its parameters are gathered by cp.async one batch ahead into `raw`, and right before the walk the lane
// expands them into its record.  All ordering is warp-local: __syncwarp and cp.async.wait_group, no CTA barrier.
template <bool BWD>
struct WarpStager {
    float4* rec; float4* recb; float4* raw;
    GutParticleSource src;
    uint32_t b0, n, nb, lane;
    float3 origin;
    uint32_t p_cur, p_next, p_after;   // this lane's particle of batch bi, bi+1, bi+2

    GUT_D uint32_t count(uint32_t bi) const { return min((uint32_t)BATCH, n - bi * BATCH); }
    GUT_D uint32_t load_idx(uint32_t bi) const {
        return (bi < nb && lane < count(bi)) ? __ldg(src.sorted_idx + b0 + bi * BATCH + lane) : 0u;
    }
    GUT_D void issue_gather(uint32_t bi, uint32_t p) {
        if (bi < nb && lane < count(bi)) {
            cp_async16(raw + lane, src.density12 + 3 * (size_t)p);
            cp_async16(raw + BATCH + lane, src.density12 + 3 * (size_t)p + 1);
            cp_async16(raw + 2 * BATCH + lane, src.density12 + 3 * (size_t)p + 2);
            cp_async16(raw + 3 * BATCH + lane, src.features + p);
        }
        cp_async_commit();
    }
    GUT_D void init(unsigned char* smem, const GutParticleSource& s, uint32_t b0_, uint32_t n_, float3 origin_) {
        rec = reinterpret_cast<float4*>(smem);
        recb = reinterpret_cast<float4*>(smem + StageLayout<BWD>::REC);
        raw = reinterpret_cast<float4*>(smem + StageLayout<BWD>::REC + StageLayout<BWD>::RECB);
        src = s; b0 = b0_; n = n_; nb = (n_ + BATCH - 1) / BATCH; origin = origin_; lane = threadIdx.x;
        p_cur = load_idx(0); p_next = load_idx(1); p_after = 0;
        issue_gather(0, p_cur);
    }
    // gathered parameters of batch bi -> records; then put batch bi+1 (gather) and bi+2 (list entry) in flight.
    // Lanes past the end of the list pad it to a multiple of WALK_UNROLL with records that can never be accepted:
    // identity frame (so that u = d is never zero), density 0.
    GUT_D void stage(uint32_t bi) {
        cp_async_wait_all();
        __syncwarp();   // the walk of batch bi-1 is over in every lane: rec may be overwritten
        const uint32_t cnt = count(bi);
        const uint32_t cnt_pad = (cnt + WALK_UNROLL - 1) & ~(WALK_UNROLL - 1);
        float4* out = rec + 5 * lane;
        if (lane < cnt) {
            const float4 r0 = raw[lane], q = raw[BATCH + lane], r2 = raw[2 * BATCH + lane], ft = raw[3 * BATCH + lane];
            float3 c0, c1, c2;
            quat_to_columns(q.x, q.y, q.z, q.w, c0, c1, c2);   // columns of R == rows of R^T
            const float ix = mufu_rcp(r2.x), iy = mufu_rcp(r2.y), iz = mufu_rcp(r2.z);
            const float3 m0 = c0 * ix, m1 = c1 * iy, m2 = c2 * iz;   // rows of ms = diag(1/s) R^T
            const float3 v = origin - f3(r0.x, r0.y, r0.z);          // frame origin - position
            out[0] = make_float4(m0.x, m0.y, m0.z, dot(m0, v));
            out[1] = make_float4(m1.x, m1.y, m1.z, dot(m1, v));
            out[2] = make_float4(m2.x, m2.y, m2.z, dot(m2, v));
            out[3] = make_float4(r2.x, r2.y, r2.z, r0.w);
            out[4] = make_float4(fmaxf(ft.x, 0.0f), fmaxf(ft.y, 0.0f), fmaxf(ft.z, 0.0f), __uint_as_float(p_cur));
            if (BWD) {   // 1 / s and s^2
                float4* outb = recb + 2 * lane;
                outb[0] = make_float4(ix, iy, iz, 0.0f);
                outb[1] = make_float4(r2.x * r2.x, r2.y * r2.y, r2.z * r2.z, 0.0f);
            }
        } else if (lane < cnt_pad) {
            out[0] = make_float4(1.0f, 0.0f, 0.0f, 0.0f);
            out[1] = make_float4(0.0f, 1.0f, 0.0f, 0.0f);
            out[2] = make_float4(0.0f, 0.0f, 1.0f, 0.0f);
            out[3] = make_float4(1.0f, 1.0f, 1.0f, 0.0f);
            out[4] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        }
        __syncwarp();   // raw is consumed, rec is published
        issue_gather(bi + 1, p_next);
        p_after = load_idx(bi + 2);
    }
    // Forward only: stage + block-level cull.  The lane decides whether ANY ray of the 8x8 block can accept its record,
    // from a cone that contains the block's ray directions (axis cone.xyz, tan of the half angle cone.w; w < 0: no cull).
    // A ray accepts when its canonical distance^2 g2 = |go|^2 sin^2(angle(u, go)) is below
    // g2max = 2 min(ln(1 / min_response), ln(density / min_alpha)) (degree-2 kernel).  Every ray direction of the block is
    // axis + tan(a) e with a <= half angle and e a unit vector, so u = ms d deviates from uc = ms axis by at most
    // asin(rho), rho = tan(half angle) |ms|_2 / |uc|, |ms|_2 = 1 / min scale.  The smallest possible g2 is therefore
    // |go|^2 sin^2(phi_c - asin(rho)); the record is dropped when even that is above g2max (2 % margin on g2 and an
    // absolute one on the sine, far above the fp32 error of either side).  Conservative: image and hit masks are
    // bit-identical to walking every record.  Kept records are compacted to the front of the batch in list order, the
    // rest of the batch becomes padding; `keep` has a bit per list position.  Returns the number of kept records.
    GUT_D uint32_t stage_culled(uint32_t bi, const GutConfig& cfg, const float4 cone, uint32_t& keep) {
        static_assert(!BWD, "forward only");
        cp_async_wait_all();
        __syncwarp();   // the walk of batch bi-1 is over in every lane: rec may be overwritten
        const uint32_t cnt = count(bi);
        float4 o0 = make_float4(1.0f, 0.0f, 0.0f, 0.0f), o1 = make_float4(0.0f, 1.0f, 0.0f, 0.0f), o2 = make_float4(0.0f, 0.0f, 1.0f, 0.0f);
        float4 o3 = make_float4(1.0f, 1.0f, 1.0f, 0.0f), o4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        bool k = false;
        if (lane < cnt) {
            const float4 r0 = raw[lane], q = raw[BATCH + lane], r2 = raw[2 * BATCH + lane], ft = raw[3 * BATCH + lane];
            float3 c0, c1, c2;
            quat_to_columns(q.x, q.y, q.z, q.w, c0, c1, c2);   // columns of R == rows of R^T
            const float ix = mufu_rcp(r2.x), iy = mufu_rcp(r2.y), iz = mufu_rcp(r2.z);
            const float3 m0 = c0 * ix, m1 = c1 * iy, m2 = c2 * iz;   // rows of ms = diag(1/s) R^T
            const float3 v = origin - f3(r0.x, r0.y, r0.z);          // frame origin - position
            const float3 go = f3(dot(m0, v), dot(m1, v), dot(m2, v));
            k = true;
            if (cone.w >= 0.0f && !(r0.w > cfg.min_alpha)) k = false;   // araw = resp * density never passes min_alpha
            if (k && cone.w >= 0.0f) {
                const float3 ax = f3(cone.x, cone.y, cone.z);
                const float3 uc = f3(dot(m0, ax), dot(m1, ax), dot(m2, ax));
                const float l2 = dot(uc, uc), gg = dot(go, go), ug = dot(uc, go);
                const float3 cr = cross(uc, go);
                const float il = mufu_rsq(l2), inv = il * mufu_rsq(gg);
                const float sin_c = mufu_sqrt(dot(cr, cr)) * inv, cos_c = fabsf(ug) * inv;
                const float rho = cone.w * il * fmaxf(ix, fmaxf(iy, iz));
                // ln(x) = lg2(x) ln 2
                const float g2max = 1.3862943611f * fminf(__log2f(1.0f / cfg.min_kernel_response), __log2f(r0.w / cfg.min_alpha));
                const float lhs = sin_c * mufu_sqrt(fmaxf(1.0f - rho * rho, 0.0f)) - cos_c * rho - 4e-6f;
                if ((rho < 1.0f) && (lhs > 0.0f) && (lhs * lhs * gg > 1.02f * g2max + 1e-3f)) k = false;
            }
            if (k) {
                o0 = make_float4(m0.x, m0.y, m0.z, go.x);
                o1 = make_float4(m1.x, m1.y, m1.z, go.y);
                o2 = make_float4(m2.x, m2.y, m2.z, go.z);
                o3 = make_float4(r2.x, r2.y, r2.z, r0.w);
                o4 = make_float4(fmaxf(ft.x, 0.0f), fmaxf(ft.y, 0.0f), fmaxf(ft.z, 0.0f), __uint_as_float(p_cur));
            }
        }
        keep = __ballot_sync(0xffffffffu, k);
        const uint32_t kept = __popc(keep), below = (1u << lane) - 1u;
        float4* out = rec + 5 * (k ? __popc(keep & below) : kept + __popc(~keep & below));
        out[0] = o0; out[1] = o1; out[2] = o2; out[3] = o3; out[4] = o4;
        __syncwarp();   // raw is consumed, rec is published
        issue_gather(bi + 1, p_next);
        p_after = load_idx(bi + 2);
        return kept;
    }
    GUT_D void rotate() { p_cur = p_next; p_next = p_after; }
    // before the warp retires: nothing may still be landing in its shared memory
    GUT_D void drain() { cp_async_wait_all(); }
};

}  // namespace
