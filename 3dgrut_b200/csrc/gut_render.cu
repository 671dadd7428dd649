// gut_render.cu -- K4 (per-tile forward compositing) and K5 (backward) of the 3DGUT rasteriser.
//
// reference: render / renderBackward kernels (include/3dgut/kernels/cuda/renderers/gutRenderer.cuh:83-165),
//   GUTKBufferRenderer::evalKBuffer / evalBackwardNoKBuffer (renderers/gutKBufferRenderer.cuh:217-386),
//   ray set-up initializeRay / initializeBackwardRay (common/rayPayload.cuh:75-129, rayPayloadBackward.cuh:30-58),
//   hit math gaussianParticle.hit / integrateHit (kernels/slang/models/gaussianParticles.slang:96-254),
//   per-hit gradient processHitBwd (kernels/cuda/models/gaussianParticles.cuh:480-738).
//
// B200 design (DESIGN.md "K4/K5"):
//  * one CTA per 16x16 tile, 8 warps, each warp owns a compact 8x4 pixel block (more lanes hit the same
//    particle than with the reference's 16x2 rows -> better SIMT efficiency in the divergent hit branch);
//  * the tile's particles arrive as a contiguous 80-byte record stream written in sorted order by the
//    tile-sort kernel, so a batch is ONE cp.async.bulk (TMA bulk copy, UBLKCP) into shared memory, completed
//    on an mbarrier and double buffered against compute; no per-thread gathers, no quaternion->matrix math here;
//  * records carry ms = diag(1/s) R^T so a ray/particle test is 2 mat-vecs, a cross product, one MUFU.RCP and one
//    MUFU.EX2; normalisation (MUFU.RSQ) and the hit distance are computed only for accepted hits;
//  * backward: lanes that hit compute the 14 per-particle gradient values; the warp sums them with a 16-shuffle
//    transpose-reduce (not 14 x 5 butterflies) and issues ONE red.global.add.f32 instruction per (warp, particle)
//    -- and nothing at all when no lane of the warp hit (the reference reduces and does 14 atomics regardless,
//    shRadiativeGaussianParticles.cuh:409-482 called outside the isAlive branch at gutKBufferRenderer.cuh:379-383).
#include "gut_internal.h"

namespace {

constexpr int RENDER_THREADS = 256;
constexpr int BATCH = 256;  // records per shared-memory stage (20 KB)

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

// canonical-frame test shared by forward and backward
struct Canon {
    float3 gro, grdu, c;  // canonical ray origin, unnormalised canonical direction, grdu x gro
    float l2, g2, resp, alpha;
};
template <int DEG>
GUT_D bool particle_test(const GutConfig& cfg, const float4 A, const float4 B, const float4 Cm, const float4 S, float3 o, float3 d,
                         Canon& k) {
    const float3 v = f3(o.x - A.w, o.y - B.w, o.z - Cm.w);
    k.gro = f3(A.x * v.x + A.y * v.y + A.z * v.z, B.x * v.x + B.y * v.y + B.z * v.z, Cm.x * v.x + Cm.y * v.y + Cm.z * v.z);
    k.grdu = f3(A.x * d.x + A.y * d.y + A.z * d.z, B.x * d.x + B.y * d.y + B.z * d.z, Cm.x * d.x + Cm.y * d.y + Cm.z * d.z);
    k.l2 = dot(k.grdu, k.grdu);
    k.c = cross(k.grdu, k.gro);
    k.g2 = __fdividef(dot(k.c, k.c), k.l2);  // |normalize(grdu) x gro|^2
    k.resp = kernel_response<DEG>(k.g2);
    k.alpha = fminf(cfg.max_alpha, k.resp * S.w);
    return (k.resp > cfg.min_kernel_response) && (k.alpha > cfg.min_alpha);
}

// ---------------------------------------------------------------------------------------------------
// K4 forward
// ---------------------------------------------------------------------------------------------------
// COUNT = true is the instrumented variant used by gut_measure_work(): it writes no image, it counts the ray x particle
// tests and accepted hits the frame performs (the units the FP32 roofline of K4/K5 is expressed in).
template <int DEG, bool COUNT>
__global__ void __launch_bounds__(RENDER_THREADS) gut_render_kernel(GutConfig cfg, GutFrame fr, const uint32_t* __restrict__ tile_start,
                                                                    const GutRecord* __restrict__ records,
                                                                    const float* __restrict__ ray_ori, const float* __restrict__ ray_dir,
                                                                    float4* __restrict__ rgba, float* __restrict__ dist,
                                                                    float* __restrict__ hits, unsigned long long* __restrict__ counters) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr uint32_t STAGE_BYTES = BATCH * sizeof(GutRecord);
    __shared__ __align__(8) uint64_t full_bar[2];

    const uint32_t tile = blockIdx.x;
    uint32_t px, py;
    thread_pixel(fr, tile, px, py);
    const bool inside = px < fr.width && py < fr.height;
    const uint32_t pix = inside ? px + fr.width * py : 0u;
    Ray ray;
    ray.valid = false;
    if (inside) ray = make_ray(fr, ray_ori, ray_dir, pix);
    bool alive = inside && ray.valid;

    const uint32_t b0 = tile_start[tile], b1 = tile_start[tile + 1];
    const uint32_t n = b1 - b0;
    const uint32_t nb = (n + BATCH - 1) / BATCH;

    if (threadIdx.x == 0) {
        mbar_init(&full_bar[0], 1);
        mbar_init(&full_bar[1], 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (threadIdx.x == 0 && nb > 0) {
        const uint32_t cnt = min((uint32_t)BATCH, n);
        mbar_expect_tx(&full_bar[0], cnt * (uint32_t)sizeof(GutRecord));
        bulk_g2s(smem_raw, records + b0, cnt * (uint32_t)sizeof(GutRecord), &full_bar[0]);
    }

    float T = 1.0f, D = 0.0f, cr = 0.0f, cg = 0.0f, cb = 0.0f;
    uint32_t nhits = 0, ntests = 0;
    uint32_t b = 0;
    for (; b < nb; ++b) {
        const uint32_t cnt = min((uint32_t)BATCH, n - b * BATCH);
        if (threadIdx.x == 0 && b + 1 < nb) {
            // stage (b+1)&1 was last read in iteration b-1, which ended with a CTA barrier
            const uint32_t c2 = min((uint32_t)BATCH, n - (b + 1) * BATCH);
            mbar_expect_tx(&full_bar[(b + 1) & 1], c2 * (uint32_t)sizeof(GutRecord));
            bulk_g2s(smem_raw + ((b + 1) & 1) * STAGE_BYTES, records + b0 + (size_t)(b + 1) * BATCH, c2 * (uint32_t)sizeof(GutRecord),
                     &full_bar[(b + 1) & 1]);
        }
        mbar_wait(&full_bar[b & 1], (b >> 1) & 1);
        const float4* rec = reinterpret_cast<const float4*>(smem_raw + (b & 1) * STAGE_BYTES);
        for (uint32_t j = 0; j < cnt; ++j) {
            if ((j & 15u) == 0u && !__any_sync(0xffffffffu, alive)) break;
            const float4 A = rec[5 * j], B = rec[5 * j + 1], Cm = rec[5 * j + 2], S = rec[5 * j + 3];
            Canon k;
            if (COUNT && alive) ntests++;
            if (particle_test<DEG>(cfg, A, B, Cm, S, ray.o, ray.d, k) && alive) {
                // accepted: hit distance = |scale * grd * dot(grd, -gro)|   (canonicalRayDistance, gaussianParticles.slang:165-171)
                const float il = rsqrtf(k.l2);
                const float3 grd = k.grdu * il;
                const float dp = dot(grd, k.gro);
                const float3 sg = f3(S.x * grd.x, S.y * grd.y, S.z * grd.z);
                const float t = fabsf(dp) * sqrtf(dot(sg, sg));
                if (t > ray.tmin && t < ray.tmax) {
                    const float4 F = rec[5 * j + 4];
                    const float w = k.alpha * T;
                    D += t * w;
                    T *= (1.0f - k.alpha);
                    if (w > 0.0f) { cr += F.x * w; cg += F.y * w; cb += F.z * w; nhits++; }
                    if (T < cfg.min_transmittance) alive = false;
                }
            }
        }
        if (__syncthreads_and(!alive)) break;
    }
    // the tile saturated early while the next batch was already in flight: let the copy land before the CTA retires
    if (b + 1 < nb) mbar_wait(&full_bar[(b + 1) & 1], ((b + 1) >> 1) & 1);
    if (COUNT) {
        unsigned long long t = ntests, hh = nhits;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { t += __shfl_xor_sync(0xffffffffu, t, o); hh += __shfl_xor_sync(0xffffffffu, hh, o); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(&counters[0], t); atomicAdd(&counters[1], hh); }
        return;
    }
    if (inside && ray.valid) {
        rgba[pix] = make_float4(cr, cg, cb, 1.0f - T);
        dist[pix] = D;
        if (cfg.enable_hitcounts) hits[pix] = (float)nhits;
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

template <int DEG>
__global__ void __launch_bounds__(RENDER_THREADS) gut_render_backward_kernel(
    GutConfig cfg, GutFrame fr, const uint32_t* __restrict__ tile_start, const GutRecord* __restrict__ records,
    const float4* __restrict__ quats, const float* __restrict__ ray_ori, const float* __restrict__ ray_dir,
    const float4* __restrict__ rgba, const float4* __restrict__ rgba_grad, const float* __restrict__ dist,
    const float* __restrict__ dist_grad, float* __restrict__ density_grad, float4* __restrict__ feature_grad) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr uint32_t REC_BYTES = BATCH * sizeof(GutRecord), Q_BYTES = BATCH * sizeof(float4);
    constexpr uint32_t STAGE_BYTES = REC_BYTES + Q_BYTES;  // [records | quaternions] per stage
    __shared__ __align__(8) uint64_t full_bar[2];

    const uint32_t tile = blockIdx.x;
    const uint32_t lane = threadIdx.x & 31;
    uint32_t px, py;
    thread_pixel(fr, tile, px, py);
    const bool inside = px < fr.width && py < fr.height;
    const uint32_t pix = inside ? px + fr.width * py : 0u;
    Ray ray;
    ray.valid = false;
    ray.o = f3(0.f, 0.f, 0.f); ray.d = f3(0.f, 0.f, 1.f);
    if (inside) ray = make_ray(fr, ray_ori, ray_dir, pix);
    bool alive = inside && ray.valid;

    // initializeBackwardRay (rayPayloadBackward.cuh:46-55)
    float T_f = 1.0f, D_f = 0.0f, dA = 0.0f, dD = 0.0f;
    float3 C_f = f3(0.f, 0.f, 0.f), dC = f3(0.f, 0.f, 0.f);
    if (alive) {
        const float4 fd = __ldg(rgba + pix), gd = __ldg(rgba_grad + pix);
        T_f = 1.0f - fd.w;          // transmittanceBackward
        dA = gd.w;                  // transmittanceGradient = -dA; used as residualTrm * -(-dA)
        D_f = __ldg(dist + pix);
        dD = __ldg(dist_grad + pix);
        C_f = f3(fd.x, fd.y, fd.z);
        dC = f3(gd.x, gd.y, gd.z);
    }

    const uint32_t b0 = tile_start[tile], b1 = tile_start[tile + 1];
    const uint32_t n = b1 - b0;
    const uint32_t nb = (n + BATCH - 1) / BATCH;

    if (threadIdx.x == 0) {
        mbar_init(&full_bar[0], 1);
        mbar_init(&full_bar[1], 1);
        fence_mbar_init();
    }
    __syncthreads();
    auto issue = [&](uint32_t bi) {
        const uint32_t c = min((uint32_t)BATCH, n - bi * BATCH);
        uint64_t* bar = &full_bar[bi & 1];
        mbar_expect_tx(bar, c * (uint32_t)(sizeof(GutRecord) + sizeof(float4)));
        unsigned char* base = smem_raw + (bi & 1) * STAGE_BYTES;
        bulk_g2s(base, records + b0 + (size_t)bi * BATCH, c * (uint32_t)sizeof(GutRecord), bar);
        bulk_g2s(base + REC_BYTES, quats + b0 + (size_t)bi * BATCH, c * (uint32_t)sizeof(float4), bar);
    };
    if (threadIdx.x == 0 && nb > 0) issue(0);

    float T = 1.0f, D = 0.0f;
    float3 C = f3(0.f, 0.f, 0.f);
    const float minT = cfg.min_transmittance;

    uint32_t b = 0;
    for (; b < nb; ++b) {
        const uint32_t cnt = min((uint32_t)BATCH, n - b * BATCH);
        if (threadIdx.x == 0 && b + 1 < nb) issue(b + 1);
        mbar_wait(&full_bar[b & 1], (b >> 1) & 1);
        const float4* rec = reinterpret_cast<const float4*>(smem_raw + (b & 1) * STAGE_BYTES);
        const float4* qrec = reinterpret_cast<const float4*>(smem_raw + (b & 1) * STAGE_BYTES + REC_BYTES);
        for (uint32_t j = 0; j < cnt; ++j) {
            if ((j & 15u) == 0u && !__any_sync(0xffffffffu, alive)) break;
            const float4 A = rec[5 * j], B = rec[5 * j + 1], Cm = rec[5 * j + 2], S = rec[5 * j + 3];
            Canon k;
            const bool hit = particle_test<DEG>(cfg, A, B, Cm, S, ray.o, ray.d, k) && alive;  // no tMinMax test here
            if (!__any_sync(0xffffffffu, hit)) continue;

            float v[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) v[q] = 0.0f;
            const float4 F = rec[5 * j + 4];
            if (hit) {
                const float4 Q = qrec[j];  // w,x,y,z
                const float3 s = f3(S.x, S.y, S.z);
                const float3 gro = k.gro;
                const float il = rsqrtf(k.l2);
                const float3 grd = k.grdu * il;
                const float dgo = dot(grd, gro);
                const float3 grdd = grd * (-dgo);
                const float3 grds = s * grdd;
                const float gsq = dot(grds, grds);
                const float gdist = sqrtf(gsq);
                const float alpha = k.alpha;
                const float w = alpha * T;
                const float nextT = (1.0f - alpha) * T;

                D += w * gdist;
                const float resD = fmaxf(nextT <= minT ? 0.0f : __fdividef(D_f - D, nextT), 0.0f);
                const float aHit = (gdist - resD) * T * dD;
                const float hs = gsq > 0.0f ? __fdividef(w, gdist) * dD : 0.0f;
                const float3 grdsG = grds * hs;
                const float3 sclHit = grdd * grdsG;
                const float3 grdHit = f3(-s.x * (grd.x * gro.x + dgo) * grdsG.x, -s.y * (grd.y * gro.y + dgo) * grdsG.y,
                                         -s.z * (grd.z * gro.z + dgo) * grdsG.z);
                const float3 groHit = f3(-s.x * grd.x * grd.x * grdsG.x, -s.y * grd.y * grd.y * grdsG.y, -s.z * grd.z * grd.z * grdsG.z);

                const float resT = alpha < 0.999999f ? __fdividef(T_f, 1.0f - alpha) : T;
                const float aDns = resT * dA;

                C = f3(C.x + w * F.x, C.y + w * F.y, C.z + w * F.z);
                float3 resC = f3(0.f, 0.f, 0.f);
                if (!(nextT <= minT)) {
                    const float inv = __fdividef(1.0f, nextT);
                    resC = f3(fmaxf((C_f.x - C.x) * inv, 0.0f), fmaxf((C_f.y - C.y) * inv, 0.0f), fmaxf((C_f.z - C.z) * inv, 0.0f));
                }
                const float Ssum = aHit + aDns + T * ((F.x - resC.x) * dC.x + (F.y - resC.y) * dC.y + (F.z - resC.z) * dC.z);
                const float dDens = k.resp * Ssum;
                const float respG = S.w * Ssum;
                const float g2G = kernel_response_grad<DEG>(k.g2, k.resp, respG);

                // g2 = |grd x gro|^2 : adjoint of the cross product (gaussianParticles.cuh:671-689)
                const float3 gc = k.c * il;
                const float3 gcG = gc * (2.0f * g2G);
                const float3 grdG = f3(gcG.z * gro.y - gcG.y * gro.z, gcG.x * gro.z - gcG.z * gro.x, gcG.y * gro.x - gcG.x * gro.y);
                const float3 groG = f3(gcG.y * grd.z - gcG.z * grd.y, gcG.z * grd.x - gcG.x * grd.z, gcG.x * grd.y - gcG.y * grd.x);
                const float3 groTot = groG + groHit;

                const float3 gis = f3(__fdividef(1.0f, s.x), __fdividef(1.0f, s.y), __fdividef(1.0f, s.z));
                // gro = gposcr / s : d/ds = -gposcr/s^2 = -gro/s ; d/dgposcr = 1/s
                const float3 sclGro = f3(-gro.x * gis.x * groTot.x, -gro.y * gis.y * groTot.y, -gro.z * gis.z * groTot.z);
                const float3 gposcrG = gis * groTot;
                // R (gposcrG) with rows of R^T = ms_i * s_i  ==> sum_i groTot_i * ms_i
                const float3 gposcG = f3(groTot.x * A.x + groTot.y * B.x + groTot.z * Cm.x, groTot.x * A.y + groTot.y * B.y + groTot.z * Cm.y,
                                         groTot.x * A.z + groTot.y * B.z + groTot.z * Cm.z);
                // normalisation adjoint (safe_normalize_bw, mathUtils.cuh:420-430): il * (X - grd (grd . X))
                const float3 X = grdG + grdHit;
                const float gx = dot(grd, X);
                const float3 grduG = f3(il * (X.x - grd.x * gx), il * (X.y - grd.y * gx), il * (X.z - grd.z * gx));
                // grdu = rayDirR / s
                const float3 sclDir = f3(-k.grdu.x * gis.x * grduG.x, -k.grdu.y * gis.y * grduG.y, -k.grdu.z * gis.z * grduG.z);
                const float3 rayDirRG = gis * grduG;

                // dL/dR^T = gposcrG (x) (o - pos) + rayDirRG (x) d, contracted with dR^T/dq (matmul_bw_quat, mathUtils.cuh:468-533)
                const float3 vv = f3(ray.o.x - A.w, ray.o.y - B.w, ray.o.z - Cm.w);
                const float3 dd = ray.d;
                const float d00 = gposcrG.x * vv.x + rayDirRG.x * dd.x, d01 = gposcrG.x * vv.y + rayDirRG.x * dd.y, d02 = gposcrG.x * vv.z + rayDirRG.x * dd.z;
                const float d10 = gposcrG.y * vv.x + rayDirRG.y * dd.x, d11 = gposcrG.y * vv.y + rayDirRG.y * dd.y, d12 = gposcrG.y * vv.z + rayDirRG.y * dd.z;
                const float d20 = gposcrG.z * vv.x + rayDirRG.z * dd.x, d21 = gposcrG.z * vv.y + rayDirRG.z * dd.y, d22 = gposcrG.z * vv.z + rayDirRG.z * dd.z;
                const float qr = Q.x, qx = Q.y, qy = Q.z, qz = Q.w;
                const float s01 = d01 + d10, a01 = d01 - d10, s02 = d02 + d20, a20 = d20 - d02, s12 = d12 + d21, a12 = d12 - d21;

                v[0] = -gposcG.x; v[1] = -gposcG.y; v[2] = -gposcG.z;
                v[3] = dDens;
                v[4] = 2.0f * (qz * a01 + qy * a20 + qx * a12);
                v[5] = 2.0f * (qy * s01 + qz * s02 + qr * a12 - 2.0f * qx * (d11 + d22));
                v[6] = 2.0f * (qx * s01 + qz * s12 + qr * a20 - 2.0f * qy * (d00 + d22));
                v[7] = 2.0f * (qx * s02 + qy * s12 + qr * a01 - 2.0f * qz * (d00 + d11));
                v[8] = sclHit.x + sclGro.x + sclDir.x; v[9] = sclHit.y + sclGro.y + sclDir.y; v[10] = sclHit.z + sclGro.z + sclDir.z;
                v[11] = w * dC.x; v[12] = w * dC.y; v[13] = w * dC.z;

                T = nextT;
                if (T < minT) alive = false;
            }
            const float total = warp_transpose_reduce16(v, lane);
            const uint32_t slot = ((lane >> 4) & 1u) * 8u + ((lane >> 3) & 1u) * 4u + ((lane >> 2) & 1u) * 2u + ((lane >> 1) & 1u);
            if ((lane & 1u) == 0u && slot < 14u) {
                const uint32_t p = __float_as_uint(F.w);
                float* dst = slot < 11u ? density_grad + 12 * (size_t)p + slot : reinterpret_cast<float*>(feature_grad + p) + (slot - 11u);
                atomicAdd(dst, total);
            }
        }
        if (__syncthreads_and(!alive)) break;
    }
    if (b + 1 < nb) mbar_wait(&full_bar[(b + 1) & 1], ((b + 1) >> 1) & 1);
}

template <int DEG>
cudaError_t launch_render_deg(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start, const GutRecord* records,
                              const float* ray_ori, const float* ray_dir, float* rgba, float* dist, float* hits,
                              unsigned long long* counters, cudaStream_t stream) {
    const size_t smem = 2 * BATCH * sizeof(GutRecord);
    cudaError_t e;
    if (counters) {
        e = cudaFuncSetAttribute(gut_render_kernel<DEG, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        gut_render_kernel<DEG, true><<<frame.num_tiles, RENDER_THREADS, smem, stream>>>(cfg, frame, tile_start, records, ray_ori, ray_dir,
                                                                                      nullptr, nullptr, nullptr, counters);
    } else {
        e = cudaFuncSetAttribute(gut_render_kernel<DEG, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        gut_render_kernel<DEG, false><<<frame.num_tiles, RENDER_THREADS, smem, stream>>>(cfg, frame, tile_start, records, ray_ori, ray_dir,
                                                                                       reinterpret_cast<float4*>(rgba), dist, hits, nullptr);
    }
    return cudaGetLastError();
}

template <int DEG>
cudaError_t launch_render_bwd_deg(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start, const GutRecord* records,
                                  const float4* quats, const float* ray_ori, const float* ray_dir, const float* rgba,
                                  const float* rgba_grad, const float* dist, const float* dist_grad, float* density_grad,
                                  float4* feature_grad, cudaStream_t stream) {
    const size_t smem = 2 * BATCH * (sizeof(GutRecord) + sizeof(float4));
    cudaError_t e = cudaFuncSetAttribute(gut_render_backward_kernel<DEG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    gut_render_backward_kernel<DEG><<<frame.num_tiles, RENDER_THREADS, smem, stream>>>(
        cfg, frame, tile_start, records, quats, ray_ori, ray_dir, reinterpret_cast<const float4*>(rgba),
        reinterpret_cast<const float4*>(rgba_grad), dist, dist_grad, density_grad, feature_grad);
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

cudaError_t gut_launch_render(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start, const GutRecord* records,
                              const float* ray_ori, const float* ray_dir, float* rgba, float* dist, float* hits,
                              unsigned long long* counters, cudaStream_t stream) {
#define CALL_FWD(D) launch_render_deg<D>(cfg, frame, tile_start, records, ray_ori, ray_dir, rgba, dist, hits, counters, stream)
    GUT_DISPATCH_DEGREE(cfg.kernel_degree, CALL_FWD)
#undef CALL_FWD
}

cudaError_t gut_launch_render_backward(const GutConfig& cfg, const GutFrame& frame, const uint32_t* tile_start,
                                       const GutRecord* records, const float4* quats, const float* ray_ori, const float* ray_dir,
                                       const float* rgba, const float* rgba_grad, const float* dist, const float* dist_grad,
                                       float* density_grad, float4* feature_grad, cudaStream_t stream) {
#define CALL_BWD(D) \
    launch_render_bwd_deg<D>(cfg, frame, tile_start, records, quats, ray_ori, ray_dir, rgba, rgba_grad, dist, dist_grad, density_grad, feature_grad, stream)
    GUT_DISPATCH_DEGREE(cfg.kernel_degree, CALL_BWD)
#undef CALL_BWD
}
