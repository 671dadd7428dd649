// gut_project.cu -- K1 (per-particle unscented-transform projection + SH colour + tile counting) and
// K6 (SH backward).  Built with --fmad=false (see gut_math.cuh): integer results of this stage are
// bit-reproducible against the CPU oracle.
//
// reference: GUTProjector::eval / unscentedParticleProjection / computeProjectedExtentConicOpacity
//            (threedgut_tracer/include/3dgut/kernels/cuda/renderers/gutProjector.cuh:81-322),
//            projectPoint* (include/3dgut/kernels/cuda/sensors/cameraProjections.cuh:22-185),
//            radianceFromSpH / radianceFromSpHBwd (include/3dgut/kernels/cuda/models/gaussianParticles.cuh:64-187),
//            GUTProjector::evalBackward (gutProjector.cuh:390-430).
#include <float.h>

#include <type_traits>

#include "gut_internal.h"

namespace {

struct Quat { float w, x, y, z; };

GUT_D Quat pose_quat(const float* p) { return Quat{p[6], p[3], p[4], p[5]}; }
GUT_D float3 pose_t(const float* p) { return f3(p[0], p[1], p[2]); }
GUT_D float mixf(float a, float b, float c) { return a * (1.0f - c) + b * c; }

// tcnn::slerp semantics (thirdparty/tiny-cuda-nn/include/tiny-cuda-nn/vec.h:1133-1153): shortest arc, lerp when nearly parallel,
// result NOT renormalised.
GUT_D Quat slerp(Quat a, Quat b, float t) {
    float c = (a.w * b.w + a.x * b.x) + (a.y * b.y + a.z * b.z);
    if (c < 0.0f) { b.w = -b.w; b.x = -b.x; b.y = -b.y; b.z = -b.z; c = -c; }
    Quat r;
    if (c > 1.0f - FLT_EPSILON) {
        const float u = 1.0f - t;
        r.w = a.w * u + b.w * t; r.x = a.x * u + b.x * t; r.y = a.y * u + b.y * t; r.z = a.z * u + b.z * t;
    } else {
        const float ang = acosf(c);
        const float s0 = sinf((1.0f - t) * ang), s1 = sinf(t * ang), sd = sinf(ang);
        r.w = (s0 * a.w + s1 * b.w) / sd; r.x = (s0 * a.x + s1 * b.x) / sd;
        r.y = (s0 * a.y + s1 * b.y) / sd; r.z = (s0 * a.z + s1 * b.z) / sd;
    }
    return r;
}

// world -> sensor as a matrix: the shutter-start pose is the same for every sigma point of every particle, so a thread
// forms it once instead of once per projected point
struct PoseMat { float3 c0, c1, c2, t; };
GUT_D PoseMat pose_matrix(Quat q, float3 t) {
    PoseMat m;
    quat_to_columns(q.w, q.x, q.y, q.z, m.c0, m.c1, m.c2);
    m.t = t;
    return m;
}
GUT_D float3 to_sensor(const PoseMat& m, float3 p) { return rotate_cols(m.c0, m.c1, m.c2, p) + m.t; }
GUT_D float3 to_sensor(Quat q, float3 t, float3 p) { return to_sensor(pose_matrix(q, t), p); }

GUT_D bool within_resolution(float rx, float ry, float tol, float2 p) {
    const float mx = rx * tol, my = ry * tol;
    return (p.x > -mx) && (p.y > -my) && (p.x < rx + mx) && (p.y < ry + my);
}

// OpenCV pinhole with rational radial, tangential and thin-prism terms (cameraProjections.cuh:57-103)
// `ideal`: every distortion coefficient is zero (frame-uniform).  Then num = den = icd = 1 and dx = dy = 0 exactly for any
// finite r2, and (1 * u + 0) * f + c rounds like u * f + c: the short path below returns the same bits.
GUT_D bool project_pinhole(const GutCamera& cam, float fw, float fh, float3 pos, float tol, float2& out, bool ideal) {
    if (pos.z <= 0.0f) { out = make_float2(0.0f, 0.0f); return false; }
    const float ux = pos.x / pos.z, uy = pos.y / pos.z;
    const float sx = ux * ux, sy = uy * uy;
    const float r2 = sx + sy;
    if (ideal && r2 <= 1e18f) {   // r2^3 stays finite
        out.x = (ux + 0.0f) * cam.focal_length[0] + cam.principal_point[0];   // + 0: -0 -> +0 as the long path does
        out.y = (uy + 0.0f) * cam.focal_length[1] + cam.principal_point[1];
        return within_resolution(fw, fh, tol, out);
    }
    const float a1 = 2.0f * ux * uy, a2 = r2 + 2.0f * sx, a3 = r2 + 2.0f * sy;
    const float* k = cam.radial_coeffs; const float* p = cam.tangential_coeffs; const float* s = cam.thin_prism_coeffs;
    const float num = 1.0f + r2 * (k[0] + r2 * (k[1] + r2 * k[2]));
    const float den = 1.0f + r2 * (k[3] + r2 * (k[4] + r2 * k[5]));
    const float icd = num / den;
    const float dx = p[0] * a1 + p[1] * a2 + r2 * (s[0] + r2 * s[1]);
    const float dy = p[0] * a3 + p[1] * a1 + r2 * (s[2] + r2 * s[3]);
    const bool radial_ok = (icd > 0.8f) && (icd < 1.2f);
    if (radial_ok) {
        out.x = (icd * ux + dx) * cam.focal_length[0] + cam.principal_point[0];
        out.y = (icd * uy + dy) * cam.focal_length[1] + cam.principal_point[1];
    } else {
        const float f = sqrtf(fw * fw + fh * fh) / sqrtf(r2);
        out.x = f * ux + cam.principal_point[0];
        out.y = f * uy + cam.principal_point[1];
    }
    return radial_ok && within_resolution(fw, fh, tol, out);
}

// OpenCV fisheye (equidistant + 4-term polynomial), angle clamped at max_angle (cameraProjections.cuh:105-128)
GUT_D bool project_fisheye(const GutCamera& cam, float fw, float fh, float3 pos, float tol, float2& out) {
    const float rho = fmaxf(sqrtf(pos.x * pos.x + pos.y * pos.y), FLT_EPSILON);
    const float theta = fminf(atan2f(rho, pos.z), cam.max_angle);
    const float t2 = theta * theta;
    const float* c = cam.radial_coeffs;
    float y = c[3];
    y = t2 * y + c[2];
    y = t2 * y + c[1];
    y = t2 * y + c[0];
    const float delta = (theta * (y * t2 + 1.0f)) / rho;
    out.x = cam.focal_length[0] * pos.x * delta + cam.principal_point[0];
    out.y = cam.focal_length[1] * pos.y * delta + cam.principal_point[1];
    return (theta < cam.max_angle) && within_resolution(fw, fh, tol, out);
}

GUT_D bool pinhole_is_ideal(const GutCamera& cam) {
    bool z = cam.model == GUT_CAMERA_OPENCV_PINHOLE;
#pragma unroll
    for (int k = 0; k < 6; ++k) z = z && cam.radial_coeffs[k] == 0.0f;
#pragma unroll
    for (int k = 0; k < 2; ++k) z = z && cam.tangential_coeffs[k] == 0.0f;
#pragma unroll
    for (int k = 0; k < 4; ++k) z = z && cam.thin_prism_coeffs[k] == 0.0f;
    return z;
}

GUT_D bool project_point(const GutCamera& cam, float fw, float fh, float3 pos, float tol, float2& out, bool ideal) {
    if (cam.model == GUT_CAMERA_OPENCV_PINHOLE) return project_pinhole(cam, fw, fh, pos, tol, out, ideal);
    if (cam.model == GUT_CAMERA_OPENCV_FISHEYE) return project_fisheye(cam, fw, fh, pos, tol, out);
    out = make_float2(0.0f, 0.0f);
    return false;
}

GUT_D float shutter_time(const GutCamera& cam, float fw, float fh, float2 p) {
    switch (cam.shutter) {
    case GUT_SHUTTER_ROLLING_TOP_TO_BOTTOM: return floorf(p.y) / (fh - 1.0f);
    case GUT_SHUTTER_ROLLING_LEFT_TO_RIGHT: return floorf(p.x) / (fw - 1.0f);
    case GUT_SHUTTER_ROLLING_BOTTOM_TO_TOP: return (fh - ceilf(p.y)) / (fh - 1.0f);
    case GUT_SHUTTER_ROLLING_RIGHT_TO_LEFT: return (fw - ceilf(p.x)) / (fw - 1.0f);
    default: return 0.5f;
    }
}

// Project with the start pose; under a rolling shutter refine the capture time of the point by fixed-point
// iteration on its own image row/column (cameraProjections.cuh:146-185).
// Project with the start pose; under a rolling shutter refine the capture time of the point by fixed-point
// iteration on its own image row/column (cameraProjections.cuh:146-185).
template <bool ROLLING>
GUT_D bool project_with_shutter(const GutCamera& cam, float fw, float fh, int n_iter, float3 pos, float tol, float2& out,
                                const PoseMat& m0, bool ideal) {
    bool ok = project_point(cam, fw, fh, to_sensor(m0, pos), tol, out, ideal);
    if (!ROLLING) return ok;
    const Quat q0 = pose_quat(cam.start_pose);
    const float3 t0 = pose_t(cam.start_pose);
    const Quat q1 = pose_quat(cam.end_pose);
    const float3 t1 = pose_t(cam.end_pose);
    if (!ok) {
        ok = project_point(cam, fw, fh, to_sensor(q1, t1, pos), tol, out, ideal);
        if (!ok) return false;
    }
    for (int i = 0; i < n_iter; ++i) {
        const float a = shutter_time(cam, fw, fh, out);
        const float3 t = f3(mixf(t0.x, t1.x, a), mixf(t0.y, t1.y, a), mixf(t0.z, t1.z, a));
        ok = project_point(cam, fw, fh, to_sensor(slerp(q0, q1, a), t, pos), tol, out, ideal);
    }
    return ok;
}

GUT_D float view_depth(const GutMat4x3& v, float3 p) { return p.x * v.c[0].z + p.y * v.c[1].z + p.z * v.c[2].z + v.c[3].z; }

// Dilated conic, mip-splatting opacity, culling bound and screen extent (gutProjector.cuh:81-116).
GUT_D bool conic_extent(const GutConfig& cfg, float3 cov, float opacity, float2& extent, float4& co, float& max_power) {
    const float dil = 0.3f;
    const float dx = cov.x + dil, dy = cov.y, dz = cov.z + dil;
    const float ddet = dx * dz - dy * dy;
    if (ddet == 0.0f) return false;
    co.x = dz / ddet; co.y = -dy / ddet; co.z = dx / ddet;
    const float det = cov.x * cov.z - cov.y * cov.y;
    co.w = opacity * sqrtf(fmaxf(0.000025f, det / ddet));
    if (co.w < cfg.min_alpha) return false;
    max_power = gut_logf(co.w / cfg.min_alpha);
    const float ef = cfg.tight_opacity_bounding ? fminf(3.33f, sqrtf(2.0f * max_power)) : 3.33f;
    const float mid = 0.5f * (dx + dz);
    const float lambda = mid + sqrtf(fmaxf(0.01f, mid * mid - ddet));
    const float radius = ef * sqrtf(lambda);
    if (cfg.rect_bounding) {
        extent.x = fminf(ef * sqrtf(dx), radius);
        extent.y = fminf(ef * sqrtf(dz), radius);
    } else {
        extent.x = radius; extent.y = radius;
    }
    return radius > 0.0f;
}

// ---------------------------------------------------------------------------------------------------
// K1.  One thread per particle.  The 48-byte parameter record is fetched as three 128-bit loads.
// ---------------------------------------------------------------------------------------------------
#ifndef GUT_K1_MIN_BLOCKS
#define GUT_K1_MIN_BLOCKS 3          // rolling-shutter variant: 70 registers
#define GUT_K1_MIN_BLOCKS_GLOBAL 4   // global-shutter variant: 64 registers, four CTAs per SM
#endif
// ROLLING is the frame's shutter kind (launch-level): a global-shutter frame runs a kernel that holds no rolling-shutter
// code at all, with the seven sigma points unrolled (0.277 ms at 3M particles; rolled: 0.306); a rolling-shutter frame
// runs them through ONE copy of the projection code in a rolled loop -- seven inlined copies of the shutter iteration with
// its slerp / acosf / sinf made a 13 000-instruction kernel that waited 12 % of its warp time for the instruction cache
// (C4, 1M particles: 0.659 -> 0.573 ms).
template <bool ROLLING>
__global__ void __launch_bounds__(256, ROLLING ? GUT_K1_MIN_BLOCKS : GUT_K1_MIN_BLOCKS_GLOBAL) gut_project_kernel(GutConfig cfg, GutCamera cam, GutFrame fr, uint32_t N, int sph_degree,
                                                          const float4* __restrict__ density12, const float* __restrict__ sph,
                                                          GutProjBuffers out, float* __restrict__ visibility,
                                                          uint32_t* __restrict__ tile_hist) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool in_range = i < N;                              // out-of-range lanes stay for the warp-cooperative tile walk
    const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    const float4 r0 = in_range ? __ldg(density12 + 3 * (size_t)i) : zero4;      // pos.xyz, density
    const float4 r1 = in_range ? __ldg(density12 + 3 * (size_t)i + 1) : zero4;  // quat w,x,y,z
    const float4 r2 = in_range ? __ldg(density12 + 3 * (size_t)i + 2) : zero4;  // scale.xyz, pad
    const float3 pos = f3(r0.x, r0.y, r0.z);
    const float opacity = r0.w;
    const float fw = (float)fr.width, fh = (float)fr.height;

    // UT weights (threedgut.cuh:62-71, gutProjector.cuh:150,163,201)
    const float D = 3.0f;
    const float lambda = cfg.ut_alpha * cfg.ut_alpha * (D + cfg.ut_kappa) - D;
    const float w0_mean = lambda / (D + lambda);
    const float wi = 1.0f / (2.0f * (D + lambda));
    const float w0_cov = lambda / (D + lambda) + (1.0f - cfg.ut_alpha * cfg.ut_alpha + cfg.ut_beta);
    const float tol = cfg.ut_in_image_margin_factor;
    const int n_iter = cfg.n_rolling_shutter_iterations;

    bool valid_proj = false;
    float2 centre = make_float2(0.0f, 0.0f);
    float3 cov = f3(0.0f, 0.0f, 0.0f);
    float3 sensor_ray = f3(0.0f, 0.0f, 0.0f);
    if (in_range && opacity >= cfg.min_alpha && !(view_depth(fr.view, pos) < 0.2f)) {
        float3 c[3];
        quat_to_columns(r1.x, r1.y, r1.z, r1.w, c[0], c[1], c[2]);
        const float sc[3] = {r2.x, r2.y, r2.z};
        sensor_ray = pos - fr.cam_centre;
        float2 sp[7];
        const PoseMat m0{fr.start.c[0], fr.start.c[1], fr.start.c[2], fr.start.c[3]};   // formed on the host once per frame
        const bool ideal = pinhole_is_ideal(cam);
        int nvalid = 0;
        if (ROLLING) {   // point 0: the mean; 1..3: + delta along the particle's axes; 4..6: - delta
#pragma unroll 1
            for (int k = 0; k < 7; ++k) {
                const int a = k == 0 ? 0 : (k - 1) % 3;
                const float3 ca = a == 0 ? c[0] : (a == 1 ? c[1] : c[2]);
                const float sa = a == 0 ? sc[0] : (a == 1 ? sc[1] : sc[2]);
                const float3 d = (fr.ut_delta * sa) * ca;
                const float3 pk = k == 0 ? pos : (k < 4 ? pos + d : pos - d);
                float2 o;
                nvalid += project_with_shutter<true>(cam, fw, fh, n_iter, pk, tol, o, m0, ideal) ? 1 : 0;
                sp[k] = o;
            }
        } else {
            nvalid = project_with_shutter<false>(cam, fw, fh, n_iter, pos, tol, sp[0], m0, ideal) ? 1 : 0;
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const float3 d = (fr.ut_delta * sc[a]) * c[a];
                nvalid += project_with_shutter<false>(cam, fw, fh, n_iter, pos + d, tol, sp[a + 1], m0, ideal) ? 1 : 0;
                nvalid += project_with_shutter<false>(cam, fw, fh, n_iter, pos - d, tol, sp[a + 4], m0, ideal) ? 1 : 0;
            }
        }
        // the mean in the reference's order of summation (0, then +a, -a for a = 0, 1, 2)
        centre.x = sp[0].x * w0_mean; centre.y = sp[0].y * w0_mean;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            centre.x += wi * sp[a + 1].x; centre.y += wi * sp[a + 1].y;
            centre.x += wi * sp[a + 4].x; centre.y += wi * sp[a + 4].y;
        }
        const bool reject = cfg.ut_require_all_sigma_points_valid ? (nvalid < 7) : (nvalid == 0);
        if (!reject) {
            float cx = sp[0].x - centre.x, cy = sp[0].y - centre.y;
            cov = f3(w0_cov * (cx * cx), w0_cov * (cx * cy), w0_cov * (cy * cy));
#pragma unroll
            for (int k = 1; k < 7; ++k) {
                cx = sp[k].x - centre.x; cy = sp[k].y - centre.y;
                cov.x += wi * (cx * cx); cov.y += wi * (cx * cy); cov.z += wi * (cy * cy);
            }
            valid_proj = true;
        }
    }

    float2 extent = make_float2(0.0f, 0.0f);
    float4 co = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    float max_power = 0.0f;
    const bool valid_conic = valid_proj && conic_extent(cfg, cov, opacity, extent, co, max_power);
    // reference stores int 1 into float storage (gutProjector.cuh:275); 1.0f has the same truthiness and survives FTZ
    if (in_range) visibility[i] = valid_conic ? 1.0f : 0.0f;
#ifndef GUT_K1_NO_SH_PREFETCH
    // the SH row (192 B) is read after the tile walk, by the particles that kept a tile: start it towards L2 now
    if (valid_conic) {
        const char* row = reinterpret_cast<const char*>(sph + (size_t)i * ((cfg.sph_max_degree + 1) * (cfg.sph_max_degree + 1)) * 3);
        asm volatile("prefetch.global.L2 [%0];" ::"l"(row));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(row + 128));
    }
#endif

    // Tile walk (gutProjector.cuh:284-322), warp-cooperative: box sizes differ a lot between neighbouring particles, so the
    // boxes of the warp's 32 particles are flattened into one list of (particle, tile) items and every lane tests one item
    // per round, whatever particle it belongs to.  Each lane publishes its box in shared memory; an item finds its owner by
    // binary search over the warp's exclusive box-size prefix.
    __shared__ float4 s_box[256 / 32][32][3];     // {conic, max_power} {centre, rcx, rcy} {x0, y0, width, magic}
    __shared__ uint32_t s_res[256 / 32][32][3];   // tiles kept, mask of kept tiles (boxes of up to 64 tiles)
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    GutTileBox bb{0u, 0u, 0u, 0u};
    if (valid_conic) bb = gut_tile_bbox(fr.grid_x, fr.grid_y, centre, extent);
    const uint32_t bw = bb.x1 - bb.x0;
    const uint32_t total = bw * (bb.y1 - bb.y0);
    uint32_t incl = total;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (uint32_t)o) incl += v;
    }
    const uint32_t excl = incl - total;
    const uint32_t warp_total = __shfl_sync(0xffffffffu, incl, 31);
    if (total) {
        const float ts = (float)GUT_TILE;
        s_box[wid][lane][0] = make_float4(co.x, co.y, co.z, max_power);
        s_box[wid][lane][1] = make_float4(centre.x, centre.y, 1.0f / (ts * ts * co.x), 1.0f / (ts * ts * co.z));
        // row = umulhi(item, magic) is exact for item * width < 2^32 (width 1: the item is the row)
        s_box[wid][lane][2] = make_float4(__uint_as_float(bb.x0), __uint_as_float(bb.y0), __uint_as_float(bw),
                                          __uint_as_float(bw > 1u ? 0xffffffffu / bw + 1u : 0u));
    }
    s_res[wid][lane][0] = 0u; s_res[wid][lane][1] = 0u; s_res[wid][lane][2] = 0u;
    __syncwarp();
    for (uint32_t q0 = 0; q0 < warp_total; q0 += 32) {
        const uint32_t q = q0 + lane;
        uint32_t owner = 0;   // last lane whose prefix is <= q: lanes with empty boxes share their successor's prefix and lose to it
#pragma unroll
        for (uint32_t step = 16; step; step >>= 1) {
            const uint32_t v = __shfl_sync(0xffffffffu, excl, owner + step);
            if (v <= q) owner += step;
        }
        const uint32_t first = __shfl_sync(0xffffffffu, excl, owner);
        if (q < warp_total) {
            const float4 b0 = s_box[wid][owner][0], b1 = s_box[wid][owner][1], b2 = s_box[wid][owner][2];
            const uint32_t item = q - first, w_o = __float_as_uint(b2.z), magic = __float_as_uint(b2.w);
            const uint32_t row = magic ? __umulhi(item, magic) : item;
            const uint32_t x = __float_as_uint(b2.x) + (item - row * w_o), y = __float_as_uint(b2.y) + row;
            if (!cfg.tile_based_culling ||
                gut_tile_min_power_rc((float)x, (float)y, make_float4(b0.x, b0.y, b0.z, 0.0f), make_float2(b1.x, b1.y), b1.z, b1.w) < b0.w) {
                atomicAdd(&s_res[wid][owner][0], 1u);
                if (item < GUT_EMIT_MASK_TILES) atomicOr(&s_res[wid][owner][1 + (item >> 5)], 1u << (item & 31u));
                // RED: per-tile (sub-)list length for the bucket scan; the owner's particle index has the owner's lane in its low bits
                atomicAdd(&tile_hist[GUT_SUB_SLOT(y * fr.grid_x + x, owner & (GUT_TILE_SUB - 1), fr.num_tiles)], 1u);
            }
        }
    }
    __syncwarp();
    if (!in_range) return;
    const uint32_t ntiles = s_res[wid][lane][0];
    out.tiles_count[i] = ntiles;
    if (ntiles == 0) {
        out.proj_pos[i] = make_float2(0.0f, 0.0f);
        out.conic_opacity[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        out.extent[i] = make_float2(0.0f, 0.0f);
        out.depth[i] = 0.0f;
        out.features[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        out.emit_rec[i] = make_uint4(0u, 0u, 0u, 0u);
        return;
    }
    const uint32_t box_word = bb.x0 | (bb.y0 << 12) | ((total > GUT_EMIT_MASK_TILES ? 0u : bw) << 24) | (total > GUT_EMIT_MASK_TILES ? GUT_EMIT_BIG : 0u);
    const uint32_t mask_lo = s_res[wid][lane][1], mask_hi = s_res[wid][lane][2];

    // SH colour seen from the sensor centre, +0.5, unclamped (gutProjector.cuh:304-310, gaussianParticles.cuh:64-96)
    const float dist = sqrtf(dot(sensor_ray, sensor_ray));
    const float3 dir = f3(sensor_ray.x / dist, sensor_ray.y / dist, sensor_ray.z / dist);
    float b[16];
    gut_sh_basis(sph_degree, dir, b);
    const int ncoef_all = (cfg.sph_max_degree + 1) * (cfg.sph_max_degree + 1);
    const int nact = (sph_degree + 1) * (sph_degree + 1);
    const float* cf = sph + (size_t)i * ncoef_all * 3;
    float3 rad;
    if (ncoef_all == 16) {  // 192 B per particle: twelve 128-bit loads
        const float4* cf4 = reinterpret_cast<const float4*>(cf);
        float v[48];
#pragma unroll
        for (int q = 0; q < 12; ++q) {
            const float4 t = (q * 4 < nact * 3) ? __ldg(cf4 + q) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
        }
        rad = f3(b[0] * v[0], b[0] * v[1], b[0] * v[2]);
#pragma unroll
        for (int k = 1; k < 16; ++k)
            if (k < nact) { rad.x = rad.x + b[k] * v[3 * k]; rad.y = rad.y + b[k] * v[3 * k + 1]; rad.z = rad.z + b[k] * v[3 * k + 2]; }
    } else {
        rad = f3(b[0] * cf[0], b[0] * cf[1], b[0] * cf[2]);
        for (int k = 1; k < nact; ++k) { rad.x = rad.x + b[k] * cf[3 * k]; rad.y = rad.y + b[k] * cf[3 * k + 1]; rad.z = rad.z + b[k] * cf[3 * k + 2]; }
    }
    out.features[i] = make_float4(rad.x + 0.5f, rad.y + 0.5f, rad.z + 0.5f, 0.0f);
    out.proj_pos[i] = centre;
    out.conic_opacity[i] = co;
    out.extent[i] = extent;
    const float depth = cfg.global_z_order ? view_depth(fr.view, pos) : dist;
    out.depth[i] = depth;
    out.emit_rec[i] = make_uint4(mask_lo, mask_hi, box_word, __float_as_uint(depth));
}

// ---------------------------------------------------------------------------------------------------
// K6.  Per particle: (i) turn the ray-summed canonical-frame accumulators of K5 (layout: gut_render.cu, backward_walk)
// into the [N,12] parameter gradient -- the per-particle tail of processHitBwd (gaussianParticles.cuh:690-738:
// matmul_bw_vec, matmul_bw_quat, the 1/s factors), done once per particle instead of once per ray; (ii) dL/dSH =
// basis(dir) x (dL/dfeature masked by feature > 0)  (gutProjector.cuh:390-430, gaussianParticles.cuh:120-187).
// Particles the frame did not touch get zeros (accumulate == 0), so no memset pass over [N,12] / [N,48] is needed; the
// accumulators of touched particles are reset to zero here for the next backward.
// ---------------------------------------------------------------------------------------------------
constexpr int K6_THREADS = 128;
constexpr int K6_WARPS = K6_THREADS / 32;

// ---------------------------------------------------------------------------------------------------
// Colour-gradient slot of one view (what K6 / the fold pass hand to the step's SH expansion), in one of two layouts:
//   dense    [N][3] floats, zero rows for the particles the view did not touch;
//   compact  (gut_set_colour_layout(h, GUT_COLOUR_COMPACT)) only the rows the view's rays pushed on, 12 B each, in the order the
//            warps of the writing kernel claimed them, behind a table with one (mask, first row) pair per 32 particles:
//            word 0 = rows handed out (the writers' cursor, zeroed by the host before the pass), words 4 .. 4 + table_words the
//            table, rows behind it.  A view touches about a sixth of the particles at the bench workload: a reader on another
//            GPU pulls ~3 B per particle over NVLink instead of 12.
// ---------------------------------------------------------------------------------------------------
template <int WARPS>
GUT_D void colour_store(float* __restrict__ slot, uint32_t N, uint32_t i, const float3 g, const bool compact) {
    if (!compact) {
        if (i < N) { slot[3 * (size_t)i] = g.x; slot[3 * (size_t)i + 1] = g.y; slot[3 * (size_t)i + 2] = g.z; }
        return;
    }
    // (every thread of the CTA gets here exactly once: the callers guard their per-particle work with `i < N`, they do not
    // return early.)  One claim per CTA: the cursor is one address, and same-address atomics retire about one per clock.
    __shared__ uint32_t s_rows[WARPS + 1];
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const bool moved = i < N && (g.x != 0.0f || g.y != 0.0f || g.z != 0.0f);
    const uint32_t m = __ballot_sync(0xffffffffu, moved);
    if (lane == 0) s_rows[wid] = (uint32_t)__popc(m);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t total = 0;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) { const uint32_t c = s_rows[w]; s_rows[w] = total; total += c; }
        s_rows[WARPS] = total ? atomicAdd(reinterpret_cast<uint32_t*>(slot), total) : 0u;
    }
    __syncthreads();
    const uint32_t base = s_rows[WARPS] + s_rows[wid];
    if (lane == 0 && i < N) reinterpret_cast<uint2*>(slot + 4)[i >> 5] = make_uint2(m, base);
    if (moved) {
        float* row = slot + 4 + gut_colour_table_words(N) + 3 * (size_t)(base + __popc(m & ((1u << lane) - 1u)));
        row[0] = g.x; row[1] = g.y; row[2] = g.z;
    }
}
// Reading a slot, in two steps so that a caller can put the table loads of several views in flight before the first row load
// (the second round trip depends on the first; across NVLink each is a few microseconds).
GUT_D uint2 colour_table_load(const float* __restrict__ slot, uint32_t i, const bool compact) {
    return compact ? __ldg(reinterpret_cast<const uint2*>(slot + 4) + (i >> 5)) : make_uint2(0u, 0u);   // one 8-byte request per warp
}
GUT_D float3 colour_load(const float* __restrict__ slot, uint32_t N, uint32_t i, const uint2 t, const bool compact) {
    if (!compact) {
        const float* gp = slot + 3 * (size_t)i;
        return f3(__ldg(gp), __ldg(gp + 1), __ldg(gp + 2));
    }
    const uint32_t lane = i & 31u;
    if (!((t.x >> lane) & 1u)) return f3(0.0f, 0.0f, 0.0f);
    const float* gp = slot + 4 + gut_colour_table_words(N) + 3 * (size_t)(t.y + __popc(t.x & ((1u << lane) - 1u)));
    return f3(__ldg(gp), __ldg(gp + 1), __ldg(gp + 2));
}

// The "origin fold": K5 leaves out the go (x) A halves of the torque, the trace and the scale terms when go is the same
// for every ray of the view (gut_render.cu, backward_walk).  With A summed over the view's rays they are
//   tau' += A x (s^2 go) = A x (s * R^T v),   tr += sum_i A_i v_i / s_i,   Z_i += go_i A_i,   W += go . A,      v = view centre - position,
// go = diag(1/s) R^T v.  c0..c2 are the columns of R, is = 1 / s.
GUT_D void origin_fold(const float3 A, const float3 v, const float3 c0, const float3 c1, const float3 c2, const float3 sc,
                       const float3 is, float3& tq, float& tr, float3& Z, float& W) {
    const float3 rv = f3(dot(c0, v), dot(c1, v), dot(c2, v));
    const float3 go = rv * is;
    const float3 G = sc * rv;
    tq = tq + cross(A, G);
    tr = tr + (A.x * (v.x * is.x) + A.y * (v.y * is.y) + A.z * (v.z * is.z));
    Z = Z + go * A;
    W = W + dot(go, A);
}

// Phase A: one thread per particle computes the [N,12] gradient row and the masked colour gradient + SH basis.
// Phase B: each warp writes its 32 rows of [N,12] (3 float4 per row) and [N,48] (12 float4 per row) as contiguous,
// fully coalesced 128-bit accesses (12 independent loads in flight per lane when accumulating) instead of every
// thread walking its own 192-byte row.
// DENSITY: convert the accumulators into the [N,12] gradient (and reset them); COLOUR: the SH / colour part (and reset
// its three accumulators).  Both: one backward call of the reference API.  COLOUR only: one view of a deferred step
// (gut_backward_colour).  DENSITY only: end of a deferred step, over ALL particles (tiles_count == nullptr): the
// accumulators then hold the sum over the step's views, which is legitimate because they are view-independent.
template <bool DENSITY, bool COLOUR>
__global__ void __launch_bounds__(K6_THREADS) gut_project_backward_kernel(GutConfig cfg, GutFrame fr, uint32_t N, int sph_degree,
                                                                          const float4* __restrict__ density12,
                                                                          const uint32_t* __restrict__ tiles_count,
                                                                          const float4* __restrict__ features,
                                                                          float4* __restrict__ raw_grad, float4* __restrict__ density_grad,
                                                                          float* __restrict__ sph_grad, float* __restrict__ colour_out,
                                                                          float4* __restrict__ a_sum, int accumulate,
                                                                          int colour_compact /* layout of colour_out, colour_store above */) {
    __shared__ float4 s_dg[K6_WARPS][32 * 3];      // parameter-gradient rows of the warp
    __shared__ float s_basis[K6_WARPS][32][17];    // SH basis per row (+1 pad: conflict-free column reads)
    __shared__ float4 s_g[K6_WARPS][32];           // masked colour gradient per row, w = 1 if the row was touched
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t warp_base = i - lane;           // first particle of this warp
    const int ncoef_all = (cfg.sph_max_degree + 1) * (cfg.sph_max_degree + 1);
    const int nact = (sph_degree + 1) * (sph_degree + 1);

    bool touched = false;
    float4 o0 = make_float4(0.f, 0.f, 0.f, 0.f), o1 = o0, o2 = o0;
    float3 g = f3(0.f, 0.f, 0.f);
    float b[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) b[k] = 0.0f;
    // All loads are issued before tiles_count is known (an untouched particle has zero accumulators, so computing on it is
    // harmless): the kernel is bound by the latency of its load -> store chains, and a dependent first load doubles them.
    if (i < N) {
        touched = tiles_count ? tiles_count[i] != 0 : true;
        const float4 r0 = __ldg(density12 + 3 * (size_t)i);      // pos, density
        float4* rg = raw_grad + 4 * (size_t)i;
        float4 g0, g1, g2 = rg[2], g3 = rg[3];
        const float4 z = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (DENSITY) { g0 = rg[0]; g1 = rg[1]; }
        if (touched) {
            if (DENSITY) { rg[0] = z; rg[1] = z; }
            if (DENSITY && COLOUR) { rg[2] = z; rg[3] = z; }
            else if (DENSITY) { rg[2] = make_float4(0.0f, 0.0f, 0.0f, g2.w); rg[3] = make_float4(g3.x, g3.y, 0.0f, 0.0f); }
            else { rg[2] = make_float4(g2.x, g2.y, g2.z, 0.0f); rg[3] = make_float4(0.0f, 0.0f, g3.z, g3.w); }
        }
        if (DENSITY) {
        const float4 qv = __ldg(density12 + 3 * (size_t)i + 1);  // quat w,x,y,z
        const float4 r2 = __ldg(density12 + 3 * (size_t)i + 2);  // scale
        // ---- (i) parameter gradient ----
        float3 A = f3(g0.x, g0.y, g0.z);
        float3 tq = f3(g1.x, g1.y, g1.z);
        float tr_dir = g1.w;
        float3 Z = f3(g2.x, g2.y, g2.z);
        float W = g3.z;   // sum_i P_ii (the trace of dL/dR^T R^T)
        const float3 is = f3(1.0f / r2.x, 1.0f / r2.y, 1.0f / r2.z);
        float3 c0, c1, c2;
        quat_to_columns(qv.x, qv.y, qv.z, qv.w, c0, c1, c2);
        if (a_sum) {   // end of a deferred step: every view folded its own origin (gut_view_fold_kernel) and moved its A here
            const float4 as = a_sum[i];
            A = f3(as.x, as.y, as.z);
            if (as.x != 0.0f || as.y != 0.0f || as.z != 0.0f) a_sum[i] = z;
        } else {
            origin_fold(A, fr.cam_centre - f3(r0.x, r0.y, r0.z), c0, c1, c2, f3(r2.x, r2.y, r2.z), is, tq, tr_dir, Z, W);
        }
        const float3 tau = f3(tq.x * (is.y * is.z), tq.y * (is.x * is.z), tq.z * (is.x * is.y));
        const float3 As = A * is;
        // dL/dpos = -R (A / s)
        o0 = make_float4(-(c0.x * As.x + c1.x * As.y + c2.x * As.z), -(c0.y * As.x + c1.y * As.y + c2.y * As.z),
                         -(c0.z * As.x + c1.z * As.y + c2.z * As.z), g0.w);
        o2 = make_float4(-Z.x * is.x, -Z.y * is.y, -Z.z * is.z, 0.0f);
        // radial part: q . dL/dq = 2 <dL/dR^T, R^T - I> ; <dL/dR^T, R^T> = sum_i P_ii = go . A (W: origin fold, the direction
        // half u . B vanishes) ; <dL/dR^T, I> = tr (direction half summed per ray by K5, origin half from the fold)
        const float rho = 2.0f * (W - tr_dir);
        // tangential part: body-frame torque -> 2 q (x) (0, tau)
        const float3 rv = f3(qv.y, qv.z, qv.w);
        const float3 rxt = cross(rv, tau);
        o1 = make_float4(rho * qv.x - 2.0f * dot(rv, tau), rho * rv.x + 2.0f * (qv.x * tau.x + rxt.x),
                         rho * rv.y + 2.0f * (qv.x * tau.y + rxt.y), rho * rv.z + 2.0f * (qv.x * tau.z + rxt.z));
        }
        if (COLOUR) {
        // ---- (ii) SH: masked colour gradient and basis of the view direction ----
        const float4 ft = __ldg(features + i);
        const float3 v = f3(r0.x, r0.y, r0.z) - fr.cam_centre;
        const float len = sqrtf(dot(v, v));
        const float3 dir = len <= 0.0f ? f3(1.0f, 0.0f, 0.0f) : f3(v.x / len, v.y / len, v.z / len);
        gut_sh_basis(sph_degree, dir, b);
        g = f3(g2.w * (ft.x > 0.0f ? 1.0f : 0.0f), g3.x * (ft.y > 0.0f ? 1.0f : 0.0f), g3.y * (ft.z > 0.0f ? 1.0f : 0.0f));
        }
    }
    s_dg[wid][3 * lane] = o0; s_dg[wid][3 * lane + 1] = o1; s_dg[wid][3 * lane + 2] = o2;
#pragma unroll
    for (int k = 0; k < 16; ++k) s_basis[wid][lane][k] = (k < nact) ? b[k] : 0.0f;
    s_g[wid][lane] = make_float4(g.x, g.y, g.z, touched ? 1.0f : 0.0f);
    __syncwarp();

    const uint32_t rows = warp_base < N ? min(32u, N - warp_base) : 0u;
    // [N,12]: 3 float4 per row, rows*3 contiguous float4
    if (DENSITY) {
        float4* dst = density_grad + 3 * (size_t)warp_base;
#pragma unroll
        for (int it = 0; it < 3; ++it) {
            const uint32_t f = it * 32 + lane;
            const uint32_t row = f / 3;
            if (row < rows) {
                const bool t_row = s_g[wid][row].w != 0.0f;
                float4 val = s_dg[wid][f];
                if (accumulate) {
                    if (t_row) { const float4 p = dst[f]; val.x += p.x; val.y += p.y; val.z += p.z; val.w += p.w; dst[f] = val; }
                } else {
                    dst[f] = val;
                }
            }
        }
    }
    if (!COLOUR) return;
    // deferred SH gradient: hand out the masked colour gradient of this view, gut_sh_expand_kernel sums the views later
    if (colour_out) {
        colour_store<K6_WARPS>(colour_out, N, i, g, colour_compact != 0);
        return;
    }
    // [N,48]
    if (ncoef_all == 16) {
        float4* dst = reinterpret_cast<float4*>(sph_grad) + 12 * (size_t)warp_base;
        float4 prev[12];
#pragma unroll
        for (int it = 0; it < 12; ++it) {   // all loads first: 12 independent coalesced requests per lane
            const uint32_t f = it * 32 + lane;
            const uint32_t row = f / 12;
            prev[it] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (accumulate && row < rows && s_g[wid][row].w != 0.0f) prev[it] = dst[f];
        }
#pragma unroll
        for (int it = 0; it < 12; ++it) {
            const uint32_t f = it * 32 + lane;
            const uint32_t row = f / 12, e0 = (f % 12) * 4;   // elements e0..e0+3 of the row: coefficient e/3, channel e%3
            if (row < rows) {
                const float4 gg = s_g[wid][row];
                if (!accumulate || gg.w != 0.0f) {
                    float val[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint32_t e = e0 + q, kk = e / 3, ch = e - 3 * kk;
                        const float bk = s_basis[wid][row][kk];
                        val[q] = bk * (ch == 0 ? gg.x : (ch == 1 ? gg.y : gg.z));
                    }
                    dst[f] = make_float4(val[0] + prev[it].x, val[1] + prev[it].y, val[2] + prev[it].z, val[3] + prev[it].w);
                }
            }
        }
    } else if (i < N) {   // generic coefficient count: per-thread rows
        float* out = sph_grad + (size_t)i * ncoef_all * 3;
        if (touched || !accumulate)
            for (int k = 0; k < ncoef_all; ++k) {
                const float bk = (k < nact) ? b[k] : 0.0f;
                for (int ch = 0; ch < 3; ++ch) {
                    const float gv = bk * (ch == 0 ? g.x : (ch == 1 ? g.y : g.z));
                    out[3 * k + ch] = accumulate ? out[3 * k + ch] + gv : gv;
                }
            }
    }
}

__global__ void gut_copy_features3_kernel(uint32_t N, const float4* __restrict__ f4, float* __restrict__ f3out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const float4 v = f4[i];
    f3out[3 * (size_t)i] = v.x; f3out[3 * (size_t)i + 1] = v.y; f3out[3 * (size_t)i + 2] = v.z;
}

}  // namespace

cudaError_t gut_launch_project(const GutConfig& cfg, const GutCamera& cam, const GutFrame& frame, uint32_t N, int sph_degree,
                               const float* density12, const float* sph, GutProjBuffers out, float* visibility,
                               uint32_t* tile_hist, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    if (cam.shutter == GUT_SHUTTER_GLOBAL)
        gut_project_kernel<false><<<(N + 255) / 256, 256, 0, stream>>>(cfg, cam, frame, N, sph_degree,
                                                                        reinterpret_cast<const float4*>(density12), sph, out, visibility, tile_hist);
    else
        gut_project_kernel<true><<<(N + 255) / 256, 256, 0, stream>>>(cfg, cam, frame, N, sph_degree,
                                                                       reinterpret_cast<const float4*>(density12), sph, out, visibility, tile_hist);
    return cudaGetLastError();
}

// One view of a deferred step (gut_backward_colour): hand out the view's masked colour gradient and reset its three
// accumulators; fold the view's origin into the torque / trace / scale accumulators (origin_fold above: it needs the
// A of THIS view alone and this view's centre) and move A to the step's sum `a_sum`, leaving [0..2] zero for the next
// view.  Everything else waits for the end of the step.
template <bool COMPACT>
__global__ void __launch_bounds__(256) gut_view_fold_kernel(uint32_t N, float3 centre, const float4* __restrict__ density12,
                                                            const float4* __restrict__ features, float4* __restrict__ raw_grad,
                                                            float4* __restrict__ a_sum, float* __restrict__ colour_out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float3 g = f3(0.0f, 0.0f, 0.0f);
    // (no early return: in the compact layout the CTA's row claim below is a barrier and a ballot every thread takes part in,
    // from this one call site)
    if (i < N) {
    // no test of tiles_count: the accumulators of an untouched particle are zero, and skipping the dependent load matters
    // more than the bytes (the kernel is bound by the latency of its load -> store chain, not by bandwidth)
    float4* rg = raw_grad + 4 * (size_t)i;
    const float4 a = rg[0], g2 = rg[2], g3 = rg[3];
    const float4 ft = __ldg(features + i);
    g = f3(g2.w * (ft.x > 0.0f ? 1.0f : 0.0f), g3.x * (ft.y > 0.0f ? 1.0f : 0.0f), g3.y * (ft.z > 0.0f ? 1.0f : 0.0f));
    float3 Z = f3(g2.x, g2.y, g2.z);
    float W = g3.z;
    const bool moved = a.x != 0.0f || a.y != 0.0f || a.z != 0.0f;
    if (moved) {   // second round trip for the particles the view's rays pushed on
        const float4 r0 = __ldg(density12 + 3 * (size_t)i), qv = __ldg(density12 + 3 * (size_t)i + 1), r2 = __ldg(density12 + 3 * (size_t)i + 2);
        const float4 g1 = rg[1], as = a_sum[i];
        float3 c0, c1, c2;
        quat_to_columns(qv.x, qv.y, qv.z, qv.w, c0, c1, c2);
        float3 tq = f3(g1.x, g1.y, g1.z);
        float tr = g1.w;
        origin_fold(f3(a.x, a.y, a.z), centre - f3(r0.x, r0.y, r0.z), c0, c1, c2, f3(r2.x, r2.y, r2.z),
                    f3(1.0f / r2.x, 1.0f / r2.y, 1.0f / r2.z), tq, tr, Z, W);
        rg[1] = make_float4(tq.x, tq.y, tq.z, tr);
        a_sum[i] = make_float4(as.x + a.x, as.y + a.y, as.z + a.z, 0.0f);
        rg[0] = make_float4(0.0f, 0.0f, 0.0f, a.w);
    }
    if (moved || g2.w != 0.0f) rg[2] = make_float4(Z.x, Z.y, Z.z, 0.0f);
    if (moved || g3.x != 0.0f || g3.y != 0.0f) rg[3] = make_float4(0.0f, 0.0f, W, g3.w);
    }
    colour_store<8>(colour_out, N, i, g, COMPACT);
}

// Row cursor of a compact slot back to zero.  A kernel, not cudaMemsetAsync: the slot may live in memory that peers have mapped
// (gut_peer_buffer_create), and a memset there does not stay confined to its stream -- it cost 2 ms per view with two handles
// on two streams.
__global__ void gut_colour_cursor_reset_kernel(uint4* __restrict__ slot) { *slot = make_uint4(0u, 0u, 0u, 0u); }

// a_sum [N,4]: the deferred step's sum of A over its views.  Required with colour_out (one view of a deferred step) and
// with density_grad alone (end of the step); NULL for the one-view backward, which folds its own origin.
cudaError_t gut_launch_project_backward(const GutConfig& cfg, const GutFrame& frame, uint32_t N, int sph_degree,
                                        const float* density12, const uint32_t* tiles_count, const float4* features,
                                        float* raw_grad, float* density_grad, float* sph_grad, float* colour_out, int colour_compact,
                                        float* a_sum, int accumulate, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    if (colour_out && colour_compact) gut_colour_cursor_reset_kernel<<<1, 1, 0, stream>>>(reinterpret_cast<uint4*>(colour_out));
    const dim3 grid((N + K6_THREADS - 1) / K6_THREADS);
    const float4* d12 = reinterpret_cast<const float4*>(density12);
    float4* rg = reinterpret_cast<float4*>(raw_grad);
    float4* dg = reinterpret_cast<float4*>(density_grad);
    float4* as = reinterpret_cast<float4*>(a_sum);
    if (density_grad && (sph_grad || colour_out))
        gut_project_backward_kernel<true, true><<<grid, K6_THREADS, 0, stream>>>(cfg, frame, N, sph_degree, d12, tiles_count, features, rg, dg,
                                                                               sph_grad, colour_out, nullptr, accumulate, colour_compact);
    else if (density_grad) {
        if (!a_sum) return cudaErrorInvalidValue;
        gut_project_backward_kernel<true, false><<<grid, K6_THREADS, 0, stream>>>(cfg, frame, N, sph_degree, d12, tiles_count, features, rg, dg,
                                                                                nullptr, nullptr, as, accumulate, 0);
    } else if (colour_out) {
        if (!a_sum) return cudaErrorInvalidValue;
        if (colour_compact)
            gut_view_fold_kernel<true><<<(N + 255) / 256, 256, 0, stream>>>(N, frame.cam_centre, d12, features, rg, as, colour_out);
        else
            gut_view_fold_kernel<false><<<(N + 255) / 256, 256, 0, stream>>>(N, frame.cam_centre, d12, features, rg, as, colour_out);
    } else
        gut_project_backward_kernel<false, true><<<grid, K6_THREADS, 0, stream>>>(cfg, frame, N, sph_degree, d12, tiles_count, features, rg,
                                                                                nullptr, sph_grad, colour_out, nullptr, accumulate, colour_compact);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------
// Deferred SH backward of a multi-view step: dL/dSH[i] (+)= sum_v basis(normalize(pos_i - centre_v)) (x) colour_grad_v[i].
// K6 hands out 12 bytes per particle and view instead of read-modify-writing the 192-byte SH gradient row once per
// view; this kernel expands all views at once (and, across GPUs, the views of all ranks after an all-gather of the
// colour gradients -- 12 bytes per particle and view on the wire instead of 192 per particle, 3dgrut_b200/dist.py).
// Phase A: one thread per particle accumulates its 48 values over the views; phase B: the warp writes its 32 rows as
// contiguous 128-bit accesses.
// ---------------------------------------------------------------------------------------------------
// The colour gradient of view v is read through views.colour[v]: a slice of a local buffer, or -- across GPUs -- a buffer
// of a PEER GPU mapped into this process (gut_peer_buffer_open), so that the expansion reads the other ranks' gradients
// straight over NVLink while it computes and no all-gather pass exists (3dgrut_b200/dist.py, training_step_peer).  All
// loads of a chunk of views are issued before the first is used: one NVLink round trip per chunk, not per view.
struct ExpandViews {
    const float* colour[GUT_EXPAND_MAX_VIEWS];   // one slot each: [N][3], or the compact layout (colour_store / colour_load above)
    int wide;                                    // dense slots, all 16-byte aligned: a warp may fetch its 32 rows as 24 x 16 bytes
};
GUT_D void copy16_async(void* sdst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(sdst)), "l"(gsrc) : "memory");
}
GUT_D void copy_async_wait() { asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory"); }
constexpr int EXPAND_CHUNK = 8;
// The colour gradients of views v0 .. v0 + EXPAND_CHUNK of particle i, all loads in flight together.  The layout is a template
// parameter of the two expansion kernels: the dense instantiation is the code it was before the compact layout existed (a
// run-time flag cost it 0.14 ms per step at 3M particles and 8 views).
template <bool COMPACT>
GUT_D void load_colour_chunk(const ExpandViews& views, uint32_t v0, uint32_t n_views, uint32_t N, uint32_t i, float3 (&gq)[EXPAND_CHUNK]) {
    if (!COMPACT) {
#pragma unroll
        for (int q = 0; q < EXPAND_CHUNK; ++q) {
            gq[q] = f3(0.f, 0.f, 0.f);
            if (v0 + q < n_views) {
                const float* gp = views.colour[v0 + q] + 3 * (size_t)i;
                gq[q] = f3(__ldg(gp), __ldg(gp + 1), __ldg(gp + 2));
            }
        }
    } else {
        uint2 tq[EXPAND_CHUNK];
#pragma unroll
        for (int q = 0; q < EXPAND_CHUNK; ++q)
            tq[q] = (v0 + q < n_views) ? colour_table_load(views.colour[v0 + q], i, true) : make_uint2(0u, 0u);
#pragma unroll
        for (int q = 0; q < EXPAND_CHUNK; ++q) {
            gq[q] = f3(0.f, 0.f, 0.f);
            if (v0 + q < n_views) gq[q] = colour_load(views.colour[v0 + q], N, i, tq[q], true);
        }
    }
}
// -DGUT_EXPAND_SMEM_ACC=1 keeps the 48 sums in the lane's row of the shared tile instead of registers (117 -> 64 registers, 4 -> 8
// CTAs per SM; same bits).  Measured on one GPU (3M particles, 8 local views): 0.35 -> 0.40 ms, so it is off; whether twice the
// loads in flight shorten the NVLink-latency-bound expansion at 8 GPUs is not measured yet.
#ifndef GUT_EXPAND_SMEM_ACC
#define GUT_EXPAND_SMEM_ACC 0
#endif
template <bool COMPACT>
__global__ void __launch_bounds__(K6_THREADS, GUT_EXPAND_SMEM_ACC ? 8 : 1) gut_sh_expand_kernel(uint32_t N, int sph_degree, const float4* __restrict__ density12,
                                                                   uint32_t n_views, const float* __restrict__ centres /* [n_views][3] */,
                                                                   ExpandViews views,
                                                                   float4* __restrict__ sph_grad /* [N][12] float4 */, int accumulate,
                                                                   int ncoef_all /* (sph_max_degree + 1)^2: the row holds 3 ncoef_all floats */) {
    __shared__ __align__(16) float s_out[K6_WARPS][32][49];
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t warp_base = i - lane;
#if GUT_EXPAND_SMEM_ACC
    float* const acc = &s_out[wid][lane][0];
    const bool wide = false;
#else
    float acc[48];
    // A full warp's 32 dense rows of one view are 384 contiguous bytes: 24 lanes copy 16 bytes each straight into the warp's part
    // of the (not yet used) output tile -- 8 asynchronous wide requests per chunk of views instead of 24 four-byte loads per lane,
    // which is what a peer's memory across NVLink wants -- and every lane then picks its row out of shared memory.
    const bool wide = !COMPACT && views.wide != 0 && warp_base + 32u <= N;   // warp-uniform
#endif
#pragma unroll
    for (int k = 0; k < 48; ++k) acc[k] = 0.0f;
    float4 r0 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < N) r0 = __ldg(density12 + 3 * (size_t)i);
#pragma unroll
    for (uint32_t v0 = 0; v0 < GUT_EXPAND_MAX_VIEWS; v0 += EXPAND_CHUNK) {   // static indices into the pointer table
        if (v0 >= n_views) break;
        float3 gq[EXPAND_CHUNK];
        if (wide) {
            float* const st = &s_out[wid][0][0];   // 96 floats per view of the chunk
            __syncwarp();                          // the previous chunk has been read
#pragma unroll
            for (int q = 0; q < EXPAND_CHUNK; ++q)
                if (v0 + q < n_views && lane < 24u) copy16_async(st + 96 * q + 4 * lane, views.colour[v0 + q] + 3 * (size_t)warp_base + 4 * lane);
            copy_async_wait();
            __syncwarp();
#pragma unroll
            for (int q = 0; q < EXPAND_CHUNK; ++q)
                gq[q] = (v0 + q < n_views) ? f3(st[96 * q + 3 * lane], st[96 * q + 3 * lane + 1], st[96 * q + 3 * lane + 2]) : f3(0.f, 0.f, 0.f);
        } else if (i < N) {
            load_colour_chunk<COMPACT>(views, v0, n_views, N, i, gq);
        }
        if (i < N) {
#pragma unroll
            for (int q = 0; q < EXPAND_CHUNK; ++q) {
                const float3 g = gq[q];
                if (g.x == 0.0f && g.y == 0.0f && g.z == 0.0f) continue;   // the view did not touch the particle
                const uint32_t v = v0 + q;
                const float3 d = f3(r0.x, r0.y, r0.z) - f3(__ldg(centres + 3 * v), __ldg(centres + 3 * v + 1), __ldg(centres + 3 * v + 2));
                const float len = sqrtf(dot(d, d));
                const float3 dir = len <= 0.0f ? f3(1.0f, 0.0f, 0.0f) : f3(d.x / len, d.y / len, d.z / len);
                float b[16];
#pragma unroll
                for (int k = 0; k < 16; ++k) b[k] = 0.0f;
                gut_sh_basis(sph_degree, dir, b);
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    acc[3 * k] += b[k] * g.x; acc[3 * k + 1] += b[k] * g.y; acc[3 * k + 2] += b[k] * g.z;
                }
            }
        }
    }
    if (ncoef_all != 16) {   // sph_max_degree 0..2: per-thread rows of 3, 12 or 27 floats (coefficients beyond the active degree are zero)
        if (i < N) {
            float* row = reinterpret_cast<float*>(sph_grad) + (size_t)i * ncoef_all * 3;
#pragma unroll
            for (int k = 0; k < 27; ++k)
                if (k < 3 * ncoef_all) row[k] = accumulate ? row[k] + acc[k] : acc[k];
        }
        return;
    }
#if !GUT_EXPAND_SMEM_ACC
    __syncwarp();   // (wide: the staged rows have been read by every lane before the tile takes the sums)
#pragma unroll
    for (int k = 0; k < 48; ++k) s_out[wid][lane][k] = acc[k];
#endif
    __syncwarp();
    const uint32_t rows = warp_base < N ? min(32u, N - warp_base) : 0u;
    float4* dst = sph_grad + 12 * (size_t)warp_base;
#pragma unroll
    for (int it = 0; it < 12; ++it) {
        const uint32_t f = it * 32 + lane;
        const uint32_t row = f / 12, e0 = (f % 12) * 4;
        if (row < rows) {
            float4 val = make_float4(s_out[wid][row][e0], s_out[wid][row][e0 + 1], s_out[wid][row][e0 + 2], s_out[wid][row][e0 + 3]);
            if (accumulate) { const float4 p = dst[f]; val.x += p.x; val.y += p.y; val.z += p.z; val.w += p.w; }
            dst[f] = val;
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// SH expansion with SelectiveAdam as its epilogue (gut_sh_expand_adam).  Phase A is gut_sh_expand_kernel's
// (same arithmetic in the same order: the 48 values are the bits the expansion would have written), phase B applies them:
// the warp walks its 32 rows of features_albedo (96 contiguous floats) and features_specular (1440 contiguous floats) one
// float per lane and round, parameter / exp_avg / exp_avg_sq read and written coalesced, rows of invisible particles left alone.
// ---------------------------------------------------------------------------------------------------
#ifndef GUT_SHADAM_MIN_BLOCKS
#define GUT_SHADAM_MIN_BLOCKS 8
#endif
template <bool COMPACT>
__global__ void __launch_bounds__(K6_THREADS, GUT_SHADAM_MIN_BLOCKS) gut_sh_expand_adam_kernel(uint32_t N, int sph_degree, const float4* __restrict__ density12,
                                                                        uint32_t n_views, const float* __restrict__ centres, ExpandViews views,
                                                                        GutAdamTensors albedo, GutAdamTensors specular,
                                                                        const float* __restrict__ visibility, float b1, float b2, float eps,
                                                                        uint32_t vec_mask /* bit 0 / 1: albedo / specular tensors 16-byte aligned */) {
    __shared__ float s_out[K6_WARPS][32][49];
    __shared__ float s_vis[K6_WARPS][32];
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t warp_base = i - lane;
    const bool vis = i < N && (visibility ? __ldg(visibility + i) != 0.0f : true);
    if (!__any_sync(0xffffffffu, vis)) return;   // warp-uniform: none of the 32 rows is updated
    // the 48 sums live in the lane's row of the shared tile, not in registers: the Adam phase below is what the kernel spends
    // its time in (IEEE sqrt and division per element), and it runs at twice the occupancy without 48 accumulators in the register
    // budget: 116 -> 64 registers, 4 -> 8 CTAs per SM, 1.34 -> 1.05 ms at 3M particles and 8 views (the unfused chain: 1.27 ms)
    float* const acc = &s_out[wid][lane][0];
#pragma unroll
    for (int k = 0; k < 48; ++k) acc[k] = 0.0f;
    if (vis) {
        const float4 r0 = __ldg(density12 + 3 * (size_t)i);
#pragma unroll
        for (uint32_t v0 = 0; v0 < GUT_EXPAND_MAX_VIEWS; v0 += EXPAND_CHUNK) {
            if (v0 >= n_views) break;
            float3 gq[EXPAND_CHUNK];
            load_colour_chunk<COMPACT>(views, v0, n_views, N, i, gq);
#pragma unroll
            for (int q = 0; q < EXPAND_CHUNK; ++q) {
                const float3 g = gq[q];
                if (g.x == 0.0f && g.y == 0.0f && g.z == 0.0f) continue;
                const uint32_t v = v0 + q;
                const float3 d = f3(r0.x, r0.y, r0.z) - f3(__ldg(centres + 3 * v), __ldg(centres + 3 * v + 1), __ldg(centres + 3 * v + 2));
                const float len = sqrtf(dot(d, d));
                const float3 dir = len <= 0.0f ? f3(1.0f, 0.0f, 0.0f) : f3(d.x / len, d.y / len, d.z / len);
                float b[16];
#pragma unroll
                for (int k = 0; k < 16; ++k) b[k] = 0.0f;
                gut_sh_basis(sph_degree, dir, b);
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    acc[3 * k] += b[k] * g.x; acc[3 * k + 1] += b[k] * g.y; acc[3 * k + 2] += b[k] * g.z;
                }
            }
        }
    }
    s_vis[wid][lane] = vis ? 1.0f : 0.0f;
    __syncwarp();
    const uint32_t rows = warp_base < N ? min(32u, N - warp_base) : 0u;
    // The warp's rows are ONE contiguous span of each tensor (32 x 3 and 32 x 45 floats, starting 16-byte aligned when the
    // tensors are): walked as 128-bit vectors, four consecutive floats per lane and round -- a vector may straddle two rows,
    // so row, column and visibility are looked up per component.
    auto update_span = [&](const GutAdamTensors& t, auto width_c, const uint32_t col0, const bool vec_ok) {
        constexpr uint32_t width = decltype(width_c)::value;
        const uint32_t total = rows * width;
        float* p = t.param + (size_t)width * warp_base; float* m = t.exp_avg + (size_t)width * warp_base;
        float* v = t.exp_avg_sq + (size_t)width * warp_base;
        constexpr uint32_t rounds = (32 * width / 4 + 31) / 32;
#pragma unroll
        for (uint32_t it = 0; it < rounds; ++it) {   // unrolled: all of a lane's loads are in flight before the first update
            const uint32_t e0 = 4 * (it * 32 + lane);
            if (e0 >= total) continue;
            if (vec_ok && e0 + 3 < total) {
                float gk[4]; bool on[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const uint32_t e = e0 + c, row = e / width, col = e - row * width;
                    on[c] = s_vis[wid][row] != 0.0f;
                    gk[c] = s_out[wid][row][col0 + col];
                }
                if (on[0] || on[1] || on[2] || on[3]) {
                    float4 pp = *reinterpret_cast<float4*>(p + e0), mm = *reinterpret_cast<float4*>(m + e0), vv = *reinterpret_cast<float4*>(v + e0);
                    if (on[0]) adam_scalar(pp.x, gk[0], mm.x, vv.x, t.lr, b1, b2, eps);
                    if (on[1]) adam_scalar(pp.y, gk[1], mm.y, vv.y, t.lr, b1, b2, eps);
                    if (on[2]) adam_scalar(pp.z, gk[2], mm.z, vv.z, t.lr, b1, b2, eps);
                    if (on[3]) adam_scalar(pp.w, gk[3], mm.w, vv.w, t.lr, b1, b2, eps);
                    *reinterpret_cast<float4*>(p + e0) = pp; *reinterpret_cast<float4*>(m + e0) = mm; *reinterpret_cast<float4*>(v + e0) = vv;
                }
            } else {
                for (uint32_t e = e0; e < min(e0 + 4u, total); ++e) {
                    const uint32_t row = e / width, col = e - row * width;
                    if (s_vis[wid][row] != 0.0f) {
                        float pp = p[e], mm = m[e], vv = v[e];
                        adam_scalar(pp, s_out[wid][row][col0 + col], mm, vv, t.lr, b1, b2, eps);
                        p[e] = pp; m[e] = mm; v[e] = vv;
                    }
                }
            }
        }
    };
    update_span(albedo, std::integral_constant<uint32_t, 3u>{}, 0u, (vec_mask & 1u) != 0u);
    update_span(specular, std::integral_constant<uint32_t, 45u>{}, 3u, (vec_mask & 2u) != 0u);
}

cudaError_t gut_launch_sh_expand_adam(uint32_t N, int sph_degree, const float* density12, uint32_t n_views, const float* centres,
                                      const float* const* colour_ptrs, int colour_compact, const GutAdamTensors& albedo, const GutAdamTensors& specular,
                                      const float* visibility, float b1, float b2, float eps, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    if (n_views > GUT_EXPAND_MAX_VIEWS) return cudaErrorInvalidValue;
    ExpandViews views = {};
    for (uint32_t v = 0; v < n_views; ++v) views.colour[v] = colour_ptrs[v];
    auto aligned = [](const GutAdamTensors& t) { return (((uintptr_t)t.param | (uintptr_t)t.exp_avg | (uintptr_t)t.exp_avg_sq) & 15u) == 0; };
    const uint32_t vec_mask = (aligned(albedo) ? 1u : 0u) | (aligned(specular) ? 2u : 0u);   // a warp's span starts at 32 x width floats
    const dim3 grid((N + K6_THREADS - 1) / K6_THREADS);
    if (colour_compact)
        gut_sh_expand_adam_kernel<true><<<grid, K6_THREADS, 0, stream>>>(
            N, sph_degree, reinterpret_cast<const float4*>(density12), n_views, centres, views, albedo, specular, visibility, b1, b2, eps, vec_mask);
    else
        gut_sh_expand_adam_kernel<false><<<grid, K6_THREADS, 0, stream>>>(
            N, sph_degree, reinterpret_cast<const float4*>(density12), n_views, centres, views, albedo, specular, visibility, b1, b2, eps, vec_mask);
    return cudaGetLastError();
}

// colour_ptrs: HOST array of n_views device pointers.  More than GUT_EXPAND_MAX_VIEWS views are expanded in several
// launches, the later ones accumulating.
cudaError_t gut_launch_sh_expand_ptrs(uint32_t N, int sph_degree, int ncoef_all, const float* density12, uint32_t n_views, const float* centres,
                                      const float* const* colour_ptrs, int colour_compact, float* sph_grad, int accumulate, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    uint32_t done = 0;
    do {
        const uint32_t nv = n_views - done < GUT_EXPAND_MAX_VIEWS ? n_views - done : GUT_EXPAND_MAX_VIEWS;
        ExpandViews views = {};
        // the wide fetch (cp.async into the output tile) is used for 16-byte aligned slots in THIS device's memory: it is measured
        // there (0.35 -> 0.33 ms); a peer's slots keep the per-lane loads the multi-GPU figures were taken with until the wide form
        // has been timed across NVLink
        views.wide = colour_compact ? 0 : 1;
        int here = 0;
        if (cudaGetDevice(&here) != cudaSuccess) views.wide = 0;
        for (uint32_t v = 0; v < nv; ++v) {
            views.colour[v] = colour_ptrs[done + v];
            if ((uintptr_t)colour_ptrs[done + v] & 15u) views.wide = 0;
            if (views.wide) {
                cudaPointerAttributes at;
                if (cudaPointerGetAttributes(&at, colour_ptrs[done + v]) != cudaSuccess) { (void)cudaGetLastError(); views.wide = 0; }
                else if (at.type != cudaMemoryTypeDevice || at.device != here) views.wide = 0;
            }
        }
        const dim3 grid((N + K6_THREADS - 1) / K6_THREADS);
        if (colour_compact)
            gut_sh_expand_kernel<true><<<grid, K6_THREADS, 0, stream>>>(
                N, sph_degree, reinterpret_cast<const float4*>(density12), nv, centres + 3 * (size_t)done, views,
                reinterpret_cast<float4*>(sph_grad), (accumulate || done > 0) ? 1 : 0, ncoef_all);
        else
            gut_sh_expand_kernel<false><<<grid, K6_THREADS, 0, stream>>>(
                N, sph_degree, reinterpret_cast<const float4*>(density12), nv, centres + 3 * (size_t)done, views,
                reinterpret_cast<float4*>(sph_grad), (accumulate || done > 0) ? 1 : 0, ncoef_all);
        done += nv;
    } while (done < n_views);
    return cudaGetLastError();
}

cudaError_t gut_launch_sh_expand(uint32_t N, int sph_degree, int ncoef_all, const float* density12, uint32_t n_views, const float* centres,
                                 const float* colour_grads, float* sph_grad, int accumulate, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    if (n_views == 0) return gut_launch_sh_expand_ptrs(N, sph_degree, ncoef_all, density12, 0, centres, nullptr, 0, sph_grad, accumulate, stream);
    uint32_t done = 0;
    cudaError_t e = cudaSuccess;
    while (done < n_views && e == cudaSuccess) {   // contiguous [n_views][N][3] buffer -> pointer table, one chunk per launch
        const float* ptrs[GUT_EXPAND_MAX_VIEWS];
        const uint32_t nv = n_views - done < GUT_EXPAND_MAX_VIEWS ? n_views - done : GUT_EXPAND_MAX_VIEWS;
        for (uint32_t v = 0; v < nv; ++v) ptrs[v] = colour_grads + 3 * (size_t)(done + v) * N;
        e = gut_launch_sh_expand_ptrs(N, sph_degree, ncoef_all, density12, nv, centres + 3 * (size_t)done, ptrs, 0, sph_grad, (accumulate || done > 0) ? 1 : 0, stream);
        done += nv;
    }
    return e;
}

cudaError_t gut_launch_copy_features3(uint32_t N, const float4* features4, float* features3, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    gut_copy_features3_kernel<<<(N + 255) / 256, 256, 0, stream>>>(N, features4, features3);
    return cudaGetLastError();
}


// ---------------------------------------------------------------------------------------------------
// Packing of the model's parameter tensors into the [N,12] record the kernels read, and the inverse for its gradient
// (reference: torch.cat in threedgut_tracer/tracer.py:176-178 and the column split autograd derives from it).
// One thread per particle; a warp's reads of a [N,3] / [N,4] tensor cover one contiguous span.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) gut_pack_particles_kernel(uint32_t N, const float* __restrict__ pos, const float* __restrict__ dens,
                                                                 const float4* __restrict__ rot, const float* __restrict__ scale,
                                                                 float4* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const float px = __ldg(pos + 3 * (size_t)i), py = __ldg(pos + 3 * (size_t)i + 1), pz = __ldg(pos + 3 * (size_t)i + 2);
    const float d = __ldg(dens + i);
    const float4 q = __ldg(rot + i);
    const float sx = __ldg(scale + 3 * (size_t)i), sy = __ldg(scale + 3 * (size_t)i + 1), sz = __ldg(scale + 3 * (size_t)i + 2);
    out[3 * (size_t)i] = make_float4(px, py, pz, d);
    out[3 * (size_t)i + 1] = q;
    out[3 * (size_t)i + 2] = make_float4(sx, sy, sz, 0.0f);
}

__global__ void __launch_bounds__(256) gut_unpack_gradient_kernel(uint32_t N, const float4* __restrict__ grad12, float* __restrict__ pos,
                                                                  float* __restrict__ dens, float4* __restrict__ rot, float* __restrict__ scale) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const float4 a = __ldg(grad12 + 3 * (size_t)i), q = __ldg(grad12 + 3 * (size_t)i + 1), c = __ldg(grad12 + 3 * (size_t)i + 2);
    pos[3 * (size_t)i] = a.x; pos[3 * (size_t)i + 1] = a.y; pos[3 * (size_t)i + 2] = a.z;
    dens[i] = a.w;
    rot[i] = q;
    scale[3 * (size_t)i] = c.x; scale[3 * (size_t)i + 1] = c.y; scale[3 * (size_t)i + 2] = c.z;
}

cudaError_t gut_launch_pack_particles(uint32_t N, const float* pos, const float* dens, const float* rot, const float* scale, float* out12,
                                      cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    gut_pack_particles_kernel<<<(N + 255) / 256, 256, 0, stream>>>(N, pos, dens, reinterpret_cast<const float4*>(rot), scale,
                                                                  reinterpret_cast<float4*>(out12));
    return cudaGetLastError();
}

cudaError_t gut_launch_unpack_gradient(uint32_t N, const float* grad12, float* pos, float* dens, float* rot, float* scale, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    gut_unpack_gradient_kernel<<<(N + 255) / 256, 256, 0, stream>>>(N, reinterpret_cast<const float4*>(grad12), pos, dens,
                                                                   reinterpret_cast<float4*>(rot), scale);
    return cudaGetLastError();
}
