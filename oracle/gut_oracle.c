/*
 * gut_oracle.c -- CPU restatement of the reference's 3DGUT rasteriser
 * (threedgut_tracer) in plain C, IEEE fp32, no FMA contraction.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
 * may load it.  The product path (3dgrut_b200/) never links or calls it.
 *
 * PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures
 * for this path and its CUDA tracer cannot be built in
 * this image (it needs `slangc`, threedgut_tracer/setup_3dgut.py:99-121).  The
 * oracle is therefore anchored on (i) a line-by-line restatement of the
 * reference sources cited at every function below, (ii) a second, independent
 * PyTorch transcription with autograd (oracle/gut_torch_ref.py) that must
 * agree with it, (iii) closed-form cases in tests/.
 *
 * Paths below are relative to /root/reference/threedgut_tracer/ unless they
 * start with thirdparty/.
 *
 * Build: see oracle/Makefile  (gcc -O2 -ffp-contract=off -fopenmp).
 */
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#include "../include/gut_b200.h"

#define ORACLE_API __attribute__((visibility("default")))

typedef struct { float x, y; } v2;
typedef struct { float x, y, z; } v3;
typedef struct { float x, y, z, w; } v4;
/* column-major 3x3: c[i] is column i (thirdparty/tiny-cuda-nn/include/tiny-cuda-nn/vec.h:520-614) */
typedef struct { v3 c[3]; } m3;
typedef struct { float w, x, y, z; } quat; /* tcnn::quat{w,x,y,z} vec.h:1065 */

static inline v3 V3(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 add3(v3 a, v3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 sub3(v3 a, v3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 mul3(v3 a, v3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 scl3(float s, v3 a) { return V3(s * a.x, s * a.y, s * a.z); }
/* mathUtils.cuh:313-315 (left-to-right sum) */
static inline float dot3(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
/* mathUtils.cuh:334-340 */
static inline v3 cross3(v3 a, v3 b) {
    return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
/* tcnn mat*vec: result[j] = 0 + m[0][j]*v[0] + m[1][j]*v[1] + m[2][j]*v[2]  (vec.h:588-598) */
static inline v3 m3_mul(const m3* m, v3 v) {
    v3 r;
    r.x = 0.0f + m->c[0].x * v.x; r.x += m->c[1].x * v.y; r.x += m->c[2].x * v.z;
    r.y = 0.0f + m->c[0].y * v.x; r.y += m->c[1].y * v.y; r.y += m->c[2].y * v.z;
    r.z = 0.0f + m->c[0].z * v.x; r.z += m->c[1].z * v.y; r.z += m->c[2].z * v.z;
    return r;
}

/* tcnn::to_mat3(quat) vec.h:1172-1183 */
static m3 quat_to_mat3(quat q) {
    float qxx = q.x * q.x, qyy = q.y * q.y, qzz = q.z * q.z;
    float qxz = q.x * q.z, qxy = q.x * q.y, qyz = q.y * q.z;
    float qwx = q.w * q.x, qwy = q.w * q.y, qwz = q.w * q.z;
    m3 m;
    m.c[0] = V3(1.0f - 2.0f * (qyy + qzz), 2.0f * (qxy + qwz), 2.0f * (qxz - qwy));
    m.c[1] = V3(2.0f * (qxy - qwz), 1.0f - 2.0f * (qxx + qzz), 2.0f * (qyz + qwx));
    m.c[2] = V3(2.0f * (qxz + qwy), 2.0f * (qyz - qwx), 1.0f - 2.0f * (qxx + qyy));
    return m;
}

/* tcnn::tquat(const mat3&) vec.h:1066-1097; m[i][j] = column i, row j */
static quat mat3_to_quat(const m3* mm) {
    float m00 = mm->c[0].x, m01 = mm->c[0].y, m02 = mm->c[0].z;
    float m10 = mm->c[1].x, m11 = mm->c[1].y, m12 = mm->c[1].z;
    float m20 = mm->c[2].x, m21 = mm->c[2].y, m22 = mm->c[2].z;
    quat q;
    float tr = m00 + m11 + m22;
    if (tr > 0.0f) {
        float S = sqrtf(tr + 1.0f) * 2.0f;
        q.w = 0.25f * S; q.x = (m12 - m21) / S; q.y = (m20 - m02) / S; q.z = (m01 - m10) / S;
    } else if (m00 > m11 && m00 > m22) {
        float S = sqrtf(1.0f + m00 - m11 - m22) * 2.0f;
        q.w = (m12 - m21) / S; q.x = 0.25f * S; q.y = (m10 + m01) / S; q.z = (m20 + m02) / S;
    } else if (m11 > m22) {
        float S = sqrtf(1.0f + m11 - m00 - m22) * 2.0f;
        q.w = (m20 - m02) / S; q.x = (m10 + m01) / S; q.y = 0.25f * S; q.z = (m21 + m12) / S;
    } else {
        float S = sqrtf(1.0f + m22 - m00 - m11) * 2.0f;
        q.w = (m01 - m10) / S; q.x = (m20 + m02) / S; q.y = (m21 + m12) / S; q.z = 0.25f * S;
    }
    return q;
}

/* tcnn::slerp(quat,quat,t) vec.h:1133-1153; dot is (w*w + x*x) + (y*y + z*z) vec.h:1107 */
static quat quat_slerp(quat x, quat y, float t) {
    quat z = y;
    float cos_theta = (x.w * y.w + x.x * y.x) + (x.y * y.y + x.z * y.z);
    if (cos_theta < 0.0f) {
        z.w = -y.w; z.x = -y.x; z.y = -y.y; z.z = -y.z;
        cos_theta = -cos_theta;
    }
    quat r;
    if (cos_theta > 1.0f - FLT_EPSILON) {
        /* mix(x, z, t) = x*(1-t) + z*t  vec.h:1111 */
        float a = 1.0f - t;
        r.w = x.w * a + z.w * t; r.x = x.x * a + z.x * t; r.y = x.y * a + z.y * t; r.z = x.z * a + z.z * t;
    } else {
        float angle = acosf(cos_theta);
        float s0 = sinf((1.0f - t) * angle), s1 = sinf(t * angle), sd = sinf(angle);
        r.w = (s0 * x.w + s1 * z.w) / sd; r.x = (s0 * x.x + s1 * z.x) / sd;
        r.y = (s0 * x.y + s1 * z.y) / sd; r.z = (s0 * x.z + s1 * z.z) / sd;
    }
    return r;
}

/* pose = [tx,ty,tz,qx,qy,qz,qw] -> tcnn::quat{pose[6],pose[3],pose[4],pose[5]} (sensors.h:33, cameraProjections.cuh:155) */
static inline quat pose_quat(const float* p) { quat q = {p[6], p[3], p[4], p[5]}; return q; }
static inline v3 pose_t(const float* p) { return V3(p[0], p[1], p[2]); }
/* tcnn::mix(a,b,c) = a*(1-c) + b*c  vec.h:183,277 */
static inline float mixf(float a, float b, float c) { return a * (1.0f - c) + b * c; }
static inline v3 mix3(v3 a, v3 b, float c) { return V3(mixf(a.x, b.x, c), mixf(a.y, b.y, c), mixf(a.z, b.z, c)); }

/* interpolatedSensorPose, include/3dgut/sensors/sensors.h:53-66 */
static void pose_interpolate(const float* a, const float* b, float t, float* out) {
    quat q = quat_slerp(pose_quat(a), pose_quat(b), t);
    v3 tr = mix3(pose_t(a), pose_t(b), t);
    out[0] = tr.x; out[1] = tr.y; out[2] = tr.z;
    out[3] = q.x; out[4] = q.y; out[5] = q.z; out[6] = q.w;
}

/* sensorPoseInverse, sensors.h:44-51 */
static void pose_inverse(const float* p, float* out) {
    m3 r = quat_to_mat3(pose_quat(p));
    m3 inv; /* transpose */
    inv.c[0] = V3(r.c[0].x, r.c[1].x, r.c[2].x);
    inv.c[1] = V3(r.c[0].y, r.c[1].y, r.c[2].y);
    inv.c[2] = V3(r.c[0].z, r.c[1].z, r.c[2].z);
    quat iq = mat3_to_quat(&inv);
    /* -1.0f * invRotation * t : scalar*matrix first (vec.h:693), then mat*vec */
    m3 neg;
    for (int i = 0; i < 3; ++i) neg.c[i] = scl3(-1.0f, inv.c[i]);
    v3 t = m3_mul(&neg, pose_t(p));
    out[0] = t.x; out[1] = t.y; out[2] = t.z;
    out[3] = iq.x; out[4] = iq.y; out[5] = iq.z; out[6] = iq.w;
}

/* 4x3 matrix = [R | t] ; sensorPoseToMat sensors.h:68-73 */
typedef struct { m3 r; v3 t; } m4x3;
static m4x3 pose_to_mat(const float* p) {
    m4x3 m; m.r = quat_to_mat3(pose_quat(p)); m.t = pose_t(p); return m;
}

/*
 * Deterministic natural log shared bit-for-bit with the CUDA kernels
 * (3dgrut_b200/csrc/gut_math.cuh::gut_logf).  The reference calls logf under
 * -use_fast_math (gutProjector.cuh:107,361), i.e. the approximate __logf;
 * both that and libm differ from each other in the last bits, which would make
 * the tile-culling comparison irreproducible.  This version uses only IEEE
 * +,*,/ and fmaf, so the oracle and the GPU agree exactly.  |err| < 2 ulp.
 */
static float gut_logf(float x) {
    union { float f; uint32_t u; } v; v.f = x;
    /* split x = m * 2^e with m in [sqrt(0.5), sqrt(2)) */
    int32_t ix = (int32_t)(v.u - 0x3f3504f3u);
    int32_t e = ix >> 23; /* arithmetic shift */
    v.u = (uint32_t)(ix & 0x007fffff) + 0x3f3504f3u;
    float m = v.f;
    float s = (m - 1.0f) / (m + 1.0f);
    float z = s * s;
    /* 2*atanh(s) = 2s(1 + z/3 + z^2/5 + z^3/7 + z^4/9 + z^5/11) */
    float p = 0.18181818181818182f;
    p = fmaf(p, z, 0.22222222222222222f);
    p = fmaf(p, z, 0.2857142857142857f);
    p = fmaf(p, z, 0.4f);
    p = fmaf(p, z, 0.6666666666666666f);
    p = p * z;
    float r = fmaf(p, s, 2.0f * s);
    float fe = (float)e;
    return fmaf(fe, 0.6931471805599453f, r);
}
ORACLE_API float gut_oracle_logf(float x) { return gut_logf(x); }

/* ------------------------------------------------------------------------------------------
 * Camera projection: include/3dgut/kernels/cuda/sensors/cameraProjections.cuh
 * ---------------------------------------------------------------------------------------- */

/* cameraProjections.cuh:52-55 */
static int within_resolution(float rx, float ry, float tol, v2 p) {
    float mx = rx * tol, my = ry * tol;
    return (p.x > -mx) && (p.y > -my) && (p.x < rx + mx) && (p.y < ry + my);
}

/* cameraProjections.cuh:57-103 */
static int project_pinhole(const GutCamera* cam, int W, int H, v3 pos, float tol, v2* out) {
    if (pos.z <= 0.0f) { out->x = 0.0f; out->y = 0.0f; return 0; }
    float ux = pos.x / pos.z, uy = pos.y / pos.z;
    float sx = ux * ux, sy = uy * uy;
    float r2 = sx + sy;
    float a1 = 2.0f * ux * uy;
    float a2 = r2 + 2.0f * sx;
    float a3 = r2 + 2.0f * sy;
    const float* k = cam->radial_coeffs; const float* p = cam->tangential_coeffs; const float* s = cam->thin_prism_coeffs;
    float num = 1.0f + r2 * (k[0] + r2 * (k[1] + r2 * k[2]));
    float den = 1.0f + r2 * (k[3] + r2 * (k[4] + r2 * k[5]));
    float icD = num / den;
    float dx = p[0] * a1 + p[1] * a2 + r2 * (s[0] + r2 * s[1]);
    float dy = p[0] * a3 + p[1] * a1 + r2 * (s[2] + r2 * s[3]);
    float ndx = icD * ux + dx, ndy = icD * uy + dy;
    int valid_radial = (icD > 0.8f) && (icD < 1.2f);
    if (valid_radial) {
        out->x = ndx * cam->focal_length[0] + cam->principal_point[0];
        out->y = ndy * cam->focal_length[1] + cam->principal_point[1];
    } else {
        /* hypotf(resolution.x, resolution.y): restated as sqrtf(x*x+y*y) on both sides (coordinates in this
         * branch "cannot be trusted", cameraProjections.cuh:94-99) */
        float fw = (float)W, fh = (float)H;
        float clip = sqrtf(fw * fw + fh * fh);
        float f = clip / sqrtf(r2);
        out->x = f * ux + cam->principal_point[0];
        out->y = f * uy + cam->principal_point[1];
    }
    return valid_radial && within_resolution((float)W, (float)H, tol, *out);
}

/* cameraProjections.cuh:22-33 + 105-128 */
static int project_fisheye(const GutCamera* cam, int W, int H, v3 pos, float tol, v2* out) {
    const float eps = FLT_EPSILON;
    float rho = fmaxf(sqrtf(pos.x * pos.x + pos.y * pos.y), eps);
    float theta_full = atan2f(rho, pos.z);
    float theta = fminf(theta_full, cam->max_angle);
    float theta2 = theta * theta;
    const float* c = cam->radial_coeffs;
    float y = c[3];
    y = theta2 * y + c[2];
    y = theta2 * y + c[1];
    y = theta2 * y + c[0];
    float delta = (theta * (y * theta2 + 1.0f)) / rho;
    out->x = cam->focal_length[0] * pos.x * delta + cam->principal_point[0];
    out->y = cam->focal_length[1] * pos.y * delta + cam->principal_point[1];
    return (theta < cam->max_angle) && within_resolution((float)W, (float)H, tol, *out);
}

/* cameraProjections.cuh:130-144 */
static int project_point(const GutCamera* cam, int W, int H, v3 pos, float tol, v2* out) {
    switch (cam->model) {
    case GUT_CAMERA_OPENCV_PINHOLE: return project_pinhole(cam, W, H, pos, tol, out);
    case GUT_CAMERA_OPENCV_FISHEYE: return project_fisheye(cam, W, H, pos, tol, out);
    default: out->x = 0.0f; out->y = 0.0f; return 0;
    }
}

/* cameraProjections.cuh:35-50 */
static float relative_shutter_time(const GutCamera* cam, int W, int H, v2 p) {
    switch (cam->shutter) {
    case GUT_SHUTTER_ROLLING_TOP_TO_BOTTOM: return floorf(p.y) / ((float)H - 1.0f);
    case GUT_SHUTTER_ROLLING_LEFT_TO_RIGHT: return floorf(p.x) / ((float)W - 1.0f);
    case GUT_SHUTTER_ROLLING_BOTTOM_TO_TOP: return ((float)H - ceilf(p.y)) / ((float)H - 1.0f);
    case GUT_SHUTTER_ROLLING_RIGHT_TO_LEFT: return ((float)W - ceilf(p.x)) / ((float)W - 1.0f);
    default: return 0.5f;
    }
}

/* to_mat3(q) * position + t   (cameraProjections.cuh:157,166,179) */
static v3 transform_point(quat q, v3 t, v3 p) {
    m3 r = quat_to_mat3(q);
    return add3(m3_mul(&r, p), t);
}

/* cameraProjections.cuh:146-185 */
static int project_point_with_shutter(const GutCamera* cam, int W, int H, int n_iter, v3 pos, float tol, v2* out) {
    v3 t0 = pose_t(cam->start_pose); quat q0 = pose_quat(cam->start_pose);
    int valid = project_point(cam, W, H, transform_point(q0, t0, pos), tol, out);
    if (cam->shutter == GUT_SHUTTER_GLOBAL) return valid;
    v3 t1 = pose_t(cam->end_pose); quat q1 = pose_quat(cam->end_pose);
    if (!valid) {
        valid = project_point(cam, W, H, transform_point(q1, t1, pos), tol, out);
        if (!valid) return 0;
    }
    for (int i = 0; i < n_iter; ++i) {
        float alpha = relative_shutter_time(cam, W, H, *out);
        valid = project_point(cam, W, H, transform_point(quat_slerp(q0, q1, alpha), mix3(t0, t1, alpha), pos), tol, out);
    }
    return valid;
}

/* ------------------------------------------------------------------------------------------
 * Particle helpers
 * ---------------------------------------------------------------------------------------- */

/* ParticleDensity, include/3dgut/kernels/cuda/models/gaussianParticles.cuh:20-26 : 12 floats / particle */
typedef struct { v3 pos; float density; float qw, qx, qy, qz; v3 scale; float pad; } raw_particle;

/* rotationT rows (== columns of R): transforms.slang:21-38 / quaternionWXYZToMatrix gaussianParticles.cuh:35-55 */
static void particle_rotT(const raw_particle* p, v3 rt[3]) {
    float r = p->qw, x = p->qx, y = p->qy, z = p->qz;
    float xx = x * x, yy = y * y, zz = z * z, xy = x * y, xz = x * z, yz = y * z, rx = r * x, ry = r * y, rz = r * z;
    rt[0] = V3(1.0f - 2.0f * (yy + zz), 2.0f * (xy + rz), 2.0f * (xz - ry));
    rt[1] = V3(2.0f * (xy - rz), 1.0f - 2.0f * (xx + zz), 2.0f * (yz + rx));
    rt[2] = V3(2.0f * (xz + ry), 2.0f * (yz - rx), 1.0f - 2.0f * (xx + yy));
}

static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                               -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

/* SH basis value for coefficient k, shared by forward (gaussianParticles.cuh:64-96) and backward (:120-187) */
static void sh_basis(int deg, v3 d, float b[16]) {
    float x = d.x, y = d.y, z = d.z;
    b[0] = SH_C0;
    if (deg > 0) {
        b[1] = -SH_C1 * y; b[2] = SH_C1 * z; b[3] = -SH_C1 * x;
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            b[4] = SH_C2[0] * xy; b[5] = SH_C2[1] * yz; b[6] = SH_C2[2] * (2.0f * zz - xx - yy);
            b[7] = SH_C2[3] * xz; b[8] = SH_C2[4] * (xx - yy);
            if (deg > 2) {
                b[9]  = SH_C3[0] * y * (3.0f * xx - yy);
                b[10] = SH_C3[1] * xy * z;
                b[11] = SH_C3[2] * y * (4.0f * zz - xx - yy);
                b[12] = SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
                b[13] = SH_C3[4] * x * (4.0f * zz - xx - yy);
                b[14] = SH_C3[5] * z * (xx - yy);
                b[15] = SH_C3[6] * x * (xx - 3.0f * yy);
            }
        }
    }
}

/* radianceFromSpH(deg, coeffs, dir, clamped=false)  gaussianParticles.cuh:64-96: sum in coefficient order, +0.5, unclamped */
static v3 sh_to_radiance(int deg, const float* coeffs /* [ncoef][3] */, v3 dir) {
    float b[16];
    sh_basis(deg, dir, b);
    int n = (deg + 1) * (deg + 1);
    v3 rad = V3(b[0] * coeffs[0], b[0] * coeffs[1], b[0] * coeffs[2]);
    for (int k = 1; k < n; ++k) {
        rad.x = rad.x + b[k] * coeffs[3 * k + 0];
        rad.y = rad.y + b[k] * coeffs[3 * k + 1];
        rad.z = rad.z + b[k] * coeffs[3 * k + 2];
    }
    rad.x += 0.5f; rad.y += 0.5f; rad.z += 0.5f;
    return rad;
}

#define TILE 16 /* GUTParameters::Tiling::BlockX/BlockY, renderer/gutRendererParameters.h:22-23 */

typedef struct { uint32_t minx, miny, maxx, maxy; } tile_bbox;

/* computeTileSpaceBBox, gutProjector.cuh:32-43 */
static tile_bbox tile_space_bbox(uint32_t gx, uint32_t gy, v2 pos, v2 ext) {
    tile_bbox b;
    int v;
    v = (int)floorf((pos.x - 0.5f - ext.x) / (float)TILE); v = v < 0 ? 0 : v; b.minx = (uint32_t)v < gx ? (uint32_t)v : gx;
    v = (int)floorf((pos.y - 0.5f - ext.y) / (float)TILE); v = v < 0 ? 0 : v; b.miny = (uint32_t)v < gy ? (uint32_t)v : gy;
    v = (int)ceilf((pos.x - 0.5f + ext.x) / (float)TILE);  v = v < 0 ? 0 : v; b.maxx = (uint32_t)v < gx ? (uint32_t)v : gx;
    v = (int)ceilf((pos.y - 0.5f + ext.y) / (float)TILE);  v = v < 0 ? 0 : v; b.maxy = (uint32_t)v < gy ? (uint32_t)v : gy;
    return b;
}

static inline float saturatef(float x) { return (x != x) ? 0.0f : (x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x)); } /* __saturatef: NaN -> 0 */

/* tileMinParticlePowerResponse, gutProjector.cuh:49-78 (evaluation order kept; every * and + rounds separately) */
static float tile_min_power(float tx_, float ty_, v4 co, v2 mean) {
    const float ts = (float)TILE;
    float tminx = ts * tx_, tminy = ts * ty_;
    float tmaxx = ts + tminx, tmaxy = ts + tminy;
    float mox = tminx - mean.x, moy = tminy - mean.y;
    float lax = mox > 0.0f ? 1.0f : 0.0f, lay = moy > 0.0f ? 1.0f : 0.0f;
    float nrx = lax + (mean.x > tmaxx ? 1.0f : 0.0f), nry = lay + (mean.y > tmaxy ? 1.0f : 0.0f);
    if ((nrx + nry) > 0.0f) {
        float px = mixf(tmaxx, tminx, lax), py = mixf(tmaxy, tminy, lay);
        float dxx = copysignf(ts, mox), dxy = copysignf(ts, moy);
        float dfx = mean.x - px, dfy = mean.y - py;
        float rcx = 1.0f / (ts * ts * co.x), rcy = 1.0f / (ts * ts * co.z); /* __frcp_rn */
        float tx = nry * saturatef((dxx * co.x * dfx + dxx * co.y * dfy) * rcx);
        float ty = nrx * saturatef((dxy * co.y * dfx + dxy * co.z * dfy) * rcy);
        float mx = mean.x - (px + tx * dxx), my = mean.y - (py + ty * dxy);
        return 0.5f * (co.x * mx * mx + co.z * my * my) + co.y * mx * my;
    }
    return 0.0f;
}

/* computeProjectedExtentConicOpacity, gutProjector.cuh:81-116 */
static int extent_conic_opacity(const GutConfig* cfg, v3 cov, float opacity, v2* extent, v4* co, float* max_power) {
    const float dil = 0.3f; /* TGUTProjectorParams::CovarianceDilation threedgut.cuh:49 */
    v3 d = V3(cov.x + dil, cov.y, cov.z + dil);
    float ddet = d.x * d.z - d.y * d.y;
    if (ddet == 0.0f) return 0;
    co->x = d.z / ddet; co->y = -d.y / ddet; co->z = d.x / ddet;
    {   /* MipSplattingScaling = true threedgut.cuh:59 */
        float det = cov.x * cov.z - cov.y * cov.y;
        float conv = sqrtf(fmaxf(0.000025f, det / ddet));
        co->w = opacity * conv;
    }
    if (co->w < cfg->min_alpha) return 0;
    *max_power = gut_logf(co->w / cfg->min_alpha);
    float ef = cfg->tight_opacity_bounding ? fminf(3.33f, sqrtf(2.0f * *max_power)) : 3.33f;
    float mid = 0.5f * (d.x + d.z);
    float lambda = mid + sqrtf(fmaxf(0.01f, mid * mid - ddet));
    float radius = ef * sqrtf(lambda);
    if (cfg->rect_bounding) {
        extent->x = fminf(ef * sqrtf(d.x), radius);
        extent->y = fminf(ef * sqrtf(d.z), radius);
    } else {
        extent->x = radius; extent->y = radius;
    }
    return radius > 0.0f;
}

/* per-frame host constants, gutRenderer.cu:259-267,282-285,385 */
typedef struct {
    uint32_t gx, gy;
    float pose_mid[7], pose_mid_inv[7];
    m4x3 view;      /* sensorPoseToMat(sensorPose)    : world -> sensor  */
    m4x3 c2w;       /* sensorPoseToMat(sensorPoseInv) : sensor -> world  */
    v3 cam_centre;  /* sensorPoseInverse(sensorPose).slice<0,3>()        */
} frame_consts;

static frame_consts make_frame(const GutCamera* cam, uint32_t W, uint32_t H) {
    frame_consts f;
    f.gx = (W + TILE - 1) / TILE; f.gy = (H + TILE - 1) / TILE;
    pose_interpolate(cam->start_pose, cam->end_pose, 0.5f, f.pose_mid);
    pose_inverse(f.pose_mid, f.pose_mid_inv);
    f.view = pose_to_mat(f.pose_mid);
    f.c2w = pose_to_mat(f.pose_mid_inv);
    f.cam_centre = pose_t(f.pose_mid_inv);
    return f;
}

/* exported so tests can pin the host-side pose algebra against the CUDA library's */
ORACLE_API void gut_oracle_frame_consts(const GutCamera* cam, uint32_t W, uint32_t H, float* out /* 7+7+12+12+3 */) {
    frame_consts f = make_frame(cam, W, H);
    memcpy(out, f.pose_mid, 28); memcpy(out + 7, f.pose_mid_inv, 28);
    const m4x3* ms[2] = {&f.view, &f.c2w};
    for (int k = 0; k < 2; ++k) {
        float* o = out + 14 + 12 * k;
        for (int c = 0; c < 3; ++c) { o[3 * c] = ms[k]->r.c[c].x; o[3 * c + 1] = ms[k]->r.c[c].y; o[3 * c + 2] = ms[k]->r.c[c].z; }
        o[9] = ms[k]->t.x; o[10] = ms[k]->t.y; o[11] = ms[k]->t.z;
    }
    out[38] = f.cam_centre.x; out[39] = f.cam_centre.y; out[40] = f.cam_centre.z;
}

/* view-space z: sensorMatrix[0][2]*x + [1][2]*y + [2][2]*z + [3][2]  gutProjector.cuh:137-138,317-318 */
static inline float view_z(const m4x3* v, v3 p) {
    return p.x * v->r.c[0].z + p.y * v->r.c[1].z + p.z * v->r.c[2].z + v->t.z;
}

/* ------------------------------------------------------------------------------------------
 * K1: GUTProjector::eval  gutProjector.cuh:217-322 with unscentedParticleProjection :118-215
 * ---------------------------------------------------------------------------------------- */
ORACLE_API void gut_oracle_project(const GutConfig* cfg, const GutCamera* cam, uint32_t N, int sph_degree,
                                   const float* density12, const float* sph, uint32_t W, uint32_t H,
                                   uint32_t* tiles_count, float* proj_pos, float* conic_opacity, float* proj_extent,
                                   float* depth, float* features, float* visibility) {
    const frame_consts f = make_frame(cam, W, H);
    const int ncoef = (cfg->sph_max_degree + 1) * (cfg->sph_max_degree + 1);
    /* UT constants, threedgut.cuh:62-71 ; delta computed in double by setup_3dgut.py:45 then narrowed */
    const float ua = cfg->ut_alpha, ub = cfg->ut_beta, uk = cfg->ut_kappa;
    const float D = 3.0f;
    const float lambda = ua * ua * (D + uk) - D;
    const float delta = (float)sqrt((double)ua * (double)ua * (3.0 + (double)uk));
    const float w0_mean = lambda / (D + lambda);
    const float wi = 1.0f / (2.0f * (D + lambda));
    const float w0_cov = lambda / (D + lambda) + (1.0f - ua * ua + ub);
    const float tol = cfg->ut_in_image_margin_factor;
    const int n_iter = cfg->n_rolling_shutter_iterations;

#pragma omp parallel for schedule(dynamic, 1024)
    for (int64_t i = 0; i < (int64_t)N; ++i) {
        const raw_particle* p = (const raw_particle*)(density12 + 12 * i);
        int valid_proj = 0;
        v2 centre = {0.0f, 0.0f};
        v3 cov = {0.0f, 0.0f, 0.0f};
        float opacity = p->density;
        v3 sensor_ray = {0.0f, 0.0f, 0.0f};
        do {
            if (opacity < cfg->min_alpha) break;
            if (view_z(&f.view, p->pos) < 0.2f) break; /* ParticleMinSensorZ threedgut.cuh:48 */
            v3 rt[3];
            particle_rotT(p, rt); /* rotation()[i] == column i of R == row i of rotationT, shRadiativeGaussianParticles.cuh:84-87 */
            sensor_ray = sub3(p->pos, f.cam_centre);
            v2 sp[7];
            int nvalid = 0;
            nvalid += project_point_with_shutter(cam, (int)W, (int)H, n_iter, p->pos, tol, &sp[0]);
            centre.x = sp[0].x * w0_mean; centre.y = sp[0].y * w0_mean;
            const float sc[3] = {p->scale.x, p->scale.y, p->scale.z};
            for (int a = 0; a < 3; ++a) {
                float ds = delta * sc[a];
                v3 d = scl3(ds, rt[a]); /* Delta * scale[i] * rotation[i] */
                nvalid += project_point_with_shutter(cam, (int)W, (int)H, n_iter, add3(p->pos, d), tol, &sp[a + 1]);
                centre.x += wi * sp[a + 1].x; centre.y += wi * sp[a + 1].y;
                nvalid += project_point_with_shutter(cam, (int)W, (int)H, n_iter, sub3(p->pos, d), tol, &sp[a + 4]);
                centre.x += wi * sp[a + 4].x; centre.y += wi * sp[a + 4].y;
            }
            if (cfg->ut_require_all_sigma_points_valid ? (nvalid < 7) : (nvalid == 0)) break;
            {
                float cx = sp[0].x - centre.x, cy = sp[0].y - centre.y;
                cov = V3(w0_cov * (cx * cx), w0_cov * (cx * cy), w0_cov * (cy * cy));
            }
            for (int k = 0; k < 6; ++k) {
                float cx = sp[k + 1].x - centre.x, cy = sp[k + 1].y - centre.y;
                cov.x += wi * (cx * cx); cov.y += wi * (cx * cy); cov.z += wi * (cy * cy);
            }
            valid_proj = 1;
        } while (0);

        v2 extent = {0.0f, 0.0f}; v4 co = {0.0f, 0.0f, 0.0f, 0.0f}; float max_power = 0.0f;
        /* the reference runs this on an uninitialised covariance when the projection was rejected
         * (gutProjector.cuh:242-273): visibility of rejected particles is undefined there; 0 here. */
        int valid_conic = valid_proj ? extent_conic_opacity(cfg, cov, opacity, &extent, &co, &max_power) : 0;
        visibility[i] = valid_conic ? 1.0f : 0.0f;
        int valid = valid_proj && valid_conic;

        uint32_t ntiles = 0;
        if (valid) {
            tile_bbox bb = tile_space_bbox(f.gx, f.gy, centre, extent);
            if (cfg->tile_based_culling) {
                for (uint32_t y = bb.miny; y < bb.maxy; ++y)
                    for (uint32_t x = bb.minx; x < bb.maxx; ++x)
                        if (tile_min_power((float)x, (float)y, co, centre) < max_power) ntiles++;
            } else {
                ntiles = (bb.maxx - bb.minx) * (bb.maxy - bb.miny);
            }
        }
        tiles_count[i] = ntiles;
        if (ntiles == 0) {
            proj_pos[2 * i] = proj_pos[2 * i + 1] = 0.0f;
            conic_opacity[4 * i] = conic_opacity[4 * i + 1] = conic_opacity[4 * i + 2] = conic_opacity[4 * i + 3] = 0.0f;
            proj_extent[2 * i] = proj_extent[2 * i + 1] = 0.0f;
            depth[i] = 0.0f;
            /* reference leaves features[i] untouched (stale); the oracle writes 0 and tests ignore these rows */
            features[3 * i] = features[3 * i + 1] = features[3 * i + 2] = 0.0f;
            continue;
        }
        float dist = sqrtf(dot3(sensor_ray, sensor_ray));
        v3 dir = V3(sensor_ray.x / dist, sensor_ray.y / dist, sensor_ray.z / dist);
        v3 rad = sh_to_radiance(sph_degree, sph + (size_t)i * ncoef * 3, dir);
        features[3 * i] = rad.x; features[3 * i + 1] = rad.y; features[3 * i + 2] = rad.z;
        proj_pos[2 * i] = centre.x; proj_pos[2 * i + 1] = centre.y;
        conic_opacity[4 * i] = co.x; conic_opacity[4 * i + 1] = co.y; conic_opacity[4 * i + 2] = co.z; conic_opacity[4 * i + 3] = co.w;
        proj_extent[2 * i] = extent.x; proj_extent[2 * i + 1] = extent.y;
        depth[i] = cfg->global_z_order ? view_z(&f.view, p->pos) : dist;
    }
}

/* ------------------------------------------------------------------------------------------
 * L1: inclusive prefix sum (cub::DeviceScan::InclusiveSum, src/gutRenderer.cu:302-310)
 * ---------------------------------------------------------------------------------------- */
ORACLE_API uint32_t gut_oracle_scan(uint32_t N, const uint32_t* tiles_count, uint32_t* tiles_offset) {
    uint32_t acc = 0;
    for (uint32_t i = 0; i < N; ++i) { acc += tiles_count[i]; tiles_offset[i] = acc; }
    return N ? acc : 0;
}

/* ------------------------------------------------------------------------------------------
 * K2: GUTProjector::expand  gutProjector.cuh:324-388
 * keys[M] u64, values[M] u32 must be sized M = tiles_offset[N-1]
 * ---------------------------------------------------------------------------------------- */
ORACLE_API void gut_oracle_expand(const GutConfig* cfg, uint32_t N, uint32_t W, uint32_t H,
                                  const uint32_t* tiles_offset, const float* proj_pos, const float* conic_opacity,
                                  const float* proj_extent, const float* depth, uint64_t* keys, uint32_t* values) {
    const uint32_t gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(dynamic, 1024)
    for (int64_t i = 0; i < (int64_t)N; ++i) {
        v2 ext = {proj_extent[2 * i], proj_extent[2 * i + 1]};
        if (ext.x <= 1e-06f) continue;
        uint32_t depth_key; memcpy(&depth_key, &depth[i], 4);
        uint32_t off = (i == 0) ? 0 : tiles_offset[i - 1];
        v2 c = {proj_pos[2 * i], proj_pos[2 * i + 1]};
        tile_bbox bb = tile_space_bbox(gx, gy, c, ext);
        if (cfg->tile_based_culling) {
            const uint32_t max_off = tiles_offset[i];
            v4 co = {conic_opacity[4 * i], conic_opacity[4 * i + 1], conic_opacity[4 * i + 2], conic_opacity[4 * i + 3]};
            const float max_power = gut_logf(co.w / cfg->min_alpha);
            for (uint32_t y = bb.miny; (y < bb.maxy) && (off < max_off); ++y)
                for (uint32_t x = bb.minx; (x < bb.maxx) && (off < max_off); ++x)
                    if (tile_min_power((float)x, (float)y, co, c) < max_power) {
                        keys[off] = ((uint64_t)(y * gx + x) << 32) | depth_key;
                        values[off] = (uint32_t)i;
                        off++;
                    }
            for (; off < max_off; ++off) { /* pad: InvalidTileIdx, MaxDepthValue, InvalidParticleIdx */
                const float maxd = 3.4028235e+38f; uint32_t mb; memcpy(&mb, &maxd, 4);
                keys[off] = ((uint64_t)0xFFFFFFFFu << 32) | mb;
                values[off] = 0xFFFFFFFFu;
            }
        } else {
            for (uint32_t y = bb.miny; y < bb.maxy; ++y)
                for (uint32_t x = bb.minx; x < bb.maxx; ++x) {
                    keys[off] = ((uint64_t)(y * gx + x) << 32) | depth_key;
                    values[off] = (uint32_t)i;
                    off++;
                }
        }
    }
}

/* higherMsb, src/gutRenderer.cu:79-94 */
ORACLE_API uint32_t gut_oracle_higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/* ------------------------------------------------------------------------------------------
 * L2: cub::DeviceRadixSort::SortPairs on bits [0, 32+higherMsb(T))  src/gutRenderer.cu:356-365
 * restated as a stable LSD radix sort (16-bit digits) on the masked key; bits above end_bit are ignored
 * for ordering but carried through, exactly as CUB does.
 * ---------------------------------------------------------------------------------------- */
ORACLE_API int gut_oracle_sort(uint64_t M, uint32_t end_bit, const uint64_t* keys_in, const uint32_t* vals_in,
                               uint64_t* keys_out, uint32_t* vals_out) {
    if (M == 0) return 0;
    uint64_t* kb = (uint64_t*)malloc(M * 8); uint32_t* vb = (uint32_t*)malloc(M * 4);
    uint64_t* ka = (uint64_t*)malloc(M * 8); uint32_t* va = (uint32_t*)malloc(M * 4);
    size_t* hist = (size_t*)malloc(65537 * sizeof(size_t));
    if (!kb || !vb || !ka || !va || !hist) { free(kb); free(vb); free(ka); free(va); free(hist); return 1; }
    memcpy(ka, keys_in, M * 8); memcpy(va, vals_in, M * 4);
    for (uint32_t shift = 0; shift < end_bit; shift += 16) {
        uint32_t nb = end_bit - shift < 16 ? end_bit - shift : 16;
        uint64_t mask = ((uint64_t)1 << nb) - 1;
        memset(hist, 0, 65537 * sizeof(size_t));
        for (uint64_t i = 0; i < M; ++i) hist[((ka[i] >> shift) & mask) + 1]++;
        for (uint32_t d = 0; d < 65536; ++d) hist[d + 1] += hist[d];
        for (uint64_t i = 0; i < M; ++i) { size_t d = (ka[i] >> shift) & mask; size_t o = hist[d]++; kb[o] = ka[i]; vb[o] = va[i]; }
        uint64_t* tk = ka; ka = kb; kb = tk; uint32_t* tv = va; va = vb; vb = tv;
    }
    memcpy(keys_out, ka, M * 8); memcpy(vals_out, va, M * 4);
    free(ka); free(va); free(kb); free(vb); free(hist);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * K3: computeSortedTileRangeIndices  src/gutRenderer.cu:46-76 ; ranges[T][2] pre-zeroed (:161)
 * ---------------------------------------------------------------------------------------- */
ORACLE_API void gut_oracle_tile_ranges(uint64_t M, const uint64_t* sorted_keys, uint32_t T, uint32_t* ranges) {
    memset(ranges, 0, (size_t)T * 8);
    for (uint64_t k = 0; k < M; ++k) {
        uint32_t tile = (uint32_t)(sorted_keys[k] >> 32);
        int valid = tile != 0xFFFFFFFFu;
        if (k == 0) {
            if (valid) ranges[2 * tile] = (uint32_t)k;
        } else {
            uint32_t prev = (uint32_t)(sorted_keys[k - 1] >> 32);
            if (prev != tile) {
                if (prev != 0xFFFFFFFFu) ranges[2 * prev + 1] = (uint32_t)k;
                if (valid) ranges[2 * tile] = (uint32_t)k;
            }
        }
        if (valid && (k == M - 1)) ranges[2 * tile + 1] = (uint32_t)M;
    }
}

/* ------------------------------------------------------------------------------------------
 * Ray set-up: initializeRay  include/3dgut/kernels/cuda/common/rayPayload.cuh:75-108
 *             BoundingBox::ray_intersect  include/3dgut/utils/bounding_box.h:79-128
 * ---------------------------------------------------------------------------------------- */
typedef struct { v3 o, d; float tmin, tmax; int valid; } ray_t;

static void aabb_intersect(v3 pos, v3 dir, float lo, float hi, float* tmin_out, float* tmax_out) {
    float tmin = (lo - pos.x) / dir.x, tmax = (hi - pos.x) / dir.x;
    if (tmin > tmax) { float t = tmin; tmin = tmax; tmax = t; }
    float tymin = (lo - pos.y) / dir.y, tymax = (hi - pos.y) / dir.y;
    if (tymin > tymax) { float t = tymin; tymin = tymax; tymax = t; }
    if (tmin > tymax || tymin > tmax) { *tmin_out = FLT_MAX; *tmax_out = FLT_MAX; return; }
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    float tzmin = (lo - pos.z) / dir.z, tzmax = (hi - pos.z) / dir.z;
    if (tzmin > tzmax) { float t = tzmin; tzmin = tzmax; tzmax = t; }
    if (tmin > tzmax || tzmin > tmax) { *tmin_out = FLT_MAX; *tmax_out = FLT_MAX; return; }
    if (tzmin > tmin) tmin = tzmin;
    if (tzmax < tmax) tmax = tzmax;
    *tmin_out = tmin; *tmax_out = tmax;
}

static ray_t init_ray(const frame_consts* f, const float* ray_ori, const float* ray_dir, size_t idx) {
    ray_t r;
    v3 o = V3(ray_ori[3 * idx], ray_ori[3 * idx + 1], ray_ori[3 * idx + 2]);
    v3 d = V3(ray_dir[3 * idx], ray_dir[3 * idx + 1], ray_dir[3 * idx + 2]);
    /* mat4x3 * vec4(o,1): result[j] = 0 + m[0][j]*o.x + m[1][j]*o.y + m[2][j]*o.z + m[3][j]*1 */
    r.o = m3_mul(&f->c2w.r, o);
    r.o.x += f->c2w.t.x * 1.0f; r.o.y += f->c2w.t.y * 1.0f; r.o.z += f->c2w.t.z * 1.0f;
    r.d = m3_mul(&f->c2w.r, d);
    aabb_intersect(r.o, r.d, -1e06f, 1e06f, &r.tmin, &r.tmax); /* objectAABB, src/splatRaster.cpp:223 */
    r.tmin = fmaxf(r.tmin, 0.0f);
    r.valid = r.tmax > r.tmin;
    return r;
}

/* ------------------------------------------------------------------------------------------
 * Particle response: canonicalRayMaxKernelResponse  kernels/slang/models/gaussianParticles.slang:119-163
 *                    == particleResponse            kernels/cuda/models/gaussianParticles.cuh:263-304
 * ---------------------------------------------------------------------------------------- */
static float particle_response(int degree, float g) {
    switch (degree) {
    case 8: { const float s = -0.000685871056241f; float g2 = g * g; return expf(s * g2 * g2); }
    case 5: { const float s = -0.0185185185185f; return expf(s * g * g * sqrtf(g)); }
    case 4: { const float s = -0.0555555555556f; return expf(s * g * g); }
    case 3: { const float s = -0.166666666667f; return expf(s * g * sqrtf(g)); }
    case 1: { const float s = -1.5f; return expf(s * sqrtf(g)); }
    case 0: { const float s = -0.329630334487f; return fmaxf(1.0f + s * sqrtf(g), 0.0f); }
    default: { const float s = -0.5f; return expf(s * g); }
    }
}

/* particleResponseGrd  gaussianParticles.cuh:219-261 */
static float particle_response_grd(int degree, float g, float gres, float gres_grd) {
    switch (degree) {
    case 8: { const float s = (float)(-0.000685871056241 * (double)(0.5f * 8)); float g2 = g * g; return s * g2 * g * gres * gres_grd; }
    case 5: { const float s = (float)(-0.0185185185185 * (double)(0.5f * 5)); return s * g * sqrtf(g) * gres * gres_grd; }
    case 4: { const float s = (float)(-0.0555555555556 * (double)(0.5f * 4)); return s * g * gres * gres_grd; }
    case 3: { const float s = (float)(-0.166666666667 * (double)(0.5f * 3)); return s * sqrtf(g) * gres * gres_grd; }
    case 1: { const float s = -1.5f * (0.5f * 1); return s * sqrtf(g) * gres * gres_grd; }
    case 0: { const float s = (float)-0.329630334487; return gres > 0.0f ? (0.5f * s * (1.0f / sqrtf(g))) * gres_grd : 0.0f; }
    default: { const float s = -0.5f; return s * gres * gres_grd; }
    }
}

typedef struct { v3 gro, grd, grdu, gposc, gposcr, ray_dir_r, giscl; float g, resp, alpha; } canon_t;

/* cannonicalRay + canonicalRayMaxKernelResponse + alpha, gaussianParticles.slang:96-117,188-222
 * (same arithmetic as processHitFwd, gaussianParticles.cuh:373-387) */
static int particle_hit(const GutConfig* cfg, const raw_particle* p, const v3 rt[3], v3 o, v3 d, canon_t* c) {
    c->giscl = V3(1.0f / p->scale.x, 1.0f / p->scale.y, 1.0f / p->scale.z);
    c->gposc = sub3(o, p->pos);
    c->gposcr = V3(dot3(rt[0], c->gposc), dot3(rt[1], c->gposc), dot3(rt[2], c->gposc));
    c->gro = mul3(c->giscl, c->gposcr);
    c->ray_dir_r = V3(dot3(rt[0], d), dot3(rt[1], d), dot3(rt[2], d));
    c->grdu = mul3(c->giscl, c->ray_dir_r);
    float l2 = dot3(c->grdu, c->grdu);
    float il = 1.0f / sqrtf(l2); /* normalize / safe_normalize(rsqrtf), mathUtils.cuh:390-393 */
    c->grd = l2 > 0.0f ? scl3(il, c->grdu) : c->grdu;
    v3 gc = cross3(c->grd, c->gro);
    c->g = dot3(gc, gc);
    c->resp = particle_response(cfg->kernel_degree, c->g);
    c->alpha = fminf(cfg->max_alpha, c->resp * p->density);
    return (c->resp > cfg->min_kernel_response) && (c->alpha > cfg->min_alpha);
}

/* canonicalRayDistance, gaussianParticles.slang:165-171 */
static float canon_distance(const raw_particle* p, const canon_t* c) {
    float dp = dot3(c->grd, scl3(-1.0f, c->gro));
    v3 grds = scl3(dp, mul3(p->scale, c->grd));
    return sqrtf(dot3(grds, grds));
}

#define MAX_K 64
typedef struct { int idx; float t, alpha; } hit_t;

/* ------------------------------------------------------------------------------------------
 * K4: render  kernels/cuda/renderers/gutRenderer.cuh:83-115
 *     GUTKBufferRenderer::evalKBuffer  gutKBufferRenderer.cuh:217-292, processHitParticle fwd :153-170,
 *     HitParticleKBuffer :28-85, finalizeRay rayPayload.cuh:110-129
 * Outputs must be pre-initialised by the caller like splatRaster.cpp:196-198 (rgba 0, dist 1e6, hits 0).
 * tests_out / (optional) counts ray x particle tests for the roofline accounting.
 * ---------------------------------------------------------------------------------------- */
ORACLE_API void gut_oracle_render(const GutConfig* cfg, const GutCamera* cam, uint32_t N, uint32_t W, uint32_t H,
                                  const float* density12, const float* features /* [N,3] unclamped */,
                                  const float* ray_ori, const float* ray_dir,
                                  const uint32_t* tile_ranges, const uint32_t* sorted_idx,
                                  float* rgba, float* dist, float* hits, uint64_t* tests_out,
                                  const uint32_t* tile_list /* optional: render only these tiles (bounded CPU-baseline sample) */,
                                  uint32_t n_tile_list,
                                  float* margin /* optional [H,W]: smallest relative distance of any accept/kill decision of the
                                                   ray to its threshold -- pixels with a tiny margin may legitimately differ between two
                                                   fp32 implementations by one whole hit; the parity tests treat them separately */) {
    (void)N;
    const frame_consts f = make_frame(cam, W, H);
    const int K = cfg->k_buffer_size;
    uint64_t tests = 0;
    const int64_t n_work = tile_list ? (int64_t)n_tile_list : (int64_t)f.gx * f.gy;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : tests)
    for (int64_t work = 0; work < n_work; ++work) {
        const int64_t tile = tile_list ? (int64_t)tile_list[work] : work;
        const uint32_t tx = (uint32_t)(tile % f.gx), ty = (uint32_t)(tile / f.gx);
        const uint32_t r0 = tile_ranges[2 * tile], r1 = tile_ranges[2 * tile + 1];
        for (uint32_t py = ty * TILE; py < ty * TILE + TILE && py < H; ++py)
            for (uint32_t px = tx * TILE; px < tx * TILE + TILE && px < W; ++px) {
                const size_t idx = (size_t)px + (size_t)W * py;
                ray_t ray = init_ray(&f, ray_ori, ray_dir, idx);
                if (!ray.valid) continue;
                float T = 1.0f, D = 0.0f; v3 C = {0.0f, 0.0f, 0.0f}; uint32_t n = 0; int alive = 1;
                float mrg = FLT_MAX;
                hit_t kb[MAX_K]; int nk = 0;
                for (int k = 0; k < K; ++k) { kb[k].idx = -1; kb[k].t = -1.0f; kb[k].alpha = 0.0f; }
#define PROCESS_HIT(HT)                                                                                         \
    do {                                                                                                       \
        const hit_t* h_ = (HT);                                                                                \
        float w_ = h_->alpha * T; /* integrateHit<false> gaussianParticles.slang:224-254 */                   \
        D += h_->t * w_;                                                                                       \
        T *= (1.0f - h_->alpha);                                                                               \
        if (w_ > 0.0f) { /* integrateRadiance<false> shRadiativeParticles.slang:83-99 */                      \
            const float* ft_ = features + 3 * (size_t)h_->idx;                                                 \
            C.x += fmaxf(ft_[0], 0.0f) * w_; C.y += fmaxf(ft_[1], 0.0f) * w_; C.z += fmaxf(ft_[2], 0.0f) * w_; \
            n++;                                                                                               \
        }                                                                                                      \
        mrg = fminf(mrg, fabsf(T - cfg->min_transmittance) / cfg->min_transmittance);                          \
        if (T < cfg->min_transmittance) alive = 0;                                                             \
    } while (0)
                for (uint32_t s = r0; alive && s < r1; ++s) {
                    const uint32_t pi = sorted_idx[s];
                    if (pi == 0xFFFFFFFFu) break;
                    const raw_particle* p = (const raw_particle*)(density12 + 12 * (size_t)pi);
                    v3 rt[3]; particle_rotT(p, rt);
                    canon_t c; tests++;
                    const int accepted = particle_hit(cfg, p, rt, ray.o, ray.d, &c);
                    mrg = fminf(mrg, fabsf(c.resp - cfg->min_kernel_response) / cfg->min_kernel_response);
                    mrg = fminf(mrg, fabsf(c.resp * p->density - cfg->min_alpha) / cfg->min_alpha);
                    if (!accepted) continue;
                    hit_t h = {(int)pi, canon_distance(p, &c), c.alpha};
                    if (!(h.t > ray.tmin && h.t < ray.tmax)) continue;
                    if (K == 0) { PROCESS_HIT(&h); continue; }
                    if (nk == K) { hit_t closest = kb[0]; PROCESS_HIT(&closest); kb[0].t = -1.0f; } else nk++;
                    /* a near tie in hit distance may be ordered the other way by another fp32 implementation: distances agree to
                       ~1e-6, and libgut_b200 orders buffered hits on the distance bits without the low 4 mantissa bits (1.9e-6) */
                    for (int i = 0; i < K; ++i)
                        if (kb[i].t >= 0.0f) mrg = fminf(mrg, 5.0f * fabsf(h.t - kb[i].t) / fmaxf(fabsf(h.t), 1e-30f)); /* flags |dt|/t < 4e-6 */
                    for (int i = K - 1; i >= 0; --i)
                        if (h.t > kb[i].t) { hit_t tmp = kb[i]; kb[i] = h; h = tmp; }
                }
                if (K > 0)
                    for (int i = 0; alive && i < nk; ++i) PROCESS_HIT(&kb[K - nk + i]);
#undef PROCESS_HIT
                rgba[4 * idx] = C.x; rgba[4 * idx + 1] = C.y; rgba[4 * idx + 2] = C.z; rgba[4 * idx + 3] = 1.0f - T;
                dist[idx] = D;
                if (cfg->enable_hitcounts) hits[idx] = (float)n;
                if (margin) margin[idx] = mrg;
            }
    }
    if (tests_out) *tests_out = tests;
}

/* ------------------------------------------------------------------------------------------
 * K5 (K = 0): renderBackward  gutRenderer.cuh:117-165 -> evalBackwardNoKBuffer gutKBufferRenderer.cuh:294-386
 *   per-hit math: processHitBwd<degree,false,false>  kernels/cuda/models/gaussianParticles.cuh:480-738
 *   helpers: safe_normalize_bw mathUtils.cuh:420-430, matmul_bw_vec :461-466, matmul_bw_quat :468-533
 *   ray state: initializeBackwardRay rayPayloadBackward.cuh:30-58
 * The reference accumulates with fp32 atomics in arbitrary order; the oracle accumulates in fp64 and the
 * caller narrows, so it is deterministic.  density_grad [N,12] (col 11 untouched), feature_grad [N,3].
 * ---------------------------------------------------------------------------------------- */
static v3 safe_normalize_bw(v3 v, v3 d_out) {
    float l = v.x * v.x + v.y * v.y + v.z * v.z;
    if (l > 0.0f) {
        float il = 1.0f / sqrtf(l);
        float il3 = il * il * il;
        v3 t = V3(d_out.x * (v.x * v.x) + d_out.y * (v.y * v.x) + d_out.z * (v.z * v.x),
                  d_out.x * (v.x * v.y) + d_out.y * (v.y * v.y) + d_out.z * (v.z * v.y),
                  d_out.x * (v.x * v.z) + d_out.y * (v.y * v.z) + d_out.z * (v.z * v.z));
        return sub3(scl3(il, d_out), scl3(il3, t));
    }
    return V3(0.0f, 0.0f, 0.0f);
}

static v4 matmul_bw_quat(v3 p, v3 g, float r, float x, float y, float z) {
    v3 d0 = scl3(g.x, p), d1 = scl3(g.y, p), d2 = scl3(g.z, p);
    float dr = 0, dx = 0, dy = 0, dz = 0;
    dy += -4 * y * d0.x; dz += -4 * z * d0.x;
    dr += 2 * z * d0.y; dx += 2 * y * d0.y; dy += 2 * x * d0.y; dz += 2 * r * d0.y;
    dr += -2 * y * d0.z; dx += 2 * z * d0.z; dy += -2 * r * d0.z; dz += 2 * x * d0.z;
    dr += -2 * z * d1.x; dx += 2 * y * d1.x; dy += 2 * x * d1.x; dz += -2 * r * d1.x;
    dx += -4 * x * d1.y; dz += -4 * z * d1.y;
    dr += 2 * x * d1.z; dx += 2 * r * d1.z; dy += 2 * z * d1.z; dz += 2 * y * d1.z;
    dr += 2 * y * d2.x; dx += 2 * z * d2.x; dy += 2 * r * d2.x; dz += 2 * x * d2.x;
    dr += -2 * x * d2.y; dx += -2 * r * d2.y; dy += 2 * z * d2.y; dz += 2 * y * d2.y;
    dx += -4 * x * d2.z; dy += -4 * y * d2.z;
    v4 o = {dr, dx, dy, dz};
    return o;
}

ORACLE_API void gut_oracle_render_bwd(const GutConfig* cfg, const GutCamera* cam, uint32_t N, uint32_t W, uint32_t H,
                                      const float* density12, const float* features,
                                      const float* ray_ori, const float* ray_dir,
                                      const uint32_t* tile_ranges, const uint32_t* sorted_idx,
                                      const float* rgba, const float* rgba_grad, const float* dist, const float* dist_grad,
                                      double* density_grad /* [N,12] += */, double* feature_grad /* [N,3] += */,
                                      const uint32_t* tile_list, uint32_t n_tile_list) {
    (void)N;
    const frame_consts f = make_frame(cam, W, H);
    const float minT = cfg->min_transmittance;
    const int64_t n_work = tile_list ? (int64_t)n_tile_list : (int64_t)f.gx * f.gy;
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t work = 0; work < n_work; ++work) {
        const int64_t tile = tile_list ? (int64_t)tile_list[work] : work;
        const uint32_t tx = (uint32_t)(tile % f.gx), ty = (uint32_t)(tile / f.gx);
        const uint32_t r0 = tile_ranges[2 * tile], r1 = tile_ranges[2 * tile + 1];
        for (uint32_t py = ty * TILE; py < ty * TILE + TILE && py < H; ++py)
            for (uint32_t px = tx * TILE; px < tx * TILE + TILE && px < W; ++px) {
                const size_t idx = (size_t)px + (size_t)W * py;
                ray_t ray = init_ray(&f, ray_ori, ray_dir, idx);
                if (!ray.valid) continue;
                const float T_f = 1.0f - rgba[4 * idx + 3];           /* transmittanceBackward */
                const float T_grad = -1.0f * rgba_grad[4 * idx + 3];  /* transmittanceGradient */
                const float D_f = dist[idx], D_grad = dist_grad[idx];
                const v3 C_f = V3(rgba[4 * idx], rgba[4 * idx + 1], rgba[4 * idx + 2]);
                const v3 C_grad = V3(rgba_grad[4 * idx], rgba_grad[4 * idx + 1], rgba_grad[4 * idx + 2]);
                float T = 1.0f, D = 0.0f; v3 C = {0.0f, 0.0f, 0.0f};
                for (uint32_t s = r0; s < r1; ++s) {
                    const uint32_t pi = sorted_idx[s];
                    if (pi == 0xFFFFFFFFu) break;
                    const raw_particle* p = (const raw_particle*)(density12 + 12 * (size_t)pi);
                    v3 rt[3]; particle_rotT(p, rt);
                    canon_t c;
                    if (!particle_hit(cfg, p, rt, ray.o, ray.d, &c)) continue; /* no tMinMax test in the K=0 backward */
                    const v3 gscl = p->scale;
                    const float galpha = c.alpha, gres = c.resp;
                    const v3 grd = c.grd, gro = c.gro;
                    const v3 feat = V3(fmaxf(features[3 * (size_t)pi], 0.0f), fmaxf(features[3 * (size_t)pi + 1], 0.0f),
                                       fmaxf(features[3 * (size_t)pi + 2], 0.0f)); /* staged clamped, gutKBufferRenderer.cuh:324 */

                    const v3 grdd = scl3(dot3(grd, scl3(-1.0f, gro)), grd);
                    const v3 grds = mul3(gscl, grdd);
                    const float gsqdist = dot3(grds, grds);
                    const float gdist = sqrtf(gsqdist);
                    const float weight = galpha * T;
                    const float nextT = (1 - galpha) * T;

                    /* CALL-SITE QUIRK (found by the reference + shim build, tests/test_gpu_reference.py): the wrapper
                     * shRadiativeGaussianParticles.cuh:369-406 hands processHitBwd its last three arguments as
                     * (hitT, hitTBackward, hitTGradient) where the callee (gaussianParticles.cuh:499-501) declares
                     * (integratedDepth, depth&, depthGrad): `integratedDepth` is the ray's RUNNING distance ray.hitT -- 0 from
                     * initializeRay (rayPayload.cuh:88) and never updated, because `depth&` binds to the wrapper's by-value
                     * copy of the forward's final distance.  So per call: depth = D_f + weight * gdist (discarded afterwards),
                     * integratedDepth = 0, and the residual is max(-(D_f + w gdist) / nextT, 0) = 0 for every non-negative
                     * forward distance.  The hit-distance part of the alpha gradient is therefore gdist * T * dD. */
                    const float integratedDepth = 0.0f;          /* ray.hitT */
                    float depth_local = D_f;                       /* by-value copy of ray.hitTBackward */
                    depth_local += weight * gdist;                 /* :542 */
                    D += weight * gdist;                           /* what the callee's author meant; kept for reference only */
                    const float resD = fmaxf((nextT <= minT ? 0 : (integratedDepth - depth_local) / nextT), 0);   /* :543-545 */
                    const float galphaRayHitGrd = (gdist - resD) * T * D_grad;
                    const v3 grdsRayHitGrd = gsqdist > 0.0f
                        ? scl3(D_grad, V3((2 * grds.x * weight) / (2 * gdist), (2 * grds.y * weight) / (2 * gdist), (2 * grds.z * weight) / (2 * gdist)))
                        : V3(0.0f, 0.0f, 0.0f);
                    const v3 gsclRayHitGrd = mul3(grdd, grdsRayHitGrd);
                    const v3 grdRayHitGrd = mul3(mul3(V3(-gscl.x, -gscl.y, -gscl.z),
                                                      V3(2 * grd.x * gro.x + grd.y * gro.y + grd.z * gro.z,
                                                         grd.x * gro.x + 2 * grd.y * gro.y + grd.z * gro.z,
                                                         grd.x * gro.x + grd.y * gro.y + 2 * grd.z * gro.z)), grdsRayHitGrd);
                    const v3 groRayHitGrd = mul3(mul3(mul3(V3(-gscl.x, -gscl.y, -gscl.z), grd), grd), grdsRayHitGrd);

                    const float resT = galpha < 0.999999f ? T_f / (1 - galpha) : T;
                    const float galphaRayDnsGrd = resT * -T_grad; /* transmittanceGrad argument is -dA already: see below */
                    /* NB: processHitBwd receives transmittanceGrad = ray.transmittanceGradient = -dA and computes
                     * residualTrm * -transmittanceGrad (gaussianParticles.cuh:574). T_grad here == -dA. */

                    /* feature gradient: particleRadianceGradPtr[k] = radianceGrad[k] * weight (:590-592) */
                    const v3 fgrad = scl3(weight, C_grad);
                    const v3 rayRad = scl3(weight, feat);
                    C = add3(C, rayRad);
                    v3 resC;
                    if (nextT <= minT) resC = V3(0.0f, 0.0f, 0.0f);
                    else resC = V3(fmaxf((C_f.x - C.x) / nextT, 0.0f), fmaxf((C_f.y - C.y) / nextT, 0.0f), fmaxf((C_f.z - C.z) / nextT, 0.0f));

                    const float S = galphaRayHitGrd + galphaRayDnsGrd + T * (feat.x - resC.x) * C_grad.x +
                                    T * (feat.y - resC.y) * C_grad.y + T * (feat.z - resC.z) * C_grad.z;
                    const float dns_grad = gres * S;
                    const float gresGrd = p->density * S;
                    const float grayDistGrd = particle_response_grd(cfg->kernel_degree, c.g, gres, gresGrd);

                    const v3 gcrod = cross3(grd, gro);
                    const v3 gcrodGrd = scl3(grayDistGrd, scl3(2.0f, gcrod));
                    const v3 grdGrd = V3(gcrodGrd.z * gro.y - gcrodGrd.y * gro.z, gcrodGrd.x * gro.z - gcrodGrd.z * gro.x,
                                         gcrodGrd.y * gro.x - gcrodGrd.x * gro.y);
                    const v3 groGrd = V3(gcrodGrd.y * grd.z - gcrodGrd.z * grd.y, gcrodGrd.z * grd.x - gcrodGrd.x * grd.z,
                                         gcrodGrd.x * grd.y - gcrodGrd.y * grd.x);

                    const v3 groTot = add3(groGrd, groRayHitGrd);
                    const v3 gsclGrdGro = mul3(V3(-c.gposcr.x / (gscl.x * gscl.x), -c.gposcr.y / (gscl.y * gscl.y), -c.gposcr.z / (gscl.z * gscl.z)), groTot);
                    const v3 gposcrGrd = mul3(c.giscl, groTot);
                    /* matmul_bw_vec(R^T rows, g): g.x*m[0] + g.y*m[1] + g.z*m[2] */
                    const v3 gposcGrd = V3(gposcrGrd.x * rt[0].x + gposcrGrd.y * rt[1].x + gposcrGrd.z * rt[2].x,
                                           gposcrGrd.x * rt[0].y + gposcrGrd.y * rt[1].y + gposcrGrd.z * rt[2].y,
                                           gposcrGrd.x * rt[0].z + gposcrGrd.y * rt[1].z + gposcrGrd.z * rt[2].z);
                    const v4 grotGrdPoscr = matmul_bw_quat(c.gposc, gposcrGrd, p->qw, p->qx, p->qy, p->qz);
                    const v3 pos_grad = V3(-gposcGrd.x, -gposcGrd.y, -gposcGrd.z);

                    const v3 grduGrd = safe_normalize_bw(c.grdu, add3(grdGrd, grdRayHitGrd));
                    const v3 scl_grad = add3(add3(gsclRayHitGrd, gsclGrdGro),
                                             mul3(V3(-c.ray_dir_r.x / (gscl.x * gscl.x), -c.ray_dir_r.y / (gscl.y * gscl.y), -c.ray_dir_r.z / (gscl.z * gscl.z)), grduGrd));
                    const v3 rayDirRGrd = mul3(c.giscl, grduGrd);
                    const v4 grotGrdRayDirR = matmul_bw_quat(ray.d, rayDirRGrd, p->qw, p->qx, p->qy, p->qz);

                    double* dg = density_grad + 12 * (size_t)pi;
                    double* fg = feature_grad + 3 * (size_t)pi;
                    const float vals[11] = {pos_grad.x, pos_grad.y, pos_grad.z, dns_grad,
                                            grotGrdPoscr.x + grotGrdRayDirR.x, grotGrdPoscr.y + grotGrdRayDirR.y,
                                            grotGrdPoscr.z + grotGrdRayDirR.z, grotGrdPoscr.w + grotGrdRayDirR.w,
                                            scl_grad.x, scl_grad.y, scl_grad.z};
                    for (int k = 0; k < 11; ++k) {
#pragma omp atomic
                        dg[k] += (double)vals[k];
                    }
#pragma omp atomic
                    fg[0] += (double)fgrad.x;
#pragma omp atomic
                    fg[1] += (double)fgrad.y;
#pragma omp atomic
                    fg[2] += (double)fgrad.z;

                    T = nextT;
                    if (T < minT) break; /* ray.kill(), gutKBufferRenderer.cuh:374-376 */
                }
            }
    }
}

/* ------------------------------------------------------------------------------------------
 * K5 (K > 0): renderBackward -> evalKBuffer<Backward>  gutKBufferRenderer.cuh:217-292 with the backward branch of
 *   processHitParticle :116-152.  The hits leave the k-buffer in the forward's order; each one is "un-blended" from the
 *   ray's final colour / distance / transmittance (back-to-front formulation walked front to back):
 *     particleFeaturesIntegrateBwd          kernels/slang/models/shRadiativeParticles.slang:179-207
 *       (bwd_diff of integrateRadiance<true> = lerp, :83-99; NB it is handed the UNclamped particle feature,
 *        gutKBufferRenderer.cuh:131-136, while the forward composited max(feature, 0), :162)
 *     particleDensityProcessHitBwdToBuffer  kernels/slang/models/gaussianParticles.slang:394-451
 *       (bwd_diff of processHitFromBuffer<true> :254-283 = hit :188-222 + integrateHit<true> :224-252): Slang autodiff,
 *        i.e. the TRUE derivative of alpha = min(MaxAlpha, response * density) (no gradient when clamped) and of the hit
 *        distance |scale * grd * dot(grd, -gro)| -- unlike the hand-written K = 0 backward above.
 *   Gradients are added with one atomic per value per hit (gaussianParticles.slang:61-75); fp64 accumulation here.
 * ---------------------------------------------------------------------------------------- */
typedef struct { float Tb, Tg, Db, Dg; v3 Fb, Fg; } bwd_state;

static void kbuffer_hit_bwd(const GutConfig* cfg, const float* density12, const float* features, const ray_t* ray, const hit_t* h,
                            bwd_state* st, double* density_grad, double* feature_grad) {
    const size_t pi = (size_t)h->idx;
    const float alpha = h->alpha, t = h->t;
    float alpha_grad = 0.0f;
    v3 fgrad = V3(0.0f, 0.0f, 0.0f);
    if (alpha > 0.0f) { /* particleFeaturesIntegrateBwd */
        const v3 feat = V3(features[3 * pi], features[3 * pi + 1], features[3 * pi + 2]);
        const float w = 1.0f / (1.0f - alpha);
        st->Fb = scl3(w, sub3(st->Fb, scl3(alpha, feat)));
        /* d lerp(I, f, a): da = dI . (f - I), df = dI a, dI_in = dI (1 - a) */
        alpha_grad = st->Fg.x * (feat.x - st->Fb.x) + st->Fg.y * (feat.y - st->Fb.y) + st->Fg.z * (feat.z - st->Fb.z);
        fgrad = scl3(alpha, st->Fg);
        st->Fg = scl3(1.0f - alpha, st->Fg);
    }
    double* fg = feature_grad + 3 * pi;
#pragma omp atomic
    fg[0] += (double)fgrad.x;
#pragma omp atomic
    fg[1] += (double)fgrad.y;
#pragma omp atomic
    fg[2] += (double)fgrad.z;
    if (!(alpha > 0.0f)) return;

    /* particleDensityProcessHitBwdToBuffer */
    const float w = 1.0f / (1.0f - alpha);
    st->Tb *= w;
    st->Db = (st->Db - t * alpha) * w;
    const raw_particle* p = (const raw_particle*)(density12 + 12 * pi);
    v3 rt[3]; particle_rotT(p, rt);
    canon_t c;
    if (!particle_hit(cfg, p, rt, ray->o, ray->d, &c)) return; /* processHitFromBuffer re-evaluates hit(): same result as the forward */
    const float depth = canon_distance(p, &c);
    /* integrateHit<true>: D_out = lerp(D_in, depth, alpha), T_out = T_in (1 - alpha), returns alpha */
    const float d_alpha = alpha_grad + st->Dg * (depth - st->Db) - st->Tg * st->Tb;
    const float d_depth = st->Dg * c.alpha;
    st->Dg *= (1.0f - c.alpha);
    st->Tg *= (1.0f - c.alpha);
    /* hit(): alpha = min(MaxAlpha, resp * density) -- no gradient when the clamp is active */
    const float araw = c.resp * p->density;
    const int unclamped = !(cfg->max_alpha < araw);
    const float d_resp = unclamped ? d_alpha * p->density : 0.0f;
    const float d_dns = unclamped ? d_alpha * c.resp : 0.0f;
    const float d_g = particle_response_grd(cfg->kernel_degree, c.g, c.resp, d_resp);
    /* g = |grd x gro|^2 */
    const v3 gc = cross3(c.grd, c.gro);
    const v3 gcg = scl3(2.0f * d_g, gc);
    v3 d_grd = cross3(c.gro, gcg);
    v3 d_gro = cross3(gcg, c.grd);
    /* depth = |grds|, grds = scale * grd * dp, dp = dot(grd, -gro) */
    const float dp = dot3(c.grd, scl3(-1.0f, c.gro));
    const v3 grds = scl3(dp, mul3(p->scale, c.grd));
    v3 d_scl = V3(0.0f, 0.0f, 0.0f);
    if (depth > 0.0f) {
        const v3 d_grds = scl3(d_depth / depth, grds);
        d_scl = scl3(dp, mul3(d_grds, c.grd));
        d_grd = add3(d_grd, scl3(dp, mul3(d_grds, p->scale)));
        const float d_dp = dot3(d_grds, mul3(p->scale, c.grd));
        d_grd = sub3(d_grd, scl3(d_dp, c.gro));
        d_gro = sub3(d_gro, scl3(d_dp, c.grd));
    }
    /* cannonicalRay: grd = normalize(giscl * R^T d), gro = giscl * R^T (o - pos) */
    const v3 d_grdu = safe_normalize_bw(c.grdu, d_grd);
    const v3 d_rdr = mul3(c.giscl, d_grdu);
    const v3 d_gposcr = mul3(c.giscl, d_gro);
    const v3 d_giscl = add3(mul3(c.ray_dir_r, d_grdu), mul3(c.gposcr, d_gro));
    d_scl = sub3(d_scl, V3(d_giscl.x / (p->scale.x * p->scale.x), d_giscl.y / (p->scale.y * p->scale.y), d_giscl.z / (p->scale.z * p->scale.z)));
    const v3 d_gposc = V3(d_gposcr.x * rt[0].x + d_gposcr.y * rt[1].x + d_gposcr.z * rt[2].x,
                          d_gposcr.x * rt[0].y + d_gposcr.y * rt[1].y + d_gposcr.z * rt[2].y,
                          d_gposcr.x * rt[0].z + d_gposcr.y * rt[1].z + d_gposcr.z * rt[2].z);
    const v4 q1 = matmul_bw_quat(c.gposc, d_gposcr, p->qw, p->qx, p->qy, p->qz);
    const v4 q2 = matmul_bw_quat(ray->d, d_rdr, p->qw, p->qx, p->qy, p->qz);
    const float vals[11] = {-d_gposc.x, -d_gposc.y, -d_gposc.z, d_dns, q1.x + q2.x, q1.y + q2.y, q1.z + q2.z, q1.w + q2.w,
                            d_scl.x, d_scl.y, d_scl.z};
    double* dg = density_grad + 12 * pi;
    for (int k = 0; k < 11; ++k) {
#pragma omp atomic
        dg[k] += (double)vals[k];
    }
}

ORACLE_API void gut_oracle_render_bwd_kbuffer(const GutConfig* cfg, const GutCamera* cam, uint32_t N, uint32_t W, uint32_t H,
                                              const float* density12, const float* features,
                                              const float* ray_ori, const float* ray_dir,
                                              const uint32_t* tile_ranges, const uint32_t* sorted_idx,
                                              const float* rgba, const float* rgba_grad, const float* dist, const float* dist_grad,
                                              double* density_grad /* [N,12] += */, double* feature_grad /* [N,3] += */,
                                              const uint32_t* tile_list, uint32_t n_tile_list) {
    (void)N;
    const frame_consts f = make_frame(cam, W, H);
    const int K = cfg->k_buffer_size;
    const float minT = cfg->min_transmittance;
    const int64_t n_work = tile_list ? (int64_t)n_tile_list : (int64_t)f.gx * f.gy;
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t work = 0; work < n_work; ++work) {
        const int64_t tile = tile_list ? (int64_t)tile_list[work] : work;
        const uint32_t tx = (uint32_t)(tile % f.gx), ty = (uint32_t)(tile / f.gx);
        const uint32_t r0 = tile_ranges[2 * tile], r1 = tile_ranges[2 * tile + 1];
        for (uint32_t py = ty * TILE; py < ty * TILE + TILE && py < H; ++py)
            for (uint32_t px = tx * TILE; px < tx * TILE + TILE && px < W; ++px) {
                const size_t idx = (size_t)px + (size_t)W * py;
                ray_t ray = init_ray(&f, ray_ori, ray_dir, idx);
                if (!ray.valid) continue;
                bwd_state st; /* initializeBackwardRay rayPayloadBackward.cuh:46-55 */
                st.Tb = 1.0f - rgba[4 * idx + 3];
                st.Tg = -1.0f * rgba_grad[4 * idx + 3];
                st.Db = dist[idx]; st.Dg = dist_grad[idx];
                st.Fb = V3(rgba[4 * idx], rgba[4 * idx + 1], rgba[4 * idx + 2]);
                st.Fg = V3(rgba_grad[4 * idx], rgba_grad[4 * idx + 1], rgba_grad[4 * idx + 2]);
                float T = 1.0f; int alive = 1;
                hit_t kb[MAX_K]; int nk = 0;
                for (int k = 0; k < K; ++k) { kb[k].idx = -1; kb[k].t = -1.0f; kb[k].alpha = 0.0f; }
#define PROCESS_HIT_BWD(HT)                                                                    \
    do {                                                                                      \
        const hit_t* h_ = (HT);                                                               \
        kbuffer_hit_bwd(cfg, density12, features, &ray, h_, &st, density_grad, feature_grad); \
        T *= (1.0f - h_->alpha);                                                              \
        if (T < minT) alive = 0;                                                              \
    } while (0)
                for (uint32_t s = r0; alive && s < r1; ++s) {
                    const uint32_t pi = sorted_idx[s];
                    if (pi == 0xFFFFFFFFu) break;
                    const raw_particle* p = (const raw_particle*)(density12 + 12 * (size_t)pi);
                    v3 rt[3]; particle_rotT(p, rt);
                    canon_t c;
                    if (!particle_hit(cfg, p, rt, ray.o, ray.d, &c)) continue;
                    hit_t h = {(int)pi, canon_distance(p, &c), c.alpha};
                    if (!(h.t > ray.tmin && h.t < ray.tmax)) continue;
                    if (K == 0) { PROCESS_HIT_BWD(&h); continue; }
                    if (nk == K) { hit_t closest = kb[0]; PROCESS_HIT_BWD(&closest); kb[0].t = -1.0f; } else nk++;
                    for (int i = K - 1; i >= 0; --i)
                        if (h.t > kb[i].t) { hit_t tmp = kb[i]; kb[i] = h; h = tmp; }
                }
                if (K > 0)
                    for (int i = 0; alive && i < nk; ++i) PROCESS_HIT_BWD(&kb[K - nk + i]);
#undef PROCESS_HIT_BWD
            }
    }
}

/* ------------------------------------------------------------------------------------------
 * K6: GUTProjector::evalBackward gutProjector.cuh:390-430 ; radianceFromSpHBwd gaussianParticles.cuh:120-187
 * sph_grad [N,ncoef,3] += ; the view-direction -> position term is a no-op (gaussianParticles.slang:307-318).
 * ---------------------------------------------------------------------------------------- */
ORACLE_API void gut_oracle_project_bwd(const GutConfig* cfg, const GutCamera* cam, uint32_t N, int sph_degree,
                                       uint32_t W, uint32_t H, const float* density12, const uint32_t* tiles_count,
                                       const float* features, const float* feature_grad, float* sph_grad) {
    const frame_consts f = make_frame(cam, W, H);
    const int ncoef = (cfg->sph_max_degree + 1) * (cfg->sph_max_degree + 1);
    const int nact = (sph_degree + 1) * (sph_degree + 1);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)N; ++i) {
        if (tiles_count[i] == 0) continue;
        const raw_particle* p = (const raw_particle*)(density12 + 12 * i);
        v3 v = sub3(p->pos, f.cam_centre);
        float len = sqrtf(dot3(v, v));
        v3 dir = len <= 0.0f ? V3(1.0f, 0.0f, 0.0f) : V3(v.x / len, v.y / len, v.z / len); /* tcnn::normalize vec.h:466-474 */
        float b[16]; sh_basis(sph_degree, dir, b);
        v3 g = V3(feature_grad[3 * i], feature_grad[3 * i + 1], feature_grad[3 * i + 2]);
        g.x *= (features[3 * i] > 0.0f ? 1 : 0); g.y *= (features[3 * i + 1] > 0.0f ? 1 : 0); g.z *= (features[3 * i + 2] > 0.0f ? 1 : 0);
        float* out = sph_grad + (size_t)i * ncoef * 3;
        for (int k = 0; k < nact; ++k) { out[3 * k] += b[k] * g.x; out[3 * k + 1] += b[k] * g.y; out[3 * k + 2] += b[k] * g.z; }
    }
}

/* ------------------------------------------------------------------------------------------
 * SelectiveAdam: selective_adam_update_kernel, threedgrut/optimizers/optimizers.cu:47-84, one scalar per
 * thread.  PINNED on the reference's own kernel: the rounding below is that of the reference's COMPILED kernel (default
 * nvcc flags, setup_optimizers.py:27-44: FMA contraction on, IEEE sqrt and division), read from its SASS
 * (oracle/_ref/optimizers_cu.o):   FMUL t = m * b1 ; FADD c = 1 - b1 ; FFMA m' = g * c + t
 *                                  FMUL u = v * b2 ; FMUL w = g * (1 - b2) ; FFMA v' = g * w + u
 *                                  FMUL s = m' * lr ; sqrt_rn ; FADD d = sqrt + eps ; div_rn ; FADD param - s / d
 * and tests/test_gpu_model.py compares it bit for bit with that kernel on the GPU box; tests/golden/reference_adam.npz is
 * the same comparison frozen for the CPU suite. */
ORACLE_API void gut_oracle_selective_adam(uint64_t n_rows, uint32_t width, float* param, const float* grad, float* exp_avg,
                                          float* exp_avg_sq, const uint8_t* visibility /* per row, may be NULL */,
                                          float lr, float b1, float b2, float eps) {
    for (uint64_t r = 0; r < n_rows; ++r) {
        if (visibility && !visibility[r]) continue;
        for (uint32_t c = 0; c < width; ++c) {
            const uint64_t i = r * width + c;
            const float g = grad[i];
            const float m = fmaf(g, 1.0f - b1, b1 * exp_avg[i]);
            const float v = fmaf(g, g * (1.0f - b2), b2 * exp_avg_sq[i]);
            const float step = (lr * m) / (sqrtf(v) + eps);
            param[i] = param[i] - step;
            exp_avg[i] = m;
            exp_avg_sq[i] = v;
        }
    }
}

/* Host threads the OpenMP loops of this library use.  torchrun exports OMP_NUM_THREADS=1 to its workers, so bench.py's
 * reference arm sets the count explicitly and reports what the library really used (gut_oracle_threads_used runs a
 * parallel region and counts). */
ORACLE_API void gut_oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
ORACLE_API int gut_oracle_threads_used(void) {
    int n = 1;
#ifdef _OPENMP
#pragma omp parallel
    {
#pragma omp single
        n = omp_get_num_threads();
    }
#endif
    return n;
}
