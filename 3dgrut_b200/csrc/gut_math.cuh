// gut_math.cuh -- small fp32 vector helpers and the *deterministic* scalar math shared by the
// projection / binning kernels of libgut_b200.
//
// Two numeric regimes live in this library (DESIGN.md "Numerics"):
//   * projection + binning (gut_project.cu, gut_binning.cu) are compiled with --fmad=false and use
//     only IEEE-rounded +,*,/,sqrt plus explicit fmaf, so tile counts / keys / ranges are
//     bit-reproducible against the CPU oracle (integer results must be exact);
//   * compositing (gut_render.cu) is free to contract FMAs and use MUFU approximations, like the
//     reference which is built with -use_fast_math (threedgut_tracer/setup_3dgut.py:83).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define GUT_TILE 16  // reference: GUTParameters::Tiling::BlockX/BlockY (renderer/gutRendererParameters.h:22-23)
#define GUT_INVALID_PARTICLE 0xFFFFFFFFu
// Every tile's bucket counter / cursor is split into GUT_TILE_SUB sub-counters selected by the low bits of the particle
// index: the returning L2 atomics of the emit pass then spread over 8x more addresses (same-address atomics serialise).
#define GUT_TILE_SUB 16
// sub-counter k of tile t (T tiles): sub-counter major, so that the sub-counters of one hot tile sit in different L2 sectors
// (atomics to one 32-byte sector serialise just like atomics to one address)
#define GUT_SUB_SLOT(t, k, T) ((size_t)(k) * (T) + (t))

#define GUT_HD __host__ __device__ __forceinline__
#define GUT_D __device__ __forceinline__

struct GutMat4x3 {  // column-major: c[0..2] rotation columns, c[3] translation (tcnn::mat4x3)
    float3 c[4];
};

// compact colour-gradient slot (gut_project.cu, colour_store): words of the (mask, first row) table, a multiple of four
GUT_HD uint32_t gut_colour_table_words(uint32_t N) { return (2u * ((N + 31u) >> 5) + 3u) & ~3u; }
GUT_HD float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
GUT_HD float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
GUT_HD float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
GUT_HD float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
GUT_HD float3 operator*(float s, float3 a) { return f3(s * a.x, s * a.y, s * a.z); }
GUT_HD float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
GUT_HD float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
GUT_HD float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
GUT_HD float3 cross(float3 a, float3 b) {
    return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

// rotation-matrix columns of a unit quaternion (w,x,y,z); column i = R e_i.  (tcnn::to_mat3, vec.h:1172-1183)
GUT_HD void quat_to_columns(float w, float x, float y, float z, float3& c0, float3& c1, float3& c2) {
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, xz = x * z, yz = y * z, wx = w * x, wy = w * y, wz = w * z;
    c0 = f3(1.0f - 2.0f * (yy + zz), 2.0f * (xy + wz), 2.0f * (xz - wy));
    c1 = f3(2.0f * (xy - wz), 1.0f - 2.0f * (xx + zz), 2.0f * (yz + wx));
    c2 = f3(2.0f * (xz + wy), 2.0f * (yz - wx), 1.0f - 2.0f * (xx + yy));
}

// R * p accumulated column by column starting from 0 (tcnn mat*vec, vec.h:588-598)
GUT_HD float3 rotate_cols(float3 c0, float3 c1, float3 c2, float3 p) {
    float3 r;
    r.x = 0.0f + c0.x * p.x; r.x += c1.x * p.y; r.x += c2.x * p.z;
    r.y = 0.0f + c0.y * p.x; r.y += c1.y * p.y; r.y += c2.y * p.z;
    r.z = 0.0f + c0.z * p.x; r.z += c1.z * p.y; r.z += c2.z * p.z;
    return r;
}

// Natural log from IEEE ops + explicit fma only, so that host and device agree to the bit.
// Replaces the reference's fast-math logf in the culling bound (gutProjector.cuh:107,361).
GUT_HD float gut_logf(float x) {
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(x);
#else
    union { float f; uint32_t u; } cv; cv.f = x; uint32_t u = cv.u;
#endif
    const int32_t ix = (int32_t)(u - 0x3f3504f3u);
    const int32_t e = ix >> 23;
    const uint32_t mu = (uint32_t)(ix & 0x007fffff) + 0x3f3504f3u;
#ifdef __CUDA_ARCH__
    const float m = __uint_as_float(mu);
#else
    cv.u = mu; const float m = cv.f;
#endif
    const float s = (m - 1.0f) / (m + 1.0f);
    const float z = s * s;
    float p = 0.18181818181818182f;
    p = fmaf(p, z, 0.22222222222222222f);
    p = fmaf(p, z, 0.2857142857142857f);
    p = fmaf(p, z, 0.4f);
    p = fmaf(p, z, 0.6666666666666666f);
    p = p * z;
    const float r = fmaf(p, s, 2.0f * s);
    return fmaf((float)e, 0.6931471805599453f, r);
}

struct GutTileBox {
    uint32_t x0, y0, x1, y1;  // half-open
};

// reference: computeTileSpaceBBox, gutProjector.cuh:32-43
GUT_HD GutTileBox gut_tile_bbox(uint32_t gx, uint32_t gy, float2 pos, float2 ext) {
    GutTileBox b;
    int v;
    v = (int)floorf((pos.x - 0.5f - ext.x) / (float)GUT_TILE); v = v < 0 ? 0 : v; b.x0 = (uint32_t)v < gx ? (uint32_t)v : gx;
    v = (int)floorf((pos.y - 0.5f - ext.y) / (float)GUT_TILE); v = v < 0 ? 0 : v; b.y0 = (uint32_t)v < gy ? (uint32_t)v : gy;
    v = (int)ceilf((pos.x - 0.5f + ext.x) / (float)GUT_TILE);  v = v < 0 ? 0 : v; b.x1 = (uint32_t)v < gx ? (uint32_t)v : gx;
    v = (int)ceilf((pos.y - 0.5f + ext.y) / (float)GUT_TILE);  v = v < 0 ? 0 : v; b.y1 = (uint32_t)v < gy ? (uint32_t)v : gy;
    return b;
}

// Lowest conic "power" the particle reaches inside tile (tx,ty): the tile-culling bound.
// reference: tileMinParticlePowerResponse, gutProjector.cuh:49-78.  Written so that every product and
// sum rounds once, in the reference's evaluation order (the TU is built with --fmad=false), which makes
// the count pass (K1) and the emit pass (K2) agree with each other and with the oracle exactly.
// rcx, rcy = 1 / (ts^2 co.x), 1 / (ts^2 co.z): per-particle IEEE quotients, passed in so that a caller that tests many
// tiles of one particle divides once (same bits as dividing per tile).
GUT_HD float gut_tile_min_power_rc(float tx_, float ty_, float4 co, float2 mean, float rcx, float rcy) {
    const float ts = (float)GUT_TILE;
    const float tminx = ts * tx_, tminy = ts * ty_;
    const float tmaxx = ts + tminx, tmaxy = ts + tminy;
    const float mox = tminx - mean.x, moy = tminy - mean.y;
    const float lax = mox > 0.0f ? 1.0f : 0.0f, lay = moy > 0.0f ? 1.0f : 0.0f;
    const float nrx = lax + (mean.x > tmaxx ? 1.0f : 0.0f), nry = lay + (mean.y > tmaxy ? 1.0f : 0.0f);
    if ((nrx + nry) > 0.0f) {
        const float px = tmaxx * (1.0f - lax) + tminx * lax, py = tmaxy * (1.0f - lay) + tminy * lay;
        const float dxx = copysignf(ts, mox), dxy = copysignf(ts, moy);
        const float dfx = mean.x - px, dfy = mean.y - py;
        float ax = (dxx * co.x * dfx + dxx * co.y * dfy) * rcx;
        float ay = (dxy * co.y * dfx + dxy * co.z * dfy) * rcy;
        ax = ax < 0.0f ? 0.0f : (ax > 1.0f ? 1.0f : ax);  // __saturatef, NaN -> 0 below
        ay = ay < 0.0f ? 0.0f : (ay > 1.0f ? 1.0f : ay);
        if (!(ax == ax)) ax = 0.0f;
        if (!(ay == ay)) ay = 0.0f;
        const float tx = nry * ax, ty = nrx * ay;
        const float mx = mean.x - (px + tx * dxx), my = mean.y - (py + ty * dxy);
        return 0.5f * (co.x * mx * mx + co.z * my * my) + co.y * mx * my;
    }
    return 0.0f;
}
GUT_HD float gut_tile_min_power(float tx_, float ty_, float4 co, float2 mean) {
    const float ts = (float)GUT_TILE;
    return gut_tile_min_power_rc(tx_, ty_, co, mean, 1.0f / (ts * ts * co.x), 1.0f / (ts * ts * co.z));
}

// K1 -> emit hand-off (16 B per particle): the culling decisions of a tile box of up to 64 tiles as a bit mask (bit =
// row-major position inside the box), so the emit pass appends without repeating the tests.
//   x, y = mask bits 0-31, 32-63;  z = x0 | y0 << 12 | box width << 24 | (box above 64 tiles: emit re-walks it) << 31;  w = depth bits
#define GUT_EMIT_BIG 0x80000000u
#define GUT_EMIT_MASK_TILES 64u

// SH basis (degree <= 3), sign convention of the reference: gaussianParticles.cuh:57-96
GUT_HD void gut_sh_basis(int deg, float3 d, float* b) {
    const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
    const float x = d.x, y = d.y, z = d.z;
    b[0] = C0;
    if (deg > 0) {
        b[1] = -C1 * y; b[2] = C1 * z; b[3] = -C1 * x;
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            b[4] = 1.0925484305920792f * xy;
            b[5] = -1.0925484305920792f * yz;
            b[6] = 0.31539156525252005f * (2.0f * zz - xx - yy);
            b[7] = -1.0925484305920792f * xz;
            b[8] = 0.5462742152960396f * (xx - yy);
            if (deg > 2) {
                b[9]  = -0.5900435899266435f * y * (3.0f * xx - yy);
                b[10] = 2.890611442640554f * xy * z;
                b[11] = -0.4570457994644658f * y * (4.0f * zz - xx - yy);
                b[12] = 0.3731763325901154f * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
                b[13] = -0.4570457994644658f * x * (4.0f * zz - xx - yy);
                b[14] = 1.445305721320277f * z * (xx - yy);
                b[15] = -0.5900435899266435f * x * (xx - 3.0f * yy);
            }
        }
    }
}

// SelectiveAdam's scalar update (threedgrut/optimizers/optimizers.cu:47-84; no bias correction).  The rounding is that of
// the reference's compiled kernel (default nvcc flags: the moment updates contract into one FMA each, IEEE sqrt /
// division; SASS of oracle/_ref/optimizers_cu.o), spelled out because the translation units that use it are built with
// --fmad=false: bit-identical parameters, exp_avg and exp_avg_sq (tests/test_gpu_model.py, against that kernel).
GUT_D void adam_scalar(float& p, float g, float& m, float& v, float lr, float b1, float b2, float eps) {
    m = fmaf(g, 1.0f - b1, b1 * m);
    v = fmaf(g, g * (1.0f - b2), b2 * v);
    p = p - (lr * m) / (sqrtf(v) + eps);
}
