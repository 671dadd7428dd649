// threedgutSlang.cuh -- hand-written stand-in for the header `slangc` generates at
// /root/reference/threedgut_tracer/setup_3dgut.py:99-121 from include/3dgut/threedgut.slang.
//
// TEST INFRASTRUCTURE ONLY.  It lets the reference's own CUDA/C++ sources (src/gutRenderer.cu,
// src/splatRaster.cpp, src/cudaBuffer.cpp, bindings.cpp and every header below include/3dgut) compile
// UNCHANGED in an image without slangc.  Results obtained with it are "reference + shim", never
// "reference".  Nothing under 3dgrut_b200/ includes, links or loads anything built from this file.
//
// What the generated header provides, and what this file restates (slang file:line beside each):
//   * the CUDA prelude's element-wise vector operators the reference headers rely on (they are
//     commented out in kernels/cuda/common/mathUtils.cuh:26-34,81-91,168-178,204,263-300);
//   * the structs `gaussianParticle_*_0`, `shRadiativeParticle_*_0` (slangc's name mangling:
//     namespace_name_0) with the layouts of models/gaussianParticles.slang:20-35,78-83 and
//     models/shRadiativeParticles.slang:21-38;
//   * the [CudaDeviceExport] entry points called from
//     kernels/cuda/models/shRadiativeGaussianParticles.cuh:65-70,93-122,124-138,140-175,177-229,246-253,
//     265-274,288-343.
// The forward functions are literal restatements of the Slang source, operation by operation.
// The two backward entry points used when k_buffer_size > 0 are Slang *autodiff* output in the real
// build (bwd_diff(...), gaussianParticles.slang:425-434, shRadiativeParticles.slang:191-194); here they are
// hand-written reverse-mode adjoints of the same forward functions and are marked OWN MATH below.
// The K = 0 paths (forward through the functions here, backward through the reference's own
// processHitBwd, gaussianParticles.cuh:480-738) contain none of it.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

// --------------------------------------------------------------------------------------------------
// Prelude: element-wise vector arithmetic (slang-cuda-prelude.h provides these in the real build).
// Only the operators the reference headers do NOT define themselves.
static __host__ __device__ inline float2 operator+(const float2& a, const float2& b) { return make_float2(a.x + b.x, a.y + b.y); }
static __host__ __device__ inline float2 operator-(const float2& a, const float2& b) { return make_float2(a.x - b.x, a.y - b.y); }
static __host__ __device__ inline float2 operator*(const float2& a, const float2& b) { return make_float2(a.x * b.x, a.y * b.y); }
static __host__ __device__ inline float2 operator/(const float2& a, const float2& b) { return make_float2(a.x / b.x, a.y / b.y); }
static __host__ __device__ inline float2 operator-(const float2& a) { return make_float2(-a.x, -a.y); }
static __host__ __device__ inline float3 operator+(const float3& a, const float3& b) { return make_float3(a.x + b.x, a.y + b.y, a.z + b.z); }
static __host__ __device__ inline float3 operator-(const float3& a, const float3& b) { return make_float3(a.x - b.x, a.y - b.y, a.z - b.z); }
static __host__ __device__ inline float3 operator*(const float3& a, const float3& b) { return make_float3(a.x * b.x, a.y * b.y, a.z * b.z); }
static __host__ __device__ inline float3 operator/(const float3& a, const float3& b) { return make_float3(a.x / b.x, a.y / b.y, a.z / b.z); }
static __host__ __device__ inline float3 operator-(const float3& a) { return make_float3(-a.x, -a.y, -a.z); }
static __host__ __device__ inline float3 make_float3(float a) { return make_float3(a, a, a); }
static __host__ __device__ inline float3 make_float3(const float4& a) { return make_float3(a.x, a.y, a.z); }
static __host__ __device__ inline float4 make_float4(float a) { return make_float4(a, a, a, a); }

// --------------------------------------------------------------------------------------------------
// Types.  Slang matrices are row-major on the CUDA target (-matrix-layout-row-major,
// setup_3dgut.py:107): `rows[i]` is row i.
struct GutShimFloat3x3 {
    float3 rows[3];
};

// gaussianParticles.slang:20-26
struct gaussianParticle_RawParameters_0 {
    float3 position_0;
    float density_0;
    float4 quaternion_0;
    float3 scale_0;
    float padding_0;
};
// gaussianParticles.slang:28-32
struct gaussianParticle_RawParametersBuffer_0 {
    const gaussianParticle_RawParameters_0* _dataPtr_0;
    gaussianParticle_RawParameters_0* _gradPtr_0;
    bool exclusiveGradient_0;
};
// gaussianParticles.slang:34-36
struct gaussianParticle_CommonParameters_0 {
    gaussianParticle_RawParametersBuffer_0 parametersBuffer_0;
};
// gaussianParticles.slang:78-83
struct gaussianParticle_Parameters_0 {
    float3 position_0;
    float3 scale_0;
    GutShimFloat3x3 rotationT_0;
    float density_0;
};
// shRadiativeParticles.slang:21-26
struct shRadiativeParticle_ParametersBuffer_0 {
    const float3* _dataPtr_1;
    float3* _gradPtr_1;
    bool exclusiveGradient_1;
};
// shRadiativeParticles.slang:28-32
struct shRadiativeParticle_CommonParameters_0 {
    shRadiativeParticle_ParametersBuffer_0 parametersBuffer_1;
    int sphDegree_0;
};
// shRadiativeParticles.slang:34-37
struct shRadiativeParticle_Parameters_0 {
    float3 sphCoefficients_0[PARTICLE_RADIANCE_NUM_COEFFS];
};

namespace gutshim {

// compile-time constants of threedgut.slang:16-24,28-32
constexpr int KernelDegree              = GAUSSIAN_PARTICLE_KERNEL_DEGREE;
constexpr float MinParticleKernelDensity = GAUSSIAN_PARTICLE_MIN_KERNEL_DENSITY;
constexpr float MaxParticleAlpha         = GAUSSIAN_PARTICLE_MAX_ALPHA;
constexpr float MinParticleAlpha         = GAUSSIAN_PARTICLE_MIN_ALPHA;
constexpr int NumSph                     = PARTICLE_RADIANCE_NUM_COEFFS;

// scalar-vector products: the reference declares its own global versions later (mathUtils.cuh:180-201), after this
// header is included, so the functions of this file use these namespace-local ones
__device__ inline float3 operator*(float a, const float3& b) { return make_float3(a * b.x, a * b.y, a * b.z); }
__device__ inline float3 operator*(const float3& a, float b) { return make_float3(a.x * b, a.y * b, a.z * b); }
__device__ inline float3& operator+=(float3& a, const float3& b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
__device__ inline float4& operator+=(float4& a, const float4& b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; return a; }
__device__ inline float dot3(const float3& a, const float3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ inline float3 cross3(const float3& a, const float3& b) {
    return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
// Slang mul(M, v) with row-major M: out_i = dot(row_i, v)
__device__ inline float3 mulMatVec(const GutShimFloat3x3& m, const float3& v) {
    return make_float3(dot3(m.rows[0], v), dot3(m.rows[1], v), dot3(m.rows[2], v));
}
// Slang core library: normalize(x) = x / length(x), length(x) = sqrt(dot(x, x))
__device__ inline float3 normalize3(const float3& v) {
    const float len = sqrtf(dot3(v, v));
    return make_float3(v.x / len, v.y / len, v.z / len);
}

// transforms.slang:21-38
__device__ inline GutShimFloat3x3 rotationMatrixTranspose(const float4& quaternion) {
    const float xx = quaternion.y * quaternion.y;
    const float yy = quaternion.z * quaternion.z;
    const float zz = quaternion.w * quaternion.w;
    const float xy = quaternion.y * quaternion.z;
    const float xz = quaternion.y * quaternion.w;
    const float yz = quaternion.z * quaternion.w;
    const float rx = quaternion.x * quaternion.y;
    const float ry = quaternion.x * quaternion.z;
    const float rz = quaternion.x * quaternion.w;
    GutShimFloat3x3 m;
    m.rows[0] = make_float3((1.f - 2.f * (yy + zz)), 2.f * (xy + rz), 2.f * (xz - ry));
    m.rows[1] = make_float3(2.f * (xy - rz), (1.f - 2.f * (xx + zz)), 2.f * (yz + rx));
    m.rows[2] = make_float3(2.f * (xz + ry), 2.f * (yz - rx), (1.f - 2.f * (xx + yy)));
    return m;
}

// gaussianParticles.slang:85-95 (fetchParameters; Surfel = false, threedgut.slang:23)
__device__ inline gaussianParticle_Parameters_0 fetchParameters(uint32_t particleIdx,
                                                                const gaussianParticle_RawParametersBuffer_0& buffer) {
    const gaussianParticle_RawParameters_0 raw = buffer._dataPtr_0[particleIdx];
    gaussianParticle_Parameters_0 p;
    p.position_0  = raw.position_0;
    p.scale_0     = raw.scale_0;
    p.rotationT_0 = rotationMatrixTranspose(raw.quaternion_0);
    p.density_0   = raw.density_0;
    return p;
}

// gaussianParticles.slang:96-111
__device__ inline void cannonicalRay(const float3& rayOrigin, const float3& rayDirection,
                                     const gaussianParticle_Parameters_0& parameters,
                                     float3& particleRayOrigin, float3& particleRayDirection) {
    const float3 giscl  = make_float3(1.0f / parameters.scale_0.x, 1.0f / parameters.scale_0.y, 1.0f / parameters.scale_0.z);
    const float3 gposc  = rayOrigin - parameters.position_0;
    const float3 gposcr = mulMatVec(parameters.rotationT_0, gposc);
    particleRayOrigin   = giscl * gposcr;
    const float3 rayDirR = mulMatVec(parameters.rotationT_0, rayDirection);
    const float3 grdu    = giscl * rayDirR;
    particleRayDirection = normalize3(grdu);
}

// gaussianParticles.slang:113-118
__device__ inline float canonicalRayMinSquaredDistance(const float3& o, const float3& d) {
    const float3 gcrod = cross3(d, o);
    return dot3(gcrod, gcrod);
}

// gaussianParticles.slang:120-166
template <int Degree>
__device__ inline float canonicalRayMaxKernelResponse(const float3& o, const float3& d) {
    const float grayDist = canonicalRayMinSquaredDistance(o, d);
    if constexpr (Degree == 8) {
        const float s  = -0.000685871056241f;
        const float sq = grayDist * grayDist;
        return expf(s * sq * sq);
    } else if constexpr (Degree == 5) {
        const float s = -0.0185185185185f;
        return expf(s * grayDist * grayDist * sqrtf(grayDist));
    } else if constexpr (Degree == 4) {
        const float s = -0.0555555555556f;
        return expf(s * grayDist * grayDist);
    } else if constexpr (Degree == 3) {
        const float s = -0.166666666667f;
        return expf(s * grayDist * sqrtf(grayDist));
    } else if constexpr (Degree == 1) {
        const float s = -1.5f;
        return expf(s * sqrtf(grayDist));
    } else if constexpr (Degree == 0) {
        const float s = -0.329630334487f;
        return fmaxf(1 + s * sqrtf(grayDist), 0.f);
    } else {
        const float s = -0.5f;
        return expf(s * grayDist);
    }
}

// gaussianParticles.slang:168-174
__device__ inline float canonicalRayDistance(const float3& o, const float3& d, const float3& scale) {
    const float k     = dot3(d, make_float3(-1 * o.x, -1 * o.y, -1 * o.z));
    const float3 grds = make_float3(scale.x * d.x * k, scale.y * d.y * k, scale.z * d.z * k);
    return sqrtf(dot3(grds, grds));
}

// gaussianParticles.slang:188-222 (EnableNormal = false: the normal branch is never taken by the 3DGUT renderer,
// gutKBufferRenderer.cuh:245-249 passes no normal pointer)
__device__ inline bool hit(const float3& rayOrigin, const float3& rayDirection,
                           const gaussianParticle_Parameters_0& parameters, float& alpha, float& depth) {
    float3 o, d;
    cannonicalRay(rayOrigin, rayDirection, parameters, o, d);
    const float maxResponse = canonicalRayMaxKernelResponse<KernelDegree>(o, d);
    alpha                   = fminf(MaxParticleAlpha, maxResponse * parameters.density_0);
    const bool acceptHit    = ((maxResponse > MinParticleKernelDensity) && (alpha > MinParticleAlpha));
    if (acceptHit) {
        depth = canonicalRayDistance(o, d, parameters.scale_0);
    }
    return acceptHit;
}

// gaussianParticles.slang:224-254
template <bool backToFront>
__device__ inline float integrateHit(float alpha, float& transmittance, float depth, float& integratedDepth) {
    const float weight = backToFront ? alpha : alpha * transmittance;
    if (backToFront) {
        integratedDepth = integratedDepth + (depth - integratedDepth) * alpha; // lerp(x, y, s) = x + (y - x) * s
    } else {
        integratedDepth += depth * weight;
    }
    transmittance *= (1 - alpha);
    return weight;
}

// sphericalHarmonics.slang:20-63
__device__ inline float3 shDecode(int degree, const float3* c, const float3& dir) {
    const float C0   = 0.28209479177387814f;
    const float C1   = 0.4886025119029199f;
    const float C2[] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f};
    const float C3[] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                        -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};
    float3 f = C0 * c[0];
    if (degree > 0) {
        const float x = dir.x, y = dir.y, z = dir.z;
        f = f - C1 * y * c[1] + C1 * z * c[2] - C1 * x * c[3];
        if (degree > 1) {
            const float xx = x * x, yy = y * y, zz = z * z;
            const float xy = x * y, yz = y * z, xz = x * z;
            f = f + C2[0] * xy * c[4] + C2[1] * yz * c[5] + C2[2] * (2.0f * zz - xx - yy) * c[6] + C2[3] * xz * c[7] +
                C2[4] * (xx - yy) * c[8];
            if (degree > 2) {
                f = f + C3[0] * y * (3.0f * xx - yy) * c[9] + C3[1] * xy * z * c[10] + C3[2] * y * (4.0f * zz - xx - yy) * c[11] +
                    C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * c[12] + C3[4] * x * (4.0f * zz - xx - yy) * c[13] +
                    C3[5] * z * (xx - yy) * c[14] + C3[6] * x * (xx - 3.0f * yy) * c[15];
            }
        }
    }
    return make_float3(fmaxf(f.x + 0.5f, 0.f), fmaxf(f.y + 0.5f, 0.f), fmaxf(f.z + 0.5f, 0.f));
}

// shRadiativeParticles.slang:83-99
template <bool backToFront>
__device__ inline void integrateRadiance(float weight, const float3& radiance, float3& integratedRadiance) {
    if (weight > 0.0f) {
        if (backToFront) {
            // lerp(x, y, s) = x + (y - x) * s, written per component: this is a template, and at its point of
            // instantiation the reference's own global float3 operators are visible as well (ambiguous otherwise)
            integratedRadiance.x = integratedRadiance.x + (radiance.x - integratedRadiance.x) * weight;
            integratedRadiance.y = integratedRadiance.y + (radiance.y - integratedRadiance.y) * weight;
            integratedRadiance.z = integratedRadiance.z + (radiance.z - integratedRadiance.z) * weight;
        } else {
            integratedRadiance.x += radiance.x * weight;
            integratedRadiance.y += radiance.y * weight;
            integratedRadiance.z += radiance.z * weight;
        }
    }
}

} // namespace gutshim

// --------------------------------------------------------------------------------------------------
// Entry points ([CudaDeviceExport], gaussianParticles.slang:322-528)

// gaussianParticles.slang:322-329
__device__ inline gaussianParticle_Parameters_0 particleDensityParameters(uint32_t particleIdx,
                                                                          gaussianParticle_CommonParameters_0 commonParameters) {
    return gutshim::fetchParameters(particleIdx, commonParameters.parametersBuffer_0);
}

// gaussianParticles.slang:331-349
__device__ inline bool particleDensityHit(float3 rayOrigin, float3 rayDirection, gaussianParticle_Parameters_0 parameters,
                                          float* alpha, float* depth, bool /*enableNormal*/, float3* /*normal*/) {
    float a = 0.f;
    float d = *depth; // `inout` in gaussianParticle.hit: untouched when the hit is rejected
    const bool r = gutshim::hit(rayOrigin, rayDirection, parameters, a, d);
    *alpha = a;
    *depth = d;
    return r;
}

// gaussianParticles.slang:351-368
__device__ inline float particleDensityIntegrateHit(float alpha, float* transmittance, float depth, float* integratedDepth,
                                                    bool /*enableNormal*/, float3 /*normal*/, float3* /*integratedNormal*/) {
    return gutshim::integrateHit<false>(alpha, *transmittance, depth, *integratedDepth);
}

// gaussianParticles.slang:370-391 / 256-286
__device__ inline float particleDensityProcessHitFwdFromBuffer(float3 rayOrigin, float3 rayDirection, uint32_t particleIdx,
                                                               gaussianParticle_CommonParameters_0 commonParameters,
                                                               float* transmittance, float* integratedDepth,
                                                               bool /*enableNormal*/, float3* /*integratedNormal*/) {
    float alpha = 0.0f;
    float depth = 0.0f;
    if (gutshim::hit(rayOrigin, rayDirection, gutshim::fetchParameters(particleIdx, commonParameters.parametersBuffer_0), alpha, depth)) {
        return gutshim::integrateHit<false>(alpha, *transmittance, depth, *integratedDepth);
    }
    return 0.0f;
}

// gaussianParticles.slang:394-451.  OWN MATH: the call bwd_diff(processHitFromBuffer<true>) at :425-434 is Slang
// autodiff; below is the reverse-mode adjoint of processHitFromBuffer<true> (:256-286) written by hand, one
// adjoint statement per forward statement, in the order autodiff would visit them.
__device__ inline void particleDensityProcessHitBwdToBuffer(float3 rayOrigin, float3 rayDirection, uint32_t particleIdx,
                                                            gaussianParticle_CommonParameters_0 commonParameters,
                                                            float alpha, float alphaGrad,
                                                            float* transmittance, float* transmittanceGrad,
                                                            float depth, float* integratedDepth, float* integratedDepthGrad,
                                                            bool /*enableNormal*/, float3 /*normal*/,
                                                            float3* /*integratedNormal*/, float3* /*integratedNormalGrad*/) {
    using namespace gutshim;
    if (!(alpha > 0.0f)) return;
    // :413-419 un-blend the ray state to what it was before this hit (back-to-front formulation)
    const float weight = 1.0f / (1.0f - alpha);
    *transmittance *= weight;
    *integratedDepth = (*integratedDepth - depth * alpha) * weight;
    const float T0  = *transmittance;   // primal inputs of processHitFromBuffer<true>
    const float D0  = *integratedDepth;
    const float dT1 = *transmittanceGrad;   // adjoints of its outputs
    const float dD1 = *integratedDepthGrad;
    const float dW  = alphaGrad;            // adjoint of the returned weight (= alpha when backToFront)

    // ---- forward recomputation (fetchParameters, hit) keeping the intermediates the adjoint needs
    const gaussianParticle_RawParametersBuffer_0& buf = commonParameters.parametersBuffer_0;
    const gaussianParticle_RawParameters_0 raw = buf._dataPtr_0[particleIdx];
    const GutShimFloat3x3 Rt = rotationMatrixTranspose(raw.quaternion_0);
    const float3 s      = raw.scale_0;
    const float3 giscl  = make_float3(1.0f / s.x, 1.0f / s.y, 1.0f / s.z);
    const float3 gposc  = rayOrigin - raw.position_0;
    const float3 gposcr = mulMatVec(Rt, gposc);
    const float3 o      = giscl * gposcr;
    const float3 rayDirR = mulMatVec(Rt, rayDirection);
    const float3 grdu    = giscl * rayDirR;
    const float len      = sqrtf(dot3(grdu, grdu));
    const float3 d       = make_float3(grdu.x / len, grdu.y / len, grdu.z / len);
    const float3 gcrod   = cross3(d, o);
    const float grayDist = dot3(gcrod, gcrod);
    static_assert(KernelDegree == 2, "shim adjoint written for the quadratic kernel (3dgut.yaml:8) only");
    const float maxResponse = expf(-0.5f * grayDist);
    const float prod        = maxResponse * raw.density_0;
    const float a           = fminf(MaxParticleAlpha, prod);
    const bool acceptHit    = ((maxResponse > MinParticleKernelDensity) && (a > MinParticleAlpha));

    float dT0 = dT1, dD0 = dD1;
    if (acceptHit) {
        const float k      = -dot3(d, o);
        const float3 grds  = make_float3(s.x * d.x * k, s.y * d.y * k, s.z * d.z * k);
        const float dist   = sqrtf(dot3(grds, grds));
        // adjoint of integrateHit<true>: D1 = D0 + (dist - D0) a ; T1 = T0 (1 - a) ; w = a
        float dA    = dW + dD1 * (dist - D0) - dT1 * T0;
        float dDist = dD1 * a;
        dD0         = dD1 * (1.0f - a);
        dT0         = dT1 * (1.0f - a);
        // adjoint of hit(): dist = sqrt(dot(grds, grds))
        float3 dgrds = make_float3(0.f);
        if (dist > 0.f) {
            const float c = dDist / dist; // d sqrt(x) = 1 / (2 sqrt x), d dot(g,g) = 2 g
            dgrds = make_float3(c * grds.x, c * grds.y, c * grds.z);
        }
        // grds = s * d * k
        float3 ds  = make_float3(dgrds.x * d.x * k, dgrds.y * d.y * k, dgrds.z * d.z * k);
        float3 dd  = make_float3(dgrds.x * s.x * k, dgrds.y * s.y * k, dgrds.z * s.z * k);
        const float dk = dgrds.x * s.x * d.x + dgrds.y * s.y * d.y + dgrds.z * s.z * d.z;
        // k = -dot(d, o)
        dd = dd - dk * o;
        float3 dO = make_float3(-dk * d.x, -dk * d.y, -dk * d.z);
        // alpha = min(MaxParticleAlpha, maxResponse * density): no gradient through the clamped branch
        const float dProd = (MaxParticleAlpha < prod) ? 0.f : dA; // Slang d min(x, y) = x < y ? dx : dy
        const float dResp = dProd * raw.density_0;
        const float dDens = dProd * maxResponse;
        // maxResponse = exp(-0.5 grayDist)
        const float dGray = dResp * maxResponse * -0.5f;
        // grayDist = dot(gcrod, gcrod), gcrod = cross(d, o)
        const float3 dgcrod = make_float3(2.f * dGray * gcrod.x, 2.f * dGray * gcrod.y, 2.f * dGray * gcrod.z);
        dd = dd + cross3(o, dgcrod);      // d/da cross(a,b)^T g = b x g
        dO = dO + cross3(dgcrod, d);      // d/db cross(a,b)^T g = g x a
        // d = grdu / len
        const float ddotd = dot3(dd, d);
        const float3 dgrdu = make_float3((dd.x - d.x * ddotd) / len, (dd.y - d.y * ddotd) / len, (dd.z - d.z * ddotd) / len);
        // grdu = giscl * rayDirR ; o = giscl * gposcr
        float3 dgiscl   = dgrdu * rayDirR + dO * gposcr;
        const float3 drayDirR = dgrdu * giscl;
        const float3 dgposcr  = dO * giscl;
        // giscl = 1 / s
        ds = ds - dgiscl * giscl * giscl;
        // gposcr = Rt gposc ; rayDirR = Rt rayDirection  -> dRt rows, dgposc = Rt^T dgposcr
        GutShimFloat3x3 dRt;
        dRt.rows[0] = dgposcr.x * gposc + drayDirR.x * rayDirection;
        dRt.rows[1] = dgposcr.y * gposc + drayDirR.y * rayDirection;
        dRt.rows[2] = dgposcr.z * gposc + drayDirR.z * rayDirection;
        const float3 dgposc = make_float3(Rt.rows[0].x * dgposcr.x + Rt.rows[1].x * dgposcr.y + Rt.rows[2].x * dgposcr.z,
                                          Rt.rows[0].y * dgposcr.x + Rt.rows[1].y * dgposcr.y + Rt.rows[2].y * dgposcr.z,
                                          Rt.rows[0].z * dgposcr.x + Rt.rows[1].z * dgposcr.y + Rt.rows[2].z * dgposcr.z);
        const float3 dpos = -dgposc;
        // adjoint of rotationMatrixTranspose (transforms.slang:21-38); quaternion = (r, x, y, z) in (.x,.y,.z,.w)
        const float qr = raw.quaternion_0.x, qx = raw.quaternion_0.y, qy = raw.quaternion_0.z, qz = raw.quaternion_0.w;
        const float m00 = dRt.rows[0].x, m01 = dRt.rows[0].y, m02 = dRt.rows[0].z;
        const float m10 = dRt.rows[1].x, m11 = dRt.rows[1].y, m12 = dRt.rows[1].z;
        const float m20 = dRt.rows[2].x, m21 = dRt.rows[2].y, m22 = dRt.rows[2].z;
        const float dxx = -2.f * (m11 + m22), dyy = -2.f * (m00 + m22), dzz = -2.f * (m00 + m11);
        const float dxy = 2.f * (m01 + m10), dxz = 2.f * (m02 + m20), dyz = 2.f * (m12 + m21);
        const float drx = 2.f * (m12 - m21), dry = 2.f * (m20 - m02), drz = 2.f * (m01 - m10);
        float4 dq;
        dq.x = drx * qx + dry * qy + drz * qz;
        dq.y = 2.f * dxx * qx + dxy * qy + dxz * qz + drx * qr;
        dq.z = 2.f * dyy * qy + dxy * qx + dyz * qz + dry * qr;
        dq.w = 2.f * dzz * qz + dxz * qx + dyz * qy + drz * qr;
        // fetchParametersFromBufferBwd, gaussianParticles.slang:49-76
        gaussianParticle_RawParameters_0* g = &buf._gradPtr_0[particleIdx];
        if (buf.exclusiveGradient_0) {
            g->density_0 += dDens;
            g->position_0 += dpos;
            g->quaternion_0 += dq;
            g->scale_0 += ds;
        } else {
            atomicAdd(&g->density_0, dDens);
            atomicAdd(&g->position_0.x, dpos.x);
            atomicAdd(&g->position_0.y, dpos.y);
            atomicAdd(&g->position_0.z, dpos.z);
            atomicAdd(&g->quaternion_0.x, dq.x);
            atomicAdd(&g->quaternion_0.y, dq.y);
            atomicAdd(&g->quaternion_0.z, dq.z);
            atomicAdd(&g->quaternion_0.w, dq.w);
            atomicAdd(&g->scale_0.x, ds.x);
            atomicAdd(&g->scale_0.y, ds.y);
            atomicAdd(&g->scale_0.z, ds.z);
        }
    }
    // :436-441
    *transmittanceGrad   = dT0;
    *integratedDepthGrad = dD0;
}

// gaussianParticles.slang:453-488
__device__ inline bool particleDensityHitCustom(float3 rayOrigin, float3 rayDirection, int32_t particleIdx,
                                                gaussianParticle_CommonParameters_0 commonParameters,
                                                float minHitDistance, float maxHitDistance, float maxParticleSquaredDistance,
                                                float* hitDistance) {
    const gaussianParticle_Parameters_0 parameters = gutshim::fetchParameters(particleIdx, commonParameters.parametersBuffer_0);
    float3 o, d;
    gutshim::cannonicalRay(rayOrigin, rayDirection, parameters, o, d);
    *hitDistance = gutshim::canonicalRayDistance(o, d, parameters.scale_0);
    return (*hitDistance > minHitDistance) && (*hitDistance < maxHitDistance) &&
           (gutshim::canonicalRayMinSquaredDistance(o, d) < maxParticleSquaredDistance);
}

// gaussianParticles.slang:490-509
__device__ inline bool particleDensityHitInstance(float3 canonicalRayOrigin, float3 canonicalUnormalizedRayDirection,
                                                  float minHitDistance, float maxHitDistance, float maxParticleSquaredDistance,
                                                  float* hitDistance) {
    const float numerator   = -gutshim::dot3(canonicalRayOrigin, canonicalUnormalizedRayDirection);
    const float denominator = 1.0f / gutshim::dot3(canonicalUnormalizedRayDirection, canonicalUnormalizedRayDirection);
    *hitDistance            = numerator * denominator;
    return (*hitDistance > minHitDistance) && (*hitDistance < maxHitDistance) &&
           (gutshim::canonicalRayMinSquaredDistance(canonicalRayOrigin, gutshim::normalize3(canonicalUnormalizedRayDirection)) <
            maxParticleSquaredDistance);
}

// gaussianParticles.slang:511-516 / 288-296
__device__ inline float3 particleDensityIncidentDirection(gaussianParticle_Parameters_0 parameters, float3 sourcePosition) {
    return gutshim::normalize3(parameters.position_0 - sourcePosition);
}

// gaussianParticles.slang:518-528: bwd_diff of incidentDirectionFromBuffer, whose result is declared `no_diff`
// (:298-310) -- the generated function propagates nothing.
__device__ inline void particleDensityIncidentDirectionBwdToBuffer(uint32_t /*particleIdx*/,
                                                                   gaussianParticle_CommonParameters_0 /*commonParameters*/,
                                                                   float3 /*sourcePosition*/) {
}

// --------------------------------------------------------------------------------------------------
// Entry points (shRadiativeParticles.slang:136-253)

// shRadiativeParticles.slang:136-146
__device__ inline float3 particleFeaturesFromBuffer(uint32_t particleIdx, shRadiativeParticle_CommonParameters_0 commonParameters,
                                                    float3 incidentDirection) {
    return gutshim::shDecode(commonParameters.sphDegree_0,
                             commonParameters.parametersBuffer_1._dataPtr_1 + particleIdx * gutshim::NumSph, incidentDirection);
}

// shRadiativeParticles.slang:148-158
__device__ inline void particleFeaturesIntegrateFwd(float weight, float3 features, float3* integratedFeatures) {
    gutshim::integrateRadiance<false>(weight, features, *integratedFeatures);
}

// shRadiativeParticles.slang:160-173
__device__ inline void particleFeaturesIntegrateFwdFromBuffer(float3 incidentDirection, float weight, uint32_t particleIdx,
                                                              shRadiativeParticle_CommonParameters_0 commonParameters,
                                                              float3* integratedFeatures) {
    if (weight > 0.0f) {
        gutshim::integrateRadiance<false>(weight, particleFeaturesFromBuffer(particleIdx, commonParameters, incidentDirection),
                                          *integratedFeatures);
    }
}

// shRadiativeParticles.slang:175-202.  OWN MATH: hand-written adjoint of integrateRadiance<true> (:83-99)
// where the real build calls bwd_diff:  I1 = I0 + (f - I0) w  (only when w > 0).
__device__ inline void particleFeaturesIntegrateBwd(float alpha, float* alphaGrad, float3 features, float3* featuresGrad,
                                                    float3* integratedFeatures, float3* integratedFeaturesGrad) {
    using namespace gutshim;
    if (alpha > 0.0f) {
        const float weight   = 1.0f / (1.0f - alpha);
        *integratedFeatures  = (*integratedFeatures - features * alpha) * weight;
        const float3 I0      = *integratedFeatures;
        const float3 dI1     = *integratedFeaturesGrad;
        // weight argument = alpha > 0 here, so the branch of integrateRadiance is taken
        // `in` parameters of a bwd_diff call receive their gradient (no accumulation); the caller passes zeros
        *alphaGrad    = dot3(dI1, features - I0);
        *featuresGrad = dI1 * alpha;
        *integratedFeaturesGrad = dI1 * (1.0f - alpha);
    }
}

// shRadiativeParticles.slang:204-235 (per-ray features; PerRayParticleFeatures = false in threedgut.cuh:55, never
// instantiated by the 3DGUT renderer) -- declared so that the header is complete.
__device__ inline void particleFeaturesIntegrateBwdToBuffer(float3, float, float*, uint32_t, shRadiativeParticle_CommonParameters_0,
                                                            float3, float3*, float3*) {
    asm volatile("trap;"); // not reachable from the 3DGUT configuration
}

// shRadiativeParticles.slang:237-253 (used only with per-ray features, as above)
__device__ inline void particleFeaturesBwdToBuffer(uint32_t, shRadiativeParticle_CommonParameters_0, float3, float3) {
    asm volatile("trap;"); // not reachable from the 3DGUT configuration
}
