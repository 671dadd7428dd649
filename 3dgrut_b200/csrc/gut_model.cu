// gut_model.cu -- the per-particle passes either side of the rasteriser.
//
//  (1) activation + packing at the Tracer boundary.  The reference model applies its activations as separate torch
//      kernels every forward -- normalize(rotation), exp(scale), sigmoid(density), cat(albedo, specular)
//      (threedgrut/model/model.py:74-93, activations from threedgrut/utils/misc.py:48-53) -- and the tracer then
//      concatenates the results into [N,12] (threedgut_tracer/tracer.py:176-178); autograd runs the mirror image.
//      Here raw nn.Parameters go in and the packed records come out of ONE pass (44 B read + 48 B written per
//      particle), and the backward turns the [N,12] / [N,48] gradients into raw-parameter gradients in one pass each.
//  (2) the optimiser step after the gradient exchange: SelectiveAdam (threedgrut/optimizers/optimizers.cu:47-84, no
//      bias correction, invisible particles untouched) for ALL parameter groups in one launch, and the densification
//      statistics of GSStrategy.update_gradient_buffer (threedgrut/strategy/gs.py:104-115).
//
// Every kernel here is HBM bound: one pass, 128-bit accesses where the row pitch allows, no intermediate tensors.
// Built WITHOUT -use_fast_math: expf / sqrtf / division are the IEEE-accurate ones torch uses.
#include "gut_internal.h"

namespace {

constexpr int MODEL_THREADS = 256;

GUT_D float activate(int kind, float x) {
    if (kind == GUT_ACT_EXP) return expf(x);
    if (kind == GUT_ACT_SIGMOID) return 1.0f / (1.0f + expf(-x));
    return x;
}
// derivative of the activation expressed through its OUTPUT y
GUT_D float activate_grad(int kind, float y, float g) {
    if (kind == GUT_ACT_EXP) return g * y;
    if (kind == GUT_ACT_SIGMOID) return g * (y * (1.0f - y));   // torch: grad * (1 - y) * y
    return g;
}

// torch.nn.functional.normalize(x, p=2, dim=1, eps=1e-12): x / max(|x|, eps)
constexpr float NORMALIZE_EPS = 1e-12f;

__global__ void __launch_bounds__(MODEL_THREADS) gut_activate_pack_kernel(uint32_t N, const float* __restrict__ pos,
                                                                         const float4* __restrict__ rot, const float* __restrict__ scale,
                                                                         const float* __restrict__ dens, int scale_act, int density_act,
                                                                         float4* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const size_t i3 = 3 * (size_t)i;
    const float px = __ldg(pos + i3), py = __ldg(pos + i3 + 1), pz = __ldg(pos + i3 + 2);
    const float4 q = __ldg(rot + i);
    const float sx = __ldg(scale + i3), sy = __ldg(scale + i3 + 1), sz = __ldg(scale + i3 + 2);
    const float d = __ldg(dens + i);
    const float n = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), NORMALIZE_EPS);
    out[i3] = make_float4(px, py, pz, activate(density_act, d));
    out[i3 + 1] = make_float4(q.x / n, q.y / n, q.z / n, q.w / n);
    out[i3 + 2] = make_float4(activate(scale_act, sx), activate(scale_act, sy), activate(scale_act, sz), 0.0f);
}

// [N,12] gradient (wrt the ACTIVATED record) -> gradients of the raw parameters.  The activated scale / density are read
// back from the packed record the forward wrote (saved for the backward anyway), the quaternion norm is recomputed.
__global__ void __launch_bounds__(MODEL_THREADS) gut_activate_backward_kernel(uint32_t N, const float4* __restrict__ rot,
                                                                             const float4* __restrict__ packed, const float4* __restrict__ grad12,
                                                                             int scale_act, int density_act, float* __restrict__ pos_g,
                                                                             float4* __restrict__ rot_g, float* __restrict__ scale_g,
                                                                             float* __restrict__ dens_g) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const size_t i3 = 3 * (size_t)i;
    const float4 ga = __ldg(grad12 + i3), gq = __ldg(grad12 + i3 + 1), gs = __ldg(grad12 + i3 + 2);
    const float4 pa = __ldg(packed + i3), ps = __ldg(packed + i3 + 2);
    const float4 q = __ldg(rot + i);
    pos_g[i3] = ga.x; pos_g[i3 + 1] = ga.y; pos_g[i3 + 2] = ga.z;
    dens_g[i] = activate_grad(density_act, pa.w, ga.w);
    scale_g[i3] = activate_grad(scale_act, ps.x, gs.x);
    scale_g[i3 + 1] = activate_grad(scale_act, ps.y, gs.y);
    scale_g[i3 + 2] = activate_grad(scale_act, ps.z, gs.z);
    // y = x / max(|x|, eps):  |x| > eps -> (g - y (y.g)) / |x| ;  else the clamp passes no gradient to the norm -> g / eps
    const float nn = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    float4 r;
    if (nn > NORMALIZE_EPS) {
        const float inv = 1.0f / nn;
        const float yx = q.x * inv, yy = q.y * inv, yz = q.z * inv, yw = q.w * inv;
        const float dt = yx * gq.x + yy * gq.y + yz * gq.z + yw * gq.w;
        r = make_float4((gq.x - yx * dt) * inv, (gq.y - yy * dt) * inv, (gq.z - yz * dt) * inv, (gq.w - yw * dt) * inv);
    } else {
        r = make_float4(gq.x / NORMALIZE_EPS, gq.y / NORMALIZE_EPS, gq.z / NORMALIZE_EPS, gq.w / NORMALIZE_EPS);
    }
    rot_g[i] = r;
}

// cat(albedo [N,3], specular [N,45]) -> [N,48]: one thread per 16-byte piece of the output, so the stores are full
// 128-bit and a warp's reads of either source cover one contiguous span.
__global__ void __launch_bounds__(MODEL_THREADS) gut_pack_features_kernel(uint64_t n_vec, const float* __restrict__ albedo,
                                                                         const float* __restrict__ specular, float4* __restrict__ out) {
    const uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_vec) return;
    const uint64_t row = k / 12u;
    const uint32_t c = (uint32_t)(k - row * 12u);
    float4 v;
    if (c == 0) {
        const float* a = albedo + 3 * row;
        v = make_float4(__ldg(a), __ldg(a + 1), __ldg(a + 2), __ldg(specular + 45 * row));
    } else {
        const float* s = specular + 45 * row + (4 * c - 3);
        v = make_float4(__ldg(s), __ldg(s + 1), __ldg(s + 2), __ldg(s + 3));
    }
    out[k] = v;
}

__global__ void __launch_bounds__(MODEL_THREADS) gut_unpack_features_kernel(uint64_t n_vec, const float4* __restrict__ grad48,
                                                                           float* __restrict__ albedo_g, float* __restrict__ specular_g) {
    const uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_vec) return;
    const uint64_t row = k / 12u;
    const uint32_t c = (uint32_t)(k - row * 12u);
    const float4 v = __ldg(grad48 + k);
    if (c == 0) {
        float* a = albedo_g + 3 * row;
        a[0] = v.x; a[1] = v.y; a[2] = v.z;
        specular_g[45 * row] = v.w;
    } else {
        float* s = specular_g + 45 * row + (4 * c - 3);
        s[0] = v.x; s[1] = v.y; s[2] = v.z; s[3] = v.w;
    }
}

// Any coefficient count (sph_max_degree 0..2: 1, 4, 9 coefficients): one thread per float of the [N, 3 ncoef] side, both sides
// read / written as contiguous spans.  PACK = true: cat; false: the split of the gradient.
template <bool PACK>
__global__ void __launch_bounds__(MODEL_THREADS) gut_features_generic_kernel(uint64_t total, uint32_t width, float* __restrict__ albedo,
                                                                            float* __restrict__ specular, float* __restrict__ packed) {
    const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= total) return;
    const uint64_t row = e / width;
    const uint32_t c = (uint32_t)(e - row * width);
    float* side = c < 3 ? albedo + 3 * row + c : specular + (uint64_t)(width - 3) * row + (c - 3);
    if (PACK) packed[e] = *side;
    else *side = packed[e];
}

// ---------------------------------------------------------------------------------------------------
// SelectiveAdam for every parameter group in one launch.
// reference: selective_adam_update_kernel (threedgrut/optimizers/optimizers.cu:47-84) is launched once per group with
// one thread per scalar; SelectiveAdam.step (optimizers/__init__.py:88-133) loops over the groups.  Same arithmetic
// (no bias correction; `step = -lr * m / (sqrt(v) + eps)`; invisible particles keep param, m and v), but: the CTA index
// selects the group, a thread owns four consecutive scalars (128-bit accesses when the group's base pointers allow),
// the visibility of the rows it touches is read first so that invisible rows cost no parameter traffic, and the
// gradient can be zeroed in the same pass (optimizer.zero_grad(), trainer.py:751).
// ---------------------------------------------------------------------------------------------------
struct AdamGroups {
    GutAdamGroup g[GUT_ADAM_MAX_GROUPS];
    uint32_t block_start[GUT_ADAM_MAX_GROUPS + 1];   // first CTA of each group in the flattened 1-D grid
    uint32_t n_groups;
    uint32_t vec_ok;     // bit i: group i's four pointers are 16-byte aligned and its element count is a multiple of 4
};

// adam_scalar: gut_math.cuh (shared with the fused SH expansion + Adam kernel of gut_project.cu)
__global__ void __launch_bounds__(MODEL_THREADS) gut_selective_adam_kernel(AdamGroups groups, uint32_t N, const float* __restrict__ visibility,
                                                                          float b1, float b2, float eps, int zero_grad) {
    uint32_t gi = 0;
    while (gi + 1 < groups.n_groups && blockIdx.x >= groups.block_start[gi + 1]) ++gi;
    const GutAdamGroup grp = groups.g[gi];
    const uint64_t total = (uint64_t)N * grp.width;
    const uint64_t e0 = 4ull * ((uint64_t)(blockIdx.x - groups.block_start[gi]) * blockDim.x + threadIdx.x);
    if (e0 >= total) return;
    const uint32_t cnt = (uint32_t)min((uint64_t)4, total - e0);
    // rows of the four scalars
    bool vis[4];
    bool any = false;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        vis[k] = false;
        if ((uint32_t)k < cnt) {
            const uint64_t row = (e0 + k) / grp.width;
            vis[k] = visibility ? (__ldg(visibility + row) != 0.0f) : true;
        }
        any = any || vis[k];
    }
    float* gradw = const_cast<float*>(grp.grad);
    if ((groups.vec_ok >> gi) & 1u) {
        if (any) {
            float4 p = *reinterpret_cast<float4*>(grp.param + e0);
            const float4 g = *reinterpret_cast<const float4*>(grp.grad + e0);
            float4 m = *reinterpret_cast<float4*>(grp.exp_avg + e0);
            float4 v = *reinterpret_cast<float4*>(grp.exp_avg_sq + e0);
            if (vis[0]) adam_scalar(p.x, g.x, m.x, v.x, grp.lr, b1, b2, eps);
            if (vis[1]) adam_scalar(p.y, g.y, m.y, v.y, grp.lr, b1, b2, eps);
            if (vis[2]) adam_scalar(p.z, g.z, m.z, v.z, grp.lr, b1, b2, eps);
            if (vis[3]) adam_scalar(p.w, g.w, m.w, v.w, grp.lr, b1, b2, eps);
            *reinterpret_cast<float4*>(grp.param + e0) = p;
            *reinterpret_cast<float4*>(grp.exp_avg + e0) = m;
            *reinterpret_cast<float4*>(grp.exp_avg_sq + e0) = v;
        }
        if (zero_grad) *reinterpret_cast<float4*>(gradw + e0) = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
        for (uint32_t k = 0; k < cnt; ++k) {
            if (vis[k]) {
                float p = grp.param[e0 + k], m = grp.exp_avg[e0 + k], v = grp.exp_avg_sq[e0 + k];
                adam_scalar(p, grp.grad[e0 + k], m, v, grp.lr, b1, b2, eps);
                grp.param[e0 + k] = p; grp.exp_avg[e0 + k] = m; grp.exp_avg_sq[e0 + k] = v;
            }
            if (zero_grad) gradw[e0 + k] = 0.0f;
        }
    }
}

// GSStrategy.update_gradient_buffer (strategy/gs.py:104-115): for particles whose position gradient is non-zero,
//   accum += |grad * |pos - sensor|| / 2 ;  denom += 1
__global__ void __launch_bounds__(MODEL_THREADS) gut_densify_stats_kernel(uint32_t N, const float* __restrict__ pos, const float* __restrict__ pos_g,
                                                                         float3 sensor, float* __restrict__ accum, float* __restrict__ denom) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const size_t i3 = 3 * (size_t)i;
    const float gx = __ldg(pos_g + i3), gy = __ldg(pos_g + i3 + 1), gz = __ldg(pos_g + i3 + 2);
    if (gx == 0.0f && gy == 0.0f && gz == 0.0f) return;
    const float dx = __ldg(pos + i3) - sensor.x, dy = __ldg(pos + i3 + 1) - sensor.y, dz = __ldg(pos + i3 + 2) - sensor.z;
    const float dist = sqrtf(dx * dx + dy * dy + dz * dz);
    const float ax = gx * dist, ay = gy * dist, az = gz * dist;
    accum[i] += sqrtf(ax * ax + ay * ay + az * az) / 2.0f;
    denom[i] += 1.0f;
}

// ---------------------------------------------------------------------------------------------------
// The [N,12] half of the optimiser step as the EPILOGUE of the gradient exchange: the summed [N,12]
// gradient goes through the activation adjoints (gut_activate_backward_kernel's arithmetic) straight into SelectiveAdam for
// positions, rotation, scale and density (adam_scalar: the reference kernel's rounding), and the densification statistics
// of GSStrategy.update_gradient_buffer are taken from the position gradient on the way -- one pass, no raw-gradient tensors.
// Bit-identical to gut_activate_pack_backward -> gut_selective_adam_step -> gut_densify_stats on the same inputs.
// ---------------------------------------------------------------------------------------------------
struct DensityAdam {
    GutAdamTensors pos, rot, scale, dens;
};
__global__ void __launch_bounds__(MODEL_THREADS) gut_density_adam_kernel(uint32_t N, const float4* __restrict__ packed,
                                                                        const float4* __restrict__ grad12, DensityAdam t, int scale_act,
                                                                        int density_act, const float* __restrict__ visibility, float b1,
                                                                        float b2, float eps, int stats, float3 sensor,
                                                                        float* __restrict__ accum, float* __restrict__ denom) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const size_t i3 = 3 * (size_t)i;
    const float4 ga = __ldg(grad12 + i3);
    float px = 0.f, py = 0.f, pz = 0.f;
    bool have_pos = false;
    if (stats && !(ga.x == 0.0f && ga.y == 0.0f && ga.z == 0.0f)) {   // positions BEFORE the update, as post_backward sees them
        px = t.pos.param[i3]; py = t.pos.param[i3 + 1]; pz = t.pos.param[i3 + 2];
        have_pos = true;
        const float dx = px - sensor.x, dy = py - sensor.y, dz = pz - sensor.z;
        const float dist = sqrtf(dx * dx + dy * dy + dz * dz);
        const float ax = ga.x * dist, ay = ga.y * dist, az = ga.z * dist;
        accum[i] += sqrtf(ax * ax + ay * ay + az * az) / 2.0f;
        denom[i] += 1.0f;
    }
    if (visibility && __ldg(visibility + i) == 0.0f) return;   // invisible rows keep param, exp_avg, exp_avg_sq
    const float4 gq = __ldg(grad12 + i3 + 1), gs = __ldg(grad12 + i3 + 2);
    const float4 pa = __ldg(packed + i3), ps = __ldg(packed + i3 + 2);
    // positions
    {
        if (!have_pos) { px = t.pos.param[i3]; py = t.pos.param[i3 + 1]; pz = t.pos.param[i3 + 2]; }
        float m0 = t.pos.exp_avg[i3], m1 = t.pos.exp_avg[i3 + 1], m2 = t.pos.exp_avg[i3 + 2];
        float v0 = t.pos.exp_avg_sq[i3], v1 = t.pos.exp_avg_sq[i3 + 1], v2 = t.pos.exp_avg_sq[i3 + 2];
        adam_scalar(px, ga.x, m0, v0, t.pos.lr, b1, b2, eps);
        adam_scalar(py, ga.y, m1, v1, t.pos.lr, b1, b2, eps);
        adam_scalar(pz, ga.z, m2, v2, t.pos.lr, b1, b2, eps);
        t.pos.param[i3] = px; t.pos.param[i3 + 1] = py; t.pos.param[i3 + 2] = pz;
        t.pos.exp_avg[i3] = m0; t.pos.exp_avg[i3 + 1] = m1; t.pos.exp_avg[i3 + 2] = m2;
        t.pos.exp_avg_sq[i3] = v0; t.pos.exp_avg_sq[i3 + 1] = v1; t.pos.exp_avg_sq[i3 + 2] = v2;
    }
    // density
    {
        float p = t.dens.param[i], m = t.dens.exp_avg[i], v = t.dens.exp_avg_sq[i];
        adam_scalar(p, activate_grad(density_act, pa.w, ga.w), m, v, t.dens.lr, b1, b2, eps);
        t.dens.param[i] = p; t.dens.exp_avg[i] = m; t.dens.exp_avg_sq[i] = v;
    }
    // scale
    {
        const float g0 = activate_grad(scale_act, ps.x, gs.x), g1 = activate_grad(scale_act, ps.y, gs.y), g2 = activate_grad(scale_act, ps.z, gs.z);
        float p0 = t.scale.param[i3], p1 = t.scale.param[i3 + 1], p2 = t.scale.param[i3 + 2];
        float m0 = t.scale.exp_avg[i3], m1 = t.scale.exp_avg[i3 + 1], m2 = t.scale.exp_avg[i3 + 2];
        float v0 = t.scale.exp_avg_sq[i3], v1 = t.scale.exp_avg_sq[i3 + 1], v2 = t.scale.exp_avg_sq[i3 + 2];
        adam_scalar(p0, g0, m0, v0, t.scale.lr, b1, b2, eps);
        adam_scalar(p1, g1, m1, v1, t.scale.lr, b1, b2, eps);
        adam_scalar(p2, g2, m2, v2, t.scale.lr, b1, b2, eps);
        t.scale.param[i3] = p0; t.scale.param[i3 + 1] = p1; t.scale.param[i3 + 2] = p2;
        t.scale.exp_avg[i3] = m0; t.scale.exp_avg[i3 + 1] = m1; t.scale.exp_avg[i3 + 2] = m2;
        t.scale.exp_avg_sq[i3] = v0; t.scale.exp_avg_sq[i3 + 1] = v1; t.scale.exp_avg_sq[i3 + 2] = v2;
    }
    // rotation: adjoint of y = x / max(|x|, eps) (gut_activate_backward_kernel), then Adam on the raw quaternion
    {
        float4 q = *reinterpret_cast<const float4*>(t.rot.param + 4 * (size_t)i);
        const float nn = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
        float4 r;
        if (nn > NORMALIZE_EPS) {
            const float inv = 1.0f / nn;
            const float yx = q.x * inv, yy = q.y * inv, yz = q.z * inv, yw = q.w * inv;
            const float dt = yx * gq.x + yy * gq.y + yz * gq.z + yw * gq.w;
            r = make_float4((gq.x - yx * dt) * inv, (gq.y - yy * dt) * inv, (gq.z - yz * dt) * inv, (gq.w - yw * dt) * inv);
        } else {
            r = make_float4(gq.x / NORMALIZE_EPS, gq.y / NORMALIZE_EPS, gq.z / NORMALIZE_EPS, gq.w / NORMALIZE_EPS);
        }
        float4 m = *reinterpret_cast<float4*>(t.rot.exp_avg + 4 * (size_t)i), v = *reinterpret_cast<float4*>(t.rot.exp_avg_sq + 4 * (size_t)i);
        adam_scalar(q.x, r.x, m.x, v.x, t.rot.lr, b1, b2, eps);
        adam_scalar(q.y, r.y, m.y, v.y, t.rot.lr, b1, b2, eps);
        adam_scalar(q.z, r.z, m.z, v.z, t.rot.lr, b1, b2, eps);
        adam_scalar(q.w, r.w, m.w, v.w, t.rot.lr, b1, b2, eps);
        *reinterpret_cast<float4*>(t.rot.param + 4 * (size_t)i) = q;
        *reinterpret_cast<float4*>(t.rot.exp_avg + 4 * (size_t)i) = m;
        *reinterpret_cast<float4*>(t.rot.exp_avg_sq + 4 * (size_t)i) = v;
    }
}

inline uint32_t blocks_for(uint64_t n) { return (uint32_t)((n + MODEL_THREADS - 1) / MODEL_THREADS); }

}  // namespace

cudaError_t gut_launch_activate_pack(uint32_t N, const float* pos, const float* rot, const float* scale, const float* dens, int scale_act,
                                     int density_act, float* out12, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    gut_activate_pack_kernel<<<blocks_for(N), MODEL_THREADS, 0, stream>>>(N, pos, reinterpret_cast<const float4*>(rot), scale, dens, scale_act,
                                                                          density_act, reinterpret_cast<float4*>(out12));
    return cudaGetLastError();
}

cudaError_t gut_launch_activate_backward(uint32_t N, const float* rot, const float* packed12, const float* grad12, int scale_act, int density_act,
                                         float* pos_g, float* rot_g, float* scale_g, float* dens_g, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    gut_activate_backward_kernel<<<blocks_for(N), MODEL_THREADS, 0, stream>>>(
        N, reinterpret_cast<const float4*>(rot), reinterpret_cast<const float4*>(packed12), reinterpret_cast<const float4*>(grad12), scale_act,
        density_act, pos_g, reinterpret_cast<float4*>(rot_g), scale_g, dens_g);
    return cudaGetLastError();
}

cudaError_t gut_launch_pack_features(uint32_t N, uint32_t ncoef, const float* albedo, const float* specular, float* out, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    if (ncoef == 16) {
        const uint64_t n_vec = 12ull * N;
        gut_pack_features_kernel<<<blocks_for(n_vec), MODEL_THREADS, 0, stream>>>(n_vec, albedo, specular, reinterpret_cast<float4*>(out));
    } else {
        const uint64_t total = 3ull * ncoef * N;
        gut_features_generic_kernel<true><<<blocks_for(total), MODEL_THREADS, 0, stream>>>(total, 3 * ncoef, const_cast<float*>(albedo),
                                                                                          const_cast<float*>(specular), out);
    }
    return cudaGetLastError();
}

cudaError_t gut_launch_unpack_features(uint32_t N, uint32_t ncoef, const float* grad, float* albedo_g, float* specular_g, cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    if (ncoef == 16) {
        const uint64_t n_vec = 12ull * N;
        gut_unpack_features_kernel<<<blocks_for(n_vec), MODEL_THREADS, 0, stream>>>(n_vec, reinterpret_cast<const float4*>(grad), albedo_g, specular_g);
    } else {
        const uint64_t total = 3ull * ncoef * N;
        gut_features_generic_kernel<false><<<blocks_for(total), MODEL_THREADS, 0, stream>>>(total, 3 * ncoef, albedo_g, specular_g,
                                                                                           const_cast<float*>(grad));
    }
    return cudaGetLastError();
}

cudaError_t gut_launch_selective_adam(uint32_t N, const GutAdamGroup* groups, uint32_t n_groups, const float* visibility, float b1, float b2,
                                      float eps, int zero_grad, cudaStream_t stream) {
    if (N == 0 || n_groups == 0) return cudaSuccess;
    AdamGroups a = {};
    a.n_groups = n_groups;
    uint64_t blocks = 0;
    for (uint32_t i = 0; i < n_groups; ++i) {
        a.g[i] = groups[i];
        const uintptr_t bits = (uintptr_t)groups[i].param | (uintptr_t)groups[i].grad | (uintptr_t)groups[i].exp_avg | (uintptr_t)groups[i].exp_avg_sq;
        if ((bits & 15) == 0 && (((uint64_t)N * groups[i].width) & 3) == 0) a.vec_ok |= 1u << i;
        a.block_start[i] = (uint32_t)blocks;
        blocks += blocks_for(((uint64_t)N * groups[i].width + 3) / 4);
    }
    if (blocks > 0x7fffffffull) return cudaErrorInvalidConfiguration;
    a.block_start[n_groups] = (uint32_t)blocks;
    gut_selective_adam_kernel<<<(uint32_t)blocks, MODEL_THREADS, 0, stream>>>(a, N, visibility, b1, b2, eps, zero_grad);
    return cudaGetLastError();
}

cudaError_t gut_launch_densify_stats(uint32_t N, const float* pos, const float* pos_g, float3 sensor, float* accum, float* denom,
                                     cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    gut_densify_stats_kernel<<<blocks_for(N), MODEL_THREADS, 0, stream>>>(N, pos, pos_g, sensor, accum, denom);
    return cudaGetLastError();
}

cudaError_t gut_launch_density_adam(uint32_t N, const float* packed12, const float* grad12, const GutAdamTensors& pos, const GutAdamTensors& rot,
                                    const GutAdamTensors& scale, const GutAdamTensors& dens, int scale_act, int density_act,
                                    const float* visibility, float b1, float b2, float eps, const float* sensor, float* accum, float* denom,
                                    cudaStream_t stream) {
    if (N == 0) return cudaSuccess;
    DensityAdam t{pos, rot, scale, dens};
    const int stats = (sensor && accum && denom) ? 1 : 0;
    const float3 sp = stats ? make_float3(sensor[0], sensor[1], sensor[2]) : make_float3(0.f, 0.f, 0.f);
    gut_density_adam_kernel<<<blocks_for(N), MODEL_THREADS, 0, stream>>>(N, reinterpret_cast<const float4*>(packed12),
                                                                       reinterpret_cast<const float4*>(grad12), t, scale_act, density_act,
                                                                       visibility, b1, b2, eps, stats, sp, accum, denom);
    return cudaGetLastError();
}
